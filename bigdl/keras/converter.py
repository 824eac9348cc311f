from bigdl_b200.interop.keras_converter import DefinitionLoader, WeightLoader, load_keras  # noqa: F401


LayerConverter = DefinitionLoader          # the reference splits per-layer creation / weight conversion into these names
WeightsConverter = WeightLoader


def unsupport_exp(name):
    raise Exception("We don't support %s for now" % name)
