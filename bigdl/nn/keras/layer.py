from bigdl_b200.nn.keras.layers import *  # noqa: F401,F403
from bigdl_b200.nn.keras.layers import Input, KerasLayer, merge  # noqa: F401


class InferShape:
    """Mixin of the reference's Keras layers: shapes in Keras convention (batch dimension None)."""

    @staticmethod
    def _keras(shape):
        return tuple([None] + list(shape)[1:])

    def _shapes(self, s):
        if s and isinstance(s[0], (list, tuple)):
            return [self._keras(x) for x in s]
        return self._keras(s)

    def get_input_shape(self):
        return self._shapes(self.getInputShape())

    def get_output_shape(self):
        return self._shapes(self.getOutputShape())


class KerasCreator:
    """The reference derives the JVM factory name from the class name; there is no JVM here."""

    def jvm_class_constructor(self):
        return "createKeras" + self.__class__.__name__


for _n in ("get_input_shape", "get_output_shape"):
    if not hasattr(KerasLayer, _n):
        setattr(KerasLayer, _n, getattr(InferShape, _n))
for _n in ("_keras", "_shapes"):
    if not hasattr(KerasLayer, _n):
        setattr(KerasLayer, _n, InferShape.__dict__[_n])
