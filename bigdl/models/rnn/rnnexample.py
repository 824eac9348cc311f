from bigdl_b200.models.rnn import SimpleRNN as build_model  # noqa: F401
