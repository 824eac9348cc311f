"""Language models: SimpleRNN (S/models/rnn/SimpleRNN.scala) and the PTB word-level LSTM
(S/example/languagemodel/PTBModel.scala)."""
from ..nn.layer import (LSTM, Dropout, Linear, LogSoftMax, LookupTable, Recurrent, RnnCell, Sequential, Tanh,
                        TimeDistributed)


def SimpleRNN(input_size: int, hidden_size: int, output_size: int):
    model = Sequential()
    model.add(Recurrent().add(RnnCell(input_size, hidden_size, Tanh())))
    model.add(TimeDistributed(Linear(hidden_size, output_size)))
    return model


def PTBModel(input_size: int, hidden_size: int, output_size: int, num_layers: int = 2, keep_prob: float = 2.0):
    """Embedding → num_layers × LSTM → TimeDistributed(Linear) → TimeDistributed(LogSoftMax)."""
    model = Sequential()
    model.add(LookupTable(input_size, hidden_size))
    if keep_prob < 1:
        model.add(Dropout(1 - keep_prob))
    for _ in range(num_layers):
        model.add(Recurrent().add(LSTM(hidden_size, hidden_size)))
        if keep_prob < 1:
            model.add(Dropout(1 - keep_prob))
    model.add(TimeDistributed(Linear(hidden_size, output_size)))
    model.add(TimeDistributed(LogSoftMax()))
    return model
