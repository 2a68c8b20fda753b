"""Fused sm_100a forward/backward of the ICA-LSTM (filled in by ops.lstm / ops.head)."""


def supported(model, x):
    return False


def icalstm_forward(model, x):
    raise NotImplementedError
