from .msannet import MSANNet
from .icalstm import ICALstm, LSTM, LSTMCell

__all__ = ["MSANNet", "ICALstm", "LSTM", "LSTMCell"]
