from .lstm import LSTM, LSTMCellBase, LSTMCellKFAC, LSTMCell, LSTMLayer

__all__ = ['LSTM', 'LSTMCellBase', 'LSTMCellKFAC', 'LSTMCell', 'LSTMLayer']
