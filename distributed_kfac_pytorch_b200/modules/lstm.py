"""LSTM assembled from ``nn.Linear`` so that K-FAC's module hooks can see every
matrix multiply.  Class names, constructor arguments and parameter names follow
the reference's ``kfac/modules/lstm.py`` (cites into /root/reference) so that
checkpoints and the examples' ``LSTMModel`` interchange; the model code itself is
not part of the accelerated path."""
import torch
import torch.nn as nn


class LSTMCellBase(nn.Module):
    """Marker base class: K-FAC registers every child ``nn.Linear`` of a subclass
    as a LinearMultiLayer (kfac/layers/__init__.py:29-30)."""

    def __init__(self, input_size, hidden_size, bias=True):
        super().__init__()
        self.input_size = input_size
        self.hidden_size = hidden_size
        self.bias = bias

    def forward(self, input, hidden):
        raise NotImplementedError

    def __repr__(self):
        return 'LSTMCell(input_size={}, hidden_size={}, bias={})'.format(
            self.input_size, self.hidden_size, self.bias)


def _lstm_pointwise(i, f, g, o, c):
    c_next = torch.sigmoid(f) * c + torch.sigmoid(i) * torch.tanh(g)
    return torch.sigmoid(o) * torch.tanh(c_next), c_next


class LSTMCell(LSTMCellBase):
    """Two fused-gate Linears, ``linear_ih`` and ``linear_hh`` (lstm.py:71-88)."""

    def __init__(self, *args, **kwargs):
        super().__init__(*args, **kwargs)
        self.linear_ih = nn.Linear(self.input_size, 4 * self.hidden_size, bias=self.bias)
        self.linear_hh = nn.Linear(self.hidden_size, 4 * self.hidden_size, bias=self.bias)

    def forward(self, input, hidden):
        h, c = hidden
        i, f, g, o = (self.linear_ih(input) + self.linear_hh(h)).chunk(4, 1)
        return _lstm_pointwise(i, f, g, o, c)


class LSTMCellKFAC(LSTMCellBase):
    """One Linear per gate and source: eight small factors instead of two large
    ones (lstm.py:41-68)."""

    def __init__(self, *args, **kwargs):
        super().__init__(*args, **kwargs)
        for gate in 'ifgo':
            setattr(self, 'linear_%s_i' % gate,
                    nn.Linear(self.input_size, self.hidden_size, bias=self.bias))
            setattr(self, 'linear_%s_h' % gate,
                    nn.Linear(self.hidden_size, self.hidden_size, bias=self.bias))

    def forward(self, input, hidden=None):
        h, c = hidden
        pre = [getattr(self, 'linear_%s_i' % gate)(input) + getattr(self, 'linear_%s_h' % gate)(h)
               for gate in 'ifgo']
        return _lstm_pointwise(pre[0], pre[1], pre[2], pre[3], c)


class LSTMLayer(nn.Module):
    """Unrolls one cell over the sequence.  Like the reference (lstm.py:107) the
    step index always addresses dim 0 of ``input``."""

    def __init__(self, input_size, hidden_size, bias=True, batch_first=False, reverse=False):
        super().__init__()
        self.input_size = input_size
        self.hidden_size = hidden_size
        self.bias = bias
        self.batch_first = batch_first
        self.reverse = reverse
        self.cell = LSTMCell(input_size, hidden_size, bias=bias)
        self.seq_dim = 1 if batch_first else 0

    def forward(self, input, hidden):
        n = input.size(self.seq_dim)
        order = range(n - 1, -1, -1) if self.reverse else range(n)
        outs = []
        for t in order:
            hidden = self.cell(input[t], hidden)
            outs.append(hidden[0])
        if self.reverse:
            outs.reverse()
        return torch.stack(outs, self.seq_dim), hidden

    def __repr__(self):
        return ('LSTMLayer(input_size={}, hidden_size={}, bias={}, batch_first={}, '
                'reverse={})').format(self.input_size, self.hidden_size, self.bias,
                                      self.batch_first, self.reverse)


class LSTM(nn.Module):
    """Multi-layer (optionally bidirectional) LSTM over LSTMLayer (lstm.py:120-225).
    ``self.layers[i]`` is a ModuleList of one or two directions, as in the reference,
    so parameter names match."""

    def __init__(self, input_size, hidden_size, num_layers=1, bias=True, batch_first=False,
                 dropout=0.0, bidirectional=False):
        super().__init__()
        self.input_size = input_size
        self.hidden_size = hidden_size
        self.num_layers = num_layers
        self.bias = bias
        self.batch_first = batch_first
        self.dropout = dropout
        self.bidirectional = bidirectional
        self.num_directions = 2 if bidirectional else 1
        stacks = []
        for _ in range(num_layers):
            dirs = [LSTMLayer(input_size, hidden_size, bias=bias, batch_first=batch_first)]
            if bidirectional:
                dirs.append(LSTMLayer(input_size, hidden_size, bias=bias,
                                      batch_first=batch_first, reverse=True))
            stacks.append(nn.ModuleList(dirs))
        self.layers = nn.ModuleList(stacks)
        self.drop = nn.Dropout(dropout) if dropout > 0 and num_layers > 1 else None

    def forward(self, input, hx=None):
        if isinstance(input, nn.utils.rnn.PackedSequence):
            raise NotImplementedError('PackedSequence input is not supported')
        batch = input.size(0) if self.batch_first else input.size(1)
        if hx is None:
            z = input.new_zeros(self.num_layers * self.num_directions, batch, self.hidden_size)
            hx = (z, z)
        states = list(zip(*hx))
        finals = []
        x = input
        for i, dirs in enumerate(self.layers):
            outs = []
            for j, direction in enumerate(dirs):
                out, last = direction(x, states[i * self.num_directions + j])
                outs.append(out)
                finals.append(last)
            x = torch.cat(outs, -1)
            if self.drop is not None and i < self.num_layers - 1:
                x = self.drop(x)
        hs, cs = zip(*finals)
        return x, (torch.stack(hs), torch.stack(cs))
