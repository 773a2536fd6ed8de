"""Small seeded models shared by the golden generator (runs against the reference)
and by the tests (run against this repo).  No dependency on either K-FAC package
except for the LSTM case, which takes the LSTM class as an argument because the
reference and this repo each ship their own ``kfac.modules.LSTM``."""
import torch
import torch.nn as nn
import torch.nn.functional as F


class SmallCNN(nn.Module):
    """Covers every conv geometry class of ResNet-50 in miniature:
    3x3 s1 p1 with bias, 3x3 s2 p1, 1x1 s1, 1x1 s2, 5x5 s2 p2 (conv1-like) and a
    Linear with bias.  Non-square input (H != W) catches transposition bugs."""

    def __init__(self):
        super().__init__()
        self.c1 = nn.Conv2d(3, 8, 5, stride=2, padding=2, bias=False)     # da=75
        self.c2 = nn.Conv2d(8, 16, 3, stride=1, padding=1, bias=True)     # da=73
        self.c3 = nn.Conv2d(16, 12, 1, stride=1, bias=False)              # da=16
        self.c4 = nn.Conv2d(12, 20, 3, stride=2, padding=1, bias=False)   # da=108
        self.c5 = nn.Conv2d(20, 24, 1, stride=2, bias=False)              # da=20
        self.fc = nn.Linear(24, 10)                                       # da=25

    def forward(self, x):
        x = F.relu(self.c1(x))
        x = F.relu(self.c2(x))
        x = F.relu(self.c3(x))
        x = F.relu(self.c4(x))
        x = F.relu(self.c5(x))
        x = F.adaptive_avg_pool2d(x, 1).flatten(1)
        return self.fc(x)


class _BasicBlock(nn.Module):
    def __init__(self, cin, cout, stride):
        super().__init__()
        self.conv1 = nn.Conv2d(cin, cout, 3, stride=stride, padding=1, bias=False)
        self.bn1 = nn.BatchNorm2d(cout)
        self.conv2 = nn.Conv2d(cout, cout, 3, stride=1, padding=1, bias=False)
        self.bn2 = nn.BatchNorm2d(cout)
        self.pad = (cout - cin) // 2 if (stride != 1 or cin != cout) else None
        self.stride = stride

    def forward(self, x):
        out = F.relu(self.bn1(self.conv1(x)))
        out = self.bn2(self.conv2(out))
        if self.pad is not None:      # option-A shortcut: subsample and zero-pad channels
            x = F.pad(x[:, :, ::self.stride, ::self.stride], (0, 0, 0, 0, self.pad, self.pad))
        return F.relu(out + x)


class CifarResNet32(nn.Module):
    """The standard 32-layer CIFAR ResNet (He et al. 2015, 3 stages of 5 basic blocks, 16/32/64
    channels, option-A shortcuts, 0.46 M parameters): BASELINE.json configs[0].  32 K-FAC layers
    (31 convs without bias + the biased classifier), factor dims from (27,16) to (576,64)+(65,10)."""

    def __init__(self, num_classes=10):
        super().__init__()
        self.conv1 = nn.Conv2d(3, 16, 3, stride=1, padding=1, bias=False)
        self.bn1 = nn.BatchNorm2d(16)
        blocks, cin = [], 16
        for cout, stride in ((16, 1), (32, 2), (64, 2)):
            for i in range(5):
                blocks.append(_BasicBlock(cin, cout, stride if i == 0 else 1))
                cin = cout
        self.blocks = nn.Sequential(*blocks)
        self.linear = nn.Linear(64, num_classes)
        for m in self.modules():
            if isinstance(m, (nn.Conv2d, nn.Linear)):
                nn.init.kaiming_normal_(m.weight)

    def forward(self, x):
        x = F.relu(self.bn1(self.conv1(x)))
        x = self.blocks(x)
        x = F.adaptive_avg_pool2d(x, 1).flatten(1)
        return self.linear(x)


def make_resnet32(seed=0):
    torch.manual_seed(seed)
    return CifarResNet32()


def cifar_batch(step, n=128):
    g = torch.Generator().manual_seed(3000 + step)
    return torch.randn(n, 3, 32, 32, generator=g), torch.randint(0, 10, (n,), generator=g)


CNN_INPUT = (6, 3, 24, 20)      # N, C, H, W
CNN_CLASSES = 10


class SmallLSTMModel(nn.Module):
    """Embedding -> K-FAC-friendly LSTM -> decoder, seq-first, like
    examples/rnn_utils/lstm.py but tiny."""

    def __init__(self, lstm_cls, ntoken=30, ninp=12, nhid=12):
        super().__init__()
        self.encoder = nn.Embedding(ntoken, ninp)
        self.rnn = lstm_cls(ninp, nhid, num_layers=2, batch_first=False)
        self.decoder = nn.Linear(nhid, ntoken)
        self.ntoken = ntoken

    def forward(self, tokens):
        x = self.encoder(tokens)
        x, _ = self.rnn(x)
        return F.log_softmax(self.decoder(x).view(-1, self.ntoken), dim=1)


LSTM_SEQ, LSTM_BATCH, LSTM_NTOKEN = 5, 4, 30


class FixedScaler(object):
    """Stand-in for torch.cuda.amp.GradScaler with a constant loss scale (K-FAC only calls
    get_scale(): kfac/layers/base.py:374-375)."""

    def __init__(self, scale):
        self.scale = scale

    def get_scale(self):
        return self.scale


class TiedLinear(nn.Linear):
    """A Linear whose class name K-FAC's registration does not know, so that it is only
    registered by hand through register_shared_module."""


class TiedModel(nn.Module):
    """Two square Linear modules sharing ONE weight (and bias), then an ordinary classifier."""

    def __init__(self, width=12, classes=10):
        super().__init__()
        self.first = TiedLinear(width, width)
        self.second = TiedLinear(width, width)
        self.second.weight = self.first.weight
        self.second.bias = self.first.bias
        self.fc = nn.Linear(width, classes)

    def forward(self, x):
        return self.fc(F.relu(self.second(F.relu(self.first(x)))))


def make_tied(seed=0):
    torch.manual_seed(seed)
    return TiedModel()


def tied_batch(step, n=8):
    g = torch.Generator().manual_seed(4000 + step)
    return torch.randn(n, 12, generator=g), torch.randint(0, 10, (n,), generator=g)


def poison_captures(layers):
    """What an fp16 overflow does to K-FAC's captures: layer 1 loses the FIRST of its two accumulated
    output-gradient captures (one inf), layer 3 loses BOTH (inf and nan), so its G must stay as it is.
    Works on the reference's KFACLayer and on this repository's alike: g_outputs holds (tensor, scale)."""
    def spoil(layer, idx, value):
        g, scale = layer.g_outputs[idx]
        g = g.clone()
        g.view(-1)[3] = value
        layer.g_outputs[idx] = (g, scale)
    spoil(layers[1], 0, float('inf'))
    spoil(layers[3], 0, float('inf'))
    spoil(layers[3], 1, float('nan'))


def make_cnn(seed=0):
    torch.manual_seed(seed)
    return SmallCNN()


def make_lstm(lstm_cls, seed=0):
    torch.manual_seed(seed)
    return SmallLSTMModel(lstm_cls)


def cnn_batch(step, rank=0):
    g = torch.Generator().manual_seed(1000 + 17 * step + 100 * rank)
    x = torch.randn(*CNN_INPUT, generator=g)
    y = torch.randint(0, CNN_CLASSES, (CNN_INPUT[0],), generator=g)
    return x, y


def lstm_batch(step, rank=0):
    g = torch.Generator().manual_seed(2000 + 17 * step + 100 * rank)
    x = torch.randint(0, LSTM_NTOKEN, (LSTM_SEQ, LSTM_BATCH), generator=g)
    y = torch.randint(0, LSTM_NTOKEN, (LSTM_SEQ * LSTM_BATCH,), generator=g)
    return x, y


def resnet_factor_dims(name):
    """(da, dg) per K-FAC layer of torchvision resnet50/152 in registration order
    (kfac/preconditioner.py:382-402 walks named_children in definition order)."""
    import torchvision
    model = getattr(torchvision.models, name)()
    dims = []

    def walk(m):
        for child in m.children():
            cname = child.__class__.__name__.lower()
            if cname == 'conv2d':
                da = child.in_channels * child.kernel_size[0] * child.kernel_size[1]
                dims.append((da + (child.bias is not None), child.out_channels))
            elif cname == 'linear':
                dims.append((child.in_features + (child.bias is not None), child.out_features))
            else:
                walk(child)
    walk(model)
    return dims
