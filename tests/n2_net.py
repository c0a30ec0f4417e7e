"""Small quantized network shared by oracle/gen_golden_n2.py (built on the unmodified reference's modules) and
tests/test_helpers_gpu.py (built on the CUDA modules): QConv2d 3x3 + bias -> RangeBN -> ReLU -> QConv2d 1x1 -> avg-pool -> QLinear."""
import torch.nn as nn


class N2Net(nn.Module):
    def __init__(self, q):
        super().__init__()
        kw = dict(num_bits=8, num_bits_weight=8, num_bits_grad=8, biprecision=True)
        self.conv1 = q.QConv2d(16, 32, 3, padding=1, bias=True, **kw)
        self.bn1 = q.RangeBN(32, num_bits=8, num_bits_grad=8)
        self.relu = nn.ReLU()
        self.conv2 = q.QConv2d(32, 16, 1, bias=False, **kw)
        self.pool = nn.AvgPool2d(8)
        self.fc = q.QLinear(16, 10, **kw)

    def forward(self, x):
        x = self.relu(self.bn1(self.conv1(x)))
        x = self.pool(self.conv2(x))
        return self.fc(x.view(x.size(0), -1))
