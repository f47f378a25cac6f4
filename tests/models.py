"""Small models shared by the golden-vector generator and the parity tests (weights travel in the fixtures)."""
from __future__ import annotations

import torch
from torch import nn


class SmallConvNet(nn.Module):
    """Conv(bias) -> Conv(stride 2, no bias) -> 1x1 conv -> Linear(bias): covers every factor kernel variant."""

    def __init__(self) -> None:
        super().__init__()
        self.conv1 = nn.Conv2d(3, 8, 3, padding=1, bias=True)
        self.conv2 = nn.Conv2d(8, 12, 3, stride=2, padding=1, bias=False)
        self.conv3 = nn.Conv2d(12, 16, 1, bias=True)
        self.fc = nn.Linear(16, 10, bias=True)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = torch.relu(self.conv1(x))
        x = torch.relu(self.conv2(x))
        x = torch.relu(self.conv3(x))
        x = x.mean(dim=(2, 3))
        return self.fc(x)


class SmallMLP(nn.Module):
    """Linear without bias -> Linear with bias, 3-D input (tokens) so the leading dims get flattened."""

    def __init__(self) -> None:
        super().__init__()
        self.l1 = nn.Linear(12, 24, bias=False)
        self.l2 = nn.Linear(24, 7, bias=True)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return self.l2(torch.tanh(self.l1(x)))


class _BasicBlock(nn.Module):
    """3x3 conv - BN - ReLU - 3x3 conv - BN, identity shortcut (parameter-free: stride-2 subsampling + zero
    channel padding where the shape changes), as in the CIFAR ResNets of He et al. 2015, section 4.2."""

    def __init__(self, cin: int, cout: int, stride: int) -> None:
        super().__init__()
        self.conv1 = nn.Conv2d(cin, cout, 3, stride=stride, padding=1, bias=False)
        self.bn1 = nn.BatchNorm2d(cout)
        self.conv2 = nn.Conv2d(cout, cout, 3, padding=1, bias=False)
        self.bn2 = nn.BatchNorm2d(cout)
        self.pad = (cout - cin) // 2 if (stride != 1 or cin != cout) else 0
        self.stride = stride

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        out = torch.relu(self.bn1(self.conv1(x)))
        out = self.bn2(self.conv2(out))
        sc = x
        if self.pad or self.stride != 1:
            sc = nn.functional.pad(x[:, :, ::self.stride, ::self.stride], (0, 0, 0, 0, self.pad, self.pad))
        return torch.relu(out + sc)


class CifarResNet32(nn.Module):
    """BASELINE.json configs[0]: ResNet-32 for 32x32 inputs (3 stages x 5 basic blocks, 16/32/64 channels): 31
    bias-free 3x3 convolutions + one Linear(64, 10) = the 32 K-FAC layers of SURVEY.md section 8d, config 1:
    (A, G) = (27,16) x1, (144,16) x10, (144,32) x1, (288,32) x9, (288,64) x1, (576,64) x9, (65,10) x1."""

    def __init__(self) -> None:
        super().__init__()
        self.conv1 = nn.Conv2d(3, 16, 3, padding=1, bias=False)
        self.bn1 = nn.BatchNorm2d(16)
        blocks, cin = [], 16
        for cout, stride in ((16, 1), (32, 2), (64, 2)):
            for b in range(5):
                blocks.append(_BasicBlock(cin, cout, stride if b == 0 else 1))
                cin = cout
        self.blocks = nn.Sequential(*blocks)
        self.fc = nn.Linear(64, 10)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = torch.relu(self.bn1(self.conv1(x)))
        x = self.blocks(x)
        return self.fc(x.mean(dim=(2, 3)))


MODELS = {'convnet': SmallConvNet, 'mlp': SmallMLP, 'resnet32': CifarResNet32}


def make_batch(name: str, step: int, rank: int = 0) -> tuple[torch.Tensor, torch.Tensor]:
    """Deterministic synthetic batch for (model, step, rank)."""
    g = torch.Generator().manual_seed(1000 * step + 17 * rank + 5)
    if name == 'convnet':
        x = torch.randn(6, 3, 10, 10, generator=g)
        y = torch.randint(0, 10, (6,), generator=g)
    elif name == 'resnet32':   # config 1: synthetic 32x32, batch 128
        x = torch.randn(128, 3, 32, 32, generator=g)
        y = torch.randint(0, 10, (128,), generator=g)
    else:
        x = torch.randn(4, 5, 12, generator=g)
        y = torch.randint(0, 7, (4, 5), generator=g)
    return x, y


def loss_fn(name: str, out: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
    if name in ('convnet', 'resnet32'):
        return nn.functional.cross_entropy(out, y)
    return nn.functional.cross_entropy(out.reshape(-1, out.shape[-1]), y.reshape(-1))
