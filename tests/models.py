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


MODELS = {'convnet': SmallConvNet, 'mlp': SmallMLP}


def make_batch(name: str, step: int, rank: int = 0) -> tuple[torch.Tensor, torch.Tensor]:
    """Deterministic synthetic batch for (model, step, rank)."""
    g = torch.Generator().manual_seed(1000 * step + 17 * rank + 5)
    if name == 'convnet':
        x = torch.randn(6, 3, 10, 10, generator=g)
        y = torch.randint(0, 10, (6,), generator=g)
    else:
        x = torch.randn(4, 5, 12, generator=g)
        y = torch.randint(0, 7, (4, 5), generator=g)
    return x, y


def loss_fn(name: str, out: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
    if name == 'convnet':
        return nn.functional.cross_entropy(out, y)
    return nn.functional.cross_entropy(out.reshape(-1, out.shape[-1]), y.reshape(-1))
