"""Q-network hook of the RL loop: a 1-D circular-convolution residual U-Net, one Q-value per "flip cell i" action.

Stays plain PyTorch/cuDNN (north_star: the UNet surrogate hook is unchanged; it is not on the accelerated path).  The module
is built from a stage table, but its parameter names and shapes equal the reference's ``ShallowUQNet`` (model.py:31-224:
``conv{s}_1`` plain Conv1d, ``conv{s}_{2..6}.convrelu.{0,1}`` Conv1d + BatchNorm1d, ``conv11_fin``), so the pretrained
``state_dict`` the reference loads in ``Callbacks.on_create_policy`` (main.py:48-53) loads here unchanged.
1 974 113 parameters for the 256-cell default.
"""
from __future__ import annotations

import torch
from torch import nn

# (stage id, input channels, output channels); encoder stages pool afterwards, decoder stages up-sample + concatenate before
_ENCODER = [(1, 1, 16), (2, 16, 32), (3, 32, 64), (4, 64, 128)]
_BOTTOM = (6, 128, 256)
_DECODER = [(8, 128 + 256, 128), (9, 64 + 128, 64), (10, 32 + 64, 32), (11, 16 + 32, 16)]


def _conv(cin, cout):
    return nn.Conv1d(cin, cout, 3, padding="same", bias=True, padding_mode="circular")


class _ConvBNReLU(nn.Module):
    def __init__(self, cin, cout):
        super().__init__()
        self.convrelu = nn.Sequential(_conv(cin, cout), nn.BatchNorm1d(cout), nn.ReLU(inplace=True))

    def forward(self, x):
        return self.convrelu(x)


class ShallowUQNet(nn.Module):
    """obs (B, 1, n) or (B, n) float -> (B, n) Q-values."""

    def __init__(self, n_cells: int = 256):
        super().__init__()
        self.ncells = n_cells
        for sid, cin, cout in _ENCODER + [_BOTTOM] + _DECODER:
            setattr(self, f"conv{sid}_1", _conv(cin, cout))
            for k in range(2, 7):
                setattr(self, f"conv{sid}_{k}", _ConvBNReLU(cout, cout))
        self.conv11_fin = _conv(16, 1)
        self.pool = nn.MaxPool1d(2)
        self.up = nn.Upsample(scale_factor=2)

    def _stage(self, sid, x):
        """conv_1 -> (conv_2, conv_3) + skip -> conv_4 -> (conv_5, conv_6) + skip   (model.py:135-140 pattern)"""
        layer = lambda k: getattr(self, f"conv{sid}_{k}")
        r1 = layer(1)(x)
        t = layer(3)(layer(2)(r1)) + r1
        r2 = layer(4)(t)
        return layer(6)(layer(5)(r2)) + r2

    def forward(self, obs):
        x = obs if obs.dim() == 3 else obs.unsqueeze(1)
        x = x.to(self.conv1_1.weight.dtype)
        skips = []
        for sid, _, _ in _ENCODER:
            x = self._stage(sid, x)
            skips.append(x)
            x = self.pool(x)
        x = self._stage(_BOTTOM[0], x)
        for sid, _, _ in _DECODER:
            x = torch.cat([self.up(x), skips.pop()], dim=1)
            x = self._stage(sid, x)
        return self.conv11_fin(x).flatten(1)


class FCNQNet(nn.Module):
    """The reference's MLP alternatives (model.py:227-313): n -> hidden -> hidden -> n with LeakyReLU(0.1); ``hidden`` is 2n for
    ``FCNQNet`` and 1200 for ``FCNQNet_heavy``.  Parameter names fc1/fc2/fc3 as in the reference."""

    def __init__(self, n_cells: int = 256, hidden: int | None = None):
        super().__init__()
        hidden = hidden or 2 * n_cells
        self.fc1 = nn.Linear(n_cells, hidden)
        self.fc2 = nn.Linear(hidden, hidden)
        self.fc3 = nn.Linear(hidden, n_cells)
        self.m = nn.LeakyReLU(0.1)

    def forward(self, obs):
        x = obs.to(self.fc1.weight.dtype)
        return self.fc3(self.m(self.fc2(self.m(self.fc1(x))))).flatten(1)


def load_pretrained(model: nn.Module, path: str | None) -> bool:
    """main.py:48-53: load the pretrained state_dict when the file exists (it is not shipped with the reference tree:
    .MISSING_LARGE_BLOBS); otherwise keep the random initialisation.  Returns whether weights were loaded."""
    import os
    if not path or not os.path.exists(path):
        return False
    sd = torch.load(path, map_location="cpu")
    if isinstance(sd, dict) and "state_dict" in sd:
        sd = sd["state_dict"]
    model.load_state_dict(sd)
    return True
