"""DeepSDF auto-decoder module (host-side container for the decoder parameters).

Mirrors the interface of the reference's ``deep_sdf/deep_sdf_decoder.py:9-110``
(`Decoder(latent_size, dims, ...)`, sub-modules named ``lin0 .. linK`` so that a
DeepSDF checkpoint's ``state_dict`` loads unchanged).  On the B200 path this
module is *only* a parameter container: `reconstruct.optimizer.Optimizer`
reads its ``state_dict`` once, folds the weight-norm and hands packed tiles to
the CUDA library.  ``forward`` exists so the module can be trained / evaluated
with stock PyTorch (it is what `tools/make_wfit.py` trains and what the tests
use as the torch-fp32 cross-check); it is never called by the product path.

Architecture (reference `deep_sdf_decoder.py:28-72`):
  widths = [latent+3, *dims, 1]; a layer whose *successor* index is listed in
  ``latent_in`` emits ``width - (latent+3)`` features so that the successor's
  input is ``cat([hidden, net_input])``; ReLU after every layer but the last;
  one final tanh.
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F


class Decoder(nn.Module):
    def __init__(self, latent_size, dims, dropout=None, dropout_prob=0.0,
                 norm_layers=(), latent_in=(), weight_norm=False,
                 xyz_in_all=None, use_tanh=False, latent_dropout=False):
        super().__init__()
        in_width = latent_size + 3
        widths = [in_width] + list(dims) + [1]
        self.num_layers = len(widths)
        self.latent_size = latent_size
        self.norm_layers = tuple(norm_layers) if norm_layers is not None else None
        self.latent_in = tuple(latent_in)
        self.weight_norm = bool(weight_norm)
        self.xyz_in_all = xyz_in_all
        self.use_tanh = use_tanh
        self.latent_dropout = latent_dropout
        self.dropout = tuple(dropout) if dropout is not None else None
        self.dropout_prob = dropout_prob

        for k in range(self.num_layers - 1):
            fan_out = widths[k + 1]
            if (k + 1) in self.latent_in:
                fan_out -= in_width
            elif xyz_in_all and k != self.num_layers - 2:
                fan_out -= 3
            lin = nn.Linear(widths[k], fan_out)
            normed = self.norm_layers is not None and k in self.norm_layers
            if normed and self.weight_norm:
                lin = nn.utils.weight_norm(lin)
            self.add_module("lin%d" % k, lin)
            if normed and not self.weight_norm:
                self.add_module("bn%d" % k, nn.LayerNorm(fan_out))

    def forward(self, net_input):
        xyz = net_input[..., -3:]
        h = net_input
        if self.latent_dropout and net_input.shape[-1] > 3:
            lat = F.dropout(net_input[..., :-3], p=0.2, training=self.training)
            h = torch.cat([lat, xyz], dim=-1)
        last = self.num_layers - 2
        for k in range(self.num_layers - 1):
            if k in self.latent_in:
                h = torch.cat([h, net_input], dim=-1)
            elif k != 0 and self.xyz_in_all:
                h = torch.cat([h, xyz], dim=-1)
            h = getattr(self, "lin%d" % k)(h)
            if k == last:
                if self.use_tanh:
                    h = torch.tanh(h)
                break
            if (self.norm_layers is not None and k in self.norm_layers
                    and not self.weight_norm):
                h = getattr(self, "bn%d" % k)(h)
            h = F.relu(h)
            if self.dropout is not None and k in self.dropout:
                h = F.dropout(h, p=self.dropout_prob, training=self.training)
        return torch.tanh(h)
