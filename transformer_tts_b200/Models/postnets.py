"""Drop-in for ``PostConvNet`` of the reference's ``Models/postnets.py:13-79``."""
from __future__ import annotations

import copy

import torch
import torch.nn as nn

from . import _engine as E


def clones(module, N):
    return nn.ModuleList([copy.deepcopy(module) for _ in range(N)])


class PostConvNet(nn.Module):
    """Linear d->mel, then 5 causal conv1d(k=5) with BatchNorm+tanh, residual (postnets.py:64-79)."""

    def __init__(self, hp, num_hidden, mel_dim, reduction_rate, dropout=0.5, prev_version=True):
        super().__init__()
        self.prev_version = prev_version
        self.conv1 = nn.Conv1d(in_channels=mel_dim * reduction_rate, out_channels=num_hidden, kernel_size=5, padding=4)
        self.conv_list = clones(nn.Conv1d(in_channels=num_hidden, out_channels=num_hidden, kernel_size=5, padding=4), 3)
        self.conv2 = nn.Conv1d(in_channels=num_hidden, out_channels=mel_dim * reduction_rate, kernel_size=5, padding=4)
        if self.prev_version:  # Transformer-TTS builds it with prev_version=False: no `out` (postnets.py:40-41)
            self.out = nn.Linear(num_hidden, mel_dim * reduction_rate)
        self.batch_norm_list = clones(nn.BatchNorm1d(num_hidden), 3)
        self.pre_batchnorm = nn.BatchNorm1d(num_hidden)
        self.dropout1 = nn.Dropout(p=dropout)
        self.dropout_list = nn.ModuleList([nn.Dropout(p=dropout) for _ in range(3)])
        self._ttsb_cache = E.Cache()

    def run(self, d_bf16, tile_map=None):
        return E.run_postnet(E.postnet_cache(self), d_bf16, tile_map=tile_map)

    def forward(self, input_, mask=None):
        E.require_inference(self, "PostConvNet")
        E.require_cuda(input_, "PostConvNet")
        if not self.prev_version:
            # postnets.py:64-79 with prev_version=False computes the five convolutions and then returns
            # `mel_pred.transpose(1, 2)`, i.e. its INPUT: the result does not depend on them, so none is launched
            return input_
        return self.run(input_.to(torch.bfloat16).contiguous())
