"""Drop-in for ``DecoderPreNet`` of the reference's ``Models/prenets.py:8-44`` (parameter holder; the kernels run from
``Models/decoder.py``)."""
from __future__ import annotations

from collections import OrderedDict

import torch.nn as nn


class DecoderPreNet(nn.Module):
    """fc1 -> ReLU -> dropout -> fc2 -> ReLU -> dropout (prenets.py:24-44).  The inference script builds the model with
    dropout_prenet = 0 (test.py:107) and nn.Dropout is inactive in eval(): the decode kernels apply none."""

    def __init__(self, input_size, output_size, hidden_size=256, p=0.5, output_type=None):
        super().__init__()
        if output_type:
            raise NotImplementedError("output_type (discrete targets) is out of scope (SURVEY §2a: default hparams only)")
        self.input_size = input_size
        self.output_size = output_size
        self.hidden_size = hidden_size
        self.layer = nn.Sequential(OrderedDict([
            ("fc1", nn.Linear(self.input_size, self.hidden_size)),
            ("relu1", nn.ReLU()),
            ("dropout1", nn.Dropout(p)),
            ("fc2", nn.Linear(self.hidden_size, self.output_size)),
            ("relu2", nn.ReLU()),
            ("dropout2", nn.Dropout(p)),
        ]))

    def forward(self, input_):
        raise NotImplementedError("DecoderPreNet runs inside Decoder (Models/decoder.py); it has no standalone kernel path")
