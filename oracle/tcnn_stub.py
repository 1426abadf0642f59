"""Stand-in for the ``tinycudann`` module so that the reference's ``model/`` package imports on a box
without tiny-cuda-nn.  Only ``Encoding(OneBlob)`` is functional -- it is the only
tcnn object the live reference path builds (``model/scene_rep.py:157``); ``Network`` exists because
``model/decoder.py:4`` imports the module, and is never constructed (``tcnn_network: False``).

Used by ``oracle/make_golden.py`` only.  TEST INFRASTRUCTURE.
"""
import sys
import types

import torch

from .oneblob import oneblob


class Encoding(torch.nn.Module):
    def __init__(self, n_input_dims, encoding_config, dtype=torch.float):
        super().__init__()
        if encoding_config.get('otype') != 'OneBlob':
            raise NotImplementedError('tcnn stub: only OneBlob is restated')
        self.n_bins = int(encoding_config['n_bins'])
        self.n_input_dims = n_input_dims
        self.n_output_dims = n_input_dims * self.n_bins

    def forward(self, x):
        return oneblob(x.float(), self.n_bins)


class Network(torch.nn.Module):
    def __init__(self, *a, **k):
        raise NotImplementedError('tcnn stub: FullyFusedMLP is dormant in the reference')


def install():
    mod = types.ModuleType('tinycudann')
    mod.Encoding = Encoding
    mod.Network = Network
    sys.modules['tinycudann'] = mod
    return mod
