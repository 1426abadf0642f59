"""Golden-case registry shared by the oracle tests (CPU) and the parity tests (GPU).

Each case names the config the fixture was generated with by ``oracle/make_golden.py`` (keep the two
in sync) and rebuilds the seeded planes / decoder weights the same way."""
import os

import numpy as np
import torch

from mneslam_b200 import configs
from oracle import joint_encoding as J

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')


def _cfg(name):
    small = configs.small_test_config
    if name == 'small_onegrid':
        return small(True)
    if name == 'small_twogrid':
        return small(False)
    if name == 'small_cosdf':
        c = small(True); c['is_co_sdf'] = True; return c
    if name == 'small_importance':
        return small(True, n_importance=4)
    if name == 'small_nodepth_axis':
        return small(True)
    if name == 'small_nodepth_quat':
        return small(False)
    if name == 'small_noperturb':
        c = small(True); c['training']['perturb'] = 0; return c
    if name == 'r0_256':
        return configs.replica_room0()
    if name == 'sn_128':
        return configs.scannet_scene0000()
    raise KeyError(name)


SMALL_DEPTH = ['small_onegrid', 'small_twogrid', 'small_cosdf', 'small_noperturb']
SMALL_NODEPTH = ['small_nodepth_axis', 'small_nodepth_quat']
FULL = ['r0_256', 'sn_128']
ALL = SMALL_DEPTH + ['small_importance'] + SMALL_NODEPTH + FULL


class Case:
    def __init__(self, name):
        self.name = name
        self.cfg = _cfg(name)
        self.g = dict(np.load(os.path.join(GOLDEN, name + '.npz')))
        self.with_depth = 'nodepth' not in name
        seed, std = int(self.g['seed']), float(self.g['plane_std'])
        self.bound = J.enlarged_bound(self.cfg)
        self.bounding_box = torch.from_numpy(self.g['bounding_box'])
        self.planes = J.init_planes(self.cfg, self.bound, seed=seed, std=std)
        self.weights = J.init_weights(self.cfg, seed=seed)
        if 'plane0' in self.g:      # small cases carry their planes: check the seeded rebuild
            flat = [p for group in self.planes for p in group]
            for i, p in enumerate(flat):
                assert np.array_equal(p.numpy(), self.g[f'plane{i}'])

    def t(self, key):
        return torch.from_numpy(self.g[key])

    def flat_planes(self):
        return [p for group in self.planes for p in group]
