"""Host-side description of the multi-resolution hash grid encoder (the encoder the reference's configs name:
``grid.enc: HashGrid``, ``model/encodings.py:31-46``; never instantiated by the reference itself).

Level geometry follows tiny-cuda-nn's grid encoding: ``scale_l = 2^(l * log2(per_level_scale)) * base - 1``,
``resolution_l = ceil(scale_l) + 1``, ``entries_l = min(2^log2_hashmap_size, round_up(resolution_l^3, 8))``.
The float32 scales are computed here once and handed to the kernels in the descriptor.
"""
import math

import numpy as np
import torch


class HashGridMeta:
    def __init__(self, desired_resolution, log2_hashmap_size, n_levels=16, n_features=2, base_resolution=16):
        if (n_levels, n_features) != (16, 2):
            raise NotImplementedError('the kernels are built for 16 levels x 2 features (model/encodings.py:8-10 defaults)')
        self.n_levels, self.n_features, self.base_resolution = n_levels, n_features, base_resolution
        self.log2_hashmap_size = log2_hashmap_size
        self.per_level_scale = float(np.exp2(np.log2(desired_resolution / base_resolution) / (n_levels - 1)))
        log2_pls = np.float32(math.log2(self.per_level_scale))
        self.scales, self.resolutions, self.offsets = [], [], [0]
        for l in range(n_levels):
            scale = np.float32(np.float32(np.exp2(np.float32(l) * log2_pls)) * np.float32(base_resolution) - np.float32(1.0))
            res = int(math.ceil(float(scale))) + 1
            entries = min((res ** 3 + 7) // 8 * 8, 1 << log2_hashmap_size)
            self.scales.append(float(scale))
            self.resolutions.append(res)
            self.offsets.append(self.offsets[-1] + entries)
        self.n_entries = self.offsets[-1]
        self.n_output = n_levels * n_features
        self.color = None        # HashGridMeta of the separate colour table (grid.oneGrid: False), attached by the model

    @classmethod
    def from_config(cls, cfg, key='voxel_sdf'):
        """``desired_resolution`` as the reference's (commented) ``get_resolution`` derives it
        (model/scene_rep.py:55-70): the value itself when > 10, else int(longest box side / voxel size)."""
        v = cfg['grid'][key]
        if v > 10:
            res = int(v)
        else:
            b = np.array(cfg['mapping']['bound'], dtype=np.float64)
            res = int((b[:, 1] - b[:, 0]).max() / v)
        return cls(res, cfg['grid']['hash_size'])

    def new_table(self, device, scale=1e-4):
        """[n_entries, 2] float32, U(-scale, scale) (tiny-cuda-nn's grid initialisation)."""
        return (torch.rand(self.n_entries, self.n_features, device=device) * 2 - 1) * scale

    def fill_desc(self, desc, table, grad=None, table_h=None):
        if not (table.is_cuda and table.dtype == torch.float32 and table.is_contiguous()
                and tuple(table.shape) == (self.n_entries, self.n_features)):
            raise TypeError('hash table must be a contiguous float32 CUDA tensor of shape [%d, %d]' % (self.n_entries, self.n_features))
        desc.table = table.data_ptr()
        desc.table_h = None
        if table_h is not None:
            if not (table_h.is_cuda and table_h.dtype == torch.float16 and table_h.is_contiguous() and table_h.shape == table.shape):
                raise TypeError('the fp16 copy of the hash table must be a contiguous float16 CUDA tensor of the table\'s shape')
            desc.table_h = table_h.data_ptr()
        desc.grad = None
        if grad is not None:
            if grad.shape != table.shape or not grad.is_contiguous() or not grad.is_cuda:
                raise TypeError('hash gradient buffer must match the table')
            desc.grad = grad.data_ptr()
        desc.n_levels, desc.n_features, desc.base_resolution = self.n_levels, self.n_features, self.base_resolution
        desc.per_level_scale = self.per_level_scale
        for l in range(self.n_levels + 1):
            desc.offsets[l] = self.offsets[l]
        for l in range(self.n_levels):
            desc.level_scale[l] = self.scales[l]
            desc.level_res[l] = self.resolutions[l]
