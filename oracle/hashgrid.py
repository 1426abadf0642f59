"""Multi-resolution hash grid -- CPU restatement (torch, differentiable).  PARITY UNPINNED.

The reference would build ``tcnn.Encoding({"otype": "HashGrid", n_levels 16, n_features_per_level 2,
log2_hashmap_size, base_resolution 16, per_level_scale}, dtype=float)`` (``model/encodings.py:31-46``) with
``desired_resolution = int(dim_max / voxel_sdf)`` (``model/scene_rep.py:55-70``, commented out) -- but never
calls it (``model/scene_rep.py:160`` is commented), and holds no test vector for it.  The arithmetic lives in
tiny-cuda-nn (``include/tiny-cuda-nn/encodings/grid.h``, un-vendored, ``requirements.txt:120``); this file
restates its published algorithm ("pure-torch hash grid with tcnn's hashing constants", BASELINE.json):

* level ``l``: ``scale = 2^(l * log2(per_level_scale)) * base - 1``, ``resolution = ceil(scale) + 1``,
  entries ``= min(2^log2_hashmap_size, round_up(resolution^3, 8))``, levels stored back to back;
* a point ``x`` in ``[0,1]^3``: ``pos = fmaf(x, scale, 0.5)`` (ONE rounding, as tiny-cuda-nn's ``pos_fract``; restated here
  as the float32 rounding of the float64 product-and-sum: the product of two float32 is exact in float64);
  cell = ``floor(pos)`` (as uint32), ``frac = pos - floor``;
* corner index: ``x + y*res + z*res^2`` while the running stride fits the level, otherwise the coherent prime hash
  ``x ^ (y * 2654435761) ^ (z * 805459861)`` in uint32; then ``% entries``;
* output = trilinear blend of the ``n_features`` values of the 8 corners, levels concatenated (level-major).

"Hash-grid indices bit-exact" (BASELINE.json) is defined against ``corner_indices`` below.
"""
import math

import numpy as np
import torch

PRIMES = (1, 2654435761, 805459861)
MASK = 0xFFFFFFFF


def desired_resolution(cfg, key='voxel_sdf'):
    """``get_resolution`` (model/scene_rep.py:55-70): int(dim_max / voxel) unless the config gives a resolution."""
    b = np.array(cfg['mapping']['bound'], dtype=np.float64)
    v = cfg['grid'][key]
    return int(v) if v > 10 else int((b[:, 1] - b[:, 0]).max() / v)


def level_table(n_levels, base_resolution, per_level_scale, log2_hashmap_size):
    """Per-level (scale float32, resolution, entries, offset) exactly as tiny-cuda-nn derives them."""
    log2_pls = np.float32(math.log2(per_level_scale))
    scales, res, sizes, offsets = [], [], [], [0]
    for l in range(n_levels):
        scale = np.float32(np.exp2(np.float32(l) * log2_pls)) * np.float32(base_resolution) - np.float32(1.0)
        scale = np.float32(scale)
        r = int(math.ceil(float(scale))) + 1
        n = r ** 3
        n = (n + 7) // 8 * 8
        n = min(n, 1 << log2_hashmap_size)
        scales.append(scale); res.append(r); sizes.append(n); offsets.append(offsets[-1] + n)
    return np.array(scales, np.float32), res, sizes, offsets


def make_meta(cfg, resolution=None, n_levels=16, n_features=2, base_resolution=16):
    res = resolution or desired_resolution(cfg)
    pls = float(np.exp2(np.log2(res / base_resolution) / (n_levels - 1)))
    scales, lres, sizes, offsets = level_table(n_levels, base_resolution, pls, cfg['grid']['hash_size'])
    return {'n_levels': n_levels, 'n_features': n_features, 'base_resolution': base_resolution, 'per_level_scale': pls,
            'scales': scales, 'resolutions': lres, 'sizes': sizes, 'offsets': offsets}


def init_table(meta, seed=0, scale=1e-4):
    """tiny-cuda-nn initialises grid parameters U(-1e-4, 1e-4)."""
    gen = torch.Generator().manual_seed(seed)
    return (torch.rand(meta['offsets'][-1], meta['n_features'], generator=gen) * 2 - 1) * scale


def fused_pos(x, scale):
    """fmaf(x, scale, 0.5) in float32: product and sum in float64 (the product is exact there), one rounding to float32."""
    return (x.double() * scale.double() + 0.5).float()


def corner_indices(meta, x, level):
    """Entry indices (within the level) of the 8 corners and the fractional position.  x: [M,3] float32."""
    scale = torch.tensor(meta['scales'][level])
    res, size = meta['resolutions'][level], meta['sizes'][level]
    pos = fused_pos(x, scale)
    cell_f = torch.floor(pos)
    frac = pos - cell_f
    cell = cell_f.to(torch.int64) & MASK      # (uint32_t)(int) wrap-around
    idx = []
    for c in range(8):
        g = [(cell[:, d] + ((c >> d) & 1)) & MASK for d in range(3)]
        stride, index, d = 1, torch.zeros_like(g[0]), 0
        while d < 3 and stride <= size:
            index = (index + g[d] * stride) & MASK
            stride *= res
            d += 1
        if size < stride:
            index = ((g[0] * PRIMES[0]) & MASK) ^ ((g[1] * PRIMES[1]) & MASK) ^ ((g[2] * PRIMES[2]) & MASK)
        idx.append(index % size)
    return torch.stack(idx, -1), frac


def encode(meta, table, x):
    """x: [M,3] float32 in [0,1] (the OneBlob input of the same point) -> [M, n_levels * n_features]."""
    out = []
    for l in range(meta['n_levels']):
        idx, frac = corner_indices(meta, x.detach(), l)
        idx = idx + meta['offsets'][l]
        # fractional position with gradient (pose path); indices carry none
        scale = torch.tensor(meta['scales'][l])
        xs = x * scale
        pos = fused_pos(x.detach(), scale) + (xs - xs.detach())     # value: the fused rounding; gradient: d pos / d x = scale
        f = pos - torch.floor(pos.detach())
        acc = 0
        for c in range(8):
            w = 1.0
            for d in range(3):
                w = w * (f[:, d] if (c >> d) & 1 else (1.0 - f[:, d]))
            acc = acc + w[:, None] * table[idx[:, c]]
        out.append(acc)
    return torch.cat(out, -1)
