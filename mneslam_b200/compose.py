"""Rendering along explicitly given sample positions (no sampling stage)."""
import torch

from . import ops


def composite_along(model, rays_o, rays_d, z_row):
    """Shade ``rays_o + rays_d * z`` for the shared sample row ``z_row`` [S] and composite with sdf2weights
    (reference ``model/scene_rep.py:333-349`` uses this with the surface normal as direction)."""
    ops.require_cuda()
    n = rays_o.shape[0]
    z_vals = z_row.float().reshape(1, -1).repeat(n, 1).contiguous()
    rcfg = ops.make_render_cfg(model.config, norm_rays=n)
    out = ops.render_forward(model._scene(), rcfg, rays_o.float(), rays_d.float(), z_vals)
    out['z_vals'] = z_vals
    return out
