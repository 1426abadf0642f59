"""Two-pass rendering with importance sampling (``training.n_importance > 0``; reference
``model/scene_rep.py:388-417`` + ``sample_pdf`` ``model/utils.py:65-104``; dormant in every shipped config).

coarse pass on the depth-guided samples -> inverse-CDF samples from its weights (detached, ``:395``) -> merge ->
fine pass on all samples.  Both passes receive loss gradient (``rgb0`` / ``depth0`` enter ``rgb_loss`` /
``depth_loss``, ``:579-581``); the SDF losses use the fine pass only.
"""
import ctypes as C

import torch

from . import _lib, ops
from ._lib import LOSS_KEYS


def _det_u(n_imp, device):
    return torch.linspace(0.5 / n_imp, 1. - 0.5 / n_imp, steps=n_imp).to(device)


class _RenderImportance(torch.autograd.Function):
    @staticmethod
    def forward(ctx, model, rays_o, rays_d, target_rgb, target_d, noise, u_imp, with_losses, *params):
        cfg, dev = model.config, rays_o.device
        n = rays_o.shape[0]
        n_imp = cfg['training']['n_importance']
        with_depth = target_d is not None
        n_planes = len(params) - 4
        planes, weights = list(params[:n_planes]), list(params[n_planes:])
        tables = ops.SampleTables.get(cfg, dev, with_depth)
        s0 = tables[0].numel() + tables[1].numel()
        perturb = cfg['training']['perturb'] > 0.
        if perturb and noise is None:
            noise = torch.rand(n, s0, device=dev)
        if perturb and u_imp is None:
            u_imp = torch.rand(n, n_imp, device=dev)
        rc0 = ops.make_render_cfg(cfg, norm_rays=n * model.norm_scale, first_pass=True)
        rc1 = ops.make_render_cfg(cfg, norm_rays=n * model.norm_scale)
        scfg = ops.make_sample_cfg(tables, perturb)
        td = target_d.reshape(-1) if with_depth else None
        acc = ops.new_acc(dev) if with_losses else None
        z0 = ops.sample_rays(scfg, rc0, n, s0, td, noise if perturb else None, None, dev)   # counts come from the final samples
        scene = ops.make_scene(planes, weights, model.bound, model.bounding_box, cfg['pos']['n_bins'], hash_meta=model.hash_meta)
        first = ops.render_forward(scene, rc0, rays_o, rays_d, z0, target_rgb if with_losses else None,
                                   td if with_losses else None, acc, want_weights=True)
        u = u_imp if perturb else _det_u(n_imp, dev)
        u = ops._f32(u, 'u_importance')
        z1 = torch.empty(n, s0 + n_imp, dtype=torch.float32, device=dev)
        inds = torch.empty(n, n_imp, dtype=torch.int32, device=dev)
        z_std = torch.empty(n, dtype=torch.float32, device=dev)
        _lib.check(_lib.lib().mne_sample_importance(C.byref(rc1), n, s0, n_imp, z0.data_ptr(), first['weights'].data_ptr(),
                                                    u.data_ptr(), 1 if perturb else 0, ops._ptr(td) if with_losses else None,
                                                    z1.data_ptr(), inds.data_ptr(), z_std.data_ptr(), ops._ptr(acc),
                                                    ops._stream()), 'mne_sample_importance')
        out = ops.render_forward(scene, rc1, rays_o, rays_d, z1, target_rgb if with_losses else None,
                                 td if with_losses else None, acc)
        if with_losses:
            if model.acc_allreduce is not None:
                model.acc_allreduce(acc)
            losses = ops.loss_finalize(rc1, s0 + n_imp, acc)
        else:
            losses = torch.zeros(len(LOSS_KEYS), device=dev)
        ctx.model, ctx.rc0, ctx.rc1, ctx.with_losses, ctx.n_planes = model, rc0, rc1, with_losses, n_planes
        ctx.save_for_backward(rays_o, rays_d, z0, first['raw'], z1, out['raw'], target_rgb if with_losses else None,
                              td if with_losses else None, acc, *params)
        ctx.set_materialize_grads(False)
        nondiff = (z1, out['disp_map'], out['acc_map'], out['depth_var'], out['inds'], first['disp_map'], first['acc_map'],
                   first['depth_var'], z_std, inds)
        ctx.mark_non_differentiable(*nondiff)
        return (out['rgb'], out['depth'], out['raw'], losses, first['rgb'], first['depth']) + nondiff

    @staticmethod
    def backward(ctx, g_rgb, g_depth, g_raw, g_losses, g_rgb0, g_depth0, *unused):
        rays_o, rays_d, z0, raw0, z1, raw1, target_rgb, td, acc = ctx.saved_tensors[:9]
        params = ctx.saved_tensors[9:]
        model, n_planes = ctx.model, ctx.n_planes
        planes, weights = list(params[:n_planes]), list(params[n_planes:])
        need = ctx.needs_input_grad
        need_params = need[8:]
        plane_grads = [torch.zeros_like(p) if nd else None for p, nd in zip(planes, need_params[:n_planes])]
        weight_grads = [torch.zeros_like(w) if nd else None for w, nd in zip(weights, need_params[n_planes:])]
        want_rays = bool(need[1] or need[2])
        scene = ops.make_scene(planes, weights, model.bound, model.bounding_box, model.config['pos']['n_bins'],
                               plane_grads=plane_grads, weight_grads=weight_grads, hash_meta=model.hash_meta)
        lg = g_losses if (ctx.with_losses and g_losses is not None) else None
        go1, gd1 = ops.render_backward(scene, ctx.rc1, rays_o, rays_d, z1, raw1, target_rgb, td, acc, loss_grads=lg,
                                       g_rgb=g_rgb, g_depth=g_depth, g_raw=g_raw, want_ray_grads=want_rays)
        go0, gd0 = ops.render_backward(scene, ctx.rc0, rays_o, rays_d, z0, raw0, target_rgb, td, acc, loss_grads=lg,
                                       g_rgb=g_rgb0, g_depth=g_depth0, want_ray_grads=want_rays)
        g_o = (go1 + go0) if want_rays else None
        g_d = (gd1 + gd0) if want_rays else None
        return (None, g_o if need[1] else None, g_d if need[2] else None, None, None, None, None, None,
                *plane_grads, *weight_grads)


def render_with_importance(model, rays_o, rays_d, target_rgb, target_d, with_losses, noise, u_importance=None):
    noise = noise if noise is not None else model.noise_override
    u_importance = u_importance if u_importance is not None else getattr(model, 'u_importance_override', None)
    params = model._flat_planes() + model.decoder.weight_list()
    res = _RenderImportance.apply(model, rays_o, rays_d, target_rgb, target_d, noise, u_importance, with_losses, *params)
    rgb, depth, raw, losses, rgb0, depth0, z1, disp, acc_map, depth_var, inds, disp0, acc0, depth_var0, z_std, pdf_inds = res
    rend = {'rgb': rgb, 'depth': depth, 'disp_map': disp, 'acc_map': acc_map, 'depth_var': depth_var, 'z_vals': z1,
            'raw': raw, 'rgb0': rgb0, 'disp0': disp0, 'acc0': acc0, 'depth0': depth0, 'depth_var0': depth_var0,
            'z_std': z_std, 'pdf_inds': pdf_inds}
    return rend, losses, inds
