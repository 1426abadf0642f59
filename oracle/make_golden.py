"""Generate the golden fixtures under ``tests/golden/`` by running the REFERENCE'S OWN modules.

Run in the build container only (``/root/reference`` does not exist on the GPU box):

    python -m oracle.make_golden

It imports ``model.scene_rep.JointEncoding`` from ``/root/reference`` with ``oracle.tcnn_stub`` standing
in for tiny-cuda-nn, installs seeded planes / decoder weights, runs ``forward`` (+ the caller's loss
weighting, ``mneslam_mp.py:350-372``, restated in ``oracle.joint_encoding.total_loss``) and ``backward``,
and stores inputs, outputs and gradients as ``.npz``.  The stratified noise of
``model/scene_rep.py:381`` is reproduced by seeding ``torch.manual_seed`` right before the call and
drawing the same ``torch.rand`` sequence here.

TEST INFRASTRUCTURE ONLY.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF = '/root/reference'
sys.path.insert(0, ROOT)

from oracle import tcnn_stub, joint_encoding as J          # noqa: E402
from mneslam_b200 import configs, synthetic                # noqa: E402


def reference_model(cfg, planes, weights):
    """Build the reference JointEncoding and load our seeded state into it."""
    tcnn_stub.install()
    if REF not in sys.path:
        sys.path.insert(0, REF)
    from model.scene_rep import JointEncoding      # the reference's own class
    bb = torch.from_numpy(np.array(cfg['mapping']['bound']))          # float64, mneslam_mp.py:223
    model = JointEncoding(cfg, bb)
    assert len(model.all_planes) == len(planes)
    params = []
    for dst, src in zip(model.all_planes, planes):
        assert len(dst) == len(src)
        for i, p in enumerate(src):
            assert tuple(dst[i].shape) == tuple(p.shape), (dst[i].shape, p.shape)
            dst[i] = torch.nn.Parameter(p.clone())      # as create_optimizer does, mneslam_mp.py:451-457
            params.append(dst[i])
    with torch.no_grad():
        model.decoder.sdf_net.model[0].weight.copy_(weights['sdf0'])
        model.decoder.sdf_net.model[2].weight.copy_(weights['sdf1'])
        model.decoder.color_net.model[0].weight.copy_(weights['col0'])
        model.decoder.color_net.model[2].weight.copy_(weights['col1'])
    return model, params, bb


def probe(t, seed, k=2048):
    """Compact summary of a large gradient tensor: values at seeded flat indices + sum + L2 norm."""
    flat = t.detach().reshape(-1)
    g = torch.Generator().manual_seed(seed)
    idx = torch.randint(0, flat.numel(), (min(k, flat.numel()),), generator=g)
    return idx.numpy(), flat[idx].numpy(), np.array([flat.double().sum().item(), flat.double().norm().item()])


def run_case(name, cfg, n_rays, seed, plane_std, store_planes, with_depth=True, pose_rep=None,
             zero_depth_frac=0.1):
    torch.manual_seed(0)
    bound = J.enlarged_bound(cfg)
    planes = J.init_planes(cfg, bound, seed=seed, std=plane_std)
    weights = J.init_weights(cfg, seed=seed)
    model, plane_params, bb = reference_model(cfg, planes, weights)
    assert torch.equal(model.bound, bound), (model.bound, bound)
    model.train()

    batch = synthetic.make_ray_batch(cfg, n_rays, seed=seed, zero_depth_frac=zero_depth_frac)
    S = synthetic.n_samples_total(cfg, with_depth)
    n_imp = cfg['training']['n_importance']
    noise_seed = 1234 + seed
    torch.manual_seed(noise_seed)
    noise = torch.rand(n_rays, S)
    u_imp = torch.rand(n_rays, n_imp) if n_imp > 0 else None

    out = {'n_rays': n_rays, 'seed': seed, 'plane_std': plane_std, 'noise': noise.numpy(),
           'target_rgb': batch['target_rgb'].numpy(), 'target_d': batch['target_d'].numpy(),
           'dirs_cam': batch['dirs_cam'].numpy(), 'c2w': batch['c2w'].numpy(),
           'bound': bound.numpy(), 'bounding_box': bb.numpy()}
    if u_imp is not None:
        out['u_importance'] = u_imp.numpy()

    if pose_rep is None:
        rays_o = batch['rays_o'].clone().requires_grad_(True)
        rays_d = batch['rays_d'].clone().requires_grad_(True)
        pose_params = []
    else:
        # loop-closure alignment path (mp_slam/mapper.py:388-410): rays from a differentiable pose
        R0, t0 = batch['c2w'][:3, :3], batch['c2w'][:3, 3]
        trans = t0.clone().requires_grad_(True)
        if pose_rep == 'axis_angle':
            rot = torch.tensor([0.02, -0.05, 0.03], requires_grad=True)
            R = J.axis_angle_to_matrix(rot[None])[0] @ R0
        else:
            rot = torch.tensor([0.99, 0.02, -0.05, 0.03], requires_grad=True)
            R = J.quat_to_matrix(rot[None])[0] @ R0
        rays_o, rays_d = J.rays_from_pose(R, trans, batch['dirs_cam'])
        rays_o.retain_grad(); rays_d.retain_grad()
        pose_params = [rot, trans]
        out['pose_rot'] = rot.detach().numpy()
    out['rays_o'] = rays_o.detach().numpy()
    out['rays_d'] = rays_d.detach().numpy()

    # record the bin indices of the reference's own torch.searchsorted call (model/utils.py:91)
    captured, real_searchsorted = [], torch.searchsorted

    def spy(*a, **k):
        r = real_searchsorted(*a, **k)
        captured.append(r.clone())
        return r
    torch.searchsorted = spy

    torch.manual_seed(noise_seed)                      # the reference draws the same sequence now
    if with_depth:
        ret = model.forward(rays_o, rays_d, batch['target_rgb'], batch['target_d'])
        loss = J.total_loss(cfg, ret)
        for k in ('rgb_loss', 'depth_loss', 'co_sdf_loss', 'co_fs_loss', 'e_fs_loss', 'e_center_loss',
                  'e_tail_loss', 'psnr'):
            out[k] = ret[k].detach().numpy()
        # the render dict (z_vals, raw) is not in the training-mode return: take it from a second,
        # identically seeded call of render_rays
        torch.manual_seed(noise_seed)
        with torch.no_grad():
            rend = model.render_rays(rays_o.detach(), rays_d.detach(), target_d=batch['target_d'])
    else:
        rend_g = model.render_rays(rays_o, rays_d, target_d=None)
        tgt_rgb, tgt_d = batch['target_rgb'], batch['target_d'].squeeze(-1)
        loss = (cfg['training']['rgb_weight'] * torch.nn.functional.mse_loss(rend_g['rgb'], tgt_rgb)
                + cfg['training']['depth_weight'] * torch.nn.functional.mse_loss(rend_g['depth'], tgt_d))
        rend = {k: v.detach() for k, v in rend_g.items()}
    out['loss'] = loss.detach().numpy()
    torch.searchsorted = real_searchsorted
    if captured:
        out['pdf_inds'] = captured[0].numpy()
    for k in ('rgb', 'depth', 'disp_map', 'acc_map', 'depth_var', 'z_vals', 'raw', 'rgb0', 'depth0', 'z_std'):
        if k in rend:
            out[k] = rend[k].numpy()
    # sdf2weights sign-change indices (model/scene_rep.py:195-197), recomputed with the reference's ops
    sdf = rend['raw'][..., 3]
    out['inds'] = torch.argmax(torch.where(sdf[:, 1:] * sdf[:, :-1] < 0.0, 1.0, 0.0), axis=1).numpy()

    loss.backward()
    out['g_rays_o'] = rays_o.grad.numpy()
    out['g_rays_d'] = rays_d.grad.numpy()
    for p, nm in zip(pose_params, ('g_pose_rot', 'g_pose_trans')):
        out[nm] = p.grad.numpy()
    dec = model.decoder
    out['g_sdf0'] = dec.sdf_net.model[0].weight.grad.numpy()
    out['g_sdf1'] = dec.sdf_net.model[2].weight.grad.numpy()
    out['g_col0'] = dec.color_net.model[0].weight.grad.numpy()
    out['g_col1'] = dec.color_net.model[2].weight.grad.numpy()
    for k, w in weights.items():
        out['w_' + k] = w.numpy()
    for i, p in enumerate(plane_params):
        if store_planes:
            out[f'plane{i}'] = p.detach().numpy()
            out[f'g_plane{i}'] = p.grad.numpy()
        else:
            idx, val, stats = probe(p.grad, seed=100 + i)
            out[f'gp_idx{i}'], out[f'gp_val{i}'], out[f'gp_stats{i}'] = idx, val, stats
    os.makedirs(os.path.join(ROOT, 'tests', 'golden'), exist_ok=True)
    path = os.path.join(ROOT, 'tests', 'golden', name + '.npz')
    np.savez_compressed(path, **out)
    print(f'{name}: loss={float(loss):.6f}  -> {path}  ({os.path.getsize(path) / 1024:.0f} KiB)')


def main():
    os.chdir(REF)      # the reference's YAML inherit_from is CWD-relative (config.py:18-23); harmless here
    small = configs.small_test_config
    run_case('small_onegrid', small(True), 96, seed=1, plane_std=0.4, store_planes=True)
    run_case('small_twogrid', small(False), 96, seed=2, plane_std=0.4, store_planes=True)
    cfg = small(True); cfg['is_co_sdf'] = True
    run_case('small_cosdf', cfg, 96, seed=3, plane_std=0.4, store_planes=True)
    run_case('small_importance', small(True, n_importance=4), 64, seed=4, plane_std=0.4, store_planes=True)
    run_case('small_nodepth_axis', small(True), 64, seed=5, plane_std=0.4, store_planes=True,
             with_depth=False, pose_rep='axis_angle')
    run_case('small_nodepth_quat', small(False), 64, seed=6, plane_std=0.4, store_planes=True,
             with_depth=False, pose_rep='quat')
    cfg = small(True); cfg['training']['perturb'] = 0
    run_case('small_noperturb', cfg, 64, seed=7, plane_std=0.4, store_planes=True)
    # full-size scenes: planes are re-created from the seed by the tests, gradients stored as probes
    run_case('r0_256', configs.replica_room0(), 256, seed=8, plane_std=0.3, store_planes=False)
    run_case('sn_128', configs.scannet_scene0000(), 128, seed=9, plane_std=0.3, store_planes=False)


if __name__ == '__main__':
    main()
