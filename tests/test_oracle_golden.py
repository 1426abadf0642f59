"""The oracle restatement against the fixtures produced by the reference's own modules (CPU)."""
import numpy as np
import pytest
import torch

from oracle import joint_encoding as J
from tests.cases import ALL, Case


def rel(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return np.abs(a - b).max() / (np.abs(b).max() + 1e-12)


def cosine(a, b):
    a, b = np.asarray(a, np.float64).ravel(), np.asarray(b, np.float64).ravel()
    return float(a @ b / (np.linalg.norm(a) * np.linalg.norm(b) + 1e-300))


def run_oracle(case, fast=False):
    g, cfg = case.g, case.cfg
    planes = tuple([p.clone().requires_grad_(True) for p in grp] for grp in case.planes)
    weights = {k: v.clone().requires_grad_(True) for k, v in case.weights.items()}
    rays_o = case.t('rays_o').requires_grad_(True)
    rays_d = case.t('rays_d').requires_grad_(True)
    noise = case.t('noise')
    u_imp = case.t('u_importance') if 'u_importance' in g else None
    if case.with_depth:
        ret, rend = J.forward(cfg, planes, weights, case.bound, case.bounding_box, rays_o, rays_d,
                              case.t('target_rgb'), case.t('target_d'), noise, u_imp, fast=fast)
        loss = J.total_loss(cfg, ret)
    else:
        rend = J.render_rays(cfg, planes, weights, case.bound, case.bounding_box, rays_o, rays_d, None,
                             noise, fast=fast)
        ret = {}
        t = cfg['training']
        loss = (t['rgb_weight'] * ((rend['rgb'] - case.t('target_rgb')) ** 2).mean()
                + t['depth_weight'] * ((rend['depth'] - case.t('target_d').squeeze(-1)) ** 2).mean())
    loss.backward()
    return ret, rend, loss, planes, weights, rays_o, rays_d


@pytest.mark.parametrize('name', ALL)
@pytest.mark.parametrize('fast', [False, True])
def test_oracle_matches_reference(name, fast):
    case = Case(name)
    g = case.g
    ret, rend, loss, planes, weights, rays_o, rays_d = run_oracle(case, fast)
    # sample positions and sign-change indices: bit-exact
    if 'pdf_inds' in g:     # importance samples depend on network outputs: bins exact, positions close
        assert np.array_equal(rend['pdf_inds'].numpy(), g['pdf_inds'])
        np.testing.assert_allclose(rend['z_vals'].detach().numpy(), g['z_vals'], rtol=0, atol=2e-6)
    else:
        assert np.array_equal(rend['z_vals'].detach().numpy(), g['z_vals'])
    assert np.array_equal(rend['inds'].numpy(), g['inds'])
    for k in ('rgb', 'depth', 'disp_map', 'acc_map', 'depth_var', 'raw', 'rgb0', 'depth0', 'z_std'):
        if k in g:
            assert rel(rend[k].detach().numpy(), g[k]) < 2e-5, k
    for k in ('rgb_loss', 'depth_loss', 'co_sdf_loss', 'co_fs_loss', 'e_fs_loss', 'e_center_loss',
              'e_tail_loss', 'psnr'):
        if k in g:
            np.testing.assert_allclose(ret[k].detach().numpy(), g[k], rtol=2e-5, atol=1e-7, err_msg=k)
    np.testing.assert_allclose(loss.detach().numpy(), g['loss'], rtol=2e-5)
    # gradients
    for k in ('sdf0', 'sdf1', 'col0', 'col1'):
        assert cosine(weights[k].grad.numpy(), g['g_' + k]) > 0.999999, k
        assert rel(weights[k].grad.numpy(), g['g_' + k]) < 1e-3, k
    assert rel(rays_o.grad.numpy(), g['g_rays_o']) < 1e-3
    assert rel(rays_d.grad.numpy(), g['g_rays_d']) < 1e-3
    flat = [p for grp in planes for p in grp]
    for i, p in enumerate(flat):
        if f'g_plane{i}' in g:
            assert cosine(p.grad.numpy(), g[f'g_plane{i}']) > 0.999999, i
            assert rel(p.grad.numpy(), g[f'g_plane{i}']) < 1e-3, i
        else:
            got = p.grad.reshape(-1)[torch.from_numpy(g[f'gp_idx{i}'])].numpy()
            assert rel(got, g[f'gp_val{i}']) < 1e-3 or np.abs(g[f'gp_val{i}']).max() == 0, i
            stats = np.array([p.grad.double().sum().item(), p.grad.double().norm().item()])
            np.testing.assert_allclose(stats[1], g[f'gp_stats{i}'][1], rtol=1e-4)


def test_pose_gradients_chain():
    """Ray gradients chained through the pose parameterisations reproduce the reference's pose grads."""
    for name, rep in (('small_nodepth_axis', 'axis_angle'), ('small_nodepth_quat', 'quat')):
        case = Case(name)
        g = case.g
        c2w = case.t('c2w')
        rot = case.t('pose_rot').clone().requires_grad_(True)
        trans = c2w[:3, 3].clone().requires_grad_(True)
        Rm = (J.axis_angle_to_matrix(rot[None])[0] if rep == 'axis_angle' else J.quat_to_matrix(rot[None])[0])
        ro, rd = J.rays_from_pose(Rm @ c2w[:3, :3], trans, case.t('dirs_cam'))
        assert rel(ro.detach().numpy(), g['rays_o']) < 1e-6 and rel(rd.detach().numpy(), g['rays_d']) < 1e-6
        torch.autograd.backward([ro, rd], [case.t('g_rays_o'), case.t('g_rays_d')])
        assert rel(rot.grad.numpy(), g['g_pose_rot']) < 1e-4
        assert rel(trans.grad.numpy(), g['g_pose_trans']) < 1e-4


def test_oneblob_properties():
    """OneBlob self-checks (the restatement is unpinned by the reference, see oracle/oneblob.py)."""
    from oracle.oneblob import oneblob
    x = torch.rand(4096, 3)
    y = oneblob(x, 16)
    assert y.shape == (4096, 48)
    assert float(y.min()) >= -1e-6 and float(y.max()) <= 1 + 1e-6
    np.testing.assert_allclose(y.reshape(-1, 3, 16).sum(-1).numpy(), 1.0, atol=1e-5)
    # the power-of-two shortcut equals the direct two-edge formula on the periodic range
    xw = torch.rand(4096, 3) * 2.8 - 0.9
    left = torch.arange(16.) / 16
    from oracle.oneblob import _wrapped_cdf
    direct = (_wrapped_cdf((left + 1 / 16)[None, None] - xw[..., None], 16)
              - _wrapped_cdf(left[None, None] - xw[..., None], 16)).reshape(-1, 48)
    np.testing.assert_allclose(oneblob(xw, 16).numpy(), direct.numpy(), atol=1e-6)
