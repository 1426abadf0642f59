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


def _inference_fixture():
    import os
    from oracle.make_golden_inference import inference_config
    from tests.cases import GOLDEN
    g = np.load(os.path.join(GOLDEN, 'small_inference.npz'))
    cfg = inference_config()
    bound = J.enlarged_bound(cfg)
    seed = int(g['seed'])
    planes, weights = J.init_planes(cfg, bound, seed=seed, std=0.4), J.init_weights(cfg, seed=seed)
    bb = torch.tensor(cfg['mapping']['bound'], dtype=torch.float64)
    return g, cfg, bound, bb, planes, weights


def test_axis_angle_restatement_equals_the_reference_function():
    """oracle.joint_encoding.axis_angle_to_matrix against the matrices the reference's own axis_angle_to_matrix /
    at_to_transform_matrix (optimization/utils.py:161-197, executed from its source) produced."""
    g = _inference_fixture()[0]
    rot = torch.from_numpy(g['pose_rot'])
    R = J.axis_angle_to_matrix(rot)
    np.testing.assert_allclose(R.numpy(), g['pose_R'], rtol=0, atol=2e-7)
    np.testing.assert_allclose(g['pose_T'][:, :3, :3], g['pose_R'], rtol=0, atol=0)
    np.testing.assert_allclose(g['pose_T'][:, :3, 3], g['pose_trans'], rtol=0, atol=0)


def test_oracle_renders_the_reference_image_and_eval_forward():
    """render_img (model/scene_rep.py:422-473) and the eval-mode forward (:566-567) of the reference against the oracle's
    render_rays on rays built the reference's way (get_rays with the raw cam.cx / cam.cy)."""
    g, cfg, bound, bb, planes, weights = _inference_fixture()
    cam = cfg['cam']
    H, W = int(g['H']), int(g['W'])
    c2w = torch.from_numpy(g['c2w'])
    i, j = torch.meshgrid(torch.arange(W, dtype=torch.float32), torch.arange(H, dtype=torch.float32), indexing='xy')
    dirs = torch.stack([(i - cam['cx']) / cam['fx'], -(j - cam['cy']) / cam['fy'], -torch.ones_like(i)], -1)
    rays_d = torch.sum(dirs[..., None, :] * c2w[:3, :3], -1).reshape(-1, 3)
    rays_o = c2w[:3, -1].expand(rays_d.shape)
    with torch.no_grad():
        a = J.render_rays(cfg, planes, weights, bound, bb, rays_o, rays_d, None, None)
        b = J.render_rays(cfg, planes, weights, bound, bb, rays_o, rays_d, torch.from_numpy(g['gt_depth']).reshape(-1, 1), None)
    assert rel(a['depth'].reshape(H, W).numpy(), g['img_depth_nodepth']) < 1e-4
    assert rel(a['rgb'].reshape(H, W, 3).numpy(), g['img_color_nodepth']) < 1e-4
    assert rel(b['depth'].reshape(H, W).numpy(), g['img_depth']) < 1e-4
    assert rel(b['rgb'].reshape(H, W, 3).numpy(), g['img_color']) < 1e-4
    with torch.no_grad():
        e = J.render_rays(cfg, planes, weights, bound, bb, torch.from_numpy(g['eval_rays_o']), torch.from_numpy(g['eval_rays_d']),
                          torch.from_numpy(g['eval_target_d']), None)
    assert np.array_equal(e['z_vals'].numpy(), g['eval_ret_z_vals'])
    for k in ('rgb', 'depth', 'raw'):
        assert rel(e[k].numpy(), g['eval_ret_' + k]) < 1e-4, k


def test_draw_oracle_is_a_bijection_of_the_row_range():
    """oracle/draw.py (the function the device draw must compute): for every population size the first n elements of the keyed
    sequence are a permutation of range(n) -- so any prefix holds distinct rows, like ``random.sample`` (model/keyframe.py:98);
    different keys give different orders, and the two draws of a batch (keyframes / current frame) use different round keys."""
    from oracle import draw
    for n in (1, 2, 3, 5, 16, 17, 480, 1000, 4097):
        for key in (0, 1, 0x123456789abcdef, (1 << 62) - 1):
            rows = draw.draw_rows(key, n, n)
            assert sorted(rows) == list(range(n)), (n, key)
    a, b = draw.draw_rows(7, 1000, 64), draw.draw_rows(8, 1000, 64)
    assert a != b and len(set(a)) == 64
    both = draw.draw_rows(7, 1000, 32, 1000, 32)
    assert both[:32] == a[:32] and both[32:] != both[:32] and len(set(both[32:])) == 32
    # spread: a short prefix is not clustered at one end of the range
    many = [r for key in range(50) for r in draw.draw_rows(key, 100000, 40)]
    mean = sum(many) / len(many) / 100000
    assert 0.45 < mean < 0.55
