"""GPU parity: the CUDA path (through the C ABI) against the golden fixtures of the reference and against the
CPU oracle on fresh seeded inputs.  Tolerances are BASELINE.json's: sample positions and sign-change
indices bit-exact; rendered maps and losses <= 1e-3 relative (fp32 parameters); gradient cosine >= 0.999."""
import numpy as np
import pytest
import torch

from mneslam_b200 import configs, synthetic
from oracle import joint_encoding as J
from tests.cases import FULL, SMALL_DEPTH, SMALL_NODEPTH, Case
from tests.gpu_util import build_model, cosine, flat_planes, rel, total_loss

pytestmark = pytest.mark.gpu

REL_TOL = 1e-3
COS_TOL = 0.999


def check_grads_against_golden(case, model, rays_o=None, rays_d=None):
    g = case.g
    for w, k in zip(model.decoder.weight_list(), ('sdf0', 'sdf1', 'col0', 'col1')):
        got = w.grad.cpu().numpy()
        assert cosine(got, g['g_' + k]) >= COS_TOL, k
        assert rel(got, g['g_' + k]) < 5e-3, (k, rel(got, g['g_' + k]))
    for i, p in enumerate(flat_planes(model)):
        got = p.grad.cpu()
        if f'g_plane{i}' in g:
            assert cosine(got.numpy(), g[f'g_plane{i}']) >= COS_TOL, i
            assert rel(got.numpy(), g[f'g_plane{i}']) < 5e-3, (i, rel(got.numpy(), g[f'g_plane{i}']))
        else:
            vals = got.contiguous().reshape(-1)[torch.from_numpy(g[f'gp_idx{i}'])].numpy()
            if np.abs(g[f'gp_val{i}']).max() > 0:
                assert cosine(vals, g[f'gp_val{i}']) >= COS_TOL, i
                assert rel(vals, g[f'gp_val{i}']) < 5e-3, (i, rel(vals, g[f'gp_val{i}']))     # same bar as the small cases
            np.testing.assert_allclose(got.double().norm().item(), g[f'gp_stats{i}'][1], rtol=2e-3)
    if rays_o is not None:
        assert cosine(rays_o.grad.cpu().numpy(), g['g_rays_o']) >= COS_TOL
        assert cosine(rays_d.grad.cpu().numpy(), g['g_rays_d']) >= COS_TOL
        assert rel(rays_o.grad.cpu().numpy(), g['g_rays_o']) < 5e-3
        assert rel(rays_d.grad.cpu().numpy(), g['g_rays_d']) < 5e-3


@pytest.mark.parametrize('channels_last', [True, False])
@pytest.mark.parametrize('name', SMALL_DEPTH + FULL)
def test_forward_backward_vs_reference_golden(name, channels_last):
    if name in FULL and not channels_last:
        pytest.skip('strided-layout path is covered by the small cases')
    case = Case(name)
    g, cfg = case.g, case.cfg
    model = build_model(cfg, case.bounding_box, case.planes, case.weights, channels_last)
    rays_o = case.t('rays_o').cuda().requires_grad_(True)
    rays_d = case.t('rays_d').cuda().requires_grad_(True)
    trgb, td, noise = case.t('target_rgb').cuda(), case.t('target_d').cuda(), case.t('noise').cuda()

    rend = model.render_rays(rays_o.detach(), rays_d.detach(), target_d=td, noise=noise)
    assert np.array_equal(rend['z_vals'].detach().cpu().numpy(), g['z_vals']), 'sample positions must be bit-exact'
    for k in ('rgb', 'depth', 'raw', 'acc_map', 'depth_var', 'disp_map'):
        assert rel(rend[k].detach().cpu().numpy(), g[k]) < REL_TOL, (k, rel(rend[k].detach().cpu().numpy(), g[k]))
    _, _, inds = model._render(rays_o.detach(), rays_d.detach(), None, td, False, noise)
    assert np.array_equal(inds.cpu().numpy().astype(np.int64), g['inds']), 'sign-change indices must be bit-exact'

    ret = model.forward(rays_o, rays_d, trgb, td, noise=noise)
    for k in ('rgb_loss', 'depth_loss', 'co_sdf_loss', 'co_fs_loss', 'e_fs_loss', 'e_center_loss', 'e_tail_loss', 'psnr'):
        np.testing.assert_allclose(ret[k].detach().cpu().numpy(), g[k], rtol=REL_TOL, err_msg=k)
    loss = total_loss(cfg, ret)
    np.testing.assert_allclose(loss.item(), g['loss'], rtol=REL_TOL)
    loss.backward()
    check_grads_against_golden(case, model, rays_o, rays_d)


@pytest.mark.parametrize('name', SMALL_NODEPTH)
def test_pose_path_vs_reference_golden(name):
    """Loop-closure alignment path: render_rays(target_d=None), MSE on the rendered maps, gradients reach the
    pose parameters through d loss / d rays (mp_slam/mapper.py:388-410)."""
    case = Case(name)
    g, cfg = case.g, case.cfg
    model = build_model(cfg, case.bounding_box, case.planes, case.weights)
    c2w = case.t('c2w').cuda()
    rot = case.t('pose_rot').cuda().requires_grad_(True)
    trans = c2w[:3, 3].clone().requires_grad_(True)
    from mneslam_b200 import pose
    Rm = pose.axis_angle_to_matrix(rot[None])[0] if 'axis' in name else pose.quaternion_to_matrix(rot[None])[0]
    pose_c2w = torch.cat([Rm @ c2w[:3, :3], trans[:, None]], -1)
    rays_o, rays_d = pose.rays_from_c2w(pose_c2w, case.t('dirs_cam').cuda())
    rays_o.retain_grad(); rays_d.retain_grad()
    rend = model.render_rays(rays_o, rays_d, target_d=None, noise=case.t('noise').cuda())
    assert np.array_equal(rend['z_vals'].detach().cpu().numpy(), g['z_vals'])
    t = cfg['training']
    loss = (t['rgb_weight'] * torch.nn.functional.mse_loss(rend['rgb'], case.t('target_rgb').cuda())
            + t['depth_weight'] * torch.nn.functional.mse_loss(rend['depth'], case.t('target_d').cuda().squeeze(-1)))
    np.testing.assert_allclose(loss.item(), g['loss'], rtol=REL_TOL)
    loss.backward()
    check_grads_against_golden(case, model, rays_o, rays_d)
    assert rel(rot.grad.cpu().numpy(), g['g_pose_rot']) < 5e-3
    assert rel(trans.grad.cpu().numpy(), g['g_pose_trans']) < 5e-3


@pytest.mark.parametrize('two_grid', [False, True])
@pytest.mark.parametrize('n_rays', [1, 37, 300])
def test_fresh_inputs_vs_oracle(two_grid, n_rays):
    """Fresh seeds and ragged sizes against the CPU oracle run on the spot."""
    cfg = configs.small_test_config(one_grid=not two_grid, n_samples_d=13, n_range_d=7)
    cfg['is_co_sdf'] = two_grid
    seed = 100 + n_rays
    bound = J.enlarged_bound(cfg)
    bb = torch.tensor(cfg['mapping']['bound'], dtype=torch.float64)
    planes = J.init_planes(cfg, bound, seed=seed, std=0.5)
    weights = J.init_weights(cfg, seed=seed)
    batch = synthetic.make_ray_batch(cfg, n_rays, seed=seed, zero_depth_frac=0.2, depth_range=(0.3, 2.5))
    S = synthetic.n_samples_total(cfg)
    noise = synthetic.make_noise(n_rays, S, seed)

    op = tuple([p.clone().requires_grad_(True) for p in grp] for grp in planes)
    ow = {k: v.clone().requires_grad_(True) for k, v in weights.items()}
    oret, orend = J.forward(cfg, op, ow, bound, bb, batch['rays_o'], batch['rays_d'], batch['target_rgb'],
                            batch['target_d'], noise)
    oloss = J.total_loss(cfg, oret)

    model = build_model(cfg, bb, planes, weights)
    ret = model.forward(batch['rays_o'].cuda(), batch['rays_d'].cuda(), batch['target_rgb'].cuda(),
                        batch['target_d'].cuda(), noise=noise.cuda())
    for k in ('rgb', 'depth'):
        assert rel(ret[k].detach().cpu().numpy(), orend[k].detach().numpy()) < REL_TOL, k
    finite = True
    for k in ('rgb_loss', 'depth_loss', 'co_sdf_loss', 'co_fs_loss', 'e_fs_loss', 'e_center_loss', 'e_tail_loss'):
        a, b = ret[k].item(), oret[k].item()
        if np.isnan(b):                       # empty mask: the reference's mean() of nothing is nan
            assert np.isnan(a), k
            finite = False
        else:
            np.testing.assert_allclose(a, b, rtol=REL_TOL, atol=1e-7, err_msg=k)
    if not finite:
        return
    oloss.backward()
    total_loss(cfg, ret).backward()
    for p, q in zip(flat_planes(model), [p for grp in op for p in grp]):
        if q.grad.abs().max() > 0:
            assert cosine(p.grad.cpu().numpy(), q.grad.numpy()) >= COS_TOL
    for w, k in zip(model.decoder.weight_list(), ('sdf0', 'sdf1', 'col0', 'col1')):
        assert cosine(w.grad.cpu().numpy(), ow[k].grad.numpy()) >= COS_TOL, k


def test_empty_masks_give_nan_loss_and_finite_grads():
    """All depths invalid: depth_loss and the ESLAM terms are means over empty selections (nan in the reference),
    while the gradients of the remaining terms stay finite."""
    cfg = configs.small_test_config()
    bound = J.enlarged_bound(cfg)
    bb = torch.tensor(cfg['mapping']['bound'], dtype=torch.float64)
    model = build_model(cfg, bb, J.init_planes(cfg, bound, seed=5, std=0.5), J.init_weights(cfg, seed=5))
    batch = synthetic.make_ray_batch(cfg, 64, seed=5, zero_depth_frac=1.1)
    assert float(batch['target_d'].max()) == 0.0
    ret = model.forward(batch['rays_o'].cuda(), batch['rays_d'].cuda(), batch['target_rgb'].cuda(),
                        batch['target_d'].cuda())
    assert torch.isnan(ret['depth_loss']) and torch.isnan(ret['e_fs_loss'])
    assert torch.isfinite(ret['rgb_loss'])
    ret['rgb_loss'].backward()
    for p in flat_planes(model):
        assert torch.isfinite(p.grad).all()


def test_zero_rays_and_query_entry_points():
    cfg = configs.small_test_config(one_grid=False)
    bound = J.enlarged_bound(cfg)
    bb = torch.tensor(cfg['mapping']['bound'], dtype=torch.float64)
    planes, weights = J.init_planes(cfg, bound, seed=3, std=0.5), J.init_weights(cfg, seed=3)
    model = build_model(cfg, bb, planes, weights, as_params=False)
    model.eval()
    pts = (torch.rand(1000, 3) * 2.4 - 1.2)              # some outside the box: border clamp + OneBlob wrap
    raw = model.run_network_flat(pts.cuda()).detach()
    oraw = J.query_points(cfg, planes, weights, bound, bb, pts)
    assert rel(raw.cpu().numpy(), oraw.numpy()) < REL_TOL
    sdf, geo = model.query_sdf(pts.cuda().reshape(100, 10, 3), return_geo=True)
    assert sdf.shape == (100, 10) and geo.shape == (100, 10, 15)
    assert rel(sdf.reshape(-1).cpu().numpy(), oraw[:, 3].numpy()) < REL_TOL
    emb = model.query_sdf(pts.cuda().reshape(1000, 1, 3), embed=True)
    ofeat = J.triplane_features(J.normalize_plane_coords(pts, bound), planes[0], planes[1], planes[2])
    assert emb.shape == (1000, 1, 64) and rel(emb.reshape(-1, 64).cpu().numpy(), ofeat.numpy()) < REL_TOL
    col = model.query_color(pts.cuda().reshape(1000, 1, 3))
    assert col.shape == (1000, 3) and rel(col.cpu().numpy(), torch.sigmoid(oraw[:, :3]).numpy()) < REL_TOL
    empty = model.run_network_flat(torch.zeros(0, 3).cuda())
    assert empty.shape == (0, 4)


def test_full_size_properties_r0():
    """BASELINE config 1/2 shape (Replica room0, 2048 rays x 43 samples): properties that need no oracle."""
    cfg = configs.replica_room0()
    torch.manual_seed(0)
    bb = torch.tensor(cfg['mapping']['bound'], dtype=torch.float64)
    from mneslam_b200 import JointEncoding
    model = JointEncoding(cfg, bb).cuda()
    for grp in model.all_planes:
        for i, p in enumerate(grp):
            grp[i] = torch.nn.Parameter(p.mul_(30.0))
    model.train()
    batch = {k: v.cuda() for k, v in synthetic.make_ray_batch(cfg, 2048, seed=11).items()}
    rend = model.render_rays(batch['rays_o'], batch['rays_d'], target_d=batch['target_d'])
    z = rend['z_vals']
    assert z.shape == (2048, 43) and bool((z[:, 1:] >= z[:, :-1]).all()), 'perturbed samples stay sorted'
    acc = rend['acc_map']
    assert bool(((acc - 1).abs() < 1e-4).all()), 'normalised weights sum to one'
    assert bool((rend['rgb'] >= 0).all()) and bool((rend['rgb'] <= 1).all())
    noise = torch.rand(2048, 43, device='cuda')

    def grads(scale):
        for p in flat_planes(model) + list(model.decoder.parameters()):
            p.grad = None
        ret = model.forward(batch['rays_o'], batch['rays_d'], batch['target_rgb'], batch['target_d'], noise=noise)
        (scale * total_loss(cfg, ret)).backward()
        return [p.grad.clone() for p in flat_planes(model) + list(model.decoder.parameters())], ret
    g1, ret1 = grads(1.0)
    g2, ret2 = grads(2.0)
    assert ret1['rgb_loss'].item() == pytest.approx(ret2['rgb_loss'].item(), rel=1e-6)
    for a, b in zip(g1, g2):                       # linearity in the upstream gradient
        assert torch.allclose(2 * a, b, rtol=1e-3, atol=1e-9 + 1e-4 * float(b.abs().max()))
    # the same step through the fused (non-autograd) iteration driver gives the same gradients
    from mneslam_b200.mapping import FusedMapper
    mapper = FusedMapper(model)
    losses = mapper.forward_backward(batch['rays_o'], batch['rays_d'], batch['target_rgb'], batch['target_d'], noise=noise)
    assert mapper.total_loss(losses).item() == pytest.approx(total_loss(cfg, ret1).item(), rel=1e-5)
    params = flat_planes(model) + list(model.decoder.parameters())
    by_id = {id(p): g for p, g in zip(params, g1)}
    for (p, *_), g in zip(mapper.params, mapper.grads):
        ref = by_id[id(p)]
        assert torch.allclose(g, ref, rtol=1e-3, atol=1e-9 + 1e-4 * float(ref.abs().max()))


def test_importance_sampling_vs_reference_golden():
    """training.n_importance > 0 (dormant in the shipped configs, live code in the reference): searchsorted bin
    indices bit-exact, both passes' maps and the summed losses within 1e-3, gradients of both passes."""
    case = Case('small_importance')
    g, cfg = case.g, case.cfg
    model = build_model(cfg, case.bounding_box, case.planes, case.weights)
    model.u_importance_override = case.t('u_importance').cuda()
    rays_o = case.t('rays_o').cuda().requires_grad_(True)
    rays_d = case.t('rays_d').cuda().requires_grad_(True)
    trgb, td, noise = case.t('target_rgb').cuda(), case.t('target_d').cuda(), case.t('noise').cuda()
    rend = model.render_rays(rays_o.detach(), rays_d.detach(), target_d=td, noise=noise)
    assert np.array_equal(rend['pdf_inds'].cpu().numpy().astype(np.int64), g['pdf_inds']), 'sample-bin indices must be bit-exact'
    np.testing.assert_allclose(rend['z_vals'].cpu().numpy(), g['z_vals'], rtol=0, atol=2e-6)
    for k in ('rgb', 'depth', 'raw', 'rgb0', 'depth0', 'z_std', 'acc_map', 'depth_var'):
        assert rel(rend[k].detach().cpu().numpy(), g[k]) < REL_TOL, (k, rel(rend[k].detach().cpu().numpy(), g[k]))
    ret = model.forward(rays_o, rays_d, trgb, td, noise=noise)
    for k in ('rgb_loss', 'depth_loss', 'e_fs_loss', 'e_center_loss', 'e_tail_loss', 'co_sdf_loss', 'co_fs_loss', 'psnr'):
        np.testing.assert_allclose(ret[k].detach().cpu().numpy(), g[k], rtol=REL_TOL, err_msg=k)
    loss = total_loss(cfg, ret)
    np.testing.assert_allclose(loss.item(), g['loss'], rtol=REL_TOL)
    loss.backward()
    check_grads_against_golden(case, model, rays_o, rays_d)


def test_raw2outputs_and_sdf2weights_entry_points():
    """The compositing helpers the reference exposes as methods, on the golden raw / z_vals of the reference."""
    case = Case('small_onegrid')
    g = case.g
    model = build_model(case.cfg, case.bounding_box, case.planes, case.weights, as_params=False)
    raw, z = case.t('raw').cuda(), case.t('z_vals').cuda()
    rgb, disp, acc, w, depth, var = model.raw2outputs(raw, z)
    for got, k in ((rgb, 'rgb'), (disp, 'disp_map'), (acc, 'acc_map'), (depth, 'depth'), (var, 'depth_var')):
        assert rel(got.cpu().numpy(), g[k]) < REL_TOL, k
    w2 = model.sdf2weights(raw[..., 3].contiguous(), z)
    assert torch.allclose(w, w2) and rel(w.sum(-1).cpu().numpy(), g['acc_map']) < REL_TOL


@pytest.mark.parametrize('two_grid', [False, True])
@pytest.mark.parametrize('mode', ['param', 'ray', 'both'])
def test_backward_with_kept_activations_equals_recompute(two_grid, mode):
    """mne_render_backward with the forward's activation buffer (no recompute) against the same call without it: same
    gradients up to summation order, for parameter gradients, ray gradients and both (two grids with both run on the
    tensor-pipe kernel only with the buffer, on the fp32 v1 kernel without it)."""
    from mneslam_b200 import ops
    cfg = configs.small_test_config(one_grid=not two_grid, n_samples_d=13, n_range_d=7)
    cfg['is_co_sdf'] = two_grid
    bound = J.enlarged_bound(cfg)
    bb = torch.tensor(cfg['mapping']['bound'], dtype=torch.float64)
    planes = J.init_planes(cfg, bound, seed=9, std=0.5)
    weights = J.init_weights(cfg, seed=9)
    model = build_model(cfg, bb, planes, weights, as_params=False)
    n, s = 333, synthetic.n_samples_total(cfg)
    b = {k: v.cuda() for k, v in synthetic.make_ray_batch(cfg, n, seed=21, zero_depth_frac=0.1, depth_range=(0.3, 2.5)).items()}
    noise = synthetic.make_noise(n, s, 21).cuda()
    fp, ws = flat_planes(model), model.decoder.weight_list()
    rcfg = ops.make_render_cfg(cfg, norm_rays=n)
    tables = ops.SampleTables.get(cfg, 'cuda', True)
    td = b['target_d'].reshape(-1)
    acc = ops.new_acc('cuda')
    z = ops.sample_rays(ops.make_sample_cfg(tables, True), rcfg, n, s, td, noise, acc, 'cuda')
    sc = ops.make_scene(fp, [w.detach() for w in ws], model.bound, model.bounding_box)
    out = ops.render_forward(sc, rcfg, b['rays_o'], b['rays_d'], z, b['target_rgb'], td, acc, act=True)
    assert out["act"].numel() == n * s * ops.act_floats(sc)
    lg = torch.tensor([1.0, 0.5, 2.0, 3.0, 2.0, 4.0, 1.5, 0.0], device='cuda')
    want_p, want_r = mode in ('param', 'both'), mode in ('ray', 'both')
    res = []
    for act in (out['act'], None):
        pg = [torch.zeros_like(p) for p in fp] if want_p else None
        wg = [torch.zeros_like(w) for w in ws] if want_p else None
        scb = ops.make_scene(fp, [w.detach() for w in ws], model.bound, model.bounding_box, plane_grads=pg, weight_grads=wg)
        g_o, g_d = ops.render_backward(scb, rcfg, b['rays_o'], b['rays_d'], z, out['raw'], b['target_rgb'], td, acc,
                                       loss_grads=lg, want_ray_grads=want_r, act=act)
        res.append(((pg or []) + (wg or []) + ([g_o, g_d] if want_r else [])))
    torch.cuda.synchronize()
    for a, c in zip(*res):
        a, c = a.double(), c.double()
        assert float(a.abs().max()) > 0
        # same arithmetic, different accumulation order (and v1's fp32 contraction for two grids with both gradient kinds)
        tol = 2e-3 if (two_grid and mode == 'both') else 2e-5
        assert float((a - c).norm() / c.norm()) < tol


@pytest.mark.parametrize('n_rays', [1, 37, 333])
def test_workspaces_are_not_overrun(n_rays):
    """Guard words behind every buffer the kernels write (raw, act, the backward workspace with its sample list, ray
    gradients, z_vals) stay intact for ragged sizes -- compute-sanitizer is not available on the pool, so the bounds are
    checked this way."""
    from mneslam_b200 import _lib, ops
    cfg = configs.small_test_config(one_grid=True, n_samples_d=13, n_range_d=7)
    bound = J.enlarged_bound(cfg)
    bb = torch.tensor(cfg['mapping']['bound'], dtype=torch.float64)
    model = build_model(cfg, bb, J.init_planes(cfg, bound, seed=3, std=0.5), J.init_weights(cfg, seed=3), as_params=False)
    s = synthetic.n_samples_total(cfg)
    n = n_rays
    b = {k: v.cuda() for k, v in synthetic.make_ray_batch(cfg, n, seed=5, zero_depth_frac=0.2, depth_range=(0.3, 2.5)).items()}
    noise = synthetic.make_noise(n, s, 5).cuda()
    GUARD, MAGIC = 64, 12345.0

    def guarded(numel):
        t = torch.full((numel + GUARD,), MAGIC, dtype=torch.float32, device='cuda')
        return t, t[:numel]
    fp, ws = flat_planes(model), [w.detach() for w in model.decoder.weight_list()]
    rcfg = ops.make_render_cfg(cfg, norm_rays=n)
    tables = ops.SampleTables.get(cfg, 'cuda', True)
    td = b['target_d'].reshape(-1)
    acc = ops.new_acc('cuda')
    z_full, z = guarded(n * s)
    z = z.view(n, s)
    ops.sample_rays(ops.make_sample_cfg(tables, True), rcfg, n, s, td, noise, acc, 'cuda', out=z)
    sc = ops.make_scene(fp, ws, model.bound, model.bounding_box)
    act_full, act = guarded(n * s * ops.act_floats(sc))
    out = ops.render_forward(sc, rcfg, b['rays_o'], b['rays_d'], z, b['target_rgb'], td, acc, act=act)
    nbytes = _lib.lib().mne_backward_scratch_bytes(n, s)
    assert nbytes % 4 == 0
    scr_full, scr = guarded(nbytes // 4)
    pg, wg = [torch.zeros_like(p) for p in fp], [torch.zeros_like(w) for w in ws]
    scb = ops.make_scene(fp, ws, model.bound, model.bounding_box, plane_grads=pg, weight_grads=wg)
    lg = torch.tensor([1.0, 0.5, 2.0, 3.0, 2.0, 4.0, 1.5, 0.0], device='cuda')
    g_o, g_d = ops.render_backward(scb, rcfg, b['rays_o'], b['rays_d'], z, out['raw'], b['target_rgb'], td, acc,
                                   loss_grads=lg, want_ray_grads=True, scratch=scr, act=act)
    torch.cuda.synchronize()
    for name, full in (('z_vals', z_full), ('act', act_full), ('scratch', scr_full)):
        assert bool((full[-GUARD:] == MAGIC).all()), name
    n_active = ops.backward_active_count(scr, n, s)
    assert 0 < n_active <= n * s
    lst = scr.view(torch.int32)[n * s * 4: n * s * 4 + n_active]
    assert int(lst.min()) >= 0 and int(lst.max()) < n * s and lst.unique().numel() == n_active
    assert torch.isfinite(g_o).all() and torch.isfinite(g_d).all()


@pytest.mark.parametrize('with_depth', [True, False])
def test_render_img_matches_render_rays(with_depth):
    """render_img (reference model/scene_rep.py:422-473): all pixels of the cropped image through render_rays in chunks;
    (depth [H,W] float64, colour [H,W,3]) must equal one render_rays call over the same rays, whatever the chunking."""
    from mneslam_b200.scene_rep import get_rays
    cfg = configs.small_test_config(one_grid=True, n_samples_d=13, n_range_d=7, n_samples=24)
    cfg['training']['perturb'] = 0
    cfg['cam'].update({'H': 44, 'W': 64, 'fx': 50.0, 'fy': 50.0, 'cx': 31.5, 'cy': 21.5, 'crop_edge': 2})
    bound = J.enlarged_bound(cfg)
    bb = torch.tensor(cfg['mapping']['bound'], dtype=torch.float64)
    model = build_model(cfg, bb, J.init_planes(cfg, bound, seed=3, std=0.5), J.init_weights(cfg, seed=3), as_params=False)
    model.eval()
    H, W = 40, 60
    c2w = torch.eye(4, device='cuda')
    c2w[:3, 3] = torch.tensor([0.1, 0.0, 0.05])
    gt = (torch.rand(H, W, device='cuda') * 2.0 + 0.3) if with_depth else None
    model.render_chunk = 700                                    # ragged chunks (default: at least 64 Ki rays per launch)
    depth, colour = model.render_img(c2w, 'cuda', gt_depth=gt)
    model.render_chunk = None
    depth1, colour1 = model.render_img(c2w, 'cuda', gt_depth=gt)
    assert torch.equal(depth, depth1) and torch.equal(colour, colour1)
    assert depth.shape == (H, W) and depth.dtype == torch.float64 and colour.shape == (H, W, 3)
    ro, rd = get_rays(H, W, 50.0, 50.0, 31.5, 21.5, c2w, 'cuda')
    with torch.no_grad():
        ref = model.render_rays(ro.reshape(-1, 3).contiguous(), rd.reshape(-1, 3).contiguous(),
                                target_d=None if gt is None else gt.reshape(-1, 1))
    assert torch.equal(depth.float().reshape(-1), ref['depth']) and torch.equal(colour.reshape(-1, 3), ref['rgb'])


@pytest.mark.parametrize('one_grid', [True, False])
@pytest.mark.parametrize('n_rays', [1, 333])
def test_fused_render_equals_point_kernel_plus_ray_kernels(one_grid, n_rays):
    """mne_render_forward_fused (one launch: the warp that stores a ray's last sample composites it) against the same
    work as separate launches: mne_render_forward (point kernel + ray kernel) and the ray kernel of mne_render_backward
    (d loss / d raw + the list of samples where it is non-zero).  Same arithmetic, so the values must agree exactly; the
    list may come out in another order."""
    from mneslam_b200 import _lib, ops
    cfg = configs.small_test_config(one_grid=one_grid)
    bound = J.enlarged_bound(cfg)
    bb = torch.tensor(cfg['mapping']['bound'], dtype=torch.float64)
    model = build_model(cfg, bb, J.init_planes(cfg, bound, seed=4, std=0.5), J.init_weights(cfg, seed=4), as_params=False)
    n, s = n_rays, synthetic.n_samples_total(cfg)
    b = {k: v.cuda() for k, v in synthetic.make_ray_batch(cfg, n, seed=8, zero_depth_frac=0.1, depth_range=(0.3, 2.5)).items()}
    noise = synthetic.make_noise(n, s, 8).cuda()
    fp, ws = flat_planes(model), [w.detach() for w in model.decoder.weight_list()]
    rcfg = ops.make_render_cfg(cfg, norm_rays=n)
    td = b['target_d'].reshape(-1)
    tables = ops.SampleTables.get(cfg, 'cuda', True)
    acc0 = ops.new_acc('cuda')
    z = ops.sample_rays(ops.make_sample_cfg(tables, True), rcfg, n, s, td, noise, acc0, 'cuda')
    sc = ops.make_scene(fp, ws, model.bound, model.bounding_box)
    lg = torch.tensor([1.0, 0.5, 2.0, 3.0, 2.0, 4.0, 1.5, 0.0], device='cuda')

    acc_a = acc0.clone()
    ref = ops.render_forward_unfused(sc, rcfg, b['rays_o'], b['rays_d'], z, b['target_rgb'], td, acc_a)
    scr_a = ops.backward_scratch(n, s, 'cuda')
    pg = [torch.zeros_like(p) for p in fp]
    ops.render_backward(ops.make_scene(fp, ws, model.bound, model.bounding_box, plane_grads=pg), rcfg, b['rays_o'], b['rays_d'], z,
                        ref['raw'], b['target_rgb'], td, acc_a, loss_grads=lg, scratch=scr_a)

    acc_b = acc0.clone()
    counters = torch.zeros(n + 7, dtype=torch.int32, device='cuda')
    scr_b = ops.backward_scratch(n, s, 'cuda')
    m = n * s
    cnt_a = ops.backward_active_count(scr_a, n, s)
    la = scr_a.view(torch.int32)[4 * m:4 * m + cnt_a].sort().values
    for single_launch in (True, False):               # with per-ray counters: one launch; without: point kernel + one ray kernel
        acc_b = acc0.clone()
        scr_b = ops.backward_scratch(n, s, 'cuda')
        out = ops.render_forward(sc, rcfg, b['rays_o'], b['rays_d'], z, b['target_rgb'], td, acc_b, want_weights=True,
                                 ray_counters=counters if single_launch else None, loss_grads=lg, scratch=scr_b)
        torch.cuda.synchronize()
        for k in ('raw', 'rgb', 'depth', 'weights', 'disp_map', 'acc_map', 'depth_var', 'inds'):
            assert torch.equal(out[k], ref[k]), (k, single_launch)
        np.testing.assert_allclose(acc_b.cpu().numpy(), acc_a.cpu().numpy(), rtol=1e-6)     # per-ray float sums, then double
        assert int(counters.abs().sum()) == 0, 'the per-ray counters must be left at zero'
        assert torch.equal(scr_b[:4 * m], scr_a[:4 * m]), 'd loss / d raw'
        cnt_b = ops.backward_active_count(scr_b, n, s)
        assert cnt_a == cnt_b > 0
        lb = scr_b.view(torch.int32)[4 * m:4 * m + cnt_b].sort().values
        assert torch.equal(la, lb)
    # a second launch on the same counters (they reset themselves)
    out2 = ops.render_forward(sc, rcfg, b['rays_o'], b['rays_d'], z, ray_counters=counters)
    assert torch.equal(out2['rgb'], ref['rgb']) and int(counters.abs().sum()) == 0


def _inference_model():
    import os
    from oracle.make_golden_inference import inference_config
    from tests.cases import GOLDEN
    g = np.load(os.path.join(GOLDEN, 'small_inference.npz'))
    cfg = inference_config()
    bound = J.enlarged_bound(cfg)
    seed = int(g['seed'])
    bb = torch.tensor(cfg['mapping']['bound'], dtype=torch.float64)
    model = build_model(cfg, bb, J.init_planes(cfg, bound, seed=seed, std=0.4), J.init_weights(cfg, seed=seed), as_params=False)
    return g, cfg, model


def test_inference_entry_points_vs_reference_golden():
    """render_img (with and without gt_depth), render_surface_color and the eval-mode forward against what the REFERENCE'S
    own methods returned (tests/golden/small_inference.npz, oracle/make_golden_inference.py): pins get_rays, the raw
    cam.cx / cam.cy of model/scene_rep.py:438-441, the float64 depth image, ragged chunks."""
    g, cfg, model = _inference_model()
    H, W = int(g['H']), int(g['W'])
    c2w = torch.from_numpy(g['c2w'])
    model.eval()
    model.render_chunk = 100                      # several ragged chunks, like the fixture's ray_batch_size
    d0, c0 = model.render_img(c2w, 'cuda', gt_depth=None)
    d1, c1 = model.render_img(c2w, 'cuda', gt_depth=torch.from_numpy(g['gt_depth']))
    assert d0.dtype == torch.float64 and tuple(d0.shape) == (H, W) and tuple(c0.shape) == (H, W, 3)
    assert rel(d0.cpu().numpy(), g['img_depth_nodepth']) < REL_TOL
    assert rel(c0.cpu().numpy(), g['img_color_nodepth']) < REL_TOL
    assert rel(d1.cpu().numpy(), g['img_depth']) < REL_TOL
    assert rel(c1.cpu().numpy(), g['img_color']) < REL_TOL
    rgb = model.render_surface_color(torch.from_numpy(g['surf_pts']).cuda()[:, None, :], torch.from_numpy(g['surf_normal']).cuda())
    assert tuple(rgb.shape) == tuple(g['surf_rgb'].shape)
    assert rel(rgb.cpu().numpy(), g['surf_rgb']) < REL_TOL
    ret = model.forward(*(torch.from_numpy(g['eval_' + k]).cuda() for k in ('rays_o', 'rays_d', 'target_rgb', 'target_d')))
    assert sorted(ret) == sorted(['rgb', 'depth', 'disp_map', 'acc_map', 'depth_var', 'z_vals', 'raw'])
    assert np.array_equal(ret['z_vals'].cpu().numpy(), g['eval_ret_z_vals'])
    for k in ('rgb', 'depth', 'raw', 'acc_map', 'depth_var', 'disp_map'):
        assert rel(ret[k].detach().cpu().numpy(), g['eval_ret_' + k]) < REL_TOL, k


def test_pose_kernels_vs_reference_axis_angle():
    """mne_pose_rays (rotation parameters -> R -> rays) and pose.axis_angle_to_matrix against the matrices the reference's
    own axis_angle_to_matrix produced (optimization/utils.py:161-177, executed from its source by the fixture script)."""
    from mneslam_b200 import _lib, ops, pose
    g = _inference_model()[0]
    rot, trans, R_ref = torch.from_numpy(g['pose_rot']).cuda(), torch.from_numpy(g['pose_trans']).cuda(), g['pose_R']
    np.testing.assert_allclose(pose.axis_angle_to_matrix(rot).cpu().numpy(), R_ref, rtol=0, atol=5e-7)
    np.testing.assert_allclose(pose.transform_matrix(rot, trans).cpu().numpy(), g['pose_T'], rtol=0, atol=5e-7)
    dirs = torch.eye(3, device='cuda').contiguous()               # rays_d[k] = sum_j e_k[j] R[:, j] = column k of R
    for b in range(rot.shape[0]):
        ro, rd = torch.empty(3, 3, device='cuda'), torch.empty(3, 3, device='cuda')
        _lib.check(_lib.lib().mne_pose_rays(0, rot[b].contiguous().data_ptr(), trans[b].contiguous().data_ptr(), dirs.data_ptr(), 3,
                                            ro.data_ptr(), rd.data_ptr(), ops._stream()), 'mne_pose_rays')
        np.testing.assert_allclose(rd.t().cpu().numpy(), R_ref[b], rtol=0, atol=2e-6)
        np.testing.assert_allclose(ro.cpu().numpy(), np.tile(g['pose_trans'][b], (3, 1)), rtol=0, atol=0)
