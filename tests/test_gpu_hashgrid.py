"""Hash-grid encoder (the encoder BASELINE.json's north star names; dormant in the reference, restated from
tiny-cuda-nn in oracle/hashgrid.py -- parity unpinned by the reference): corner indices bit-exact against the
oracle, forward / losses within 1e-3, table and decoder gradients cosine >= 0.999, and a short training run."""
import ctypes as C

import numpy as np
import pytest
import torch

from mneslam_b200 import JointEncoding, _lib, configs, ops, synthetic
from mneslam_b200.mapping import FusedMapper
from oracle import hashgrid as H
from oracle import joint_encoding as J
from tests.gpu_util import cosine, rel, total_loss

pytestmark = pytest.mark.gpu


def _hash_cfg(hash_size=12, voxel=0.04):
    cfg = configs.small_test_config(n_samples_d=13, n_range_d=7)
    cfg['grid'].update({'use_hash': True, 'hash_size': hash_size, 'voxel_sdf': voxel})
    return cfg


def _model(cfg, seed=0, table_scale=0.3):
    torch.manual_seed(seed)
    bb = torch.tensor(cfg['mapping']['bound'], dtype=torch.float64)
    model = JointEncoding(cfg, bb).cuda()
    meta = H.make_meta(cfg)
    assert meta['offsets'] == model.hash_meta.offsets and meta['resolutions'] == model.hash_meta.resolutions
    table = H.init_table(meta, seed=seed, scale=table_scale)
    weights = J.init_weights(cfg, seed=seed)
    model.hash_tables[0] = torch.nn.Parameter(table.cuda())
    with torch.no_grad():
        for w, k in zip(model.decoder.weight_list(), ('sdf0', 'sdf1', 'col0', 'col1')):
            w.copy_(weights[k])
    model.train()
    return model, meta, table, weights, bb


@pytest.mark.parametrize('hash_size', [8, 13, 16])
def test_corner_indices_bit_exact(hash_size):
    cfg = _hash_cfg(hash_size=hash_size, voxel=0.02)
    model, meta, table, _, _ = _model(cfg)
    x = (torch.rand(3000, 3) * 1.6 - 0.3)          # inside and outside the unit cube (negative cells wrap as uint32)
    desc = _lib.HashDesc()
    model.hash_meta.fill_desc(desc, model.hash_tables[0].detach())
    out = torch.empty(3000, 16, 8, dtype=torch.int32, device='cuda')
    xd = x.cuda()
    _lib.check(_lib.lib().mne_hash_indices(C.byref(desc), 3000, xd.data_ptr(), out.data_ptr(), ops._stream()), 'mne_hash_indices')
    got = out.cpu().numpy()
    hashed_levels = 0
    for l in range(16):
        want, _ = H.corner_indices(meta, x, l)
        assert np.array_equal(got[:, l, :].astype(np.int64), want.numpy()), l
        hashed_levels += int(meta['sizes'][l] < meta['resolutions'][l] ** 3)
    assert hashed_levels > 0 and (hash_size < 12 or hashed_levels < 16), 'dense and hashed levels must both be covered'


def test_forward_backward_vs_oracle():
    cfg = _hash_cfg()
    model, meta, table, weights, bb = _model(cfg, seed=3)
    n = 300
    batch = synthetic.make_ray_batch(cfg, n, seed=8, zero_depth_frac=0.1, depth_range=(0.3, 2.5))
    noise = synthetic.make_noise(n, synthetic.n_samples_total(cfg), seed=8)
    ot = table.clone().requires_grad_(True)
    ow = {k: v.clone().requires_grad_(True) for k, v in weights.items()}
    oret, orend = J.forward(cfg, ([ot],), ow, J.enlarged_bound(cfg), bb, batch['rays_o'], batch['rays_d'], batch['target_rgb'],
                            batch['target_d'], noise)
    J.total_loss(cfg, oret).backward()

    dev = {k: v.cuda() for k, v in batch.items()}
    rend = model.render_rays(dev['rays_o'], dev['rays_d'], target_d=dev['target_d'], noise=noise.cuda())
    assert np.array_equal(rend['z_vals'].cpu().numpy(), orend['z_vals'].numpy())
    for k in ('rgb', 'depth', 'raw'):
        assert rel(rend[k].detach().cpu().numpy(), orend[k].detach().numpy()) < 1e-3, k
    ret = model.forward(dev['rays_o'], dev['rays_d'], dev['target_rgb'], dev['target_d'], noise=noise.cuda())
    for k in ('rgb_loss', 'depth_loss', 'e_fs_loss', 'e_center_loss', 'e_tail_loss'):
        np.testing.assert_allclose(ret[k].item(), oret[k].item(), rtol=1e-3, err_msg=k)
    total_loss(cfg, ret).backward()
    assert cosine(model.hash_tables[0].grad.cpu().numpy(), ot.grad.numpy()) >= 0.999
    assert rel(model.hash_tables[0].grad.cpu().numpy(), ot.grad.numpy()) < 5e-3
    for w, k in zip(model.decoder.weight_list(), ('sdf0', 'sdf1', 'col0', 'col1')):
        assert cosine(w.grad.cpu().numpy(), ow[k].grad.numpy()) >= 0.999, k
    # point queries: 32 hash features for embed=True
    pts = torch.rand(500, 3) * 1.6 - 0.8
    emb = model.query_sdf(pts.cuda().reshape(500, 1, 3), embed=True)
    want = H.encode(meta, table, J.blob_inputs(pts, bb))
    assert emb.shape == (500, 1, 32) and rel(emb.reshape(500, 32).detach().cpu().numpy(), want.numpy()) < 1e-4


@pytest.mark.parametrize('with_params', [True, False])
def test_ray_gradients_through_the_hash_grid(with_params):
    """Pose path: d loss / d rays_o, d rays_d through the table lookup (fractional position) and OneBlob, alone (tracking:
    map frozen) and together with the table / decoder gradients (joint optimisation), against autograd on the oracle."""
    cfg = _hash_cfg()
    model, meta, table, weights, bb = _model(cfg, seed=5)
    n = 200
    batch = synthetic.make_ray_batch(cfg, n, seed=11, zero_depth_frac=0.1, depth_range=(0.3, 2.5))
    noise = synthetic.make_noise(n, synthetic.n_samples_total(cfg), seed=11)
    ot = table.clone().requires_grad_(with_params)
    ow = {k: v.clone().requires_grad_(with_params) for k, v in weights.items()}
    oo, od = batch['rays_o'].clone().requires_grad_(True), batch['rays_d'].clone().requires_grad_(True)
    oret, _ = J.forward(cfg, ([ot],), ow, J.enlarged_bound(cfg), bb, oo, od, batch['target_rgb'], batch['target_d'], noise)
    J.total_loss(cfg, oret).backward()

    if not with_params:
        model.hash_tables[0].requires_grad_(False)
        for w in model.decoder.weight_list():
            w.requires_grad_(False)
    ro, rd = batch['rays_o'].cuda().requires_grad_(True), batch['rays_d'].cuda().requires_grad_(True)
    ret = model.forward(ro, rd, batch['target_rgb'].cuda(), batch['target_d'].cuda(), noise=noise.cuda())
    total_loss(cfg, ret).backward()
    for got, want, name in ((ro.grad, oo.grad, 'rays_o'), (rd.grad, od.grad, 'rays_d')):
        assert cosine(got.cpu().numpy(), want.numpy()) >= 0.999, name
        assert rel(got.cpu().numpy(), want.numpy()) < 5e-3, name
    if with_params:
        assert cosine(model.hash_tables[0].grad.cpu().numpy(), ot.grad.numpy()) >= 0.999
        for w, k in zip(model.decoder.weight_list(), ('sdf0', 'sdf1', 'col0', 'col1')):
            assert cosine(w.grad.cpu().numpy(), ow[k].grad.numpy()) >= 0.999, k
    else:
        assert model.hash_tables[0].grad is None


def test_fused_mapper_trains_the_hash_grid():
    cfg = _hash_cfg(hash_size=14)
    model, *_ = _model(cfg, seed=1, table_scale=1e-4)
    model.hash_tables[0] = model.hash_tables[0].detach()
    mapper = FusedMapper(model)
    gen = torch.Generator().manual_seed(0)
    hist = []
    from mneslam_b200 import pose
    for it in range(150):
        c2w = synthetic.make_pose(cfg, gen, frac=(0.45 + 0.1 * (it % 5) / 4, 0.5, 0.5))
        b = synthetic.make_ray_batch(cfg, 512, seed=it, c2w=c2w)
        rgb, d = synthetic.render_box_scene(cfg, c2w, b['dirs_cam'])
        losses = mapper.step(b['rays_o'].cuda(), b['rays_d'].cuda(), rgb.cuda(), d.cuda())
        hist.append(mapper.total_loss(losses).item())
    assert np.isfinite(hist).all() and np.mean(hist[-10:]) < 0.5 * np.mean(hist[:10]), (hist[:3], hist[-3:])


def _two_table_model(cfg, seed=0, table_scale=0.3):
    torch.manual_seed(seed)
    bb = torch.tensor(cfg['mapping']['bound'], dtype=torch.float64)
    model = JointEncoding(cfg, bb).cuda()
    meta = H.make_meta(cfg)
    cmeta = H.make_meta(cfg, resolution=H.desired_resolution(cfg, 'voxel_color'))
    assert cmeta['offsets'] == model.hash_meta.color.offsets and cmeta['resolutions'] == model.hash_meta.color.resolutions
    assert cmeta['per_level_scale'] != meta['per_level_scale']
    table, ctable = H.init_table(meta, seed=seed, scale=table_scale), H.init_table(cmeta, seed=seed + 7, scale=table_scale)
    weights = J.init_weights(cfg, seed=seed)
    assert tuple(weights['col0'].shape) == (32, 95)
    model.hash_tables[0] = torch.nn.Parameter(table.cuda())
    model.hash_tables[1] = torch.nn.Parameter(ctable.cuda())
    with torch.no_grad():
        for w, k in zip(model.decoder.weight_list(), ('sdf0', 'sdf1', 'col0', 'col1')):
            w.copy_(weights[k])
    model.train()
    return model, table, ctable, weights, bb


def test_two_tables_forward_backward_and_ray_gradients_vs_oracle():
    """grid.oneGrid: False on the hash grid: an sdf table and a colour table (resolution from voxel_color, the commented
    model/scene_rep.py:169), colour decoder fed [blob | colour features | geo].  Forward, losses, both table gradients, decoder
    gradients and ray gradients against autograd on the oracle; then a short FusedMapper run."""
    cfg = _hash_cfg()
    cfg['grid']['oneGrid'] = False
    cfg['is_co_sdf'] = True
    model, table, ctable, weights, bb = _two_table_model(cfg, seed=3)
    n = 300
    batch = synthetic.make_ray_batch(cfg, n, seed=8, zero_depth_frac=0.1, depth_range=(0.3, 2.5))
    noise = synthetic.make_noise(n, synthetic.n_samples_total(cfg), seed=8)
    ot, oc = table.clone().requires_grad_(True), ctable.clone().requires_grad_(True)
    ow = {k: v.clone().requires_grad_(True) for k, v in weights.items()}
    oo, od = batch['rays_o'].clone().requires_grad_(True), batch['rays_d'].clone().requires_grad_(True)
    oret, orend = J.forward(cfg, ([ot], [oc]), ow, J.enlarged_bound(cfg), bb, oo, od, batch['target_rgb'], batch['target_d'], noise)
    J.total_loss(cfg, oret).backward()

    ro, rd = batch['rays_o'].cuda().requires_grad_(True), batch['rays_d'].cuda().requires_grad_(True)
    ret = model.forward(ro, rd, batch['target_rgb'].cuda(), batch['target_d'].cuda(), noise=noise.cuda())
    for k in ('rgb', 'depth'):
        assert rel(ret[k].detach().cpu().numpy(), orend[k].detach().numpy()) < 1e-3, k
    for k in ('rgb_loss', 'depth_loss', 'co_sdf_loss', 'co_fs_loss'):
        np.testing.assert_allclose(ret[k].item(), oret[k].item(), rtol=1e-3, err_msg=k)
    total_loss(cfg, ret).backward()
    for got, want, name in ((model.hash_tables[0].grad, ot.grad, 'sdf table'), (model.hash_tables[1].grad, oc.grad, 'colour table'),
                            (ro.grad, oo.grad, 'rays_o'), (rd.grad, od.grad, 'rays_d')):
        assert cosine(got.cpu().numpy(), want.numpy()) >= 0.999, name
        assert rel(got.cpu().numpy(), want.numpy()) < 5e-3, name
    for w, k in zip(model.decoder.weight_list(), ('sdf0', 'sdf1', 'col0', 'col1')):
        assert cosine(w.grad.cpu().numpy(), ow[k].grad.numpy()) >= 0.999, k

    mapper = FusedMapper(model)
    first = last = None
    for it in range(40):
        b = synthetic.make_ray_batch(cfg, 256, seed=100 + it % 4)
        losses = mapper.step(b['rays_o'].cuda(), b['rays_d'].cuda(), b['target_rgb'].cuda(), b['target_d'].cuda())
        v = float(mapper.total_loss(losses))
        first = v if first is None else first
        last = v
    assert np.isfinite(last) and last < first


def test_full_size_scannet_hash_grid_vs_oracle():
    """BASELINE.json configs[3] at its real table size: ScanNet scene0000 bound and sampling (96 + 21 samples per ray), 16
    levels, T = 2^19 -- 256 rays against the oracle: corner indices bit-exact at every level (dense and hashed), forward and
    losses within 1e-3, table / decoder / ray gradients cosine >= 0.999."""
    cfg = configs.scannet_scene0000_onegrid()
    cfg['grid']['use_hash'] = True
    model, meta, table, weights, bb = _model(cfg, seed=4, table_scale=0.3)
    assert cfg['grid']['hash_size'] == 19 and max(meta['sizes']) == 1 << 19
    n, s = 256, synthetic.n_samples_total(cfg)
    assert s == 117
    batch = synthetic.make_ray_batch(cfg, n, seed=14, zero_depth_frac=0.1)
    noise = synthetic.make_noise(n, s, seed=14)
    ot = table.clone().requires_grad_(True)
    ow = {k: v.clone().requires_grad_(True) for k, v in weights.items()}
    oo, od = batch['rays_o'].clone().requires_grad_(True), batch['rays_d'].clone().requires_grad_(True)
    oret, orend = J.forward(cfg, ([ot],), ow, J.enlarged_bound(cfg), bb, oo, od, batch['target_rgb'], batch['target_d'], noise)
    J.total_loss(cfg, oret).backward()
    # corner indices of the batch's own sample positions
    pts = (batch['rays_o'][:, None, :] + batch['rays_d'][:, None, :] * orend['z_vals'][..., None]).reshape(-1, 3)
    x01 = J.blob_inputs(pts, bb)[:4000].contiguous()
    desc = _lib.HashDesc()
    model.hash_meta.fill_desc(desc, model.hash_tables[0].detach())
    out = torch.empty(x01.shape[0], 16, 8, dtype=torch.int32, device='cuda')
    xd = x01.cuda()
    _lib.check(_lib.lib().mne_hash_indices(C.byref(desc), x01.shape[0], xd.data_ptr(), out.data_ptr(), ops._stream()), 'mne_hash_indices')
    got = out.cpu().numpy()
    for l in range(16):
        want, _ = H.corner_indices(meta, x01, l)
        assert np.array_equal(got[:, l, :].astype(np.int64), want.numpy()), l
    ro, rd = batch['rays_o'].cuda().requires_grad_(True), batch['rays_d'].cuda().requires_grad_(True)
    ret = model.forward(ro, rd, batch['target_rgb'].cuda(), batch['target_d'].cuda(), noise=noise.cuda())
    for k in ('rgb_loss', 'depth_loss', 'e_fs_loss', 'e_center_loss', 'e_tail_loss'):
        np.testing.assert_allclose(ret[k].item(), oret[k].item(), rtol=1e-3, err_msg=k)
    assert rel(ret['rgb'].detach().cpu().numpy(), oret['rgb'].detach().numpy()) < 1e-3
    assert rel(ret['depth'].detach().cpu().numpy(), oret['depth'].detach().numpy()) < 1e-3
    total_loss(cfg, ret).backward()
    assert cosine(model.hash_tables[0].grad.cpu().numpy(), ot.grad.numpy()) >= 0.999
    for w, k in zip(model.decoder.weight_list(), ('sdf0', 'sdf1', 'col0', 'col1')):
        assert cosine(w.grad.cpu().numpy(), ow[k].grad.numpy()) >= 0.999, k
    assert cosine(ro.grad.cpu().numpy(), oo.grad.numpy()) >= 0.999 and cosine(rd.grad.cpu().numpy(), od.grad.numpy()) >= 0.999


def test_fp16_table_option():
    """grid.hash_fp16: the kernels read the features from an fp16 copy of the table (half2 loads).  Against the oracle run on
    the fp16-ROUNDED table the usual 1e-3 gates hold (the arithmetic after the load is unchanged); against the fp32 table the
    BASELINE.json gate for fp16 tables, 1e-2.  Gradients stay fp32 and flow to the fp32 master table."""
    cfg = _hash_cfg()
    cfg['grid']['hash_fp16'] = True
    model, meta, table, weights, bb = _model(cfg, seed=6)
    assert model.hash_fp16
    n = 300
    batch = synthetic.make_ray_batch(cfg, n, seed=9, zero_depth_frac=0.1, depth_range=(0.3, 2.5))
    noise = synthetic.make_noise(n, synthetic.n_samples_total(cfg), seed=9)
    results = {}
    for name, tab in (('rounded', table.half().float()), ('fp32', table)):
        ot = tab.clone().requires_grad_(True)
        ow = {k: v.clone().requires_grad_(True) for k, v in weights.items()}
        oret, orend = J.forward(cfg, ([ot],), ow, J.enlarged_bound(cfg), bb, batch['rays_o'], batch['rays_d'], batch['target_rgb'],
                                batch['target_d'], noise)
        J.total_loss(cfg, oret).backward()
        results[name] = (oret, orend, ot, ow)
    dev = {k: v.cuda() for k, v in batch.items()}
    ret = model.forward(dev['rays_o'], dev['rays_d'], dev['target_rgb'], dev['target_d'], noise=noise.cuda())
    assert model.hash_tables_h is not None and model.hash_tables_h[0].dtype == torch.float16
    total_loss(cfg, ret).backward()
    oret, orend, ot, ow = results['rounded']
    for k in ('rgb_loss', 'depth_loss', 'e_fs_loss', 'e_center_loss', 'e_tail_loss'):
        np.testing.assert_allclose(ret[k].item(), oret[k].item(), rtol=1e-3, err_msg=k)
    assert rel(ret['rgb'].detach().cpu().numpy(), oret['rgb'].detach().numpy()) < 1e-3
    assert cosine(model.hash_tables[0].grad.cpu().numpy(), ot.grad.numpy()) >= 0.999
    for w, k in zip(model.decoder.weight_list(), ('sdf0', 'sdf1', 'col0', 'col1')):
        assert cosine(w.grad.cpu().numpy(), ow[k].grad.numpy()) >= 0.999, k
    oret32 = results['fp32'][0]
    for k in ('rgb_loss', 'depth_loss', 'e_fs_loss', 'e_center_loss', 'e_tail_loss'):
        np.testing.assert_allclose(ret[k].item(), oret32[k].item(), rtol=1e-2, err_msg=k)
    assert rel(ret['depth'].detach().cpu().numpy(), oret32['depth'].detach().numpy()) < 1e-2
    # the fused iteration driver keeps the copy in step with the master table
    mapper = FusedMapper(model)
    before = model.hash_tables[0].detach().clone()
    mapper.step(dev['rays_o'], dev['rays_d'], dev['target_rgb'], dev['target_d'], noise=noise.cuda())
    mapper.step(dev['rays_o'], dev['rays_d'], dev['target_rgb'], dev['target_d'], noise=noise.cuda())
    torch.cuda.synchronize()
    assert not torch.equal(before, model.hash_tables[0])
    model.refresh_hash_half()
    assert torch.equal(model.hash_tables_h[0], model.hash_tables[0].detach().half())


def test_level0_in_shared_memory_option_gives_the_same_features():
    """mne_set_option("hash_level0_smem", 1): level 0 of the table staged in shared memory by the forward kernel (off by
    default -- measured slower on B200) must not change a bit of the result."""
    cfg = _hash_cfg(hash_size=14, voxel=0.02)
    model, meta, table, weights, bb = _model(cfg, seed=2)
    pts = (torch.rand(4000, 3) * 1.8 - 0.9).cuda()
    lib = _lib.lib()
    try:
        a = model.query_color_sdf(pts).clone()
        _lib.check(lib.mne_set_option(b'hash_level0_smem', 1), 'set_option')
        b = model.query_color_sdf(pts).clone()
    finally:
        _lib.check(lib.mne_set_option(b'hash_level0_smem', 0), 'set_option')
    assert torch.equal(a, b)
