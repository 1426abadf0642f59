"""GPU tests of the iteration driver: fused Adam against torch.optim.Adam, device ray assembly, CUDA-graph replay."""
import numpy as np
import pytest
import torch

from mneslam_b200 import configs, synthetic
from oracle import joint_encoding as J
from mneslam_b200._lib import LOSS_KEYS
from tests.gpu_util import build_model, cosine, flat_planes, rel, total_loss

pytestmark = pytest.mark.gpu


def _small_model(seed=3, one_grid=True):
    cfg = configs.small_test_config(one_grid=one_grid)
    bound = J.enlarged_bound(cfg)
    bb = torch.tensor(cfg['mapping']['bound'], dtype=torch.float64)
    planes, weights = J.init_planes(cfg, bound, seed=seed, std=0.3), J.init_weights(cfg, seed=seed)
    return cfg, bb, planes, weights


def _five_iterations(one_grid):
    """Bulk relative parameter differences (planes, decoder weights) between FusedMapper and the drop-in class driven
    the way the reference drives it, after five mapping iterations from the same start."""
    from mneslam_b200.mapping import FusedMapper
    cfg, bb, planes, weights = _small_model(one_grid=one_grid)
    a = build_model(cfg, bb, planes, weights)
    b = build_model(cfg, bb, planes, weights, as_params=False)
    mp = cfg['mapping']
    groups = [{'params': a.decoder.parameters(), 'weight_decay': 1e-6, 'lr': mp['lr_decoder']},
              {'params': flat_planes(a)[:6], 'eps': 1e-15, 'lr': mp['lr_embed']}]
    if not one_grid:
        groups.append({'params': flat_planes(a)[6:], 'eps': 1e-15, 'lr': mp['lr_embed_color']})
    opt = torch.optim.Adam(groups, betas=(0.9, 0.99))
    mapper = FusedMapper(b)
    n, s = 200, synthetic.n_samples_total(cfg)
    for it in range(5):
        batch = {k: v.cuda() for k, v in synthetic.make_ray_batch(cfg, n, seed=50 + it).items()}
        noise = synthetic.make_noise(n, s, seed=it).cuda()
        ret = a.forward(batch['rays_o'], batch['rays_d'], batch['target_rgb'], batch['target_d'], noise=noise)
        loss = total_loss(cfg, ret)
        loss.backward()
        opt.step()
        opt.zero_grad()
        losses = mapper.step(batch['rays_o'], batch['rays_d'], batch['target_rgb'], batch['target_d'], noise=noise)
        assert mapper.total_loss(losses).item() == pytest.approx(loss.item(), rel=2e-4)
    assert float(mapper.flat_grad.abs().max()) == 0.0, 'the Adam kernel leaves the gradients zeroed'

    def bulk(x, y):
        x, y = x.detach().double().reshape(-1), y.detach().double().reshape(-1)
        return float((x - y).norm() / (y.norm() + 1e-30))
    return (max(bulk(pb, pa) for pa, pb in zip(flat_planes(a), flat_planes(b))),
            max(bulk(wb, wa) for wa, wb in zip(a.decoder.weight_list(), b.decoder.weight_list())))


@pytest.mark.parametrize('one_grid', [True, False])
def test_fused_iterations_match_autograd_plus_torch_adam(one_grid):
    """Five mapping iterations: FusedMapper (sample/fwd/bwd/Adam kernels) against the drop-in class driven the
    way the reference drives it (forward -> get_loss_from_ret -> backward -> torch.optim.Adam, create_optimizer's
    parameter groups).

    Both arms sum their gradients with float atomics, so two runs never see bit-identical gradients, and a trajectory
    is not a smooth function of them: Adam's early steps are sign-like (m / sqrt(v) ~ g / |g|) and the compositing
    has discrete decisions (first sign change of the sdf, truncation masks).  Measured over repeated runs
    (scripts/traj_noise.py): the bulk difference is ~1e-6, with an occasional run at ~2e-3 when one of those decisions
    falls the other way.  So: up to three attempts, the best one must be within 2e-5 (any systematic error -- moments,
    bias correction, parameter groups, zeroing -- shows in every attempt), and none may be off by more than 5e-2."""
    best = (1.0, 1.0)
    for attempt in range(3):
        pe, we = _five_iterations(one_grid)
        assert pe < 5e-2 and we < 5e-2, (pe, we)
        best = min(best, (pe, we), key=max)
        if max(best) < 2e-5:
            break
    assert max(best) < 2e-5, best


def test_adam_kernel_matches_torch_single_tensor():
    from mneslam_b200 import _lib, ops
    torch.manual_seed(0)
    n = 4096 * 3 + 8
    p = torch.randn(n, device='cuda')
    ref = torch.nn.Parameter(p.clone())
    opt = torch.optim.Adam([ref], lr=5e-3, betas=(0.9, 0.99), eps=1e-15, foreach=False)
    m, v = torch.zeros_like(p), torch.zeros_like(p)
    for t in range(1, 6):
        g = torch.randn(n, device='cuda') * (0.1 if t % 2 else 10.0)
        ref.grad = g.clone()
        opt.step()
        gg = g.clone()
        _lib.check(_lib.lib().mne_adam_step(p.data_ptr(), gg.data_ptr(), m.data_ptr(), v.data_ptr(), n, 5e-3, 0.9, 0.99,
                                            1e-15, 0.0, None, t, 1, ops._stream()), 'adam')
        assert float(gg.abs().max()) == 0.0
    torch.cuda.synchronize()
    # same formulas as torch's single-tensor Adam; fp32 rounding order differs by a few ulp over five steps
    assert rel(p.cpu().numpy(), ref.detach().cpu().numpy()) < 5e-6
    assert rel(v.cpu().numpy(), opt.state[ref]['exp_avg_sq'].cpu().numpy()) < 5e-6


def test_device_keyframe_assembly_matches_caller_code():
    """rays_d = sum_k dirs_k R[ids][:, k], rays_o = t[ids] (mp_slam/mapper.py:150-153)."""
    from mneslam_b200.mapping import DeviceKeyframeRays
    cfg = configs.small_test_config()
    cfg['cam'].update({'H': 40, 'W': 60, 'fx': 50.0, 'fy': 50.0, 'cx': 29.5, 'cy': 19.5})
    rays, poses = synthetic.make_keyframe_db(cfg, n_kf=4, seed=2)
    db = DeviceKeyframeRays(rays.cuda(), poses.cuda())
    idx = torch.randint(0, rays.shape[0] * rays.shape[1], (333,))
    ids = idx // rays.shape[1]
    r7 = rays.reshape(-1, 7)[idx]
    ro, rd, rgb, d = db.assemble(r7.cuda(), ids.cuda())
    want_d = torch.sum(r7[:, None, None, :3] * poses[ids, None, :3, :3], -1).reshape(-1, 3)
    want_o = poses[ids, None, :3, -1].repeat(1, 1, 1).reshape(-1, 3)
    assert torch.allclose(rd.cpu(), want_d, atol=1e-6) and torch.equal(ro.cpu(), want_o)
    assert torch.equal(rgb.cpu(), r7[:, 3:6]) and torch.equal(d.cpu(), r7[:, 6:7])
    ro2, rd2, rgb2, d2 = db.sample(128)
    assert ro2.shape == (128, 3) and d2.shape == (128, 1)


def test_device_draw_is_without_replacement_and_adds_the_current_frame():
    """The batch of mapping_optimize (mp_slam/mapper.py:135-153): ``random.sample`` of the stored rays (no row twice,
    model/keyframe.py:98) followed by max(sample // n_kf, min_pixels_cur) distinct pixels of the current frame posed by the
    last pose (id -1) -- drawn and assembled on the device in one launch."""
    from mneslam_b200 import _lib
    from mneslam_b200.mapping import DeviceKeyframeRays
    cfg = configs.small_test_config()
    cfg['cam'].update({'H': 40, 'W': 60, 'fx': 50.0, 'fy': 50.0, 'cx': 29.5, 'cy': 19.5})
    rays, poses = synthetic.make_keyframe_db(cfg, n_kf=4, seed=2)
    total = rays.shape[0] * rays.shape[1]
    g = torch.Generator().manual_seed(3)
    current = torch.rand(40 * 60, 7, generator=g)
    db = DeviceKeyframeRays(rays.cuda(), poses.cuda(), current=current.cuda(), min_pixels_cur=100)
    assert db.n_current(2048) == 512 and db.n_current(256) == 100 and db.batch_rays(256) == 356
    flat = rays.reshape(-1, 7)
    seen = []
    for n in (256, total):                               # a normal batch, and the whole population (a full permutation)
        ro, rd, rgb, d, rows = db.sample(n, return_rows=True)
        n_cur = db.n_current(n)
        rows = rows.cpu()
        rk, rc = rows[:n], rows[n:]
        assert ro.shape == (n + n_cur, 3) and rc.numel() == n_cur
        assert int(rk.min()) >= 0 and int(rk.max()) < total and torch.unique(rk).numel() == n
        assert int(rc.min()) >= 0 and int(rc.max()) < current.shape[0] and torch.unique(rc).numel() == n_cur
        r7 = torch.cat([flat[rk], current[rc]])
        ids = torch.cat([rk // rays.shape[1], torch.full((n_cur,), -1, dtype=torch.int64)])
        want_d = torch.sum(r7[:, None, None, :3] * poses[ids, None, :3, :3], -1).reshape(-1, 3)
        assert torch.allclose(rd.cpu(), want_d, atol=1e-6) and torch.equal(ro.cpu(), poses[ids, :3, -1])
        assert torch.equal(rgb.cpu(), r7[:, 3:6]) and torch.equal(d.cpu(), r7[:, 6:7])
        seen.append(rk)
    assert torch.equal(torch.sort(seen[1]).values, torch.arange(total))
    # the rows are exactly those of the CPU restatement for the same key (index work: bit-exact)
    from oracle import draw
    for k in (0, 12345678901234567, (1 << 62) - 1):
        key = torch.tensor([k], dtype=torch.int64, device='cuda')
        rows = db.sample(200, return_rows=True, key=key)[4].cpu().tolist()
        assert rows == draw.draw_rows(k, total, 200, current.shape[0], db.n_current(200))
    # a fresh key gives a fresh draw; the draw spreads over the whole store
    a = db.sample(256, return_rows=True)[4].cpu()
    b = db.sample(256, return_rows=True)[4].cpu()
    assert not torch.equal(a, b)
    many = torch.cat([db.sample(128, return_rows=True)[4][:128].cpu() for _ in range(160)]).double()
    assert abs(float(many.mean()) / total - 0.5) < 0.02 and abs(float(many.std()) / total - 12 ** -0.5) < 0.02
    # larger than the population: an error, as random.sample raises
    with pytest.raises(_lib.MneError):
        db.sample(total + 1)


def test_cuda_graph_replay_of_a_mapping_iteration():
    """The whole iteration (ray assembly, sample, forward, losses, backward, Adam) captured once and replayed:
    the loss keeps decreasing and no launch needs the host."""
    from mneslam_b200.mapping import DeviceKeyframeRays, FusedMapper
    cfg, bb, planes, weights = _small_model(seed=9)
    cfg['cam'].update({'H': 40, 'W': 60, 'fx': 50.0, 'fy': 50.0, 'cx': 29.5, 'cy': 19.5})
    model = build_model(cfg, bb, planes, weights, as_params=False)
    mapper = FusedMapper(model)
    rays, poses = synthetic.make_keyframe_db(cfg, n_kf=3, seed=2)
    db = DeviceKeyframeRays(rays.cuda(), poses.cuda())

    def one():
        return mapper.step(*db.sample(256))
    for _ in range(3):
        one()
    torch.cuda.synchronize()
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        one()
    torch.cuda.current_stream().wait_stream(side)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        losses = one()
    first = None
    hist = []
    for i in range(60):
        g.replay()
        hist.append(mapper.total_loss(losses).item())
    assert np.isfinite(hist).all()
    assert np.mean(hist[-10:]) < np.mean(hist[:10]), (hist[:3], hist[-3:])
    assert int(mapper.step_dev.item()) == 3 + 1 + 60      # warm-up, side-stream run, replays (capture runs nothing)


def test_two_gpu_sharded_iteration_matches_single_gpu():
    """Runs tests/multigpu_check.py under torchrun when two GPUs are visible (skipped on the one-GPU box)."""
    import os
    import subprocess
    import sys
    if torch.cuda.device_count() < 2:
        pytest.skip('needs two GPUs')
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node', '2', '--master-addr',
                        '127.0.0.1', '--master-port', '29533', os.path.join(root, 'tests', 'multigpu_check.py')],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and 'multigpu_check ok' in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]


def test_three_agents_exchange_maps_and_distill():
    """Runs tests/multiagent_check.py (BASELINE.json configs[4] in small: three agents, one per GPU, ring exchange of the maps
    over NVLink -- NCCL and peer stores -- then joint distillation) when three GPUs are visible."""
    import os
    import subprocess
    import sys
    if torch.cuda.device_count() < 3:
        pytest.skip('needs three GPUs')
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node', '3', '--master-addr',
                        '127.0.0.1', '--master-port', '29534', os.path.join(root, 'tests', 'multiagent_check.py')],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and 'multiagent_check ok' in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]


def test_touched_map_adam_is_bit_identical_to_dense():
    """Skipping the 16-byte groups that never received a gradient must not change a single bit."""
    import ctypes as C
    from mneslam_b200 import _lib, ops
    torch.manual_seed(1)
    n = 1 << 16
    p0 = torch.randn(n, device='cuda')
    ends = (C.c_int64 * 2)(n - 4096, n)
    lr, eps, wd = (C.c_float * 2)(5e-3, 1e-2), (C.c_float * 2)(1e-15, 1e-8), (C.c_float * 2)(0.0, 1e-6)
    state = []
    for use_map in (False, True):
        p, m, v = p0.clone(), torch.zeros(n, device='cuda'), torch.zeros(n, device='cuda')
        touched = torch.zeros(n // 4, dtype=torch.uint8, device='cuda') if use_map else None
        gen = torch.Generator(device='cuda').manual_seed(7)
        for t in range(1, 7):
            g = torch.randn(n, device='cuda', generator=gen)
            mask = torch.rand(n // 64, device='cuda', generator=gen) < (0.1 * t)      # growing set of touched 64-float tiles
            g = (g.reshape(-1, 64) * mask[:, None]).reshape(-1).contiguous()
            _lib.check(_lib.lib().mne_adam_step_segments(p.data_ptr(), g.data_ptr(), m.data_ptr(), v.data_ptr(), 2, ends, lr, eps,
                                                         wd, 0.9, 0.99, None, t, 1, ops._ptr(touched), ops._stream()), 'adam')
            assert float(g.abs().max()) == 0.0
        state.append((p, m, v, touched))
    (pd, md, vd, _), (ps, ms, vs, touched) = state
    assert torch.equal(pd, ps) and torch.equal(md, ms) and torch.equal(vd, vs)
    frac = touched[: (n - 4096) // 4].float().mean().item()
    assert 0.2 < frac < 0.95, frac


def test_peer_barrier_kernel_signals_and_bounded_wait():
    """mne_peer_barrier with both "ranks" on ONE device (two pads).  Two kernels that wait for each other need to be resident
    together, which one device does not promise, so the hand-shake is walked through one side at a time: a rank that arrives
    alone signals its partner, gives up after the bounded wait and raises the error word of ITS pad; the partner, arriving
    later, finds the signal and returns at once.  (Ranks that really run side by side: tests/multigpu_check.py, which asserts
    that no barrier of a sharded run timed out.)"""
    import ctypes as C
    from mneslam_b200 import _lib
    lib = _lib.lib()
    n_ch = 4
    words = _lib.peer_pad_words(n_ch)
    pads = [torch.zeros(words, dtype=torch.int32, device='cuda') for _ in range(2)]
    ptrs = (C.c_void_p * 2)(*[p.data_ptr() for p in pads])
    one = (C.c_void_p * 1)(pads[0].data_ptr())
    n_arr = n_ch * _lib.MAX_PEERS
    # a single rank: the barrier is a self-signal; epochs count the calls per channel
    for it in range(6):
        _lib.check(lib.mne_peer_barrier(one, 1, 0, it % 2, n_ch, None), 'mne_peer_barrier')
    torch.cuda.synchronize()
    assert pads[0][n_arr:n_arr + n_ch].tolist() == [3, 3, 0, 0] and int(pads[0][-1]) == 0
    assert int(pads[0][0]) == 3 and int(pads[0][_lib.MAX_PEERS]) == 3
    pads[0].zero_()
    _lib.check(lib.mne_set_option(b'peer_barrier_polls_log2', 14), 'mne_set_option')
    try:
        for epoch in (1, 2):
            # rank 1 arrives alone on channel 2: signals rank 0, waits in vain, raises its own error word
            _lib.check(lib.mne_peer_barrier(ptrs, 2, 1, 2, n_ch, None), 'mne_peer_barrier')
            torch.cuda.synchronize()
            assert int(pads[1][-1]) == 1 and int(pads[0][-1]) == 0
            assert int(pads[0][2 * _lib.MAX_PEERS + 1]) == epoch          # rank 1's arrival word in rank 0's pad
            pads[1][-1] = 0
            # rank 0 arrives: rank 1's signal of this epoch is already there
            _lib.check(lib.mne_peer_barrier(ptrs, 2, 0, 2, n_ch, None), 'mne_peer_barrier')
            torch.cuda.synchronize()
            assert int(pads[0][-1]) == 0 and int(pads[1][2 * _lib.MAX_PEERS + 0]) == epoch
            assert int(pads[0][n_arr + 2]) == epoch and int(pads[1][n_arr + 2]) == epoch
    finally:
        _lib.check(lib.mne_set_option(b'peer_barrier_polls_log2', 29), 'mne_set_option')


def test_texel_flag_adam_is_bit_identical_to_dense():
    """mne_adam_step_segments_dirty: with the backward pass's texel flags (one byte per 32 floats of the plane range) the
    optimiser reads the gradient of flagged texels only.  Same bits as the dense kernel, gradients and flags cleared; flagged
    texels whose gradient is all zero, an unflagged dense tail and a ragged end are all in the case."""
    import ctypes as C
    from mneslam_b200 import _lib, ops
    torch.manual_seed(2)
    n_planes, n = 1 << 16, (1 << 16) + 4096 + 8
    p0 = torch.randn(n, device='cuda')
    ends = (C.c_int64 * 2)(n_planes, n)
    lr, eps, wd = (C.c_float * 2)(5e-3, 1e-2), (C.c_float * 2)(1e-15, 1e-8), (C.c_float * 2)(0.0, 1e-6)
    state = []
    for use_flags in (False, True):
        p, m, v = p0.clone(), torch.zeros(n, device='cuda'), torch.zeros(n, device='cuda')
        touched = torch.zeros(n // 4, dtype=torch.uint8, device='cuda')
        dirty = torch.zeros(n_planes // 32, dtype=torch.uint8, device='cuda')
        gen = torch.Generator(device='cuda').manual_seed(7)
        for t in range(1, 7):
            g = torch.randn(n, device='cuda', generator=gen)
            mask = torch.rand(n_planes // 32, device='cuda', generator=gen) < (0.08 * t)
            extra = torch.rand(n_planes // 32, device='cuda', generator=gen) < 0.05        # flagged, but only zeros were added
            g[:n_planes] = (g[:n_planes].reshape(-1, 32) * mask[:, None]).reshape(-1)
            if use_flags:
                dirty.copy_((mask | extra).to(torch.uint8))
                _lib.check(_lib.lib().mne_adam_step_segments_dirty(p.data_ptr(), g.data_ptr(), m.data_ptr(), v.data_ptr(), 2, ends, lr, eps, wd,
                                                                   0.9, 0.99, None, t, 1, touched.data_ptr(), dirty.data_ptr(), n_planes,
                                                                   ops._stream()), 'adam')
                assert int(dirty.sum()) == 0
            else:
                _lib.check(_lib.lib().mne_adam_step_segments(p.data_ptr(), g.data_ptr(), m.data_ptr(), v.data_ptr(), 2, ends, lr, eps, wd,
                                                             0.9, 0.99, None, t, 1, touched.data_ptr(), ops._stream()), 'adam')
            assert float(g.abs().max()) == 0.0
        state.append((p, m, v, touched))
    (pd, md, vd, td), (ps, ms, vs, ts) = state
    assert torch.equal(pd, ps) and torch.equal(md, ms) and torch.equal(vd, vs) and torch.equal(td, ts)
    assert not torch.equal(pd, p0)


@pytest.mark.parametrize('push', [False, True, 'flags'])
def test_peer_optimiser_on_one_gpu_matches_dense_adam(push):
    """mne_grad_dirty_map (owners fetch the flagged groups) or mne_grad_push (non-zero texels pushed into the owners' inboxes)
    + mne_adam_step_peers with two emulated ranks on ONE device (both "peers'" buffers live on the same GPU, the two ranks'
    kernels run one after the other): block-cyclic ownership, pushed dirty flags, sparse gradient reduction, parameter
    write-back into both replicas, exact skip of never-touched groups."""
    import ctypes as C
    from mneslam_b200 import _lib, ops
    lib, st = _lib.lib(), ops._stream()
    torch.manual_seed(1)
    world, chunk = 2, 64
    n_used, n_total = 5 * 128 - 32, 5 * 128 + 128           # padding: one and a bit chunk pairs
    seg_ends = (C.c_int64 * 2)(256, n_used)
    lr, eps, wd = (C.c_float * 2)(5e-3, 1e-2), (C.c_float * 2)(1e-15, 1e-8), (C.c_float * 2)(0.0, 1e-6)
    p0 = torch.randn(n_total, device='cuda'); p0[n_used:] = 0
    params = [p0.clone(), p0.clone()]
    grads = [torch.zeros(n_total, device='cuda') for _ in range(world)]
    dirty = [torch.zeros(n_total // 32, dtype=torch.uint8, device='cuda') for _ in range(world)]
    local = n_total // world
    m = [torch.zeros(local, device='cuda') for _ in range(world)]
    v = [torch.zeros(local, device='cuda') for _ in range(world)]
    touched = [torch.zeros(local // 4, dtype=torch.uint8, device='cuda') for _ in range(world)]
    pp = (C.c_void_p * world)(*[x.data_ptr() for x in params])
    gp = (C.c_void_p * world)(*[x.data_ptr() for x in grads])
    dp = (C.c_void_p * world)(*[x.data_ptr() for x in dirty])
    inbox = [torch.full((n_total,), float('nan'), device='cuda') for _ in range(world)]     # stale contents must never be read
    ip = (C.c_void_p * world)(*[x.data_ptr() for x in inbox])
    ref = torch.nn.Parameter(p0.clone())
    opts = [torch.optim.Adam([ref], lr=5e-3, betas=(0.9, 0.99), eps=1e-15, foreach=False),
            torch.optim.Adam([ref], lr=1e-2, betas=(0.9, 0.99), eps=1e-8, weight_decay=1e-6, foreach=False)]
    ever = torch.zeros(n_used, dtype=torch.bool, device='cuda')
    for t in range(1, 5):
        total = torch.zeros(n_total, device='cuda')
        for r in range(world):
            mask = torch.rand(n_total // 32, device='cuda') < 0.3            # sparse, texel-granular like a map step
            g = torch.randn(n_total, device='cuda') * mask.repeat_interleave(32)
            g[::3] = 0; g[n_used:] = 0
            grads[r].copy_(g); total += g
        ever |= (total[:n_used] != 0)
        # reference: the two parameter groups with their own hyper-parameters; plane groups never touched stay put
        full = total.clone()
        for o, (a, b) in zip(opts, ((0, 256), (256, n_used))):
            gsel = torch.zeros_like(full); gsel[a:b] = full[a:b]
            before = ref.detach().clone()
            ref.grad = gsel
            o.step()
            with torch.no_grad():
                keep = torch.ones(n_total, dtype=torch.bool, device='cuda'); keep[a:b] = False
                ref[keep] = before[keep]
        for r in range(world):
            if push == 'flags':       # the flags come from the backward pass (here: from the gradient itself, plus a few spurious ones)
                fl = (grads[r].view(-1, 32) != 0).any(1).to(torch.uint8)
                fl[::7] = 1
                _lib.check(lib.mne_dirty_push_data(fl.data_ptr(), grads[r].data_ptr(), n_total, n_total, chunk, world, r, dp, ip, st), 'push_data')
                torch.cuda.synchronize()
                assert int(fl.sum()) == 0, 'the local flag map is cleared on the way'
            elif push:
                _lib.check(lib.mne_grad_push(grads[r].data_ptr(), n_total, n_total, chunk, world, r, dp, ip, st), 'push')
            else:
                _lib.check(lib.mne_grad_dirty_map(grads[r].data_ptr(), n_total, chunk, world, r, dp, st), 'dirty')
        for r in range(world):
            _lib.check(lib.mne_adam_step_peers(world, r, pp, gp, n_total, chunk, m[r].data_ptr(), v[r].data_ptr(), 2, seg_ends,
                                               lr, eps, wd, 0.9, 0.99, None, t, touched[r].data_ptr(), None, None,
                                               dirty[r].data_ptr(), inbox[r].data_ptr() if push else None, st), 'peers')
        torch.cuda.synchronize()
        assert all(float(g.abs().max()) == 0.0 for g in grads)               # every gradient was consumed and cleared
        assert torch.equal(params[0], params[1])                             # both replicas hold the same parameters
    # the reference above ran Adam on never-touched plane groups too (zero gradient, zero moments: identity), so the
    # results must agree everywhere to fp32 rounding
    assert rel(params[0].cpu().numpy(), ref.detach().cpu().numpy()) < 5e-6
    assert float(params[0][n_used:].abs().max()) == 0.0
    # touched flags: set exactly where a gradient ever arrived (plane segment), in ownership order
    gi = torch.arange(n_total // 4, device='cuda')
    owner, li = (gi // (chunk // 4)) % world, (gi // (chunk // 4)) // world * (chunk // 4) + gi % (chunk // 4)
    ever4 = torch.zeros(n_total // 4, dtype=torch.bool, device='cuda')
    ever4[:n_used // 4] = ever.view(-1, 4).any(1)
    for r in range(world):
        sel = (owner == r) & (gi < 256 // 4)
        assert torch.equal(touched[r][li[sel]].bool(), ever4[sel])


@pytest.mark.parametrize('use_hash', [False, True])
def test_backward_flags_exactly_the_texels_it_adds_into(use_hash):
    """MneSceneDesc.dirty: after a backward pass every 32-float texel of the plane / table gradient that holds a
    non-zero value is flagged, (almost) nothing else is, and mne_dirty_push hands the flags over in the owner layout
    mne_grad_dirty_map derives from the gradient buffer itself."""
    import ctypes as C
    from mneslam_b200 import JointEncoding, _lib, ops
    from mneslam_b200.mapping import FusedMapper
    cfg = configs.small_test_config()
    if use_hash:
        cfg['grid']['use_hash'] = True
        cfg['grid']['oneGrid'] = True
    torch.manual_seed(0)
    model = JointEncoding(cfg, torch.tensor(cfg['mapping']['bound'], dtype=torch.float64)).cuda()
    mapper = FusedMapper(model)
    n = mapper.flat_grad.numel()
    mapper.dirty_local = torch.zeros((n + 31) // 32, dtype=torch.uint8, device='cuda')
    b = synthetic.make_ray_batch(cfg, 256, seed=4)
    mapper.forward_backward(b['rays_o'].cuda(), b['rays_d'].cuda(), b['target_rgb'].cuda(), b['target_d'].cuda())
    torch.cuda.synchronize()
    g = mapper.flat_grad
    n_tab = mapper.dense_from
    full = n_tab // 32
    nz = (g[:full * 32].view(-1, 32) != 0).any(1)
    fl = mapper.dirty_local[:full].bool()
    assert int(nz.sum()) > 0
    assert bool((fl | ~nz).all())                                   # non-zero => flagged
    # flagged => (almost always) non-zero; the hash scatter also flags corners it adds an exact zero to (points whose
    # sdf gradient vanished), which costs a little NVLink traffic and nothing else
    slack = 1.2 if use_hash else 1.02
    assert int(fl.sum()) <= int(nz.sum()) * slack + 2
    assert int(mapper.dirty_local[(n_tab + 31) // 32:].sum()) == 0  # the decoder weights are not flagged by the kernels
    if n % 128 == 0:
        # push (flags from the kernels) and scan (flags from the buffer) produce the same owner-side arrays where the
        # gradient is non-zero; two emulated ranks on one device
        world, chunk = 2, 64
        lib, st = _lib.lib(), ops._stream()
        a = [torch.zeros(n // 32, dtype=torch.uint8, device='cuda') for _ in range(world)]
        b = [torch.zeros(n // 32, dtype=torch.uint8, device='cuda') for _ in range(world)]
        pa, pb = (C.c_void_p * world)(*[x.data_ptr() for x in a]), (C.c_void_p * world)(*[x.data_ptr() for x in b])
        _lib.check(lib.mne_grad_dirty_map(g.data_ptr(), n, chunk, world, 1, pa, st), 'scan')
        _lib.check(lib.mne_dirty_push(mapper.dirty_local.data_ptr(), n, n_tab, chunk, world, 1, pb, st), 'push')
        torch.cuda.synchronize()
        for x, y in zip(a, b):
            assert bool((y.bool() | ~x.bool()).all())
            assert int(y.sum()) <= int(x.sum()) * slack + 2 + (n - n_tab + 31) // 32
        assert int(mapper.dirty_local.sum()) == 0                   # cleared for the next iteration


def test_host_fed_graph_iteration_matches_the_eager_driver():
    """HostFedMapper (host batch -> pinned staging -> one graph launch with the copies inside -> loss as a float) against
    FusedMapper.step on the same batches copied to the device by hand: same losses, same parameters."""
    from mneslam_b200.mapping import DeviceKeyframeRays, FusedMapper, HostFedMapper
    cfg, bb, planes, weights = _small_model(seed=9)
    cfg['cam'].update({'H': 40, 'W': 60, 'fx': 50.0, 'fy': 50.0, 'cx': 29.5, 'cy': 19.5})
    cfg['training']['perturb'] = 0                      # no device-side random numbers: the two arms see identical samples
    rays, poses = synthetic.make_keyframe_db(cfg, n_kf=3, seed=2)
    flat = rays.reshape(-1, 7)
    gen = torch.Generator().manual_seed(5)
    batches = []
    for _ in range(6):
        idx = torch.randint(0, flat.shape[0], (256,), generator=gen)
        batches.append((flat[idx].contiguous(), (idx // rays.shape[1]).contiguous()))
    arms = []
    for hosted in (True, False):
        model = build_model(cfg, bb, planes, weights, as_params=False)
        mapper = FusedMapper(model)
        db = DeviceKeyframeRays(rays.cuda(), poses.cuda())
        losses = []
        if hosted:
            fed = HostFedMapper(mapper, db, 256)
            # capture() warms up with two real iterations and the captured one does not run: replay them on the other arm
            fed.capture(*batches[0])
            for r7, ids in batches:
                losses.append(fed.step(r7, ids))
        else:
            for _ in range(2):
                mapper.step(*db.assemble(batches[0][0].cuda(), batches[0][1].cuda()))
            for r7, ids in batches:
                losses.append(float(mapper.total_loss(mapper.step(*db.assemble(r7.cuda(), ids.cuda())))))
        torch.cuda.synchronize()
        arms.append((losses, [p.detach().clone() for p in flat_planes(model)] + [w.detach().clone() for w in model.decoder.weight_list()]))
    (la, pa), (lb, pb) = arms
    assert la == pytest.approx(lb, rel=1e-4)
    for x, y in zip(pa, pb):
        assert float((x.double() - y.double()).norm() / (y.double().norm() + 1e-30)) < 5e-2     # see the trajectory test above
    assert min(float((x.double() - y.double()).norm() / (y.double().norm() + 1e-30)) for x, y in zip(pa, pb)) < 2e-5


def test_pipelined_iterations_do_the_same_work():
    """PipelinedMapper draws and samples batch t+1 on a second stream while the optimiser of iteration t runs.  Fed the same
    batches it must give the same losses and parameters as FusedMapper.step; captured as a graph it must keep training."""
    from mneslam_b200.mapping import DeviceKeyframeRays, FusedMapper, PipelinedMapper
    cfg, bb, planes, weights = _small_model(seed=9)
    cfg['cam'].update({'H': 40, 'W': 60, 'fx': 50.0, 'fy': 50.0, 'cx': 29.5, 'cy': 19.5})
    cfg['training']['perturb'] = 0
    rays, poses = synthetic.make_keyframe_db(cfg, n_kf=3, seed=2)

    class FixedDb(DeviceKeyframeRays):
        """Hands out a fixed sequence of batches."""
        def __init__(self, rays, poses, n, count):
            super().__init__(rays, poses)
            g = torch.Generator().manual_seed(11)
            self.queue = [torch.randint(0, self.n_kf * self.per_kf, (n,), generator=g).cuda() for _ in range(count)]
            self.at = 0

        def sample(self, n_rays, generator=None):
            idx = self.queue[self.at]
            self.at += 1
            return self.assemble(self.rays.reshape(-1, 7)[idx], torch.div(idx, self.per_kf, rounding_mode='trunc'))

    results = []
    for pipelined in (True, False):
        model = build_model(cfg, bb, planes, weights, as_params=False)
        mapper = FusedMapper(model)
        db = FixedDb(rays.cuda(), poses.cuda(), 256, 8)
        losses = []
        if pipelined:
            pipe = PipelinedMapper(mapper, db, 256)
            pipe._draw()
            for _ in range(6):
                pipe._iteration()                     # eager: the same stream choreography as the captured graph
                torch.cuda.synchronize()
                losses.append(float(mapper.total_loss(pipe.losses)))
            pipe.close()
        else:
            for _ in range(6):
                losses.append(float(mapper.total_loss(mapper.step(*db.sample(256)))))
        torch.cuda.synchronize()
        results.append((losses, [p.detach().clone() for p in flat_planes(model)]))
    (la, pa), (lb, pb) = results
    assert la == pytest.approx(lb, rel=1e-4)
    assert min(float((x.double() - y.double()).norm() / (y.double().norm() + 1e-30)) for x, y in zip(pa, pb)) < 2e-5

    # graph mode on a random database: the loss keeps going down
    model = build_model(cfg, bb, planes, weights, as_params=False)
    mapper = FusedMapper(model)
    cur = rays[-1].cuda()                                  # the current frame: here, the rays of the last keyframe
    pipe = PipelinedMapper(mapper, DeviceKeyframeRays(rays.cuda(), poses.cuda(), current=cur, min_pixels_cur=20), 256, spare_sms=6).capture()
    assert pipe.ro.shape[0] == 256 + 85
    hist = []
    for i in range(60):
        pipe.step()
        if i % 10 == 9:
            torch.cuda.synchronize()
            hist.append(float(mapper.total_loss(pipe.losses)))
    assert all(np.isfinite(hist)) and hist[-1] < hist[0], hist
    # the library-wide option the pipeline sets is put back when it is closed
    import ctypes
    val = ctypes.c_int(-1)
    _lib_mod = __import__('mneslam_b200._lib', fromlist=['lib'])
    _lib_mod.lib().mne_get_option(b'adam_spare_sms', ctypes.addressof(val))
    assert val.value == 6
    pipe.close()
    _lib_mod.lib().mne_get_option(b'adam_spare_sms', ctypes.addressof(val))
    assert val.value == 0


def test_distillation_step_pulls_the_student_towards_the_teacher():
    """FusedMapper.distill_step (mp_slam/mapper.py:598-642): the student's rendering of the shared rays approaches the
    teacher's."""
    from mneslam_b200.mapping import FusedMapper
    cfg, bb, planes, weights = _small_model(seed=9)
    teacher = build_model(cfg, bb, planes, weights, as_params=False)
    cfg2, _, planes2, weights2 = _small_model(seed=21)
    student = build_model(cfg, bb, planes2, weights2, as_params=False)
    mapper = FusedMapper(student)
    b = {k: v.cuda() for k, v in synthetic.make_ray_batch(cfg, 400, seed=3).items()}

    def gap():
        with torch.no_grad():
            a = teacher.render_rays(b['rays_o'], b['rays_d'], target_d=None)
            c = student.render_rays(b['rays_o'], b['rays_d'], target_d=None)
        return float((a['depth'] - c['depth']).abs().mean()), float((a['rgb'] - c['rgb']).abs().mean())
    d0, c0 = gap()
    for _ in range(80):
        losses = mapper.distill_step(teacher, b['rays_o'], b['rays_d'])
    torch.cuda.synchronize()
    d1, c1 = gap()
    assert torch.isfinite(losses[:2]).all()
    assert d1 < 0.7 * d0 or c1 < 0.7 * c0, (d0, d1, c0, c1)


def test_optim_adam_is_a_drop_in_for_torch_adam():
    """mneslam_b200.optim.Adam in place of torch.optim.Adam in create_optimizer's two parameter groups (decoder with
    weight decay, channels-last planes with eps 1e-15): same parameters after five steps, zero_grad / state_dict work."""
    from mneslam_b200.optim import Adam
    torch.manual_seed(3)
    shapes = [((32, 112), False), ((1, 32, 20, 30), True), ((3, 32), False), ((1, 32, 7, 9), True)]

    def make():
        torch.manual_seed(3)
        ps = []
        for shp, cl in shapes:
            t = torch.randn(*shp, device='cuda')
            if cl:
                t = t.contiguous(memory_format=torch.channels_last)
            ps.append(torch.nn.Parameter(t))
        return ps
    pa, pb = make(), make()
    groups = lambda ps: [{'params': [ps[0], ps[2]], 'weight_decay': 1e-6, 'lr': 1e-2}, {'params': [ps[1], ps[3]], 'eps': 1e-15, 'lr': 5e-3}]
    oa = torch.optim.Adam(groups(pa), betas=(0.9, 0.99), foreach=False)
    ob = Adam(groups(pb), betas=(0.9, 0.99))
    for t in range(5):
        torch.manual_seed(100 + t)
        for x, y in zip(pa, pb):
            g = torch.randn_like(x) * (0.1 if t % 2 else 5.0)
            x.grad, y.grad = g.clone(), g.clone()
        oa.step(); ob.step()
        oa.zero_grad(); ob.zero_grad()
    torch.cuda.synchronize()
    for x, y in zip(pa, pb):
        assert y.stride() == x.stride()
        assert rel(y.detach().cpu().numpy(), x.detach().cpu().numpy()) < 5e-6
    sd = ob.state_dict()
    assert len(sd['state']) == 4 and sd['state'][0]['step'] == 5


@pytest.mark.parametrize('async_loss', [False, True])
def test_prefetching_host_mapper_trains_on_the_previous_batch(async_loss):
    """PrefetchingHostMapper.step(next) = FusedMapper.step on the batch handed over one call earlier: same losses, same
    parameters (prime() spends two real iterations per captured graph on the first batch, as documented).  ``async_loss``:
    the same iterations, each loss handed back one call later (two graphs alternate; the last loss comes from flush())."""
    from mneslam_b200.mapping import DeviceKeyframeRays, FusedMapper, PrefetchingHostMapper
    cfg, bb, planes, weights = _small_model(seed=9)
    cfg['cam'].update({'H': 40, 'W': 60, 'fx': 50.0, 'fy': 50.0, 'cx': 29.5, 'cy': 19.5})
    cfg['training']['perturb'] = 0
    rays, poses = synthetic.make_keyframe_db(cfg, n_kf=3, seed=2)
    flat = rays.reshape(-1, 7)
    gen = torch.Generator().manual_seed(5)
    batches = []
    for _ in range(7):
        idx = torch.randint(0, flat.shape[0], (256,), generator=gen)
        batches.append((flat[idx].contiguous(), (idx // rays.shape[1]).contiguous()))
    arms = []
    for prefetch in (True, False):
        model = build_model(cfg, bb, planes, weights, as_params=False)
        mapper = FusedMapper(model)
        db = DeviceKeyframeRays(rays.cuda(), poses.cuda())
        losses = []
        if prefetch:
            fed = PrefetchingHostMapper(mapper, db, 256, async_loss=async_loss).prime(*batches[0])
            for nxt in batches[1:]:
                losses.append(fed.step(*nxt))                       # the loss of the batch given one call earlier
            if async_loss:
                assert losses[0] is None                            # nothing had been launched before the first call
                losses = losses[1:] + [fed.flush()]
                assert fed.flush() is None
            fed.close()
        else:
            for _ in range(4 if async_loss else 2):
                mapper.step(*db.assemble(batches[0][0].cuda(), batches[0][1].cuda()))
            for r7, ids in batches[:-1]:
                losses.append(float(mapper.total_loss(mapper.step(*db.assemble(r7.cuda(), ids.cuda())))))
        torch.cuda.synchronize()
        arms.append((losses, [p.detach().clone() for p in flat_planes(model)]))
    (la, pa), (lb, pb) = arms
    assert la == pytest.approx(lb, rel=1e-4)
    assert min(float((x.double() - y.double()).norm() / (y.double().norm() + 1e-30)) for x, y in zip(pa, pb)) < 2e-5


def test_dropin_with_fused_adam_trains_like_torch_adam_and_survives_plane_replacement():
    """The reference's call sites with one line changed (mneslam_b200.optim.Adam in create_optimizer, mneslam_mp.py:469):
    planes wrapped in nn.Parameter inside all_planes (mneslam_mp.py:451-457), forward / loss.backward() / step() / zero_grad().
    The gradients are views of the model's persistent arena and the planes lie back to back, so the optimiser is two
    launches; the parameters must follow torch.optim.Adam's.  Half way, all_planes is replaced wholesale (what
    load_foreign_model does, mp_slam/mapper.py:719) and new optimisers are built: both keep training, identically."""
    from mneslam_b200 import JointEncoding
    from mneslam_b200.optim import Adam
    from mneslam_b200.scene_rep import GradArena
    cfg = configs.small_test_config(one_grid=True)
    bb = torch.tensor(cfg['mapping']['bound'], dtype=torch.float64)
    mp = cfg['mapping']

    def make():
        torch.manual_seed(11)
        m = JointEncoding(cfg, bb).cuda()
        m.train()
        for grp in m.all_planes:
            for i, p in enumerate(grp):
                grp[i] = torch.nn.Parameter(p)
        return m

    def optimiser(m, cls, **kw):
        return cls([{'params': m.decoder.parameters(), 'weight_decay': 1e-6, 'lr': mp['lr_decoder']},
                    {'params': flat_planes(m), 'eps': 1e-15, 'lr': mp['lr_embed']}], betas=(0.9, 0.99), **kw)
    a, b = make(), make()
    oa, ob = optimiser(a, torch.optim.Adam, foreach=False), optimiser(b, Adam)
    n, s = 150, synthetic.n_samples_total(cfg)

    def iteration(it):
        batch = {k: v.cuda() for k, v in synthetic.make_ray_batch(cfg, n, seed=70 + it).items()}
        noise = synthetic.make_noise(n, s, seed=it).cuda()
        out = []
        for m, o in ((a, oa), (b, ob)):
            ret = m.forward(batch['rays_o'], batch['rays_d'], batch['target_rgb'], batch['target_d'], noise=noise)
            loss = total_loss(cfg, ret)
            loss.backward()
            o.step()
            o.zero_grad()
            out.append(loss.item())
        assert out[1] == pytest.approx(out[0], rel=2e-4)
    for it in range(3):
        iteration(it)
    arena = b._grad_arena
    assert arena is not None and arena.clean and float(arena.flat.abs().max()) == 0.0, 'the update kernel cleared the gradients'
    assert GradArena.owning(arena.flat.data_ptr() + 64) is arena
    fp = flat_planes(b)
    assert all(fp[i + 1].data_ptr() == fp[i].data_ptr() + 4 * fp[i].numel() for i in range(5)), 'planes lie back to back'
    # a peer's map arrives: every plane list is replaced by fresh (NCHW-contiguous, as a reference checkpoint holds them) tensors
    torch.manual_seed(5)
    fresh = [[torch.randn_like(p).contiguous() * 0.3 for p in grp] for grp in a.all_planes]
    for m in (a, b):
        m.all_planes = tuple([torch.nn.Parameter(t.clone()) for t in grp] for grp in fresh)
    oa, ob = optimiser(a, torch.optim.Adam, foreach=False), optimiser(b, Adam)
    for it in range(3, 6):
        iteration(it)
    for x, y in zip(flat_planes(a) + a.decoder.weight_list(), flat_planes(b) + b.decoder.weight_list()):
        assert rel(y.detach().cpu().numpy(), x.detach().cpu().numpy()) < 2e-3


def test_distillation_step_matches_the_oracle():
    """One iteration of the joint distillation (mp_slam/mapper.py:594-644) against the CPU oracle: the teacher renders the
    shared rays without depth guidance (render_rays(target_d=None)), the student takes a mapping iteration on that
    rendering -- losses within 1e-3, the student's parameter update direction as the oracle's (autograd + torch Adam).
    training.perturb = 0 so that both sides sample the same positions."""
    from mneslam_b200.mapping import FusedMapper
    cfg, bb, planes_t, weights_t = _small_model(seed=9)
    cfg['training']['perturb'] = 0
    _, _, planes_s, weights_s = _small_model(seed=21)
    teacher = build_model(cfg, bb, planes_t, weights_t, as_params=False)
    student = build_model(cfg, bb, planes_s, weights_s, as_params=False)
    before = [p.detach().clone().cpu() for p in flat_planes(student) + student.decoder.weight_list()]
    mapper = FusedMapper(student)
    b = synthetic.make_ray_batch(cfg, 300, seed=3)
    losses = mapper.distill_step(teacher, b['rays_o'].cuda(), b['rays_d'].cuda())
    torch.cuda.synchronize()
    # ---- the same on the CPU
    bound = J.enlarged_bound(cfg)
    with torch.no_grad():
        t = J.render_rays(cfg, planes_t, weights_t, bound, bb, b['rays_o'], b['rays_d'], None, None)
    op = tuple([torch.nn.Parameter(p.clone()) for p in grp] for grp in planes_s)
    ow = {k: torch.nn.Parameter(v.clone()) for k, v in weights_s.items()}
    ret, _ = J.forward(cfg, op, ow, bound, bb, b['rays_o'], b['rays_d'], t['rgb'], t['depth'].unsqueeze(-1), None)
    mp = cfg['mapping']
    opt = torch.optim.Adam([{'params': list(ow.values()), 'weight_decay': 1e-6, 'lr': mp['lr_decoder']},
                            {'params': [p for g in op for p in g], 'eps': 1e-15, 'lr': mp['lr_embed']}], betas=(0.9, 0.99))
    J.total_loss(cfg, ret).backward()
    opt.step()
    for i, k in enumerate(LOSS_KEYS[:7]):
        ref = ret[k].item()
        if np.isfinite(ref):
            assert losses[i].item() == pytest.approx(ref, rel=1e-3, abs=1e-7), k
    after = [p.detach().cpu() for p in flat_planes(student) + student.decoder.weight_list()]
    ref_after = [p.detach() for g in op for p in g] + [ow[k].detach() for k in ('sdf0', 'sdf1', 'col0', 'col1')]
    for a0, a1, r1 in zip(before, after, ref_after):
        da, dr = (a1 - a0).double().reshape(-1), (r1 - a0).double().reshape(-1)
        if float(dr.norm()) > 0:
            assert cosine(da.numpy(), dr.numpy()) >= 0.99, 'update direction'       # Adam's first step is lr * sign(g): sign flips of ~0 gradients only
