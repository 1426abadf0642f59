"""The allocation-free backward entry point (``mne_render_backward_static``) against the regular one.

Some users of the library cannot call ``cudaMalloc`` where the backward runs (the host application forbids
allocation around the call, or only hands over input and output buffers), so one variant of the backward keeps
its 20-byte-per-sample workspace in an array built into the library instead of taking it from the caller.
This file checks, through the C ABI, that the variant gives the same gradients as ``mne_render_backward`` on
two calls in a row (different batches and sizes, so leftovers of the first call would show), and that batches
larger than the built-in workspace are refused with a message.  It is named so that pytest collects it last.
"""
import ctypes as C

import pytest
import torch

from mneslam_b200 import _lib, configs, ops, synthetic
from oracle import joint_encoding as J
from tests.gpu_util import build_model, flat_planes

pytestmark = pytest.mark.gpu


def _forward(model, cfg, batch, noise):
    """sample + forward through the ABI; returns everything a backward call needs."""
    dev = batch['rays_o'].device
    n = batch['rays_o'].shape[0]
    tables = ops.SampleTables.get(cfg, dev, True)
    s = tables[0].numel() + tables[1].numel()
    rcfg = ops.make_render_cfg(cfg, norm_rays=n)
    scfg = ops.make_sample_cfg(tables, True)
    td = batch['target_d'].reshape(-1).contiguous()
    acc = ops.new_acc(dev)
    z = ops.sample_rays(scfg, rcfg, n, s, td, noise, acc, dev)
    scene = ops.make_scene(flat_planes(model), model.decoder.weight_list(), model.bound, model.bounding_box)
    out = ops.render_forward(scene, rcfg, batch['rays_o'], batch['rays_d'], z, batch['target_rgb'], td, acc)
    return rcfg, td, acc, z, out['raw'], n, s


def _backward(model, cfg, batch, fwd, static):
    rcfg, td, acc, z, raw, n, s = fwd
    dev = z.device
    planes, weights = flat_planes(model), model.decoder.weight_list()
    pg = [torch.zeros_like(p) for p in planes]
    wg = [torch.zeros_like(w) for w in weights]
    scene = ops.make_scene(planes, weights, model.bound, model.bounding_box, plane_grads=pg, weight_grads=wg)
    lg = torch.tensor([5.0, 0.1, 0.0, 0.0, 5.0, 200.0, 30.0, 0.0], device=dev)
    g_o, g_d = torch.empty(n, 3, device=dev), torch.empty(n, 3, device=dev)
    common = [C.byref(scene), C.byref(rcfg), n, s, batch['rays_o'].data_ptr(), batch['rays_d'].data_ptr(), z.data_ptr(),
              raw.data_ptr(), batch['target_rgb'].data_ptr(), td.data_ptr(), acc.data_ptr(), lg.data_ptr(), None, None, None]
    stream = torch.cuda.current_stream().cuda_stream
    if static:
        st = _lib.lib().mne_render_backward_static(*common, g_o.data_ptr(), g_d.data_ptr(), stream)
    else:
        scratch = ops.backward_scratch(n, s, dev)
        st = _lib.lib().mne_render_backward(*common, scratch.data_ptr(), g_o.data_ptr(), g_d.data_ptr(), None, stream)
    _lib.check(st, 'backward')
    torch.cuda.synchronize()
    return pg + wg + [g_o, g_d]


def _close(a, b):
    # same kernel, same inputs: only the order of the floating-point atomics differs
    return torch.allclose(a, b, rtol=1e-3, atol=1e-9 + 1e-4 * float(b.abs().max()))


def test_static_scratch_backward_matches_regular_backward_twice():
    cfg = configs.small_test_config(n_samples_d=13, n_range_d=7)
    bound = J.enlarged_bound(cfg)
    bb = torch.tensor(cfg['mapping']['bound'], dtype=torch.float64)
    model = build_model(cfg, bb, J.init_planes(cfg, bound, seed=31, std=0.5), J.init_weights(cfg, seed=31), as_params=False)
    s = synthetic.n_samples_total(cfg)
    for call, n_rays in enumerate((257, 96)):            # second call is smaller: stale entries stay in the array
        batch = {k: v.cuda().contiguous() for k, v in synthetic.make_ray_batch(cfg, n_rays, seed=70 + call).items()}
        noise = synthetic.make_noise(n_rays, s, seed=call).cuda()
        fwd = _forward(model, cfg, batch, noise)
        want = _backward(model, cfg, batch, fwd, static=False)
        got = _backward(model, cfg, batch, fwd, static=True)
        assert any(float(w.abs().max()) > 0 for w in want)
        for g, w in zip(got, want):
            assert _close(g, w), (call, float((g - w).abs().max()), float(w.abs().max()))


def test_static_scratch_rejects_batches_beyond_its_maximum():
    cfg = configs.small_test_config()
    rcfg = ops.make_render_cfg(cfg, norm_rays=1)
    scene = _lib.SceneDesc()
    too_many = _lib.STATIC_SCRATCH_SAMPLES // 64 + 1
    st = _lib.lib().mne_render_backward_static(C.byref(scene), C.byref(rcfg), too_many, 64, *([None] * 14))
    assert st != 0
    msg = _lib.lib().mne_last_error().decode()
    assert 'exceed' in msg and str(_lib.STATIC_SCRATCH_SAMPLES) in msg
