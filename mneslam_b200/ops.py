"""Thin wrappers: torch tensors in, C-ABI calls out.  PyTorch is used for device memory and streams only;
every computation below happens inside ``libmneslam_b200.so``."""
import ctypes as C

import torch

from . import _lib
from ._lib import ACC_SLOTS, N_LOSSES, RenderCfg, SampleCfg, SceneDesc


def _ptr(t):
    return None if t is None else t.data_ptr()


_raw_stream = getattr(torch._C, '_cuda_getCurrentRawStream', None)


def _stream():
    """The raw handle of torch's current stream on the current device.  (``torch.cuda.current_stream()`` builds a Python
    Stream object on every call: ~25 us, several times per iteration in the eager paths.)"""
    if _raw_stream is not None:
        return _raw_stream(torch.cuda.current_device())
    return torch.cuda.current_stream().cuda_stream


def _f32(t, name):
    if t is None:
        return None
    if not (t.is_cuda and t.dtype == torch.float32):
        raise TypeError('%s must be a float32 CUDA tensor, got %s on %s' % (name, t.dtype, t.device))
    return t if t.is_contiguous() else t.contiguous()


def require_cuda():
    if not torch.cuda.is_available():
        raise RuntimeError('mneslam_b200 needs a CUDA device (there is no CPU fallback)')


# --------------------------------------------------------------------------------------------------
# descriptors
# --------------------------------------------------------------------------------------------------
def make_render_cfg(config, norm_rays, first_pass=False):
    """Float32 roundings of the Python constants, as the reference's tensor-scalar ops round them."""
    t = config['training']
    rc = RenderCfg()
    rc.trunc = t['trunc']
    rc.sc_trunc = config['data']['sc_factor'] * t['trunc']
    rc.model_trunc = config['model']['truncation']
    rc.center_trunc = 0.4 * config['model']['truncation']
    rc.depth_trunc = config['cam']['depth_trunc']
    rc.white_bkgd = 1 if t['white_bkgd'] else 0
    rc.first_pass = 1 if first_pass else 0
    rc.norm_rays = float(norm_rays)
    return rc


class SampleTables:
    """The reference rebuilds its linspace tables on the CPU every call (``model/scene_rep.py:363-373``); here
    they are built once per (config values, device) with the same torch.linspace calls and kept on the device."""
    _cache = {}

    @classmethod
    def get(cls, config, device, with_depth):
        t, cam = config['training'], config['cam']
        if with_depth:
            key = (str(device), True, cam['near'], cam['far'], t['n_samples_d'], t['n_range_d'], t['range_d'])
        else:
            key = (str(device), False, cam['near'], cam['far'], t['n_samples'])
        tab = cls._cache.get(key)
        if tab is None:
            if with_depth:
                uni = torch.linspace(cam['near'], cam['far'], t['n_samples_d']) if t['n_samples_d'] > 0 \
                    else torch.empty(0)
                off = torch.linspace(-t['range_d'], t['range_d'], steps=t['n_range_d'])
                fb = torch.linspace(cam['near'], cam['far'], steps=t['n_range_d'])
            else:
                uni = torch.linspace(cam['near'], cam['far'], t['n_samples'])
                off = fb = torch.empty(0)
            tab = tuple(x.float().to(device) for x in (uni, off, fb))
            cls._cache[key] = tab
        return tab


def make_sample_cfg(tables, perturb):
    uni, off, fb = tables
    sc = SampleCfg()
    sc.n_uniform = uni.numel()
    sc.n_range = off.numel()
    sc.uniform = _ptr(uni) if uni.numel() else None
    sc.range_offsets = _ptr(off) if off.numel() else None
    sc.range_fallback = _ptr(fb) if fb.numel() else None
    sc.perturb = 1 if perturb else 0
    return sc


_bounds_cache = {}


def bounds_to_host(bound, bounding_box):
    """Host floats of the two boxes.  ``bounding_box`` is a float64 CUDA tensor in the reference
    (mneslam_mp.py:223); reading it costs a device sync, so the result is cached per (tensor, version)."""
    key = (id(bound), bound._version, id(bounding_box), bounding_box._version)
    hit = _bounds_cache.get(key)
    if hit is None:
        b = bound.detach().to('cpu', torch.float32)
        ext = b[:, 1] - b[:, 0]                       # float32 subtraction, as model/utils.py:38-40
        bb = bounding_box.detach().to('cpu', torch.float64)
        bext = bb[:, 1] - bb[:, 0]
        hit = ([float(v) for v in b[:, 0]], [float(v) for v in ext], [float(v) for v in bb[:, 0]],
               [float(v) for v in bext], bound, bounding_box)      # keep the tensors alive: ids stay unique
        if len(_bounds_cache) > 64:
            _bounds_cache.clear()
        _bounds_cache[key] = hit
    return hit[:4]


def make_scene(planes, weights, bound, bounding_box, n_bins=16, plane_grads=None, weight_grads=None, hash_meta=None,
               dirty=None, dirty_base=None, hash_half=None):
    """planes: flat list of 6 or 12 ``[1,32,H,W]`` tensors in ``all_planes`` order (group, axis, level) -- or, with
    ``hash_meta`` (a ``hashgrid.HashGridMeta``; its ``.color`` is the colour table's meta or None), ``[table]`` /
    ``[table, colour table]`` of the hash grid encoder.
    weights: [sdf0, sdf1, col0, col1].  bound: float32 [3,2] (CPU ok).  bounding_box: float64 [3,2].
    dirty / dirty_base: uint8 flag map and the flat gradient buffer it indexes (``MneSceneDesc.dirty``)."""
    sc = SceneDesc()
    sc.n_bins = n_bins
    keep = []
    if hash_meta is not None:
        color_meta = getattr(hash_meta, 'color', None)        # second table (grid.oneGrid: False)
        if len(planes) != (1 if color_meta is None else 2):
            raise ValueError('the hash grid encoder takes [table] or, with a colour table, [table, colour table]')
        sc.encoder, sc.two_grid, sc.c_dim = 1, 0 if color_meta is None else 1, 32
        # hash_half: fp16 copies of the tables ([table] / [table, colour table]); the kernels then read the features from them
        hash_meta.fill_desc(sc.hash[0], planes[0], plane_grads[0] if plane_grads is not None else None,
                            hash_half[0] if hash_half is not None else None)
        if color_meta is not None:
            color_meta.fill_desc(sc.hash[1], planes[1], plane_grads[1] if plane_grads is not None else None,
                                 hash_half[1] if hash_half is not None else None)
        if hash_half is not None:
            keep.extend(hash_half)
        planes, plane_grads = [], None
        feat = hash_meta.n_output
    else:
        if len(planes) not in (6, 12):
            raise ValueError('expected 6 or 12 planes, got %d' % len(planes))
        sc.encoder = 0
        sc.two_grid = 1 if len(planes) == 12 else 0
        sc.c_dim = planes[0].shape[1]
        feat = 64
    for i, p in enumerate(planes):
        if not (p.is_cuda and p.dtype == torch.float32 and p.dim() == 4 and p.shape[0] == 1):
            raise TypeError('plane %d must be a float32 CUDA tensor of shape [1,C,H,W]' % i)
        d = sc.planes[i]
        d.data = p.data_ptr()
        d.H, d.W = p.shape[2], p.shape[3]
        d.stride_c, d.stride_h, d.stride_w = p.stride(1), p.stride(2), p.stride(3)
        g = plane_grads[i] if plane_grads is not None else None
        if g is not None:
            if g.shape != p.shape or g.stride() != p.stride() or g.dtype != torch.float32 or not g.is_cuda:
                raise TypeError('gradient buffer of plane %d must match its shape and strides' % i)
            d.grad = g.data_ptr()
        else:
            d.grad = None
    blo, bext_f, bblo, bbext = bounds_to_host(bound, bounding_box)
    for a in range(3):
        sc.bound_lo[a], sc.bound_ext[a] = blo[a], bext_f[a]
        sc.bbox_lo[a], sc.bbox_ext[a] = bblo[a], bbext[a]
    ws = [_f32(w, 'decoder weight') for w in weights]
    keep.extend(ws)
    sc.w_sdf0, sc.w_sdf1, sc.w_col0, sc.w_col1 = [w.data_ptr() for w in ws]
    in_sdf = feat + 48
    in_col = 48 + 15 + (feat if sc.two_grid else 0)
    shapes = [(32, in_sdf), (16, 32), (32, in_col), (3, 32)]
    for w, s in zip(ws, shapes):
        if tuple(w.shape) != s:
            raise ValueError('decoder weight shape %s, expected %s' % (tuple(w.shape), s))
    if weight_grads is not None:
        for g, w in zip(weight_grads, ws):
            if g is not None and (g.shape != w.shape or not g.is_contiguous()):
                raise TypeError('decoder gradient buffers must be contiguous and match the weights')
        sc.g_sdf0, sc.g_sdf1, sc.g_col0, sc.g_col1 = [_ptr(g) for g in weight_grads]
    if dirty is not None:
        if dirty.dtype != torch.uint8 or dirty_base is None or dirty.numel() * 32 < dirty_base.numel():
            raise TypeError('dirty must be a uint8 map with one byte per 32 floats of dirty_base')
        sc.dirty, sc.dirty_base = dirty.data_ptr(), dirty_base.data_ptr()
        keep.extend([dirty, dirty_base])
    sc._keep = keep
    return sc


# --------------------------------------------------------------------------------------------------
# entry points
# --------------------------------------------------------------------------------------------------
def new_acc(device):
    return torch.zeros(ACC_SLOTS, dtype=torch.float64, device=device)


def sample_rays(sample_cfg, render_cfg, n_rays, n_samples, target_d, noise, acc, device, out=None):
    if out is not None:
        if tuple(out.shape) != (n_rays, n_samples) or out.dtype != torch.float32 or not out.is_contiguous():
            raise TypeError('out must be a contiguous float32 [n_rays, n_samples] tensor')
        z = out
    else:
        z = torch.empty(n_rays, n_samples, dtype=torch.float32, device=device)
    td = _f32(target_d, 'target_d')
    nz = _f32(noise, 'noise')
    if nz is not None and tuple(nz.shape) != (n_rays, n_samples):
        raise ValueError('noise must have shape [%d, %d], got %s' % (n_rays, n_samples, tuple(nz.shape)))
    _lib.check(_lib.lib().mne_sample_rays(C.byref(sample_cfg), C.byref(render_cfg), n_rays, _ptr(td), _ptr(nz),
                                          _ptr(z), _ptr(acc), _stream()), 'mne_sample_rays')
    return z


def act_floats(scene):
    """Floats per sample of the record the forward keeps for the backward (``mne_act_floats``)."""
    return _lib.lib().mne_act_floats(C.byref(scene))


def render_forward(scene, render_cfg, rays_o, rays_d, z_vals, target_rgb=None, target_d=None, acc=None,
                   want_weights=False, want_aux=True, act=None, ray_counters=None, loss_grads=None, scratch=None):
    """The fused render (``mne_render_forward_fused``): encode -> decode -> compositing -> loss sums, as ONE launch when
    ``ray_counters`` is given, else as the point kernel followed by one register-resident ray kernel.
    ``act``: True = allocate, or a float32 buffer of n * s * ``act_floats(scene)`` elements: the per-sample record (hidden
    activations, features, ReLU masks) is kept there for the backward (returned as ``out['act']``).
    ``ray_counters``: int32 [>= n] zeros (the kernel leaves them zero), True = a fresh buffer, None = two launches.
    ``loss_grads`` + ``scratch`` (training): the kernel also writes d loss / d raw and the list of samples where it is
    non-zero into ``scratch`` (``backward_scratch``), so that ``render_backward_points`` can follow directly."""
    n, s = z_vals.shape
    dev = z_vals.device
    ro, rd = _f32(rays_o, 'rays_o'), _f32(rays_d, 'rays_d')
    trgb, td = _f32(target_rgb, 'target_rgb'), _f32(target_d, 'target_d')
    out = {'raw': torch.empty(n, s, 4, dtype=torch.float32, device=dev),
           'rgb': torch.empty(n, 3, dtype=torch.float32, device=dev),
           'depth': torch.empty(n, dtype=torch.float32, device=dev)}
    if want_aux:
        for k in ('disp_map', 'acc_map', 'depth_var'):
            out[k] = torch.empty(n, dtype=torch.float32, device=dev)
        out['inds'] = torch.empty(n, dtype=torch.int32, device=dev)
    if want_weights:
        out['weights'] = torch.empty(n, s, dtype=torch.float32, device=dev)
    if act is True:
        act = torch.empty(n * s * act_floats(scene), dtype=torch.float32, device=dev)
    if act is not None:
        if act.dtype != torch.float32 or act.numel() < n * s * act_floats(scene) or not act.is_contiguous() or act.data_ptr() % 32:
            raise TypeError('act must be a contiguous, 32-byte aligned float32 buffer of at least n_rays * n_samples * %d elements' % act_floats(scene))
        out['act'] = act
    if ray_counters is True:
        ray_counters = torch.zeros(n, dtype=torch.int32, device=dev)
    elif ray_counters is None:
        pass
    elif ray_counters.dtype != torch.int32 or ray_counters.numel() < n or not ray_counters.is_contiguous():
        raise TypeError('ray_counters must be a contiguous int32 buffer of at least n_rays zeros')
    lg = _f32(loss_grads, 'loss_grads')
    if lg is not None and scratch is None:
        raise ValueError('loss_grads needs a scratch buffer (ops.backward_scratch)')
    _lib.check(_lib.lib().mne_render_forward_fused(
        C.byref(scene), C.byref(render_cfg), n, s, _ptr(ro), _ptr(rd), _ptr(z_vals), _ptr(trgb), _ptr(td),
        _ptr(out['raw']), _ptr(out.get('weights')), _ptr(out['rgb']), _ptr(out['depth']), _ptr(out.get('disp_map')),
        _ptr(out.get('acc_map')), _ptr(out.get('depth_var')), _ptr(out.get('inds')), _ptr(acc), _ptr(act),
        _ptr(ray_counters), _ptr(lg), _ptr(scratch) if lg is not None else None, _stream()),
        'mne_render_forward_fused')
    return out


def render_forward_unfused(scene, render_cfg, rays_o, rays_d, z_vals, target_rgb=None, target_d=None, acc=None, act=None):
    """``mne_render_forward``: the point kernel followed by the stand-alone ray kernel (two launches, no counters);
    kept as the cross-check of the fused entry point and for impl 1."""
    n, s = z_vals.shape
    dev = z_vals.device
    ro, rd = _f32(rays_o, 'rays_o'), _f32(rays_d, 'rays_d')
    trgb, td = _f32(target_rgb, 'target_rgb'), _f32(target_d, 'target_d')
    out = {'raw': torch.empty(n, s, 4, dtype=torch.float32, device=dev),
           'rgb': torch.empty(n, 3, dtype=torch.float32, device=dev),
           'depth': torch.empty(n, dtype=torch.float32, device=dev),
           'weights': torch.empty(n, s, dtype=torch.float32, device=dev),
           'inds': torch.empty(n, dtype=torch.int32, device=dev)}
    for k in ('disp_map', 'acc_map', 'depth_var'):
        out[k] = torch.empty(n, dtype=torch.float32, device=dev)
    _lib.check(_lib.lib().mne_render_forward(
        C.byref(scene), C.byref(render_cfg), n, s, _ptr(ro), _ptr(rd), _ptr(z_vals), _ptr(trgb), _ptr(td),
        _ptr(out['raw']), _ptr(out['weights']), _ptr(out['rgb']), _ptr(out['depth']), _ptr(out['disp_map']),
        _ptr(out['acc_map']), _ptr(out['depth_var']), _ptr(out['inds']), _ptr(acc), _ptr(act), _stream()),
        'mne_render_forward')
    return out


def loss_finalize(render_cfg, n_samples, acc):
    losses = torch.empty(N_LOSSES, dtype=torch.float32, device=acc.device)
    _lib.check(_lib.lib().mne_loss_finalize(C.byref(render_cfg), n_samples, _ptr(acc), _ptr(losses), _stream()),
               'mne_loss_finalize')
    return losses


def backward_scratch(n_rays, n_samples, device):
    nbytes = _lib.lib().mne_backward_scratch_bytes(n_rays, n_samples)
    return torch.empty(max(nbytes // 4, 4), dtype=torch.float32, device=device)


def backward_active_count(scratch, n_rays, n_samples):
    """How many samples the last mne_render_backward on this scratch buffer processed (those with a non-zero loss
    gradient); layout of the buffer: d_raw float4[n] | sample list int32[round4(n)] | count.  Synchronises."""
    n = n_rays * n_samples
    return int(scratch.view(torch.int32)[n * 4 + (n + 3) // 4 * 4].item())


def render_backward(scene, render_cfg, rays_o, rays_d, z_vals, raw, target_rgb=None, target_d=None, acc=None,
                    loss_grads=None, g_rgb=None, g_depth=None, g_raw=None, want_ray_grads=False, scratch=None, act=None):
    n, s = z_vals.shape
    dev = z_vals.device
    ro, rd = _f32(rays_o, 'rays_o'), _f32(rays_d, 'rays_d')
    trgb, td = _f32(target_rgb, 'target_rgb'), _f32(target_d, 'target_d')
    lg, grgb, gdep, graw = (_f32(loss_grads, 'loss_grads'), _f32(g_rgb, 'g_rgb'), _f32(g_depth, 'g_depth'),
                            _f32(g_raw, 'g_raw'))
    if scratch is None:
        scratch = backward_scratch(n, s, dev)
    g_o = g_d = None
    if want_ray_grads:
        g_o = torch.empty(n, 3, dtype=torch.float32, device=dev)
        g_d = torch.empty(n, 3, dtype=torch.float32, device=dev)
    _lib.check(_lib.lib().mne_render_backward(
        C.byref(scene), C.byref(render_cfg), n, s, _ptr(ro), _ptr(rd), _ptr(z_vals), _ptr(raw), _ptr(trgb), _ptr(td),
        _ptr(acc), _ptr(lg), _ptr(grgb), _ptr(gdep), _ptr(graw), _ptr(scratch), _ptr(g_o), _ptr(g_d), _ptr(act), _stream()),
        'mne_render_backward')
    return g_o, g_d


def render_backward_points(scene, rays_o, rays_d, z_vals, scratch, want_ray_grads=False, act=None):
    """The point stage of the backward over a scratch buffer the fused training forward filled (``mne_render_backward_points``):
    plane / table gradients (and decoder gradients when the scene carries their buffers) are accumulated, ray gradients
    returned on request."""
    n, s = z_vals.shape
    dev = z_vals.device
    ro, rd = _f32(rays_o, 'rays_o'), _f32(rays_d, 'rays_d')
    g_o = g_d = None
    if want_ray_grads:
        g_o = torch.empty(n, 3, dtype=torch.float32, device=dev)
        g_d = torch.empty(n, 3, dtype=torch.float32, device=dev)
    _lib.check(_lib.lib().mne_render_backward_points(C.byref(scene), n, s, _ptr(ro), _ptr(rd), _ptr(z_vals), _ptr(scratch),
                                                     _ptr(g_o), _ptr(g_d), _ptr(act), _stream()), 'mne_render_backward_points')
    return g_o, g_d


def decoder_wgrad(scene, n_rays, n_samples, scratch, act, max_ctas=0):
    """Decoder weight gradients from the records a ``render_backward_points`` call completed (``mne_decoder_wgrad``); the
    call before must have been made WITHOUT decoder gradient buffers in its scene, this one with them."""
    _lib.check(_lib.lib().mne_decoder_wgrad(C.byref(scene), n_rays, n_samples, _ptr(scratch), _ptr(act), int(max_ctas), _stream()),
               'mne_decoder_wgrad')


def query_points(scene, pts, want_raw=True, want_geo=False, want_feat=False):
    p = _f32(pts, 'pts').reshape(-1, 3)
    m, dev = p.shape[0], p.device
    raw = torch.empty(m, 4, dtype=torch.float32, device=dev) if want_raw else None
    geo = torch.empty(m, 15, dtype=torch.float32, device=dev) if want_geo else None
    feat = torch.empty(m, 32 if scene.encoder == 1 else 64, dtype=torch.float32, device=dev) if want_feat else None
    _lib.check(_lib.lib().mne_query_points(C.byref(scene), m, _ptr(p), _ptr(raw), _ptr(geo), _ptr(feat), _stream()),
               'mne_query_points')
    return raw, geo, feat
