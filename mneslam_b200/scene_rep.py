"""Drop-in replacement for the reference's ``model/scene_rep.py::JointEncoding``.

Same constructor, attributes (``all_planes``, ``bound``, ``bounding_box``, ``decoder``, ``config``), method
names, argument meaning and return keys as the reference class; the work is done by the
fused sm_100a kernels in ``libmneslam_b200.so`` behind ``torch.autograd.Function`` wrappers.  There is no
torch / CPU fallback: without a CUDA device or without the library every compute method raises.

Differences a caller can observe (all documented in DESIGN.md):
* the feature planes are allocated in ``torch.channels_last`` memory format (same logical ``[1,32,H,W]``
  shape, so checkpoints round-trip); planes in any other layout (e.g. loaded from a reference checkpoint)
  still work through strides, only slower;
* the stratified noise is drawn with ``torch.rand`` on the device instead of on the CPU
  (``model/scene_rep.py:381``); pass ``noise=`` (or set ``self.noise_override``) for bit-identical samples;
* ``disp_map`` / ``acc_map`` / ``depth_var`` / ``z_vals`` are returned without a gradient path (they enter no
  loss in the reference, ``model/scene_rep.py:597-609``).
"""
import numpy as np
import torch
import torch.nn as nn

from . import ops
from ._lib import LOSS_KEYS
from .decoder import ColorSDFNet, ColorSDFNet_v2


def get_rays(H, W, fx, fy, cx, cy, c2w, device):
    """Rays of a whole image, OpenGL convention (reference ``model/utils.py:6-25``)."""
    if isinstance(c2w, np.ndarray):
        c2w = torch.from_numpy(c2w)
    c2w = c2w.to(device=device, dtype=torch.float32)
    i = torch.arange(W, dtype=torch.float32, device=device)[None, :].expand(H, W)
    j = torch.arange(H, dtype=torch.float32, device=device)[:, None].expand(H, W)
    dirs = torch.stack([(i - cx) / fx, -(j - cy) / fy, -torch.ones_like(i)], -1)
    rays_d = torch.sum(dirs[..., None, :] * c2w[:3, :3], -1)
    rays_o = c2w[:3, -1].expand(rays_d.shape)
    return rays_o, rays_d


class GradArena:
    """One flat gradient buffer for a fixed list of parameter tensors, with a view per parameter (its shape and strides):
    ``_Render.backward`` accumulates into it and returns the views, instead of allocating and zero-filling 129 MB of fresh
    gradient tensors per call.  ``clean`` says the buffer is all zero: ``mneslam_b200.optim.Adam`` clears the gradients inside
    its update kernel and sets it; otherwise the next backward clears the buffer itself (one memset, no allocation)."""
    registry = []            # weak references, looked up by mneslam_b200.optim.Adam

    def __init__(self, params):
        import weakref
        self.key = GradArena.key_of(params)
        self.numel = sum(p.numel() for p in params)
        self.flat = torch.zeros(self.numel, dtype=torch.float32, device=params[0].device)
        self.layout, off = [], 0
        for p in params:
            self.layout.append((tuple(p.shape), p.stride(), off))
            off += p.numel()
        self.clean = True
        GradArena.registry[:] = [r for r in GradArena.registry if r() is not None] + [weakref.ref(self)]

    def make_views(self):
        """Fresh tensor objects over the buffer, one per parameter.  (Fresh on purpose: autograd adopts an incoming gradient as
        ``.grad`` without copying it only when nobody else references that tensor object.)"""
        return [torch.as_strided(self.flat, shape, stride, off) for shape, stride, off in self.layout]

    @staticmethod
    def key_of(params):
        return tuple((tuple(p.shape), p.stride()) for p in params)

    @staticmethod
    def owning(ptr):
        """The live arena whose buffer contains device address ``ptr``, or None."""
        for r in GradArena.registry:
            a = r()
            if a is not None and a.flat.data_ptr() <= ptr < a.flat.data_ptr() + 4 * a.numel:
                return a
        return None


class _Render(torch.autograd.Function):
    """sample -> fused forward (+ loss sums, hidden activations kept when a backward can follow) ; backward = fused backward
    kernels over the samples with a non-zero loss gradient."""

    @staticmethod
    def forward(ctx, model, rays_o, rays_d, target_rgb, target_d, noise, with_losses, *params):
        cfg = model.config
        dev = rays_o.device
        n = rays_o.shape[0]
        with_depth = target_d is not None
        n_planes = len(params) - 4
        planes, weights = list(params[:n_planes]), list(params[n_planes:])
        tables = ops.SampleTables.get(cfg, dev, with_depth)
        n_samples = tables[0].numel() + tables[1].numel()
        perturb = cfg['training']['perturb'] > 0.
        if perturb and noise is None:
            noise = torch.rand(n, n_samples, device=dev)
        rcfg = ops.make_render_cfg(cfg, norm_rays=n * model.norm_scale)
        scfg = ops.make_sample_cfg(tables, perturb)
        td = target_d.reshape(-1) if with_depth else None
        acc = ops.new_acc(dev) if with_losses else None
        z_vals = ops.sample_rays(scfg, rcfg, n, n_samples, td, noise if perturb else None, acc, dev)
        scene = model._cached_scene(planes, weights)
        # what autograd would save: the decoders' hidden activations (80 floats per sample), only when a backward can follow
        want_act = any(ctx.needs_input_grad)
        out = ops.render_forward(scene, rcfg, rays_o, rays_d, z_vals, target_rgb if with_losses else None,
                                 td if with_losses else None, acc, act=True if want_act else None)
        if with_losses:
            if model.acc_allreduce is not None:
                model.acc_allreduce(acc)
            losses = ops.loss_finalize(rcfg, n_samples, acc)
        else:
            losses = torch.zeros(len(LOSS_KEYS), device=dev)
        ctx.model, ctx.rcfg, ctx.with_losses, ctx.n_planes = model, rcfg, with_losses, n_planes
        ctx.save_for_backward(rays_o, rays_d, z_vals, out['raw'], target_rgb if with_losses else None,
                              td if with_losses else None, acc, out.get('act'), *params)
        ctx.set_materialize_grads(False)
        nondiff = (z_vals, out['disp_map'], out['acc_map'], out['depth_var'], out['inds'])
        ctx.mark_non_differentiable(*nondiff)
        # the eight losses leave as eight 0-dim outputs: the caller weights and sums them (get_loss_from_ret), and their
        # gradients come straight back to backward() -- no select / scatter nodes (and their small kernels) in between
        return (out['rgb'], out['depth'], out['raw']) + tuple(losses.unbind(0)) + nondiff

    @staticmethod
    def backward(ctx, g_rgb, g_depth, g_raw, *rest):
        g_each = rest[:len(LOSS_KEYS)]
        g_losses = None
        if ctx.with_losses and any(g is not None for g in g_each):
            zero = g_rgb.new_zeros(()) if g_rgb is not None else next(g for g in g_each if g is not None).new_zeros(())
            g_losses = torch.stack([g if g is not None else zero for g in g_each])
        rays_o, rays_d, z_vals, raw, target_rgb, td, acc, act = ctx.saved_tensors[:8]
        params = ctx.saved_tensors[8:]
        model, n_planes = ctx.model, ctx.n_planes
        planes, weights = list(params[:n_planes]), list(params[n_planes:])
        need = ctx.needs_input_grad
        need_params = need[7:]
        plane_grads, weight_grads = model._gradient_buffers(planes, weights, need_params)
        want_rays = bool(need[1] or need[2])
        scene = model._cached_scene(planes, weights, plane_grads, weight_grads)
        lg = g_losses if (ctx.with_losses and g_losses is not None) else None
        g_o, g_d = ops.render_backward(scene, ctx.rcfg, rays_o, rays_d, z_vals, raw, target_rgb, td, acc,
                                       loss_grads=lg, g_rgb=g_rgb, g_depth=g_depth, g_raw=g_raw,
                                       want_ray_grads=want_rays, act=act)
        return (None, g_o if need[1] else None, g_d if need[2] else None, None, None, None, None,
                *plane_grads, *weight_grads)


class JointEncoding(nn.Module):
    def __init__(self, config, bound_box, ray_batch_size=4096):
        super().__init__()
        self.config = config
        self.bounding_box = bound_box
        self.device = torch.device('cuda' if torch.cuda.is_available() else 'cpu')
        self.load_bound(config)
        self.get_encoding(config)
        self.get_decoder(config)
        self.ray_batch_size = ray_batch_size
        # extensions (not in the reference): explicit noise for parity runs, multi-GPU hooks
        self.noise_override = None
        self._grad_arena = None
        self._scene_cache = {}
        self.norm_scale = 1          # world size when the ray batch is sharded over GPUs
        self.acc_allreduce = None    # callable(acc) all-reducing the loss accumulators across ranks

    # ------------------------------------------------------------------ construction
    def load_bound(self, cfg):
        """Enlarged scene box for the planes (reference ``model/scene_rep.py:72-83``)."""
        self.bound = torch.from_numpy(np.array(cfg['mapping']['bound']) * cfg['scale']).float()
        q = cfg['planes_res']['bound_dividable']
        self.bound[:, 1] = (((self.bound[:, 1] - self.bound[:, 0]) / q).int() + 1) * q + self.bound[:, 0]

    def _init_planes(self, res_key):
        c_dim = self.config['model']['c_dim']
        if c_dim != 32:
            raise NotImplementedError('the fused kernels are built for model.c_dim = 32 (every reference config)')
        ext = self.bound[:, 1] - self.bound[:, 0]
        groups = ([], [], [])
        # All planes of the set live back to back in ONE buffer, in all_planes order (group, level), each channels-last (a
        # texel's 32 channels are one 128-byte line): an optimiser that finds its parameters adjacent in memory
        # (mneslam_b200.optim.Adam) can then update the whole map with one launch.  Values are drawn in NCHW order, plane by
        # plane in the reference's order (same stream of normals as model/scene_rep.py:105-109).
        shapes = []
        for li, res in enumerate((self.config[res_key]['coarse'], self.config[res_key]['fine'])):
            nx, ny, nz = [int(v) for v in (ext / res).tolist()]
            for gi, hw in enumerate(((ny, nx), (nz, nx), (nz, ny))):
                shapes.append((gi, li, hw))
        order = sorted(range(len(shapes)), key=lambda k: (shapes[k][0], shapes[k][1]))       # all_planes order
        offsets, off = {}, 0
        for k in order:
            offsets[k] = off
            off += c_dim * shapes[k][2][0] * shapes[k][2][1]
        arena = torch.empty(off, dtype=torch.float32, device=self.device)
        made = {}
        for k, (gi, li, hw) in enumerate(shapes):
            p = torch.empty([1, c_dim, *hw]).normal_(mean=0, std=0.01)
            view = arena[offsets[k]:offsets[k] + p.numel()].view(1, hw[0], hw[1], c_dim).permute(0, 3, 1, 2)
            view.copy_(p)
            made[(gi, li)] = view
        for gi in range(3):
            for li in range(2):
                groups[gi].append(made[(gi, li)])
        return groups

    def init_all_planes(self, config):
        self.coarse_planes_res = config['planes_res']['coarse']
        self.fine_planes_res = config['planes_res']['fine']
        return self._init_planes('planes_res')

    def init_all_c_planes(self, config):
        self.coarse_c_planes_res = config['c_planes_res']['coarse']
        self.fine_c_planes_res = config['c_planes_res']['fine']
        return self._init_planes('c_planes_res')

    def get_encoding(self, config):
        if 'blob' not in config['pos']['enc'].lower() or config['pos']['n_bins'] != 16:
            raise NotImplementedError('the fused kernels are built for pos.enc = OneBlob with 16 bins')
        self.input_ch_pos = config['model']['input_ch_pos']
        self.hash_meta, self.hash_tables = None, None
        self.hash_fp16, self.hash_tables_h = False, None
        if config['grid'].get('use_hash', False):
            # the encoder the configs NAME (grid.enc: HashGrid + hash_size + voxel_sdf, model/scene_rep.py:55-70,160)
            # but the reference never builds; opt-in through the extension key grid.use_hash
            from .hashgrid import HashGridMeta
            self.hash_meta = HashGridMeta.from_config(config)
            self.hash_tables = [self.hash_meta.new_table(self.device)]      # mutable list, like all_planes
            if not config['grid']['oneGrid']:
                # the colour encoder of the (commented) model/scene_rep.py:169: same grid, resolution from voxel_color
                self.hash_meta.color = HashGridMeta.from_config(config, key='voxel_color')
                self.hash_tables.append(self.hash_meta.color.new_table(self.device))
            self.input_ch = self.hash_meta.n_output
            # grid.hash_fp16 (extension key): the kernels read the features from an fp16 copy of the table(s), refreshed from
            # the fp32 master before every render -- tiny-cuda-nn's default parameter precision (parity gate 1e-2)
            self.hash_fp16 = bool(config['grid'].get('hash_fp16', False))
            self.all_planes = ()
            if self.input_ch_pos != 48:
                raise NotImplementedError('model.input_ch_pos must be 48 (3 x n_bins)')
            return
        self.input_ch = config['model']['input_ch']
        if (self.input_ch, self.input_ch_pos) != (64, 48):
            raise NotImplementedError('model.input_ch / input_ch_pos must be 64 / 48 (2 x c_dim, 3 x n_bins)')
        self.all_planes = self.init_all_planes(config)
        if not config['grid']['oneGrid']:
            self.all_planes = self.all_planes + self.init_all_c_planes(config)

    def get_decoder(self, config):
        cls = ColorSDFNet_v2 if config['grid']['oneGrid'] else ColorSDFNet
        self.decoder = cls(config, input_ch=self.input_ch, input_ch_pos=self.input_ch_pos)
        self.color_net = self.decoder.color_net
        self.sdf_net = self.decoder.sdf_net
        self._pack_decoder()

    def _pack_decoder(self):
        """The four decoder matrices back to back in one buffer (each keeps its identity, shape and state_dict key; only its
        storage moves), so that an optimiser which looks for adjacent parameters updates them with one launch."""
        ws = self.decoder.weight_list()
        flat = torch.empty(sum(w.numel() for w in ws), dtype=ws[0].dtype, device=ws[0].device)
        off = 0
        with torch.no_grad():
            for w in ws:
                view = flat[off:off + w.numel()].view_as(w)
                view.copy_(w)
                w.data = view
                off += w.numel()

    def _apply(self, fn, *args, **kwargs):
        """``.to(device)`` / ``.cuda()`` (mneslam_mp.py:114) move every registered parameter on its own: pack the decoder again."""
        out = super()._apply(fn, *args, **kwargs)
        self._pack_decoder()
        return out

    # ------------------------------------------------------------------ helpers
    def _gradient_buffers(self, planes, weights, need):
        """Zeroed gradient tensors for ``_Render.backward``: views of the model's persistent ``GradArena`` when every
        parameter wants a gradient and nobody still holds last step's views as ``.grad`` (gradient accumulation across
        backward calls) -- else freshly allocated ones, as autograd itself would produce."""
        params = planes + weights
        if not all(need) or any(p.grad is not None for p in params):
            return ([torch.zeros_like(p) if nd else None for p, nd in zip(planes, need[:len(planes)])],
                    [torch.zeros_like(w) if nd else None for w, nd in zip(weights, need[len(planes):])])
        arena = self._grad_arena
        if arena is None or arena.key != GradArena.key_of(params) or arena.flat.device != params[0].device:
            arena = self._grad_arena = GradArena(params)
        if not arena.clean:
            arena.flat.zero_()
        arena.clean = False
        views = arena.make_views()
        return views[:len(planes)], views[len(planes):]

    def _cached_scene(self, planes, weights, plane_grads=None, weight_grads=None):
        """``ops.make_scene`` memoised on what the descriptor is made of (addresses, shapes and strides of every tensor, the
        two boxes): the reference's callers may swap planes or weights between calls (mneslam_mp.py:451-457,
        mp_slam/mapper.py:719), so nothing is assumed -- but an unchanged model costs one tuple comparison instead of
        rebuilding a 2.5 KB ctypes structure several times per iteration."""
        def sig(ts):
            return tuple((t.data_ptr(), t.shape, t.stride()) if t is not None else None for t in ts)
        key = (sig(planes), sig(weights), sig(plane_grads) if plane_grads is not None else None,
               sig(weight_grads) if weight_grads is not None else None,
               id(self.bound), self.bound._version, id(self.bounding_box), self.bounding_box._version, id(self.hash_meta),
               sig(self.hash_tables_h) if self.hash_tables_h is not None else None)
        sc = self._scene_cache.get(key)
        if sc is None:
            if len(self._scene_cache) > 8:
                self._scene_cache.clear()
            sc = ops.make_scene(planes, weights, self.bound, self.bounding_box, self.config['pos']['n_bins'],
                                plane_grads=plane_grads, weight_grads=weight_grads, hash_meta=self.hash_meta,
                                hash_half=self.hash_tables_h)
            self._scene_cache[key] = sc
        return sc

    def refresh_hash_half(self):
        """grid.hash_fp16: bring the fp16 copies of the hash tables in step with the fp32 masters (one conversion pass per
        table; called before every render, and by FusedMapper after its optimiser step)."""
        if not self.hash_fp16:
            self.hash_tables_h = None
            return
        if self.hash_tables_h is None or len(self.hash_tables_h) != len(self.hash_tables) or \
                any(h.shape != t.shape or h.device != t.device for h, t in zip(self.hash_tables_h, self.hash_tables)):
            self.hash_tables_h = [torch.empty(t.shape, dtype=torch.float16, device=t.device) for t in self.hash_tables]
        with torch.no_grad():
            for h, t in zip(self.hash_tables_h, self.hash_tables):
                h.copy_(t)

    def _flat_planes(self):
        """The encoder's parameter tensors: the planes in ``all_planes`` order, or the hash table(s)."""
        if self.hash_meta is not None:
            return list(self.hash_tables)
        expect = 3 if self.config['grid']['oneGrid'] else 6
        if len(self.all_planes) != expect:
            raise ValueError('all_planes has %d groups, config expects %d' % (len(self.all_planes), expect))
        planes = [p for group in self.all_planes for p in group]
        for p in planes:
            # a map that arrived from outside (a reference checkpoint, mp_slam/mapper.py:719) is NCHW-contiguous: lay it out
            # channels-last ONCE, in place (same tensor object, same logical shape), instead of serving it through the strided
            # kernels from then on.  Left alone while a gradient or an optimiser state of the old layout may exist.
            if p.is_cuda and p.dim() == 4 and p.stride(1) != 1 and p.grad is None and not getattr(p, '_mne_keep_layout', False):
                with torch.no_grad():
                    p.data = p.data.contiguous(memory_format=torch.channels_last)
        return planes

    def _scene(self):
        self.refresh_hash_half()
        return self._cached_scene(self._flat_planes(), self.decoder.weight_list())

    def _render(self, rays_o, rays_d, target_rgb, target_d, with_losses, noise=None):
        ops.require_cuda()
        if self.config['training']['n_importance'] > 0:
            from .importance import render_with_importance
            return render_with_importance(self, rays_o, rays_d, target_rgb, target_d, with_losses, noise)
        noise = noise if noise is not None else self.noise_override
        self.refresh_hash_half()
        params = self._flat_planes() + self.decoder.weight_list()
        res = _Render.apply(self, rays_o, rays_d, target_rgb, target_d, noise, with_losses, *params)
        rgb, depth, raw = res[:3]
        losses = res[3:3 + len(LOSS_KEYS)]
        z_vals, disp, acc_map, depth_var, inds = res[3 + len(LOSS_KEYS):]
        rend = {'rgb': rgb, 'depth': depth, 'disp_map': disp, 'acc_map': acc_map, 'depth_var': depth_var,
                'z_vals': z_vals, 'raw': raw}
        return rend, losses, inds

    # ------------------------------------------------------------------ reference API: rendering
    def render_rays(self, rays_o, rays_d, target_d=None, noise=None):
        """reference ``model/scene_rep.py:351-419``: dict with rgb, depth, disp_map, acc_map, depth_var, z_vals,
        raw (+ rgb0, disp0, acc0, depth0, depth_var0, z_std when ``training.n_importance > 0``)."""
        rend, _, _ = self._render(rays_o, rays_d, None, target_d, False, noise)
        return rend

    def forward(self, rays_o, rays_d, target_rgb, target_d, global_step=0, noise=None):
        """reference ``model/scene_rep.py:549-611``."""
        if not self.training:
            return self.render_rays(rays_o, rays_d, target_d=target_d, noise=noise)
        rend, losses, _ = self._render(rays_o, rays_d, target_rgb, target_d, True, noise)
        ret = {'rgb': rend['rgb'], 'depth': rend['depth']}
        for i, k in enumerate(LOSS_KEYS[:-1]):
            ret[k] = losses[i]
        ret['psnr'] = losses[7].detach().reshape(1)
        return ret

    def render_img(self, c2w, device, gt_depth=None):
        """reference ``model/scene_rep.py:422-473``: (depth [H,W] float64, colour [H,W,3])."""
        with torch.no_grad():
            cam = self.config['cam']
            H = cam['H'] - 2 * cam['crop_edge']
            W = cam['W'] - 2 * cam['crop_edge']
            rays_o, rays_d = get_rays(H, W, cam['fx'], cam['fy'], cam['cx'], cam['cy'], c2w, device)
            rays_o = rays_o.reshape(-1, 3).contiguous()
            rays_d = rays_d.reshape(-1, 3).contiguous()
            if gt_depth is not None:
                gt_depth = gt_depth.reshape(-1).unsqueeze(1).to(device)
            depths, colours = [], []
            # rays are rendered independently, so the chunk size only sets the launch count: at least 64 Ki rays per launch
            # (the reference's 4096-ray chunks are 200 launches for a 1200 x 680 image)
            bs = getattr(self, 'render_chunk', None) or max(self.ray_batch_size, 65536)
            for i in range(0, rays_d.shape[0], bs):
                td = None if gt_depth is None else gt_depth[i:i + bs]
                ret = self.render_rays(rays_o[i:i + bs], rays_d[i:i + bs], target_d=td)
                depths.append(ret['depth'].double())
                colours.append(ret['rgb'])
            return torch.cat(depths, 0).reshape(H, W), torch.cat(colours, 0).reshape(H, W, 3)

    # ------------------------------------------------------------------ reference API: point queries
    def query_color_sdf(self, query_points):
        """raw (rgb logits, sdf) at points ``[..., 3]`` -> ``[M, 4]`` flattened like the reference
        (``model/scene_rep.py:273-301``)."""
        ops.require_cuda()
        raw, _, _ = ops.query_points(self._scene(), query_points.reshape(-1, 3).float())
        return raw

    def run_network(self, inputs):
        flat = self.query_color_sdf(inputs)
        return flat.reshape(list(inputs.shape[:-1]) + [4])

    def run_network_flat(self, inputs_flat):
        return self.query_color_sdf(inputs_flat)

    def query_sdf(self, query_points, return_geo=False, embed=False):
        """reference ``model/scene_rep.py:232-268``."""
        ops.require_cuda()
        lead = list(query_points.shape[:-1])
        pts = query_points.reshape(-1, 3).float()
        if embed:
            _, _, feat = ops.query_points(self._scene(), pts, want_raw=False, want_feat=True)
            return feat.reshape(lead + [feat.shape[-1]])
        raw, geo, _ = ops.query_points(self._scene(), pts, want_geo=return_geo)
        sdf = raw[:, 3].reshape(lead)
        if not return_geo:
            return sdf
        return sdf, geo.reshape(lead + [geo.shape[-1]])

    def query_color(self, query_points):
        return torch.sigmoid(self.query_color_sdf(query_points)[..., :3])

    def raw2outputs(self, raw, z_vals, white_bkgd=False):
        """reference ``model/scene_rep.py:205-230`` (forward values only): rgb_map, disp_map, acc_map, weights,
        depth_map, depth_var from raw ``[N,S,4]`` and z_vals ``[N,S]``."""
        import ctypes as C
        from . import _lib
        ops.require_cuda()
        raw, z = ops._f32(raw, 'raw'), ops._f32(z_vals, 'z_vals')
        n, s = z.shape
        dev = z.device
        cfg = dict(self.config)
        cfg['training'] = dict(self.config['training'], white_bkgd=bool(white_bkgd))
        rcfg = ops.make_render_cfg(cfg, norm_rays=max(n, 1))
        o = {k: torch.empty(n, dtype=torch.float32, device=dev) for k in ('depth', 'disp', 'acc', 'var')}
        rgb = torch.empty(n, 3, dtype=torch.float32, device=dev)
        w = torch.empty(n, s, dtype=torch.float32, device=dev)
        _lib.check(_lib.lib().mne_composite(C.byref(rcfg), n, s, z.data_ptr(), raw.data_ptr(), w.data_ptr(), rgb.data_ptr(),
                                            o['depth'].data_ptr(), o['disp'].data_ptr(), o['acc'].data_ptr(), o['var'].data_ptr(),
                                            None, ops._stream()), 'mne_composite')
        return rgb, o['disp'], o['acc'], w, o['depth'], o['var']

    def sdf2weights(self, sdf, z_vals, args=None):
        """reference ``model/scene_rep.py:183-203`` (forward values only)."""
        raw = torch.zeros(*sdf.shape, 4, dtype=torch.float32, device=sdf.device)
        raw[..., 3] = sdf
        return self.raw2outputs(raw, z_vals)[3]

    def render_surface_color(self, rays_o, normal):
        """reference ``model/scene_rep.py:333-349``: composite colour over a short normal-aligned segment."""
        from .compose import composite_along
        trunc = self.config['training']['trunc']
        z = torch.linspace(-trunc, trunc, steps=self.config['training']['n_range_d']).to(rays_o)
        return composite_along(self, rays_o.reshape(-1, 3), normal.reshape(-1, 3), z)['rgb']
