"""Device-resident mapping iteration: the body of the reference's hot loop
(``mp_slam/mapper.py:134-162`` / ``:73-89``) as five kernel launches with no host synchronisation:

    sample (z_vals + mask counts) -> fused forward (+ loss sums) -> loss finalize
      -> fused backward (upstream = the caller's loss weights, mneslam_mp.py:350-372)
      -> fused Adam + gradient zero-fill (mneslam_mp.py:459-469, mapper.py:161-162)

so that a whole iteration (or several) can be captured in one CUDA graph.  ``JointEncoding.forward`` +
``loss.backward()`` + ``torch.optim.Adam`` remain available and give the same numbers through autograd; this
class is the fast path ``bench.py`` measures.

Multi-GPU (one process per GPU, ``torch.distributed`` / NCCL): every rank holds the full map, renders its own
shard of the ray batch, and the loss accumulators (16 doubles, before the losses are finalised) and the flat
gradient buffer are all-reduced once per iteration; the Adam step is then identical on every rank.
"""
import ctypes as C
import os

import torch

from . import _lib, ops
from ._lib import LOSS_KEYS


def loss_weight_vector(cfg, is_co_sdf=None):
    """d total / d each loss of ``get_loss_from_ret`` (mneslam_mp.py:350-372, rgb + depth + sdf terms)."""
    t, m = cfg['training'], cfg['mapping']
    if is_co_sdf is None:
        is_co_sdf = cfg['is_co_sdf']
    w = [t['rgb_weight'], t['depth_weight'], 0., 0., 0., 0., 0., 0.]
    if is_co_sdf:
        w[2], w[3] = t['sdf_weight'], t['fs_weight']
    else:
        w[4], w[5], w[6] = m['w_sdf_fs'], m['w_sdf_center'], m['w_sdf_tail']
    return w


PEER_CHUNK = 1024      # floats per ownership chunk (4 KB) of the peer-memory optimiser


N_BARRIER_CHANNELS = 8


class FusedMapper:
    def __init__(self, model, is_co_sdf=None, process_group=None, world_size=1, adopt_parameters=True, skip_untouched=True,
                 peer_memory=True, use_multicast=True, mark_dirty=None):
        """``adopt_parameters``: move the planes and decoder weights into ONE flat buffer (each tensor keeps its
        Python identity, shape and strides; only its storage changes), with gradients and Adam moments laid out
        the same way, so that the optimiser is a single streaming launch.  ``skip_untouched``: exact skip of parameter
        groups that never received a gradient.  Several GPUs (``world_size`` > 1, rays sharded over ranks):
        ``peer_memory`` puts the flat buffers into NVLink peer-mapped memory and runs the optimiser as ONE kernel that
        reduces the gradient over the peers and writes the parameters back (else NCCL reduce-scatter / all-gather);
        ``use_multicast`` lets that kernel store through the NVLink switch from 8 ranks up; the texels of the gradient that
        are non-zero this step are pushed to their owners together with one flag byte each.  ``mark_dirty``: the flags come
        from the backward kernels themselves (``MneSceneDesc.dirty``: +4.5 us on the backward kernel, then a pass over the
        1 MB flag map, 22 us) instead of a pass over the 129 MB gradient buffer (``mne_grad_push``, 42 us); None = wherever
        the kernels support it (channels-last planes; the hash grid adds 8-byte entries and takes the pass)."""
        ops.require_cuda()
        self.model, self.cfg = model, model.config
        self.world_size, self.group = world_size, process_group
        planes = model._flat_planes()
        weights = model.decoder.weight_list()
        self.dev = planes[0].device
        one_grid = self.cfg['grid']['oneGrid']
        mp = self.cfg['mapping']
        n_sdf_planes = 6 if model.hash_meta is None else 1
        # parameter groups of create_optimizer (mneslam_mp.py:459-469): (tensor, lr, eps, weight_decay)
        self.params = [(p, mp['lr_embed'], 1e-15, 0.0) for p in planes[:n_sdf_planes]]
        if not one_grid:
            self.params += [(p, mp['lr_embed_color'], 1e-15, 0.0) for p in planes[n_sdf_planes:]]
        self.params += [(w, mp['lr_decoder'], 1e-8, 1e-6) for w in weights]
        self.betas = (0.9, 0.99)
        for p, *_ in self.params:
            if not (p.is_contiguous() or p.is_contiguous(memory_format=torch.channels_last)) or p.numel() % 4:
                raise ValueError('map parameters must be dense tensors with a multiple of 4 elements')
        # the decoder's weights start on a 32-float (128-byte texel) boundary of the flat buffers, so that the texel-granular
        # gradient exchange of the encoder never straddles into them (hash tables are not a multiple of 32 floats long)
        enc_len = sum(p.numel() for p in planes)
        self._dec_pad = (-enc_len) % 32
        total = sum(p.numel() for p, *_ in self.params) + self._dec_pad
        if world_size > 1:
            if not adopt_parameters:
                raise ValueError('multi-GPU mapping needs adopt_parameters=True (flat buffers for the collectives)')
            q = world_size * PEER_CHUNK  # whole ownership chunks per rank (block-cyclic ownership of the peer optimiser)
            padded = (total + q - 1) // q * q
        else:
            padded = total
        self.total = total
        # several GPUs: gradient and parameter buffers in NVLink peer-mapped (symmetric) memory, so that the optimiser
        # can be ONE kernel that reduces over the peers and writes the result back to them (mne_adam_step_peers);
        # if that memory cannot be set up, the NCCL reduce-scatter / all-gather path is used instead
        self.peers = None
        if world_size > 1 and peer_memory:
            try:
                import torch.distributed._symmetric_memory as symm
                group = process_group if process_group is not None else torch.distributed.group.WORLD
                g_buf = symm.empty(padded, dtype=torch.float32, device=self.dev)
                p_buf = symm.empty(padded, dtype=torch.float32, device=self.dev)
                d_buf = symm.empty(padded // 32, dtype=torch.uint8, device=self.dev)     # [rank][owned texel] "non-zero this step" flags, pushed by the peers
                hg, hp = symm.rendezvous(g_buf, group.group_name), symm.rendezvous(p_buf, group.group_name)
                hd = symm.rendezvous(d_buf, group.group_name)
                d_buf.zero_()
                # inbox: [rank][owned elements] -- the peers push their non-zero gradient texels of this rank's slice here
                i_buf = symm.empty(padded, dtype=torch.float32, device=self.dev)
                hi = symm.rendezvous(i_buf, group.group_name)
                i_buf.zero_()
                # exchange buffer of the small all-reduces: per region [rank][payload]; regions: mask counts (6 doubles), loss
                # sums (10 doubles), decoder weight gradients
                n_dec = sum(w.numel() for w in weights)
                self._x_regions = {'counts': (0, 12), 'sums': (12 * world_size, 20), 'dec': (32 * world_size, n_dec)}
                x_buf = symm.empty(world_size * (32 + n_dec), dtype=torch.float32, device=self.dev)
                hx = symm.rendezvous(x_buf, group.group_name)
                x_buf.zero_()
                # signal pad of the library's own barrier kernel (mne_peer_barrier): channels 0-2 = the small all-reduces,
                # 3 = gradients pushed, 4 = parameters landed
                b_buf = symm.empty(_lib.peer_pad_words(N_BARRIER_CHANNELS), dtype=torch.int32, device=self.dev)
                hb = symm.rendezvous(b_buf, group.group_name)
                b_buf.zero_()
                torch.cuda.synchronize(self.dev)
                hb.barrier(channel=0)                  # every pad is zero before anybody signals
                torch.cuda.synchronize(self.dev)
                self._own_barrier = os.environ.get('MNE_OWN_BARRIER', '1') == '1'
                self.peers = {'xchg': x_buf, 'hx': hx, 'pad': b_buf, 'hb': hb,
                              'pad_ptrs': (C.c_void_p * world_size)(*[int(x) for x in hb.buffer_ptrs]), 'x_ptrs': (C.c_void_p * world_size)(*[int(x) for x in hx.buffer_ptrs]),
                              'grad': g_buf, 'param': p_buf, 'dirty': d_buf, 'hg': hg, 'hp': hp, 'hd': hd, 'inbox': i_buf, 'hi': hi,
                              'inbox_ptrs': (C.c_void_p * world_size)(*[int(x) for x in hi.buffer_ptrs]),
                              'dirty_ptrs': (C.c_void_p * world_size)(*[int(x) for x in hd.buffer_ptrs]),
                              'grad_ptrs': (C.c_void_p * world_size)(*[int(x) for x in hg.buffer_ptrs]),
                              'param_ptrs': (C.c_void_p * world_size)(*[int(x) for x in hp.buffer_ptrs]),
                              # NVLink-switch multicast addresses (0 when the fabric has no multicast support)
                              'mc_grad': int(getattr(hg, 'multicast_ptr', 0) or 0) if use_multicast else 0,
                              'mc_param': int(getattr(hp, 'multicast_ptr', 0) or 0) if use_multicast else 0}
                # measured (r02, room0): four ranks 0.264 ms per iteration with plain peer stores, 0.276 through the switch's multicast
                # address; eight ranks 0.287 / 0.288 (a tie: the inbound side is the bound either way, multicast sends 1/7 of the bytes)
                if world_size < int(os.environ.get('MNE_MULTICAST_FROM', '8')):
                    self.peers['mc_param'] = 0
                self.peers['mc_grad'] = 0             # the gradient side uses the sparse (dirty-map) peer reduction
            except Exception as e:      # no peer access / no symmetric-memory support: NCCL collectives instead
                import warnings
                warnings.warn('mneslam_b200: peer-memory optimiser unavailable (%r); using NCCL collectives' % (e,))
                self.peers = None
        if self.peers is not None:
            self.flat_grad = self.peers['grad'].zero_()
        else:
            self.flat_grad = torch.zeros(padded, dtype=torch.float32, device=self.dev)
        # single GPU: full-length Adam moments; multi GPU: each rank keeps the moments of its own slice only
        mv_len = padded // world_size
        self.flat_m = torch.zeros(mv_len, dtype=torch.float32, device=self.dev)
        self.flat_v = torch.zeros(mv_len, dtype=torch.float32, device=self.dev)
        if self.peers is not None:
            self.flat_param = self.peers['param'].zero_()
        else:
            self.flat_param = torch.zeros(padded, dtype=torch.float32, device=self.dev) if adopt_parameters else None
        if world_size > 1:            # the decoder's moments, replicated (its Adam step runs on every rank)
            n_dec = sum(w.numel() for w in weights)
            self.dec_m = torch.zeros(n_dec, dtype=torch.float32, device=self.dev)
            self.dec_v = torch.zeros(n_dec, dtype=torch.float32, device=self.dev)
        # one byte per 4 parameters of this rank's slice: groups never hit by a gradient are skipped by Adam (exact)
        self.touched = torch.zeros(mv_len // 4, dtype=torch.uint8, device=self.dev) if (adopt_parameters and skip_untouched) else None
        self.grads, self.exp_avg, self.exp_avg_sq = [], [], []
        seg_ends, seg_lr, seg_eps, seg_wd = [], [], [], []
        off = 0
        for k, (p, lr, eps, wd) in enumerate(self.params):
            n = p.numel()
            if k == len(planes) and self._dec_pad:
                off += self._dec_pad                # padding floats belong to the last encoder group (never touched, never updated)
                seg_ends[-1] = off
            self.grads.append(torch.as_strided(self.flat_grad, p.shape, p.stride(), off))
            if world_size == 1:
                self.exp_avg.append(torch.as_strided(self.flat_m, p.shape, p.stride(), off))
                self.exp_avg_sq.append(torch.as_strided(self.flat_v, p.shape, p.stride(), off))
            if adopt_parameters:
                view = torch.as_strided(self.flat_param, p.shape, p.stride(), off)
                with torch.no_grad():
                    view.copy_(p)
                p.data = view                       # same tensor object for every outside holder, new storage
            off += n
            if seg_lr and (seg_lr[-1], seg_eps[-1], seg_wd[-1]) == (lr, eps, wd):
                seg_ends[-1] = off                  # merge neighbours with identical hyper-parameters
            else:
                seg_ends.append(off); seg_lr.append(lr); seg_eps.append(eps); seg_wd.append(wd)
        self._segments = self._pack_segments(seg_ends, list(zip(seg_lr, seg_eps, seg_wd)))
        # the same split in two: the encoder's groups (planes / tables), and the decoder's group at the tail
        dec0 = enc_len + self._dec_pad
        self._seg_head = self._pack_segments(seg_ends[:-1], list(zip(seg_lr, seg_eps, seg_wd))[:-1])
        self._seg_tail = self._pack_segments([total - dec0], [(seg_lr[-1], seg_eps[-1], seg_wd[-1])])
        if len(seg_ends) < 2 or seg_ends[-2] != dec0:
            raise ValueError('the decoder weights must form the last parameter group')
        self.shards = None
        if world_size > 1:
            from .distributed import FlatShards
            rank = torch.distributed.get_rank(process_group)
            self.shards = FlatShards(total, world_size, rank, seg_ends, list(zip(seg_lr, seg_eps, seg_wd)), align=PEER_CHUNK)
            self.grad_shard = torch.zeros(self.shards.shard_len, dtype=torch.float32, device=self.dev)
            self._local_segments = self._pack_segments(self.shards.local_ends, self.shards.local_hparams)
        n_planes = len(planes)
        # several GPUs: the backward pass flags the 128-byte texels of the plane gradient it adds into, so that the owners
        # fetch only those over NVLink (needs the vector kernels: channels-last planes).  The hash grid adds 8-byte entries,
        # 128 per sample -- one flag store per addition cost more than the pass over the gradient buffer that derives the
        # flags afterwards (measured: +0.6 ms at 8192 rays), so hash tables take that pass (mne_grad_dirty_map).
        self.dirty_local = None
        self.dense_from = enc_len + self._dec_pad
        if os.environ.get('MNE_MARK_DIRTY') is not None:
            mark_dirty = os.environ['MNE_MARK_DIRTY'] == '1'
        if mark_dirty is None:
            mark_dirty = True
        vector_planes = model.hash_meta is None and len(planes) > 0 and all(p.stride(1) == 1 and p.shape[1] == 32 for p in planes)
        if self.peers is not None and mark_dirty and vector_planes:
            self.dirty_local = torch.zeros(padded // 32, dtype=torch.uint8, device=self.dev)
        # one GPU: the same flags can tell the optimiser which texels of the plane gradient are non-zero, so that it does not read
        # (nor clear) the other ~80 % of the 129 MB gradient buffer (mne_adam_step_segments_dirty).  Off by default -- measured on
        # B200 (room0, 2150 rays): the optimiser pass 111 -> 105 us, the flag stores cost the backward kernel 34 -> 38.5 us (one
        # store per tap; 52 us when every corner stored its own flag): the iteration stays at 0.235 ms
        self.adam_dirty = world_size == 1 and vector_planes and os.environ.get('MNE_ADAM_DIRTY', '0') == '1' and enc_len % 32 == 0
        if self.adam_dirty:
            self.dirty_local = torch.zeros((padded + 31) // 32, dtype=torch.uint8, device=self.dev)
            self._dirty_floats = enc_len
        self.plane_grads = self.grads[:n_planes]
        self.weight_grads = self.grads[n_planes:]
        self.step_dev = torch.zeros(1, dtype=torch.int64, device=self.dev)
        self.loss_grads = torch.tensor(loss_weight_vector(self.cfg, is_co_sdf), dtype=torch.float32, device=self.dev)
        self.loss_weights = self.loss_grads.clone()
        self._scratch = None
        self._ray_counters = None     # per-ray counters of the fused forward's tail (always left at zero)
        self._pending_wgrad = None    # the batch whose decoder weight gradients are still to be computed
        self._wgrad_stream = None     # side stream: decoder weight gradients + decoder Adam beside the planes' Adam
        self._loss_stream = None      # side stream (several GPUs): all-reduce of the loss sums + loss finalisation
        self._loss_pending = False
        self._tick_issued = False
        self.losses_on_side = False   # one GPU too: finalise the losses and advance the step counter beside the backward (PipelinedMapper)
        self._act = None              # the per-sample record, forward -> backward (ops.act_floats per sample)
        self.last_losses = None
        self.global_rays_override = None
        # CTAs of the decoder weight-gradient kernel when it runs beside the planes' Adam pass (the optimiser grids leave
        # `adam_spare_sms` SMs free; 0 = a full grid)
        self.wgrad_ctas = int(os.environ.get('MNE_WGRAD_CTAS', '0'))
        self.wgrad_inline = os.environ.get('MNE_WGRAD_INLINE', '0') == '1'
        # the render as ONE launch (the warp that stores a ray's last sample composites it; needs per-ray counters) or as the
        # point kernel + one ray kernel.  Measured on B200 (room0, 2048 rays): 82 us against 55 + 8 us -- finished rays
        # queue up behind the tiles of the warp that finished them, on the kernel's critical path -- so two launches.
        self.single_launch_render = False
        # several GPUs: push the non-zero gradient texels to their owners (mne_grad_push) instead of letting the owners fetch them
        self.push_gradients = os.environ.get('MNE_PUSH_GRADIENTS', '1') == '1'

    @staticmethod
    def _pack_segments(ends, hparams):
        n = len(ends)
        return (n, (C.c_int64 * n)(*ends), (C.c_float * n)(*[h[0] for h in hparams]),
                (C.c_float * n)(*[h[1] for h in hparams]), (C.c_float * n)(*[h[2] for h in hparams]))

    def _adam_segments(self, param, grad, m, v, segments, zero_grad, touched='own', dirty=None, dirty_floats=0):
        n_seg, ends, lr, eps, wd = segments
        if isinstance(touched, str):
            touched = self.touched
        _lib.check(_lib.lib().mne_adam_step_segments_dirty(param.data_ptr(), grad.data_ptr(), m.data_ptr(), v.data_ptr(), n_seg,
                                                           ends, lr, eps, wd, self.betas[0], self.betas[1],
                                                           self.step_dev.data_ptr(), 0, 1 if zero_grad else 0,
                                                           ops._ptr(touched), ops._ptr(dirty), dirty_floats, ops._stream()),
                   'mne_adam_step_segments')

    # ------------------------------------------------------------------------------------------
    def _scene(self, with_grads, plane_grads=True, weight_grads=True):
        m = self.model
        return ops.make_scene(m._flat_planes(), m.decoder.weight_list(), m.bound, m.bounding_box,
                              self.cfg['pos']['n_bins'],
                              plane_grads=self.plane_grads if (with_grads and plane_grads) else None,
                              weight_grads=self.weight_grads if (with_grads and weight_grads) else None, hash_meta=m.hash_meta,
                              hash_half=m.hash_tables_h,
                              dirty=self.dirty_local if with_grads else None,
                              dirty_base=self.flat_grad if (with_grads and self.dirty_local is not None) else None)

    def sample_stage(self, target_d, noise=None, z_out=None, acc_out=None):
        """The part of an iteration that does not read the map: sample positions along the rays and the mask counts
        (``render_rays``, model/scene_rep.py:361-381).  Returns (z_vals [N,S], acc[16]); with ``z_out`` / ``acc_out`` the
        results are written into those buffers (``PipelinedMapper`` keeps them across graph replays)."""
        cfg, dev = self.cfg, self.dev
        n = target_d.shape[0]
        tables = ops.SampleTables.get(cfg, dev, True)
        s = tables[0].numel() + tables[1].numel()
        perturb = cfg['training']['perturb'] > 0.
        if perturb and noise is None:
            noise = torch.rand(n, s, device=dev)
        rcfg = ops.make_render_cfg(cfg, norm_rays=self.global_rays_override or n * self.world_size)
        scfg = ops.make_sample_cfg(tables, perturb)
        acc = ops.new_acc(dev) if acc_out is None else acc_out.zero_()
        z_vals = ops.sample_rays(scfg, rcfg, n, s, target_d.reshape(-1), noise if perturb else None, acc, dev, out=z_out)
        if self.world_size > 1:
            # the loss gradients are scaled by the GLOBAL mask counts (slots 0..5), which are known as soon as the batch is
            # sampled: reduce them here, off the render's critical path when the batch is sampled ahead (PipelinedMapper)
            self._small_allreduce('counts', acc[:6], 0)
        return z_vals, acc

    def forward_backward(self, rays_o, rays_d, target_rgb, target_d, noise=None, sampled=None, defer_wgrad=False):
        """sample -> fused forward (encode, decode, compositing, loss sums, d loss / d raw, sample compaction: ONE launch) ->
        loss finalize -> point backward -> decoder weight gradients; gradients are ADDED to the flat buffer.  Returns
        losses[8].  ``sampled``: the (z_vals, acc) of an earlier ``sample_stage`` call on the same ``target_d``.
        ``defer_wgrad`` (one GPU): leave the decoder weight gradients to ``optimizer_step`` -- they only feed the decoder's
        own Adam step, so ``step`` runs them beside the planes' optimiser pass."""
        cfg, dev = self.cfg, self.dev
        n = rays_o.shape[0]
        tables = ops.SampleTables.get(cfg, dev, True)
        s = tables[0].numel() + tables[1].numel()
        # the means of the losses run over the GLOBAL batch; ragged shards set global_rays_override
        rcfg = ops.make_render_cfg(cfg, norm_rays=self.global_rays_override or n * self.world_size)
        td = target_d.reshape(-1)
        z_vals, acc = sampled if sampled is not None else self.sample_stage(target_d, noise)
        self.model.refresh_hash_half()            # (grid.hash_fp16: the kernels read an fp16 copy of the tables)
        scene = self._scene(False)
        if self._act is None or self._act.numel() < n * s * ops.act_floats(scene):
            self._act = torch.empty(n * s * ops.act_floats(scene), dtype=torch.float32, device=dev)
        if self._scratch is None or self._scratch_for != (n, s):
            self._scratch, self._scratch_for = ops.backward_scratch(n, s, dev), (n, s)
        if self.single_launch_render and (self._ray_counters is None or self._ray_counters.numel() < n):
            self._ray_counters = torch.zeros(n, dtype=torch.int32, device=dev)
        out = ops.render_forward(scene, rcfg, rays_o, rays_d, z_vals, target_rgb, td, acc, want_aux=False, act=self._act,
                                 ray_counters=self._ray_counters if self.single_launch_render else None,
                                 loss_grads=self.loss_grads, scratch=self._scratch)
        if self.world_size > 1 or self.losses_on_side:
            # the loss VALUES (and, several GPUs, the global sums they need: slots 6..) are not on the gradient path: their
            # reduction and finalisation run on a side stream, forked here -- before the backward kernel is queued -- and joined
            # in optimizer_step.  With ``losses_on_side`` (PipelinedMapper) the optimiser's step counter is advanced there too.
            main = torch.cuda.current_stream()
            if self._loss_stream is None:
                self._loss_stream = torch.cuda.Stream(device=dev)
            self._loss_stream.wait_stream(main)
            with torch.cuda.stream(self._loss_stream):
                if self.world_size > 1:
                    self._small_allreduce('sums', acc[6:], 1)
                losses = ops.loss_finalize(rcfg, s, acc)
                if self.losses_on_side and not self._tick_issued:
                    _lib.check(_lib.lib().mne_adam_tick(self.step_dev.data_ptr(), ops._stream()), 'mne_adam_tick')
                    self._tick_issued = True
            self._loss_pending = True
        else:
            losses = ops.loss_finalize(rcfg, s, acc)
        ops.render_backward_points(self._scene(True, weight_grads=False), rays_o, rays_d, z_vals, self._scratch, act=self._act)
        self._pending_wgrad = (n, s)
        if not defer_wgrad or self.flat_param is None or (self.world_size > 1 and self.peers is None):
            self._decoder_wgrad()     # (NCCL fallback: the reduce-scatter moves the whole flat buffer, nothing to overlap)
        self.last_render = out
        return losses

    def _small_allreduce(self, name, tensor, channel):
        """In-place sum of a small device vector over the ranks, on the current stream: through the peer-memory exchange buffer
        (scatter to every rank's slot, signal-pad barrier, local sum, barrier) -- or NCCL when peer memory is unavailable."""
        if self.peers is None:
            torch.distributed.all_reduce(tensor, group=self.group)
            return
        lib, st, pr = _lib.lib(), ops._stream(), self.peers
        off, n_f = self._x_regions[name]
        rank = self.shards.rank
        is_double = tensor.dtype == torch.float64
        assert tensor.is_contiguous() and tensor.numel() * (2 if is_double else 1) == n_f
        _lib.check(lib.mne_peer_scatter(tensor.data_ptr(), n_f, pr['x_ptrs'], off + rank * n_f, self.world_size, st), 'mne_peer_scatter')
        self._barrier(channel)
        _lib.check(lib.mne_peer_sum(pr['xchg'].data_ptr() + 4 * off, tensor.numel(), self.world_size, tensor.data_ptr(),
                                    1 if is_double else 0, st), 'mne_peer_sum')
        self._barrier(channel)     # nobody overwrites a slot that a slower rank is still summing

    def _barrier(self, channel):
        """Cross-rank barrier on the current stream: the library's own kernel over a signal pad in peer memory (stores to the
        peers, polls of local memory), or torch's symmetric-memory barrier (MNE_OWN_BARRIER=0)."""
        pr = self.peers
        if self._own_barrier:
            _lib.check(_lib.lib().mne_peer_barrier(pr['pad_ptrs'], self.world_size, self.shards.rank, channel, N_BARRIER_CHANNELS,
                                                   ops._stream()), 'mne_peer_barrier')
        elif channel < 3:
            pr['hx'].barrier(channel=channel)
        elif channel == 3:
            pr['hg'].barrier(channel=0)
        else:
            pr['hp'].barrier(channel=1)

    def barrier_timeouts(self):
        """Non-zero if one of the library's barriers gave up waiting for a rank (the error word of the signal pad)."""
        return 0 if self.peers is None else int(self.peers['pad'][-1].item())

    def _decoder_wgrad(self, max_ctas=0):
        if self._pending_wgrad is None:
            return
        n, s = self._pending_wgrad
        self._pending_wgrad = None
        ops.decoder_wgrad(self._scene(True, plane_grads=False), n, s, self._scratch, self._act, max_ctas)

    def optimizer_step(self):
        """Adam + zero_grad (``_optimizer_step``), then the join with the side stream that finalises the losses."""
        self._optimizer_step()
        if self._loss_pending:
            torch.cuda.current_stream().wait_stream(self._loss_stream)
            self._loss_pending = False

    def _optimizer_step(self):
        """Adam + zero_grad.  One GPU: one launch over the flat buffers (one per tensor when the parameters were not
        adopted).  Several GPUs: flags of the non-zero gradient texels handed to their owners, barrier, ONE kernel over
        NVLink peer memory (reduce the flagged gradients, Adam on the owned groups, parameters written back to every
        rank), barrier -- or, without peer memory, NCCL reduce-scatter / Adam on the slice / all-gather
        (``mneslam_b200.distributed``)."""
        lib = _lib.lib()
        st = ops._stream()
        if self._tick_issued:        # the step counter was advanced on the loss stream, under the backward kernel
            torch.cuda.current_stream().wait_stream(self._loss_stream)
            self._tick_issued = False
        else:
            _lib.check(lib.mne_adam_tick(self.step_dev.data_ptr(), st), 'mne_adam_tick')
        if self.world_size > 1 and self.peers is not None:
            # The decoder (6 208 weights) stays out of the peer exchange: on a side stream its weight gradients are computed,
            # all-reduced (25 KB) and applied by every rank's own Adam step (identical sums and moments -> identical weights),
            # all of it beside the exchange of the planes below.
            main = torch.cuda.current_stream()
            if self._wgrad_stream is None:
                self._wgrad_stream = torch.cuda.Stream(device=self.dev)
            side = self._wgrad_stream
            side.wait_stream(main)
            with torch.cuda.stream(side):
                self._decoder_wgrad()
                tail = slice(self.dense_from, self.total)
                self._small_allreduce('dec', self.flat_grad[tail], 2)
                self._adam_segments(self.flat_param[tail], self.flat_grad[tail], self.dec_m, self.dec_v, self._seg_tail, True, touched=None)
            n_seg, ends, lr, eps, wd = self._seg_head          # global segment ends of the planes / tables; ownership is block-cyclic
            pr = self.peers
            if self.dirty_local is not None and self.push_gradients:
                # flags set by the backward kernels: only the flagged texels are read, pushed to their owners and cleared
                _lib.check(lib.mne_dirty_push_data(self.dirty_local.data_ptr(), self.flat_grad.data_ptr(), self.flat_grad.numel(),
                                                   self.dense_from, PEER_CHUNK, self.world_size, self.shards.rank,
                                                   pr['dirty_ptrs'], pr['inbox_ptrs'], st), 'mne_dirty_push_data')
            elif self.dirty_local is not None:        # flags set by the backward kernels: hand them to the owners
                _lib.check(lib.mne_dirty_push(self.dirty_local.data_ptr(), self.flat_grad.numel(), self.dense_from, PEER_CHUNK,
                                              self.world_size, self.shards.rank, pr['dirty_ptrs'], st), 'mne_dirty_push')
            elif self.push_gradients:                 # one pass over the gradient buffer: flags AND the non-zero texels go to their owners
                _lib.check(lib.mne_grad_push(self.flat_grad.data_ptr(), self.flat_grad.numel(), self.dense_from, PEER_CHUNK, self.world_size,
                                             self.shards.rank, pr['dirty_ptrs'], pr['inbox_ptrs'], st), 'mne_grad_push')
            else:                                     # flags only; the owners fetch the flagged groups themselves
                _lib.check(lib.mne_grad_dirty_map(self.flat_grad.data_ptr(), self.flat_grad.numel(), PEER_CHUNK, self.world_size,
                                                  self.shards.rank, pr['dirty_ptrs'], st), 'mne_grad_dirty_map')
            self._barrier(3)                          # every rank's gradient and dirty map are complete
            _lib.check(lib.mne_adam_step_peers(self.world_size, self.shards.rank, pr['param_ptrs'], pr['grad_ptrs'],
                                               self.flat_grad.numel(), PEER_CHUNK, self.flat_m.data_ptr(), self.flat_v.data_ptr(), n_seg, ends, lr,
                                               eps, wd, self.betas[0], self.betas[1], self.step_dev.data_ptr(), 0,
                                               ops._ptr(self.touched), pr['mc_param'] or None, pr['mc_grad'] or None, pr['dirty'].data_ptr(),
                                               pr['inbox'].data_ptr() if self.push_gradients else None, st),
                       'mne_adam_step_peers')
            self._barrier(4)                          # every rank's slice has landed in every parameter buffer
            main.wait_stream(side)
            return
        if self.world_size > 1:
            from .distributed import sharded_optimizer_step

            def adam_fn(param_slice, grad_slice, _ends, _hp):
                self._adam_segments(param_slice, grad_slice, self.flat_m, self.flat_v, self._local_segments, False)
            sharded_optimizer_step(self.flat_param, self.flat_grad, self.grad_shard, self.shards, adam_fn, self.group)
            return
        if self.flat_param is not None:
            self._adam_split()
            return
        self._decoder_wgrad()
        for (p, lr, eps, wd), g, m, v in zip(self.params, self.grads, self.exp_avg, self.exp_avg_sq):
            _lib.check(lib.mne_adam_step(p.data_ptr(), g.data_ptr(), m.data_ptr(), v.data_ptr(), p.numel(), lr,
                                         self.betas[0], self.betas[1], eps, wd, self.step_dev.data_ptr(), 0, 1, st),
                       'mne_adam_step')

    def _adam_split(self):
        """One GPU: the planes' Adam pass (a pure HBM stream over 129 MB) on the current stream; on a side stream, beside it,
        the decoder's weight gradients (which only the decoder's own Adam step consumes) and that step."""
        dec0 = self.dense_from
        flags = dict(dirty=self.dirty_local, dirty_floats=self._dirty_floats) if self.adam_dirty else {}
        if self.wgrad_inline:
            self._decoder_wgrad()
            self._adam_segments(self.flat_param, self.flat_grad, self.flat_m, self.flat_v, self._segments, True, **flags)
            return
        main = torch.cuda.current_stream()
        if self._wgrad_stream is None:
            self._wgrad_stream = torch.cuda.Stream(device=self.dev)
        side = self._wgrad_stream
        side.wait_stream(main)
        with torch.cuda.stream(side):
            self._decoder_wgrad(max_ctas=self.wgrad_ctas)
            tail = slice(dec0, self.total)
            self._adam_segments(self.flat_param[tail], self.flat_grad[tail], self.flat_m[tail], self.flat_v[tail], self._seg_tail, True,
                                touched=self.touched[dec0 // 4:] if self.touched is not None else None)
        self._adam_segments(self.flat_param, self.flat_grad, self.flat_m, self.flat_v, self._seg_head, True, **flags)
        main.wait_stream(side)

    def step(self, rays_o, rays_d, target_rgb, target_d, noise=None, sampled=None):
        """One mapping iteration.  Returns the eight losses (device tensor, order ``LOSS_KEYS``)."""
        losses = self.forward_backward(rays_o, rays_d, target_rgb, target_d, noise, sampled, defer_wgrad=True)
        self.optimizer_step()
        self.last_losses = losses
        return losses

    def distill_step(self, teacher, rays_o, rays_d):
        """One iteration of the joint distillation of mp_slam/mapper.py:598-642: the teacher (a foreign agent's model)
        renders the rays without depth guidance, the student takes a mapping iteration on those renderings."""
        with torch.no_grad():
            t = teacher.render_rays(rays_o, rays_d, target_d=None)
        return self.step(rays_o, rays_d, t['rgb'].detach(), t['depth'].detach().unsqueeze(-1))

    def total_loss(self, losses):
        """``get_loss_from_ret`` (mneslam_mp.py:350-372): only the selected loss family enters the sum -- a loss with weight
        zero is never touched there, so its value (``nan`` when its mask is empty) must not reach the total."""
        w = self.loss_weights[:7]
        return torch.where(w != 0, losses[:7] * w, torch.zeros_like(w)).sum()

    def losses_dict(self, losses):
        return {k: losses[i] for i, k in enumerate(LOSS_KEYS)}


class DeviceKeyframeRays:
    """Keyframe ray database kept on the device with on-device ray assembly: replaces the CPU
    gather + three H2D copies per iteration of ``KeyFrameDatabase.sample_global_rays`` (model/keyframe.py:91-103)
    and the ray assembly of ``mapping_optimize`` (mp_slam/mapper.py:135-153)."""

    def __init__(self, rays, poses, current=None, min_pixels_cur=0):
        self.rays = rays          # [n_kf, R, 7] (dir_cam, rgb, depth)
        self.poses = poses        # [n_kf, 4, 4] camera-to-world
        self.n_kf, self.per_kf = rays.shape[0], rays.shape[1]
        self.min_pixels_cur = int(min_pixels_cur)
        self.set_current(current)

    def set_current(self, current):
        """The rays of the current frame ([H * W, 7], every pixel: mapper.py:131-132), or None.  With a current frame every
        batch carries ``max(n // n_kf, min_pixels_cur)`` of its pixels after the keyframe rays, posed by the LAST pose (id -1)."""
        if current is not None:
            current = current.reshape(-1, 7)
            if current.dtype != torch.float32 or not current.is_contiguous() or current.device != self.rays.device:
                raise TypeError('the current frame must be a contiguous float32 [H * W, 7] tensor on the device of the store')
        self.current = current

    def n_current(self, n_rays):
        """Rays of the current frame in a batch of ``n_rays`` keyframe rays (mp_slam/mapper.py:137-139)."""
        return 0 if self.current is None else max(n_rays // self.n_kf, self.min_pixels_cur)

    def batch_rays(self, n_rays):
        return n_rays + self.n_current(n_rays)

    def sample(self, n_rays, generator=None, return_rows=False, key=None):
        """``n_rays`` distinct stored rays (``random.sample`` of model/keyframe.py:98: without replacement) followed by
        ``n_current(n_rays)`` distinct pixels of the current frame, drawn and assembled in one launch.  The draw is keyed by one
        device-side random integer, so a CUDA-graph replay draws a fresh batch (``key``: that integer, a one-element int64
        device tensor, for reproducible draws)."""
        dev = self.rays.device
        n_cur = self.n_current(n_rays)
        n = n_rays + n_cur
        if self.rays.dtype != torch.float32 or not self.rays.is_contiguous() or self.poses.dtype != torch.float32 or not self.poses.is_contiguous():
            raise TypeError('rays and poses must be contiguous float32 tensors')
        if key is None:
            key = torch.randint(0, 1 << 62, (1,), device=dev, generator=generator)
        elif key.dtype != torch.int64 or key.device != dev or key.numel() != 1:
            raise TypeError('key must be a one-element int64 tensor on the device of the store')
        rows = torch.empty(n, dtype=torch.int64, device=dev) if return_rows else None
        ro, rd, rgb, d = (torch.empty(n, 3, dtype=torch.float32, device=dev), torch.empty(n, 3, dtype=torch.float32, device=dev),
                          torch.empty(n, 3, dtype=torch.float32, device=dev), torch.empty(n, 1, dtype=torch.float32, device=dev))
        _lib.check(_lib.lib().mne_draw_rays(key.data_ptr(), self.rays.data_ptr(), self.n_kf * self.per_kf, self.per_kf, n_rays,
                                            ops._ptr(self.current), 0 if self.current is None else self.current.shape[0], n_cur,
                                            self.poses.data_ptr(), self.n_kf, ro.data_ptr(), rd.data_ptr(), rgb.data_ptr(), d.data_ptr(),
                                            ops._ptr(rows), ops._stream()), 'mne_draw_rays')
        return (ro, rd, rgb, d, rows) if return_rows else (ro, rd, rgb, d)

    def assemble(self, rays7, kf_ids):
        """rays7 [n,7] and the keyframe of each ray (-1 = the last pose) -> rays_o, rays_d, target_rgb [n,3], target_d [n,1]."""
        return self._assemble(rays7, None, kf_ids, rays7.shape[0])

    def _assemble(self, rays7, idx, kf_ids, n):
        dev = self.rays.device
        if rays7.dtype != torch.float32 or not rays7.is_contiguous() or self.poses.dtype != torch.float32 or not self.poses.is_contiguous():
            raise TypeError('rays and poses must be contiguous float32 tensors')
        if kf_ids is not None and (kf_ids.dtype != torch.int64 or not kf_ids.is_contiguous()):
            kf_ids = kf_ids.to(torch.int64).contiguous()
        ro, rd, rgb, d = (torch.empty(n, 3, dtype=torch.float32, device=dev), torch.empty(n, 3, dtype=torch.float32, device=dev),
                          torch.empty(n, 3, dtype=torch.float32, device=dev), torch.empty(n, 1, dtype=torch.float32, device=dev))
        _lib.check(_lib.lib().mne_assemble_rays(rays7.data_ptr(), ops._ptr(idx), self.per_kf, ops._ptr(kf_ids), self.poses.data_ptr(),
                                                self.n_kf, n, ro.data_ptr(), rd.data_ptr(), rgb.data_ptr(), d.data_ptr(), ops._stream()),
                   'mne_assemble_rays')
        return ro, rd, rgb, d


class HostFedMapper:
    """Mapping iterations fed from HOST memory as ONE CUDA-graph launch each: the ray batch the caller hands over (the
    keyframe database lives on the host in the reference, model/keyframe.py:91-103) is copied into a pinned staging buffer,
    and the graph holds the H2D copies, the on-device ray assembly, the whole iteration (sample, forward, losses,
    backward, optimiser) and the D2H copy of the loss.  ``step`` returns the loss as a Python float, i.e. it includes the
    read-back the reference loop does when it logs the loss."""

    def __init__(self, mapper, db, n_rays):
        self.mapper, self.db, self.n = mapper, db, n_rays
        dev = mapper.dev
        self.h_rays = torch.empty(n_rays, 7, dtype=torch.float32).pin_memory()
        self.h_ids = torch.empty(n_rays, dtype=torch.int64).pin_memory()
        self.h_loss = torch.zeros(1, dtype=torch.float32).pin_memory()
        self.d_rays = torch.empty(n_rays, 7, dtype=torch.float32, device=dev)
        self.d_ids = torch.empty(n_rays, dtype=torch.int64, device=dev)
        self.h_ids.zero_(); self.h_rays.zero_(); self.h_rays[:, 2] = -1.0
        self.graph = None
        self.stream = torch.cuda.Stream(device=dev)

    def _iteration(self):
        self.d_rays.copy_(self.h_rays, non_blocking=True)
        self.d_ids.copy_(self.h_ids, non_blocking=True)
        ro, rd, rgb, d = self.db.assemble(self.d_rays, self.d_ids)
        losses = self.mapper.step(ro, rd, rgb, d)
        self.h_loss.copy_(self.mapper.total_loss(losses).reshape(1), non_blocking=True)

    def capture(self, rays7, kf_ids):
        """Warm up on a real batch (kernel attributes, workspaces; these are real optimiser steps) and capture the graph."""
        self.h_rays.copy_(rays7); self.h_ids.copy_(kf_ids)
        self.stream.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(self.stream):
            for _ in range(2):
                self._iteration()
        self.stream.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph, stream=self.stream):
            self._iteration()
        return self

    def step(self, rays7, kf_ids):
        """rays7 [n,7] (camera-frame direction, rgb, depth) and kf_ids [n] int64: HOST tensors (any memory)."""
        if self.graph is None:
            self.capture(rays7, kf_ids)
        self.h_rays.copy_(rays7)
        self.h_ids.copy_(kf_ids)
        with torch.cuda.stream(self.stream):        # replay() launches on the CURRENT stream
            self.graph.replay()
        self.stream.synchronize()
        return float(self.h_loss[0])


class PipelinedMapper:
    """Mapping iterations on a device-resident keyframe database as ONE CUDA-graph launch each, software-pipelined: the
    part of iteration t+1 that does not read the map -- drawing the ray batch, assembling the rays, sampling along them --
    runs on a second stream while the optimiser of iteration t (a pure HBM stream) is in flight.  Every iteration does
    exactly the work of ``FusedMapper.step(*db.sample(n))``; only the order across iterations changes."""

    def __init__(self, mapper, db, n_rays, spare_sms=8):
        self.mapper, self.db, self.n = mapper, db, n_rays
        dev = mapper.dev
        # the optimiser kernels are persistent grids over every SM: keep a few SMs out of them, or the side stream's
        # small kernels only start when the optimiser has finished.  The option is library-wide and read at LAUNCH time, i.e.
        # when the graph is captured: it is put back in close() / when this object goes away
        import ctypes
        prev = ctypes.c_int(0)
        _lib.check(_lib.lib().mne_get_option(b'adam_spare_sms', ctypes.addressof(prev)), 'mne_get_option')
        self._spare_before = prev.value
        _lib.check(_lib.lib().mne_set_option(b'adam_spare_sms', int(spare_sms)), 'mne_set_option')
        n_rays = self._batch_size(n_rays)
        tables = ops.SampleTables.get(mapper.cfg, dev, True)
        s = tables[0].numel() + tables[1].numel()
        f32 = dict(dtype=torch.float32, device=dev)
        # the batch in flight: written at the tail of one replay, read at the head of the next
        self.ro, self.rd, self.rgb = torch.empty(n_rays, 3, **f32), torch.empty(n_rays, 3, **f32), torch.empty(n_rays, 3, **f32)
        self.d = torch.empty(n_rays, 1, **f32)
        self.z = torch.empty(n_rays, s, **f32)
        self.acc = torch.zeros(_lib.ACC_SLOTS, dtype=torch.float64, device=dev)
        self.losses = torch.zeros(_lib.N_LOSSES, **f32)
        self.side = torch.cuda.Stream(device=dev)
        self.graph = None
        self.stream = None

    def _batch_size(self, n_rays):
        return self.db.batch_rays(n_rays)            # + the rays of the current frame, when the store has one

    def close(self):
        """Put the library-wide ``adam_spare_sms`` option back (idempotent)."""
        if self._spare_before is not None:
            _lib.lib().mne_set_option(b'adam_spare_sms', int(self._spare_before))
            self._spare_before = None
            self.mapper.losses_on_side = False

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _draw(self):
        ro, rd, rgb, d = self.db.sample(self.n)
        self.ro.copy_(ro); self.rd.copy_(rd); self.rgb.copy_(rgb); self.d.copy_(d)
        self.mapper.sample_stage(self.d, z_out=self.z, acc_out=self.acc)

    def _iteration(self):
        m = self.mapper
        main = torch.cuda.current_stream()
        # acc (mask counts from the sampling stage; the forward adds the loss sums) is read by the forward on this stream and by
        # the loss finalisation on the mapper's loss stream: the side stream refills it only after both (below)
        m.losses_on_side = True
        losses = m.forward_backward(self.ro, self.rd, self.rgb, self.d, sampled=(self.z, self.acc), defer_wgrad=True)
        with torch.cuda.stream(m._loss_stream):      # (the losses are finalised there, beside the backward kernel)
            self.losses.copy_(losses)
            self._losses_ready()
        self.side.wait_stream(main)                 # the batch has been consumed: refill it under the optimiser
        self.side.wait_stream(m._loss_stream)
        with torch.cuda.stream(self.side):
            self._after_losses()
            self._draw()
        m.optimizer_step()
        main.wait_stream(self.side)
        m.last_losses = self.losses

    def _losses_ready(self):
        """Hook on the loss stream, once the iteration's losses exist (subclasses queue their read-back here)."""

    def _after_losses(self):
        """Hook on the side stream, once the iteration's losses exist (subclasses copy them to the host here)."""

    def capture(self):
        stream = self.stream if self.stream is not None else torch.cuda.Stream(device=self.mapper.dev)
        stream.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(stream):
            self._draw()                             # prologue: the first batch
            for _ in range(2):
                self._iteration()
        stream.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph, stream=stream):
            self._iteration()
        self.stream = stream
        return self

    def step(self):
        """One iteration; returns the eight losses (device tensor, valid once the launch stream has run)."""
        if self.graph is None:
            self.capture()
        self.graph.replay()
        return self.losses


class PrefetchingHostMapper(PipelinedMapper):
    """``PipelinedMapper`` fed from HOST memory, data-loader style: every ``step(next_batch)`` trains on the batch handed over
    by the PREVIOUS call and, under its optimiser, uploads, assembles and samples ``next_batch`` (one CUDA-graph launch
    holding the H2D copies, the iteration and the D2H copy of the loss).  The ray batches of the reference come from the
    host keyframe store (model/keyframe.py:91-103) and do not depend on the map, so the caller can always be one ahead.

    ``async_loss``: the loss is read back one call late as well -- ``step`` launches its iteration and returns the loss of
    the iteration launched by the previous call (None on the first call; ``flush()`` returns the last one).  The host then
    never waits for the launch it has just made, so the device runs the iterations back to back (the reference loop itself
    never reads the loss between ``backward()`` and ``step()``, mp_slam/mapper.py:155-162).  Two graphs with their own
    pinned staging and loss buffers alternate, so that a buffer is only rewritten after its graph has finished."""

    def __init__(self, mapper, db, n_rays, spare_sms=8, async_loss=False):
        super().__init__(mapper, db, n_rays, spare_sms)
        dev = mapper.dev
        self.async_loss = bool(async_loss)
        self.slots = []
        for _ in range(2 if async_loss else 1):
            self.slots.append({'h_rays': torch.empty(n_rays, 7, dtype=torch.float32).pin_memory(),
                               'h_ids': torch.empty(n_rays, dtype=torch.int64).pin_memory(),
                               'h_loss': torch.zeros(1, dtype=torch.float32).pin_memory(),
                               'graph': None, 'done': torch.cuda.Event()})
        self._use(0)
        self.d_rays = torch.empty(n_rays, 7, dtype=torch.float32, device=dev)
        self.d_ids = torch.empty(n_rays, dtype=torch.int64, device=dev)
        self._calls, self._in_flight = 0, None

    def _use(self, i):
        slot = self.slots[i]
        self.h_rays, self.h_ids, self.h_loss = slot['h_rays'], slot['h_ids'], slot['h_loss']
        return slot

    def _batch_size(self, n_rays):
        return n_rays                                # the caller hands over whole batches (current-frame rays included)

    def _draw(self):
        self.d_rays.copy_(self.h_rays, non_blocking=True)
        self.d_ids.copy_(self.h_ids, non_blocking=True)
        ro, rd, rgb, d = self.db.assemble(self.d_rays, self.d_ids)
        self.ro.copy_(ro); self.rd.copy_(rd); self.rgb.copy_(rgb); self.d.copy_(d)
        self.mapper.sample_stage(self.d, z_out=self.z, acc_out=self.acc)

    def _losses_ready(self):
        # the weighted sum of the losses and its D2H copy, on the loss stream beside the backward kernel and the optimiser.  (On the
        # side stream that uploads and samples the next batch they measured 25 us SLOWER per iteration -- that work then outlasts
        # the optimiser; at the end of the main stream they sit on the critical path.)
        self.h_loss.copy_(self.mapper.total_loss(self.losses).reshape(1), non_blocking=True)

    def prime(self, rays7, kf_ids):
        """Hand over the first batch (uploaded and sampled right away; also warms up and captures the graph(s), which costs
        two real iterations on this batch per graph)."""
        for i in range(len(self.slots)):
            slot = self._use(i)
            self.h_rays.copy_(rays7); self.h_ids.copy_(kf_ids)
            self.graph = None
            self.capture()
            slot['graph'] = self.graph
        return self

    def step(self, next_rays7, next_kf_ids):
        """Train on the batch of the previous call (or of ``prime``); upload ``next_*`` (HOST tensors) under its optimiser.
        Returns the loss of the trained batch as a Python float -- with ``async_loss`` the loss of the previous call's
        iteration (None on the first call)."""
        if self.graph is None:
            raise RuntimeError('PrefetchingHostMapper: call prime(first_batch) first')
        slot = self.slots[self._calls % len(self.slots)]
        self._calls += 1
        slot['h_rays'].copy_(next_rays7)
        slot['h_ids'].copy_(next_kf_ids)
        with torch.cuda.stream(self.stream):
            slot['graph'].replay()
            slot['done'].record()
        if not self.async_loss:
            slot['done'].synchronize()
            return float(slot['h_loss'][0])
        before, self._in_flight = self._in_flight, slot
        if before is None:
            return None
        before['done'].synchronize()
        return float(before['h_loss'][0])

    def flush(self):
        """``async_loss``: wait for the iteration still in flight and return its loss (None if there is none)."""
        slot, self._in_flight = self._in_flight, None
        if slot is None:
            return None
        slot['done'].synchronize()
        return float(slot['h_loss'][0])
