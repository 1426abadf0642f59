"""Render-based pose optimisation on a frozen map (BASELINE.json config 3).

The reference's only live render-based pose optimisation is the loop-closure alignment of
``mp_slam/mapper.py:364-410``: pose parameters from ``get_pose_param_optim`` (``mneslam_mp.py:577-584``),
``matrix_from_tensor`` -> rays -> ``render_rays`` -> losses -> ``backward`` -> Adam on the pose only.  This class
is that loop with the map's gradients switched off, so the fused backward runs its ray-gradient-only variant
(no plane scatter, no weight gradients), and with the sensor depth guiding the samples as in
``JointEncoding.forward`` (``PoseArray``-style per-frame tracking).
"""
import torch

from . import _lib, ops
from . import pose as pose_lib
from .mapping import loss_weight_vector


def matrix_to_quaternion(R):
    """Rotation matrix -> real-first unit quaternion (single matrix, stable branch selection)."""
    m = R.double()
    t = m.trace()
    if t > 0:
        s = torch.sqrt(t + 1.0) * 2
        q = torch.stack([0.25 * s, (m[2, 1] - m[1, 2]) / s, (m[0, 2] - m[2, 0]) / s, (m[1, 0] - m[0, 1]) / s])
    else:
        i = int(torch.argmax(torch.diagonal(m)))
        j, k = (i + 1) % 3, (i + 2) % 3
        s = torch.sqrt(1.0 + m[i, i] - m[j, j] - m[k, k]) * 2
        q = torch.zeros(4, dtype=torch.float64, device=R.device)
        q[0] = (m[k, j] - m[j, k]) / s
        q[1 + i] = 0.25 * s
        q[1 + j] = (m[j, i] + m[i, j]) / s
        q[1 + k] = (m[k, i] + m[i, k]) / s
    return q.to(R.dtype)


def matrix_to_axis_angle(R):
    q = matrix_to_quaternion(R).double()
    n = torch.linalg.norm(q[1:])
    ang = 2.0 * torch.atan2(n, q[0])
    axis = q[1:] / n if float(n) > 1e-12 else torch.tensor([1.0, 0.0, 0.0], dtype=torch.float64, device=R.device)
    if float(ang) < 1e-8:         # Rodrigues divides by the angle: keep it away from exactly zero
        ang = torch.tensor(1e-8, dtype=torch.float64, device=R.device)
    return (axis * ang).to(R.dtype)


class frozen_map:
    """Context manager: switch off ``requires_grad`` of the planes and the decoder so that the fused backward
    runs its ray-gradient-only variant (no plane scatter, no weight gradients)."""

    def __init__(self, model):
        self.params = [p for grp in model.all_planes for p in grp] + list(model.decoder.parameters())

    def __enter__(self):
        self.saved = [p.requires_grad for p in self.params]
        for p in self.params:
            p.requires_grad_(False)

    def __exit__(self, *exc):
        for p, r in zip(self.params, self.saved):
            p.requires_grad_(r)


class PoseTracker:
    def __init__(self, model, c2w_init, rot_rep=None, lr_rot=1e-3, lr_trans=1e-3, is_co_sdf=None):
        self.model, self.cfg = model, model.config
        self.rot_rep = rot_rep or self.cfg['training']['rot_rep']
        dev = c2w_init.device
        R = c2w_init[:3, :3]
        rot = matrix_to_axis_angle(R) if self.rot_rep == 'axis_angle' else matrix_to_quaternion(R)
        self.rot = torch.nn.Parameter(rot[None].clone())
        self.trans = torch.nn.Parameter(c2w_init[None, :3, 3].clone())
        self.optimizer = torch.optim.Adam([{'params': [self.rot], 'lr': lr_rot}, {'params': [self.trans], 'lr': lr_trans}],
                                          capturable=True)
        self.weights = torch.tensor(loss_weight_vector(self.cfg, is_co_sdf)[:7], device=dev)

    def c2w(self):
        return pose_lib.transform_matrix(self.rot, self.trans, self.rot_rep)[0]

    def step(self, dirs_cam, target_rgb, target_d, noise=None):
        """One pose iteration on a frozen map.  Returns the weighted loss (device scalar)."""
        model = self.model
        was_training = model.training
        model.train()
        self.optimizer.zero_grad(set_to_none=True)
        with frozen_map(model):
            rays_o, rays_d = pose_lib.rays_from_c2w(self.c2w(), dirs_cam)
            ret = model.forward(rays_o.contiguous(), rays_d.contiguous(), target_rgb, target_d, noise=noise)
            losses = torch.stack([ret[k] for k in ('rgb_loss', 'depth_loss', 'co_sdf_loss', 'co_fs_loss', 'e_fs_loss',
                                                   'e_center_loss', 'e_tail_loss')])
            loss = (losses * self.weights).sum()
            loss.backward()
        self.optimizer.step()
        model.train(was_training)
        return loss.detach()


class FusedTracker:
    """``PoseTracker`` without autograd and without torch's optimiser: pose -> rays (``mne_pose_rays``), sample / forward /
    losses / ray-gradient backward on the frozen map, then ``mne_pose_step`` (ray gradients -> pose gradient through the
    parameterisation -> Adam on the 6 or 7 pose parameters).  Nine launches per iteration instead of about a hundred
    element-wise torch kernels; same update (tested against ``PoseTracker``)."""

    def __init__(self, model, c2w_init, rot_rep=None, lr_rot=1e-3, lr_trans=1e-3, is_co_sdf=None,
                 betas=(0.9, 0.999), eps=1e-8):
        ops.require_cuda()
        self.model, self.cfg = model, model.config
        self.rot_rep = rot_rep or self.cfg['training']['rot_rep']
        dev = c2w_init.device
        R = c2w_init[:3, :3]
        rot = matrix_to_axis_angle(R) if self.rot_rep == 'axis_angle' else matrix_to_quaternion(R)
        self.rot = rot.detach().to(torch.float32).contiguous().clone()          # [3] or [4]
        self.trans = c2w_init[:3, 3].detach().to(torch.float32).contiguous().clone()
        n_par = self.rot.numel() + 3
        self.exp_avg = torch.zeros(n_par, dtype=torch.float32, device=dev)
        self.exp_avg_sq = torch.zeros(n_par, dtype=torch.float32, device=dev)
        self.step_dev = torch.zeros(1, dtype=torch.int64, device=dev)
        self.pose_grad = torch.zeros(n_par, dtype=torch.float32, device=dev)
        self.lr_rot, self.lr_trans, self.betas, self.eps = lr_rot, lr_trans, betas, eps
        w = loss_weight_vector(self.cfg, is_co_sdf)
        self.loss_grads = torch.tensor(w, dtype=torch.float32, device=dev)
        self.weights = self.loss_grads[:7].clone()
        self._act = self._scratch = None
        self._for = None

    def c2w(self):
        return pose_lib.transform_matrix(self.rot[None], self.trans[None], self.rot_rep)[0]

    def step(self, dirs_cam, target_rgb, target_d, noise=None):
        """One pose iteration on the frozen map.  Returns the weighted loss (device scalar)."""
        model, cfg = self.model, self.cfg
        dev = self.rot.device
        n = dirs_cam.shape[0]
        lib, st = _lib.lib(), ops._stream()
        rep = 0 if self.rot_rep == 'axis_angle' else 1
        dirs = ops._f32(dirs_cam, 'dirs_cam')
        rays_o = torch.empty(n, 3, dtype=torch.float32, device=dev)
        rays_d = torch.empty(n, 3, dtype=torch.float32, device=dev)
        _lib.check(lib.mne_pose_rays(rep, self.rot.data_ptr(), self.trans.data_ptr(), dirs.data_ptr(), n, rays_o.data_ptr(),
                                     rays_d.data_ptr(), st), 'mne_pose_rays')
        tables = ops.SampleTables.get(cfg, dev, True)
        s = tables[0].numel() + tables[1].numel()
        perturb = cfg['training']['perturb'] > 0.
        if perturb and noise is None:
            noise = torch.rand(n, s, device=dev)
        rcfg = ops.make_render_cfg(cfg, norm_rays=n)
        td = target_d.reshape(-1)
        acc = ops.new_acc(dev)
        z_vals = ops.sample_rays(ops.make_sample_cfg(tables, perturb), rcfg, n, s, td, noise if perturb else None, acc, dev)
        if self._for != (n, s):
            self._act = torch.empty(n * s * _lib.ACT_FLOATS_MAX, dtype=torch.float32, device=dev)
            self._scratch = ops.backward_scratch(n, s, dev)
            self._for = (n, s)
        scene = ops.make_scene(model._flat_planes(), [w.detach() for w in model.decoder.weight_list()], model.bound,
                               model.bounding_box, cfg['pos']['n_bins'], hash_meta=model.hash_meta)
        out = ops.render_forward(scene, rcfg, rays_o, rays_d, z_vals, target_rgb, td, acc, want_aux=False, act=self._act)
        losses = ops.loss_finalize(rcfg, s, acc)
        g_o, g_d = ops.render_backward(scene, rcfg, rays_o, rays_d, z_vals, out['raw'], target_rgb, td, acc,
                                       loss_grads=self.loss_grads, want_ray_grads=True, scratch=self._scratch, act=self._act)
        _lib.check(lib.mne_pose_step(rep, self.rot.data_ptr(), self.trans.data_ptr(), dirs.data_ptr(), g_o.data_ptr(),
                                     g_d.data_ptr(), n, self.exp_avg.data_ptr(), self.exp_avg_sq.data_ptr(),
                                     self.step_dev.data_ptr(), self.lr_rot, self.lr_trans, self.betas[0], self.betas[1],
                                     self.eps, self.pose_grad.data_ptr(), st), 'mne_pose_step')
        self.last_render = out
        return (losses[:7] * self.weights).sum()


    def step_align(self, dirs_cam, target_rgb, target_depth, noise=None):
        """One iteration of the loop-closure alignment (mp_slam/mapper.py:388-410): rays from the current pose estimate,
        ``render_rays(..., target_d=None)`` (uniform samples), ``rgb_weight * mse(rgb) + depth_weight * mse(depth)`` against
        the teacher's rendering, pose update.  Returns the loss (device scalar)."""
        model, cfg = self.model, self.cfg
        dev = self.rot.device
        n = dirs_cam.shape[0]
        lib, st = _lib.lib(), ops._stream()
        rep = 0 if self.rot_rep == 'axis_angle' else 1
        dirs = ops._f32(dirs_cam, 'dirs_cam')
        rays_o = torch.empty(n, 3, dtype=torch.float32, device=dev)
        rays_d = torch.empty(n, 3, dtype=torch.float32, device=dev)
        _lib.check(lib.mne_pose_rays(rep, self.rot.data_ptr(), self.trans.data_ptr(), dirs.data_ptr(), n, rays_o.data_ptr(),
                                     rays_d.data_ptr(), st), 'mne_pose_rays')
        tables = ops.SampleTables.get(cfg, dev, False)
        s = tables[0].numel() + tables[1].numel()
        perturb = cfg['training']['perturb'] > 0.
        if perturb and noise is None:
            noise = torch.rand(n, s, device=dev)
        rcfg = ops.make_render_cfg(cfg, norm_rays=n)
        z_vals = ops.sample_rays(ops.make_sample_cfg(tables, perturb), rcfg, n, s, None, noise if perturb else None, None, dev)
        if self._for != (n, s):
            self._act = torch.empty(n * s * _lib.ACT_FLOATS_MAX, dtype=torch.float32, device=dev)
            self._scratch = ops.backward_scratch(n, s, dev)
            self._for = (n, s)
        scene = ops.make_scene(model._flat_planes(), [w.detach() for w in model.decoder.weight_list()], model.bound,
                               model.bounding_box, cfg['pos']['n_bins'], hash_meta=model.hash_meta)
        out = ops.render_forward(scene, rcfg, rays_o, rays_d, z_vals, want_aux=False, act=self._act)
        w_c, w_d = cfg['training']['rgb_weight'], cfg['training']['depth_weight']
        e_c, e_d = out['rgb'] - target_rgb, out['depth'] - target_depth.reshape(-1)
        g_rgb, g_depth = e_c * (2.0 * w_c / (3 * n)), e_d * (2.0 * w_d / n)          # d loss / d rendered maps (F.mse_loss means)
        g_o, g_d = ops.render_backward(scene, rcfg, rays_o, rays_d, z_vals, out['raw'], g_rgb=g_rgb, g_depth=g_depth,
                                       want_ray_grads=True, scratch=self._scratch, act=self._act)
        _lib.check(lib.mne_pose_step(rep, self.rot.data_ptr(), self.trans.data_ptr(), dirs.data_ptr(), g_o.data_ptr(),
                                     g_d.data_ptr(), n, self.exp_avg.data_ptr(), self.exp_avg_sq.data_ptr(),
                                     self.step_dev.data_ptr(), self.lr_rot, self.lr_trans, self.betas[0], self.betas[1],
                                     self.eps, self.pose_grad.data_ptr(), st), 'mne_pose_step')
        self.last_render = out
        return w_c * (e_c * e_c).mean() + w_d * (e_d * e_d).mean()


class GraphedPoseTracker:
    """One pose iteration (pose -> rays -> fused forward -> losses -> ray-gradient backward -> Adam) captured in a
    CUDA graph for a fixed frame: the per-frame optimisation is ``iters`` replays with no Python in between."""

    def __init__(self, tracker, dirs_cam, target_rgb, target_d):
        self.tracker = tracker
        self.dirs, self.rgb, self.d = dirs_cam.clone(), target_rgb.clone(), target_d.clone()
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(3):                                  # warm-up outside capture (allocations, kernel attributes)
                tracker.step(self.dirs, self.rgb, self.d)
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        if hasattr(tracker, 'optimizer'):
            tracker.optimizer.zero_grad(set_to_none=True)
        with torch.cuda.graph(self.graph):
            self.loss = tracker.step(self.dirs, self.rgb, self.d)

    def set_frame(self, dirs_cam, target_rgb, target_d):
        self.dirs.copy_(dirs_cam); self.rgb.copy_(target_rgb); self.d.copy_(target_d)

    def run(self, iters):
        for _ in range(iters):
            self.graph.replay()
        return self.loss


def ate_rmse(est_xyz, gt_xyz):
    """Absolute trajectory error after the closed-form rigid alignment of Horn (the definition the reference's
    ``tools/eval_ate.py:67-101`` uses): RMSE of ``R est + t - gt``."""
    e, g = est_xyz.double().t(), gt_xyz.double().t()          # [3, n]
    ec, gc = e - e.mean(1, keepdim=True), g - g.mean(1, keepdim=True)
    W = ec @ gc.t()
    U, _, Vh = torch.linalg.svd(W.t())
    S = torch.eye(3, dtype=torch.float64)
    if torch.det(U) * torch.det(Vh) < 0:
        S[2, 2] = -1
    R = U @ S @ Vh
    t = g.mean(1, keepdim=True) - R @ e.mean(1, keepdim=True)
    err = R @ e + t - g
    return float(torch.sqrt((err * err).sum(0).mean()))
