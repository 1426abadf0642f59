"""Render-based pose optimisation on a frozen map (BASELINE.json config 3).

The reference's only live render-based pose optimisation is the loop-closure alignment of
``mp_slam/mapper.py:364-410``: pose parameters from ``get_pose_param_optim`` (``mneslam_mp.py:577-584``),
``matrix_from_tensor`` -> rays -> ``render_rays`` -> losses -> ``backward`` -> Adam on the pose only.  This class
is that loop with the map's gradients switched off, so the fused backward runs its ray-gradient-only variant
(no plane scatter, no weight gradients), and with the sensor depth guiding the samples as in
``JointEncoding.forward`` (``PoseArray``-style per-frame tracking).
"""
import torch

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
