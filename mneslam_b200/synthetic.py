"""Seeded synthetic Replica/ScanNet-shaped RGB-D ray batches.

There is no dataset on the box, so every test and benchmark draws its rays from here.  The ray
convention is the reference's OpenGL one (``datasets/utils.py:24-57``): camera-frame directions
``((i-cx)/fx, -(j-cy)/fy, -1)`` (not normalised, so ``z_vals`` are depths along the optical axis),
intrinsics floor-divided by ``data.downsample`` and crop-shifted (``datasets/dataset.py:41-44``),
``rays_d = sum_k dirs_k * R[:, k]`` and ``rays_o = t`` (``mp_slam/mapper.py:82-83``).
"""
import math

import torch


def intrinsics(cfg):
    cam, ds = cfg['cam'], cfg['data']['downsample']
    crop = cam.get('crop_edge', 0)
    H, W = cam['H'] // ds - 2 * crop, cam['W'] // ds - 2 * crop
    fx, fy = cam['fx'] // ds, cam['fy'] // ds
    cx, cy = cam['cx'] // ds - crop, cam['cy'] // ds - crop
    return H, W, float(fx), float(fy), float(cx), float(cy)


def small_rotation(gen, max_angle=0.2):
    """Rodrigues rotation about a random axis by a small random angle (float32)."""
    axis = torch.randn(3, generator=gen)
    axis = axis / axis.norm()
    ang = (torch.rand(1, generator=gen).item() * 2 - 1) * max_angle
    K = torch.tensor([[0, -axis[2], axis[1]], [axis[2], 0, -axis[0]], [-axis[1], axis[0], 0]])
    return torch.eye(3) + math.sin(ang) * K + (1 - math.cos(ang)) * (K @ K)


def make_pose(cfg, gen, frac=(0.5, 0.46, 0.55)):
    """A camera-to-world matrix whose translation lies inside ``mapping.bound``."""
    b = torch.tensor(cfg['mapping']['bound'], dtype=torch.float32)
    t = b[:, 0] + torch.tensor(frac) * (b[:, 1] - b[:, 0])
    c2w = torch.eye(4)
    c2w[:3, :3] = small_rotation(gen)
    c2w[:3, 3] = t
    return c2w


def make_ray_batch(cfg, n_rays, seed=0, zero_depth_frac=0.05, depth_range=(0.5, 3.5), c2w=None):
    """Returns a dict of float32 CPU tensors: rays_o [N,3], rays_d [N,3], target_rgb [N,3],
    target_d [N,1], dirs_cam [N,3], c2w [4,4]."""
    gen = torch.Generator().manual_seed(seed)
    H, W, fx, fy, cx, cy = intrinsics(cfg)
    pix = torch.randperm(H * W, generator=gen)[:n_rays]
    if pix.numel() < n_rays:   # more rays than pixels: sample with replacement
        pix = torch.randint(0, H * W, (n_rays,), generator=gen)
    j, i = (pix // W).float(), (pix % W).float()
    dirs = torch.stack([(i - cx) / fx, -(j - cy) / fy, -torch.ones_like(i)], -1)
    if c2w is None:
        c2w = make_pose(cfg, gen)
    rays_d = torch.sum(dirs[..., None, :] * c2w[:3, :3], -1)
    rays_o = c2w[:3, 3].expand(n_rays, 3).contiguous()
    depth = depth_range[0] + (depth_range[1] - depth_range[0]) * torch.rand(n_rays, 1, generator=gen)
    depth[torch.rand(n_rays, 1, generator=gen) < zero_depth_frac] = 0.0
    rgb = torch.rand(n_rays, 3, generator=gen)
    return {'rays_o': rays_o, 'rays_d': rays_d.contiguous(), 'target_rgb': rgb, 'target_d': depth,
            'dirs_cam': dirs, 'c2w': c2w}


def n_samples_total(cfg, with_depth=True):
    t = cfg['training']
    if with_depth:
        return t['n_samples_d'] + t['n_range_d']
    return t['n_samples']


def make_noise(n_rays, n_samples, seed=0):
    """Stratified-perturbation noise U[0,1) [N,S]; the reference draws it with ``torch.rand`` on the
    CPU (``model/scene_rep.py:381``) so parity runs feed the same tensor to both sides."""
    gen = torch.Generator().manual_seed(seed + 7919)
    return torch.rand(n_rays, n_samples, generator=gen)


def make_keyframe_db(cfg, n_kf=20, seed=0):
    """Synthetic keyframe ray database [n_kf, rays_per_kf, 7] = (dir_cam, rgb, depth) with one pose
    per keyframe along a smooth curve (shape of ``model/keyframe.py`` ``KeyFrameDatabase.rays``)."""
    gen = torch.Generator().manual_seed(seed + 101)
    H, W, fx, fy, cx, cy = intrinsics(cfg)
    per_kf = int(H * W * cfg['mapping']['n_pixels'])
    rays = torch.empty(n_kf, per_kf, 7)
    poses = torch.empty(n_kf, 4, 4)
    for k in range(n_kf):
        a = k / max(n_kf - 1, 1)
        poses[k] = make_pose(cfg, gen, frac=(0.35 + 0.3 * a, 0.4 + 0.1 * math.sin(3 * a), 0.5))
        b = make_ray_batch(cfg, per_kf, seed=seed + 1000 + k, c2w=poses[k])
        rays[k, :, 0:3] = b['dirs_cam']
        rays[k, :, 3:6] = b['target_rgb']
        rays[k, :, 6:7] = b['target_d']
    return rays, poses


def make_current_frame(cfg, c2w, seed=0):
    """Every pixel of a synthetic current frame as [H * W, 7] = (dir_cam, rgb, depth) rows -- what mapping_optimize
    (mp_slam/mapper.py:131-132) builds from the tracked frame before it draws min_pixels_cur of them."""
    H, W = intrinsics(cfg)[:2]
    b = make_ray_batch(cfg, H * W, seed=seed + 5000, c2w=c2w)
    return torch.cat([b['dirs_cam'], b['target_rgb'], b['target_d']], -1).contiguous()


def render_box_scene(cfg, c2w, dirs_cam, margin=0.15):
    """Analytic RGB-D of the inside of an axis-aligned box (the scene bound shrunk by ``margin`` per side), seen
    along the given camera-frame directions: z-depth of the first wall hit (directions have z = -1, so the ray
    parameter IS the sensor depth) and a smooth position-dependent colour.  Used by the tracking test."""
    b = torch.tensor(cfg['mapping']['bound'], dtype=torch.float32)
    lo, hi = b[:, 0] + margin, b[:, 1] - margin
    rays_d = torch.sum(dirs_cam[..., None, :] * c2w[:3, :3], -1)
    o = c2w[:3, 3]
    t_lo, t_hi = (lo - o) / rays_d, (hi - o) / rays_d
    t_exit = torch.where(rays_d > 0, t_hi, t_lo)
    t_exit = torch.where(rays_d.abs() < 1e-9, torch.full_like(t_exit, 1e9), t_exit)
    depth = t_exit.min(-1).values
    hit = o + rays_d * depth[:, None]
    u = (hit - lo) / (hi - lo)
    rgb = torch.stack([0.5 + 0.4 * torch.sin(6.0 * u[:, 0] + 1.0 * u[:, 2]), 0.5 + 0.4 * torch.sin(5.0 * u[:, 1] + 2.0),
                       0.5 + 0.4 * torch.cos(4.0 * u[:, 2] + 3.0 * u[:, 0])], -1)
    return rgb.clamp(0, 1), depth[:, None]
