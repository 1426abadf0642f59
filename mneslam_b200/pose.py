"""Pose parameterisations that consume the ray gradients of the fused backward.

The fused kernels return ``d loss / d rays_o`` and ``d loss / d rays_d``; a camera pose enters the path only
through ``rays_o = t`` and ``rays_d = sum_k dirs_k R[:, k]``, so any parameterisation chains by autograd on a
handful of 3x3 products.  Provided: axis-angle / translation and quaternion / translation, the two
``training.rot_rep`` values of the reference (``optimization/utils.py:161-210``, selected at
``mneslam_mp.py:195-208``), and the 7-vector of ``optimization/pose_array.py:6-29``.
"""
import torch


def axis_angle_to_matrix(r):
    """Rodrigues: ``I + sin(t) K + (1 - cos(t)) K^2`` with ``K`` the cross-product matrix of ``r / |r|``."""
    theta = torch.linalg.norm(r, dim=-1, keepdim=True)
    k = r / theta
    zero = torch.zeros_like(k[..., 0])
    K = torch.stack([zero, -k[..., 2], k[..., 1],
                     k[..., 2], zero, -k[..., 0],
                     -k[..., 1], k[..., 0], zero], -1).reshape(r.shape[:-1] + (3, 3))
    eye = torch.eye(3, dtype=r.dtype, device=r.device).expand_as(K)
    return eye + torch.sin(theta)[..., None] * K + (1.0 - torch.cos(theta))[..., None] * (K @ K)


def quaternion_to_matrix(q):
    """Real-part-first quaternion (not assumed unit) to rotation matrix."""
    w, x, y, z = q.unbind(-1)
    s = 2.0 / (q * q).sum(-1)
    rows = [1 - s * (y * y + z * z), s * (x * y - z * w), s * (x * z + y * w),
            s * (x * y + z * w), 1 - s * (x * x + z * z), s * (y * z - x * w),
            s * (x * z - y * w), s * (y * z + x * w), 1 - s * (x * x + y * y)]
    return torch.stack(rows, -1).reshape(q.shape[:-1] + (3, 3))


def transform_matrix(rot, trans, rot_rep='axis_angle'):
    """``at_to_transform_matrix`` / ``qt_to_transform_matrix``: [B,3|4] + [B,3] -> [B,4,4]."""
    R = axis_angle_to_matrix(rot) if rot_rep == 'axis_angle' else quaternion_to_matrix(rot)
    T = torch.eye(4, dtype=rot.dtype, device=rot.device).repeat(rot.shape[0], 1, 1)
    T[:, :3, :3] = R
    T[:, :3, 3] = trans
    return T


def pose7_to_matrix(p):
    """``PoseArray.get_camera_from_tensor``: 7-vector (quaternion wxyz, translation) -> [3,4]."""
    return torch.cat([quaternion_to_matrix(p[..., :4]), p[..., 4:, None]], -1)


def rays_from_c2w(c2w, dirs_cam):
    """``rays_o = t`` repeated, ``rays_d = sum_k dirs_k R[:, k]`` (mp_slam/mapper.py:392-393)."""
    rays_d = torch.sum(dirs_cam[:, None, :] * c2w[:3, :3], -1)
    return c2w[:3, 3].expand(dirs_cam.shape[0], 3), rays_d
