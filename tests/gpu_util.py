"""Helpers for the GPU parity tests: build the product module from a golden case."""
import numpy as np
import torch

from mneslam_b200 import JointEncoding


def rel(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return float(np.abs(a - b).max() / (np.abs(b).max() + 1e-12))


def cosine(a, b):
    a, b = np.asarray(a, np.float64).ravel(), np.asarray(b, np.float64).ravel()
    return float(a @ b / (np.linalg.norm(a) * np.linalg.norm(b) + 1e-300))


def build_model(cfg, bounding_box, planes, weights, channels_last=True, as_params=True):
    """``planes``: tuple of lists (all_planes layout, CPU tensors); ``weights``: dict sdf0/sdf1/col0/col1."""
    torch.manual_seed(0)
    model = JointEncoding(cfg, bounding_box).cuda()
    assert len(model.all_planes) == len(planes)
    for dst, src in zip(model.all_planes, planes):
        for i, p in enumerate(src):
            assert tuple(dst[i].shape) == tuple(p.shape), (dst[i].shape, p.shape)
            t = p.cuda()
            t = t.contiguous(memory_format=torch.channels_last) if channels_last else t.contiguous()
            dst[i] = torch.nn.Parameter(t) if as_params else t      # as mneslam_mp.py:451-457
            if not channels_last:
                dst[i]._mne_keep_layout = True      # keep the NCHW layout: these cases exist to exercise the strided kernels
    with torch.no_grad():
        for w, src in zip(model.decoder.weight_list(), ('sdf0', 'sdf1', 'col0', 'col1')):
            w.copy_(weights[src])
    model.train()
    return model


def flat_planes(model):
    return [p for grp in model.all_planes for p in grp]


def total_loss(cfg, ret):
    """mneslam_mp.py:350-372 (restated for the test, like oracle.joint_encoding.total_loss)."""
    t, m = cfg['training'], cfg['mapping']
    loss = t['rgb_weight'] * ret['rgb_loss'] + t['depth_weight'] * ret['depth_loss']
    if cfg['is_co_sdf']:
        return loss + t['sdf_weight'] * ret['co_sdf_loss'] + t['fs_weight'] * ret['co_fs_loss']
    return (loss + m['w_sdf_fs'] * ret['e_fs_loss'] + m['w_sdf_center'] * ret['e_center_loss']
            + m['w_sdf_tail'] * ret['e_tail_loss'])
