"""CPU-side checks: the C-ABI library loads and exports every symbol the header declares, host descriptors
are laid out as the C structs, and the product path refuses to run without a GPU (no fallback)."""
import ctypes
import os
import re

import pytest
import torch

from mneslam_b200 import _lib, configs, ops

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, 'include', 'mneslam_b200.h')).read()
    text = re.sub(r'/\*.*?\*/', '', text, flags=re.S)
    return sorted(set(re.findall(r'\b(mne_[a-z_0-9]+)\s*\(', text)))


def test_library_exports_every_declared_symbol():
    handle = _lib.lib()
    names = declared_symbols()
    assert len(names) >= 8
    for n in names:
        assert hasattr(handle, n), n
    assert sorted(_lib.exported_symbols()) == names, 'ctypes signatures and header are out of sync'
    assert handle.mne_version() >= 100


def test_struct_layout_matches_header():
    assert ctypes.sizeof(_lib.PlaneDesc) == 48
    assert ctypes.sizeof(_lib.HashDesc) == 432
    assert ctypes.sizeof(_lib.SceneDesc) == 1608
    assert _lib.SceneDesc.bound_lo.offset == 1456 and _lib.SceneDesc.bbox_lo.offset == 1480
    assert _lib.SceneDesc.w_sdf0.offset == 1528
    assert ctypes.sizeof(_lib.RenderCfg) == 40 and _lib.RenderCfg.norm_rays.offset == 32
    assert ctypes.sizeof(_lib.SampleCfg) == 40


def test_argument_errors_are_reported_without_a_gpu():
    handle = _lib.lib()
    rc = ops.make_render_cfg(configs.replica_room0(), 16)
    sc = _lib.SampleCfg()
    assert handle.mne_sample_rays(ctypes.byref(sc), ctypes.byref(rc), 4, None, None, None, None, None) != 0
    assert b'null' in handle.mne_last_error()
    scene = _lib.SceneDesc()
    scene.n_bins, scene.c_dim = 8, 32
    assert handle.mne_query_points(ctypes.byref(scene), 4, None, None, None, None, None) != 0
    assert b'n_bins' in handle.mne_last_error()
    assert handle.mne_backward_scratch_bytes(2048, 43) == 2048 * 43 * 20 + 16   # d_raw, active-sample list, its length
    assert handle.mne_assemble_rays(None, None, 0, None, None, 3, 16, None, None, None, None, None) != 0
    assert b'mne_assemble_rays' in handle.mne_last_error()
    assert handle.mne_set_option(b'adam_spare_sms', 1000) != 0 and handle.mne_set_option(b'adam_spare_sms', 0) == 0


def test_render_cfg_uses_float32_roundings():
    rc = ops.make_render_cfg(configs.replica_room0(), 2048)
    import numpy as np
    assert rc.trunc == np.float32(0.1) and rc.center_trunc == np.float32(0.4 * 0.1)
    assert rc.norm_rays == 2048.0


@pytest.mark.skipif(torch.cuda.is_available(), reason='checks the no-GPU behaviour')
def test_no_cpu_fallback():
    from mneslam_b200 import JointEncoding
    cfg = configs.small_test_config()
    model = JointEncoding(cfg, torch.tensor(cfg['mapping']['bound'], dtype=torch.float64))
    with pytest.raises(RuntimeError, match='CUDA'):
        model.render_rays(torch.zeros(4, 3), torch.ones(4, 3), target_d=torch.ones(4, 1))
    with pytest.raises(RuntimeError):
        model.run_network_flat(torch.zeros(4, 3))


def test_pose_parameterisations_match_oracle():
    from mneslam_b200 import pose
    from oracle import joint_encoding as J
    r = torch.tensor([[0.3, -0.2, 0.5], [1e-3, 2e-3, -1e-3]])
    q = torch.tensor([[0.9, 0.1, -0.3, 0.2], [2.0, 0.0, 0.1, 0.0]])
    assert torch.allclose(pose.axis_angle_to_matrix(r), J.axis_angle_to_matrix(r), atol=1e-6)
    assert torch.allclose(pose.quaternion_to_matrix(q), J.quat_to_matrix(q), atol=1e-6)
    R = pose.quaternion_to_matrix(q)
    assert torch.allclose(R @ R.transpose(-1, -2), torch.eye(3).expand(2, 3, 3), atol=1e-5)
    T = pose.transform_matrix(r, torch.ones(2, 3), 'axis_angle')
    assert T.shape == (2, 4, 4) and torch.equal(T[:, 3], torch.tensor([[0., 0, 0, 1]] * 2))


def test_hash_grid_checkpoint_round_trip(tmp_path):
    """With grid.use_hash the encoder's tables are plain tensors outside state_dict() and all_planes is empty: the checkpoint
    must carry them (key hash_tables), load_foreign_model must restore them and refuse a checkpoint that lacks them."""
    import torch
    from mneslam_b200 import JointEncoding, configs, exchange
    cfg = configs.small_test_config(one_grid=False)
    cfg['grid'].update({'use_hash': True, 'hash_size': 10, 'voxel_sdf': 0.08, 'voxel_color': 0.16})
    bb = torch.tensor(cfg['mapping']['bound'], dtype=torch.float64)
    torch.manual_seed(1)
    a = JointEncoding(cfg, bb)
    torch.manual_seed(2)
    b = JointEncoding(cfg, bb)
    assert len(a.hash_tables) == 2 and not torch.equal(a.hash_tables[0], b.hash_tables[0])
    path = str(tmp_path / 'latest_checkpoint.pt')
    exchange.save_checkpoint(a, path)
    assert 'hash_tables' in torch.load(path, weights_only=False)
    exchange.load_foreign_model(b, path, device='cpu')
    assert all(torch.equal(x, y) for x, y in zip(a.hash_tables, b.hash_tables))
    assert all(torch.equal(x, y) for x, y in zip(a.decoder.weight_list(), b.decoder.weight_list()))
    ckpt = exchange.checkpoint_dict(a)
    del ckpt['hash_tables']
    torch.save(ckpt, path)
    import pytest
    with pytest.raises(KeyError):
        exchange.load_foreign_model(b, path, device='cpu')
