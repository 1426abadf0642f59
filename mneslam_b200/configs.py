"""Hot-path configuration dictionaries.

The scene model reads every constant from ``config[...][...]`` at call time, exactly as the
reference does (reference ``model/scene_rep.py`` and ``mneslam_mp.py:350-469``).
The two dictionaries below restate the hot-path keys of the reference's shipped YAML files so that
nothing has to read ``/root/reference`` at run time:

* ``replica_room0()``      <- ``configs/Replica/replica.yaml`` + ``configs/Replica/room0.yaml``
* ``scannet_scene0000()``  <- ``configs/ScanNet/scannet.yaml`` + ``configs/ScanNet/scene0000.yaml``
  (+ the keys those files lack and the code requires: ``is_co_sdf``, ``training.n_samples``,
  values taken from the Replica file, SURVEY.md section 5 "config / flags").

``load_config`` reads a YAML file with recursive ``inherit_from`` like the reference's
``config.py:4-49`` so a user's existing config files keep working.
"""
import copy
import os


def _base():
    return {
        'scale': 1,
        'is_co_sdf': False,
        'data': {'downsample': 1, 'sc_factor': 1, 'translation': 0},
        'grid': {'enc': 'HashGrid', 'tcnn_encoding': True, 'hash_size': 16,
                 'voxel_color': 0.08, 'voxel_sdf': 0.02, 'oneGrid': True},
        'pos': {'enc': 'OneBlob', 'n_bins': 16},
        'decoder': {'geo_feat_dim': 15, 'hidden_dim': 32, 'num_layers': 2,
                    'num_layers_color': 2, 'hidden_dim_color': 32, 'tcnn_network': False},
        'planes_res': {'coarse': 0.02, 'fine': 0.01, 'bound_dividable': 0.02},
        'c_planes_res': {'coarse': 0.08, 'fine': 0.02},
        'model': {'c_dim': 32, 'truncation': 0.1, 'input_ch': 64, 'input_ch_pos': 48},
    }


def replica_room0():
    cfg = _base()
    cfg.update({
        'dataset': 'replica',
        'mapping': {
            'bound': [[-1.0, 7.0], [-1.3, 3.7], [-1.7, 1.4]],
            'marching_cubes_bound': [[-1.0, 7.0], [-1.3, 3.7], [-1.7, 1.4]],
            'sample': 2048, 'iters': 50, 'first_iters': 500, 'loop_iters': 100, 'distill_iters': 100,
            'lr_embed': 0.005, 'lr_embed_color': 0.005, 'lr_decoder': 0.01,
            'lr_rot': 0.001, 'lr_trans': 0.001, 'n_pixels': 0.05, 'min_pixels_cur': 100,
            'w_sdf_fs': 5, 'w_sdf_center': 200, 'w_sdf_tail': 30,
        },
        'tracking': {'lr_rot': 0.00001, 'lr_trans': 0.00001, 'iter': 50, 'sample': 4096},
        'cam': {'H': 680, 'W': 1200, 'fx': 600.0, 'fy': 600.0, 'cx': 599.5, 'cy': 339.5,
                'crop_edge': 0, 'near': 0, 'far': 10, 'depth_trunc': 100.0},
        'training': {'rgb_weight': 5.0, 'depth_weight': 0.1, 'sdf_weight': 1200, 'fs_weight': 10,
                     'n_samples': 256, 'n_samples_d': 32, 'range_d': 0.1, 'n_range_d': 11,
                     'n_importance': 0, 'perturb': 1, 'white_bkgd': False, 'trunc': 0.1,
                     'rot_rep': 'axis_angle'},
    })
    return cfg


def scannet_scene0000():
    cfg = _base()
    cfg['grid'].update({'hash_size': 19, 'voxel_sdf': 0.04, 'oneGrid': False})
    cfg.update({
        'dataset': 'scannet',
        'mapping': {
            'bound': [[-0.1, 8.6], [-0.1, 8.9], [-0.3, 3.3]],
            'marching_cubes_bound': [[-0.1, 8.6], [-0.1, 8.9], [-0.3, 3.3]],
            'sample': 2048, 'iters': 50, 'first_iters': 500, 'loop_iters': 100, 'distill_iters': 100,
            'lr_embed': 0.005, 'lr_embed_color': 0.005, 'lr_decoder': 0.01,
            'lr_rot': 0.001, 'lr_trans': 0.001, 'n_pixels': 0.05, 'min_pixels_cur': 20,
            'w_sdf_fs': 5, 'w_sdf_center': 200, 'w_sdf_tail': 10,
        },
        'tracking': {'lr_rot': 0.001, 'lr_trans': 0.001, 'iter': 20, 'sample': 1024},
        'cam': {'H': 480, 'W': 640, 'fx': 577.590698, 'fy': 578.729797, 'cx': 318.905426,
                'cy': 242.683609, 'crop_edge': 10, 'near': 0, 'far': 8, 'depth_trunc': 100.0},
        'training': {'rgb_weight': 5.0, 'depth_weight': 0.1, 'sdf_weight': 1000, 'fs_weight': 10,
                     'n_samples': 256, 'n_samples_d': 96, 'range_d': 0.25, 'n_range_d': 21,
                     'n_importance': 0, 'perturb': 1, 'white_bkgd': False, 'trunc': 0.1,
                     'rot_rep': 'quat'},
    })
    return cfg


def scannet_scene0000_onegrid():
    """ScanNet-shaped sampling (S = 96 + 21) on a single set of planes (``scannet.yaml`` default)."""
    cfg = scannet_scene0000()
    cfg['grid']['oneGrid'] = True
    return cfg


def small_test_config(one_grid=True, n_samples_d=8, n_range_d=5, n_importance=0, n_samples=16):
    """A tiny scene (few thousand texels) with the same code paths, for unit tests."""
    cfg = replica_room0()
    cfg['mapping']['bound'] = [[-1.0, 1.0], [-0.8, 0.9], [-0.5, 0.6]]
    cfg['planes_res'] = {'coarse': 0.2, 'fine': 0.1, 'bound_dividable': 0.2}
    cfg['c_planes_res'] = {'coarse': 0.4, 'fine': 0.2}
    cfg['cam'].update({'near': 0, 'far': 3})
    cfg['training'].update({'n_samples_d': n_samples_d, 'n_range_d': n_range_d,
                            'n_importance': n_importance, 'n_samples': n_samples})
    cfg['grid']['oneGrid'] = one_grid
    return cfg


def update_recursive(dst, src):
    for k, v in src.items():
        if isinstance(v, dict):
            if not isinstance(dst.get(k), dict):
                dst[k] = {}
            update_recursive(dst[k], v)
        else:
            dst[k] = v


def load_config(path, default_path=None):
    """YAML loader with recursive ``inherit_from`` (same semantics as reference ``config.py:4-49``:
    the parent path is resolved relative to the current working directory)."""
    import yaml
    with open(path, 'r') as f:
        special = yaml.full_load(f) or {}
    parent = special.get('inherit_from')
    if parent is not None:
        cfg = load_config(parent, default_path)
    elif default_path is not None:
        with open(default_path, 'r') as f:
            cfg = yaml.full_load(f) or {}
    else:
        cfg = {}
    update_recursive(cfg, special)
    return cfg


def clone(cfg):
    return copy.deepcopy(cfg)
