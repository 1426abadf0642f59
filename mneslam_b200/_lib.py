"""ctypes binding of ``libmneslam_b200.so`` (declarations mirror ``include/mneslam_b200.h``).

There is no CPU fallback: if the library is missing, ``lib()`` raises.  The library is built in-tree by
``mneslam_b200.build`` / ``__graft_entry__.build()``.
"""
import ctypes as C
import os

from . import build as _build

MAX_PLANES = 12
ACC_SLOTS = 16
N_LOSSES = 8
STATIC_SCRATCH_SAMPLES = 131072   # MNE_STATIC_SCRATCH_SAMPLES
ACT_FLOATS_MAX = 296              # MNE_ACT_FLOATS_MAX

# accumulator / loss slots (keep in sync with csrc/common.cuh)
ACC = {'n_valid': 0, 'n_co_front': 1, 'n_co_near': 2, 'n_e_front': 3, 'n_e_center': 4, 'n_e_tail': 5,
       's_rgb': 6, 's_depth': 7, 's_co_fs': 8, 's_co_sdf': 9, 's_e_fs': 10, 's_e_center': 11, 's_e_tail': 12,
       's_rgb0': 13, 's_depth0': 14}
LOSS_KEYS = ('rgb_loss', 'depth_loss', 'co_sdf_loss', 'co_fs_loss', 'e_fs_loss', 'e_center_loss', 'e_tail_loss',
             'psnr')


class PlaneDesc(C.Structure):
    _fields_ = [('data', C.c_void_p), ('grad', C.c_void_p), ('H', C.c_int32), ('W', C.c_int32),
                ('stride_c', C.c_int64), ('stride_h', C.c_int64), ('stride_w', C.c_int64)]


class HashDesc(C.Structure):
    _fields_ = [('table', C.c_void_p), ('grad', C.c_void_p), ('table_h', C.c_void_p), ('n_levels', C.c_int32), ('n_features', C.c_int32),
                ('base_resolution', C.c_int32), ('per_level_scale', C.c_float), ('offsets', C.c_uint32 * 33),
                ('level_scale', C.c_float * 32), ('level_res', C.c_uint32 * 32)]


class SceneDesc(C.Structure):
    _fields_ = [('encoder', C.c_int32), ('two_grid', C.c_int32), ('n_bins', C.c_int32), ('c_dim', C.c_int32),
                ('planes', PlaneDesc * MAX_PLANES), ('hash', HashDesc * 2),
                ('bound_lo', C.c_float * 3), ('bound_ext', C.c_float * 3),
                ('bbox_lo', C.c_double * 3), ('bbox_ext', C.c_double * 3),
                ('w_sdf0', C.c_void_p), ('w_sdf1', C.c_void_p), ('w_col0', C.c_void_p), ('w_col1', C.c_void_p),
                ('g_sdf0', C.c_void_p), ('g_sdf1', C.c_void_p), ('g_col0', C.c_void_p), ('g_col1', C.c_void_p),
                ('dirty', C.c_void_p), ('dirty_base', C.c_void_p)]


class SampleCfg(C.Structure):
    _fields_ = [('n_uniform', C.c_int32), ('n_range', C.c_int32), ('uniform', C.c_void_p),
                ('range_offsets', C.c_void_p), ('range_fallback', C.c_void_p), ('perturb', C.c_int32)]


class RenderCfg(C.Structure):
    _fields_ = [('trunc', C.c_float), ('sc_trunc', C.c_float), ('model_trunc', C.c_float),
                ('center_trunc', C.c_float), ('depth_trunc', C.c_float), ('white_bkgd', C.c_int32),
                ('first_pass', C.c_int32), ('norm_rays', C.c_double)]


_SIGNATURES = {
    'mne_version': (C.c_int, []),
    'mne_last_error': (C.c_char_p, []),
    'mne_set_option': (C.c_int, [C.c_char_p, C.c_int]),
    'mne_get_option': (C.c_int, [C.c_char_p, C.c_void_p]),
    'mne_assemble_rays': (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32] + [C.c_void_p] * 5),
    'mne_draw_rays': (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int32, C.c_void_p, C.c_int64, C.c_int32, C.c_void_p, C.c_int32]
                      + [C.c_void_p] * 6),
    'mne_pose_rays': (C.c_int, [C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p]),
    'mne_pose_step': (C.c_int, [C.c_int32] + [C.c_void_p] * 5 + [C.c_int32] + [C.c_void_p] * 3 + [C.c_float] * 5 + [C.c_void_p, C.c_void_p]),
    'mne_sample_rays': (C.c_int, [C.POINTER(SampleCfg), C.POINTER(RenderCfg), C.c_int32, C.c_void_p, C.c_void_p,
                                  C.c_void_p, C.c_void_p, C.c_void_p]),
    'mne_sample_importance': (C.c_int, [C.POINTER(RenderCfg), C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p,
                                        C.c_int32] + [C.c_void_p] * 6),
    'mne_render_forward': (C.c_int, [C.POINTER(SceneDesc), C.POINTER(RenderCfg), C.c_int32, C.c_int32] +
                           [C.c_void_p] * 16),
    'mne_act_floats': (C.c_int, [C.POINTER(SceneDesc)]),
    'mne_render_forward_fused': (C.c_int, [C.POINTER(SceneDesc), C.POINTER(RenderCfg), C.c_int32, C.c_int32] + [C.c_void_p] * 19),
    'mne_render_backward_points': (C.c_int, [C.POINTER(SceneDesc), C.c_int32, C.c_int32] + [C.c_void_p] * 8),
    'mne_decoder_wgrad': (C.c_int, [C.POINTER(SceneDesc), C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p]),
    'mne_composite': (C.c_int, [C.POINTER(RenderCfg), C.c_int32, C.c_int32] + [C.c_void_p] * 10),
    'mne_loss_finalize': (C.c_int, [C.POINTER(RenderCfg), C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p]),
    'mne_backward_scratch_bytes': (C.c_size_t, [C.c_int32, C.c_int32]),
    'mne_render_backward': (C.c_int, [C.POINTER(SceneDesc), C.POINTER(RenderCfg), C.c_int32, C.c_int32] +
                            [C.c_void_p] * 16),
    'mne_render_backward_static': (C.c_int, [C.POINTER(SceneDesc), C.POINTER(RenderCfg), C.c_int32, C.c_int32] +
                                   [C.c_void_p] * 14),
    'mne_hash_indices': (C.c_int, [C.POINTER(HashDesc), C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]),
    'mne_adam_tick': (C.c_int, [C.c_void_p, C.c_void_p]),
    'mne_adam_step': (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_float, C.c_float,
                                C.c_float, C.c_float, C.c_float, C.c_void_p, C.c_int64, C.c_int32, C.c_void_p]),
    'mne_adam_step_segments': (C.c_int, [C.c_void_p] * 4 + [C.c_int32] + [C.c_void_p] * 4 + [C.c_float, C.c_float, C.c_void_p,
                                                                               C.c_int64, C.c_int32, C.c_void_p, C.c_void_p]),
    'mne_adam_step_segments_dirty': (C.c_int, [C.c_void_p] * 4 + [C.c_int32] + [C.c_void_p] * 4 + [C.c_float, C.c_float, C.c_void_p,
                                                                                     C.c_int64, C.c_int32, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]),
    'mne_adam_step_peers': (C.c_int, [C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_void_p, C.c_void_p, C.c_int32,
                                      C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_float, C.c_void_p, C.c_int64,
                                      C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    'mne_grad_push': (C.c_int, [C.c_void_p, C.c_int64, C.c_int64, C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p]),
    'mne_dirty_push_data': (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p]),
    'mne_peer_scatter': (C.c_int, [C.c_void_p, C.c_int32, C.c_void_p, C.c_int64, C.c_int32, C.c_void_p]),
    'mne_peer_sum': (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_int32, C.c_void_p]),
    'mne_peer_barrier': (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_void_p]),
    'mne_bench_gather': (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_void_p, C.c_void_p]),
    'mne_grad_dirty_map': (C.c_int, [C.c_void_p, C.c_int64, C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p]),
    'mne_dirty_push': (C.c_int, [C.c_void_p, C.c_int64, C.c_int64, C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p]),
    'mne_query_points': (C.c_int, [C.POINTER(SceneDesc), C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                   C.c_void_p]),
}

_lib = None


MAX_PEERS = 16


def peer_pad_words(n_channels):
    """MNE_PEER_PAD_WORDS of include/mneslam_b200.h."""
    return n_channels * (MAX_PEERS + 1) + 1


def exported_symbols():
    """Every entry point ``include/mneslam_b200.h`` declares."""
    return sorted(_SIGNATURES)


def lib():
    global _lib
    if _lib is None:
        path = _build.LIB_PATH
        if not os.path.exists(path):
            raise RuntimeError(
                'mneslam_b200: %s is missing.  Build it with `python -m mneslam_b200.build` (needs nvcc); '
                'there is no CPU fallback.' % path)
        handle = C.CDLL(path)
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(handle, name)          # AttributeError here = header / library mismatch
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


class MneError(RuntimeError):
    pass


def check(status, what):
    if status != 0:
        raise MneError('%s failed (%d): %s' % (what, status, lib().mne_last_error().decode()))
