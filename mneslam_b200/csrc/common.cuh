// Shared device/host helpers for the mneslam_b200 kernels (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <atomic>

#include "../../include/mneslam_b200.h"

namespace mne {

constexpr int kWarp = 32;
constexpr int kHidden = 32;    // decoder.hidden_dim == hidden_dim_color
constexpr int kGeo = 15;       // decoder.geo_feat_dim
constexpr int kBins = 16;      // pos.n_bins
constexpr int kBlob = 3 * kBins;
constexpr int kCDim = 32;      // model.c_dim

// accumulator slots (include/mneslam_b200.h)
enum AccSlot {
    A_N_VALID = 0, A_N_CO_FRONT, A_N_CO_NEAR, A_N_E_FRONT, A_N_E_CENTER, A_N_E_TAIL,
    A_S_RGB, A_S_DEPTH, A_S_CO_FS, A_S_CO_SDF, A_S_E_FS, A_S_E_CENTER, A_S_E_TAIL, A_S_RGB0, A_S_DEPTH0
};
// loss slots
enum LossSlot { L_RGB = 0, L_DEPTH, L_CO_SDF, L_CO_FS, L_E_FS, L_E_CENTER, L_E_TAIL, L_PSNR };

void set_error(const char* fmt, ...);

#define MNE_CHECK_ARG(cond, ...)            \
    do {                                    \
        if (!(cond)) {                      \
            mne::set_error(__VA_ARGS__);    \
            return 1;                       \
        }                                   \
    } while (0)

#define MNE_CHECK_CUDA(expr)                                                                   \
    do {                                                                                       \
        cudaError_t _e = (expr);                                                               \
        if (_e != cudaSuccess) {                                                               \
            mne::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
            return 2;                                                                          \
        }                                                                                      \
    } while (0)

// ---------------------------------------------------------------------------------------------
// warp helpers
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ int warp_min(int v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = min(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

__device__ __forceinline__ float sigmoidf(float x) { return 1.0f / (1.0f + expf(-x)); }

// ---------------------------------------------------------------------------------------------
// per-sample loss masks: one definition shared by the counting (sample), summing (forward) and
// gradient (backward) kernels so the three always agree.  Follows model/utils.py:117-145 (Co-SLAM pair,
// every ray) and model/scene_rep.py:489-499 (ESLAM triple, rays with depth > 0 only).
// ---------------------------------------------------------------------------------------------
enum SampleMask { M_CO_FRONT = 1, M_CO_NEAR = 2, M_E_FRONT = 4, M_E_CENTER = 8, M_E_TAIL = 16 };

__device__ __forceinline__ unsigned sample_masks(float z, float d, const MneRenderCfg& c) {
    unsigned m = 0;
    const bool has_d = d > 0.0f;
    const bool co_front = z < __fsub_rn(d, c.sc_trunc);
    const bool co_back = z > __fadd_rn(d, c.sc_trunc);
    if (co_front) m |= M_CO_FRONT;
    if (!co_front && !co_back && has_d) m |= M_CO_NEAR;
    if (has_d) {
        const bool f = z < __fsub_rn(d, c.model_trunc);
        const bool b = z > __fadd_rn(d, c.model_trunc);
        const bool ctr = (z > __fsub_rn(d, c.center_trunc)) && (z < __fadd_rn(d, c.center_trunc));
        if (f) m |= M_E_FRONT;
        if (ctr) m |= M_E_CENTER;
        if (!f && !b && !ctr) m |= M_E_TAIL;
    }
    return m;
}

__device__ __forceinline__ bool depth_valid(float d, const MneRenderCfg& c) { return d > 0.0f && d < c.depth_trunc; }

}  // namespace mne
