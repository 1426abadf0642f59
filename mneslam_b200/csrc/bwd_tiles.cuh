// Pieces shared by the backward tile kernels (render_bwd_v2.cu: forward recompute + in-loop weight gradients;
// render_bwd_v3.cu: kept record, barrier-free tiles + a separate weight-gradient kernel).
#pragma once
#include "tile16.cuh"

namespace mne {

constexpr int kBwdWarps = 8;
constexpr int kActStride = 40;     // hidden-wide tiles [16][32 + 8]: == 8 (mod 32)
constexpr int kGeoStride = 24;     // 16-wide tiles
constexpr int kRgbStride = 8;

// hi-only transposed weights for the "d input = d output . W" products (B fragment = W[k = out][n = in]),
// stored [in][out] so that a lane's (b0, b1) = outputs (k0 + 2t, k0 + 2t + 1) of input n0 + g is one 64-bit load
template <bool TWO>
struct NetT {
    static constexpr int S1 = 40;                        // W1T  [64 feat][32 j]
    static constexpr int OFF_W1T = 0;
    static constexpr int OFF_VGT = OFF_W1T + 64 * S1;    // V1T geo [16][32 j] (row 15 = 0)
    static constexpr int OFF_VCT = OFF_VGT + 16 * S1;    // V1T colour feat [64][32 j] (two grids)
    static constexpr int OFF_W2T = OFF_VCT + (TWO ? 64 * S1 : 0);   // W2T [32 j][16 i] stride 24
    static constexpr int OFF_V2T = OFF_W2T + 32 * kGeoStride;       // V2T [32 j][8 c]
    static constexpr int OFF_W1BT = OFF_V2T + 32 * kRgbStride;      // W1T blob rows [48][32 j]   (ray gradients only)
    static constexpr int OFF_V1BT = OFF_W1BT + 48 * S1;             // V1T blob rows [48][32 j]
    static constexpr int FLOATS_NORAY = OFF_W1BT;
    static constexpr int FLOATS_RAY = OFF_V1BT + 48 * S1;
};

template <bool TWO, bool RAY>
__device__ __forceinline__ void load_weights_t(const MneSceneDesc& sc, float* WT, int tid, int nthreads) {
    using N = NetV2<TWO>;
    using T = NetT<TWO>;
    const bool hash = sc.encoder == 1;
    if (RAY)
        for (int i = tid; i < 48 * 32; i += nthreads) {
            const int k = i >> 5, j = i & 31;
            WT[T::OFF_W1BT + k * T::S1 + j] = sc.w_sdf0[j * sdf0_in(hash) + sdf0_col(64 + k, hash)];
            WT[T::OFF_V1BT + k * T::S1 + j] = sc.w_col0[j * col0_in(TWO, hash) + k];
        }
    for (int i = tid; i < 64 * 32; i += nthreads) {
        const int k = i >> 5, j = i & 31;
        const int col = sdf0_col(k, hash);
        WT[T::OFF_W1T + k * T::S1 + j] = (col >= 0) ? sc.w_sdf0[j * sdf0_in(hash) + col] : 0.0f;
        if (TWO) WT[T::OFF_VCT + k * T::S1 + j] = (!hash || k < 32) ? sc.w_col0[j * col0_in(TWO, hash) + kBlob + k] : 0.0f;
    }
    for (int i = tid; i < 16 * 32; i += nthreads) {
        const int m = i >> 5, j = i & 31;
        WT[T::OFF_VGT + m * T::S1 + j] = (m < kGeo) ? sc.w_col0[j * col0_in(TWO, hash) + (col0_in(TWO, hash) - kGeo) + m] : 0.0f;
        const int jj = i >> 4, c = i & 15;                     // W2T[j][c], c: 0..14 geo (out c+1), 15 sdf (out 0)
        WT[T::OFF_W2T + jj * kGeoStride + c] = sc.w_sdf1[((c == 15) ? 0 : c + 1) * 32 + jj];
    }
    for (int i = tid; i < 32 * 8; i += nthreads) {
        const int j = i >> 3, c = i & 7;
        WT[T::OFF_V2T + j * kRgbStride + c] = (c < 3) ? sc.w_col1[c * 32 + j] : 0.0f;
    }
}

struct PtsBwdParams {
    MneSceneDesc sc;
    int64_t n_pts;
    int n_samples;
    const float* rays_o;
    const float* rays_d;
    const float* z_vals;
    const float4* draw;      // [n_pts] d loss / d raw
    float* g_rays_o;         // [N,3] accumulated into (zeroed by the entry point), or NULL
    float* g_rays_d;
    const int* list;         // samples with a non-zero draw (compacted by rays_bwd_kernel)
    const int* n_active;     // how many
    const float* act;        // unused here (calls that carry the forward's record go to render_bwd_v3.cu)
};

__device__ __forceinline__ void mma_1x(float (&d)[4], float a0, float a1, float a2, float a3, float b0, float b1) {
    const uint32_t a[4] = {__float_as_uint(a0), __float_as_uint(a1), __float_as_uint(a2), __float_as_uint(a3)};
    mma_tf32(d, a, __float_as_uint(b0), __float_as_uint(b1));
}

// store an accumulator fragment block (rows g, g+8; cols n*8 + 2t, +1) into a [16][stride] tile
__device__ __forceinline__ void store_frag(float* tile, int stride, int n, int g, int t, const float (&c)[4]) {
    *reinterpret_cast<float2*>(tile + g * stride + n * 8 + 2 * t) = make_float2(c[0], c[1]);
    *reinterpret_cast<float2*>(tile + (g + 8) * stride + n * 8 + 2 * t) = make_float2(c[2], c[3]);
}

}  // namespace mne
