// Building blocks shared by the forward and backward render kernels.
//
// Work decomposition: a warp owns a group of consecutive rays; the group's (ray, sample) points are
// flattened and processed in chunks of 32.  Each chunk alternates between three lane mappings:
//   lane = channel   tri-plane gather / gradient scatter: one 128-byte line (32 channels) per texel
//   lane = point     OneBlob + the decoder MLPs, weights broadcast from shared memory
//   lane = neuron    weight-gradient outer products, reduced over the chunk's points
// Activations move between the mappings through per-warp shared-memory tiles laid out [point][feature]
// with a row stride that is 4 * odd, so that a lane reading its own row with 128-bit loads, a lane
// reading a column, and a broadcast 128-bit load of another point's row are all bank-conflict free.
#pragma once
#include "common.cuh"

namespace mne {

// ---------------------------------------------------------------------------------------------
// decoder geometry (reference model/decoder.py:110-175).  Colour-net input order: one grid
// [blob48 | geo15], two grids [blob48 | cfeat64 | geo15].  Shared-memory copies are TRANSPOSED
// ([in][out]) so that for a fixed input the 32 output weights are one contiguous 128-byte row.
// sdf1 is stored with its output columns rotated (geo 0..14 first, sdf last) so the geo block is
// 16-byte aligned; the colour net gets one zero row that lines up with that sdf slot.
// ---------------------------------------------------------------------------------------------
template <bool TWO>
struct Net {
    static constexpr int F = 64;
    static constexpr int IN_SDF = F + kBlob;                       // 112
    static constexpr int IN_COL = kBlob + (TWO ? F : 0) + kGeo;    // 63 / 127
    static constexpr int GEO_ROW0 = kBlob + (TWO ? F : 0);         // first geo row of V1T
    static constexpr int IN_COL_PAD = IN_COL + 1;                  // 64 / 128
    static constexpr int OFF_W1T = 0;                              // [IN_SDF][32]
    static constexpr int OFF_W2T = OFF_W1T + IN_SDF * 32;          // [32][16]
    static constexpr int OFF_V1T = OFF_W2T + 32 * 16;              // [IN_COL_PAD][32]
    static constexpr int OFF_V2T = OFF_V1T + IN_COL_PAD * 32;      // [32][4]
    static constexpr int W_FLOATS = OFF_V2T + 32 * 4;
    static constexpr int N_TAPS = TWO ? 12 : 6;                    // plane-levels per point
};

// tile row strides (floats); all are 4 * odd
constexpr int kInStride = 116;    // [feat64 | blob48] + pad
constexpr int kCfStride = 68;     // colour feat64 + pad
constexpr int kHStride = 36;      // hidden 32 + pad
constexpr int kGStride = 20;      // [geo15 | sdf] + pad

template <bool TWO>
__device__ __forceinline__ void load_weights_transposed(const MneSceneDesc& sc, float* WT, int tid, int nthreads) {
    using N = Net<TWO>;
    for (int i = tid; i < N::IN_SDF * 32; i += nthreads) {
        const int k = i >> 5, j = i & 31;
        WT[N::OFF_W1T + i] = sc.w_sdf0[j * N::IN_SDF + k];
    }
    for (int i = tid; i < 32 * 16; i += nthreads) {
        const int j = i >> 4, c = i & 15;                 // c: 0..14 geo (out 1..15), 15 sdf (out 0)
        const int out = (c == 15) ? 0 : c + 1;
        WT[N::OFF_W2T + i] = sc.w_sdf1[out * 32 + j];
    }
    for (int i = tid; i < N::IN_COL_PAD * 32; i += nthreads) {
        const int k = i >> 5, j = i & 31;
        WT[N::OFF_V1T + i] = (k < N::IN_COL) ? sc.w_col0[j * N::IN_COL + k] : 0.0f;
    }
    for (int i = tid; i < 32 * 4; i += nthreads) {
        const int j = i >> 2, c = i & 3;
        WT[N::OFF_V2T + i] = (c < 3) ? sc.w_col1[c * 32 + j] : 0.0f;
    }
}

// ---------------------------------------------------------------------------------------------
// bilinear taps (reference model/scene_rep.py:43-48 = F.grid_sample bilinear / border / align_corners)
// ---------------------------------------------------------------------------------------------
struct AxisTap {
    int i0;
    float f;
    bool ok1;      // i0 + 1 is inside the plane
    bool grad_on;  // pixel coordinate was not clamped: gradient flows (ATen clip_coordinates_set_grad)
};

__device__ __forceinline__ AxisTap axis_tap(float g, int size) {
    AxisTap t;
    const float hi = (float)(size - 1);
    float u = __fmul_rn(__fmul_rn(__fadd_rn(g, 1.0f), 0.5f), hi);
    t.grad_on = (u > 0.0f) && (u < hi);
    u = fminf(fmaxf(u, 0.0f), hi);
    const float fl = floorf(u);
    t.f = u - fl;
    t.i0 = (int)fl;
    t.ok1 = (t.i0 + 1) <= size - 1;
    return t;
}

// normalised plane coordinate, model/utils.py:38-40
__device__ __forceinline__ float plane_coord(float p, float lo, float ext) {
    return __fsub_rn(__fmul_rn(__fdiv_rn(__fsub_rn(p, lo), ext), 2.0f), 1.0f);
}

// packed tap: x = element offset of texel (y0, x0), y = fx bits, z = fy bits,
// w = flags: 1 x1 ok, 2 y1 ok, 4 x grad on, 8 y grad on
__device__ __forceinline__ int4 make_tap(const MnePlaneDesc& P, float gx, float gy) {
    const AxisTap tx = axis_tap(gx, P.W), ty = axis_tap(gy, P.H);
    int4 r;
    r.x = (int)((int64_t)ty.i0 * P.stride_h + (int64_t)tx.i0 * P.stride_w);
    r.y = __float_as_int(tx.f);
    r.z = __float_as_int(ty.f);
    r.w = (tx.ok1 ? 1 : 0) | (ty.ok1 ? 2 : 0) | (tx.grad_on ? 4 : 0) | (ty.grad_on ? 8 : 0);
    return r;
}

// lane = point: write the taps of this lane's point for every plane-level
template <bool TWO>
__device__ __forceinline__ void write_taps(const MneSceneDesc& sc, int4* tab_row, float px, float py, float pz) {
    const float n[3] = {plane_coord(px, sc.bound_lo[0], sc.bound_ext[0]), plane_coord(py, sc.bound_lo[1], sc.bound_ext[1]),
                        plane_coord(pz, sc.bound_lo[2], sc.bound_ext[2])};
#pragma unroll
    for (int grp = 0; grp < (TWO ? 2 : 1); ++grp)
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            const float gx = (a == 2) ? n[1] : n[0];     // xy: (x,y)  xz: (x,z)  yz: (y,z)
            const float gy = (a == 0) ? n[1] : n[2];
#pragma unroll
            for (int l = 0; l < 2; ++l) {
                const int t = grp * 6 + a * 2 + l;
                tab_row[t] = make_tap(sc.planes[t], gx, gy);
            }
        }
}

// lane = channel: gather the features of np points into the IN (and CF) tiles
template <bool TWO>
__device__ __forceinline__ void gather_chunk(const MneSceneDesc& sc, const int4* __restrict__ tab, float* __restrict__ IN,
                                             float* __restrict__ CF, int np, int lane) {
    constexpr int NT = Net<TWO>::N_TAPS;
#pragma unroll 4
    for (int p = 0; p < np; ++p) {
#pragma unroll
        for (int grp = 0; grp < (TWO ? 2 : 1); ++grp) {
#pragma unroll
            for (int l = 0; l < 2; ++l) {
                float v[3];
#pragma unroll
                for (int a = 0; a < 3; ++a) {
                    const int t = grp * 6 + a * 2 + l;
                    const MnePlaneDesc& P = sc.planes[t];
                    const int4 tp = tab[p * NT + t];
                    const float* base = P.data + (int64_t)lane * P.stride_c + tp.x;
                    const int64_t dx = (tp.w & 1) ? P.stride_w : 0, dy = (tp.w & 2) ? P.stride_h : 0;
                    const float fx = __int_as_float(tp.y), fy = __int_as_float(tp.z);
                    const float wx1 = (tp.w & 1) ? fx : 0.0f, wy1 = (tp.w & 2) ? fy : 0.0f;
                    const float wx0 = 1.0f - fx, wy0 = 1.0f - fy;
                    const float t00 = __ldg(base), t01 = __ldg(base + dx), t10 = __ldg(base + dy), t11 = __ldg(base + dy + dx);
                    v[a] = t00 * (wy0 * wx0) + t01 * (wy0 * wx1) + t10 * (wy1 * wx0) + t11 * (wy1 * wx1);
                }
                const float f = (v[0] + v[1]) + v[2];
                if (grp == 0) IN[p * kInStride + l * 32 + lane] = f;
                else CF[p * kCfStride + l * 32 + lane] = f;
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------
// OneBlob (tiny-cuda-nn encodings/oneblob.h, power-of-two bin count; see oracle/oneblob.py)
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ float quartic_cdf(float t) {
    const float u = t * (float)kBins;
    const float u2 = u * u, u4 = u2 * u2;
    return fminf(fmaxf((15.0f / 16.0f) * u * (1.0f - (2.0f / 3.0f) * u2 + (1.0f / 5.0f) * u4) + 0.5f, 0.0f), 1.0f);
}
__device__ __forceinline__ float wrapped_cdf(float t) { return quartic_cdf(t) + quartic_cdf(t - 1.0f) + quartic_cdf(t + 1.0f); }
// d cdf / dt: kBins * 15/16 * (1 - u^2)^2 inside the kernel support, 0 where the cdf is clamped
__device__ __forceinline__ float quartic_pdf(float t) {
    const float u = t * (float)kBins;
    const float w = fmaxf(1.0f - u * u, 0.0f);
    return (float)kBins * (15.0f / 16.0f) * w * w;
}
__device__ __forceinline__ float wrapped_pdf(float t) { return quartic_pdf(t) + quartic_pdf(t - 1.0f) + quartic_pdf(t + 1.0f); }

__device__ __forceinline__ float blob_input(float p, double lo, double ext) { return (float)(((double)p - lo) / ext); }

// 16 bins of one dimension into dst[0..15] (16-byte aligned shared memory)
__device__ __forceinline__ void oneblob16(float x, float* dst) {
    const float first = wrapped_cdf(0.0f - x);
    float left = first;
#pragma unroll
    for (int b4 = 0; b4 < 4; ++b4) {
        float4 o;
        float* op = &o.x;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int b = b4 * 4 + i;
            const float right = (b == kBins - 1) ? first + 1.0f : wrapped_cdf((float)(b + 1) * (1.0f / kBins) - x);
            op[i] = right - left;
            left = right;
        }
        *reinterpret_cast<float4*>(dst + b4 * 4) = o;
    }
}

// sum_b g[b] * d blob_b / dx for one dimension
__device__ __forceinline__ float oneblob16_grad(float x, const float* g) {
    // blob_b = C(e_{b+1} - x) - C(e_b - x)  ->  d/dx = C'(e_b - x) - C'(e_{b+1} - x); bin 15's right edge is e_0
    float acc = 0.0f;
    const float first = wrapped_pdf(0.0f - x);
    float left = first;
#pragma unroll
    for (int b = 0; b < kBins; ++b) {
        const float right = (b == kBins - 1) ? first : wrapped_pdf((float)(b + 1) * (1.0f / kBins) - x);
        acc += g[b] * (left - right);
        left = right;
    }
    return acc;
}

// ---------------------------------------------------------------------------------------------
// lane = point dense helpers.  WT rows are [in][32 out] in shared memory (broadcast 128-bit loads).
// ---------------------------------------------------------------------------------------------
// acc[0..31] += sum_k x[k] * WT[k][:], x = this lane's own tile row (K4 groups of 4 inputs)
template <int K4>
__device__ __forceinline__ void dense32_from_row(float (&acc)[32], const float* __restrict__ x, const float* __restrict__ WT) {
#pragma unroll 1
    for (int k4 = 0; k4 < K4; ++k4) {
        const float4 xv = *reinterpret_cast<const float4*>(x + 4 * k4);
        const float xs[4] = {xv.x, xv.y, xv.z, xv.w};
        const float4* w = reinterpret_cast<const float4*>(WT + (4 * k4) * 32);
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j4 = 0; j4 < 8; ++j4) {
                const float4 ww = w[i * 8 + j4];
                acc[4 * j4 + 0] = fmaf(ww.x, xs[i], acc[4 * j4 + 0]);
                acc[4 * j4 + 1] = fmaf(ww.y, xs[i], acc[4 * j4 + 1]);
                acc[4 * j4 + 2] = fmaf(ww.z, xs[i], acc[4 * j4 + 2]);
                acc[4 * j4 + 3] = fmaf(ww.w, xs[i], acc[4 * j4 + 3]);
            }
    }
}

// acc[0..31] += sum_{k<K} x[k] * WT[k][:], x in registers
template <int K>
__device__ __forceinline__ void dense32_from_regs(float (&acc)[32], const float (&x)[K], const float* __restrict__ WT) {
#pragma unroll
    for (int k = 0; k < K; ++k) {
        const float4* w = reinterpret_cast<const float4*>(WT + k * 32);
#pragma unroll
        for (int j4 = 0; j4 < 8; ++j4) {
            const float4 ww = w[j4];
            acc[4 * j4 + 0] = fmaf(ww.x, x[k], acc[4 * j4 + 0]);
            acc[4 * j4 + 1] = fmaf(ww.y, x[k], acc[4 * j4 + 1]);
            acc[4 * j4 + 2] = fmaf(ww.z, x[k], acc[4 * j4 + 2]);
            acc[4 * j4 + 3] = fmaf(ww.w, x[k], acc[4 * j4 + 3]);
        }
    }
}

// out16[c] = sum_j h[j] * W2T[j][c]   (c: 0..14 geo, 15 sdf)
__device__ __forceinline__ void dense16_from_regs(float (&out)[16], const float (&h)[32], const float* __restrict__ W2T) {
#pragma unroll
    for (int c = 0; c < 16; ++c) out[c] = 0.0f;
#pragma unroll
    for (int j = 0; j < 32; ++j) {
        const float4* w = reinterpret_cast<const float4*>(W2T + j * 16);
#pragma unroll
        for (int c4 = 0; c4 < 4; ++c4) {
            const float4 ww = w[c4];
            out[4 * c4 + 0] = fmaf(ww.x, h[j], out[4 * c4 + 0]);
            out[4 * c4 + 1] = fmaf(ww.y, h[j], out[4 * c4 + 1]);
            out[4 * c4 + 2] = fmaf(ww.z, h[j], out[4 * c4 + 2]);
            out[4 * c4 + 3] = fmaf(ww.w, h[j], out[4 * c4 + 3]);
        }
    }
}

// dot(WT[k][0..31], v[0..31]) with four partial sums to keep the dependency chains short
__device__ __forceinline__ float dot32_row(const float* __restrict__ WTrow, const float (&v)[32]) {
    const float4* w = reinterpret_cast<const float4*>(WTrow);
    float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll
    for (int j4 = 0; j4 < 8; ++j4) {
        const float4 ww = w[j4];
        a0 = fmaf(ww.x, v[4 * j4 + 0], a0);
        a1 = fmaf(ww.y, v[4 * j4 + 1], a1);
        a2 = fmaf(ww.z, v[4 * j4 + 2], a2);
        a3 = fmaf(ww.w, v[4 * j4 + 3], a3);
    }
    return (a0 + a1) + (a2 + a3);
}

// ---------------------------------------------------------------------------------------------
// the decoder forward for this lane's point (lane = point).
// in_row: [feat64 | blob48] tile row; cf_row: colour feature row (two grids).
// Outputs: h1 / hc post-ReLU hidden activations, out16 = [geo15 | sdf], rgb logits.
// ---------------------------------------------------------------------------------------------
template <bool TWO>
__device__ __forceinline__ void decoder_forward(const float* __restrict__ WT, const float* __restrict__ in_row,
                                                const float* __restrict__ cf_row, float (&h1)[32], float (&out16)[16],
                                                float (&hc)[32], float (&rgb)[3]) {
    using N = Net<TWO>;
#pragma unroll
    for (int j = 0; j < 32; ++j) h1[j] = 0.0f;
    dense32_from_row<N::IN_SDF / 4>(h1, in_row, WT + N::OFF_W1T);
#pragma unroll
    for (int j = 0; j < 32; ++j) h1[j] = fmaxf(h1[j], 0.0f);
    dense16_from_regs(out16, h1, WT + N::OFF_W2T);

#pragma unroll
    for (int j = 0; j < 32; ++j) hc[j] = 0.0f;
    dense32_from_row<kBlob / 4>(hc, in_row + N::F, WT + N::OFF_V1T);
    if (TWO) dense32_from_row<N::F / 4>(hc, cf_row, WT + N::OFF_V1T + kBlob * 32);
    {
        float geo[kGeo];
#pragma unroll
        for (int m = 0; m < kGeo; ++m) geo[m] = out16[m];
        dense32_from_regs<kGeo>(hc, geo, WT + N::OFF_V1T + N::GEO_ROW0 * 32);
    }
#pragma unroll
    for (int j = 0; j < 32; ++j) hc[j] = fmaxf(hc[j], 0.0f);
    rgb[0] = rgb[1] = rgb[2] = 0.0f;
#pragma unroll
    for (int j = 0; j < 32; ++j) {
        const float4 w = *reinterpret_cast<const float4*>(WT + N::OFF_V2T + j * 4);
        rgb[0] = fmaf(w.x, hc[j], rgb[0]);
        rgb[1] = fmaf(w.y, hc[j], rgb[1]);
        rgb[2] = fmaf(w.z, hc[j], rgb[2]);
    }
}

// ---------------------------------------------------------------------------------------------
// compositing (reference model/scene_rep.py:183-230), one warp per ray, lanes stride over samples.
// raw / z are read with ld.global.cg: they were written by this warp moments ago.
// ---------------------------------------------------------------------------------------------
struct RayComposite {
    int ind;          // first sign change (0 if none), sdf2weights :195-197
    float thresh;     // z[ind] + sc_factor * trunc
    float inv_norm;   // 1 / (sum u*m + 1e-8)
    float rgb[3];
    float depth;
    float acc;        // sum of weights
};

__device__ __forceinline__ float4 load_raw(const float* raw, size_t pt) {
    return __ldcg(reinterpret_cast<const float4*>(raw) + pt);
}

// u = sigmoid(s/tr) * sigmoid(-s/tr)
__device__ __forceinline__ float sdf_bell(float sdf, float trunc, float* sig_out = nullptr) {
    const float x = __fdiv_rn(sdf, trunc);
    const float sp = sigmoidf(x), sn = sigmoidf(-x);
    if (sig_out) *sig_out = sp;
    return sp * sn;
}

__device__ __forceinline__ RayComposite composite_ray(const float* raw, const float* z_vals, size_t base, int S,
                                                      const MneRenderCfg& c, int lane) {
    RayComposite r;
    int first = 0x7fffffff;
    for (int s = lane; s + 1 < S; s += 32) {
        const float a = __ldcg(raw + (base + s) * 4 + 3), b = __ldcg(raw + (base + s + 1) * 4 + 3);
        if (__fmul_rn(b, a) < 0.0f) first = min(first, s);
    }
    first = warp_min(first);
    r.ind = (first == 0x7fffffff) ? 0 : first;
    r.thresh = __fadd_rn(__ldcg(z_vals + base + r.ind), c.sc_trunc);
    float sw = 0.f, sr = 0.f, sg = 0.f, sb = 0.f, sz = 0.f;
    for (int s = lane; s < S; s += 32) {
        const float z = __ldcg(z_vals + base + s);
        if (z < r.thresh) {
            const float4 v = load_raw(raw, base + s);
            const float u = sdf_bell(v.w, c.trunc);
            sw += u;
            sr += u * sigmoidf(v.x);
            sg += u * sigmoidf(v.y);
            sb += u * sigmoidf(v.z);
            sz += u * z;
        }
    }
    sw = warp_sum(sw); sr = warp_sum(sr); sg = warp_sum(sg); sb = warp_sum(sb); sz = warp_sum(sz);
    r.inv_norm = 1.0f / (sw + 1e-8f);
    r.rgb[0] = sr * r.inv_norm;
    r.rgb[1] = sg * r.inv_norm;
    r.rgb[2] = sb * r.inv_norm;
    r.depth = sz * r.inv_norm;
    r.acc = sw * r.inv_norm;
    return r;
}

}  // namespace mne
