// v2 building blocks: a warp owns a tile of 16 points and runs the decoder on the tensor pipe.
//
//   * mma.sync.m16n8k8 TF32, fp32 accumulate.  The forward uses the 3-term split (x = hi + lo, products
//     lo*hi + hi*lo + hi*hi) which carries ~21 mantissa bits, i.e. fp32-level accuracy, so the 1e-3 parity
//     gate holds; gradients (cosine >= 0.999 gate) use single TF32 products.
//   * fragment chaining: the accumulator fragment of one layer IS the A fragment of the next layer when the
//     8 k-slots of a k-step are mapped to columns (2t, 2t+1) instead of (t, t+4); the weights (B operand) are
//     stored with the same pairing, so activations never leave registers between layers.
//   * per-warp shared memory: one [16][120] fp32 tile (feat64 | blob48); the bilinear tap table of a point
//     aliases that point's blob segment while the gather runs (192 B = 6 taps x (4 offsets + 4 weights)).
#pragma once
#include <cuda_fp16.h>
#include "render_common.cuh"

namespace mne {

constexpr int kTilePts = 16;
constexpr int kXStride = 120;      // floats per tile row; == 8 (mod 32): conflict-free 64-bit A-fragment loads
constexpr int kCStride = 72;       // colour-feature tile row (two grids)
constexpr int kBlobCol = 64;

__device__ __forceinline__ uint32_t to_tf32(float x) {
    uint32_t r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
    return r;
}

__device__ __forceinline__ void mma_tf32(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

struct SplitA {
    uint32_t hi[4], lo[4];
};
__device__ __forceinline__ SplitA split_a(float a0, float a1, float a2, float a3) {
    SplitA s;
    const float v[4] = {a0, a1, a2, a3};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        s.hi[i] = to_tf32(v[i]);
        s.lo[i] = to_tf32(v[i] - __uint_as_float(s.hi[i]));
    }
    return s;
}
// d += A * B with the 3-term split; w = (b0_hi, b1_hi, b0_lo, b1_lo)
__device__ __forceinline__ void mma_3x(float (&d)[4], const SplitA& a, const float4& w) {
    mma_tf32(d, a.lo, __float_as_uint(w.x), __float_as_uint(w.y));
    mma_tf32(d, a.hi, __float_as_uint(w.z), __float_as_uint(w.w));
    mma_tf32(d, a.hi, __float_as_uint(w.x), __float_as_uint(w.y));
}

// ---------------------------------------------------------------------------------------------
// forward weight layout in shared memory: for each output n a row of float4 pairs
//   W4[n][kp] = (hi W[n][2kp], hi W[n][2kp+1], lo W[n][2kp], lo W[n][2kp+1])
// row strides (in float4) are == 4 (mod 8): conflict-free 128-bit B-fragment loads.
// ---------------------------------------------------------------------------------------------
template <bool TWO>
struct NetV2 {
    static constexpr int K1 = 112, K1P = 56, RS1 = 60;                    // sdf0: [32][112]
    static constexpr int K2P = 16, RS2 = 20;                              // sdf1: [16][32], rows = (geo0..14, sdf)
    static constexpr int KC = kBlob + (TWO ? 64 : 0) + 16;                // col0 inputs: blob | cfeat | geo15 + 0
    static constexpr int KCP = KC / 2, RSC = KCP + 4;                     // 32+4 = 36, 64+4 = 68
    static constexpr int IN_COL = kBlob + (TWO ? 64 : 0) + kGeo;          // real columns of col0
    static constexpr int OFF1 = 0;
    static constexpr int OFF2 = OFF1 + 32 * RS1;
    static constexpr int OFFC = OFF2 + 16 * RS2;
    static constexpr int OFFO = OFFC + 32 * RSC;                          // col1: [8][32] (3 real rows)
    static constexpr int W4_COUNT = OFFO + 8 * RS2;                       // float4 elements
};

// Column of sdf0 that feeds tile column k.  Planes: identity (64 features | 48 blob).  Hash grid: the table
// yields 32 features, tile columns 32..63 are zero padding (their weights are zero), blob follows at 64.
__device__ __forceinline__ int sdf0_col(int k, bool hash) { return !hash ? k : (k < 32 ? k : (k < 64 ? -1 : k - 32)); }
__device__ __forceinline__ int sdf0_in(bool hash) { return hash ? 80 : 112; }
// The same for col0 with two grids: tile columns = blob 48 | colour features 64 | geo 15.  A colour hash table yields 32
// features: tile columns 80..111 are zero padding, geo follows at 112.
__device__ __forceinline__ int col0_col(int k, bool two, bool hash) {
    return !(two && hash) ? k : (k < 80 ? k : (k < 112 ? -1 : k - 32));
}
__device__ __forceinline__ int col0_in(bool two, bool hash) { return kBlob + kGeo + (two ? (hash ? 32 : 64) : 0); }

__device__ __forceinline__ float4 split_pair(float a, float b) {
    const uint32_t ah = to_tf32(a), bh = to_tf32(b);
    return make_float4(__uint_as_float(ah), __uint_as_float(bh), __uint_as_float(to_tf32(a - __uint_as_float(ah))),
                       __uint_as_float(to_tf32(b - __uint_as_float(bh))));
}

template <bool TWO>
__device__ __forceinline__ void load_weights_v2(const MneSceneDesc& sc, float4* W4, int tid, int nthreads) {
    using N = NetV2<TWO>;
    const bool hash = sc.encoder == 1;
    for (int i = tid; i < 32 * N::K1P; i += nthreads) {
        const int n = i / N::K1P, kp = i % N::K1P;
        const int c0 = sdf0_col(2 * kp, hash), c1 = sdf0_col(2 * kp + 1, hash);
        W4[N::OFF1 + n * N::RS1 + kp] = split_pair(c0 >= 0 ? sc.w_sdf0[n * sdf0_in(hash) + c0] : 0.0f,
                                                   c1 >= 0 ? sc.w_sdf0[n * sdf0_in(hash) + c1] : 0.0f);
    }
    for (int i = tid; i < 16 * N::K2P; i += nthreads) {
        const int n = i / N::K2P, kp = i % N::K2P;
        const int out = (n == 15) ? 0 : n + 1;                      // tile column 15 carries the sdf
        W4[N::OFF2 + n * N::RS2 + kp] = split_pair(sc.w_sdf1[out * 32 + 2 * kp], sc.w_sdf1[out * 32 + 2 * kp + 1]);
    }
    for (int i = tid; i < 32 * N::KCP; i += nthreads) {
        const int n = i / N::KCP, kp = i % N::KCP;
        const int k0 = 2 * kp, k1 = 2 * kp + 1;
        const int c0 = (k0 < N::IN_COL) ? col0_col(k0, TWO, hash) : -1, c1 = (k1 < N::IN_COL) ? col0_col(k1, TWO, hash) : -1;
        const int in = col0_in(TWO, hash);
        const float a = (c0 >= 0) ? sc.w_col0[n * in + c0] : 0.0f;
        const float b = (c1 >= 0) ? sc.w_col0[n * in + c1] : 0.0f;
        W4[N::OFFC + n * N::RSC + kp] = split_pair(a, b);
    }
    for (int i = tid; i < 8 * N::K2P; i += nthreads) {
        const int n = i / N::K2P, kp = i % N::K2P;
        const float a = (n < 3) ? sc.w_col1[n * 32 + 2 * kp] : 0.0f, b = (n < 3) ? sc.w_col1[n * 32 + 2 * kp + 1] : 0.0f;
        W4[N::OFFO + n * N::RS2 + kp] = split_pair(a, b);
    }
}

// ---------------------------------------------------------------------------------------------
// tap table: per (point, plane-level) 4 texel offsets + 4 bilinear weights (32 B)
// ---------------------------------------------------------------------------------------------
struct Tap8 {
    int4 off;      // element offsets of (y0,x0) (y0,x1) (y1,x0) (y1,x1); out-of-plane corners repeat a valid one
    float4 w;      // their weights (0 for out-of-plane corners)
};

__device__ __forceinline__ void make_tap8(const MnePlaneDesc& P, float gx, float gy, int4& off, float4& w, unsigned* flags = nullptr) {
    const AxisTap tx = axis_tap(gx, P.W), ty = axis_tap(gy, P.H);
    const int o00 = (int)((int64_t)ty.i0 * P.stride_h + (int64_t)tx.i0 * P.stride_w);
    const int dx = tx.ok1 ? (int)P.stride_w : 0, dy = ty.ok1 ? (int)P.stride_h : 0;
    const float wx1 = tx.ok1 ? tx.f : 0.0f, wy1 = ty.ok1 ? ty.f : 0.0f;
    const float wx0 = 1.0f - tx.f, wy0 = 1.0f - ty.f;
    off = make_int4(o00, o00 + dx, o00 + dy, o00 + dy + dx);
    w = make_float4(wy0 * wx0, wy0 * wx1, wy1 * wx0, wy1 * wx1);
    if (flags) *flags = (tx.grad_on ? 1u : 0u) | (ty.grad_on ? 2u : 0u);
}

// taps of one point for the planes [t_begin, t_end) into its table segment (float4 slots 2t, 2t+1)
template <bool TWO>
__device__ __forceinline__ void write_taps8(const MneSceneDesc& sc, float4* seg, float px, float py, float pz, int grp) {
    const float n[3] = {plane_coord(px, sc.bound_lo[0], sc.bound_ext[0]), plane_coord(py, sc.bound_lo[1], sc.bound_ext[1]),
                        plane_coord(pz, sc.bound_lo[2], sc.bound_ext[2])};
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const float gx = (a == 2) ? n[1] : n[0];
        const float gy = (a == 0) ? n[1] : n[2];
#pragma unroll
        for (int l = 0; l < 2; ++l) {
            const int t = a * 2 + l;
            int4 off; float4 w;
            make_tap8(sc.planes[grp * 6 + t], gx, gy, off, w);
            seg[2 * t] = make_float4(__int_as_float(off.x), __int_as_float(off.y), __int_as_float(off.z), __int_as_float(off.w));
            seg[2 * t + 1] = w;
        }
    }
}

// lane = channel: features of np points of one plane group into tile columns [0, 64)
__device__ __forceinline__ void gather16(const MneSceneDesc& sc, int grp, const float* __restrict__ tab_base, int tab_stride,
                                         float* __restrict__ dst, int dst_stride, int np, int lane) {
    const float* base[6];
#pragma unroll
    for (int t = 0; t < 6; ++t) base[t] = sc.planes[grp * 6 + t].data + (int64_t)lane * sc.planes[grp * 6 + t].stride_c;
#pragma unroll 4
    for (int p = 0; p < np; ++p) {
        const float4* seg = reinterpret_cast<const float4*>(tab_base + p * tab_stride);
#pragma unroll
        for (int l = 0; l < 2; ++l) {
            float v[3];
#pragma unroll
            for (int a = 0; a < 3; ++a) {
                const int t = a * 2 + l;
                const float4 o = seg[2 * t], w = seg[2 * t + 1];
                const float* b = base[t];
                v[a] = __ldg(b + __float_as_uint(o.x)) * w.x + __ldg(b + __float_as_uint(o.y)) * w.y +
                       __ldg(b + __float_as_uint(o.z)) * w.z + __ldg(b + __float_as_uint(o.w)) * w.w;
            }
            dst[p * dst_stride + l * 32 + lane] = (v[0] + v[1]) + v[2];
        }
    }
}

// Channels-last fast path: lane = (corner, 4-channel group).  One 128-bit load per lane fetches the four
// corner texels of a tap (4 x 128 B lines) in a single warp instruction; the per-corner partial sums of the
// three planes of a level are then reduced across the four corner lane-groups with two shuffle steps.
// Needs stride_c == 1 and 16-byte aligned texels (checked on the host: planes_vectorizable()).
__device__ __forceinline__ void gather16_vec(const MneSceneDesc& sc, int grp, const float* __restrict__ tab_base, int tab_stride,
                                             float* __restrict__ dst, int dst_stride, int np, int lane) {
    const int c = lane >> 3, q = lane & 7;
    const float* base[6];
#pragma unroll
    for (int t = 0; t < 6; ++t) base[t] = sc.planes[grp * 6 + t].data + q * 4;
#pragma unroll 4
    for (int p = 0; p < np; ++p) {
        const float* seg = tab_base + p * tab_stride;
        float4 acc[2];
#pragma unroll
        for (int l = 0; l < 2; ++l) {
            acc[l] = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int a = 0; a < 3; ++a) {
                const int t = a * 2 + l;
                const unsigned off = __float_as_uint(seg[8 * t + c]);
                const float w = seg[8 * t + 4 + c];
                const float4 v = __ldg(reinterpret_cast<const float4*>(base[t] + off));
                acc[l].x = fmaf(w, v.x, acc[l].x); acc[l].y = fmaf(w, v.y, acc[l].y);
                acc[l].z = fmaf(w, v.z, acc[l].z); acc[l].w = fmaf(w, v.w, acc[l].w);
            }
        }
#pragma unroll
        for (int l = 0; l < 2; ++l) {
#pragma unroll
            for (int o = 8; o <= 16; o <<= 1) {
                acc[l].x += __shfl_xor_sync(0xffffffffu, acc[l].x, o); acc[l].y += __shfl_xor_sync(0xffffffffu, acc[l].y, o);
                acc[l].z += __shfl_xor_sync(0xffffffffu, acc[l].z, o); acc[l].w += __shfl_xor_sync(0xffffffffu, acc[l].w, o);
            }
            if (c == 0) *reinterpret_cast<float4*>(dst + p * dst_stride + l * 32 + q * 4) = acc[l];
        }
    }
}

// taps of one point for ONE level (the three planes xy / xz / yz of that level): the two lanes of a point each take a level
__device__ __forceinline__ void write_taps8_level(const MneSceneDesc& sc, float4* seg, float px, float py, float pz, int grp, int level) {
    const float n[3] = {plane_coord(px, sc.bound_lo[0], sc.bound_ext[0]), plane_coord(py, sc.bound_lo[1], sc.bound_ext[1]),
                        plane_coord(pz, sc.bound_lo[2], sc.bound_ext[2])};
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const float gx = (a == 2) ? n[1] : n[0];
        const float gy = (a == 0) ? n[1] : n[2];
        const int t = a * 2 + level;
        int4 off; float4 w;
        make_tap8(sc.planes[grp * 6 + t], gx, gy, off, w);
        seg[2 * t] = make_float4(__int_as_float(off.x), __int_as_float(off.y), __int_as_float(off.z), __int_as_float(off.w));
        seg[2 * t + 1] = w;
    }
}

// Channels-last gather, lane = (point of a quad, 4-channel group): a lane fetches the four corner texels of every tap
// of ITS point (four 128-bit loads per tap; one warp instruction covers 4 points x one 128-byte texel line each) and
// blends them in registers in the reference's order (per plane nw, ne, sw, se; then (xy + xz) + yz) -- no cross-lane
// reduction, two 128-bit tap-table reads per tap.  `save` (optional): the features are also written to global memory,
// row p at save + p * save_stride (the backward's weight-gradient kernel reads them instead of gathering again).
__device__ __forceinline__ void gather16_pq(const MneSceneDesc& sc, int grp, const float* __restrict__ tab_base, int tab_stride,
                                            float* __restrict__ dst, int dst_stride, int np, int lane,
                                            float* __restrict__ save = nullptr, int save_stride = 0) {
    const int pp = lane >> 3, q = lane & 7;
    const float* base[6];
#pragma unroll
    for (int t = 0; t < 6; ++t) base[t] = sc.planes[grp * 6 + t].data + q * 4;
#pragma unroll 1
    for (int p0 = 0; p0 < kTilePts; p0 += 4) {
        const int p = p0 + pp;
        if (p >= np) continue;
        const float* seg = tab_base + p * tab_stride;
#pragma unroll
        for (int l = 0; l < 2; ++l) {
            float4 s[3];
#pragma unroll
            for (int a = 0; a < 3; ++a) {
                const int t = a * 2 + l;
                const uint4 off = *reinterpret_cast<const uint4*>(seg + 8 * t);
                const float4 w = *reinterpret_cast<const float4*>(seg + 8 * t + 4);
                const float4 v00 = __ldg(reinterpret_cast<const float4*>(base[t] + off.x));
                const float4 v01 = __ldg(reinterpret_cast<const float4*>(base[t] + off.y));
                const float4 v10 = __ldg(reinterpret_cast<const float4*>(base[t] + off.z));
                const float4 v11 = __ldg(reinterpret_cast<const float4*>(base[t] + off.w));
                s[a].x = fmaf(w.w, v11.x, fmaf(w.z, v10.x, fmaf(w.y, v01.x, w.x * v00.x)));
                s[a].y = fmaf(w.w, v11.y, fmaf(w.z, v10.y, fmaf(w.y, v01.y, w.x * v00.y)));
                s[a].z = fmaf(w.w, v11.z, fmaf(w.z, v10.z, fmaf(w.y, v01.z, w.x * v00.z)));
                s[a].w = fmaf(w.w, v11.w, fmaf(w.z, v10.w, fmaf(w.y, v01.w, w.x * v00.w)));
            }
            const float4 f = make_float4((s[0].x + s[1].x) + s[2].x, (s[0].y + s[1].y) + s[2].y,
                                         (s[0].z + s[1].z) + s[2].z, (s[0].w + s[1].w) + s[2].w);
            *reinterpret_cast<float4*>(dst + p * dst_stride + l * 32 + q * 4) = f;
            if (save) *reinterpret_cast<float4*>(save + (int64_t)p * save_stride + l * 32 + q * 4) = f;
        }
    }
}

// ---------------------------------------------------------------------------------------------
// multi-resolution hash grid (tiny-cuda-nn encodings/grid.h; restated in oracle/hashgrid.py).
// lane = (point parity, level): 16 levels x 2 points per trip, 8 corner loads of one float2 (2 features) each.
// Point coordinates in [0,1] (the OneBlob inputs) are read from xn[p * xn_stride + 0..2].
// ---------------------------------------------------------------------------------------------
struct HashLevel {
    float scale;
    uint32_t res, size, offset;
};
__device__ __forceinline__ HashLevel hash_level(const MneHashDesc& h, int l) {
    HashLevel L;
    L.scale = h.level_scale[l];
    L.res = h.level_res[l];
    L.offset = h.offsets[l];
    L.size = h.offsets[l + 1] - h.offsets[l];
    return L;
}
__device__ __forceinline__ uint32_t hash_index(const HashLevel& L, uint32_t gx, uint32_t gy, uint32_t gz) {
    uint64_t stride = 1;
    uint32_t index = 0;
    if (stride <= L.size) { index += gx * (uint32_t)stride; stride *= L.res; }
    if (stride <= L.size) { index += gy * (uint32_t)stride; stride *= L.res; }
    if (stride <= L.size) { index += gz * (uint32_t)stride; stride *= L.res; }
    if (L.size < stride) index = gx ^ (gy * 2654435761u) ^ (gz * 805459861u);
    return index % L.size;
}
struct HashCell {
    uint32_t g[3];
    float f[3];
};
__device__ __forceinline__ HashCell hash_cell(const HashLevel& L, const float* x) {
    HashCell c;
#pragma unroll
    for (int d = 0; d < 3; ++d) {
        const float pos = fmaf(x[d], L.scale, 0.5f);          // one rounding, as tiny-cuda-nn's pos_fract
        const float fl = floorf(pos);
        c.f[d] = pos - fl;
        c.g[d] = (uint32_t)(int)fl;
    }
    return c;
}
__device__ __forceinline__ float corner_weight(const HashCell& c, int k) {
    return ((k & 1) ? c.f[0] : 1.0f - c.f[0]) * ((k & 2) ? c.f[1] : 1.0f - c.f[1]) * ((k & 4) ? c.f[2] : 1.0f - c.f[2]);
}

// one table entry (2 features): from the fp16 copy of the table when the descriptor carries one (half2 load, 4 bytes per
// entry -- tiny-cuda-nn's default parameter precision, the 1e-2 parity gate of BASELINE.json), else from the fp32 table
__device__ __forceinline__ float2 hash_entry(const MneHashDesc& h, uint32_t level_offset, uint32_t idx) {
    if (h.table_h) return __half22float2(__ldg(reinterpret_cast<const __half2*>(h.table_h) + level_offset + idx));
    return __ldg(reinterpret_cast<const float2*>(h.table) + level_offset + idx);
}

// lvl0: the entries of level 0 (the coarsest, dense level: 17^3 entries = 39 KB) staged in shared memory by the CTA, or
// NULL.  Every sample reads 8 of them; the other levels are served by L2 / HBM.
__device__ __forceinline__ void stage_hash_level0(const MneHashDesc& h, float2* lvl0, int tid, int nthreads) {
    const uint32_t n = h.offsets[1] - h.offsets[0];
    for (uint32_t i = tid; i < n; i += nthreads) lvl0[i] = hash_entry(h, h.offsets[0], i);
}

// SRC: 0 = fp32 table, 1 = fp16 copy (half2 loads), 2 = shared-memory copy of the level (float2)
template <int SRC>
__device__ __forceinline__ void gather16_hash_from(const HashLevel& L, const void* __restrict__ entries, const float* __restrict__ xn,
                                                   int xn_stride, float* __restrict__ dst, int dst_stride, int np, int l, int pp) {
#pragma unroll 2
    for (int it = 0; it < kTilePts / 2; ++it) {
        const int p = 2 * it + pp;
        if (p < np) {
            const HashCell c = hash_cell(L, xn + p * xn_stride);
            float2 acc = make_float2(0.f, 0.f);
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                const uint32_t idx = hash_index(L, c.g[0] + (k & 1), c.g[1] + ((k >> 1) & 1), c.g[2] + ((k >> 2) & 1));
                float2 v;
                if (SRC == 0) v = __ldg(reinterpret_cast<const float2*>(entries) + idx);
                else if (SRC == 1) v = __half22float2(__ldg(reinterpret_cast<const __half2*>(entries) + idx));
                else v = reinterpret_cast<const float2*>(entries)[idx];
                const float w = corner_weight(c, k);
                acc.x = fmaf(w, v.x, acc.x); acc.y = fmaf(w, v.y, acc.y);
            }
            *reinterpret_cast<float2*>(dst + p * dst_stride + 2 * l) = acc;
        }
    }
}

__device__ __forceinline__ void gather16_hash(const MneHashDesc& h, const float* __restrict__ xn, int xn_stride,
                                              float* __restrict__ dst, int dst_stride, int np, int lane,
                                              const float2* __restrict__ lvl0 = nullptr) {
    const int l = lane & 15, pp = lane >> 4;
    const HashLevel L = hash_level(h, l);
    if (lvl0 != nullptr && l == 0) gather16_hash_from<2>(L, lvl0, xn, xn_stride, dst, dst_stride, np, l, pp);
    else if (h.table_h) gather16_hash_from<1>(L, reinterpret_cast<const __half2*>(h.table_h) + L.offset, xn, xn_stride, dst, dst_stride, np, l, pp);
    else gather16_hash_from<0>(L, reinterpret_cast<const float2*>(h.table) + L.offset, xn, xn_stride, dst, dst_stride, np, l, pp);
    for (int p = 0; p < kTilePts; ++p) dst[p * dst_stride + 32 + lane] = 0.0f;      // padding columns 32..63
}

// flag store of the multi-GPU dirty map.  Through inline PTX on purpose: a plain store through a uint8_t pointer may alias
// anything, which made the compiler reload the tile from shared memory after every flag (+120 us on the backward pass).
__device__ __forceinline__ void mark_dirty(uint8_t* map, int64_t texel) {
    asm volatile("st.global.u8 [%0], %1;" :: "l"(map + texel), "r"(1));
}

// gradient of the table: grad[corner] += w * d feat (one 64-bit vector reduction per corner).
// Warp-aggregated where it pays: the two lanes of a level hold two CONSECUTIVE samples of the list (neighbours on a ray),
// which at the coarse levels mostly fall into the same cell -- when a corner's entry is the same for both, the lower lane
// adds the two contributions and issues ONE reduction (levels >= kHashAggLevels have cells smaller than the sample
// spacing: no duplicates to find, so they skip the three shuffles per corner).
constexpr int kHashAggLevels = 8;
__device__ __forceinline__ void scatter16_hash(const MneHashDesc& h, const float* __restrict__ xn, int xn_stride,
                                               const float* __restrict__ src, int src_stride, int np, int lane,
                                               uint8_t* dirty = nullptr, const float* dirty_base = nullptr) {
    const int l = lane & 15, pp = lane >> 4;
    const HashLevel L = hash_level(h, l);
    float2* grad = reinterpret_cast<float2*>(h.grad) + L.offset;
    for (int it = 0; it < kTilePts / 2; ++it) {
        const int p = 2 * it + pp;
        const bool on = p < np;
        HashCell c;
        float2 g2 = make_float2(0.f, 0.f);
        if (on) {
            c = hash_cell(L, xn + p * xn_stride);
            g2 = *reinterpret_cast<const float2*>(src + p * src_stride + 2 * l);
        } else {
            c.g[0] = c.g[1] = c.g[2] = 0u; c.f[0] = c.f[1] = c.f[2] = 0.f;
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const uint32_t idx = hash_index(L, c.g[0] + (k & 1), c.g[1] + ((k >> 1) & 1), c.g[2] + ((k >> 2) & 1));
            const float w = corner_weight(c, k);
            float2 v = make_float2(w * g2.x, w * g2.y);
            bool mine = on;
            // (all 32 lanes take part in the shuffles; lanes of the fine levels get their own values back and ignore them)
            const uint32_t o_idx = __shfl_xor_sync(0xffffffffu, on ? idx : 0xffffffffu, 16);
            const float ox = __shfl_xor_sync(0xffffffffu, v.x, 16), oy = __shfl_xor_sync(0xffffffffu, v.y, 16);
            if (l < kHashAggLevels && on && o_idx == idx) {          // the other sample of the pair hits the same entry
                if (pp == 0) { v.x += ox; v.y += oy; } else mine = false;
            }
            if (mine) {
                atomicAdd(grad + idx, v);
                if (dirty) mark_dirty(dirty, (reinterpret_cast<const float*>(grad + idx) - dirty_base) >> 5);
            }
        }
    }
}

// d loss / d (unit-cube coordinates) through the table lookup, for the pose path: the interpolation weights are linear in
// the fractional position, so d feat / d x_d = scale * sum_k (+-1)_d(k) * prod_{d' != d} w_d'(k) * T[k]  (tiny-cuda-nn
// kernel_grid_backward_input; the indices carry no gradient).  lane = (point parity, level); the 16 levels of a point
// are reduced by shuffles and handed to the lanes whose tile row is that point (out[] is ADDED to).
__device__ __forceinline__ void hash_point_grad16(const MneHashDesc& h, const float* __restrict__ xn, int xn_stride,
                                                  const float* __restrict__ src, int src_stride, int np, int lane, float (&out)[3]) {
    const int l = lane & 15, pp = lane >> 4;
    const HashLevel L = hash_level(h, l);
    for (int it = 0; it < kTilePts / 2; ++it) {
        const int p = 2 * it + pp;
        float d[3] = {0.f, 0.f, 0.f};
        if (p < np) {
            const HashCell c = hash_cell(L, xn + p * xn_stride);
            const float2 g2 = *reinterpret_cast<const float2*>(src + p * src_stride + 2 * l);
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                const uint32_t idx = hash_index(L, c.g[0] + (k & 1), c.g[1] + ((k >> 1) & 1), c.g[2] + ((k >> 2) & 1));
                const float2 v = hash_entry(h, L.offset, idx);
                const float D = g2.x * v.x + g2.y * v.y;
                const float wx = (k & 1) ? c.f[0] : 1.0f - c.f[0], wy = (k & 2) ? c.f[1] : 1.0f - c.f[1], wz = (k & 4) ? c.f[2] : 1.0f - c.f[2];
                d[0] += ((k & 1) ? D : -D) * (wy * wz);
                d[1] += ((k & 2) ? D : -D) * (wx * wz);
                d[2] += ((k & 4) ? D : -D) * (wx * wy);
            }
            d[0] *= L.scale; d[1] *= L.scale; d[2] *= L.scale;
        }
#pragma unroll
        for (int a = 0; a < 3; ++a) {
#pragma unroll
            for (int o = 1; o <= 8; o <<= 1) d[a] += __shfl_xor_sync(0xffffffffu, d[a], o);
            const float even = __shfl_sync(0xffffffffu, d[a], 0), odd = __shfl_sync(0xffffffffu, d[a], 16);
            if (l == 2 * it) out[a] += even;
            if (l == 2 * it + 1) out[a] += odd;
        }
    }
}

__host__ inline bool planes_vectorizable(const MneSceneDesc& sc) {
    const int n = sc.two_grid ? 12 : 6;
    for (int i = 0; i < n; ++i) {
        const MnePlaneDesc& p = sc.planes[i];
        if (p.stride_c != 1 || (p.stride_w & 3) || (p.stride_h & 3) || ((uintptr_t)p.data & 15)) return false;
        if (p.grad && ((uintptr_t)p.grad & 15)) return false;
    }
    return true;
}

// ---------------------------------------------------------------------------------------------
// OneBlob, sparse evaluation: each of the three wrap images of the quartic kernel touches at most three
// bins; everything else is exactly 0.  Equal to oneblob16 (render_common.cuh) up to summation order.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void oneblob16_sparse(float x, float* o) {
#pragma unroll
    for (int i = 0; i < 4; ++i) reinterpret_cast<float4*>(o)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int s = 0; s < 3; ++s) {
        const float sh = (s == 0) ? 0.0f : ((s == 1) ? -1.0f : 1.0f);
        const float c = x - sh;                                  // centre of this image of the kernel
        const float fb = floorf(c * (float)kBins);
        if (fb < -2.0f || fb > (float)kBins + 1.0f) continue;    // image entirely outside the bins
        const int bc = (int)fb;
        float F[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) F[j] = quartic_cdf(((float)(bc - 1 + j) * (1.0f / kBins) - x) + sh);
#pragma unroll
        for (int j = 0; j < 3; ++j) {
            const int b = bc - 1 + j;
            if (b >= 0 && b < kBins) o[b] += F[j + 1] - F[j];
        }
    }
    // power-of-two variant of tiny-cuda-nn: the last bin's right edge is "left edge of bin 0, plus one"
    o[kBins - 1] += 1.0f - (quartic_cdf((1.0f - x) + 1.0f) - quartic_cdf((0.0f - x) - 1.0f));
}

// sum_b gb[b] * d blob_b / dx, sparse like oneblob16_sparse (gb: 16 upstream values in shared memory)
__device__ __forceinline__ float oneblob16_sparse_grad(float x, const float* gb) {
    float acc = 0.0f;
#pragma unroll
    for (int s = 0; s < 3; ++s) {
        const float sh = (s == 0) ? 0.0f : ((s == 1) ? -1.0f : 1.0f);
        const float c = x - sh;
        const float fb = floorf(c * (float)kBins);
        if (fb < -2.0f || fb > (float)kBins + 1.0f) continue;
        const int bc = (int)fb;
        float f[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) f[j] = quartic_pdf(((float)(bc - 1 + j) * (1.0f / kBins) - x) + sh);
#pragma unroll
        for (int j = 0; j < 3; ++j) {
            const int b = bc - 1 + j;
            if (b >= 0 && b < kBins) acc += gb[b] * (f[j] - f[j + 1]);     // d/dx cdf(e - x + sh) = -pdf
        }
    }
    acc += gb[kBins - 1] * (quartic_pdf((1.0f - x) + 1.0f) - quartic_pdf((0.0f - x) - 1.0f));
    return acc;
}

// ---------------------------------------------------------------------------------------------
// Per-sample record the forward keeps for the backward (optional; what autograd would save).  One row per sample,
// floats:   h1[32] | o16[16] = (geo0..14, sdf) | hc[32] | feat[F] | cfeat[F] (two grids) | dY[80] | masks[2] | blob x[3] | pad[3]
// with F = 64 (planes) or 32 (hash grid).  h1 / hc are post-ReLU; masks = the bit sets (h1 > 0), (hc > 0) as two 32-bit
// words (all the backward tile kernel needs of them); dY = d h1[32] | d o16[16] | d hc[32] is WRITTEN by the backward tile
// kernel and read, with everything else, by the weight-gradient kernel.  Rows are 32-byte aligned (stride % 8 == 0).
// ---------------------------------------------------------------------------------------------
template <bool TWO, bool HASH>
struct ActRow {
    static constexpr int F = HASH ? 32 : 64;
    static constexpr int H1 = 0, O16 = 32, HC = 48, FEAT = 80, CFEAT = 80 + F;
    static constexpr int DY = 80 + F + (TWO ? F : 0);          // d h1 at DY, d o16 at DY + 32, d hc at DY + 48
    static constexpr int MASK = DY + 80;
    static constexpr int BLOBX = MASK + 2;                     // the three OneBlob inputs (unit-cube coordinates) of the sample
    static constexpr int STRIDE = MASK + 8;
};
__host__ __device__ inline int act_row_floats(bool two, bool hash) {
    const int F = hash ? 32 : 64;
    return 80 + F + (two ? F : 0) + 80 + 8;
}

// fragment blocks -> row columns: lane (g, t) holds rows g / g + 8, columns n * 8 + 2t, +1 of every 8-wide block
template <int NB>
__device__ __forceinline__ void store_frag_rows(float* __restrict__ lo, float* __restrict__ hi, bool ok_lo, bool ok_hi, int t,
                                                const float (&c)[NB][4]) {
#pragma unroll
    for (int n = 0; n < NB; ++n) {
        if (ok_lo) *reinterpret_cast<float2*>(lo + n * 8 + 2 * t) = make_float2(c[n][0], c[n][1]);
        if (ok_hi) *reinterpret_cast<float2*>(hi + n * 8 + 2 * t) = make_float2(c[n][2], c[n][3]);
    }
}
// bit (n * 8 + 2t + j) of the row's mask = (c[n][row * 2 + j] > 0); the four lanes of a quad OR their parts together
__device__ __forceinline__ void relu_masks(const float (&c)[4][4], unsigned& m_lo, unsigned& m_hi, int t) {
    m_lo = 0u; m_hi = 0u;
#pragma unroll
    for (int n = 0; n < 4; ++n) {
        const int b = n * 8 + 2 * t;
        m_lo |= (c[n][0] > 0.f ? 1u : 0u) << b | (c[n][1] > 0.f ? 2u : 0u) << b;
        m_hi |= (c[n][2] > 0.f ? 1u : 0u) << b | (c[n][3] > 0.f ? 2u : 0u) << b;
    }
    m_lo |= __shfl_xor_sync(0xffffffffu, m_lo, 1); m_lo |= __shfl_xor_sync(0xffffffffu, m_lo, 2);
    m_hi |= __shfl_xor_sync(0xffffffffu, m_hi, 1); m_hi |= __shfl_xor_sync(0xffffffffu, m_hi, 2);
}
template <int STRIDE, int MASK>
__device__ __forceinline__ void store_act(float* __restrict__ act, int64_t s_lo, int64_t s_hi, bool ok_lo, bool ok_hi, int t,
                                          const float (&h1)[4][4], const float (&o16)[2][4], const float (&hc)[4][4]) {
    float* lo = act + s_lo * STRIDE;
    float* hi = act + s_hi * STRIDE;
    store_frag_rows<4>(lo, hi, ok_lo, ok_hi, t, h1);
    store_frag_rows<2>(lo + 32, hi + 32, ok_lo, ok_hi, t, o16);
    store_frag_rows<4>(lo + 48, hi + 48, ok_lo, ok_hi, t, hc);
    unsigned a_lo, a_hi, c_lo, c_hi;
    relu_masks(h1, a_lo, a_hi, t);
    relu_masks(hc, c_lo, c_hi, t);
    if (t == 0) {
        if (ok_lo) *reinterpret_cast<uint2*>(lo + MASK) = make_uint2(a_lo, c_lo);
        if (ok_hi) *reinterpret_cast<uint2*>(hi + MASK) = make_uint2(a_hi, c_hi);
    }
}

// ---------------------------------------------------------------------------------------------
// decoder forward on the tensor pipe for one 16-point tile.
// X: tile [16][kXStride] (feat | blob); CFt: colour features [16][kCStride] (two grids).
// Outputs (accumulator fragments, rows g / g+8, columns 2t, 2t+1 of each 8-wide block):
//   h1[4][4] post-ReLU, o16[2][4] = (geo0..14, sdf), hc[4][4] post-ReLU, rgb[4] (columns 0..2 real).
// ---------------------------------------------------------------------------------------------
template <bool TWO>
__device__ __forceinline__ void decoder_forward_mma(const float4* __restrict__ W4, const float* __restrict__ X,
                                                    const float* __restrict__ CFt, int lane, float (&h1)[4][4],
                                                    float (&o16)[2][4], float (&hc)[4][4], float (&rgb)[4]) {
    using N = NetV2<TWO>;
    const int g = lane >> 2, t = lane & 3;
#pragma unroll
    for (int n = 0; n < 4; ++n)
#pragma unroll
        for (int i = 0; i < 4; ++i) { h1[n][i] = 0.f; hc[n][i] = 0.f; }
    // ---- sdf layer 1: [16 x 112] . [112 x 32]
#pragma unroll 2
    for (int ks = 0; ks < N::K1 / 8; ++ks) {
        const float2 lo = *reinterpret_cast<const float2*>(X + g * kXStride + ks * 8 + 2 * t);
        const float2 hi = *reinterpret_cast<const float2*>(X + (g + 8) * kXStride + ks * 8 + 2 * t);
        const SplitA a = split_a(lo.x, hi.x, lo.y, hi.y);
#pragma unroll
        for (int n = 0; n < 4; ++n) mma_3x(h1[n], a, W4[N::OFF1 + (n * 8 + g) * N::RS1 + ks * 4 + t]);
    }
#pragma unroll
    for (int n = 0; n < 4; ++n)
#pragma unroll
        for (int i = 0; i < 4; ++i) h1[n][i] = fmaxf(h1[n][i], 0.f);
    // ---- sdf layer 2: [16 x 32] . [32 x 16], A straight from the h1 fragments
#pragma unroll
    for (int n = 0; n < 2; ++n)
#pragma unroll
        for (int i = 0; i < 4; ++i) o16[n][i] = 0.f;
#pragma unroll
    for (int ks = 0; ks < 4; ++ks) {
        const SplitA a = split_a(h1[ks][0], h1[ks][2], h1[ks][1], h1[ks][3]);
#pragma unroll
        for (int n = 0; n < 2; ++n) mma_3x(o16[n], a, W4[N::OFF2 + (n * 8 + g) * N::RS2 + ks * 4 + t]);
    }
    // ---- colour layer 1: blob (tile) | cfeat (tile) | geo (o16 fragments; the sdf column meets a zero weight row)
    int kp0 = 0;
#pragma unroll 2
    for (int ks = 0; ks < kBlob / 8; ++ks) {
        const float2 lo = *reinterpret_cast<const float2*>(X + g * kXStride + kBlobCol + ks * 8 + 2 * t);
        const float2 hi = *reinterpret_cast<const float2*>(X + (g + 8) * kXStride + kBlobCol + ks * 8 + 2 * t);
        const SplitA a = split_a(lo.x, hi.x, lo.y, hi.y);
#pragma unroll
        for (int n = 0; n < 4; ++n) mma_3x(hc[n], a, W4[N::OFFC + (n * 8 + g) * N::RSC + kp0 + ks * 4 + t]);
    }
    kp0 += kBlob / 2;
    if (TWO) {
#pragma unroll 2
        for (int ks = 0; ks < 8; ++ks) {
            const float2 lo = *reinterpret_cast<const float2*>(CFt + g * kCStride + ks * 8 + 2 * t);
            const float2 hi = *reinterpret_cast<const float2*>(CFt + (g + 8) * kCStride + ks * 8 + 2 * t);
            const SplitA a = split_a(lo.x, hi.x, lo.y, hi.y);
#pragma unroll
            for (int n = 0; n < 4; ++n) mma_3x(hc[n], a, W4[N::OFFC + (n * 8 + g) * N::RSC + kp0 + ks * 4 + t]);
        }
        kp0 += 32;
    }
#pragma unroll
    for (int ks = 0; ks < 2; ++ks) {
        const SplitA a = split_a(o16[ks][0], o16[ks][2], o16[ks][1], o16[ks][3]);
#pragma unroll
        for (int n = 0; n < 4; ++n) mma_3x(hc[n], a, W4[N::OFFC + (n * 8 + g) * N::RSC + kp0 + ks * 4 + t]);
    }
#pragma unroll
    for (int n = 0; n < 4; ++n)
#pragma unroll
        for (int i = 0; i < 4; ++i) hc[n][i] = fmaxf(hc[n][i], 0.f);
    // ---- colour layer 2: [16 x 32] . [32 x 8]
#pragma unroll
    for (int i = 0; i < 4; ++i) rgb[i] = 0.f;
#pragma unroll
    for (int ks = 0; ks < 4; ++ks) {
        const SplitA a = split_a(hc[ks][0], hc[ks][2], hc[ks][1], hc[ks][3]);
        mma_3x(rgb, a, W4[N::OFFO + g * N::RS2 + ks * 4 + t]);
    }
}

}  // namespace mne
