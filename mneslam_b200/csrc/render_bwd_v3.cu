// v3 backward for calls that carry the forward's kept record (`act`, tile16.cuh: ActRow) -- replaces the autograd pass of
// loss.backward() (reference mp_slam/mapper.py:88,159) like render_bwd_v2.cu, restructured so that nothing in the tile
// loop synchronises a CTA:
//
//   points_bwd3_kernel    one 16-sample tile per warp, free running: ReLU masks + d_raw in, decoder backward on the tensor
//                         pipe (52 mma), d feat -> shared memory -> plane / table scatter.  The per-sample d h1 | d o16 | d hc
//                         are written into the record (320 B per active sample).  No feature gather, no forward recompute,
//                         no weight-gradient accumulators, no __syncthreads.
//   decoder_wgrad_kernel  the weight gradients dW^T = X^T . dY as a split-K product over the active samples: A and B
//                         fragments are loaded straight from the records (every 32-byte sector fully used), the OneBlob
//                         inputs are re-encoded from the coordinates kept in the record; warps = (sample set, block of dW) with register
//                         accumulators, combined per CTA in shared memory and flushed with one coalesced pass of atomics.
//
// Extra traffic against v2 (Replica room0: 88 064 samples, ~35 k active): the forward writes 264 B more per sample
// (features + masks: 23 MB), the tile kernel writes 320 B per active sample (11 MB), the weight-gradient kernel reads
// ~900 B per active sample (31 MB) -- against 108 MB of algorithmic scatter traffic.
#include "bwd_tiles.cuh"

namespace mne {

template <bool TWO> __host__ __device__ constexpr int bwd3_warp_floats() { return kTilePts * (kXStride + (TWO ? kCStride : 0)); }
template <bool TWO, bool RAY, int NW>
constexpr size_t bwd3_smem_bytes() {
    return sizeof(float) * (size_t)((RAY ? NetT<TWO>::FLOATS_RAY : NetT<TWO>::FLOATS_NORAY) + NW * bwd3_warp_floats<TWO>());
}

struct PtsBwd3Params {
    PtsBwdParams b;
    float* act;              // [n_pts][ActRow::STRIDE]: masks read, dY written
};

// plane-gradient scatter of one tile, lane = (point of a quad, 4-channel group): four 128-bit vector reductions per tap
// (one per corner; a warp instruction covers 4 points x one 128-byte texel line each).  RAY: also d loss / d (normalised
// plane coordinates) of every point -- the corner dot products stay per-lane partial sums over the lane's 4 channels
// (the expression is linear in them) and are reduced over the 8 lanes of a point once per quad.
template <bool RAY, bool PARAM>
__device__ __forceinline__ void scatter16_pq(const MneSceneDesc& sc, int grp, const float* __restrict__ tab_base, int tab_stride,
                                             const float* __restrict__ gsrc, int gstride, int np, int lane, unsigned gflags,
                                             float (&dn_keep)[3]) {
    const int pp = lane >> 3, q = lane & 7;
    const int p_loc = lane & 15, half = lane >> 4;
#pragma unroll 1
    for (int p0 = 0; p0 < kTilePts; p0 += 4) {
        if (p0 >= np) break;                                     // warp-uniform
        const int p = p0 + pp;
        const bool on = p < np;
        const unsigned pflags = RAY ? __shfl_sync(0xffffffffu, gflags, p & 15) : 0u;
        float dn[3] = {0.f, 0.f, 0.f};
        if (on) {
            const float* seg = tab_base + p * tab_stride;
#pragma unroll
            for (int l = 0; l < 2; ++l) {
                const float4 gv = *reinterpret_cast<const float4*>(gsrc + p * gstride + l * 32 + q * 4);
#pragma unroll
                for (int a = 0; a < 3; ++a) {
                    const int tp = a * 2 + l;
                    const MnePlaneDesc& Pl = sc.planes[grp * 6 + tp];
                    const uint4 off = *reinterpret_cast<const uint4*>(seg + 8 * tp);
                    const float4 w = *reinterpret_cast<const float4*>(seg + 8 * tp + 4);
                    const unsigned o[4] = {off.x, off.y, off.z, off.w};
                    const float ws[4] = {w.x, w.y, w.z, w.w};
                    if (PARAM && Pl.grad) {
                        float* gp = Pl.grad + q * 4;
#pragma unroll
                        for (int c = 0; c < 4; ++c) {
                            if (ws[c] != 0.f)
                                atomicAdd(reinterpret_cast<float4*>(gp + o[c]), make_float4(gv.x * ws[c], gv.y * ws[c], gv.z * ws[c], gv.w * ws[c]));
                        }
                        // texel flags: lanes 0..3 of the point's octet flag one corner each -- ONE store instruction per tap
                        // (a store per corner under `q == 0` serialised four divergent paths: +18 us on the kernel)
                        if (sc.dirty && q < 4) {
                            const unsigned oc = q == 0 ? o[0] : q == 1 ? o[1] : q == 2 ? o[2] : o[3];
                            const float wc = q == 0 ? ws[0] : q == 1 ? ws[1] : q == 2 ? ws[2] : ws[3];
                            if (wc != 0.f) mark_dirty(sc.dirty, ((Pl.grad - sc.dirty_base) >> 5) + (oc >> 5));
                        }
                    }
                    if (RAY) {
                        const unsigned fl = (pflags >> (2 * (grp * 6 + tp))) & 3u;
                        if (fl) {
                            const float* dp = Pl.data + q * 4;
                            float d[4];
#pragma unroll
                            for (int c = 0; c < 4; ++c) {
                                const float4 v = __ldg(reinterpret_cast<const float4*>(dp + o[c]));
                                d[c] = gv.x * v.x + gv.y * v.y + gv.z * v.z + gv.w * v.w;
                            }
                            const float wy0 = w.x + w.y, wy1 = w.z + w.w, wx0 = w.x + w.z, wx1 = w.y + w.w;
                            const int ax = (a == 2) ? 1 : 0, ay = (a == 0) ? 1 : 2;
                            if (fl & 1u) dn[ax] += ((d[1] - d[0]) * wy0 + (d[3] - d[2]) * wy1) * (0.5f * (float)(Pl.W - 1));
                            if (fl & 2u) dn[ay] += ((d[2] - d[0]) * wx0 + (d[3] - d[1]) * wx1) * (0.5f * (float)(Pl.H - 1));
                        }
                    }
                }
            }
        }
        if (RAY) {
#pragma unroll
            for (int a = 0; a < 3; ++a) {
                float v = dn[a];
                v += __shfl_xor_sync(0xffffffffu, v, 1); v += __shfl_xor_sync(0xffffffffu, v, 2); v += __shfl_xor_sync(0xffffffffu, v, 4);
                const float mine = __shfl_sync(0xffffffffu, v, (lane & 3) * 8);      // octet of point p0 + (p_loc & 3)
                if (half == 0 && (p_loc >> 2) == (p0 >> 2)) dn_keep[a] += mine;
            }
        }
    }
}

// MODE: 1 = channels-last planes, 2 = hash grid.  One CTA per SM, NW warps, no CTA-wide synchronisation after the weights
// are staged.
template <bool TWO, int MODE, bool RAY, bool PARAM, int NW>
__global__ void __launch_bounds__(NW * 32, 1) points_bwd3_kernel(const __grid_constant__ PtsBwd3Params Q) {
    const PtsBwdParams& P = Q.b;
    using T = NetT<TWO>;
    using AR = ActRow<TWO, MODE == 2>;
    constexpr int WF = bwd3_warp_floats<TWO>();
    extern __shared__ __align__(16) float smem_f[];
    float* WT = smem_f;
    float* tiles = WT + (RAY ? T::FLOATS_RAY : T::FLOATS_NORAY);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float* X = tiles + warp * WF;
    float* CFt = X + kTilePts * kXStride;

    load_weights_t<TWO, RAY>(P.sc, WT, threadIdx.x, blockDim.x);
    __syncthreads();

    // samples with a non-zero d loss / d raw only (list written by the ray kernel), dealt evenly over the rounds that are
    // needed anyway (e.g. 35 k samples on 148 x 20 warps = one round of 12-sample tiles)
    const int64_t n_pts = *P.n_active;
    const int64_t tiles_per_round = (int64_t)gridDim.x * NW;
    const int64_t n_rounds = (n_pts + tiles_per_round * kTilePts - 1) / (tiles_per_round * kTilePts);
    const int ppt = n_rounds > 0 ? (int)((n_pts + n_rounds * tiles_per_round - 1) / (n_rounds * tiles_per_round)) : kTilePts;
    const int64_t n_tiles = ppt > 0 ? (n_pts + ppt - 1) / ppt : 0;
    const int p_loc = lane & 15, half = lane >> 4;
    const int g = lane >> 2, t = lane & 3;

    for (int64_t tile = (int64_t)warp * gridDim.x + blockIdx.x; tile < n_tiles; tile += tiles_per_round) {
        const int64_t pt0 = tile * ppt;
        const int np = (int)min((int64_t)ppt, n_pts - pt0);

        // ---- 1. this lane's sample (lanes p and p + 16 share one), its position
        float px = 0.f, py = 0.f, pz = 0.f, zq = 0.f;
        int q = 0;
        if (p_loc < np) {
            q = __ldg(P.list + pt0 + p_loc);
            const int ray = q / P.n_samples;
            zq = __ldg(P.z_vals + q);
            px = __fadd_rn(__ldg(P.rays_o + ray * 3 + 0), __fmul_rn(__ldg(P.rays_d + ray * 3 + 0), zq));
            py = __fadd_rn(__ldg(P.rays_o + ray * 3 + 1), __fmul_rn(__ldg(P.rays_d + ray * 3 + 1), zq));
            pz = __fadd_rn(__ldg(P.rays_o + ray * 3 + 2), __fmul_rn(__ldg(P.rays_d + ray * 3 + 2), zq));
        }
        const int q_lo = __shfl_sync(0xffffffffu, q, g), q_hi = __shfl_sync(0xffffffffu, q, g + 8);
        const bool ok_lo = g < np, ok_hi = g + 8 < np;
        float* r_lo = Q.act + (int64_t)q_lo * AR::STRIDE;
        float* r_hi = Q.act + (int64_t)q_hi * AR::STRIDE;
        uint2 m_lo = make_uint2(0u, 0u), m_hi = m_lo;            // x: (h1 > 0) bits, y: (hc > 0) bits
        float4 dr_lo = make_float4(0.f, 0.f, 0.f, 0.f), dr_hi = dr_lo;
        if (ok_lo) { m_lo = *reinterpret_cast<const uint2*>(r_lo + AR::MASK); dr_lo = __ldg(P.draw + q_lo); }
        if (ok_hi) { m_hi = *reinterpret_cast<const uint2*>(r_hi + AR::MASK); dr_hi = __ldg(P.draw + q_hi); }

        // ---- 2. colour branch backward.  A fragment of d_rgb: columns (2t, 2t+1) of (r, g, b, 0, ...)
        const float ar0 = (t == 0) ? dr_lo.x : ((t == 1) ? dr_lo.z : 0.f), ar2 = (t == 0) ? dr_lo.y : 0.f;
        const float ar1 = (t == 0) ? dr_hi.x : ((t == 1) ? dr_hi.z : 0.f), ar3 = (t == 0) ? dr_hi.y : 0.f;
        float dhc[4][4];
#pragma unroll
        for (int n = 0; n < 4; ++n) {
#pragma unroll
            for (int i = 0; i < 4; ++i) dhc[n][i] = 0.f;
            const float2 b = *reinterpret_cast<const float2*>(WT + T::OFF_V2T + (n * 8 + g) * kRgbStride + 2 * t);
            mma_1x(dhc[n], ar0, ar1, ar2, ar3, b.x, b.y);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const unsigned m = (i & 2) ? m_hi.y : m_lo.y;
                dhc[n][i] = ((m >> (n * 8 + 2 * t + (i & 1))) & 1u) ? dhc[n][i] : 0.f;
            }
        }
        float dout[2][4];                                        // [d_geo (15) | d_sdf]
#pragma unroll
        for (int n = 0; n < 2; ++n) {
#pragma unroll
            for (int i = 0; i < 4; ++i) dout[n][i] = 0.f;
#pragma unroll
            for (int ks = 0; ks < 4; ++ks) {
                const float2 b = *reinterpret_cast<const float2*>(WT + T::OFF_VGT + (n * 8 + g) * T::S1 + ks * 8 + 2 * t);
                mma_1x(dout[n], dhc[ks][0], dhc[ks][2], dhc[ks][1], dhc[ks][3], b.x, b.y);
            }
        }
        if (t == 3) { dout[1][1] = dr_lo.w; dout[1][3] = dr_hi.w; }     // column 15 = sdf

        // ---- 3. sdf branch backward
        float dh1[4][4];
#pragma unroll
        for (int n = 0; n < 4; ++n) {
#pragma unroll
            for (int i = 0; i < 4; ++i) dh1[n][i] = 0.f;
#pragma unroll
            for (int ks = 0; ks < 2; ++ks) {
                const float2 b = *reinterpret_cast<const float2*>(WT + T::OFF_W2T + (n * 8 + g) * kGeoStride + ks * 8 + 2 * t);
                mma_1x(dh1[n], dout[ks][0], dout[ks][2], dout[ks][1], dout[ks][3], b.x, b.y);
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const unsigned m = (i & 2) ? m_hi.x : m_lo.x;
                dh1[n][i] = ((m >> (n * 8 + 2 * t + (i & 1))) & 1u) ? dh1[n][i] : 0.f;
            }
        }
        if (PARAM) {            // the weight-gradient kernel's B operands
            store_frag_rows<4>(r_lo + AR::DY, r_hi + AR::DY, ok_lo, ok_hi, t, dh1);
            store_frag_rows<2>(r_lo + AR::DY + 32, r_hi + AR::DY + 32, ok_lo, ok_hi, t, dout);
            store_frag_rows<4>(r_lo + AR::DY + 48, r_hi + AR::DY + 48, ok_lo, ok_hi, t, dhc);
        }

        // ---- 4. d feat = d_h1 . W1[:, :64] -> X columns [0, 64);  d colour feat = d_hc . col0[:, 48:112] -> CF tile
#pragma unroll
        for (int n = 0; n < 8; ++n) {
            float d[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int ks = 0; ks < 4; ++ks) {
                const float2 b = *reinterpret_cast<const float2*>(WT + T::OFF_W1T + (n * 8 + g) * T::S1 + ks * 8 + 2 * t);
                mma_1x(d, dh1[ks][0], dh1[ks][2], dh1[ks][1], dh1[ks][3], b.x, b.y);
            }
            store_frag(X, kXStride, n, g, t, d);
        }
        if (TWO) {
#pragma unroll
            for (int n = 0; n < 8; ++n) {
                float d[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
                for (int ks = 0; ks < 4; ++ks) {
                    const float2 b = *reinterpret_cast<const float2*>(WT + T::OFF_VCT + (n * 8 + g) * T::S1 + ks * 8 + 2 * t);
                    mma_1x(d, dhc[ks][0], dhc[ks][2], dhc[ks][1], dhc[ks][3], b.x, b.y);
                }
                store_frag(CFt, kCStride, n, g, t, d);
            }
        }
        float* row = X + p_loc * kXStride;
        float dpt[3] = {0.f, 0.f, 0.f};                          // d loss / d point through the OneBlob (lanes 0..15)
        unsigned gflags = 0;                                     // per tap: bit 2t = x gradient on, 2t+1 = y gradient on
        const float bx0 = blob_input(px, P.sc.bbox_lo[0], P.sc.bbox_ext[0]);
        const float by0 = blob_input(py, P.sc.bbox_lo[1], P.sc.bbox_ext[1]);
        const float bz0 = blob_input(pz, P.sc.bbox_lo[2], P.sc.bbox_ext[2]);
        if (RAY) {
            // d blob = d_h1 . W1[:, 64:112] + d_hc . col0[:, 0:48] -> the blob segment of X
#pragma unroll
            for (int n = 0; n < 6; ++n) {
                float d[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
                for (int ks = 0; ks < 4; ++ks) {
                    const float2 b1 = *reinterpret_cast<const float2*>(WT + T::OFF_W1BT + (n * 8 + g) * T::S1 + ks * 8 + 2 * t);
                    mma_1x(d, dh1[ks][0], dh1[ks][2], dh1[ks][1], dh1[ks][3], b1.x, b1.y);
                    const float2 b2 = *reinterpret_cast<const float2*>(WT + T::OFF_V1BT + (n * 8 + g) * T::S1 + ks * 8 + 2 * t);
                    mma_1x(d, dhc[ks][0], dhc[ks][2], dhc[ks][1], dhc[ks][3], b2.x, b2.y);
                }
                store_frag(X + kBlobCol, kXStride, n, g, t, d);
            }
            __syncwarp();
            float d0 = 0.f, d1 = 0.f;
            if (half == 0) {
                d0 = oneblob16_sparse_grad(bx0, row + kBlobCol);
                d1 = oneblob16_sparse_grad(by0, row + kBlobCol + 16);
            } else {
                d0 = oneblob16_sparse_grad(bz0, row + kBlobCol + 32);
            }
            const float dz = __shfl_sync(0xffffffffu, d0, p_loc + 16);
            dpt[0] = (float)((double)d0 / P.sc.bbox_ext[0]);
            dpt[1] = (float)((double)d1 / P.sc.bbox_ext[1]);
            dpt[2] = (float)((double)dz / P.sc.bbox_ext[2]);
            // which taps let the coordinate gradient through (pixel coordinate not clamped)
            const float nn[3] = {plane_coord(px, P.sc.bound_lo[0], P.sc.bound_ext[0]), plane_coord(py, P.sc.bound_lo[1], P.sc.bound_ext[1]),
                                 plane_coord(pz, P.sc.bound_lo[2], P.sc.bound_ext[2])};
#pragma unroll
            for (int tp = 0; MODE != 2 && tp < (TWO ? 12 : 6); ++tp) {
                const int a = (tp % 6) >> 1;
                const MnePlaneDesc& Pl = P.sc.planes[tp];
                const AxisTap tx = axis_tap((a == 2) ? nn[1] : nn[0], Pl.W), ty = axis_tap((a == 0) ? nn[1] : nn[2], Pl.H);
                gflags |= (tx.grad_on ? 1u : 0u) << (2 * tp) | (ty.grad_on ? 1u : 0u) << (2 * tp + 1);
            }
        }
        __syncwarp();                                            // d feat tiles complete; blob segment spent

        // ---- 5. tap table (or hash coordinates) into the blob segment, then scatter
        float dn_keep[3] = {0.f, 0.f, 0.f};                      // d loss / d normalised plane (or unit-cube) coordinates of this lane's point
        if (MODE == 2) {
            if (half == 0) { row[kBlobCol] = bx0; row[kBlobCol + 1] = by0; row[kBlobCol + 2] = bz0; }
            __syncwarp();
            if (PARAM && P.sc.hash[0].grad) scatter16_hash(P.sc.hash[0], X + kBlobCol, kXStride, X, kXStride, np, lane, P.sc.dirty, P.sc.dirty_base);
            if (TWO && PARAM && P.sc.hash[1].grad) scatter16_hash(P.sc.hash[1], X + kBlobCol, kXStride, CFt, kCStride, np, lane, P.sc.dirty, P.sc.dirty_base);
            if (RAY) {
                hash_point_grad16(P.sc.hash[0], X + kBlobCol, kXStride, X, kXStride, np, lane, dn_keep);
                if (TWO) hash_point_grad16(P.sc.hash[1], X + kBlobCol, kXStride, CFt, kCStride, np, lane, dn_keep);
            }
        } else {
            write_taps8_level(P.sc, reinterpret_cast<float4*>(row + kBlobCol), px, py, pz, 0, half);
            __syncwarp();
            scatter16_pq<RAY, PARAM>(P.sc, 0, X + kBlobCol, kXStride, X, kXStride, np, lane, gflags, dn_keep);
            if (TWO) {
                __syncwarp();
                write_taps8_level(P.sc, reinterpret_cast<float4*>(row + kBlobCol), px, py, pz, 1, half);
                __syncwarp();
                scatter16_pq<RAY, PARAM>(P.sc, 1, X + kBlobCol, kXStride, CFt, kCStride, np, lane, gflags, dn_keep);
            }
        }
        if (RAY && half == 0 && p_loc < np) {
            const int ray = q / P.n_samples;
#pragma unroll
            for (int a = 0; a < 3; ++a) {
                // planes: coordinates in [-1, 1] over `bound`; hash grid: in [0, 1] over `bounding_box` (like the OneBlob input)
                const float gpnt = (MODE == 2 ? (float)((double)dn_keep[a] / P.sc.bbox_ext[a]) : dn_keep[a] * (2.0f / P.sc.bound_ext[a])) + dpt[a];
                atomicAdd(P.g_rays_o + ray * 3 + a, gpnt);
                atomicAdd(P.g_rays_d + ray * 3 + a, gpnt * zq);
            }
        }
        __syncwarp();                                            // the tiles are rewritten by the next iteration
    }
}

// ---------------------------------------------------------------------------------------------
// weight gradients: dW^T[in][out] = sum over active samples of X[sample][in] * dY[sample][out]
//   sdf0: [feat F | blob 48] x d h1        sdf1: h1 x d o16        col0: [blob 48 | cfeat F | geo 15] x d hc        col1: hc x d rgb
// mma.m16n8k8 TF32 with the SAMPLE index as the contraction dimension: a k-step is 8 samples, A[m = input][k = sample],
// B[k = sample][n = output], both read straight from the per-sample records (lanes of equal t read 8 consecutive floats
// of one record: whole 32-byte sectors).
// ---------------------------------------------------------------------------------------------
struct WGradParams {
    MneSceneDesc sc;
    const float4* draw;
    const int* list;
    const int* n_active;
    const float* act;
};

// sample sets per CTA (each a full set of roles with its own accumulators); the register file caps sets x roles
template <bool TWO> __host__ __device__ constexpr int wg_sets() { return 2; }
constexpr int kWgUnroll = 2;       // k-steps (of 8 samples) in flight per warp iteration
constexpr int kBlobTile = 56;      // floats per sample of a blob tile: == 24 (mod 32), conflict-free fragment reads

// roles: 0, 1 = feat halves x d h1;  2 = blob x d h1;  3 = blob x d hc;  4 = h1 x d o16 | geo x d hc | hc x d rgb;
//        5, 6 = colour feat halves x d hc (two grids)
template <bool TWO> __host__ __device__ constexpr int wg_roles() { return TWO ? 7 : 5; }

template <bool TWO, bool HASH>
struct WgLayout {       // the CTA's combined gradient in shared memory, in the layout of the global tensors
    static constexpr int F = HASH ? 32 : 64;
    static constexpr int IN_SDF = F + kBlob, IN_COL = kBlob + (TWO ? F : 0) + kGeo;
    static constexpr int SDF0 = 0, SDF1 = SDF0 + 32 * IN_SDF, COL0 = SDF1 + 16 * 32, COL1 = COL0 + 32 * IN_COL;
    static constexpr int TOTAL = COL1 + 3 * 32;
};

// one 16 x (NB*8) block update: acc[nb] += A(16 inputs x 8 samples) . B(8 samples x 8 outputs)
template <int NB>
__device__ __forceinline__ void wg_mma(float (&acc)[NB][4], const float (&a)[4], const float (&b)[NB][2]) {
#pragma unroll
    for (int nb = 0; nb < NB; ++nb) mma_1x(acc[nb], a[0], a[1], a[2], a[3], b[nb][0], b[nb][1]);
}

template <bool TWO, bool HASH>
__global__ void __launch_bounds__(wg_sets<TWO>() * wg_roles<TWO>() * 32, 1) decoder_wgrad_kernel(const __grid_constant__ WGradParams P) {
    using AR = ActRow<TWO, HASH>;
    using L = WgLayout<TWO, HASH>;
    constexpr int NR = wg_roles<TWO>();
    constexpr int kWgSets = wg_sets<TWO>();
    constexpr int MBF = AR::F / 32;                     // 16-row blocks per feat half-role: 2 (planes) or 1 (hash)
    extern __shared__ __align__(16) float wg_smem[];
    float* dW = wg_smem;                                // [L::TOTAL] the CTA's combined gradient
    float* blob_s = wg_smem + L::TOTAL;                 // [kWgSets * 2][kWgUnroll * 8 * kBlobTile] OneBlob tiles of the blob warps
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int set = warp / NR, role = warp % NR;
    const int g = lane >> 2, t = lane & 3;
    const int n = *P.n_active;
    const int n_ks = (n + 7) / 8;

    // accumulators: the widest role (blob x 32 outputs: 3 row blocks x 4 column blocks) sets the size
    float acc[3][4][4];
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
            for (int k = 0; k < 4; ++k) acc[i][j][k] = 0.f;

    const bool is_blob = role == 2 || role == 3;
    float* btile = blob_s + (set * 2 + (role == 3 ? 1 : 0)) * (kWgUnroll * 8 * kBlobTile);
    const int ks_stride = gridDim.x * kWgSets * kWgUnroll;
    // lane (g, t) touches samples t and t + 4 of each k-step; the sample indices of the NEXT iteration are fetched while this
    // one computes (the list -> record -> mma chain is two dependent memory round trips otherwise)
    int qA[kWgUnroll], qB[kWgUnroll];
    auto fetch_list = [&](int ks, int (&a)[kWgUnroll], int (&b)[kWgUnroll]) {
#pragma unroll
        for (int u = 0; u < kWgUnroll; ++u) {
            const int iA = (ks + u) * 8 + t, iB = iA + 4;
            a[u] = iA < n ? __ldg(P.list + iA) : -1;
            b[u] = iB < n ? __ldg(P.list + iB) : -1;
        }
    };
    const int ks_first = (blockIdx.x * kWgSets + set) * kWgUnroll;
    fetch_list(ks_first, qA, qB);
    for (int ks0 = ks_first; ks0 < n_ks; ks0 += ks_stride) {
        const float* rA[kWgUnroll];
        const float* rB[kWgUnroll];
#pragma unroll
        for (int u = 0; u < kWgUnroll; ++u) {
            rA[u] = P.act + (int64_t)max(qA[u], 0) * AR::STRIDE;
            rB[u] = P.act + (int64_t)max(qB[u], 0) * AR::STRIDE;
        }
        int nA[kWgUnroll], nB[kWgUnroll];
        fetch_list(ks0 + ks_stride, nA, nB);
        if (is_blob) {
            // OneBlob inputs of the 16 samples of this iteration, re-encoded from the coordinates the forward kept: job = (sample, dimension)
            __syncwarp();
            for (int job = lane; job < kWgUnroll * 8 * 3; job += 32) {
                const int s = job % (kWgUnroll * 8), dim = job / (kWgUnroll * 8);
                const int i = ks0 * 8 + s;
                float x = 0.5f;
                if (i < n) x = __ldg(P.act + (int64_t)__ldg(P.list + i) * AR::STRIDE + AR::BLOBX + dim);
                oneblob16_sparse(x, btile + s * kBlobTile + dim * 16);
            }
            __syncwarp();
        }
#pragma unroll
        for (int u = 0; u < kWgUnroll; ++u) {
            const bool okA = qA[u] >= 0, okB = qB[u] >= 0;
            auto ldA = [&](int col) { return okA ? __ldg(rA[u] + col) : 0.f; };
            auto ldB = [&](int col) { return okB ? __ldg(rB[u] + col) : 0.f; };
            if (role <= 1 || role >= 5) {
                // feat (role 0, 1) x d h1, or colour feat (5, 6) x d hc: MBF row blocks x 4 column blocks
                const int a_off = (role <= 1 ? AR::FEAT : AR::CFEAT) + ((role == 1 || role == 6) ? MBF * 16 : 0);
                const int b_off = AR::DY + (role <= 1 ? 0 : 48);
                float b[4][2];
#pragma unroll
                for (int nb = 0; nb < 4; ++nb) { b[nb][0] = ldA(b_off + nb * 8 + g); b[nb][1] = ldB(b_off + nb * 8 + g); }
#pragma unroll
                for (int mb = 0; mb < MBF; ++mb) {
                    const float a[4] = {ldA(a_off + mb * 16 + g), ldA(a_off + mb * 16 + g + 8), ldB(a_off + mb * 16 + g), ldB(a_off + mb * 16 + g + 8)};
                    wg_mma<4>(acc[mb], a, b);
                }
            } else if (is_blob) {
                const int b_off = AR::DY + (role == 2 ? 0 : 48);
                float b[4][2];
#pragma unroll
                for (int nb = 0; nb < 4; ++nb) { b[nb][0] = ldA(b_off + nb * 8 + g); b[nb][1] = ldB(b_off + nb * 8 + g); }
                const float* bt = btile + (u * 8 + t) * kBlobTile;
#pragma unroll
                for (int mb = 0; mb < 3; ++mb) {
                    const float a[4] = {bt[mb * 16 + g], bt[mb * 16 + g + 8], bt[4 * kBlobTile + mb * 16 + g], bt[4 * kBlobTile + mb * 16 + g + 8]};
                    wg_mma<4>(acc[mb], a, b);
                }
            } else {
                // role 4.  acc[0][0..3] = h1 (2 row blocks) x d o16 (2 column blocks): index rb * 2 + cb
                //          acc[1][0..3] = geo (1 row block) x d hc (4 column blocks)
                //          acc[2][0..1] = hc (2 row blocks) x d rgb (1 column block)
                float bo[2][2], bc[4][2], br[1][2];
#pragma unroll
                for (int nb = 0; nb < 2; ++nb) { bo[nb][0] = ldA(AR::DY + 32 + nb * 8 + g); bo[nb][1] = ldB(AR::DY + 32 + nb * 8 + g); }
#pragma unroll
                for (int nb = 0; nb < 4; ++nb) { bc[nb][0] = ldA(AR::DY + 48 + nb * 8 + g); bc[nb][1] = ldB(AR::DY + 48 + nb * 8 + g); }
                const float* dA = reinterpret_cast<const float*>(P.draw + max(qA[u], 0));
                const float* dB = reinterpret_cast<const float*>(P.draw + max(qB[u], 0));
                br[0][0] = (okA && g < 3) ? __ldg(dA + g) : 0.f;
                br[0][1] = (okB && g < 3) ? __ldg(dB + g) : 0.f;
#pragma unroll
                for (int rb = 0; rb < 2; ++rb) {
                    const float a0 = ldA(AR::H1 + rb * 16 + g), a1 = ldA(AR::H1 + rb * 16 + g + 8), a2 = ldB(AR::H1 + rb * 16 + g), a3 = ldB(AR::H1 + rb * 16 + g + 8);
#pragma unroll
                    for (int cb = 0; cb < 2; ++cb) mma_1x(acc[0][rb * 2 + cb], a0, a1, a2, a3, bo[cb][0], bo[cb][1]);
                    const float c0 = ldA(AR::HC + rb * 16 + g), c1 = ldA(AR::HC + rb * 16 + g + 8), c2 = ldB(AR::HC + rb * 16 + g), c3 = ldB(AR::HC + rb * 16 + g + 8);
                    mma_1x(acc[2][rb], c0, c1, c2, c3, br[0][0], br[0][1]);
                }
                const float a[4] = {ldA(AR::O16 + g), ldA(AR::O16 + g + 8), ldB(AR::O16 + g), ldB(AR::O16 + g + 8)};
                wg_mma<4>(acc[1], a, bc);
            }
        }
#pragma unroll
        for (int u = 0; u < kWgUnroll; ++u) { qA[u] = nA[u]; qB[u] = nB[u]; }
    }

    // ---- combine the sets in shared memory (set 0 stores, the others add in turn), in the layout of the global tensors.
    // Fragment element (block mb/nb, i): input row = 16 mb + g + (i & 2 ? 8 : 0), output column = 8 nb + 2t + (i & 1).
    for (int s = 0; s < kWgSets; ++s) {
        if (set == s) {
            auto put = [&](int idx, float v) { if (s == 0) dW[idx] = v; else dW[idx] += v; };
            if (role <= 1 || role >= 5) {
#pragma unroll
                for (int mb = 0; mb < MBF; ++mb)
#pragma unroll
                    for (int nb = 0; nb < 4; ++nb)
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            const int k = ((role == 1 || role == 6) ? MBF * 16 : 0) + mb * 16 + g + ((i & 2) ? 8 : 0), j = nb * 8 + 2 * t + (i & 1);
                            if (role <= 1) put(L::SDF0 + j * L::IN_SDF + k, acc[mb][nb][i]);
                            else put(L::COL0 + j * L::IN_COL + kBlob + k, acc[mb][nb][i]);
                        }
            } else if (is_blob) {
#pragma unroll
                for (int mb = 0; mb < 3; ++mb)
#pragma unroll
                    for (int nb = 0; nb < 4; ++nb)
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            const int k = mb * 16 + g + ((i & 2) ? 8 : 0), j = nb * 8 + 2 * t + (i & 1);
                            if (role == 2) put(L::SDF0 + j * L::IN_SDF + AR::F + k, acc[mb][nb][i]);
                            else put(L::COL0 + j * L::IN_COL + k, acc[mb][nb][i]);
                        }
            } else {
#pragma unroll
                for (int rb = 0; rb < 2; ++rb) {
#pragma unroll
                    for (int cb = 0; cb < 2; ++cb)
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            const int j = rb * 16 + g + ((i & 2) ? 8 : 0), c = cb * 8 + 2 * t + (i & 1);
                            put(L::SDF1 + ((c == 15) ? 0 : c + 1) * 32 + j, acc[0][rb * 2 + cb][i]);
                        }
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const int j = rb * 16 + g + ((i & 2) ? 8 : 0), c = 2 * t + (i & 1);
                        if (c < 3) put(L::COL1 + c * 32 + j, acc[2][rb][i]);
                    }
                }
#pragma unroll
                for (int nb = 0; nb < 4; ++nb)
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const int k = g + ((i & 2) ? 8 : 0), j = nb * 8 + 2 * t + (i & 1);
                        if (k < kGeo) put(L::COL0 + j * L::IN_COL + (L::IN_COL - kGeo) + k, acc[1][nb][i]);
                    }
            }
        }
        __syncthreads();
    }
    if ((blockIdx.x * kWgSets) * kWgUnroll >= n_ks) return;       // this CTA saw no sample: nothing to add
    // one pass of vector reductions (the four regions start at multiples of 4 floats; the entry point checked the alignment)
    for (int i = threadIdx.x * 4; i < L::TOTAL; i += blockDim.x * 4) {
        const float4 v = *reinterpret_cast<const float4*>(dW + i);
        if (v.x == 0.f && v.y == 0.f && v.z == 0.f && v.w == 0.f) continue;
        float* dst;
        if (i < L::SDF1) dst = P.sc.g_sdf0 ? P.sc.g_sdf0 + i : nullptr;
        else if (i < L::COL0) dst = P.sc.g_sdf1 ? P.sc.g_sdf1 + (i - L::SDF1) : nullptr;
        else if (i < L::COL1) dst = P.sc.g_col0 ? P.sc.g_col0 + (i - L::COL0) : nullptr;
        else dst = P.sc.g_col1 ? P.sc.g_col1 + (i - L::COL1) : nullptr;
        if (dst) atomicAdd(reinterpret_cast<float4*>(dst), v);
    }
}

int sm_count();

template <bool TWO, int MODE, bool RAY, bool PARAM, int NW>
static int launch_points_bwd3(const PtsBwd3Params& Q, cudaStream_t stream) {
    const size_t smem = bwd3_smem_bytes<TWO, RAY, NW>();
    static std::atomic<bool> configured[64] = {};
    int dev = 0;
    MNE_CHECK_CUDA(cudaGetDevice(&dev));
    if (dev >= 64 || !configured[dev]) {
        MNE_CHECK_CUDA(cudaFuncSetAttribute(points_bwd3_kernel<TWO, MODE, RAY, PARAM, NW>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        if (dev < 64) configured[dev] = true;
    }
    const int64_t n_tiles = (Q.b.n_pts + kTilePts - 1) / kTilePts;
    const int blocks = (int)max((int64_t)1, min((n_tiles + NW - 1) / NW, (int64_t)sm_count()));
    points_bwd3_kernel<TWO, MODE, RAY, PARAM, NW><<<blocks, NW * 32, smem, stream>>>(Q);
    MNE_CHECK_CUDA(cudaGetLastError());
    return 0;
}

template <bool TWO, int MODE>
static int dispatch_points_bwd3(const PtsBwd3Params& Q, bool ray, bool param, cudaStream_t st) {
    // warps per CTA: what the registers (64 K per SM) and the per-warp tiles (7.7 KB, 12.3 KB with two grids) allow
    constexpr int NW = TWO ? 12 : 16;
    if (ray && param) return launch_points_bwd3<TWO, MODE, true, true, NW>(Q, st);
    if (ray) return launch_points_bwd3<TWO, MODE, true, false, NW>(Q, st);
    return launch_points_bwd3<TWO, MODE, false, true, NW>(Q, st);
}

template <bool TWO, bool HASH>
static int launch_wgrad(const WGradParams& W, int64_t n_pts, int max_ctas, cudaStream_t st) {
    constexpr int kWgSets = wg_sets<TWO>();
    const int64_t per_cta = (int64_t)kWgSets * kWgUnroll * 8;
    const int64_t cap = max_ctas > 0 ? (int64_t)max_ctas : (int64_t)sm_count();
    const int blocks = (int)max((int64_t)1, min((n_pts + per_cta - 1) / per_cta, cap));
    const size_t smem = sizeof(float) * (size_t)(WgLayout<TWO, HASH>::TOTAL + kWgSets * 2 * kWgUnroll * 8 * kBlobTile);
    static std::atomic<bool> configured[64] = {};
    int dev = 0;
    MNE_CHECK_CUDA(cudaGetDevice(&dev));
    if (dev >= 64 || !configured[dev]) {
        MNE_CHECK_CUDA(cudaFuncSetAttribute(decoder_wgrad_kernel<TWO, HASH>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        if (dev < 64) configured[dev] = true;
    }
    decoder_wgrad_kernel<TWO, HASH><<<blocks, kWgSets * wg_roles<TWO>() * 32, smem, st>>>(W);
    MNE_CHECK_CUDA(cudaGetLastError());
    return 0;
}

int decoder_wgrad(const MneSceneDesc*, int32_t, int32_t, void*, const float*, int32_t, cudaStream_t);

// backward over the compacted sample list for calls that carry the forward's record
int points_backward_v3(const PtsBwdParams& P, bool ray, bool param, float* act, const float4* draw, cudaStream_t st) {
    MNE_CHECK_ARG(((uintptr_t)act & 31) == 0, "mne_render_backward: act must be 32-byte aligned");
    PtsBwd3Params Q{};
    Q.b = P; Q.act = act;
    const bool hash = P.sc.encoder == 1, two = P.sc.two_grid != 0;
    int e;
    if (hash) e = two ? dispatch_points_bwd3<true, 2>(Q, ray, param, st) : dispatch_points_bwd3<false, 2>(Q, ray, param, st);
    else e = two ? dispatch_points_bwd3<true, 1>(Q, ray, param, st) : dispatch_points_bwd3<false, 1>(Q, ray, param, st);
    if (e) return e;
    if (!param || !(P.sc.g_sdf0 || P.sc.g_sdf1 || P.sc.g_col0 || P.sc.g_col1)) return 0;
    return decoder_wgrad(&P.sc, (int32_t)(P.n_pts / P.n_samples), P.n_samples, const_cast<float4*>(draw), act, 0, st);
}

// weight gradients from the records the tile kernel completed (d_raw + sample list at `scratch`, layout of
// mne_backward_scratch_bytes).  max_ctas > 0 caps the grid (e.g. to run beside the optimiser on a few SMs).
int decoder_wgrad(const MneSceneDesc* scene, int32_t n_rays, int32_t n_samples, void* scratch, const float* act, int32_t max_ctas,
                  cudaStream_t st) {
    const float* gs[4] = {scene->g_sdf0, scene->g_sdf1, scene->g_col0, scene->g_col1};
    for (int i = 0; i < 4; ++i)
        MNE_CHECK_ARG(((uintptr_t)gs[i] & 15) == 0, "decoder weight-gradient buffers must be 16-byte aligned");
    const int64_t n_pts = (int64_t)n_rays * n_samples;
    WGradParams W{};
    W.sc = *scene;
    W.draw = reinterpret_cast<const float4*>(scratch);
    W.list = reinterpret_cast<const int*>(W.draw + n_pts);
    W.n_active = W.list + ((n_pts + 3) & ~(int64_t)3);
    W.act = act;
    const bool hash = scene->encoder == 1, two = scene->two_grid != 0;
    if (hash) return two ? launch_wgrad<true, true>(W, n_pts, max_ctas, st) : launch_wgrad<false, true>(W, n_pts, max_ctas, st);
    return two ? launch_wgrad<true, false>(W, n_pts, max_ctas, st) : launch_wgrad<false, false>(W, n_pts, max_ctas, st);
}

}  // namespace mne
