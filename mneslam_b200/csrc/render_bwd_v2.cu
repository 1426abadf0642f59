// v2 backward: (1) a ray kernel -- warp per ray: compositing backward, loss / map upstream -> d_raw scratch;
// (2) a point kernel -- 16-point tiles per warp, decoder forward recompute and backward on the tensor pipe.
//
// Weight gradients are the one cross-point reduction.  They are NOT pushed through shared-memory float atomics
// (a CAS loop on this hardware): the CTA's warps advance in lock-step rounds of one tile each, publish their
// activation / gradient tiles in shared memory, and every warp owns a fixed set of 16x8 blocks of the weight
// gradients which it accumulates IN REGISTERS over all tiles of all rounds with mma (the point index is the
// contraction dimension).  One pass of global atomics per CTA at the very end.
//
// Plane gradients: lane = (corner, 4-channel group), one 128-bit vector reduction (red.global.add.v4.f32) per
// lane covers the four corner texels of a tap.
#include "bwd_tiles.cuh"
#include "ray_tail.cuh"

namespace mne {

template <bool TWO> __host__ __device__ constexpr int bwd2_warp_floats() {
    return kTilePts * (kXStride + (TWO ? kCStride : 0) + 2 * kActStride + kGeoStride + kRgbStride);
}
template <bool TWO, bool RAY, int NW>
constexpr size_t bwd2_smem_bytes() {
    // forward-recompute weights: one grid keeps the split (3-term) layout -- tighter gradients; two grids only fit with
    // the single-TF32 pair layout (half the bytes)
    return sizeof(float4) * (size_t)NetV2<TWO>::W4_COUNT +
           sizeof(float) * (size_t)((RAY ? NetT<TWO>::FLOATS_RAY : NetT<TWO>::FLOATS_NORAY) + NW * bwd2_warp_floats<TWO>());
}

// acc[nb] (+)= sum over the 16 points of a source tile of  A[p][m0 + ...] (x) B[p][nb*8 + ...]   (transposed product)
template <int NB, int OFF, int NACC>
__device__ __forceinline__ void outer_mma(float (&acc)[NACC][4], const float* __restrict__ A, int a_stride, int m0,
                                          const float* __restrict__ B, int b_stride, int g, int t) {
#pragma unroll
    for (int ks = 0; ks < 2; ++ks) {
        const float* ar = A + (ks * 8 + t) * a_stride + m0 + g;
        const float a0 = ar[0], a1 = ar[8], a2 = ar[4 * a_stride], a3 = ar[4 * a_stride + 8];
#pragma unroll
        for (int nb = 0; nb < NB; ++nb) {
            const float* br = B + (ks * 8 + t) * b_stride + nb * 8 + g;
            mma_1x(acc[OFF + nb], a0, a1, a2, a3, br[0], br[4 * b_stride]);
        }
    }
}

// MODE: 0 = planes through strides, 1 = channels-last planes (vector gather / scatter), 2 = hash grid
// NW = warps per CTA: 8 for one grid; two grids need more shared memory per warp and for the weights, so 6 (or 5 with
// the ray-gradient weights); weight-gradient blocks are dealt round-robin to warps (slot = block / NW).
// (The calls that come with the forward's kept record -- `act` -- run on the barrier-free kernels of render_bwd_v3.cu;
// the kernels here recompute the forward and serve the calls without it.)
template <bool TWO, int MODE, bool RAY, bool PARAM, int NW>
__global__ void __launch_bounds__(NW * 32, 1) points_bwd_kernel(const __grid_constant__ PtsBwdParams P) {
    using N = NetV2<TWO>;
    using T = NetT<TWO>;
    constexpr int WF = bwd2_warp_floats<TWO>();
    extern __shared__ __align__(16) float4 smem4[];
    float4* W4 = smem4;                                   // split (hi, lo) pairs for the activation recompute
    float* WT = reinterpret_cast<float*>(W4 + N::W4_COUNT);
    float* tiles = WT + (RAY ? T::FLOATS_RAY : T::FLOATS_NORAY);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float* X = tiles + warp * WF;
    float* CFt = X + kTilePts * kXStride;
    float* ACT = CFt + (TWO ? kTilePts * kCStride : 0);
    float* DY = ACT + kTilePts * kActStride;
    float* GEO = DY + kTilePts * kActStride;       // o16 in phase 1, d_out in phase 2
    float* DRGB = GEO + kTilePts * kGeoStride;
    // offsets of the same tiles inside another warp's block
    constexpr int O_CF = kTilePts * kXStride, O_ACT = O_CF + (TWO ? kTilePts * kCStride : 0), O_DY = O_ACT + kTilePts * kActStride,
                  O_GEO = O_DY + kTilePts * kActStride, O_DRGB = O_GEO + kTilePts * kGeoStride;

    load_weights_v2<TWO>(P.sc, W4, threadIdx.x, blockDim.x);
    load_weights_t<TWO, RAY>(P.sc, WT, threadIdx.x, blockDim.x);
    __syncthreads();

    // only samples with a non-zero d loss / d raw are processed (list written by rays_bwd_kernel): everything behind
    // the truncation band of a ray -- about a third of a mapping batch -- would only add exact zeros
    const int64_t n_pts = *P.n_active;
    // The kernel is latency bound per tile (one tile per warp and round), so the samples are dealt evenly over the rounds
    // that are needed anyway: e.g. 35 k samples on 148 x 12 warps are two rounds of 10-sample tiles, not a full round of
    // 16-sample tiles followed by a quarter-full one of the same length.
    const int64_t tiles_per_round = (int64_t)gridDim.x * NW;
    const int64_t n_rounds = (n_pts + tiles_per_round * kTilePts - 1) / (tiles_per_round * kTilePts);
    const int ppt = n_rounds > 0 ? (int)((n_pts + n_rounds * tiles_per_round - 1) / (n_rounds * tiles_per_round)) : kTilePts;   // samples per tile, <= 16
    const int64_t n_tiles = ppt > 0 ? (n_pts + ppt - 1) / ppt : 0;
    const int p_loc = lane & 15, half = lane >> 4;
    const int g = lane >> 2, t = lane & 3;

    // weight-gradient blocks owned by this warp (transposed: rows = inputs, cols = outputs)
    //   phase 1: d col0^T  -- one 16-row block of inputs per warp (one grid: warps 0..3; two grids: 0..7), 4 col blocks
    //            d col1^T  -- rows j (2 blocks) x 8 cols: warps 4,5 (one grid) / 0,1 (two grids, extra)
    //   phase 2: d sdf0^T  -- warps 0..6, input block = warp, 4 col blocks;  d sdf1^T -- warp 7: 2 row blocks x 2 col blocks
    // block b of a phase is owned by warp b % NW, in that warp's accumulator slot b / NW
    constexpr int V1_BLOCKS = N::KC / 16;                   // col0 input blocks: 4 (one grid) or 8 (two grids)
    constexpr int NB1 = V1_BLOCKS + 2;                      // + the two row blocks of col1
    constexpr int NB2 = 8;                                  // 7 input blocks of sdf0 + sdf1 (four 16x8 blocks in one slot)
    constexpr int S1 = (NB1 + NW - 1) / NW, S2 = (NB2 + NW - 1) / NW;
    float acc1[S1][4][4], acc2[S2][4][4];
#pragma unroll
    for (int n = 0; n < 4; ++n)
#pragma unroll
        for (int i = 0; i < 4; ++i) {
#pragma unroll
            for (int sl = 0; sl < S1; ++sl) acc1[sl][n][i] = 0.f;
#pragma unroll
            for (int sl = 0; sl < S2; ++sl) acc2[sl][n][i] = 0.f;
        }

    for (int64_t round = 0; round < n_rounds; ++round) {
        // consecutive tiles of a round go to different SMs (block-minor), as in the forward
        const int64_t tile = round * tiles_per_round + (int64_t)warp * gridDim.x + blockIdx.x;
        const bool have = tile < n_tiles;
        const int64_t pt0 = tile * ppt;
        const int np = have ? (int)min((int64_t)ppt, n_pts - pt0) : 0;

        // ---- 1. position, taps, gather, OneBlob
        float px = 0.f, py = 0.f, pz = 0.f;
        int q = 0;                                               // sample index of this lane's row
        if (p_loc < np) {
            q = P.list[pt0 + p_loc];
            const int ray = q / P.n_samples;
            const float z = P.z_vals[q];
            px = __fadd_rn(P.rays_o[ray * 3 + 0], __fmul_rn(P.rays_d[ray * 3 + 0], z));
            py = __fadd_rn(P.rays_o[ray * 3 + 1], __fmul_rn(P.rays_d[ray * 3 + 1], z));
            pz = __fadd_rn(P.rays_o[ray * 3 + 2], __fmul_rn(P.rays_d[ray * 3 + 2], z));
        }
        float* row = X + p_loc * kXStride;
        const float bx0 = blob_input(px, P.sc.bbox_lo[0], P.sc.bbox_ext[0]);
        const float by0 = blob_input(py, P.sc.bbox_lo[1], P.sc.bbox_ext[1]);
        const float bz0 = blob_input(pz, P.sc.bbox_lo[2], P.sc.bbox_ext[2]);
        if (MODE == 2) {
            if (half == 0) { row[kBlobCol] = bx0; row[kBlobCol + 1] = by0; row[kBlobCol + 2] = bz0; }
            __syncwarp();
            gather16_hash(P.sc.hash[0], X + kBlobCol, kXStride, X, kXStride, np, lane);
            if (TWO) gather16_hash(P.sc.hash[1], X + kBlobCol, kXStride, CFt, kCStride, np, lane);
            __syncwarp();
        } else {
            if (half == 1) write_taps8<TWO>(P.sc, reinterpret_cast<float4*>(row + kBlobCol), px, py, pz, 0);
            __syncwarp();
            if (MODE == 1) gather16_vec(P.sc, 0, X + kBlobCol, kXStride, X, kXStride, np, lane);
            else gather16(P.sc, 0, X + kBlobCol, kXStride, X, kXStride, np, lane);
            __syncwarp();
            if (TWO) {
                if (half == 1) write_taps8<TWO>(P.sc, reinterpret_cast<float4*>(row + kBlobCol), px, py, pz, 1);
                __syncwarp();
                if (MODE == 1) gather16_vec(P.sc, 1, X + kBlobCol, kXStride, CFt, kCStride, np, lane);
                else gather16(P.sc, 1, X + kBlobCol, kXStride, CFt, kCStride, np, lane);
                __syncwarp();
            }
        }
        {
            if (half == 0) {
                oneblob16_sparse(bx0, row + kBlobCol);
                oneblob16_sparse(by0, row + kBlobCol + 16);
            } else {
                oneblob16_sparse(bz0, row + kBlobCol + 32);
            }
            // rows without a point must be finite (they are multiplied by zero gradients later)
            for (int p = np; p < kTilePts; ++p) {
                X[p * kXStride + lane] = 0.f; X[p * kXStride + 32 + lane] = 0.f;
                if (TWO) { CFt[p * kCStride + lane] = 0.f; CFt[p * kCStride + 32 + lane] = 0.f; }
            }
        }
        __syncwarp();

        // ---- 2. the forward's activations of this tile's rows g and g+8: kept by the forward, or recomputed
        float h1[4][4], o16[2][4], hc[4][4];
        const int q_lo = __shfl_sync(0xffffffffu, q, g), q_hi = __shfl_sync(0xffffffffu, q, g + 8);
        {
            // same split-precision products as the forward kernel: the ReLU masks and activations must be the
            // forward's (a single-TF32 recompute flips masks of near-zero pre-activations: measured 1e-2..5e-2
            // max-norm gradient error on the two-grid goldens, against 2e-3 with the split)
            float rgb_unused[4];
            decoder_forward_mma<TWO>(W4, X, CFt, lane, h1, o16, hc, rgb_unused);
        }
        float4 dr_lo = make_float4(0.f, 0.f, 0.f, 0.f), dr_hi = dr_lo;
        if (g < np) dr_lo = __ldg(P.draw + q_lo);
        if (g + 8 < np) dr_hi = __ldg(P.draw + q_hi);

        // ---- 3. colour branch backward
        // A fragment of d_rgb: columns (2t, 2t+1) of (r, g, b, 0, ...)
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
            for (int i = 0; i < 4; ++i) dhc[n][i] = (hc[n][i] > 0.f) ? dhc[n][i] : 0.f;
            store_frag(ACT, kActStride, n, g, t, hc[n]);
            store_frag(DY, kActStride, n, g, t, dhc[n]);
        }
#pragma unroll
        for (int n = 0; n < 2; ++n) store_frag(GEO, kGeoStride, n, g, t, o16[n]);
        {
            const float c0[4] = {ar0, ar2, ar1, ar3};            // (g,2t) (g,2t+1) (g+8,2t) (g+8,2t+1)
            store_frag(DRGB, kRgbStride, 0, g, t, c0);
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
        if (PARAM) __syncthreads();                              // A: every warp's ACT / DY / GEO / DRGB / X are published

        // ---- 4. weight gradients, phase 1 (colour net)
        if (PARAM) {
#pragma unroll
            for (int sl = 0; sl < S1; ++sl) {
                const int blk = warp + sl * NW;
                if (blk < V1_BLOCKS) {                           // 16-row block of col0's inputs
                    for (int w = 0; w < NW; ++w) {
                        const float* src = tiles + w * WF;
                        const float* A; int as, m0;
                        if (blk < 3) { A = src; as = kXStride; m0 = kBlobCol + blk * 16; }
                        else if (TWO && blk < 7) { A = src + O_CF; as = kCStride; m0 = (blk - 3) * 16; }
                        else { A = src + O_GEO; as = kGeoStride; m0 = 0; }
                        outer_mma<4, 0>(acc1[sl], A, as, m0, src + O_DY, kActStride, g, t);
                    }
                } else if (blk < NB1) {                          // 16-row block of col1's inputs (hidden j), 8 (3 real) cols
                    for (int w = 0; w < NW; ++w) {
                        const float* src = tiles + w * WF;
                        outer_mma<1, 0>(acc1[sl], src + O_ACT, kActStride, (blk - V1_BLOCKS) * 16, src + O_DRGB, kRgbStride, g, t);
                    }
                }
            }
        }
        if (PARAM) __syncthreads();                              // B: phase-1 tiles may be overwritten

        // ---- 5. sdf branch backward
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
            for (int i = 0; i < 4; ++i) dh1[n][i] = (h1[n][i] > 0.f) ? dh1[n][i] : 0.f;
            store_frag(ACT, kActStride, n, g, t, h1[n]);
            store_frag(DY, kActStride, n, g, t, dh1[n]);
        }
#pragma unroll
        for (int n = 0; n < 2; ++n) store_frag(GEO, kGeoStride, n, g, t, dout[n]);
        if (TWO) {                                               // d colour features -> CF tile (its values are spent)
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
        if (PARAM) __syncthreads();                              // C

        // ---- 6. weight gradients, phase 2 (sdf net)
        if (PARAM) {
#pragma unroll
            for (int sl = 0; sl < S2; ++sl) {
                const int blk = warp + sl * NW;
                if (blk < 7) {
                    for (int w = 0; w < NW; ++w) {
                        const float* src = tiles + w * WF;
                        outer_mma<4, 0>(acc2[sl], src, kXStride, blk * 16, src + O_DY, kActStride, g, t);
                    }
                } else if (blk == 7) {
                    // d sdf1^T [32 j][16 i]: blocks (row block rb, col block cb) -> acc2[sl][rb * 2 + cb]
                    for (int w = 0; w < NW; ++w) {
                        const float* src = tiles + w * WF;
                        outer_mma<2, 0>(acc2[sl], src + O_ACT, kActStride, 0, src + O_GEO, kGeoStride, g, t);
                        outer_mma<2, 2>(acc2[sl], src + O_ACT, kActStride, 16, src + O_GEO, kGeoStride, g, t);
                    }
                }
            }
        }
        if (PARAM) __syncthreads();                              // D: X may be overwritten with d feat

        // ---- 7. d feat = d_h1 . W1[:, :64] -> X columns [0, 64)
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
        float dpt[3] = {0.f, 0.f, 0.f};                          // d loss / d point through OneBlob (lanes 0..15)
        unsigned gflags = 0;                                     // per tap: bit 2t = x gradient on, 2t+1 = y gradient on
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
                d0 = oneblob16_sparse_grad(blob_input(px, P.sc.bbox_lo[0], P.sc.bbox_ext[0]), row + kBlobCol);
                d1 = oneblob16_sparse_grad(blob_input(py, P.sc.bbox_lo[1], P.sc.bbox_ext[1]), row + kBlobCol + 16);
            } else {
                d0 = oneblob16_sparse_grad(blob_input(pz, P.sc.bbox_lo[2], P.sc.bbox_ext[2]), row + kBlobCol + 32);
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
            __syncwarp();
        }
        // tap table (or hash coordinates) again -- the blob segment is spent -- then scatter
        if (MODE == 2) {
            if (half == 0) { row[kBlobCol] = bx0; row[kBlobCol + 1] = by0; row[kBlobCol + 2] = bz0; }
            __syncwarp();
            if (PARAM && P.sc.hash[0].grad) scatter16_hash(P.sc.hash[0], X + kBlobCol, kXStride, X, kXStride, np, lane, P.sc.dirty, P.sc.dirty_base);
            if (TWO && PARAM && P.sc.hash[1].grad) scatter16_hash(P.sc.hash[1], X + kBlobCol, kXStride, CFt, kCStride, np, lane, P.sc.dirty, P.sc.dirty_base);
        } else if (half == 1) {
            write_taps8<TWO>(P.sc, reinterpret_cast<float4*>(row + kBlobCol), px, py, pz, 0);
        }
        __syncwarp();
        float dn_keep[3] = {0.f, 0.f, 0.f};                      // d loss / d normalised plane (or unit-cube) coordinates of this lane's point
        if (MODE == 2 && RAY) {
            hash_point_grad16(P.sc.hash[0], X + kBlobCol, kXStride, X, kXStride, np, lane, dn_keep);
            if (TWO) hash_point_grad16(P.sc.hash[1], X + kBlobCol, kXStride, CFt, kCStride, np, lane, dn_keep);
        }
        // ---- 8. plane gradient scatter: lane = (corner, 4-channel group)
        for (int grp = 0; MODE != 2 && grp < (TWO ? 2 : 1); ++grp) {
            if (grp == 1) {
                __syncwarp();
                if (half == 1) write_taps8<TWO>(P.sc, reinterpret_cast<float4*>(row + kBlobCol), px, py, pz, 1);
                __syncwarp();
            }
            const float* gsrc = (grp == 0) ? X : CFt;
            const int gstride = (grp == 0) ? kXStride : kCStride;
            const int c = lane >> 3, q = lane & 7;
            for (int p = 0; p < np; ++p) {
                const float* seg = X + p * kXStride + kBlobCol;
                const unsigned pflags = RAY ? __shfl_sync(0xffffffffu, gflags, p) : 0u;
                float dn[3] = {0.f, 0.f, 0.f};
#pragma unroll
                for (int l = 0; l < 2; ++l) {
#pragma unroll
                    for (int a = 0; a < 3; ++a) {
                        const int tp = a * 2 + l;
                        const MnePlaneDesc& Pl = P.sc.planes[grp * 6 + tp];
                        float* gp = PARAM ? Pl.grad : nullptr;
                        if (MODE == 1) {
                            const float4 gv = *reinterpret_cast<const float4*>(gsrc + p * gstride + l * 32 + q * 4);
                            const unsigned off = __float_as_uint(seg[8 * tp + c]);
                            const float w = seg[8 * tp + 4 + c];
                            if (gp && w != 0.f) {
                                atomicAdd(reinterpret_cast<float4*>(gp + off) + q, make_float4(gv.x * w, gv.y * w, gv.z * w, gv.w * w));
                                if (q == 0 && P.sc.dirty) mark_dirty(P.sc.dirty, ((gp + off) - P.sc.dirty_base) >> 5);
                            }
                            const unsigned fl = (pflags >> (2 * (grp * 6 + tp))) & 3u;
                            if (RAY && fl) {
                                // D_c = <d feat, texel c>, reduced over this corner's 8 lanes, then shared with everyone
                                const float4 v = __ldg(reinterpret_cast<const float4*>(Pl.data + off) + q);
                                float dc = gv.x * v.x + gv.y * v.y + gv.z * v.z + gv.w * v.w;
                                dc += __shfl_xor_sync(0xffffffffu, dc, 1); dc += __shfl_xor_sync(0xffffffffu, dc, 2);
                                dc += __shfl_xor_sync(0xffffffffu, dc, 4);
                                const float d00 = __shfl_sync(0xffffffffu, dc, 0), d01 = __shfl_sync(0xffffffffu, dc, 8),
                                            d10 = __shfl_sync(0xffffffffu, dc, 16), d11 = __shfl_sync(0xffffffffu, dc, 24);
                                const float4 ww = *reinterpret_cast<const float4*>(seg + 8 * tp + 4);
                                const float wy0 = ww.x + ww.y, wy1 = ww.z + ww.w, wx0 = ww.x + ww.z, wx1 = ww.y + ww.w;
                                const int ax = (a == 2) ? 1 : 0, ay = (a == 0) ? 1 : 2;
                                if (fl & 1u) dn[ax] += ((d01 - d00) * wy0 + (d11 - d10) * wy1) * (0.5f * (float)(Pl.W - 1));
                                if (fl & 2u) dn[ay] += ((d10 - d00) * wx0 + (d11 - d01) * wx1) * (0.5f * (float)(Pl.H - 1));
                            }
                        } else if (RAY) {
                            // strided planes: lane = channel, corner dot products by a full-warp reduction
                            const float gch = gsrc[p * gstride + l * 32 + lane];
                            const float* vb = Pl.data + (int64_t)lane * Pl.stride_c;
                            float dcs[4];
#pragma unroll
                            for (int k = 0; k < 4; ++k) dcs[k] = warp_sum(gch * __ldg(vb + __float_as_uint(seg[8 * tp + k])));
                            if (gp) {
                                float* gb = gp + (int64_t)lane * Pl.stride_c;
#pragma unroll
                                for (int k = 0; k < 4; ++k) {
                                    const float w = seg[8 * tp + 4 + k];
                                    if (w != 0.f) atomicAdd(gb + __float_as_uint(seg[8 * tp + k]), gch * w);
                                }
                            }
                            const unsigned fl = (pflags >> (2 * (grp * 6 + tp))) & 3u;
                            const float4 ww = *reinterpret_cast<const float4*>(seg + 8 * tp + 4);
                            const float wy0 = ww.x + ww.y, wy1 = ww.z + ww.w, wx0 = ww.x + ww.z, wx1 = ww.y + ww.w;
                            const int ax = (a == 2) ? 1 : 0, ay = (a == 0) ? 1 : 2;
                            if (fl & 1u) dn[ax] += ((dcs[1] - dcs[0]) * wy0 + (dcs[3] - dcs[2]) * wy1) * (0.5f * (float)(Pl.W - 1));
                            if (fl & 2u) dn[ay] += ((dcs[2] - dcs[0]) * wx0 + (dcs[3] - dcs[1]) * wx1) * (0.5f * (float)(Pl.H - 1));
                        } else if (!gp) {
                            continue;
                        } else {
                            const float gch = gsrc[p * gstride + l * 32 + lane];
                            float* gb = gp + (int64_t)lane * Pl.stride_c;
#pragma unroll
                            for (int k = 0; k < 4; ++k) {
                                const float w = seg[8 * tp + 4 + k];
                                if (w != 0.f) atomicAdd(gb + __float_as_uint(seg[8 * tp + k]), gch * w);
                            }
                        }
                    }
                }
                if (RAY && p_loc == p) { dn_keep[0] += dn[0]; dn_keep[1] += dn[1]; dn_keep[2] += dn[2]; }
            }
        }
        if (RAY && half == 0 && p_loc < np) {
            const int ray = q / P.n_samples;
            const float z = P.z_vals[q];
#pragma unroll
            for (int a = 0; a < 3; ++a) {
                // planes: coordinates in [-1, 1] over `bound`; hash grid: in [0, 1] over `bounding_box` (like the OneBlob input)
                const float gpnt = (MODE == 2 ? (float)((double)dn_keep[a] / P.sc.bbox_ext[a]) : dn_keep[a] * (2.0f / P.sc.bound_ext[a])) + dpt[a];
                atomicAdd(P.g_rays_o + ray * 3 + a, gpnt);
                atomicAdd(P.g_rays_d + ray * 3 + a, gpnt * z);
            }
        }
        if (RAY) __syncwarp();
    }

    // ---- flush the register-resident weight gradients (once per CTA)
    if (PARAM) {
#pragma unroll
        for (int sl = 0; sl < S1; ++sl) {
            const int blk = warp + sl * NW;
            if (blk < V1_BLOCKS && P.sc.g_col0) {                // rows = inputs (blk*16 + g, +8), cols = j
#pragma unroll
                for (int nb = 0; nb < 4; ++nb)
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const int k = blk * 16 + g + ((i & 2) ? 8 : 0), j = nb * 8 + 2 * t + (i & 1);
                        const bool hash2 = MODE == 2;
                        const int col = (k < N::IN_COL) ? col0_col(k, TWO, hash2) : -1;
                        if (col >= 0) atomicAdd(P.sc.g_col0 + j * col0_in(TWO, hash2) + col, acc1[sl][nb][i]);
                    }
            } else if (blk >= V1_BLOCKS && blk < NB1 && P.sc.g_col1) {
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int j = (blk - V1_BLOCKS) * 16 + g + ((i & 2) ? 8 : 0), c = 2 * t + (i & 1);
                    if (c < 3) atomicAdd(P.sc.g_col1 + c * 32 + j, acc1[sl][0][i]);
                }
            }
        }
#pragma unroll
        for (int sl = 0; sl < S2; ++sl) {
            const int blk = warp + sl * NW;
            if (blk < 7 && P.sc.g_sdf0) {
#pragma unroll
                for (int nb = 0; nb < 4; ++nb)
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const int k = blk * 16 + g + ((i & 2) ? 8 : 0), j = nb * 8 + 2 * t + (i & 1);
                        const bool hash = MODE == 2;
                        const int col = sdf0_col(k, hash);
                        if (col >= 0) atomicAdd(P.sc.g_sdf0 + j * sdf0_in(hash) + col, acc2[sl][nb][i]);
                    }
            } else if (blk == 7 && P.sc.g_sdf1) {
#pragma unroll
                for (int rb = 0; rb < 2; ++rb)
#pragma unroll
                    for (int cb = 0; cb < 2; ++cb)
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            const int j = rb * 16 + g + ((i & 2) ? 8 : 0), c = cb * 8 + 2 * t + (i & 1);
                            atomicAdd(P.sc.g_sdf1 + ((c == 15) ? 0 : c + 1) * 32 + j, acc2[sl][rb * 2 + cb][i]);
                        }
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------
// ray kernel: d loss / d raw for every sample + compaction (ray_tail.cuh), for calls whose upstream gradient arrives
// with the backward (autograd).  The fused training forward (mne_render_forward_fused) does this in its own tail.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) rays_bwd_kernel(const RayBwdParams P) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const RayGradScales K = ray_grad_scales(P.rc, P.n_samples, P.acc, P.loss_grads);
    for (int ray = blockIdx.x * 8 + warp; ray < P.n_rays; ray += gridDim.x * 8) {
        const RayComposite c = composite_ray(P.raw, P.z_vals, (size_t)ray * P.n_samples, P.n_samples, P.rc, lane);
        ray_backward_one(P, ray, lane, K, c);
    }
}

int check_scene(const MneSceneDesc* sc, const char* who, bool need_grads);
int sm_count();
int points_backward_v3(const PtsBwdParams& P, bool ray, bool param, float* act, const float4* draw, cudaStream_t st);

template <bool TWO, int MODE, bool RAY, bool PARAM, int NW = kBwdWarps>
static int launch_points_bwd(const PtsBwdParams& P, cudaStream_t stream) {
    const size_t smem = bwd2_smem_bytes<TWO, RAY, NW>();
    static std::atomic<bool> configured[64] = {};
    int dev = 0;
    MNE_CHECK_CUDA(cudaGetDevice(&dev));
    if (dev >= 64 || !configured[dev]) {
        MNE_CHECK_CUDA(cudaFuncSetAttribute(points_bwd_kernel<TWO, MODE, RAY, PARAM, NW>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        if (dev < 64) configured[dev] = true;
    }
    const int64_t n_tiles = (P.n_pts + kTilePts - 1) / kTilePts;
    const int blocks = (int)max((int64_t)1, min((n_tiles + NW - 1) / NW, (int64_t)sm_count()));
    points_bwd_kernel<TWO, MODE, RAY, PARAM, NW><<<blocks, NW * 32, smem, stream>>>(P);
    MNE_CHECK_CUDA(cudaGetLastError());
    return 0;
}

template <int MODE>
static int dispatch_points_bwd(const PtsBwdParams& P, bool ray, bool param, cudaStream_t st) {
    if (P.sc.two_grid) {
        // (two grids + ray AND parameter gradients together exceed the shared-memory plan: render_bwd.cu keeps
        //  that one combination on the v1 kernel)
        if (ray) return launch_points_bwd<true, MODE, true, false, 5>(P, st);
        return launch_points_bwd<true, MODE, false, true, 6>(P, st);
    }
    if (ray && param) return launch_points_bwd<false, MODE, true, true>(P, st);
    if (ray) return launch_points_bwd<false, MODE, true, false>(P, st);
    return launch_points_bwd<false, MODE, false, true>(P, st);
}

bool scene_is_vectorizable(const MneSceneDesc* sc) { return planes_vectorizable(*sc); }

// scratch layout (mne_backward_scratch_bytes): d_raw float4[n] | sample list int32[round4(n)] | list length
static void scratch_views(void* scratch, int64_t n_pts, float4** draw, int** list, int** n_active) {
    *draw = reinterpret_cast<float4*>(scratch);
    *list = reinterpret_cast<int*>(*draw + n_pts);
    *n_active = *list + ((n_pts + 3) & ~(int64_t)3);
}

// the point stage of the backward over a filled scratch (d_raw + sample list): with the forward's record the barrier-free
// kernels of render_bwd_v3.cu, else the recompute kernels above
int render_backward_points(const MneSceneDesc* scene, int32_t n_rays, int32_t n_samples, const float* rays_o, const float* rays_d,
                           const float* z_vals, void* scratch, float* g_rays_o, float* g_rays_d, bool param_grad, float* act,
                           cudaStream_t st) {
    const int64_t n_pts = (int64_t)n_rays * n_samples;
    float4* draw; int* list; int* n_active;
    scratch_views(scratch, n_pts, &draw, &list, &n_active);
    PtsBwdParams P{};
    P.sc = *scene; P.n_pts = n_pts; P.n_samples = n_samples;
    P.rays_o = rays_o; P.rays_d = rays_d; P.z_vals = z_vals; P.draw = draw;
    P.g_rays_o = g_rays_o; P.g_rays_d = g_rays_d; P.list = list; P.n_active = n_active; P.act = act;
    const bool ray = g_rays_o != nullptr;
    if (act && (scene->encoder == 1 || planes_vectorizable(*scene)))
        return points_backward_v3(P, ray, param_grad, act, draw, st);
    if (scene->encoder == 1) {
        MNE_CHECK_ARG(!scene->two_grid, "mne_render_backward: the two-table hash grid needs the forward's activation buffer (act)");
        if (ray && param_grad) return launch_points_bwd<false, 2, true, true>(P, st);
        if (ray) return launch_points_bwd<false, 2, true, false>(P, st);
        return launch_points_bwd<false, 2, false, true>(P, st);
    }
    if (planes_vectorizable(*scene)) return dispatch_points_bwd<1>(P, ray, param_grad, st);
    MNE_CHECK_ARG(!scene->dirty, "mne_render_backward: MneSceneDesc.dirty needs channels-last planes or the hash grid");
    return dispatch_points_bwd<0>(P, ray, param_grad, st);
}

// v2 backward: ray kernel (d loss / d raw + compaction into scratch), then the point stage
int render_backward_v2(const MneSceneDesc* scene, const MneRenderCfg* rcfg, int32_t n_rays, int32_t n_samples,
                       const float* rays_o, const float* rays_d, const float* z_vals, const float* raw,
                       const float* target_rgb, const float* target_d, const double* acc, const float* loss_grads,
                       const float* g_rgb, const float* g_depth, const float* g_raw, void* scratch, float* g_rays_o,
                       float* g_rays_d, bool param_grad, float* act, cudaStream_t st) {
    RayBwdParams R{};
    R.rc = *rcfg; R.n_rays = n_rays; R.n_samples = n_samples; R.z_vals = z_vals; R.raw = raw; R.target_rgb = target_rgb;
    R.target_d = target_d; R.acc = acc; R.loss_grads = loss_grads; R.g_rgb = g_rgb; R.g_depth = g_depth; R.g_raw = g_raw;
    const int64_t n_pts = (int64_t)n_rays * n_samples;
    MNE_CHECK_ARG(n_pts < ((int64_t)1 << 31), "mne_render_backward: more than 2^31 samples in one call");
    scratch_views(scratch, n_pts, &R.draw, &R.list, &R.n_active);
    MNE_CHECK_CUDA(cudaMemsetAsync(R.n_active, 0, sizeof(int), st));
    const int blocks = max(1, min((n_rays + 7) / 8, sm_count() * 4));
    rays_bwd_kernel<<<blocks, 256, 0, st>>>(R);
    MNE_CHECK_CUDA(cudaGetLastError());
    return render_backward_points(scene, n_rays, n_samples, rays_o, rays_d, z_vals, scratch, g_rays_o, g_rays_d, param_grad, act, st);
}

}  // namespace mne
