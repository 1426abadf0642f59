// v2 forward: (1) a point kernel -- 16-point tiles per warp: taps + OneBlob -> tri-plane gather -> decoder on
// the tensor pipe (mma.sync TF32, 3-term split) -> raw; (2) the per-ray tail -- sdf2weights compositing, rendered
// maps, loss sums and (training) d loss / d raw + the list of samples where it is non-zero -- run by the warp that stores
// the LAST sample of a ray, inside the same kernel (a per-ray counter of stored samples decides who that is), so that the
// render is one launch and `raw` is consumed out of L2.  Replaces the same reference code as render_fwd.cu (v1).
#include "ray_tail.cuh"
#include "tile16.cuh"

namespace mne {

__device__ __forceinline__ int atom_add_release_gpu(int* p, int v) {
    int old;
    asm volatile("atom.release.gpu.global.add.s32 %0, [%1], %2;" : "=r"(old) : "l"(p), "r"(v) : "memory");
    return old;
}
__device__ __forceinline__ int atom_add_relaxed_gpu(int* p, int v) {
    int old;
    asm volatile("atom.relaxed.gpu.global.add.s32 %0, [%1], %2;" : "=r"(old) : "l"(p), "r"(v) : "memory");
    return old;
}

template <bool TWO> __host__ __device__ constexpr int pts_warps() { return TWO ? 8 : 16; }
template <bool TWO> __host__ __device__ constexpr int pts_warp_floats() { return kTilePts * kXStride + (TWO ? kTilePts * kCStride : 0); }
template <bool TWO>
constexpr size_t pts_smem_bytes() {
    return sizeof(float4) * (size_t)NetV2<TWO>::W4_COUNT + sizeof(float) * (size_t)(pts_warps<TWO>() * pts_warp_floats<TWO>());
}

struct PtsFwdParams {
    MneSceneDesc sc;
    int64_t n_pts;
    int n_samples;            // points per ray (ray = point / n_samples); unused for explicit points
    const float* rays_o;
    const float* rays_d;
    const float* z_vals;
    const float* pts;         // explicit points [M,3] (query mode) or NULL
    float* raw;               // [M,4]
    float* geo;               // [M,15] or NULL
    float* feat;              // [M,64] or NULL
    float* act;               // [M, ActRow::STRIDE] per-sample record kept for the backward (tile16.cuh: ActRow), or NULL
    int lvl0_stage;           // hash grid: 1 = level 0 of every table is staged in shared memory (the launcher sized it in)
    // ---- per-ray tail (TAIL > 0)
    RayFwdParams rf;          // compositing outputs, targets, loss accumulators
    int* ray_done;            // [n_rays] samples of each ray stored so far; zero on entry, reset by the warp that finishes the ray
    // TAIL == 2 (training): d loss / d raw from the loss weights, compacted (layout of mne_backward_scratch_bytes)
    const float* loss_grads;
    float4* draw;
    int* list;
    int* n_active;            // zeroed before the launch
};

// MODE: 0 = planes through strides, 1 = channels-last planes (vector gather), 2 = hash grid
// TAIL: 0 = points only (queries), 1 = + compositing / loss sums per finished ray, 2 = + d loss / d raw and compaction
template <bool TWO, int MODE, int TAIL>
__global__ void __launch_bounds__(pts_warps<TWO>() * 32, 1) points_fwd_kernel(const __grid_constant__ PtsFwdParams P) {
    using N = NetV2<TWO>;
    using AR = ActRow<TWO, MODE == 2>;
    constexpr int NW = pts_warps<TWO>();
    extern __shared__ __align__(16) float4 smem4[];
    float4* W4 = smem4;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float* X = reinterpret_cast<float*>(smem4 + N::W4_COUNT) + warp * pts_warp_floats<TWO>();
    float* CFt = X + kTilePts * kXStride;

    __shared__ double warp_sums[NW * RS_COUNT];          // loss sums of the rays each warp finished
    __shared__ double red_sums[RS_COUNT];
    const bool with_loss = TAIL > 0 && P.rf.acc && P.rf.target_rgb && P.rf.target_d;
    if (TAIL > 0 && lane < RS_COUNT) warp_sums[warp * RS_COUNT + lane] = 0.0;

    load_weights_v2<TWO>(P.sc, W4, threadIdx.x, blockDim.x);
    // hash grid: the coarsest level of the table(s) -- dense, 17^3 entries, read 8 times by every sample -- in shared memory
    float2* lvl0_a = nullptr;
    float2* lvl0_b = nullptr;
    if (MODE == 2 && P.lvl0_stage) {
        lvl0_a = reinterpret_cast<float2*>(reinterpret_cast<float*>(smem4 + N::W4_COUNT) + NW * pts_warp_floats<TWO>());
        stage_hash_level0(P.sc.hash[0], lvl0_a, threadIdx.x, blockDim.x);
        if (TWO) {
            lvl0_b = lvl0_a + (P.sc.hash[0].offsets[1] - P.sc.hash[0].offsets[0]);
            stage_hash_level0(P.sc.hash[1], lvl0_b, threadIdx.x, blockDim.x);
        }
    }
    __syncthreads();

    const int64_t n_tiles = (P.n_pts + kTilePts - 1) / kTilePts;
    const int64_t total_warps = (int64_t)gridDim.x * NW;
    const int p_loc = lane & 15, half = lane >> 4;
    const int g = lane >> 2, t = lane & 3;
    RayGradScales K;
    if (TAIL == 2) K = ray_grad_scales(P.rf.rc, P.n_samples, P.rf.acc, P.loss_grads);

    for (int64_t tile = (int64_t)warp * gridDim.x + blockIdx.x; tile < n_tiles; tile += total_warps) {
        const int64_t pt0 = tile * kTilePts;
        const int np = (int)min((int64_t)kTilePts, P.n_pts - pt0);
        // ---- lanes (point, half): position
        float px = 0.f, py = 0.f, pz = 0.f;
        if (p_loc < np) {
            const int64_t q = pt0 + p_loc;
            if (P.pts) {
                px = P.pts[q * 3 + 0]; py = P.pts[q * 3 + 1]; pz = P.pts[q * 3 + 2];
            } else {
                const int64_t ray = q / P.n_samples;
                const float z = P.z_vals[q];
                px = __fadd_rn(P.rays_o[ray * 3 + 0], __fmul_rn(P.rays_d[ray * 3 + 0], z));
                py = __fadd_rn(P.rays_o[ray * 3 + 1], __fmul_rn(P.rays_d[ray * 3 + 1], z));
                pz = __fadd_rn(P.rays_o[ray * 3 + 2], __fmul_rn(P.rays_d[ray * 3 + 2], z));
            }
        }
        float* row = X + p_loc * kXStride;
        const float bx = blob_input(px, P.sc.bbox_lo[0], P.sc.bbox_ext[0]);
        const float by = blob_input(py, P.sc.bbox_lo[1], P.sc.bbox_ext[1]);
        const float bz = blob_input(pz, P.sc.bbox_lo[2], P.sc.bbox_ext[2]);
        float* arow = P.act ? P.act + pt0 * AR::STRIDE : nullptr;       // this tile's rows of the kept record
        if (arow && half == 0 && p_loc < np) {                          // the OneBlob inputs, for the weight-gradient kernel
            float* xb = arow + (int64_t)p_loc * AR::STRIDE + AR::BLOBX;
            xb[0] = bx; xb[1] = by; xb[2] = bz;
        }
        if (MODE == 2) {
            // ---- hash grid: the table is addressed with the [0,1] coordinates the OneBlob uses; park them in the
            //      point's blob segment, gather, then let the OneBlob overwrite them
            if (half == 0) { row[kBlobCol] = bx; row[kBlobCol + 1] = by; row[kBlobCol + 2] = bz; }
            __syncwarp();
            gather16_hash(P.sc.hash[0], X + kBlobCol, kXStride, X, kXStride, np, lane, lvl0_a);
            if (TWO) gather16_hash(P.sc.hash[1], X + kBlobCol, kXStride, CFt, kCStride, np, lane, lvl0_b);      // colour table
            __syncwarp();
            if (arow) {
                for (int p = 0; p < np; ++p) {
                    arow[(int64_t)p * AR::STRIDE + AR::FEAT + lane] = X[p * kXStride + lane];
                    if (TWO) arow[(int64_t)p * AR::STRIDE + AR::CFEAT + lane] = CFt[p * kCStride + lane];
                }
            }
        } else {
            // ---- tap table into the point's blob segment (lane (point, half) writes level `half`), gather, then (two
            //      grids) the same for the colour planes
            write_taps8_level(P.sc, reinterpret_cast<float4*>(row + kBlobCol), px, py, pz, 0, half);
            __syncwarp();
            if (MODE == 1) gather16_pq(P.sc, 0, X + kBlobCol, kXStride, X, kXStride, np, lane, arow ? arow + AR::FEAT : nullptr, AR::STRIDE);
            else gather16(P.sc, 0, X + kBlobCol, kXStride, X, kXStride, np, lane);
            __syncwarp();
            if (TWO) {
                write_taps8_level(P.sc, reinterpret_cast<float4*>(row + kBlobCol), px, py, pz, 1, half);
                __syncwarp();
                if (MODE == 1) gather16_pq(P.sc, 1, X + kBlobCol, kXStride, CFt, kCStride, np, lane, arow ? arow + AR::CFEAT : nullptr, AR::STRIDE);
                else gather16(P.sc, 1, X + kBlobCol, kXStride, CFt, kCStride, np, lane);
                __syncwarp();
            }
            if (MODE == 0 && arow) {
                for (int p = 0; p < np; ++p) {
                    arow[(int64_t)p * AR::STRIDE + AR::FEAT + lane] = X[p * kXStride + lane];
                    arow[(int64_t)p * AR::STRIDE + AR::FEAT + 32 + lane] = X[p * kXStride + 32 + lane];
                    if (TWO) {
                        arow[(int64_t)p * AR::STRIDE + AR::CFEAT + lane] = CFt[p * kCStride + lane];
                        arow[(int64_t)p * AR::STRIDE + AR::CFEAT + 32 + lane] = CFt[p * kCStride + 32 + lane];
                    }
                }
            }
        }
        // ---- OneBlob over the (now dead) tap table: half 0 -> dims 0,1 ; half 1 -> dim 2
        if (half == 0) {
            oneblob16_sparse(bx, row + kBlobCol);
            oneblob16_sparse(by, row + kBlobCol + 16);
        } else {
            oneblob16_sparse(bz, row + kBlobCol + 32);
        }
        __syncwarp();
        if (P.feat) {
            for (int p = 0; p < np; ++p) {
                if (MODE == 2) {
                    P.feat[(pt0 + p) * 32 + lane] = X[p * kXStride + lane];
                } else {
                    P.feat[(pt0 + p) * 64 + lane] = X[p * kXStride + lane];
                    P.feat[(pt0 + p) * 64 + 32 + lane] = X[p * kXStride + 32 + lane];
                }
            }
        }
        // ---- decoder
        float h1[4][4], o16[2][4], hc[4][4], rgb[4];
        decoder_forward_mma<TWO>(W4, X, CFt, lane, h1, o16, hc, rgb);
        // rows g, g+8: rgb columns 0,1 on t = 0, column 2 on t = 1; sdf = column 15 of o16 = o16[1][1|3] on t = 3
        const int quad = lane & ~3;
        const float b_lo = __shfl_sync(0xffffffffu, rgb[0], quad | 1), b_hi = __shfl_sync(0xffffffffu, rgb[2], quad | 1);
        const float s_lo = __shfl_sync(0xffffffffu, o16[1][1], quad | 3), s_hi = __shfl_sync(0xffffffffu, o16[1][3], quad | 3);
        if (P.raw && t == 0) {
            if (g < np) reinterpret_cast<float4*>(P.raw)[pt0 + g] = make_float4(rgb[0], rgb[1], b_lo, s_lo);
            if (g + 8 < np) reinterpret_cast<float4*>(P.raw)[pt0 + g + 8] = make_float4(rgb[2], rgb[3], b_hi, s_hi);
        }
        if (P.geo) {
#pragma unroll
            for (int n = 0; n < 2; ++n)
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int col = n * 8 + 2 * t + (i & 1), r = g + ((i & 2) ? 8 : 0);
                    if (col < kGeo && r < np) P.geo[(pt0 + r) * kGeo + col] = o16[n][i];
                }
        }
        // ---- per-ray tail, part 1: every lane that stored raw rows counts them into their rays; whoever completes a ray
        //      composites it below.  The counts are RELEASE atomics: this lane's stores are performed (MEMBAR.ALL.GPU)
        //      before its count becomes visible.  A gpu-scope fence would do that too, but it also invalidates the SM's L1
        //      (CCTL.IVALL) under the gathers of every other warp -- measured +18 us on the kernel.  The finisher reads raw
        //      with ld.global.cg (L2), after its own atomic has returned, so it needs no invalidation either.
        const int S = P.n_samples;
        int ray_lo = -1, ray_hi = -1;                            // rays this lane's counts completed
        if (TAIL > 0 && t == 0 && g < np) {
            const int r_a = (int)((pt0 + g) / S);
            const int r_b = (g + 8 < np) ? (int)((pt0 + g + 8) / S) : -1;
            const int cnt = (r_b == r_a) ? 2 : 1;
            if (atom_add_release_gpu(P.ray_done + r_a, cnt) + cnt == S) ray_lo = r_a;
            if (r_b >= 0 && r_b != r_a && atom_add_relaxed_gpu(P.ray_done + r_b, 1) + 1 == S) ray_hi = r_b;
        }
        if (P.act) store_act<AR::STRIDE, AR::MASK>(P.act, pt0 + g, pt0 + g + 8, g < np, g + 8 < np, t, h1, o16, hc);
        if (TAIL > 0) {
            {
                unsigned todo_lo = __ballot_sync(0xffffffffu, ray_lo >= 0), todo_hi = __ballot_sync(0xffffffffu, ray_hi >= 0);
                while (todo_lo | todo_hi) {
                    int ray;
                    if (todo_lo) { ray = __shfl_sync(0xffffffffu, ray_lo, __ffs(todo_lo) - 1); todo_lo &= todo_lo - 1; }
                    else { ray = __shfl_sync(0xffffffffu, ray_hi, __ffs(todo_hi) - 1); todo_hi &= todo_hi - 1; }
                    if (lane == 0) P.ray_done[ray] = 0;      // nobody else touches this ray's counter in this launch
                    float sums[RS_COUNT];
                    if (S <= 64) {
                        ray_tail_regs<2, TAIL == 2>(P.rf, ray, lane, with_loss, sums, K, P.loss_grads, P.draw, P.list, P.n_active);
                    } else if (S <= 128) {
                        ray_tail_regs<4, TAIL == 2>(P.rf, ray, lane, with_loss, sums, K, P.loss_grads, P.draw, P.list, P.n_active);
                    } else {
                        const RayComposite c = ray_forward_one(P.rf, ray, lane, with_loss, sums);
                        if (TAIL == 2) {
                            RayBwdParams B;
                            B.rc = P.rf.rc; B.n_rays = P.rf.n_rays; B.n_samples = S; B.z_vals = P.rf.z_vals; B.raw = P.rf.raw;
                            B.target_rgb = P.rf.target_rgb; B.target_d = P.rf.target_d; B.acc = P.rf.acc; B.loss_grads = P.loss_grads;
                            B.g_rgb = nullptr; B.g_depth = nullptr; B.g_raw = nullptr; B.draw = P.draw; B.list = P.list; B.n_active = P.n_active;
                            ray_backward_one(B, ray, lane, K, c);
                        }
                    }
                    if (with_loss && lane < RS_COUNT) {
                        double v = 0.0;
#pragma unroll
                        for (int k = 0; k < RS_COUNT; ++k) v = (lane == k) ? (double)sums[k] : v;
                        warp_sums[warp * RS_COUNT + lane] += v;
                    }
                }
            }
        }
        __syncwarp();
    }
    if (TAIL > 0 && with_loss) flush_loss_sums(warp_sums, NW, P.rf.rc, P.rf.acc, red_sums);
}

// ---------------------------------------------------------------------------------------------
// stand-alone ray kernel: compositing + loss sums on caller-supplied raw (mne_composite, impl 1 cross-checks)
// ---------------------------------------------------------------------------------------------
constexpr int kRayWarps = 8;

__global__ void __launch_bounds__(kRayWarps * 32) rays_fwd_kernel(const RayFwdParams P) {
    __shared__ double warp_sums[kRayWarps * RS_COUNT];
    __shared__ double red_sums[RS_COUNT];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const bool with_loss = P.acc && P.target_rgb && P.target_d;
    double mine = 0.0;                                   // lane k < RS_COUNT accumulates sum k
    for (int ray = blockIdx.x * kRayWarps + warp; ray < P.n_rays; ray += gridDim.x * kRayWarps) {
        float sums[RS_COUNT];
        ray_forward_one(P, ray, lane, with_loss, sums);
#pragma unroll
        for (int k = 0; k < RS_COUNT; ++k) if (lane == k) mine += (double)sums[k];
    }
    if (with_loss) {
        if (lane < RS_COUNT) warp_sums[warp * RS_COUNT + lane] = mine;
        flush_loss_sums(warp_sums, kRayWarps, P.rc, P.acc, red_sums);
    }
}

// The whole per-ray stage in one launch, one warp per ray, the ray held in registers (ray_tail.cuh: ray_tail_regs): compositing
// + loss sums and, with loss_grads, d loss / d raw + compaction -- what rays_fwd_kernel and rays_bwd_kernel do in two.
struct RayTailParams {
    RayFwdParams rf;
    const float* loss_grads;
    float4* draw;
    int* list;
    int* n_active;
};

template <bool TRAIN>
__global__ void __launch_bounds__(kRayWarps * 32) rays_tail_kernel(const __grid_constant__ RayTailParams P) {
    __shared__ double warp_sums[kRayWarps * RS_COUNT];
    __shared__ double red_sums[RS_COUNT];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int S = P.rf.n_samples;
    const bool with_loss = P.rf.acc && P.rf.target_rgb && P.rf.target_d;
    RayGradScales K;
    if (TRAIN) K = ray_grad_scales(P.rf.rc, S, P.rf.acc, P.loss_grads);
    double mine = 0.0;                                   // lane k < RS_COUNT accumulates sum k
    for (int ray = blockIdx.x * kRayWarps + warp; ray < P.rf.n_rays; ray += gridDim.x * kRayWarps) {
        float sums[RS_COUNT];
        if (S <= 64) {
            ray_tail_regs<2, TRAIN>(P.rf, ray, lane, with_loss, sums, K, P.loss_grads, P.draw, P.list, P.n_active);
        } else if (S <= 128) {
            ray_tail_regs<4, TRAIN>(P.rf, ray, lane, with_loss, sums, K, P.loss_grads, P.draw, P.list, P.n_active);
        } else {
            const RayComposite c = ray_forward_one(P.rf, ray, lane, with_loss, sums);
            if (TRAIN) {
                RayBwdParams B;
                B.rc = P.rf.rc; B.n_rays = P.rf.n_rays; B.n_samples = S; B.z_vals = P.rf.z_vals; B.raw = P.rf.raw;
                B.target_rgb = P.rf.target_rgb; B.target_d = P.rf.target_d; B.acc = P.rf.acc; B.loss_grads = P.loss_grads;
                B.g_rgb = nullptr; B.g_depth = nullptr; B.g_raw = nullptr; B.draw = P.draw; B.list = P.list; B.n_active = P.n_active;
                ray_backward_one(B, ray, lane, K, c);
            }
        }
#pragma unroll
        for (int k = 0; k < RS_COUNT; ++k) if (lane == k) mine += (double)sums[k];
    }
    if (with_loss) {
        if (lane < RS_COUNT) warp_sums[warp * RS_COUNT + lane] = mine;
        flush_loss_sums(warp_sums, kRayWarps, P.rf.rc, P.rf.acc, red_sums);
    }
}

int check_scene(const MneSceneDesc* sc, const char* who, bool need_grads);
int sm_count();
int impl_version();
int hash_level0_smem();
int render_forward_v1(const MneSceneDesc*, const MneRenderCfg*, int32_t, int32_t, const float*, const float*, const float*,
                      const float*, const float*, float*, float*, float*, float*, float*, float*, float*, int32_t*, double*,
                      mne_stream_t);
int query_points_v1(const MneSceneDesc*, int64_t, const float*, float*, float*, float*, mne_stream_t);

template <bool TWO, int MODE, int TAIL>
static int launch_points_fwd(const PtsFwdParams& P_in, cudaStream_t stream) {
    PtsFwdParams P = P_in;
    size_t smem = pts_smem_bytes<TWO>();
    constexpr size_t kSmemMax = 227 * 1024 - 2048;       // the opt-in limit per CTA minus the kernel's static shared memory (loss sums)
    if (MODE == 2) {
        // level 0 of the table(s) goes to shared memory when it fits beside the weights and the tiles
        size_t extra = sizeof(float2) * (size_t)(P.sc.hash[0].offsets[1] - P.sc.hash[0].offsets[0]);
        if (TWO) extra += sizeof(float2) * (size_t)(P.sc.hash[1].offsets[1] - P.sc.hash[1].offsets[0]);
        // (opt-in: measured on B200 it LOSES -- 734 -> 1170 us for 958 k samples -- because the extra 39 KB leave the SM ~9 KB of
        //  L1, and the L1 is what serves the other coarse levels; option "hash_level0_smem")
        P.lvl0_stage = (hash_level0_smem() && smem + extra <= kSmemMax) ? 1 : 0;
        if (P.lvl0_stage) smem += extra;
    }
    static std::atomic<bool> configured[64] = {};
    int dev = 0;
    MNE_CHECK_CUDA(cudaGetDevice(&dev));
    if (dev >= 64 || !configured[dev]) {
        MNE_CHECK_CUDA(cudaFuncSetAttribute(points_fwd_kernel<TWO, MODE, TAIL>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                            (int)(MODE == 2 ? kSmemMax : smem)));
        if (dev < 64) configured[dev] = true;
    }
    const int64_t n_tiles = (P.n_pts + kTilePts - 1) / kTilePts;
    const int blocks = (int)max((int64_t)1, min(n_tiles, (int64_t)sm_count()));
    points_fwd_kernel<TWO, MODE, TAIL><<<blocks, pts_warps<TWO>() * 32, smem, stream>>>(P);
    MNE_CHECK_CUDA(cudaGetLastError());
    return 0;
}

template <int TAIL>
static int dispatch_points_fwd(const PtsFwdParams& P, cudaStream_t stream) {
    if (P.sc.encoder == 1) return P.sc.two_grid ? launch_points_fwd<true, 2, TAIL>(P, stream) : launch_points_fwd<false, 2, TAIL>(P, stream);
    const bool vec = planes_vectorizable(P.sc);
    if (P.sc.two_grid) return vec ? launch_points_fwd<true, 1, TAIL>(P, stream) : launch_points_fwd<true, 0, TAIL>(P, stream);
    return vec ? launch_points_fwd<false, 1, TAIL>(P, stream) : launch_points_fwd<false, 0, TAIL>(P, stream);
}

}  // namespace mne

extern "C" int mne_render_forward(const MneSceneDesc* scene, const MneRenderCfg* rcfg, int32_t n_rays, int32_t n_samples,
                                  const float* rays_o, const float* rays_d, const float* z_vals, const float* target_rgb,
                                  const float* target_d, float* raw, float* weights, float* rgb, float* depth, float* disp,
                                  float* acc_map, float* depth_var, int32_t* inds, double* acc, float* act,
                                  mne_stream_t stream) {
    using namespace mne;
    if (impl_version() == 1)
        return render_forward_v1(scene, rcfg, n_rays, n_samples, rays_o, rays_d, z_vals, target_rgb, target_d, raw, weights,
                                 rgb, depth, disp, acc_map, depth_var, inds, acc, stream);
    MNE_CHECK_ARG(scene && rcfg, "mne_render_forward: null scene/rcfg");
    if (int e = check_scene(scene, "mne_render_forward", false)) return e;
    MNE_CHECK_ARG(n_rays >= 0 && n_samples > 0, "mne_render_forward: bad sizes n_rays=%d n_samples=%d", n_rays, n_samples);
    if (n_rays == 0) return 0;
    MNE_CHECK_ARG(rays_o && rays_d && z_vals && raw, "mne_render_forward: rays_o, rays_d, z_vals and raw are required");
    MNE_CHECK_ARG(((uintptr_t)raw & 15) == 0, "mne_render_forward: raw must be 16-byte aligned");
    MNE_CHECK_ARG((int64_t)n_rays * n_samples < (int64_t)1 << 31, "mne_render_forward: n_rays * n_samples must be < 2^31");
    PtsFwdParams P{};
    P.sc = *scene; P.n_pts = (int64_t)n_rays * n_samples; P.n_samples = n_samples;
    P.rays_o = rays_o; P.rays_d = rays_d; P.z_vals = z_vals; P.raw = raw; P.act = act;
    MNE_CHECK_ARG(((uintptr_t)act & 31) == 0, "mne_render_forward: act must be 32-byte aligned");
    if (int e = dispatch_points_fwd<0>(P, (cudaStream_t)stream)) return e;
    RayFwdParams R{};
    R.rc = *rcfg; R.n_rays = n_rays; R.n_samples = n_samples; R.z_vals = z_vals; R.raw = raw;
    R.target_rgb = target_rgb; R.target_d = target_d; R.weights = weights; R.rgb = rgb; R.depth = depth; R.disp = disp;
    R.acc_map = acc_map; R.depth_var = depth_var; R.inds = inds; R.acc = acc;
    const int blocks = max(1, min((n_rays + kRayWarps - 1) / kRayWarps, sm_count() * 4));
    rays_fwd_kernel<<<blocks, kRayWarps * 32, 0, (cudaStream_t)stream>>>(R);
    MNE_CHECK_CUDA(cudaGetLastError());
    return 0;
}

extern "C" int mne_render_forward_fused(const MneSceneDesc* scene, const MneRenderCfg* rcfg, int32_t n_rays, int32_t n_samples,
                                        const float* rays_o, const float* rays_d, const float* z_vals, const float* target_rgb,
                                        const float* target_d, float* raw, float* weights, float* rgb, float* depth, float* disp,
                                        float* acc_map, float* depth_var, int32_t* inds, double* acc, float* act,
                                        int32_t* ray_counters, const float* loss_grads, void* scratch, mne_stream_t stream) {
    using namespace mne;
    MNE_CHECK_ARG(impl_version() == 2, "mne_render_forward_fused: needs the tensor-pipe kernels (option impl = 2)");
    MNE_CHECK_ARG(scene && rcfg, "mne_render_forward_fused: null scene/rcfg");
    if (int e = check_scene(scene, "mne_render_forward_fused", false)) return e;
    MNE_CHECK_ARG(n_rays >= 0 && n_samples > 0, "mne_render_forward_fused: bad sizes n_rays=%d n_samples=%d", n_rays, n_samples);
    if (n_rays == 0) return 0;
    MNE_CHECK_ARG(rays_o && rays_d && z_vals && raw, "mne_render_forward_fused: rays_o, rays_d, z_vals and raw are required");
    MNE_CHECK_ARG(((uintptr_t)raw & 15) == 0, "mne_render_forward_fused: raw must be 16-byte aligned");
    MNE_CHECK_ARG(((uintptr_t)act & 31) == 0, "mne_render_forward_fused: act must be 32-byte aligned");
    const int64_t n_pts = (int64_t)n_rays * n_samples;
    MNE_CHECK_ARG(n_pts < (int64_t)1 << 31, "mne_render_forward_fused: n_rays * n_samples must be < 2^31");
    MNE_CHECK_ARG(!loss_grads || (scratch && acc && target_rgb && target_d),
                  "mne_render_forward_fused: loss_grads needs scratch, acc, target_rgb and target_d");
    MNE_CHECK_ARG(((uintptr_t)scratch & 15) == 0, "mne_render_forward_fused: scratch must be 16-byte aligned");
    PtsFwdParams P{};
    P.sc = *scene; P.n_pts = n_pts; P.n_samples = n_samples;
    P.rays_o = rays_o; P.rays_d = rays_d; P.z_vals = z_vals; P.raw = raw; P.act = act;
    RayFwdParams& R = P.rf;
    R.rc = *rcfg; R.n_rays = n_rays; R.n_samples = n_samples; R.z_vals = z_vals; R.raw = raw;
    R.target_rgb = target_rgb; R.target_d = target_d; R.weights = weights; R.rgb = rgb; R.depth = depth; R.disp = disp;
    R.acc_map = acc_map; R.depth_var = depth_var; R.inds = inds; R.acc = acc;
    P.ray_done = ray_counters;
    if (loss_grads) {
        P.loss_grads = loss_grads;
        P.draw = reinterpret_cast<float4*>(scratch);
        P.list = reinterpret_cast<int*>(P.draw + n_pts);
        P.n_active = P.list + ((n_pts + 3) & ~(int64_t)3);
        MNE_CHECK_CUDA(cudaMemsetAsync(P.n_active, 0, sizeof(int), (cudaStream_t)stream));
    }
    if (ray_counters) return loss_grads ? dispatch_points_fwd<2>(P, (cudaStream_t)stream) : dispatch_points_fwd<1>(P, (cudaStream_t)stream);
    // without counters: the point kernel, then the whole per-ray stage as one more launch
    if (int e = dispatch_points_fwd<0>(P, (cudaStream_t)stream)) return e;
    RayTailParams T{};
    T.rf = P.rf; T.loss_grads = P.loss_grads; T.draw = P.draw; T.list = P.list; T.n_active = P.n_active;
    const int blocks = max(1, min((n_rays + kRayWarps - 1) / kRayWarps, sm_count() * 8));
    if (loss_grads) rays_tail_kernel<true><<<blocks, kRayWarps * 32, 0, (cudaStream_t)stream>>>(T);
    else rays_tail_kernel<false><<<blocks, kRayWarps * 32, 0, (cudaStream_t)stream>>>(T);
    MNE_CHECK_CUDA(cudaGetLastError());
    return 0;
}

extern "C" int mne_act_floats(const MneSceneDesc* scene) {
    if (!scene) return MNE_ACT_FLOATS_MAX;
    return mne::act_row_floats(scene->two_grid != 0, scene->encoder == 1);
}

extern "C" int mne_query_points(const MneSceneDesc* scene, int64_t n_pts, const float* pts, float* raw, float* geo,
                                float* feat, mne_stream_t stream) {
    using namespace mne;
    if (impl_version() == 1) return query_points_v1(scene, n_pts, pts, raw, geo, feat, stream);
    MNE_CHECK_ARG(scene, "mne_query_points: null scene");
    if (int e = check_scene(scene, "mne_query_points", false)) return e;
    MNE_CHECK_ARG(n_pts >= 0 && n_pts < ((int64_t)1 << 31), "mne_query_points: n_pts out of range");
    if (n_pts == 0) return 0;
    MNE_CHECK_ARG(pts, "mne_query_points: null pts");
    MNE_CHECK_ARG(((uintptr_t)raw & 15) == 0, "mne_query_points: raw must be 16-byte aligned");
    PtsFwdParams P{};
    P.sc = *scene; P.n_pts = n_pts; P.n_samples = 1; P.pts = pts; P.raw = raw; P.geo = geo; P.feat = feat;
    return dispatch_points_fwd<0>(P, (cudaStream_t)stream);
}

// ---------------------------------------------------------------------------------------------
// hash-grid corner indices of explicit [0,1]^3 coordinates (inspection / parity tests: the indices are the one
// integer quantity of the hash encoder, so they can be compared bit for bit with the oracle)
// ---------------------------------------------------------------------------------------------
namespace mne {
__global__ void hash_indices_kernel(MneHashDesc h, int64_t n, const float* __restrict__ x01, int32_t* __restrict__ out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;     // one thread per (point, level)
    if (i >= n * h.n_levels) return;
    const int64_t p = i / h.n_levels;
    const int l = (int)(i % h.n_levels);
    const HashLevel L = hash_level(h, l);
    const float x[3] = {x01[p * 3], x01[p * 3 + 1], x01[p * 3 + 2]};
    const HashCell c = hash_cell(L, x);
#pragma unroll
    for (int k = 0; k < 8; ++k)
        out[i * 8 + k] = (int32_t)hash_index(L, c.g[0] + (k & 1), c.g[1] + ((k >> 1) & 1), c.g[2] + ((k >> 2) & 1));
}
}  // namespace mne

extern "C" int mne_hash_indices(const MneHashDesc* hash, int64_t n_pts, const float* x01, int32_t* indices,
                                mne_stream_t stream) {
    using namespace mne;
    MNE_CHECK_ARG(hash && x01 && indices, "mne_hash_indices: null argument");
    MNE_CHECK_ARG(hash->n_levels >= 1 && hash->n_levels <= 32, "mne_hash_indices: n_levels out of range");
    if (n_pts <= 0) return 0;
    const int64_t total = n_pts * hash->n_levels;
    hash_indices_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(*hash, n_pts, x01, indices);
    MNE_CHECK_CUDA(cudaGetLastError());
    return 0;
}

// ---------------------------------------------------------------------------------------------
// compositing alone: raw2outputs / sdf2weights (reference model/scene_rep.py:183-230) on caller-supplied raw
// ---------------------------------------------------------------------------------------------
extern "C" int mne_composite(const MneRenderCfg* rcfg, int32_t n_rays, int32_t n_samples, const float* z_vals,
                             const float* raw, float* weights, float* rgb, float* depth, float* disp, float* acc_map,
                             float* depth_var, int32_t* inds, mne_stream_t stream) {
    using namespace mne;
    MNE_CHECK_ARG(rcfg && z_vals && raw, "mne_composite: null argument");
    MNE_CHECK_ARG(n_rays >= 0 && n_samples > 0, "mne_composite: bad sizes n_rays=%d n_samples=%d", n_rays, n_samples);
    MNE_CHECK_ARG(((uintptr_t)raw & 15) == 0, "mne_composite: raw must be 16-byte aligned");
    if (n_rays == 0) return 0;
    RayFwdParams R{};
    R.rc = *rcfg; R.n_rays = n_rays; R.n_samples = n_samples; R.z_vals = z_vals; R.raw = raw;
    R.weights = weights; R.rgb = rgb; R.depth = depth; R.disp = disp; R.acc_map = acc_map; R.depth_var = depth_var; R.inds = inds;
    const int blocks = max(1, min((n_rays + kRayWarps - 1) / kRayWarps, sm_count() * 4));
    rays_fwd_kernel<<<blocks, kRayWarps * 32, 0, (cudaStream_t)stream>>>(R);
    MNE_CHECK_CUDA(cudaGetLastError());
    return 0;
}
