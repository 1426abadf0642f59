// Fused forward: points along rays -> tri-plane gather -> OneBlob -> SDF / colour MLPs -> raw, then
// sdf2weights compositing and the loss sums, one kernel.  Also serves the point-query entry point.
// Replaces reference model/scene_rep.py:273-317 (query_color_sdf / run_network), :183-230 (sdf2weights,
// raw2outputs), :384-386 (render_rays body) and the summing halves of :569-590 (losses).
#include "render_common.cuh"

namespace mne {

template <bool TWO> __host__ __device__ constexpr int fwd_warps() { return TWO ? 6 : 8; }

struct FwdParams {
    MneSceneDesc sc;
    MneRenderCfg rc;
    int n_rays, n_samples, rays_per_group;
    const float* rays_o;
    const float* rays_d;
    const float* z_vals;
    const float* pts;         // QUERY: explicit points [M,3]
    const float* target_rgb;
    const float* target_d;
    float* raw;
    float* weights;
    float* rgb;
    float* depth;
    float* disp;
    float* acc_map;
    float* depth_var;
    int32_t* inds;
    float* geo;               // QUERY: [M,15]
    float* feat;              // QUERY: [M,64]
    double* acc;
};

template <bool TWO>
__host__ __device__ constexpr int fwd_warp_floats() {
    return 32 * kInStride + (TWO ? 32 * kCfStride : 0) + 32 * Net<TWO>::N_TAPS * 4;
}
template <bool TWO>
constexpr size_t fwd_smem_bytes() {
    return sizeof(float) * (size_t)(Net<TWO>::W_FLOATS + fwd_warps<TWO>() * fwd_warp_floats<TWO>());
}

template <bool TWO, bool QUERY>
__global__ void __launch_bounds__(fwd_warps<TWO>() * 32, 1) render_fwd_kernel(const __grid_constant__ FwdParams P) {
    using N = Net<TWO>;
    constexpr int kFwdWarps = fwd_warps<TWO>();
    extern __shared__ __align__(16) float smem[];
    float* WT = smem;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float* IN = smem + N::W_FLOATS + warp * fwd_warp_floats<TWO>();
    float* CF = IN + 32 * kInStride;
    int4* TAB = reinterpret_cast<int4*>(CF + (TWO ? 32 * kCfStride : 0));

    load_weights_transposed<TWO>(P.sc, WT, threadIdx.x, blockDim.x);
    __syncthreads();

    const int S = P.n_samples, rpg = P.rays_per_group;
    const int n_groups = (P.n_rays + rpg - 1) / rpg;
    const int total_warps = gridDim.x * kFwdWarps;
    const bool with_loss = !QUERY && P.acc && P.target_rgb && P.target_d;

    double l_rgb = 0, l_depth = 0, l_co_fs = 0, l_co_sdf = 0, l_e_fs = 0, l_e_center = 0, l_e_tail = 0;

    // block-minor order: consecutive groups land on different SMs, so a grid with spare warps stays balanced
    for (int g = warp * gridDim.x + blockIdx.x; g < n_groups; g += total_warps) {
        const int ray0 = g * rpg;
        const int nr = min(rpg, P.n_rays - ray0);
        const int npts = nr * S;
        const size_t base = (size_t)ray0 * S;

        for (int c0 = 0; c0 < npts; c0 += 32) {
            const int q = c0 + lane;
            const bool active = q < npts;
            const int np = min(32, npts - c0);
            // ---- lane = point: position, taps, OneBlob
            float px = 0.f, py = 0.f, pz = 0.f;
            if (active) {
                if (QUERY) {
                    px = P.pts[(base + q) * 3 + 0]; py = P.pts[(base + q) * 3 + 1]; pz = P.pts[(base + q) * 3 + 2];
                } else {
                    const int ray = ray0 + q / S;
                    const float z = P.z_vals[base + q];
                    px = __fadd_rn(P.rays_o[ray * 3 + 0], __fmul_rn(P.rays_d[ray * 3 + 0], z));
                    py = __fadd_rn(P.rays_o[ray * 3 + 1], __fmul_rn(P.rays_d[ray * 3 + 1], z));
                    pz = __fadd_rn(P.rays_o[ray * 3 + 2], __fmul_rn(P.rays_d[ray * 3 + 2], z));
                }
            }
            write_taps<TWO>(P.sc, TAB + lane * N::N_TAPS, px, py, pz);
            float* in_row = IN + lane * kInStride;
            oneblob16(blob_input(px, P.sc.bbox_lo[0], P.sc.bbox_ext[0]), in_row + N::F);
            oneblob16(blob_input(py, P.sc.bbox_lo[1], P.sc.bbox_ext[1]), in_row + N::F + 16);
            oneblob16(blob_input(pz, P.sc.bbox_lo[2], P.sc.bbox_ext[2]), in_row + N::F + 32);
            __syncwarp();
            // ---- lane = channel: gather
            gather_chunk<TWO>(P.sc, TAB, IN, CF, np, lane);
            __syncwarp();
            // ---- lane = point: decoder
            float h1[32], out16[16], hc[32], rgb[3];
            decoder_forward<TWO>(WT, in_row, CF + lane * kCfStride, h1, out16, hc, rgb);
            if (active) {
                if (P.raw) reinterpret_cast<float4*>(P.raw)[base + q] = make_float4(rgb[0], rgb[1], rgb[2], out16[15]);
                if (QUERY && P.geo) {
#pragma unroll
                    for (int m = 0; m < kGeo; ++m) P.geo[(base + q) * kGeo + m] = out16[m];
                }
            }
            if (QUERY && P.feat) {
                __syncwarp();
                for (int p = 0; p < np; ++p) {
                    P.feat[(base + c0 + p) * N::F + lane] = IN[p * kInStride + lane];
                    P.feat[(base + c0 + p) * N::F + 32 + lane] = IN[p * kInStride + 32 + lane];
                }
            }
            __syncwarp();
        }

        if (!QUERY) {
            for (int r = 0; r < nr; ++r) {
                const int ray = ray0 + r;
                const size_t rb = (size_t)ray * S;
                const RayComposite c = composite_ray(P.raw, P.z_vals, rb, S, P.rc, lane);
                const float d = P.target_d ? P.target_d[ray] : 0.0f;
                // second pass: weights, variance and the per-sample sdf loss sums
                float var = 0.f;
                for (int s = lane; s < S; s += 32) {
                    const float z = __ldcg(P.z_vals + rb + s);
                    const float sdf = __ldcg(P.raw + (rb + s) * 4 + 3);
                    float w = 0.f;
                    if (z < c.thresh) w = sdf_bell(sdf, P.rc.trunc) * c.inv_norm;
                    if (P.weights) P.weights[rb + s] = w;
                    var += w * (z - c.depth) * (z - c.depth);
                    if (with_loss && !P.rc.first_pass) {
                        const unsigned m = sample_masks(z, d, P.rc);
                        const float fs = (sdf - 1.0f) * (sdf - 1.0f);
                        const float rc_co = __fadd_rn(z, __fmul_rn(sdf, P.rc.sc_trunc)) - d;
                        const float rc_e = __fadd_rn(z, __fmul_rn(sdf, P.rc.model_trunc)) - d;
                        if (m & M_CO_FRONT) l_co_fs += fs;
                        if (m & M_CO_NEAR) l_co_sdf += rc_co * rc_co;
                        if (m & M_E_FRONT) l_e_fs += fs;
                        if (m & M_E_CENTER) l_e_center += rc_e * rc_e;
                        if (m & M_E_TAIL) l_e_tail += rc_e * rc_e;
                    }
                }
                var = warp_sum(var);
                if (lane == 0) {
                    float rr = c.rgb[0], gg = c.rgb[1], bb = c.rgb[2];
                    if (P.rc.white_bkgd) { rr += 1.f - c.acc; gg += 1.f - c.acc; bb += 1.f - c.acc; }
                    if (P.rgb) { P.rgb[ray * 3 + 0] = rr; P.rgb[ray * 3 + 1] = gg; P.rgb[ray * 3 + 2] = bb; }
                    if (P.depth) P.depth[ray] = c.depth;
                    if (P.disp) P.disp[ray] = 1.0f / fmaxf(1e-10f, c.depth / c.acc);
                    if (P.acc_map) P.acc_map[ray] = c.acc;
                    if (P.depth_var) P.depth_var[ray] = var;
                    if (P.inds) P.inds[ray] = c.ind;
                    if (with_loss) {
                        const float e0 = rr - P.target_rgb[ray * 3 + 0], e1 = gg - P.target_rgb[ray * 3 + 1],
                                    e2 = bb - P.target_rgb[ray * 3 + 2];
                        l_rgb += (double)(e0 * e0) + (double)(e1 * e1) + (double)(e2 * e2);
                        if (depth_valid(d, P.rc)) l_depth += (double)((c.depth - d) * (c.depth - d));
                    }
                }
            }
        }
    }

    if (with_loss) {
        l_rgb = warp_sum(l_rgb); l_depth = warp_sum(l_depth);
        l_co_fs = warp_sum(l_co_fs); l_co_sdf = warp_sum(l_co_sdf);
        l_e_fs = warp_sum(l_e_fs); l_e_center = warp_sum(l_e_center); l_e_tail = warp_sum(l_e_tail);
        if (lane == 0) {
            if (P.rc.first_pass) {
                atomicAdd(P.acc + A_S_RGB0, l_rgb);
                atomicAdd(P.acc + A_S_DEPTH0, l_depth);
            } else {
                atomicAdd(P.acc + A_S_RGB, l_rgb);
                atomicAdd(P.acc + A_S_DEPTH, l_depth);
                atomicAdd(P.acc + A_S_CO_FS, l_co_fs);
                atomicAdd(P.acc + A_S_CO_SDF, l_co_sdf);
                atomicAdd(P.acc + A_S_E_FS, l_e_fs);
                atomicAdd(P.acc + A_S_E_CENTER, l_e_center);
                atomicAdd(P.acc + A_S_E_TAIL, l_e_tail);
            }
        }
    }
}

// losses from the accumulators (single thread; the arithmetic is a dozen flops)
__global__ void loss_finalize_kernel(MneRenderCfg rc, int n_samples, const double* __restrict__ acc, float* __restrict__ out) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    const double N = rc.norm_rays, NS = rc.norm_rays * (double)n_samples;
    const float rgb_loss = (float)(acc[A_S_RGB] / (3.0 * N));
    const float psnr = -10.0f * logf(rgb_loss) / logf(10.0f);        // before the rgb0 term (scene_rep.py:576)
    float rgb_total = rgb_loss, depth_loss = (float)(acc[A_S_DEPTH] / acc[A_N_VALID]);
    if (acc[A_S_RGB0] != 0.0 || acc[A_S_DEPTH0] != 0.0) {            // importance sampling: coarse-pass terms
        rgb_total += (float)(acc[A_S_RGB0] / (3.0 * N));
        depth_loss += (float)(acc[A_S_DEPTH0] / acc[A_N_VALID]);
    }
    const float n_fs = (float)acc[A_N_CO_FRONT], n_sdf = (float)acc[A_N_CO_NEAR];
    const float fs_w = 1.0f - n_fs / (n_fs + n_sdf), sdf_w = 1.0f - n_sdf / (n_fs + n_sdf);
    out[L_RGB] = rgb_total;
    out[L_DEPTH] = depth_loss;
    out[L_CO_SDF] = (float)(acc[A_S_CO_SDF] / NS) * sdf_w;
    out[L_CO_FS] = (float)(acc[A_S_CO_FS] / NS) * fs_w;
    out[L_E_FS] = (float)(acc[A_S_E_FS] / acc[A_N_E_FRONT]);
    out[L_E_CENTER] = (float)(acc[A_S_E_CENTER] / acc[A_N_E_CENTER]);
    out[L_E_TAIL] = (float)(acc[A_S_E_TAIL] / acc[A_N_E_TAIL]);
    out[L_PSNR] = psnr;
}

int check_scene(const MneSceneDesc* sc, const char* who, bool need_grads);   // api.cu
int choose_rays_per_group(int n_rays, int n_samples, int total_warps);       // api.cu
int sm_count();                                                              // api.cu

template <bool TWO, bool QUERY>
static int launch_fwd(const FwdParams& P, cudaStream_t stream) {
    const size_t smem = fwd_smem_bytes<TWO>();
    static std::atomic<bool> configured[64] = {};      // per device; set outside any stream capture by the first (warm-up) call
    int dev = 0;
    MNE_CHECK_CUDA(cudaGetDevice(&dev));
    if (dev >= 64 || !configured[dev]) {
        MNE_CHECK_CUDA(cudaFuncSetAttribute(render_fwd_kernel<TWO, QUERY>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        if (dev < 64) configured[dev] = true;
    }
    const int n_groups = (P.n_rays + P.rays_per_group - 1) / P.rays_per_group;
    const int blocks = max(1, min(n_groups, sm_count()));
    render_fwd_kernel<TWO, QUERY><<<blocks, fwd_warps<TWO>() * 32, smem, stream>>>(P);
    MNE_CHECK_CUDA(cudaGetLastError());
    return 0;
}

}  // namespace mne

namespace mne {
int render_forward_v1(const MneSceneDesc* scene, const MneRenderCfg* rcfg, int32_t n_rays, int32_t n_samples,
                      const float* rays_o, const float* rays_d, const float* z_vals, const float* target_rgb,
                      const float* target_d, float* raw, float* weights, float* rgb, float* depth, float* disp,
                      float* acc_map, float* depth_var, int32_t* inds, double* acc, mne_stream_t stream) {
    MNE_CHECK_ARG(scene && rcfg, "mne_render_forward: null scene/rcfg");
    if (int e = check_scene(scene, "mne_render_forward", false)) return e;
    MNE_CHECK_ARG(n_rays >= 0 && n_samples > 0, "mne_render_forward: bad sizes n_rays=%d n_samples=%d", n_rays, n_samples);
    if (n_rays == 0) return 0;
    MNE_CHECK_ARG(rays_o && rays_d && z_vals && raw, "mne_render_forward: rays_o, rays_d, z_vals and raw are required");
    MNE_CHECK_ARG((int64_t)n_rays * n_samples < (int64_t)1 << 31, "mne_render_forward: n_rays * n_samples must be < 2^31");
    FwdParams P{};
    P.sc = *scene; P.rc = *rcfg; P.n_rays = n_rays; P.n_samples = n_samples;
    P.rays_per_group = choose_rays_per_group(n_rays, n_samples, sm_count() * (scene->two_grid ? fwd_warps<true>() : fwd_warps<false>()));
    P.rays_o = rays_o; P.rays_d = rays_d; P.z_vals = z_vals; P.target_rgb = target_rgb; P.target_d = target_d;
    P.raw = raw; P.weights = weights; P.rgb = rgb; P.depth = depth; P.disp = disp; P.acc_map = acc_map;
    P.depth_var = depth_var; P.inds = inds; P.acc = acc;
    return scene->two_grid ? launch_fwd<true, false>(P, (cudaStream_t)stream) : launch_fwd<false, false>(P, (cudaStream_t)stream);
}

int query_points_v1(const MneSceneDesc* scene, int64_t n_pts, const float* pts, float* raw, float* geo, float* feat,
                    mne_stream_t stream) {
    MNE_CHECK_ARG(scene, "mne_query_points: null scene");
    if (int e = check_scene(scene, "mne_query_points", false)) return e;
    MNE_CHECK_ARG(n_pts >= 0 && n_pts < ((int64_t)1 << 31), "mne_query_points: n_pts out of range");
    if (n_pts == 0) return 0;
    MNE_CHECK_ARG(pts, "mne_query_points: null pts");
    FwdParams P{};
    P.sc = *scene; P.n_rays = (int)n_pts; P.n_samples = 1; P.rays_per_group = 128;
    P.pts = pts; P.raw = raw; P.geo = geo; P.feat = feat;
    return scene->two_grid ? launch_fwd<true, true>(P, (cudaStream_t)stream) : launch_fwd<false, true>(P, (cudaStream_t)stream);
}

}  // namespace mne

extern "C" int mne_loss_finalize(const MneRenderCfg* rcfg, int32_t n_samples, const double* acc, float* losses,
                                 mne_stream_t stream) {
    using namespace mne;
    MNE_CHECK_ARG(rcfg && acc && losses, "mne_loss_finalize: null argument");
    loss_finalize_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(*rcfg, n_samples, acc, losses);
    MNE_CHECK_CUDA(cudaGetLastError());
    return 0;
}
