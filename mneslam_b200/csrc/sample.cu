// Ray sampling: depth-guided + uniform sample positions merged by rank, stratified perturbation, and the
// data-only loss-mask counts.  Replaces reference model/scene_rep.py:361-381 (z_vals), and the counting
// halves of model/utils.py:117-145, model/scene_rep.py:489-499,570.
//
// One warp per ray.  The two inputs of the reference's torch.sort(torch.cat([uniform, guided])) are each
// already ascending, so the sort is a two-way merge: the final position of an element is its index in its
// own list plus the number of elements of the other list that precede it.  Sorted VALUES are identical to
// torch.sort's whatever the tie order, so z_vals is bit-exact.  All arithmetic uses the explicit
// round-to-nearest intrinsics so that nvcc cannot contract a*b+c into an FMA (torch runs separate kernels).
#include "common.cuh"

namespace mne {

constexpr int kSampleWarps = 8;

__global__ void __launch_bounds__(kSampleWarps * 32)
sample_kernel(MneSampleCfg cfg, MneRenderCfg rcfg, int n_rays, int n_samples, const float* __restrict__ target_d,
              const float* __restrict__ noise, float* __restrict__ z_vals, double* __restrict__ acc) {
    extern __shared__ float smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float* zs = smem + (size_t)warp * n_samples;       // sorted samples of this warp's ray
    const int total_warps = gridDim.x * kSampleWarps;

    unsigned cnt_valid = 0, cnt_co_front = 0, cnt_co_near = 0, cnt_e_front = 0, cnt_e_center = 0, cnt_e_tail = 0;

    for (int ray = blockIdx.x * kSampleWarps + warp; ray < n_rays; ray += total_warps) {
        const float d = target_d ? target_d[ray] : 0.0f;
        const int nu = cfg.n_uniform, nr = cfg.n_range;
        if (nr > 0) {
            const bool guided = d > 0.0f;   // rows with target_d <= 0 take the near..far fallback (:365)
            // uniform list A: rank = i + #{B_j < A_i}
            for (int i = lane; i < nu; i += 32) {
                const float a = cfg.uniform[i];
                int lo = 0, hi = nr;        // first j with B_j >= a
                while (lo < hi) {
                    const int mid = (lo + hi) >> 1;
                    const float b = guided ? __fadd_rn(cfg.range_offsets[mid], d) : cfg.range_fallback[mid];
                    if (b < a) lo = mid + 1; else hi = mid;
                }
                zs[i + lo] = a;
            }
            // guided list B: rank = j + #{A_i <= B_j}
            for (int j = lane; j < nr; j += 32) {
                const float b = guided ? __fadd_rn(cfg.range_offsets[j], d) : cfg.range_fallback[j];
                int lo = 0, hi = nu;        // first i with A_i > b
                while (lo < hi) {
                    const int mid = (lo + hi) >> 1;
                    if (cfg.uniform[mid] <= b) lo = mid + 1; else hi = mid;
                }
                zs[j + lo] = b;
            }
        } else {
            for (int i = lane; i < nu; i += 32) zs[i] = cfg.uniform[i];
        }
        __syncwarp();

        const bool valid = target_d && depth_valid(d, rcfg);
        if (lane == 0 && valid) cnt_valid++;
        for (int s = lane; s < n_samples; s += 32) {
            float z = zs[s];
            if (cfg.perturb && noise) {
                const float up = (s + 1 < n_samples) ? __fmul_rn(0.5f, __fadd_rn(zs[s + 1], z)) : z;
                const float dn = (s > 0) ? __fmul_rn(0.5f, __fadd_rn(z, zs[s - 1])) : z;
                z = __fadd_rn(dn, __fmul_rn(__fsub_rn(up, dn), noise[(size_t)ray * n_samples + s]));
            }
            z_vals[(size_t)ray * n_samples + s] = z;
            if (acc && target_d) {
                const unsigned m = sample_masks(z, d, rcfg);
                cnt_co_front += (m & M_CO_FRONT) ? 1 : 0;
                cnt_co_near += (m & M_CO_NEAR) ? 1 : 0;
                cnt_e_front += (m & M_E_FRONT) ? 1 : 0;
                cnt_e_center += (m & M_E_CENTER) ? 1 : 0;
                cnt_e_tail += (m & M_E_TAIL) ? 1 : 0;
            }
        }
        __syncwarp();
    }

    if (acc) {
        // block-level reduction, then one atomic per counter per block
        __shared__ unsigned red[6];
        if (threadIdx.x < 6) red[threadIdx.x] = 0;
        __syncthreads();
        unsigned v[6] = {cnt_valid, cnt_co_front, cnt_co_near, cnt_e_front, cnt_e_center, cnt_e_tail};
#pragma unroll
        for (int k = 0; k < 6; ++k) {
            unsigned s = v[k];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
            if (lane == 0 && s) atomicAdd(&red[k], s);
        }
        __syncthreads();
        if (threadIdx.x < 6 && red[threadIdx.x]) atomicAdd(&acc[threadIdx.x], (double)red[threadIdx.x]);
    }
}

}  // namespace mne

extern "C" int mne_sample_rays(const MneSampleCfg* cfg, const MneRenderCfg* rcfg, int32_t n_rays,
                               const float* target_d, const float* noise, float* z_vals, double* acc,
                               mne_stream_t stream) {
    using namespace mne;
    MNE_CHECK_ARG(cfg && rcfg && z_vals, "mne_sample_rays: null cfg/rcfg/z_vals");
    MNE_CHECK_ARG(n_rays >= 0, "mne_sample_rays: n_rays < 0");
    if (n_rays == 0) return 0;
    MNE_CHECK_ARG(cfg->n_uniform >= 0 && cfg->n_range >= 0 && cfg->n_uniform + cfg->n_range > 0,
                  "mne_sample_rays: no samples requested");
    MNE_CHECK_ARG(cfg->n_uniform == 0 || cfg->uniform, "mne_sample_rays: uniform table missing");
    MNE_CHECK_ARG(cfg->n_range == 0 || (cfg->range_offsets && cfg->range_fallback && target_d),
                  "mne_sample_rays: depth-guided sampling needs target_d and both range tables");
    MNE_CHECK_ARG(!(cfg->perturb && !noise), "mne_sample_rays: perturb set but noise is null");
    const int n_samples = cfg->n_uniform + cfg->n_range;
    const size_t smem = (size_t)kSampleWarps * n_samples * sizeof(float);
    MNE_CHECK_ARG(smem <= 200 * 1024, "mne_sample_rays: %d samples per ray exceed the %d supported", n_samples,
                  (int)(200 * 1024 / (kSampleWarps * sizeof(float))));
    if (smem > 48 * 1024)
        MNE_CHECK_CUDA(cudaFuncSetAttribute(sample_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int blocks = (n_rays + kSampleWarps - 1) / kSampleWarps;
    if (blocks > 148 * 8) blocks = 148 * 8;
    sample_kernel<<<blocks, kSampleWarps * 32, smem, (cudaStream_t)stream>>>(*cfg, *rcfg, n_rays, n_samples, target_d,
                                                                              noise, z_vals, acc);
    MNE_CHECK_CUDA(cudaGetLastError());
    return 0;
}

// ---------------------------------------------------------------------------------------------------------
// Importance sampling: inverse-CDF samples from the coarse pass's weights, merged into the sample list.
// Replaces sample_pdf (reference model/utils.py:65-104) and the re-sort of render_rays (model/scene_rep.py:393-397).
// One warp per ray.  The cdf is accumulated sequentially in fp32 like torch.cumsum, the bin search is
// searchsorted(right=True); `inds` (the bin indices) is an output so parity tests can compare it bit for bit.
// ---------------------------------------------------------------------------------------------------------
namespace mne {

__global__ void __launch_bounds__(kSampleWarps * 32)
importance_kernel(MneRenderCfg rcfg, int n_rays, int s0, int n_imp, const float* __restrict__ z0,
                  const float* __restrict__ weights0, const float* __restrict__ u_in, int u_per_ray,
                  const float* __restrict__ target_d, float* __restrict__ z1, int32_t* __restrict__ inds_out,
                  float* __restrict__ z_std, double* __restrict__ acc) {
    extern __shared__ float smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n_bins = s0 - 1;                     // cdf entries = mids
    float* cdf = smem + (size_t)warp * (n_bins + n_imp);
    float* smp = cdf + n_bins;
    const int s1 = s0 + n_imp;
    unsigned cnt_valid = 0, cnt_co_front = 0, cnt_co_near = 0, cnt_e_front = 0, cnt_e_center = 0, cnt_e_tail = 0;

    for (int ray = blockIdx.x * kSampleWarps + warp; ray < n_rays; ray += gridDim.x * kSampleWarps) {
        const float* z = z0 + (size_t)ray * s0;
        const float* w = weights0 + (size_t)ray * s0;
        // pdf = (w[1:-1] + 1e-5) / sum ; cdf = [0, cumsum(pdf)]
        float part = 0.f;
        for (int i = lane; i < s0 - 2; i += 32) part += __fadd_rn(w[i + 1], 1e-5f);
        const float total = warp_sum(part);
        if (lane == 0) {
            float c = 0.f;
            cdf[0] = 0.f;
            for (int i = 0; i < s0 - 2; ++i) {
                c = __fadd_rn(c, __fdiv_rn(__fadd_rn(w[i + 1], 1e-5f), total));
                cdf[i + 1] = c;
            }
        }
        __syncwarp();
        for (int j = lane; j < n_imp; j += 32) {
            const float u = u_per_ray ? u_in[(size_t)ray * n_imp + j] : u_in[j];
            int lo = 0, hi = n_bins;                  // first index with cdf[idx] > u
            while (lo < hi) {
                const int mid = (lo + hi) >> 1;
                if (cdf[mid] <= u) lo = mid + 1; else hi = mid;
            }
            const int below = max(lo - 1, 0), above = min(n_bins - 1, lo);
            const float c0 = cdf[below], c1 = cdf[above];
            const float b0 = __fmul_rn(0.5f, __fadd_rn(z[below + 1], z[below]));
            const float b1 = __fmul_rn(0.5f, __fadd_rn(z[above + 1], z[above]));
            float den = __fsub_rn(c1, c0);
            if (den < 1e-5f) den = 1.0f;
            const float t = __fdiv_rn(__fsub_rn(u, c0), den);
            smp[j] = __fadd_rn(b0, __fmul_rn(t, __fsub_rn(b1, b0)));
            if (inds_out) inds_out[(size_t)ray * n_imp + j] = lo;
        }
        __syncwarp();
        // merge by rank: coarse samples first on ties
        float* out = z1 + (size_t)ray * s1;
        for (int i = lane; i < s0; i += 32) {
            const float a = z[i];
            int less = 0;
            for (int k = 0; k < n_imp; ++k) less += (smp[k] < a) ? 1 : 0;
            out[i + less] = a;
        }
        float mean_part = 0.f;
        for (int j = lane; j < n_imp; j += 32) {
            const float s = smp[j];
            int lo = 0, hi = s0;                      // #{z0 <= s}
            while (lo < hi) {
                const int mid = (lo + hi) >> 1;
                if (z[mid] <= s) lo = mid + 1; else hi = mid;
            }
            int r = lo;
            for (int k = 0; k < n_imp; ++k) r += (smp[k] < s || (smp[k] == s && k < j)) ? 1 : 0;
            out[r] = s;
            mean_part += s;
        }
        if (z_std) {
            const float mean = warp_sum(mean_part) / (float)n_imp;
            float var = 0.f;
            for (int j = lane; j < n_imp; j += 32) var += (smp[j] - mean) * (smp[j] - mean);
            var = warp_sum(var);
            if (lane == 0) z_std[ray] = sqrtf(var / (float)n_imp);
        }
        __syncwarp();
        if (acc && target_d) {
            const float d = target_d[ray];
            if (lane == 0 && depth_valid(d, rcfg)) cnt_valid++;
            for (int s = lane; s < s1; s += 32) {
                const unsigned m = sample_masks(out[s], d, rcfg);
                cnt_co_front += (m & M_CO_FRONT) ? 1 : 0;
                cnt_co_near += (m & M_CO_NEAR) ? 1 : 0;
                cnt_e_front += (m & M_E_FRONT) ? 1 : 0;
                cnt_e_center += (m & M_E_CENTER) ? 1 : 0;
                cnt_e_tail += (m & M_E_TAIL) ? 1 : 0;
            }
        }
        __syncwarp();
    }
    if (acc) {
        unsigned v[6] = {cnt_valid, cnt_co_front, cnt_co_near, cnt_e_front, cnt_e_center, cnt_e_tail};
#pragma unroll
        for (int k = 0; k < 6; ++k) {
            unsigned s = v[k];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
            if (lane == 0 && s) atomicAdd(&acc[k], (double)s);
        }
    }
}

}  // namespace mne

extern "C" int mne_sample_importance(const MneRenderCfg* rcfg, int32_t n_rays, int32_t n_samples, int32_t n_importance,
                                     const float* z_vals, const float* weights, const float* u, int32_t u_per_ray,
                                     const float* target_d, float* z_out, int32_t* inds, float* z_std, double* acc,
                                     mne_stream_t stream) {
    using namespace mne;
    MNE_CHECK_ARG(rcfg && z_vals && weights && u && z_out, "mne_sample_importance: null argument");
    MNE_CHECK_ARG(n_rays >= 0 && n_samples >= 3 && n_importance >= 1, "mne_sample_importance: bad sizes (%d rays, %d samples, %d importance)",
                  n_rays, n_samples, n_importance);
    if (n_rays == 0) return 0;
    const size_t smem = (size_t)kSampleWarps * (n_samples - 1 + n_importance) * sizeof(float);
    MNE_CHECK_ARG(smem <= 200 * 1024, "mne_sample_importance: %d + %d samples per ray exceed the supported size", n_samples, n_importance);
    if (smem > 48 * 1024)
        MNE_CHECK_CUDA(cudaFuncSetAttribute(importance_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int blocks = (n_rays + kSampleWarps - 1) / kSampleWarps;
    if (blocks > 148 * 8) blocks = 148 * 8;
    importance_kernel<<<blocks, kSampleWarps * 32, smem, (cudaStream_t)stream>>>(*rcfg, n_rays, n_samples, n_importance, z_vals,
                                                                                  weights, u, u_per_ray, target_d, z_out, inds,
                                                                                  z_std, acc);
    MNE_CHECK_CUDA(cudaGetLastError());
    return 0;
}
