// Ray assembly from the keyframe ray database: replaces the gather of KeyFrameDatabase.sample_global_rays
// (reference model/keyframe.py:91-103: rays[idx] from the flattened [n_kf * per_kf, 7] store, keyframe id = idx // per_kf)
// and the assembly of mapping_optimize (mp_slam/mapper.py:135-153):
//     rays_d = sum_k dirs_cam[k] * R[kf][:, k],   rays_o = t[kf],   target_s = rays[:, 3:6],   target_d = rays[:, 6:7]
// One thread per ray, one launch instead of eight small torch kernels.  The three products are summed left to right
// with explicit round-to-nearest operations (no FMA contraction), as a torch.sum over the last axis of three does.
#include "common.cuh"

namespace mne {

__global__ void __launch_bounds__(256)
assemble_rays_kernel(const float* __restrict__ store, const int64_t* __restrict__ idx, int64_t per_kf,
                     const int64_t* __restrict__ kf_ids, const float* __restrict__ poses, int n_kf, int n,
                     float* __restrict__ rays_o, float* __restrict__ rays_d, float* __restrict__ rgb, float* __restrict__ depth) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int64_t row = idx ? idx[i] : (int64_t)i;
    int64_t kf = kf_ids ? kf_ids[i] : row / per_kf;
    if (kf < 0) kf += n_kf;                        // -1 = the last pose (the current frame, mapper.py:147-150)
    const float* r = store + row * 7;
    const float* P = poses + kf * 16;              // row-major 4x4 camera-to-world
    const float d0 = r[0], d1 = r[1], d2 = r[2];
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const float s = __fadd_rn(__fadd_rn(__fmul_rn(d0, P[a * 4 + 0]), __fmul_rn(d1, P[a * 4 + 1])), __fmul_rn(d2, P[a * 4 + 2]));
        rays_d[i * 3 + a] = s;
        rays_o[i * 3 + a] = P[a * 4 + 3];
        rgb[i * 3 + a] = r[3 + a];
    }
    depth[i] = r[6];
}

}  // namespace mne

extern "C" int mne_assemble_rays(const float* rays7, const int64_t* idx, int64_t per_kf, const int64_t* kf_ids,
                                 const float* poses, int32_t n_kf, int32_t n_rays, float* rays_o, float* rays_d,
                                 float* target_rgb, float* target_d, mne_stream_t stream) {
    using namespace mne;
    MNE_CHECK_ARG(n_rays >= 0 && n_kf >= 1, "mne_assemble_rays: bad sizes n_rays=%d n_kf=%d", n_rays, n_kf);
    if (n_rays == 0) return 0;
    MNE_CHECK_ARG(rays7 && poses && rays_o && rays_d && target_rgb && target_d, "mne_assemble_rays: null argument");
    MNE_CHECK_ARG(kf_ids || per_kf > 0, "mne_assemble_rays: pass kf_ids, or per_kf > 0 so that the keyframe is idx / per_kf");
    assemble_rays_kernel<<<(n_rays + 255) / 256, 256, 0, (cudaStream_t)stream>>>(rays7, idx, per_kf, kf_ids, poses, n_kf, n_rays,
                                                                                  rays_o, rays_d, target_rgb, target_d);
    MNE_CHECK_CUDA(cudaGetLastError());
    return 0;
}
