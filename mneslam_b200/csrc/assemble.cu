// Ray assembly from the keyframe ray database: replaces the gather of KeyFrameDatabase.sample_global_rays
// (reference model/keyframe.py:91-103: rays[idx] from the flattened [n_kf * per_kf, 7] store, keyframe id = idx // per_kf)
// and the assembly of mapping_optimize (mp_slam/mapper.py:135-153):
//     rays_d = sum_k dirs_cam[k] * R[kf][:, k],   rays_o = t[kf],   target_s = rays[:, 3:6],   target_d = rays[:, 6:7]
// One thread per ray, one launch instead of eight small torch kernels.  The three products are summed left to right
// with explicit round-to-nearest operations (no FMA contraction), as a torch.sum over the last axis of three does.
#include "common.cuh"

namespace mne {

// One ray: camera-frame direction -> world, origin = the pose's translation, targets copied through.
__device__ __forceinline__ void assemble_one(const float* __restrict__ r, const float* __restrict__ P, int i, float* __restrict__ rays_o,
                                             float* __restrict__ rays_d, float* __restrict__ rgb, float* __restrict__ depth) {
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

__global__ void __launch_bounds__(256)
assemble_rays_kernel(const float* __restrict__ store, const int64_t* __restrict__ idx, int64_t per_kf,
                     const int64_t* __restrict__ kf_ids, const float* __restrict__ poses, int n_kf, int n,
                     float* __restrict__ rays_o, float* __restrict__ rays_d, float* __restrict__ rgb, float* __restrict__ depth) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int64_t row = idx ? idx[i] : (int64_t)i;
    int64_t kf = kf_ids ? kf_ids[i] : row / per_kf;
    if (kf < 0) kf += n_kf;                        // -1 = the last pose (the current frame, mapper.py:147-150)
    assemble_one(store + row * 7, poses + kf * 16 /* row-major 4x4 camera-to-world */, i, rays_o, rays_d, rgb, depth);
}

// ---- drawing the batch on the device, WITHOUT replacement (the reference draws with random.sample: keyframe.py:98 for the
// keyframe rays, mapper.py:137-139 for the pixels of the current frame).  Element i of the draw is pi(i) for a keyed
// pseudo-random BIJECTION pi of [0, n): a 4-round balanced Feistel network over the next even number of bits, walked
// along its cycle until it lands inside [0, n) (a bijection restricted to a subset by cycle walking stays a bijection).
// Distinct i give distinct rows, no sort / rejection / n-sized permutation, and a fresh key gives a fresh draw.
__device__ __forceinline__ uint32_t mix32(uint32_t x) {
    x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16;
    return x;
}

struct DrawKey { uint32_t k[4]; int half_bits; };

__device__ __forceinline__ DrawKey make_draw_key(uint64_t key, uint32_t tag, uint64_t n) {
    DrawKey d;
    int bits = 2;
    while (bits < 62 && (1ull << bits) < n) bits += 2;
    d.half_bits = bits / 2;
    const uint32_t lo = (uint32_t)key, hi = (uint32_t)(key >> 32);
#pragma unroll
    for (int r = 0; r < 4; ++r) d.k[r] = mix32(lo ^ mix32(hi + 0x9e3779b9U * (uint32_t)(r + 1) + tag));
    return d;
}

__device__ __forceinline__ uint64_t permute_index(uint64_t i, uint64_t n, const DrawKey& d) {
    const uint64_t mask = (1ull << d.half_bits) - 1;
    uint64_t x = i;
    do {
        uint64_t L = x >> d.half_bits, R = x & mask;
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            const uint64_t f = (uint64_t)mix32((uint32_t)R + d.k[r]) & mask;
            const uint64_t nr = L ^ f;
            L = R; R = nr;
        }
        x = (L << d.half_bits) | R;
    } while (x >= n);
    return x;
}

__global__ void __launch_bounds__(256)
draw_rays_kernel(const int64_t* __restrict__ key, const float* __restrict__ store, int64_t n_store, int64_t per_kf, int n_global,
                 const float* __restrict__ current, int64_t n_current, int n_cur, const float* __restrict__ poses, int n_kf,
                 float* __restrict__ rays_o, float* __restrict__ rays_d, float* __restrict__ rgb, float* __restrict__ depth,
                 int64_t* __restrict__ rows_out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_global + n_cur) return;
    const uint64_t k = (uint64_t)key[0];
    if (i < n_global) {
        const int64_t row = (int64_t)permute_index((uint64_t)i, (uint64_t)n_store, make_draw_key(k, 0u, (uint64_t)n_store));
        if (rows_out) rows_out[i] = row;
        assemble_one(store + row * 7, poses + (row / per_kf) * 16, i, rays_o, rays_d, rgb, depth);
    } else {
        const int64_t row = (int64_t)permute_index((uint64_t)(i - n_global), (uint64_t)n_current, make_draw_key(k, 0x51ed270bU, (uint64_t)n_current));
        if (rows_out) rows_out[i] = row;
        assemble_one(current + row * 7, poses + (int64_t)(n_kf - 1) * 16, i, rays_o, rays_d, rgb, depth);   // pose id -1
    }
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

extern "C" int mne_draw_rays(const int64_t* key, const float* rays7, int64_t n_stored, int64_t per_kf, int32_t n_global,
                             const float* current7, int64_t n_current, int32_t n_cur, const float* poses, int32_t n_kf,
                             float* rays_o, float* rays_d, float* target_rgb, float* target_d, int64_t* rows_out, mne_stream_t stream) {
    using namespace mne;
    MNE_CHECK_ARG(n_global >= 0 && n_cur >= 0 && n_kf >= 1 && per_kf > 0, "mne_draw_rays: bad sizes n_global=%d n_cur=%d n_kf=%d", n_global, n_cur, n_kf);
    if (n_global + n_cur == 0) return 0;
    MNE_CHECK_ARG(key && poses && rays_o && rays_d && target_rgb && target_d, "mne_draw_rays: null argument");
    // random.sample raises ValueError when the sample is larger than the population
    MNE_CHECK_ARG(n_global == 0 || (rays7 && (int64_t)n_global <= n_stored), "mne_draw_rays: %d rays wanted from a store of %lld", n_global, (long long)n_stored);
    MNE_CHECK_ARG(n_cur == 0 || (current7 && (int64_t)n_cur <= n_current), "mne_draw_rays: %d rays wanted from a current frame of %lld", n_cur, (long long)n_current);
    MNE_CHECK_ARG(n_stored < (1ll << 60) && n_current < (1ll << 60), "mne_draw_rays: store too large");
    const int n = n_global + n_cur;
    draw_rays_kernel<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(key, rays7, n_stored, per_kf, n_global, current7, n_current, n_cur, poses,
                                                                        n_kf, rays_o, rays_d, target_rgb, target_d, rows_out);
    MNE_CHECK_CUDA(cudaGetLastError());
    return 0;
}
