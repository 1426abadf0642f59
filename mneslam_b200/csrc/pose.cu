// Camera-pose side of the tracking / pose-refinement iteration, as two small kernels instead of the
// ~100 element-wise torch kernels autograd spends on a 7-parameter problem:
//   mne_pose_rays : (rotation parameters, translation) -> R -> rays_o = t, rays_d = sum_k dirs_k R[:, k]
//                   (reference optimization/utils.py:161-210 at_to_transform_matrix / qt_to_transform_matrix and the ray
//                   assembly of mp_slam/mapper.py:392-393)
//   mne_pose_step : d loss / d rays_o, d rays_d  ->  d loss / d R, d t (a reduction over the rays)  ->  chain rule through
//                   the parameterisation  ->  Adam on the 6 or 7 pose parameters (torch.optim.Adam's formulas:
//                   get_pose_param_optim, mneslam_mp.py:577-584)
// The chain rule is forward-mode automatic differentiation on a tiny dual-number type, so the derivative is that of the very
// expression that builds R: no hand-derived Jacobians per parameterisation.
#include "common.cuh"

namespace mne {

template <int N>
struct Dual {
    float v;
    float d[N];
};
template <int N> __device__ __forceinline__ Dual<N> cst(float c) { Dual<N> r; r.v = c; for (int i = 0; i < N; ++i) r.d[i] = 0.f; return r; }
template <int N> __device__ __forceinline__ Dual<N> var(float c, int k) { Dual<N> r = cst<N>(c); r.d[k] = 1.f; return r; }
template <int N> __device__ __forceinline__ Dual<N> operator+(const Dual<N>& a, const Dual<N>& b) {
    Dual<N> r; r.v = a.v + b.v; for (int i = 0; i < N; ++i) r.d[i] = a.d[i] + b.d[i]; return r;
}
template <int N> __device__ __forceinline__ Dual<N> operator-(const Dual<N>& a, const Dual<N>& b) {
    Dual<N> r; r.v = a.v - b.v; for (int i = 0; i < N; ++i) r.d[i] = a.d[i] - b.d[i]; return r;
}
template <int N> __device__ __forceinline__ Dual<N> operator*(const Dual<N>& a, const Dual<N>& b) {
    Dual<N> r; r.v = a.v * b.v; for (int i = 0; i < N; ++i) r.d[i] = a.d[i] * b.v + a.v * b.d[i]; return r;
}
template <int N> __device__ __forceinline__ Dual<N> operator/(const Dual<N>& a, const Dual<N>& b) {
    Dual<N> r; r.v = a.v / b.v; for (int i = 0; i < N; ++i) r.d[i] = (a.d[i] - r.v * b.d[i]) / b.v; return r;
}
template <int N> __device__ __forceinline__ Dual<N> dsqrt(const Dual<N>& a) {
    Dual<N> r; r.v = sqrtf(a.v); for (int i = 0; i < N; ++i) r.d[i] = a.d[i] / (2.f * r.v); return r;
}
template <int N> __device__ __forceinline__ Dual<N> dsin(const Dual<N>& a) {
    Dual<N> r; r.v = sinf(a.v); const float c = cosf(a.v); for (int i = 0; i < N; ++i) r.d[i] = c * a.d[i]; return r;
}
template <int N> __device__ __forceinline__ Dual<N> dcos(const Dual<N>& a) {
    Dual<N> r; r.v = cosf(a.v); const float s = -sinf(a.v); for (int i = 0; i < N; ++i) r.d[i] = s * a.d[i]; return r;
}

// rot_rep 0: axis-angle, Rodrigues  I + sin(t) K + (1 - cos(t)) K^2,  K = [r / |r|]x
template <int N>
__device__ __forceinline__ void rotation_axis_angle(const Dual<N> (&r)[3], Dual<N> (&R)[9]) {
    const Dual<N> theta = dsqrt(r[0] * r[0] + r[1] * r[1] + r[2] * r[2]);
    const Dual<N> k0 = r[0] / theta, k1 = r[1] / theta, k2 = r[2] / theta;
    const Dual<N> z = cst<N>(0.f);
    const Dual<N> K[9] = {z, z - k2, k1, k2, z, z - k0, z - k1, k0, z};
    const Dual<N> s = dsin(theta), c1 = cst<N>(1.f) - dcos(theta);
    for (int a = 0; a < 3; ++a)
        for (int b = 0; b < 3; ++b) {
            const Dual<N> kk = K[a * 3 + 0] * K[0 * 3 + b] + K[a * 3 + 1] * K[1 * 3 + b] + K[a * 3 + 2] * K[2 * 3 + b];
            R[a * 3 + b] = cst<N>(a == b ? 1.f : 0.f) + s * K[a * 3 + b] + c1 * kk;
        }
}
// rot_rep 1: real-first quaternion, not assumed unit (pytorch3d quaternion_to_matrix)
template <int N>
__device__ __forceinline__ void rotation_quaternion(const Dual<N> (&q)[4], Dual<N> (&R)[9]) {
    const Dual<N> w = q[0], x = q[1], y = q[2], z = q[3];
    const Dual<N> s = cst<N>(2.f) / (w * w + x * x + y * y + z * z), one = cst<N>(1.f);
    R[0] = one - s * (y * y + z * z); R[1] = s * (x * y - z * w);       R[2] = s * (x * z + y * w);
    R[3] = s * (x * y + z * w);       R[4] = one - s * (x * x + z * z); R[5] = s * (y * z - x * w);
    R[6] = s * (x * z - y * w);       R[7] = s * (y * z + x * w);       R[8] = one - s * (x * x + y * y);
}

template <int NP>
__device__ __forceinline__ void rotation_of(const float* rot, Dual<NP> (&R)[9]) {
    if (NP == 3) {
        Dual<NP> r[3];
        for (int i = 0; i < 3; ++i) r[i] = var<NP>(rot[i], i);
        rotation_axis_angle<NP>(r, R);
    } else {
        Dual<NP> q[4];
        for (int i = 0; i < 4; ++i) q[i] = var<NP>(rot[i], i % NP);
        rotation_quaternion<NP>(q, R);
    }
}

template <int NP>
__global__ void __launch_bounds__(256)
pose_rays_kernel(const float* __restrict__ rot, const float* __restrict__ trans, const float* __restrict__ dirs, int n,
                 float* __restrict__ rays_o, float* __restrict__ rays_d) {
    __shared__ float Rs[9];
    if (threadIdx.x == 0) {
        Dual<NP> R[9];
        rotation_of<NP>(rot, R);
        for (int i = 0; i < 9; ++i) Rs[i] = R[i].v;
    }
    __syncthreads();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float d0 = dirs[i * 3 + 0], d1 = dirs[i * 3 + 1], d2 = dirs[i * 3 + 2];
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        rays_d[i * 3 + a] = __fadd_rn(__fadd_rn(__fmul_rn(d0, Rs[a * 3 + 0]), __fmul_rn(d1, Rs[a * 3 + 1])), __fmul_rn(d2, Rs[a * 3 + 2]));
        rays_o[i * 3 + a] = trans[a];
    }
}

struct PoseStep {
    float* rot;                // [NP] rotation parameters, updated in place
    float* trans;              // [3]
    const float* dirs;         // [n,3] camera-frame directions
    const float* g_rays_o;     // [n,3] d loss / d rays_o
    const float* g_rays_d;     // [n,3]
    int n;
    float* exp_avg;            // [NP + 3] Adam moments (rotation first), zero-initialised by the caller
    float* exp_avg_sq;
    int64_t* step;             // device counter, incremented here
    float lr_rot, lr_trans, beta1, beta2, eps;
    float* grad_out;           // optional [NP + 3]: the pose gradient, for inspection
};

template <int NP>
__global__ void __launch_bounds__(256) pose_step_kernel(const PoseStep s) {
    // d loss / d R[a][k] = sum_i g_d[i][a] * dirs[i][k],  d loss / d t[a] = sum_i g_o[i][a]
    float acc[12];
#pragma unroll
    for (int j = 0; j < 12; ++j) acc[j] = 0.f;
    for (int i = threadIdx.x; i < s.n; i += blockDim.x) {
        const float d0 = s.dirs[i * 3 + 0], d1 = s.dirs[i * 3 + 1], d2 = s.dirs[i * 3 + 2];
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            const float g = s.g_rays_d[i * 3 + a];
            acc[a * 3 + 0] += g * d0; acc[a * 3 + 1] += g * d1; acc[a * 3 + 2] += g * d2;
            acc[9 + a] += s.g_rays_o[i * 3 + a];
        }
    }
    __shared__ float part[8][12];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
    for (int j = 0; j < 12; ++j) {
        float v = acc[j];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (lane == 0) part[warp][j] = v;
    }
    __syncthreads();
    if (threadIdx.x != 0) return;
    float G[12];
    for (int j = 0; j < 12; ++j) {
        float v = 0.f;
        for (int w = 0; w < 8; ++w) v += part[w][j];
        G[j] = v;
    }
    Dual<NP> R[9];
    rotation_of<NP>(s.rot, R);
    float grad[NP + 3];
    for (int m = 0; m < NP; ++m) {
        float v = 0.f;
        for (int j = 0; j < 9; ++j) v += G[j] * R[j].d[m];
        grad[m] = v;
    }
    for (int a = 0; a < 3; ++a) grad[NP + a] = G[9 + a];
    // Adam (torch.optim.Adam, no weight decay, no amsgrad), one group for the rotation, one for the translation
    const int64_t t = *s.step + 1;
    *s.step = t;
    const float bc1 = (float)(1.0 - pow((double)s.beta1, (double)t));
    const float bc2_sqrt = (float)sqrt(1.0 - pow((double)s.beta2, (double)t));
    for (int j = 0; j < NP + 3; ++j) {
        const float g = grad[j];
        float m = s.exp_avg[j], v = s.exp_avg_sq[j];
        m = m + (1.0f - s.beta1) * (g - m);
        v = __fmaf_rn((1.0f - s.beta2) * g, g, s.beta2 * v);
        const float lr = j < NP ? s.lr_rot : s.lr_trans;
        const float denom = sqrtf(v) / bc2_sqrt + s.eps;
        float* p = j < NP ? s.rot + j : s.trans + (j - NP);
        *p = *p - (lr / bc1) * (m / denom);
        s.exp_avg[j] = m; s.exp_avg_sq[j] = v;
        if (s.grad_out) s.grad_out[j] = g;
    }
}

}  // namespace mne

extern "C" int mne_pose_rays(int32_t rot_rep, const float* rot, const float* trans, const float* dirs_cam, int32_t n_rays,
                             float* rays_o, float* rays_d, mne_stream_t stream) {
    using namespace mne;
    MNE_CHECK_ARG(rot_rep == 0 || rot_rep == 1, "mne_pose_rays: rot_rep must be 0 (axis-angle) or 1 (quaternion)");
    MNE_CHECK_ARG(n_rays >= 0, "mne_pose_rays: n_rays < 0");
    if (n_rays == 0) return 0;
    MNE_CHECK_ARG(rot && trans && dirs_cam && rays_o && rays_d, "mne_pose_rays: null argument");
    const int blocks = (n_rays + 255) / 256;
    if (rot_rep == 0) pose_rays_kernel<3><<<blocks, 256, 0, (cudaStream_t)stream>>>(rot, trans, dirs_cam, n_rays, rays_o, rays_d);
    else pose_rays_kernel<4><<<blocks, 256, 0, (cudaStream_t)stream>>>(rot, trans, dirs_cam, n_rays, rays_o, rays_d);
    MNE_CHECK_CUDA(cudaGetLastError());
    return 0;
}

extern "C" int mne_pose_step(int32_t rot_rep, float* rot, float* trans, const float* dirs_cam, const float* g_rays_o,
                             const float* g_rays_d, int32_t n_rays, float* exp_avg, float* exp_avg_sq, int64_t* step_dev,
                             float lr_rot, float lr_trans, float beta1, float beta2, float eps, float* grad_out,
                             mne_stream_t stream) {
    using namespace mne;
    MNE_CHECK_ARG(rot_rep == 0 || rot_rep == 1, "mne_pose_step: rot_rep must be 0 (axis-angle) or 1 (quaternion)");
    MNE_CHECK_ARG(n_rays >= 0, "mne_pose_step: n_rays < 0");
    MNE_CHECK_ARG(rot && trans && exp_avg && exp_avg_sq && step_dev && (n_rays == 0 || (dirs_cam && g_rays_o && g_rays_d)),
                  "mne_pose_step: null argument");
    PoseStep s{rot, trans, dirs_cam, g_rays_o, g_rays_d, n_rays, exp_avg, exp_avg_sq, step_dev, lr_rot, lr_trans, beta1, beta2, eps, grad_out};
    if (rot_rep == 0) pose_step_kernel<3><<<1, 256, 0, (cudaStream_t)stream>>>(s);
    else pose_step_kernel<4><<<1, 256, 0, (cudaStream_t)stream>>>(s);
    MNE_CHECK_CUDA(cudaGetLastError());
    return 0;
}
