// Dense Adam step fused with the gradient zero-fill, for the planes and the decoder weights.
// Replaces torch.optim.Adam.step() + zero_grad() of the mapping loop (reference mneslam_mp.py:459-469,
// mp_slam/mapper.py:161-162).  Arithmetic follows torch's single-tensor Adam (no amsgrad, no maximize):
//   g += wd * p;  m = lerp(m, g, 1 - b1);  v = b2 * v + (1 - b2) * g * g;
//   p += -(lr / (1 - b1^t)) * (m / (sqrt(v) / sqrt(1 - b2^t) + eps))
// Pure streaming: 16 B/param read (p, g, m, v) + 16 B/param written (p, m, v, g = 0).
#include "common.cuh"

namespace mne {

struct AdamArgs {
    float* p;
    float* g;
    float* m;
    float* v;
    int64_t n;
    float lr, beta1, beta2, eps, weight_decay;
    const int64_t* step_dev;   // device step counter (already incremented) or NULL
    int64_t step_host;
    int zero_grad;
};

__device__ __forceinline__ void adam_one(float& p, float& g, float& m, float& v, const AdamArgs& a, float step_size,
                                         float bc2_sqrt) {
    float grad = g;
    if (a.weight_decay != 0.0f) grad = fmaf(a.weight_decay, p, grad);
    m = m + (1.0f - a.beta1) * (grad - m);
    v = __fmaf_rn((1.0f - a.beta2) * grad, grad, a.beta2 * v);
    const float denom = sqrtf(v) / bc2_sqrt + a.eps;
    p = p - step_size * (m / denom);
    if (a.zero_grad) g = 0.0f;
}

__global__ void __launch_bounds__(256) adam_kernel(const AdamArgs a) {
    const double t = (double)(a.step_dev ? *a.step_dev : a.step_host);
    const float step_size = (float)((double)a.lr / (1.0 - pow((double)a.beta1, t)));
    const float bc2_sqrt = (float)sqrt(1.0 - pow((double)a.beta2, t));
    const int64_t n4 = a.n >> 2;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    float4* p4 = reinterpret_cast<float4*>(a.p);
    float4* g4 = reinterpret_cast<float4*>(a.g);
    float4* m4 = reinterpret_cast<float4*>(a.m);
    float4* v4 = reinterpret_cast<float4*>(a.v);
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
        float4 p = p4[i], g = g4[i], m = m4[i], v = v4[i];
        adam_one(p.x, g.x, m.x, v.x, a, step_size, bc2_sqrt);
        adam_one(p.y, g.y, m.y, v.y, a, step_size, bc2_sqrt);
        adam_one(p.z, g.z, m.z, v.z, a, step_size, bc2_sqrt);
        adam_one(p.w, g.w, m.w, v.w, a, step_size, bc2_sqrt);
        p4[i] = p; m4[i] = m; v4[i] = v;
        if (a.zero_grad) g4[i] = g;
    }
    for (int64_t i = (n4 << 2) + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < a.n; i += stride) {
        float p = a.p[i], g = a.g[i], m = a.m[i], v = a.v[i];
        adam_one(p, g, m, v, a, step_size, bc2_sqrt);
        a.p[i] = p; a.m[i] = m; a.v[i] = v;
        if (a.zero_grad) a.g[i] = g;
    }
}

__global__ void tick_kernel(int64_t* step) { if (threadIdx.x == 0 && blockIdx.x == 0) *step += 1; }

int sm_count();

}  // namespace mne

extern "C" int mne_adam_tick(int64_t* step_dev, mne_stream_t stream) {
    using namespace mne;
    MNE_CHECK_ARG(step_dev, "mne_adam_tick: null step counter");
    tick_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(step_dev);
    MNE_CHECK_CUDA(cudaGetLastError());
    return 0;
}

extern "C" int mne_adam_step(float* param, float* grad, float* exp_avg, float* exp_avg_sq, int64_t n, float lr,
                             float beta1, float beta2, float eps, float weight_decay, const int64_t* step_dev,
                             int64_t step_host, int32_t zero_grad, mne_stream_t stream) {
    using namespace mne;
    MNE_CHECK_ARG(param && grad && exp_avg && exp_avg_sq, "mne_adam_step: null tensor");
    MNE_CHECK_ARG(n >= 0, "mne_adam_step: n < 0");
    MNE_CHECK_ARG(step_dev || step_host >= 1, "mne_adam_step: step must be >= 1");
    MNE_CHECK_ARG((((uintptr_t)param | (uintptr_t)grad | (uintptr_t)exp_avg | (uintptr_t)exp_avg_sq) & 15) == 0,
                  "mne_adam_step: tensors must be 16-byte aligned");
    if (n == 0) return 0;
    AdamArgs a{param, grad, exp_avg, exp_avg_sq, n, lr, beta1, beta2, eps, weight_decay, step_dev, step_host, zero_grad};
    int64_t blocks = (n / 4 + 255) / 256;
    const int64_t cap = (int64_t)sm_count() * 8;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    adam_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(a);
    MNE_CHECK_CUDA(cudaGetLastError());
    return 0;
}
