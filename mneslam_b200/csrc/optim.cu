// Dense Adam step fused with the gradient zero-fill, for the planes and the decoder weights.
// Replaces torch.optim.Adam.step() + zero_grad() of the mapping loop (reference mneslam_mp.py:459-469,
// mp_slam/mapper.py:161-162).  Arithmetic follows torch's single-tensor Adam (no amsgrad, no maximize):
//   g += wd * p;  m = lerp(m, g, 1 - b1);  v = b2 * v + (1 - b2) * g * g;
//   p += -(lr / (1 - b1^t)) * (m / (sqrt(v) / sqrt(1 - b2^t) + eps))
// Pure streaming: 16 B/param read (p, g, m, v) + 16 B/param written (p, m, v, g = 0).
#include <algorithm>
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
    // IEEE division and square root, as torch's kernels compute them (~50 instructions per element; with approximate
    // intrinsics the two-rank peer kernel measured 84 -> 79 us: not what bounds these kernels)
    const float denom = sqrtf(v) / bc2_sqrt + a.eps;
    p = p - step_size * (m / denom);
    if (a.zero_grad) g = 0.0f;
}

// Grid-stride kernels below are launched with exactly as many CTAs as are resident at once (occupancy API): with a fixed
// "8 per SM" the 47-register kernels fitted only 5, so the grid ran as 1.6 waves with a half-empty second one (ncu:
// 50 % of the warp slots active, profiles/r01_adam_before_after.md).
template <typename K>
static int resident_ctas(K kernel, int threads) {
    int per_sm = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, threads, 0) != cudaSuccess || per_sm < 1) per_sm = 1;
    return per_sm;
}

__global__ void __launch_bounds__(256, 6) adam_kernel(const AdamArgs a) {
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

// Several parameter groups that live back to back in four flat buffers (same offsets in param / grad / exp_avg /
// exp_avg_sq): one launch for the whole map instead of one per tensor, so the HBM pipes are not drained and
// refilled between tensors.
struct AdamSegments {
    float* p;
    float* g;
    float* m;
    float* v;
    int n_seg;
    int64_t end4[MNE_ADAM_MAX_SEGMENTS];      // exclusive end of each segment, in float4 units
    float lr[MNE_ADAM_MAX_SEGMENTS], eps[MNE_ADAM_MAX_SEGMENTS], wd[MNE_ADAM_MAX_SEGMENTS];
    float beta1, beta2;
    const int64_t* step_dev;
    int64_t step_host;
    int zero_grad;
    uint8_t* touched;      // one byte per float4 group: has this group ever received a gradient?  (NULL = dense)
    uint8_t* dirty;        // one byte per 32 floats of [0, dirty_end4 * 4): did the backward pass add into this texel this step?
    int64_t dirty_end4;
};

// The kernel is bound by memory latency (ncu: long-scoreboard stalls, 0.5 eligible warps per cycle), so every thread keeps
// two groups in flight and issues the independent loads of both before anything depends on them: flags + gradients
// first, then parameters + moments of the groups that need the update.
template <bool DIRTY, int U, int CTAS>
__global__ void __launch_bounds__(256, CTAS) adam_segments_kernel(const __grid_constant__ AdamSegments s) {
    const double t = (double)(s.step_dev ? *s.step_dev : s.step_host);
    const float bc1 = (float)(1.0 - pow((double)s.beta1, t));
    const float bc2_sqrt = (float)sqrt(1.0 - pow((double)s.beta2, t));
    const int64_t n4 = s.end4[s.n_seg - 1];
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    float4* p4 = reinterpret_cast<float4*>(s.p);
    float4* g4 = reinterpret_cast<float4*>(s.g);
    float4* m4 = reinterpret_cast<float4*>(s.m);
    float4* v4 = reinterpret_cast<float4*>(s.v);
    const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
    int seg[U] = {};
    const int lane = threadIdx.x & 31;
    // (the loop condition is the WARP's first group, so that all 32 lanes reach the __syncwarp of the DIRTY variant together)
    for (int64_t i0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i0 - lane < n4; i0 += U * stride) {
        int64_t idx[U];
        bool act[U];
        float4 g[U];
        uint8_t f[U];
#pragma unroll
        for (int k = 0; k < U; ++k) { idx[k] = i0 + k * stride; act[k] = idx[k] < n4; f[k] = 1; }
        bool nz[U];
        float4 p[U], m[U], v[U];
        if constexpr (DIRTY) {
            // The backward pass flagged (one byte per 32 floats = 8 groups = 8 neighbouring lanes) every texel it added a
            // gradient into; an unflagged texel's gradient is zero by contract and is not read.  With the two flag maps
            // known, the loads of gradient, parameter and moments are all independent of each other.
            uint8_t dq[U];
            bool want[U];
#pragma unroll
            for (int k = 0; k < U; ++k) {
                dq[k] = 1;
                if (act[k]) {
                    if (s.touched) f[k] = s.touched[idx[k]];
                    if (idx[k] < s.dirty_end4) dq[k] = s.dirty[idx[k] >> 3];
                }
            }
            __syncwarp();                     // all eight lanes of a texel have read its flag
#pragma unroll
            for (int k = 0; k < U; ++k) {
                if (act[k]) while (idx[k] >= s.end4[seg[k]]) ++seg[k];
                if (act[k] && dq[k] && (idx[k] & 7) == 0 && idx[k] < s.dirty_end4) s.dirty[idx[k] >> 3] = 0;     // for the next iteration
                want[k] = act[k] && (!s.touched || s.wd[seg[k]] != 0.0f || f[k] || dq[k]);
            }
#pragma unroll
            for (int k = 0; k < U; ++k) {
                g[k] = (act[k] && dq[k]) ? g4[idx[k]] : zero;
                if (want[k]) { p[k] = p4[idx[k]]; m[k] = m4[idx[k]]; v[k] = v4[idx[k]]; }
            }
#pragma unroll
            for (int k = 0; k < U; ++k) {
                nz[k] = g[k].x != 0.0f || g[k].y != 0.0f || g[k].z != 0.0f || g[k].w != 0.0f;
                // (flagged, but every addition was a zero and the group never saw a gradient: still the identity update)
                act[k] = want[k] && (!s.touched || s.wd[seg[k]] != 0.0f || f[k] || nz[k]);
            }
        } else {
#pragma unroll
            for (int k = 0; k < U; ++k) {
                g[k] = act[k] ? g4[idx[k]] : zero;
                if (s.touched && act[k]) f[k] = s.touched[idx[k]];
            }
#pragma unroll
            for (int k = 0; k < U; ++k) {
                if (act[k]) while (idx[k] >= s.end4[seg[k]]) ++seg[k];
                nz[k] = g[k].x != 0.0f || g[k].y != 0.0f || g[k].z != 0.0f || g[k].w != 0.0f;
                // A group that never saw a gradient has m = v = g = 0: its dense Adam update is exactly the identity, so
                // skipping it (no p / m / v traffic) gives bit-identical parameters.  Once touched, always updated.
                if (s.touched && s.wd[seg[k]] == 0.0f && !nz[k] && !f[k]) act[k] = false;
            }
#pragma unroll
            for (int k = 0; k < U; ++k)
                if (act[k]) { p[k] = p4[idx[k]]; m[k] = m4[idx[k]]; v[k] = v4[idx[k]]; }
        }
#pragma unroll
        for (int k = 0; k < U; ++k) {
            if (!act[k]) continue;
            AdamArgs a{};
            a.beta1 = s.beta1; a.beta2 = s.beta2; a.eps = s.eps[seg[k]]; a.weight_decay = s.wd[seg[k]]; a.zero_grad = s.zero_grad;
            const float step_size = (float)((double)s.lr[seg[k]] / (double)bc1);
            if (s.touched && a.weight_decay == 0.0f && !f[k]) s.touched[idx[k]] = 1;
            adam_one(p[k].x, g[k].x, m[k].x, v[k].x, a, step_size, bc2_sqrt);
            adam_one(p[k].y, g[k].y, m[k].y, v[k].y, a, step_size, bc2_sqrt);
            adam_one(p[k].z, g[k].z, m[k].z, v[k].z, a, step_size, bc2_sqrt);
            adam_one(p[k].w, g[k].w, m[k].w, v[k].w, a, step_size, bc2_sqrt);
            p4[idx[k]] = p[k]; m4[idx[k]] = m[k]; v4[idx[k]] = v[k];
            if (s.zero_grad && nz[k]) g4[idx[k]] = g[k];       // a zero gradient needs no clearing
        }
    }
}

__global__ void tick_kernel(int64_t* step) { if (threadIdx.x == 0 && blockIdx.x == 0) *step += 1; }

int sm_count();
int adam_spare_sms();
int peer_barrier_polls_log2();
static int adam_sms() { return std::max(1, sm_count() - adam_spare_sms()); }

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
    const int64_t cap = (int64_t)sm_count() * resident_ctas(adam_kernel, 256);
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    adam_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(a);
    MNE_CHECK_CUDA(cudaGetLastError());
    return 0;
}

extern "C" int mne_adam_step_segments_dirty(float* param, float* grad, float* exp_avg, float* exp_avg_sq, int32_t n_segments,
                                            const int64_t* segment_ends, const float* lr, const float* eps,
                                            const float* weight_decay, float beta1, float beta2, const int64_t* step_dev,
                                            int64_t step_host, int32_t zero_grad, uint8_t* touched, uint8_t* dirty,
                                            int64_t dirty_floats, mne_stream_t stream) {
    using namespace mne;
    MNE_CHECK_ARG(param && grad && exp_avg && exp_avg_sq && segment_ends && lr && eps && weight_decay,
                  "mne_adam_step_segments: null argument");
    MNE_CHECK_ARG(n_segments >= 1 && n_segments <= MNE_ADAM_MAX_SEGMENTS, "mne_adam_step_segments: 1..%d segments supported, got %d",
                  MNE_ADAM_MAX_SEGMENTS, n_segments);
    MNE_CHECK_ARG(step_dev || step_host >= 1, "mne_adam_step_segments: step must be >= 1");
    MNE_CHECK_ARG((((uintptr_t)param | (uintptr_t)grad | (uintptr_t)exp_avg | (uintptr_t)exp_avg_sq) & 15) == 0,
                  "mne_adam_step_segments: buffers must be 16-byte aligned");
    AdamSegments s{};
    s.p = param; s.g = grad; s.m = exp_avg; s.v = exp_avg_sq; s.n_seg = n_segments;
    int64_t prev = 0;
    for (int i = 0; i < n_segments; ++i) {
        MNE_CHECK_ARG(segment_ends[i] >= prev && (segment_ends[i] & 3) == 0,
                      "mne_adam_step_segments: segment ends must be ascending multiples of 4 elements");
        s.end4[i] = segment_ends[i] >> 2;
        s.lr[i] = lr[i]; s.eps[i] = eps[i]; s.wd[i] = weight_decay[i];
        prev = segment_ends[i];
    }
    if (prev == 0) return 0;
    s.beta1 = beta1; s.beta2 = beta2; s.step_dev = step_dev; s.step_host = step_host; s.zero_grad = zero_grad;
    s.touched = touched;
    if (dirty) {
        MNE_CHECK_ARG(dirty_floats >= 0 && (dirty_floats & 31) == 0 && dirty_floats <= prev && (((uintptr_t)grad) & 127) == 0,
                      "mne_adam_step_segments_dirty: the flagged range must be whole 128-byte texels inside the buffer (dirty_floats=%lld)",
                      (long long)dirty_floats);
        MNE_CHECK_ARG(zero_grad, "mne_adam_step_segments_dirty: flags need zero_grad (an unflagged texel's gradient must be zero)");
        s.dirty = dirty; s.dirty_end4 = dirty_floats >> 2;
    }
    int64_t blocks = ((prev >> 2) + 255) / 256;
    if (dirty) {
        const int64_t cap = (int64_t)adam_sms() * resident_ctas(adam_segments_kernel<true, 1, 6>, 256);
        if (blocks > cap) blocks = cap;
        adam_segments_kernel<true, 1, 6><<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(s);
    } else {
        // one group per thread, eight CTAs per SM (64 warps): the pass is bound by loads in flight and by the ~200 instructions
        // of IEEE division / square root per group, both of which more resident warps hide.  Measured on B200 (room0 step):
        // (U groups per thread, CTAs per SM) = (2, 4) 0.236 ms, (1, 5) 0.241, (1, 6) 0.233, (1, 8) 0.227, (2, 5, spilling) 0.264
        const int64_t cap = (int64_t)adam_sms() * resident_ctas(adam_segments_kernel<false, 1, 8>, 256);
        if (blocks > cap) blocks = cap;
        adam_segments_kernel<false, 1, 8><<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(s);
    }
    MNE_CHECK_CUDA(cudaGetLastError());
    return 0;
}

extern "C" int mne_adam_step_segments(float* param, float* grad, float* exp_avg, float* exp_avg_sq, int32_t n_segments,
                                      const int64_t* segment_ends, const float* lr, const float* eps,
                                      const float* weight_decay, float beta1, float beta2, const int64_t* step_dev,
                                      int64_t step_host, int32_t zero_grad, uint8_t* touched, mne_stream_t stream) {
    return mne_adam_step_segments_dirty(param, grad, exp_avg, exp_avg_sq, n_segments, segment_ends, lr, eps, weight_decay, beta1, beta2,
                                        step_dev, step_host, zero_grad, touched, nullptr, 0, stream);
}

// ---------------------------------------------------------------------------------------------------------
// Data-parallel optimiser step as ONE kernel over NVLink peer memory (replaces reduce-scatter + Adam + all-gather).
//
// Every rank holds the whole flat gradient and parameter buffers in peer-mapped memory.  Rank r owns the slice
// [begin, begin + len) of the parameters (and keeps the Adam moments of that slice only).  For each 16-byte group
// of its slice the kernel
//   - loads the gradient group from EVERY rank's buffer (local load + P2P loads) and sums them,
//   - clears the groups it read that were non-zero (so nobody needs a separate 129 MB memset pass),
//   - applies Adam (same arithmetic and the same never-touched skip as adam_segments_kernel),
//   - stores the updated parameters into EVERY rank's parameter buffer (local store + P2P stores).
// NVLink traffic per rank: (G-1)/G of the gradient in, (G-1)/G of the parameters out -- what reduce-scatter +
// all-gather move, but overlapped with the HBM work of the update and without staging buffers.
// The caller brackets the launch with cross-rank barriers (gradients complete before, parameters visible after).
// ---------------------------------------------------------------------------------------------------------
namespace mne {

struct AdamPeers {
    float* param[MNE_MAX_PEERS];
    float* grad[MNE_MAX_PEERS];
    int world, rank;
    int64_t total4;            // float4 groups in the whole (padded) buffer
    int64_t local4;            // float4 groups this rank owns
    int chunk4_log2;           // ownership is block-cyclic: chunk c (2^chunk4_log2 groups) belongs to rank c % world
    float* m;                  // moments of the owned groups, packed in ownership order
    float* v;
    int n_seg;
    int64_t end4[MNE_ADAM_MAX_SEGMENTS];      // GLOBAL exclusive segment ends, float4 units
    float lr[MNE_ADAM_MAX_SEGMENTS], eps[MNE_ADAM_MAX_SEGMENTS], wd[MNE_ADAM_MAX_SEGMENTS];
    float beta1, beta2;
    const int64_t* step_dev;
    int64_t step_host;
    uint8_t* touched;
    float* mc_param;           // NVLink-switch multicast addresses of the same buffers, or NULL (plain peer loads / stores)
    float* mc_grad;
    const uint8_t* dirty;      // LOCAL [world][local4 / 8]: rank r's "non-zero this step" flags of the texels this rank owns
    int use_dirty;
    const float4* inbox;       // LOCAL [world][local4] (or NULL): the peers' flagged gradient groups, pushed here by mne_grad_push
};

// one flag byte per 32 consecutive floats (a 128-byte texel): does this rank's gradient have anything there?
// A mapping iteration touches ~20 % of the texels, so the owners fetch only flagged groups from their peers.  The flag
// is PUSHED to the texel's owner (slot `rank` of the owner's flag array, packed in ownership order): peer stores are
// fire-and-forget, whereas flag loads from the peers cost the optimiser kernel 60 us of NVLink round trips (measured).
struct DirtyPush {
    uint8_t* dst[MNE_MAX_PEERS];
    float4* inbox[MNE_MAX_PEERS];   // optional: every rank's inbox [world][owned texels][8 float4]; non-zero texels are PUSHED there
    int world, rank, chunk_tx_log2;
    int push;
    int64_t n32, local32;
    int64_t scan32;                 // texels [0, scan32) are examined; the rest of the buffer is left alone
};
// With inboxes (d.push): a non-zero texel owned by ANOTHER rank is also stored into that owner's inbox (slot `rank`, packed
// in ownership order like the flags) and cleared locally, so that the owner's optimiser kernel reads nothing over NVLink:
// fire-and-forget peer stores (bandwidth bound) instead of peer loads from inside the optimiser (round-trip bound: measured
// 52 us of the 172 us kernel at eight ranks, plus as many peer stores to clear what was read).
__global__ void __launch_bounds__(256) grad_dirty_kernel(float4* __restrict__ g4, const __grid_constant__ DirtyPush d) {
    const int64_t stride = ((int64_t)gridDim.x * blockDim.x) >> 3;
    const int sub = threadIdx.x & 7;
    const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
    // four texels in flight per thread: the pass is a pure read of the gradient buffer, bound by loads in flight
    // (the loop condition is the WARP's first texel, so that all 32 lanes reach the shuffles below together; a lane whose own
    //  texel lies beyond the end only skips its load and its store)
    const int64_t lane_tx = (threadIdx.x & 31) >> 3;
    for (int64_t tw0 = ((int64_t)blockIdx.x * blockDim.x + (threadIdx.x & ~31)) >> 3; tw0 < d.scan32; tw0 += 4 * stride) {
        const int64_t tx0 = tw0 + lane_tx;
        float4 v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int64_t tx = tx0 + u * stride;
            v[u] = tx < d.scan32 ? g4[tx * 8 + sub] : zero;
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int64_t tx = tx0 + u * stride;
            unsigned nz = (v[u].x != 0.0f || v[u].y != 0.0f || v[u].z != 0.0f || v[u].w != 0.0f) ? 1u : 0u;
            nz |= __shfl_xor_sync(0xffffffffu, nz, 1);
            nz |= __shfl_xor_sync(0xffffffffu, nz, 2);
            nz |= __shfl_xor_sync(0xffffffffu, nz, 4);
            if (tx < d.scan32) {
                const uint32_t c = (uint32_t)(tx >> d.chunk_tx_log2);
                const uint32_t owner = c % (uint32_t)d.world;
                const int64_t lt = ((int64_t)(c / (uint32_t)d.world) << d.chunk_tx_log2) + (tx & ((1 << d.chunk_tx_log2) - 1));
                if (sub == 0) d.dst[owner][(int64_t)d.rank * d.local32 + lt] = (uint8_t)nz;
                if (d.push && nz && owner != (uint32_t)d.rank) {
                    d.inbox[owner][((int64_t)d.rank * d.local32 + lt) * 8 + sub] = v[u];
                    g4[tx * 8 + sub] = zero;
                }
            }
        }
    }
}

// the same hand-over when the backward pass already flagged the texels it touched (MneSceneDesc.dirty): no pass over the
// 129 MB gradient buffer, only over the 1 MB flag map, which is cleared for the next iteration on the way
__global__ void __launch_bounds__(256) dirty_push_kernel(uint8_t* __restrict__ local_map, int64_t dense_from32,
                                                         const __grid_constant__ DirtyPush d) {
    const int64_t tx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (tx >= d.n32) return;
    uint8_t f = 1;
    if (tx < dense_from32) { f = local_map[tx]; if (f) local_map[tx] = 0; }
    const uint32_t c = (uint32_t)(tx >> d.chunk_tx_log2);
    const uint32_t owner = c % (uint32_t)d.world;
    const int64_t lt = ((int64_t)(c / (uint32_t)d.world) << d.chunk_tx_log2) + (tx & ((1 << d.chunk_tx_log2) - 1));
    d.dst[owner][(int64_t)d.rank * d.local32 + lt] = f;
}

// dirty_push_kernel + mne_grad_push in one: the flags the backward pass set say where the gradient is non-zero, so only those
// texels (about a sixth of the map per iteration) are read, pushed into their owners' inboxes and cleared -- no pass over the
// whole gradient buffer.  Eight lanes per texel (one float4 each).
// A warp takes 32 texels per step: one coalesced load of their flags, the flags handed to the owners, then only the FLAGGED
// texels owned by another rank are moved -- eight at a time (lane = (slot, 16-byte part); slot j takes the j-th and the
// (j+4)-th flagged texel of the step), all their loads in flight before the first store.  (One texel per eight threads with
// the flag load, the data load and the stores in a dependent chain took 42 us for the 1 MB map of room0: 28 serial round
// trips per thread.)
__global__ void __launch_bounds__(256) dirty_push_data_kernel(uint8_t* __restrict__ local_map, float4* __restrict__ g4,
                                                              const __grid_constant__ DirtyPush d) {
    const int lane = threadIdx.x & 31, sub = lane & 7, slot = lane >> 3;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t n_warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int64_t t0 = warp * 32; t0 < d.scan32; t0 += n_warps * 32) {
        const int64_t tx = t0 + lane;
        uint8_t f = 0;
        uint32_t owner = (uint32_t)d.rank;
        int64_t lt = 0;
        if (tx < d.scan32) {
            f = local_map[tx];
            const uint32_t c = (uint32_t)(tx >> d.chunk_tx_log2);
            owner = c % (uint32_t)d.world;
            lt = ((int64_t)(c / (uint32_t)d.world) << d.chunk_tx_log2) + (tx & ((1 << d.chunk_tx_log2) - 1));
            d.dst[owner][(int64_t)d.rank * d.local32 + lt] = f;
            if (f) local_map[tx] = 0;
        }
        unsigned todo = __ballot_sync(0xffffffffu, f != 0 && owner != (uint32_t)d.rank);
        while (todo) {                                    // warp-uniform
            const unsigned src0 = __fns(todo, 0, slot + 1), src1 = __fns(todo, 0, slot + 5);     // 0xffffffff: no such bit
#pragma unroll
            for (int k = 0; k < 8; ++k) todo &= todo - 1;     // (todo - 1 of 0 wraps, but 0 & anything stays 0)
            const uint32_t o0 = __shfl_sync(0xffffffffu, owner, src0 & 31), o1 = __shfl_sync(0xffffffffu, owner, src1 & 31);
            const int64_t l0 = __shfl_sync(0xffffffffu, lt, src0 & 31), l1 = __shfl_sync(0xffffffffu, lt, src1 & 31);
            float4 v0 = zero, v1 = zero;
            float4* s0 = g4 + (t0 + (src0 & 31)) * 8 + sub;
            float4* s1 = g4 + (t0 + (src1 & 31)) * 8 + sub;
            if (src0 < 32u) v0 = *s0;
            if (src1 < 32u) v1 = *s1;
            if (src0 < 32u) { d.inbox[o0][((int64_t)d.rank * d.local32 + l0) * 8 + sub] = v0; *s0 = zero; }
            if (src1 < 32u) { d.inbox[o1][((int64_t)d.rank * d.local32 + l1) * 8 + sub] = v1; *s1 = zero; }
        }
    }
}

// in-switch reduction / broadcast (NVLS): one load returns the sum over all ranks, one store reaches all ranks
__device__ __forceinline__ float4 multimem_ld_reduce_add(const float4* p) {
    float4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void multimem_st(float4* p, const float4& v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};"
                 :: "l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

// Sparse variant (dirty maps): the NVLink round trips are what bounds this kernel, so they are batched -- all peers'
// flags of the NEXT group are fetched while the current one is processed, and the flagged peers' gradients are loaded
// four at a time before anything depends on them.
__device__ __forceinline__ int64_t owned_group(const AdamPeers& s, int64_t i) {
    return ((((i >> s.chunk4_log2) * s.world + s.rank)) << s.chunk4_log2) + (i & ((1 << s.chunk4_log2) - 1));
}
__device__ __forceinline__ uint32_t peer_flags(const AdamPeers& s, int64_t i) {
    const uint8_t* f = s.dirty + (i >> 3);
    const int64_t local32 = s.local4 >> 3;
    uint32_t mk = 0;
#pragma unroll 4
    for (int r = 0; r < s.world; ++r) mk |= (uint32_t)(f[r * local32] != 0) << r;
    return mk;
}

// U owned groups in flight per thread: the flags of all of them first, then every load that the flags call for -- the local
// parameter / moment loads and the flagged peers' gradients (NVLink round trips, the longest latency in the kernel) of ALL
// U groups are issued before anything depends on them -- then the updates, then the parameter stores to every rank
// (fire-and-forget).  With one group in flight the peer fetch, the local update and the write-back of a thread ran one
// after the other (measured r01: 38 + 49 + 80 us at eight ranks); in flight together they overlap.
constexpr int kPeerU = 2;
constexpr int kPeerFetch = 4;      // peer gradients fetched per group before the sum depends on them

// The same update when every peer gradient was PUSHED into this rank's inbox (mne_grad_push): all sources are local, only this
// rank's own gradient buffer needs clearing.  Written for occupancy -- the kernel is bound by loads in flight (ncu on the
// generic variant: 23 warps per SM, 8.7 long-scoreboard stalls per issue, 26 % of the DRAM peak): 32-bit indices, the
// per-segment step sizes computed once per CTA, F gradients fetched per group before the sum depends on them (at eight ranks
// a flagged texel is flagged by 1.4 ranks on average; more take the loop).  Sums run over the flagged ranks in ascending order.
template <int U, int CTAS>
__global__ void __launch_bounds__(256, CTAS) adam_peers_inbox_kernel(const __grid_constant__ AdamPeers s) {
    constexpr int F = 1;
    __shared__ float s_step[MNE_ADAM_MAX_SEGMENTS];
    const double t = (double)(s.step_dev ? *s.step_dev : s.step_host);
    const float bc2_sqrt = (float)sqrt(1.0 - pow((double)s.beta2, t));
    if ((int)threadIdx.x < s.n_seg) {
        const float bc1 = (float)(1.0 - pow((double)s.beta1, t));
        s_step[threadIdx.x] = (float)((double)s.lr[threadIdx.x] / (double)bc1);
    }
    __syncthreads();
    const int stride = (int)(gridDim.x * blockDim.x);
    const int local4 = (int)s.local4, local32 = local4 >> 3, end_all = (int)s.end4[s.n_seg - 1];
    const int cl = s.chunk4_log2, cmask = (1 << cl) - 1, world = s.world, rank = s.rank;
    float4* __restrict__ m4 = reinterpret_cast<float4*>(s.m);
    float4* __restrict__ v4 = reinterpret_cast<float4*>(s.v);
    float4* p_loc = reinterpret_cast<float4*>(s.param[rank]);
    float4* g_loc = reinterpret_cast<float4*>(s.grad[rank]);
    const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
    int seg[U] = {};
    for (int i0 = (int)(blockIdx.x * blockDim.x + threadIdx.x); i0 < local4; i0 += U * stride) {
        int li[U], gi[U];
        uint32_t mk[U];
        uint8_t tf[U];
        bool on[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            li[u] = i0 + u * stride;
            on[u] = li[u] < local4;
            gi[u] = on[u] ? ((((li[u] >> cl) * world + rank) << cl) + (li[u] & cmask)) : end_all;
            on[u] = on[u] && gi[u] < end_all;                 // beyond: padding
            mk[u] = 0; tf[u] = 1;
            if (on[u]) {
                const uint8_t* f = s.dirty + (li[u] >> 3);
#pragma unroll 4
                for (int r = 0; r < world; ++r) mk[u] |= (uint32_t)(f[r * local32] != 0) << r;
                if (s.touched) tf[u] = s.touched[li[u]];
            }
        }
        float4 p[U], m[U], v[U], x[U][F];
        uint32_t rest[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            if (on[u]) {
                while (gi[u] >= (int)s.end4[seg[u]]) ++seg[u];
                if (s.touched && s.wd[seg[u]] == 0.0f && !mk[u] && !tf[u]) on[u] = false;   // identity update: every replica holds this value
            }
            rest[u] = on[u] ? mk[u] : 0u;
#pragma unroll
            for (int k = 0; k < F; ++k) {
                x[u][k] = zero;
                if (rest[u]) {
                    const int r = __ffs(rest[u]) - 1;
                    rest[u] &= rest[u] - 1;
                    x[u][k] = (r != rank) ? s.inbox[(int64_t)r * local4 + li[u]] : g_loc[gi[u]];
                }
            }
            if (on[u]) { p[u] = p_loc[gi[u]]; m[u] = m4[li[u]]; v[u] = v4[li[u]]; }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            if (!on[u]) continue;
            float4 g = x[u][0];
#pragma unroll
            for (int k = 1; k < F; ++k) { g.x += x[u][k].x; g.y += x[u][k].y; g.z += x[u][k].z; g.w += x[u][k].w; }
            while (rest[u]) {
                float4 y[2];
#pragma unroll
                for (int k = 0; k < 2; ++k) {
                    y[k] = zero;
                    if (rest[u]) {
                        const int r = __ffs(rest[u]) - 1;
                        rest[u] &= rest[u] - 1;
                        y[k] = (r != rank) ? s.inbox[(int64_t)r * local4 + li[u]] : g_loc[gi[u]];
                    }
                }
#pragma unroll
                for (int k = 0; k < 2; ++k) { g.x += y[k].x; g.y += y[k].y; g.z += y[k].z; g.w += y[k].w; }
            }
            if ((mk[u] >> rank) & 1u) g_loc[gi[u]] = zero;
            AdamArgs a{};
            a.beta1 = s.beta1; a.beta2 = s.beta2; a.eps = s.eps[seg[u]]; a.weight_decay = s.wd[seg[u]]; a.zero_grad = 0;
            if (s.touched && a.weight_decay == 0.0f) {
                const bool nz = g.x != 0.0f || g.y != 0.0f || g.z != 0.0f || g.w != 0.0f;
                if (!nz && !tf[u]) continue;             // flagged texel, but this group of it carried no gradient
                if (!tf[u]) s.touched[li[u]] = 1;
            }
            const float step_size = s_step[seg[u]];
            adam_one(p[u].x, g.x, m[u].x, v[u].x, a, step_size, bc2_sqrt);
            adam_one(p[u].y, g.y, m[u].y, v[u].y, a, step_size, bc2_sqrt);
            adam_one(p[u].z, g.z, m[u].z, v[u].z, a, step_size, bc2_sqrt);
            adam_one(p[u].w, g.w, m[u].w, v[u].w, a, step_size, bc2_sqrt);
            m4[li[u]] = m[u]; v4[li[u]] = v[u];
            if (s.mc_param) multimem_st(reinterpret_cast<float4*>(s.mc_param) + gi[u], p[u]);
            else
                for (int r = 0; r < world; ++r) reinterpret_cast<float4*>(s.param[r])[gi[u]] = p[u];
        }
    }
}

// INBOX: the peers' gradients were pushed into this rank's inbox (mne_grad_push): every source but this rank's own gradient
// buffer is local, read-only and needs no clearing -- so no pointer per fetched group has to be kept (registers: the generic
// variant spilled 300 bytes per thread).
template <bool INBOX>
__global__ void __launch_bounds__(256, 3) adam_peers_sparse_kernel(const __grid_constant__ AdamPeers s) {
    // gradients fetched per group before the sum depends on them (a texel flagged by more ranks takes the loop below; at
    // eight ranks a flagged texel is flagged by 1.4 ranks on average)
    constexpr int F = INBOX ? 2 : kPeerFetch;
    const double t = (double)(s.step_dev ? *s.step_dev : s.step_host);
    const float bc1 = (float)(1.0 - pow((double)s.beta1, t));
    const float bc2_sqrt = (float)sqrt(1.0 - pow((double)s.beta2, t));
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const int64_t end_all = s.end4[s.n_seg - 1];
    float4* m4 = reinterpret_cast<float4*>(s.m);
    float4* v4 = reinterpret_cast<float4*>(s.v);
    float4* p_loc = reinterpret_cast<float4*>(s.param[s.rank]);
    float4* g_loc = reinterpret_cast<float4*>(s.grad[s.rank]);
    const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
    int seg[kPeerU] = {};
    for (int64_t i0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i0 < s.local4; i0 += kPeerU * stride) {
        int64_t li[kPeerU], gi[kPeerU];
        uint32_t mk[kPeerU];
        uint8_t tf[kPeerU];
        bool on[kPeerU];
#pragma unroll
        for (int u = 0; u < kPeerU; ++u) {
            li[u] = i0 + u * stride;
            on[u] = li[u] < s.local4;
            gi[u] = on[u] ? owned_group(s, li[u]) : end_all;
            on[u] = on[u] && gi[u] < end_all;                 // beyond: padding
            mk[u] = 0; tf[u] = 1;
            if (on[u]) { mk[u] = peer_flags(s, li[u]); if (s.touched) tf[u] = s.touched[li[u]]; }
        }
        float wd[kPeerU];
#pragma unroll
        for (int u = 0; u < kPeerU; ++u) {
            wd[u] = 0.f;
            if (on[u]) {
                while (gi[u] >= s.end4[seg[u]]) ++seg[u];
                wd[u] = s.wd[seg[u]];
                if (s.touched && wd[u] == 0.0f && !mk[u] && !tf[u]) on[u] = false;   // identity update: every replica holds this value
            }
        }
        // ---- all loads of all groups
        float4 p[kPeerU], m[kPeerU], v[kPeerU], x[kPeerU][F];
        float4* gp[INBOX ? 1 : kPeerU][INBOX ? 1 : F];
        uint32_t rest[kPeerU];
#pragma unroll
        for (int u = 0; u < kPeerU; ++u) {
            rest[u] = on[u] ? mk[u] : 0u;
#pragma unroll
            for (int k = 0; k < F; ++k) {
                x[u][k] = zero;
                if constexpr (!INBOX) gp[u][k] = nullptr;
                if (rest[u]) {
                    const int r = __ffs(rest[u]) - 1;
                    rest[u] &= rest[u] - 1;
                    if constexpr (INBOX) {
                        // pushed by rank r (nothing to clear: the flags are rewritten every step), or this rank's own buffer
                        x[u][k] = (r != s.rank) ? s.inbox[(int64_t)r * s.local4 + li[u]] : g_loc[gi[u]];
                    } else {
                        gp[u][k] = reinterpret_cast<float4*>(s.grad[r]) + gi[u];
                        x[u][k] = *gp[u][k];
                    }
                }
            }
            if (on[u]) { p[u] = p_loc[gi[u]]; m[u] = m4[li[u]]; v[u] = v4[li[u]]; }
        }
        // ---- sums (more than F flagged peers: the rest, four at a time), updates, stores
#pragma unroll
        for (int u = 0; u < kPeerU; ++u) {
            if (!on[u]) continue;
            float4 g = zero;
#pragma unroll
            for (int k = 0; k < F; ++k) {
                g.x += x[u][k].x; g.y += x[u][k].y; g.z += x[u][k].z; g.w += x[u][k].w;       // (unfetched slots hold zero)
                if constexpr (!INBOX) { if (gp[u][k]) *gp[u][k] = zero; }
            }
            if (INBOX && ((mk[u] >> s.rank) & 1u)) g_loc[gi[u]] = zero;
            while (rest[u]) {
                float4 y[4];
                float4* yp[INBOX ? 1 : 4];
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    y[k] = zero;
                    if constexpr (!INBOX) yp[k] = nullptr;
                    if (rest[u]) {
                        const int r = __ffs(rest[u]) - 1;
                        rest[u] &= rest[u] - 1;
                        if constexpr (INBOX) {
                            y[k] = (r != s.rank) ? s.inbox[(int64_t)r * s.local4 + li[u]] : g_loc[gi[u]];
                        } else {
                            yp[k] = reinterpret_cast<float4*>(s.grad[r]) + gi[u];
                            y[k] = *yp[k];
                        }
                    }
                }
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    g.x += y[k].x; g.y += y[k].y; g.z += y[k].z; g.w += y[k].w;
                    if constexpr (!INBOX) { if (yp[k]) *yp[k] = zero; }
                }
            }
            AdamArgs a{};
            a.beta1 = s.beta1; a.beta2 = s.beta2; a.eps = s.eps[seg[u]]; a.weight_decay = wd[u]; a.zero_grad = 0;
            if (s.touched && wd[u] == 0.0f) {
                const bool nz = g.x != 0.0f || g.y != 0.0f || g.z != 0.0f || g.w != 0.0f;
                if (!nz && !tf[u]) continue;             // flagged texel, but this group of it carried no gradient
                if (!tf[u]) s.touched[li[u]] = 1;
            }
            const float step_size = (float)((double)s.lr[seg[u]] / (double)bc1);
            adam_one(p[u].x, g.x, m[u].x, v[u].x, a, step_size, bc2_sqrt);
            adam_one(p[u].y, g.y, m[u].y, v[u].y, a, step_size, bc2_sqrt);
            adam_one(p[u].z, g.z, m[u].z, v[u].z, a, step_size, bc2_sqrt);
            adam_one(p[u].w, g.w, m[u].w, v[u].w, a, step_size, bc2_sqrt);
            m4[li[u]] = m[u]; v4[li[u]] = v[u];
            if (s.mc_param) multimem_st(reinterpret_cast<float4*>(s.mc_param) + gi[u], p[u]);
            else
                for (int r = 0; r < s.world; ++r) reinterpret_cast<float4*>(s.param[r])[gi[u]] = p[u];
        }
    }
}

__global__ void __launch_bounds__(256, 5) adam_peers_kernel(const __grid_constant__ AdamPeers s) {
    const double t = (double)(s.step_dev ? *s.step_dev : s.step_host);
    const float bc1 = (float)(1.0 - pow((double)s.beta1, t));
    const float bc2_sqrt = (float)sqrt(1.0 - pow((double)s.beta2, t));
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    float4* m4 = reinterpret_cast<float4*>(s.m);
    float4* v4 = reinterpret_cast<float4*>(s.v);
    int seg = 0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < s.local4; i += stride) {
        // block-cyclic ownership: contiguous slices would give one rank the densely observed planes and another the
        // barely touched ones (measured: 75 % vs 21 % touched at two ranks); 4 KB chunks dealt round-robin balance it
        const int64_t gi = owned_group(s, i);
        if (gi >= s.end4[s.n_seg - 1]) continue;          // padding
        while (gi >= s.end4[seg]) ++seg;
        float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
        const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
        if (s.mc_grad) {
            g = multimem_ld_reduce_add(reinterpret_cast<const float4*>(s.mc_grad) + gi);
            // (a sum of non-zero terms can cancel to zero; clear whenever any rank may have written: cheap test on the
            //  local copy plus the sum covers every case but exact cancellation with a zero local part, so clear on
            //  either signal and, for that last case, on the touched flag below)
            const float4 loc = reinterpret_cast<const float4*>(s.grad[s.rank])[gi];
            const bool any = g.x != 0.0f || g.y != 0.0f || g.z != 0.0f || g.w != 0.0f ||
                             loc.x != 0.0f || loc.y != 0.0f || loc.z != 0.0f || loc.w != 0.0f;
            if (any || (s.touched && s.touched[i])) multimem_st(reinterpret_cast<float4*>(s.mc_grad) + gi, zero);
        } else
        for (int r = 0; r < s.world; ++r) {
            float4* gp = reinterpret_cast<float4*>(s.grad[r]) + gi;
            const float4 x = *gp;
            if (x.x != 0.0f || x.y != 0.0f || x.z != 0.0f || x.w != 0.0f) {
                g.x += x.x; g.y += x.y; g.z += x.z; g.w += x.w;
                *gp = zero;
            }
        }
        AdamArgs a{};
        a.beta1 = s.beta1; a.beta2 = s.beta2; a.eps = s.eps[seg]; a.weight_decay = s.wd[seg]; a.zero_grad = 0;
        if (s.touched && a.weight_decay == 0.0f) {
            const bool nz = g.x != 0.0f || g.y != 0.0f || g.z != 0.0f || g.w != 0.0f;
            const uint8_t f = s.touched[i];
            if (!nz && !f) continue;                 // identity update: every replica already holds this value
            if (!f) s.touched[i] = 1;
        }
        const float step_size = (float)((double)s.lr[seg] / (double)bc1);
        float4 p = reinterpret_cast<float4*>(s.param[s.rank])[gi], m = m4[i], v = v4[i];
        adam_one(p.x, g.x, m.x, v.x, a, step_size, bc2_sqrt);
        adam_one(p.y, g.y, m.y, v.y, a, step_size, bc2_sqrt);
        adam_one(p.z, g.z, m.z, v.z, a, step_size, bc2_sqrt);
        adam_one(p.w, g.w, m.w, v.w, a, step_size, bc2_sqrt);
        m4[i] = m; v4[i] = v;
        if (s.mc_param) multimem_st(reinterpret_cast<float4*>(s.mc_param) + gi, p);
        else
            for (int r = 0; r < s.world; ++r) reinterpret_cast<float4*>(s.param[r])[gi] = p;
    }
}

}  // namespace mne

extern "C" int mne_adam_step_peers(int32_t world_size, int32_t rank, float* const* param_ptrs, float* const* grad_ptrs,
                                   int64_t n_total, int32_t chunk, float* exp_avg, float* exp_avg_sq, int32_t n_segments,
                                   const int64_t* segment_ends, const float* lr, const float* eps, const float* weight_decay,
                                   float beta1, float beta2, const int64_t* step_dev, int64_t step_host, uint8_t* touched,
                                   float* multicast_param, float* multicast_grad, const uint8_t* dirty, const float* inbox,
                                   mne_stream_t stream) {
    using namespace mne;
    MNE_CHECK_ARG(param_ptrs && grad_ptrs && exp_avg && exp_avg_sq && segment_ends && lr && eps && weight_decay,
                  "mne_adam_step_peers: null argument");
    MNE_CHECK_ARG(world_size >= 1 && world_size <= MNE_MAX_PEERS && rank >= 0 && rank < world_size,
                  "mne_adam_step_peers: world size %d / rank %d out of range (at most %d peers)", world_size, rank, MNE_MAX_PEERS);
    MNE_CHECK_ARG(n_segments >= 1 && n_segments <= MNE_ADAM_MAX_SEGMENTS, "mne_adam_step_peers: 1..%d segments supported", MNE_ADAM_MAX_SEGMENTS);
    MNE_CHECK_ARG(chunk >= 32 && (chunk & (chunk - 1)) == 0 && n_total > 0 && n_total % ((int64_t)chunk * world_size) == 0,
                  "mne_adam_step_peers: chunk must be a power of two >= 32 and n_total a multiple of chunk * world_size");
    MNE_CHECK_ARG(step_dev || step_host >= 1, "mne_adam_step_peers: step must be >= 1");
    AdamPeers s{};
    for (int r = 0; r < world_size; ++r) {
        MNE_CHECK_ARG(param_ptrs[r] && grad_ptrs[r] && ((((uintptr_t)param_ptrs[r]) | ((uintptr_t)grad_ptrs[r])) & 15) == 0,
                      "mne_adam_step_peers: peer buffers must be non-null and 16-byte aligned");
        s.param[r] = param_ptrs[r]; s.grad[r] = grad_ptrs[r];
    }
    s.world = world_size; s.rank = rank; s.total4 = n_total >> 2; s.local4 = (n_total / world_size) >> 2; s.chunk4_log2 = 31 - __builtin_clz((unsigned)chunk) - 2;
    s.m = exp_avg; s.v = exp_avg_sq; s.n_seg = n_segments;
    int64_t prev = 0;
    for (int i = 0; i < n_segments; ++i) {
        MNE_CHECK_ARG(segment_ends[i] >= prev && (segment_ends[i] & 3) == 0,
                      "mne_adam_step_peers: segment ends must be ascending multiples of 4 elements");
        s.end4[i] = segment_ends[i] >> 2; s.lr[i] = lr[i]; s.eps[i] = eps[i]; s.wd[i] = weight_decay[i];
        prev = segment_ends[i];
    }
    MNE_CHECK_ARG(prev <= n_total, "mne_adam_step_peers: segments exceed n_total");
    if (prev == 0) return 0;
    s.beta1 = beta1; s.beta2 = beta2; s.step_dev = step_dev; s.step_host = step_host; s.touched = touched;
    s.mc_param = multicast_param; s.mc_grad = multicast_grad;
    s.use_dirty = dirty != nullptr;
    s.dirty = dirty;
    MNE_CHECK_ARG(!inbox || (dirty && (((uintptr_t)inbox) & 15) == 0), "mne_adam_step_peers: inbox needs the dirty flags and 16-byte alignment");
    s.inbox = reinterpret_cast<const float4*>(inbox);
    int64_t blocks = (s.local4 + 255) / 256;
    const bool lean = s.use_dirty && s.inbox && s.total4 < ((int64_t)1 << 30);       // (32-bit group indices)
    if (lean) {
        // (U groups per thread, CTAs per SM), two emulated ranks on one B200, room0: (2, 3) 97 us, (2, 4) 84, (3, 3) 99,
        // (1, 6) 74, (1, 7) 74, (1, 8) 74 -- resident warps, not groups per thread, hide the latency
        const int64_t cap = (int64_t)adam_sms() * resident_ctas(adam_peers_inbox_kernel<1, 6>, 256);
        if (blocks > cap) blocks = cap;
        adam_peers_inbox_kernel<1, 6><<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(s);
        MNE_CHECK_CUDA(cudaGetLastError());
        return 0;
    }
    const int64_t cap = (int64_t)adam_sms() * (!s.use_dirty ? resident_ctas(adam_peers_kernel, 256)
                                                 : s.inbox ? resident_ctas(adam_peers_sparse_kernel<true>, 256) : resident_ctas(adam_peers_sparse_kernel<false>, 256));
    if (blocks > cap) blocks = cap;
    if (s.use_dirty && s.inbox) adam_peers_sparse_kernel<true><<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(s);
    else if (s.use_dirty) adam_peers_sparse_kernel<false><<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(s);
    else adam_peers_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(s);
    MNE_CHECK_CUDA(cudaGetLastError());
    return 0;
}

static int grad_flags_launch(float* grad, int64_t n_total, int64_t n_scan, int32_t chunk, int32_t world_size, int32_t rank,
                             uint8_t* const* dirty_ptrs, float* const* inbox_ptrs, mne_stream_t stream);

extern "C" int mne_grad_dirty_map(const float* grad, int64_t n_total, int32_t chunk, int32_t world_size, int32_t rank,
                                  uint8_t* const* dirty_ptrs, mne_stream_t stream) {
    return grad_flags_launch(const_cast<float*>(grad), n_total, n_total, chunk, world_size, rank, dirty_ptrs, nullptr, stream);
}

extern "C" int mne_grad_push(float* grad, int64_t n_total, int64_t n_scan, int32_t chunk, int32_t world_size, int32_t rank,
                             uint8_t* const* dirty_ptrs, float* const* inbox_ptrs, mne_stream_t stream) {
    using namespace mne;
    MNE_CHECK_ARG(inbox_ptrs, "mne_grad_push: null inbox array");
    MNE_CHECK_ARG(n_scan >= 0 && n_scan <= n_total && (n_scan & 31) == 0, "mne_grad_push: n_scan must be a multiple of 32 in [0, n_total]");
    return grad_flags_launch(grad, n_total, n_scan, chunk, world_size, rank, dirty_ptrs, inbox_ptrs, stream);
}

static int grad_flags_launch(float* grad, int64_t n_total, int64_t n_scan, int32_t chunk, int32_t world_size, int32_t rank,
                             uint8_t* const* dirty_ptrs, float* const* inbox_ptrs, mne_stream_t stream) {
    using namespace mne;
    MNE_CHECK_ARG(grad && dirty_ptrs, "mne_grad_dirty_map: null argument");
    MNE_CHECK_ARG(world_size >= 1 && world_size <= MNE_MAX_PEERS && rank >= 0 && rank < world_size, "mne_grad_dirty_map: bad world_size / rank");
    MNE_CHECK_ARG(chunk >= 32 && (chunk & (chunk - 1)) == 0 && n_total >= 0 && n_total % ((int64_t)chunk * world_size) == 0 &&
                  (((uintptr_t)grad) & 15) == 0 && (n_total >> 5) / chunk < (int64_t)1 << 31,
                  "mne_grad_dirty_map: chunk must be a power of two >= 32, n_total a multiple of chunk * world_size, grad 16-byte aligned");
    if (n_total == 0) return 0;
    DirtyPush d{};
    for (int r = 0; r < world_size; ++r) {
        MNE_CHECK_ARG(dirty_ptrs[r], "mne_grad_dirty_map: null flag array");
        d.dst[r] = dirty_ptrs[r];
        if (inbox_ptrs) {
            MNE_CHECK_ARG(inbox_ptrs[r] && (((uintptr_t)inbox_ptrs[r]) & 15) == 0, "mne_grad_push: inboxes must be non-null and 16-byte aligned");
            d.inbox[r] = reinterpret_cast<float4*>(inbox_ptrs[r]);
        }
    }
    d.push = inbox_ptrs != nullptr;
    d.world = world_size; d.rank = rank; d.chunk_tx_log2 = 31 - __builtin_clz((unsigned)chunk) - 5;
    d.n32 = n_total >> 5; d.local32 = d.n32 / world_size; d.scan32 = n_scan >> 5;
    if (d.scan32 == 0) return 0;
    int64_t blocks = (d.scan32 * 8 + 255) / 256;
    const int64_t cap = (int64_t)sm_count() * resident_ctas(grad_dirty_kernel, 256);
    if (blocks > cap) blocks = cap;
    grad_dirty_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(reinterpret_cast<float4*>(grad), d);
    MNE_CHECK_CUDA(cudaGetLastError());
    return 0;
}

extern "C" int mne_dirty_push(uint8_t* local_map, int64_t n_total, int64_t dense_from, int32_t chunk, int32_t world_size,
                              int32_t rank, uint8_t* const* dirty_ptrs, mne_stream_t stream) {
    using namespace mne;
    MNE_CHECK_ARG(local_map && dirty_ptrs, "mne_dirty_push: null argument");
    MNE_CHECK_ARG(world_size >= 1 && world_size <= MNE_MAX_PEERS && rank >= 0 && rank < world_size, "mne_dirty_push: bad world_size / rank");
    MNE_CHECK_ARG(chunk >= 32 && (chunk & (chunk - 1)) == 0 && n_total >= 0 && n_total % ((int64_t)chunk * world_size) == 0 &&
                  (n_total >> 5) / chunk < (int64_t)1 << 31 && dense_from >= 0,
                  "mne_dirty_push: chunk must be a power of two >= 32 and n_total a multiple of chunk * world_size");
    if (n_total == 0) return 0;
    DirtyPush d{};
    for (int r = 0; r < world_size; ++r) {
        MNE_CHECK_ARG(dirty_ptrs[r], "mne_dirty_push: null flag array");
        d.dst[r] = dirty_ptrs[r];
    }
    d.world = world_size; d.rank = rank; d.chunk_tx_log2 = 31 - __builtin_clz((unsigned)chunk) - 5;
    d.n32 = n_total >> 5; d.local32 = d.n32 / world_size;
    const int64_t blocks = (d.n32 + 255) / 256;
    dirty_push_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(local_map, dense_from >> 5, d);   // floor: a texel straddling dense_from is flagged
    MNE_CHECK_CUDA(cudaGetLastError());
    return 0;
}

extern "C" int mne_dirty_push_data(uint8_t* local_map, float* grad, int64_t n_total, int64_t n_scan, int32_t chunk, int32_t world_size,
                                   int32_t rank, uint8_t* const* dirty_ptrs, float* const* inbox_ptrs, mne_stream_t stream) {
    using namespace mne;
    MNE_CHECK_ARG(local_map && grad && dirty_ptrs && inbox_ptrs, "mne_dirty_push_data: null argument");
    MNE_CHECK_ARG(world_size >= 1 && world_size <= MNE_MAX_PEERS && rank >= 0 && rank < world_size, "mne_dirty_push_data: bad world_size / rank");
    MNE_CHECK_ARG(chunk >= 32 && (chunk & (chunk - 1)) == 0 && n_total >= 0 && n_total % ((int64_t)chunk * world_size) == 0 &&
                  (n_total >> 5) / chunk < (int64_t)1 << 31 && (((uintptr_t)grad) & 15) == 0,
                  "mne_dirty_push_data: chunk must be a power of two >= 32, n_total a multiple of chunk * world_size, grad 16-byte aligned");
    MNE_CHECK_ARG(n_scan >= 0 && n_scan <= n_total && (n_scan & 31) == 0, "mne_dirty_push_data: n_scan must be a multiple of 32 in [0, n_total]");
    if (n_total == 0 || n_scan == 0) return 0;
    DirtyPush d{};
    for (int r = 0; r < world_size; ++r) {
        MNE_CHECK_ARG(dirty_ptrs[r] && inbox_ptrs[r] && (((uintptr_t)inbox_ptrs[r]) & 15) == 0,
                      "mne_dirty_push_data: flag arrays and inboxes must be non-null, inboxes 16-byte aligned");
        d.dst[r] = dirty_ptrs[r];
        d.inbox[r] = reinterpret_cast<float4*>(inbox_ptrs[r]);
    }
    d.push = 1;
    d.world = world_size; d.rank = rank; d.chunk_tx_log2 = 31 - __builtin_clz((unsigned)chunk) - 5;
    d.n32 = n_total >> 5; d.local32 = d.n32 / world_size; d.scan32 = n_scan >> 5;
    int64_t blocks = (d.scan32 * 8 + 255) / 256;
    const int64_t cap = (int64_t)sm_count() * 8;
    if (blocks > cap) blocks = cap;
    dirty_push_data_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(local_map, reinterpret_cast<float4*>(grad), d);
    MNE_CHECK_CUDA(cudaGetLastError());
    return 0;
}

// ---------------------------------------------------------------------------------------------------------
// Small all-reduce over peer memory (a few dozen bytes to a few tens of KB: the loss accumulators, the decoder's weight
// gradients): every rank stores its vector into slot [rank] of every peer's exchange buffer (mne_peer_scatter), a cross-rank
// barrier follows (the caller's), then every rank sums the slots it received (mne_peer_sum).  An NCCL all-reduce of this
// size costs ~50 us at eight ranks and serialises with every other collective of the communicator; this is two tiny
// kernels and a signal-pad barrier on whatever stream needs the result.
// ---------------------------------------------------------------------------------------------------------
namespace mne {
struct PeerPtrs { float4* p[MNE_MAX_PEERS]; };

__global__ void peer_scatter_kernel(const float4* __restrict__ src, int n4, const __grid_constant__ PeerPtrs dst, int64_t off4) {
    float4* out = dst.p[blockIdx.y] + off4;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += gridDim.x * blockDim.x) out[i] = src[i];
}

template <typename T>
__global__ void peer_sum_kernel(const T* __restrict__ slots, int n, int world, T* __restrict__ out) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        T s = 0;
        for (int r = 0; r < world; ++r) s += slots[(int64_t)r * n + i];      // rank order: identical sums on every rank
        out[i] = s;
    }
}
}  // namespace mne

extern "C" int mne_peer_scatter(const float* src, int32_t n_floats, float* const* dst_ptrs, int64_t offset_floats, int32_t world_size,
                                mne_stream_t stream) {
    using namespace mne;
    MNE_CHECK_ARG(src && dst_ptrs && world_size >= 1 && world_size <= MNE_MAX_PEERS, "mne_peer_scatter: bad argument");
    MNE_CHECK_ARG(n_floats > 0 && (n_floats & 3) == 0 && (offset_floats & 3) == 0 && (((uintptr_t)src) & 15) == 0,
                  "mne_peer_scatter: sizes and offsets must be multiples of 4 floats, src 16-byte aligned");
    PeerPtrs d{};
    for (int r = 0; r < world_size; ++r) {
        MNE_CHECK_ARG(dst_ptrs[r] && (((uintptr_t)dst_ptrs[r]) & 15) == 0, "mne_peer_scatter: destination buffers must be non-null and 16-byte aligned");
        d.p[r] = reinterpret_cast<float4*>(dst_ptrs[r]);
    }
    const int n4 = n_floats >> 2;
    dim3 grid((unsigned)std::min(8, (n4 + 255) / 256), (unsigned)world_size);
    peer_scatter_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(reinterpret_cast<const float4*>(src), n4, d, offset_floats >> 2);
    MNE_CHECK_CUDA(cudaGetLastError());
    return 0;
}

extern "C" int mne_peer_sum(const void* slots, int32_t n, int32_t world_size, void* out, int32_t is_double, mne_stream_t stream) {
    using namespace mne;
    MNE_CHECK_ARG(slots && out && n > 0 && world_size >= 1, "mne_peer_sum: bad argument");
    const int blocks = std::min(16, (n + 255) / 256);
    if (is_double) peer_sum_kernel<double><<<blocks, 256, 0, (cudaStream_t)stream>>>(reinterpret_cast<const double*>(slots), n, world_size, reinterpret_cast<double*>(out));
    else peer_sum_kernel<float><<<blocks, 256, 0, (cudaStream_t)stream>>>(reinterpret_cast<const float*>(slots), n, world_size, reinterpret_cast<float*>(out));
    MNE_CHECK_CUDA(cudaGetLastError());
    return 0;
}

// ---------------------------------------------------------------------------------------------------------
// Cross-rank barrier over peer memory, one tiny kernel: lane r stores this rank's epoch of the channel into slot [rank] of
// rank r's signal pad (release, system scope: everything earlier kernels of this stream wrote -- to any GPU -- is visible
// before the signal is) and then polls slot [r] of its OWN pad until rank r's epoch has arrived (acquire).  Epochs count the
// barriers of a channel and live in device memory, so a CUDA-graph replay continues the sequence.  Stores and local polls
// only: no remote atomics, no remote loads (each costs an NVLink round trip).  The wait is bounded: after 2^29 polls (about
// two minutes; option peer_barrier_polls_log2) the kernel gives up and raises the pad's error word instead of hanging the GPU.
// ---------------------------------------------------------------------------------------------------------
namespace mne {
struct PeerBarrier {
    uint32_t* pad[MNE_MAX_PEERS];     // every rank's pad: [channels][MNE_MAX_PEERS] arrival words, then [channels] epochs, then 1 error word
    int world, rank, channel, channels;
    unsigned max_polls;
};

__global__ void peer_barrier_kernel(const __grid_constant__ PeerBarrier b) {
    __shared__ uint32_t epoch_s;
    uint32_t* mine = b.pad[b.rank];
    if (threadIdx.x == 0) {
        uint32_t* e = mine + b.channels * MNE_MAX_PEERS + b.channel;
        epoch_s = *e + 1;
        *e = epoch_s;
    }
    __syncthreads();
    const uint32_t epoch = epoch_s;
    const int r = threadIdx.x;
    if (r < b.world) {
        uint32_t* there = b.pad[r] + b.channel * MNE_MAX_PEERS + b.rank;
        asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(there), "r"(epoch) : "memory");
        const uint32_t* here = mine + b.channel * MNE_MAX_PEERS + r;
        uint32_t seen;
        unsigned polls = 0;
        do {
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(seen) : "l"(here) : "memory");
            if ((int32_t)(seen - epoch) >= 0) break;
        } while (++polls < b.max_polls);
        if ((int32_t)(seen - epoch) < 0) mine[b.channels * (MNE_MAX_PEERS + 1)] = 1u;       // timed out
    }
}
}  // namespace mne

extern "C" int mne_peer_barrier(uint32_t* const* pad_ptrs, int32_t world_size, int32_t rank, int32_t channel, int32_t n_channels,
                                mne_stream_t stream) {
    using namespace mne;
    MNE_CHECK_ARG(pad_ptrs && world_size >= 1 && world_size <= MNE_MAX_PEERS && rank >= 0 && rank < world_size,
                  "mne_peer_barrier: bad world / rank");
    MNE_CHECK_ARG(channel >= 0 && channel < n_channels && n_channels <= 64, "mne_peer_barrier: bad channel %d of %d", channel, n_channels);
    PeerBarrier b{};
    for (int r = 0; r < world_size; ++r) {
        MNE_CHECK_ARG(pad_ptrs[r] && (((uintptr_t)pad_ptrs[r]) & 3) == 0, "mne_peer_barrier: null or misaligned pad");
        b.pad[r] = pad_ptrs[r];
    }
    b.world = world_size; b.rank = rank; b.channel = channel; b.channels = n_channels;
    b.max_polls = 1u << peer_barrier_polls_log2();
    peer_barrier_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(b);
    MNE_CHECK_CUDA(cudaGetLastError());
    return 0;
}
