// Measurement aid (no counterpart in the reference): the gather rate this GPU sustains for the ACCESS PATTERN of the encoder,
// with nothing else in the kernel -- the denominator of the "fraction of the L2 / HBM gather roofline" that bench.py reports
// beside the HBM-copy-denominated one.  A forward pass serves most taps out of L2 (neighbouring samples share texels), so the
// honest ceiling is what a kernel that ONLY gathers the same addresses reaches, not the DRAM copy bandwidth.
//   mode 0: 128-byte texels (tri-planes): lane = (point of a quad, 4-channel group), one 128-bit load per lane and tap --
//           offsets[i] = float offset of a texel (a multiple of 32), 24 per sample in the forward's order
//   mode 1: 8-byte entries (hash grid): one 64-bit load per lane -- offsets[i] = float offset of an entry (a multiple of 2)
#include "common.cuh"

namespace mne {

__global__ void __launch_bounds__(256) gather_bench_kernel(const float* __restrict__ table, const uint32_t* __restrict__ offsets,
                                                           int64_t n_offsets, int mode, float* __restrict__ sink) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5, n_warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    if (mode == 0) {
        // a warp instruction covers 4 texels (8 lanes x 16 B each); 12 instructions in flight per lane, as the forward has
        const int pp = lane >> 3, q = lane & 7;
        for (int64_t base = warp * 48; base < n_offsets; base += n_warps * 48) {
            float4 v[12];
#pragma unroll
            for (int k = 0; k < 12; ++k) {
                const int64_t i = base + k * 4 + pp;
                v[k] = i < n_offsets ? __ldg(reinterpret_cast<const float4*>(table + offsets[i]) + q) : make_float4(0.f, 0.f, 0.f, 0.f);
            }
#pragma unroll
            for (int k = 0; k < 12; ++k) { acc.x += v[k].x; acc.y += v[k].y; acc.z += v[k].z; acc.w += v[k].w; }
        }
    } else {
        for (int64_t base = warp * 256; base < n_offsets; base += n_warps * 256) {
            float2 v[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                const int64_t i = base + k * 32 + lane;
                v[k] = i < n_offsets ? __ldg(reinterpret_cast<const float2*>(table + offsets[i])) : make_float2(0.f, 0.f);
            }
#pragma unroll
            for (int k = 0; k < 8; ++k) { acc.x += v[k].x; acc.y += v[k].y; }
        }
    }
    if (acc.x + acc.y + acc.z + acc.w == 123456.789f) sink[0] = acc.x;       // keeps the loads alive
}

int sm_count();

}  // namespace mne

extern "C" int mne_bench_gather(const float* table, const uint32_t* offsets, int64_t n_offsets, int32_t mode, float* sink,
                                mne_stream_t stream) {
    using namespace mne;
    MNE_CHECK_ARG(table && offsets && sink && n_offsets > 0 && (mode == 0 || mode == 1), "mne_bench_gather: bad argument");
    MNE_CHECK_ARG((((uintptr_t)table) & 15) == 0, "mne_bench_gather: table must be 16-byte aligned");
    gather_bench_kernel<<<sm_count() * 8, 256, 0, (cudaStream_t)stream>>>(table, offsets, n_offsets, mode, sink);
    MNE_CHECK_CUDA(cudaGetLastError());
    return 0;
}
