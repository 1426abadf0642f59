// Fused backward: loss / map gradients -> compositing backward -> decoder backward (activations
// recomputed from a re-gather, nothing but `raw` is carried over from the forward) -> weight gradients,
// tri-plane gradient scatter and ray (pose) gradients, one kernel.  Replaces torch autograd through
// reference model/scene_rep.py:183-230,273-317,384-386,549-611 (loss.backward() at mp_slam/mapper.py:88,159
// and :405).
#include <mutex>

#include "render_common.cuh"

namespace mne {

struct BwdParams {
    MneSceneDesc sc;
    MneRenderCfg rc;
    int n_rays, n_samples, rays_per_group;
    const float* rays_o;
    const float* rays_d;
    const float* z_vals;
    const float* raw;
    const float* target_rgb;
    const float* target_d;
    const double* acc;         // mask counts (slots 0..5)
    const float* loss_grads;   // [8] upstream of the loss scalars, or NULL
    const float* g_rgb;        // explicit upstream of the rendered maps, or NULL
    const float* g_depth;
    const float* g_raw;        // [N*S*4]
    float4* draw;              // scratch [N*S]: d loss / d raw
    float* g_rays_o;
    float* g_rays_d;
};

template <bool TWO> __host__ __device__ constexpr int bwd_warps() { return TWO ? 4 : 6; }
template <bool TWO>
__host__ __device__ constexpr int bwd_warp_floats() {
    return 32 * kInStride + (TWO ? 32 * kCfStride : 0) + 2 * 32 * kHStride + 2 * 32 * kGStride + 32 * 4;
}
template <bool TWO>
constexpr size_t bwd_smem_bytes() {
    return sizeof(float) * (size_t)(2 * Net<TWO>::W_FLOATS + bwd_warps<TWO>() * bwd_warp_floats<TWO>());
}

// lane = neuron: acc[k] += sum_p col[p] * row_p[k] for K4*4 consecutive k, then flush into the shared
// weight-gradient block at DW[(k0 + k) * ldw + lane]
template <int K4>
__device__ __forceinline__ void outer_accumulate(const float* __restrict__ col, int col_stride, const float* __restrict__ rows,
                                                 int row_stride, int np, float* __restrict__ DW, int ldw, int lane) {
    float acc[K4 * 4];
#pragma unroll
    for (int k = 0; k < K4 * 4; ++k) acc[k] = 0.0f;
    for (int p = 0; p < np; ++p) {
        const float c = col[p * col_stride + lane];
        const float4* r = reinterpret_cast<const float4*>(rows + p * row_stride);
#pragma unroll
        for (int k4 = 0; k4 < K4; ++k4) {
            const float4 x = r[k4];
            acc[4 * k4 + 0] = fmaf(c, x.x, acc[4 * k4 + 0]);
            acc[4 * k4 + 1] = fmaf(c, x.y, acc[4 * k4 + 1]);
            acc[4 * k4 + 2] = fmaf(c, x.z, acc[4 * k4 + 2]);
            acc[4 * k4 + 3] = fmaf(c, x.w, acc[4 * k4 + 3]);
        }
    }
#pragma unroll
    for (int k = 0; k < K4 * 4; ++k) atomicAdd(&DW[k * ldw + lane], acc[k]);
}

template <bool TWO, bool RAY_GRAD, bool PARAM_GRAD>
__global__ void __launch_bounds__(bwd_warps<TWO>() * 32, 1) render_bwd_kernel(const __grid_constant__ BwdParams P) {
    using N = Net<TWO>;
    constexpr int NW = bwd_warps<TWO>();
    extern __shared__ __align__(16) float smem[];
    float* WT = smem;
    float* DW = smem + N::W_FLOATS;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float* IN = smem + 2 * N::W_FLOATS + warp * bwd_warp_floats<TWO>();
    float* CF = IN + 32 * kInStride;
    float* H1 = CF + (TWO ? 32 * kCfStride : 0);
    float* HC = H1 + 32 * kHStride;
    float* GEO = HC + 32 * kHStride;
    float* DOUT = GEO + 32 * kGStride;
    float* DRGB = DOUT + 32 * kGStride;
    int4* TAB = reinterpret_cast<int4*>(H1);     // aliases H1/HC while those are dead (gather and scatter phases)

    load_weights_transposed<TWO>(P.sc, WT, threadIdx.x, blockDim.x);
    if (PARAM_GRAD)
        for (int i = threadIdx.x; i < N::W_FLOATS; i += blockDim.x) DW[i] = 0.0f;
    __syncthreads();

    const int S = P.n_samples, rpg = P.rays_per_group;
    const int n_groups = (P.n_rays + rpg - 1) / rpg;
    const int total_warps = gridDim.x * NW;

    // upstream scale factors of the loss terms (JointEncoding.forward + get_sdf_loss + sdf_losses)
    const float* lg = P.loss_grads;
    float k_rgb = 0.f, k_depth = 0.f, k_co_fs = 0.f, k_co_sdf = 0.f, k_e_fs = 0.f, k_e_center = 0.f, k_e_tail = 0.f;
    if (lg) {
        const double Nn = P.rc.norm_rays, NS = P.rc.norm_rays * (double)S;
        auto inv = [](double c) { return c > 0.0 ? 1.0 / c : 0.0; };   // empty mask: the term has no gradient
        k_rgb = (float)(2.0 * lg[L_RGB] / (3.0 * Nn));
        k_depth = (float)(2.0 * lg[L_DEPTH] * inv(P.acc[A_N_VALID]));
        if (!P.rc.first_pass) {
            const float n_fs = (float)P.acc[A_N_CO_FRONT], n_sdf = (float)P.acc[A_N_CO_NEAR];
            const float fs_w = 1.0f - n_fs / (n_fs + n_sdf), sdf_w = 1.0f - n_sdf / (n_fs + n_sdf);
            k_co_fs = (n_fs + n_sdf > 0.f) ? (float)(2.0 * lg[L_CO_FS] * fs_w / NS) : 0.f;
            k_co_sdf = (n_fs + n_sdf > 0.f) ? (float)(2.0 * lg[L_CO_SDF] * sdf_w / NS) * P.rc.sc_trunc : 0.f;
            k_e_fs = (float)(2.0 * lg[L_E_FS] * inv(P.acc[A_N_E_FRONT]));
            k_e_center = (float)(2.0 * lg[L_E_CENTER] * inv(P.acc[A_N_E_CENTER])) * P.rc.model_trunc;
            k_e_tail = (float)(2.0 * lg[L_E_TAIL] * inv(P.acc[A_N_E_TAIL])) * P.rc.model_trunc;
        }
    }

    // block-minor order: consecutive groups land on different SMs, so a grid with spare warps stays balanced
    for (int g = warp * gridDim.x + blockIdx.x; g < n_groups; g += total_warps) {
        const int ray0 = g * rpg;
        const int nr = min(rpg, P.n_rays - ray0);
        const int npts = nr * S;
        const size_t base = (size_t)ray0 * S;

        // ---------------- compositing backward: d loss / d raw for every sample of the group ----------------
        for (int r = 0; r < nr; ++r) {
            const int ray = ray0 + r;
            const size_t rb = (size_t)ray * S;
            const RayComposite c = composite_ray(P.raw, P.z_vals, rb, S, P.rc, lane);
            const float d = P.target_d ? P.target_d[ray] : 0.0f;
            float rr[3] = {c.rgb[0], c.rgb[1], c.rgb[2]};
            if (P.rc.white_bkgd) { rr[0] += 1.f - c.acc; rr[1] += 1.f - c.acc; rr[2] += 1.f - c.acc; }
            float G[3] = {0.f, 0.f, 0.f}, Gd = 0.f;
            if (lg && P.target_rgb) {
#pragma unroll
                for (int k = 0; k < 3; ++k) G[k] = k_rgb * (rr[k] - P.target_rgb[ray * 3 + k]);
                if (depth_valid(d, P.rc)) Gd = k_depth * (c.depth - d);
            }
            if (P.g_rgb) { G[0] += P.g_rgb[ray * 3 + 0]; G[1] += P.g_rgb[ray * 3 + 1]; G[2] += P.g_rgb[ray * 3 + 2]; }
            if (P.g_depth) Gd += P.g_depth[ray];
            for (int s = lane; s < S; s += 32) {
                const float z = __ldg(P.z_vals + rb + s);
                const float4 v = __ldg(reinterpret_cast<const float4*>(P.raw) + rb + s);
                float4 o = make_float4(0.f, 0.f, 0.f, 0.f);
                if (z < c.thresh) {
                    float sp;
                    const float u = sdf_bell(v.w, P.rc.trunc, &sp);
                    const float w = u * c.inv_norm;
                    const float c0 = sigmoidf(v.x), c1 = sigmoidf(v.y), c2 = sigmoidf(v.z);
                    const float dLdu = c.inv_norm * (G[0] * (c0 - rr[0]) + G[1] * (c1 - rr[1]) + G[2] * (c2 - rr[2]) + Gd * (z - c.depth));
                    o.w = dLdu * u * (1.0f - 2.0f * sp) / P.rc.trunc;
                    o.x = G[0] * w * c0 * (1.0f - c0);
                    o.y = G[1] * w * c1 * (1.0f - c1);
                    o.z = G[2] * w * c2 * (1.0f - c2);
                }
                if (lg && P.target_d && !P.rc.first_pass) {
                    const unsigned m = sample_masks(z, d, P.rc);
                    if (m & M_CO_FRONT) o.w += k_co_fs * (v.w - 1.0f);
                    if (m & M_CO_NEAR) o.w += k_co_sdf * (__fadd_rn(z, __fmul_rn(v.w, P.rc.sc_trunc)) - d);
                    if (m & M_E_FRONT) o.w += k_e_fs * (v.w - 1.0f);
                    const float re = __fadd_rn(z, __fmul_rn(v.w, P.rc.model_trunc)) - d;
                    if (m & M_E_CENTER) o.w += k_e_center * re;
                    if (m & M_E_TAIL) o.w += k_e_tail * re;
                }
                if (P.g_raw) {
                    const float4 e = __ldg(reinterpret_cast<const float4*>(P.g_raw) + rb + s);
                    o.x += e.x; o.y += e.y; o.z += e.z; o.w += e.w;
                }
                P.draw[rb + s] = o;
            }
        }
        __syncwarp();

        // ---------------- per-chunk decoder backward ----------------
        for (int c0 = 0; c0 < npts; c0 += 32) {
            const int q = c0 + lane;
            const bool active = q < npts;
            const int np = min(32, npts - c0);
            const int ray = ray0 + (active ? q / S : 0);
            float px = 0.f, py = 0.f, pz = 0.f, zq = 0.f;
            float4 dr = make_float4(0.f, 0.f, 0.f, 0.f);
            if (active) {
                zq = P.z_vals[base + q];
                px = __fadd_rn(P.rays_o[ray * 3 + 0], __fmul_rn(P.rays_d[ray * 3 + 0], zq));
                py = __fadd_rn(P.rays_o[ray * 3 + 1], __fmul_rn(P.rays_d[ray * 3 + 1], zq));
                pz = __fadd_rn(P.rays_o[ray * 3 + 2], __fmul_rn(P.rays_d[ray * 3 + 2], zq));
                dr = __ldcg(P.draw + base + q);
            }
            // P0 (lane = point): taps, OneBlob, upstream of this point
            write_taps<TWO>(P.sc, TAB + lane * N::N_TAPS, px, py, pz);
            float* in_row = IN + lane * kInStride;
            const float bx = blob_input(px, P.sc.bbox_lo[0], P.sc.bbox_ext[0]);
            const float by = blob_input(py, P.sc.bbox_lo[1], P.sc.bbox_ext[1]);
            const float bz = blob_input(pz, P.sc.bbox_lo[2], P.sc.bbox_ext[2]);
            oneblob16(bx, in_row + N::F);
            oneblob16(by, in_row + N::F + 16);
            oneblob16(bz, in_row + N::F + 32);
            *reinterpret_cast<float4*>(DRGB + lane * 4) = dr;
            __syncwarp();
            // P1 (lane = channel): gather
            gather_chunk<TWO>(P.sc, TAB, IN, CF, np, lane);
            __syncwarp();
            // P3 (lane = point): forward recompute
            unsigned m1 = 0, mc = 0;
            {
                float h1[32], out16[16], hc[32], rgb[3];
                decoder_forward<TWO>(WT, in_row, CF + lane * kCfStride, h1, out16, hc, rgb);
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    m1 |= (h1[j] > 0.0f) ? (1u << j) : 0u;
                    mc |= (hc[j] > 0.0f) ? (1u << j) : 0u;
                }
                if (PARAM_GRAD) {
#pragma unroll
                    for (int j4 = 0; j4 < 8; ++j4) {
                        *reinterpret_cast<float4*>(H1 + lane * kHStride + 4 * j4) = make_float4(h1[4 * j4], h1[4 * j4 + 1], h1[4 * j4 + 2], h1[4 * j4 + 3]);
                        *reinterpret_cast<float4*>(HC + lane * kHStride + 4 * j4) = make_float4(hc[4 * j4], hc[4 * j4 + 1], hc[4 * j4 + 2], hc[4 * j4 + 3]);
                    }
#pragma unroll
                    for (int c4 = 0; c4 < 4; ++c4)
                        *reinterpret_cast<float4*>(GEO + lane * kGStride + 4 * c4) = make_float4(out16[4 * c4], out16[4 * c4 + 1], out16[4 * c4 + 2], out16[4 * c4 + 3]);
                }
            }
            __syncwarp();
            // P4 (lane = neuron): d col1 = d_rgb (x) hc
            if (PARAM_GRAD) {
                float a0 = 0.f, a1 = 0.f, a2 = 0.f;
                for (int p = 0; p < np; ++p) {
                    const float4 g4 = *reinterpret_cast<const float4*>(DRGB + p * 4);
                    const float h = HC[p * kHStride + lane];
                    a0 = fmaf(g4.x, h, a0); a1 = fmaf(g4.y, h, a1); a2 = fmaf(g4.z, h, a2);
                }
                atomicAdd(&DW[N::OFF_V2T + lane * 4 + 0], a0);
                atomicAdd(&DW[N::OFF_V2T + lane * 4 + 1], a1);
                atomicAdd(&DW[N::OFF_V2T + lane * 4 + 2], a2);
            }
            __syncwarp();
            // P5 (lane = point): d hc, d geo -> HC tile becomes d_hc, DOUT = [d_geo | d_sdf]
            {
                float dhc[32];
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    const float4 w = *reinterpret_cast<const float4*>(WT + N::OFF_V2T + j * 4);
                    const float v = w.x * dr.x + w.y * dr.y + w.z * dr.z;
                    dhc[j] = ((mc >> j) & 1u) ? v : 0.0f;
                }
#pragma unroll
                for (int j4 = 0; j4 < 8; ++j4)
                    *reinterpret_cast<float4*>(HC + lane * kHStride + 4 * j4) = make_float4(dhc[4 * j4], dhc[4 * j4 + 1], dhc[4 * j4 + 2], dhc[4 * j4 + 3]);
                float dout[16];
#pragma unroll
                for (int m = 0; m < kGeo; ++m) dout[m] = dot32_row(WT + N::OFF_V1T + (N::GEO_ROW0 + m) * 32, dhc);
                dout[15] = dr.w;
#pragma unroll
                for (int c4 = 0; c4 < 4; ++c4)
                    *reinterpret_cast<float4*>(DOUT + lane * kGStride + 4 * c4) = make_float4(dout[4 * c4], dout[4 * c4 + 1], dout[4 * c4 + 2], dout[4 * c4 + 3]);
            }
            __syncwarp();
            // P6 (lane = neuron): d col0 = d_hc (x) [blob | cfeat | geo], d sdf1 = d_out (x) h1
            if (PARAM_GRAD) {
                outer_accumulate<kBlob / 4>(HC, kHStride, IN + N::F, kInStride, np, DW + N::OFF_V1T, 32, lane);
                if (TWO) outer_accumulate<N::F / 4>(HC, kHStride, CF, kCfStride, np, DW + N::OFF_V1T + kBlob * 32, 32, lane);
                outer_accumulate<4>(HC, kHStride, GEO, kGStride, np, DW + N::OFF_V1T + N::GEO_ROW0 * 32, 32, lane);
                // d W2T[j][c] += sum_p h1[p][j] * d_out[p][c]; lane = j
                {
                    float acc[16];
#pragma unroll
                    for (int c = 0; c < 16; ++c) acc[c] = 0.0f;
                    for (int p = 0; p < np; ++p) {
                        const float h = H1[p * kHStride + lane];
                        const float4* r = reinterpret_cast<const float4*>(DOUT + p * kGStride);
#pragma unroll
                        for (int c4 = 0; c4 < 4; ++c4) {
                            const float4 x = r[c4];
                            acc[4 * c4 + 0] = fmaf(h, x.x, acc[4 * c4 + 0]);
                            acc[4 * c4 + 1] = fmaf(h, x.y, acc[4 * c4 + 1]);
                            acc[4 * c4 + 2] = fmaf(h, x.z, acc[4 * c4 + 2]);
                            acc[4 * c4 + 3] = fmaf(h, x.w, acc[4 * c4 + 3]);
                        }
                    }
#pragma unroll
                    for (int c = 0; c < 16; ++c) atomicAdd(&DW[N::OFF_W2T + lane * 16 + c], acc[c]);
                }
            }
            __syncwarp();
            // P7 (lane = point): d h1 -> H1 tile becomes d_h1
            float dh1[32];
            {
                float dout[16];
#pragma unroll
                for (int c4 = 0; c4 < 4; ++c4) {
                    const float4 x = *reinterpret_cast<const float4*>(DOUT + lane * kGStride + 4 * c4);
                    dout[4 * c4] = x.x; dout[4 * c4 + 1] = x.y; dout[4 * c4 + 2] = x.z; dout[4 * c4 + 3] = x.w;
                }
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    const float4* w = reinterpret_cast<const float4*>(WT + N::OFF_W2T + j * 16);
                    float a0 = 0.f, a1 = 0.f;
#pragma unroll
                    for (int c4 = 0; c4 < 4; ++c4) {
                        const float4 ww = w[c4];
                        a0 = fmaf(ww.x, dout[4 * c4 + 0], a0); a1 = fmaf(ww.y, dout[4 * c4 + 1], a1);
                        a0 = fmaf(ww.z, dout[4 * c4 + 2], a0); a1 = fmaf(ww.w, dout[4 * c4 + 3], a1);
                    }
                    dh1[j] = ((m1 >> j) & 1u) ? (a0 + a1) : 0.0f;
                }
                if (PARAM_GRAD) {
#pragma unroll
                    for (int j4 = 0; j4 < 8; ++j4)
                        *reinterpret_cast<float4*>(H1 + lane * kHStride + 4 * j4) = make_float4(dh1[4 * j4], dh1[4 * j4 + 1], dh1[4 * j4 + 2], dh1[4 * j4 + 3]);
                }
            }
            __syncwarp();
            // P8 (lane = neuron): d sdf0 = d_h1 (x) [feat | blob]
            if (PARAM_GRAD) {
                outer_accumulate<N::F / 4>(H1, kHStride, IN, kInStride, np, DW + N::OFF_W1T, 32, lane);
                outer_accumulate<kBlob / 4>(H1, kHStride, IN + N::F, kInStride, np, DW + N::OFF_W1T + N::F * 32, 32, lane);
            }
            __syncwarp();
            // P9 (lane = point): d feat -> IN tile (and CF), OneBlob input gradient
            float dpt[3] = {0.f, 0.f, 0.f};      // d loss / d point, OneBlob part
            {
                float dhc[32];
                if (TWO || RAY_GRAD) {
#pragma unroll
                    for (int j4 = 0; j4 < 8; ++j4) {
                        const float4 x = *reinterpret_cast<const float4*>(HC + lane * kHStride + 4 * j4);
                        dhc[4 * j4] = x.x; dhc[4 * j4 + 1] = x.y; dhc[4 * j4 + 2] = x.z; dhc[4 * j4 + 3] = x.w;
                    }
                }
                if (RAY_GRAD) {
                    const float bxyz[3] = {bx, by, bz};
#pragma unroll
                    for (int a = 0; a < 3; ++a) {
                        float gb[16];
#pragma unroll
                        for (int b = 0; b < 16; ++b)
                            gb[b] = dot32_row(WT + N::OFF_W1T + (N::F + a * 16 + b) * 32, dh1) +
                                    dot32_row(WT + N::OFF_V1T + (a * 16 + b) * 32, dhc);
                        dpt[a] = (float)((double)oneblob16_grad(bxyz[a], gb) / P.sc.bbox_ext[a]);
                    }
                }
                if (PARAM_GRAD || RAY_GRAD) {
#pragma unroll 4
                    for (int k4 = 0; k4 < N::F / 4; ++k4) {
                        float4 o;
                        o.x = dot32_row(WT + N::OFF_W1T + (4 * k4 + 0) * 32, dh1);
                        o.y = dot32_row(WT + N::OFF_W1T + (4 * k4 + 1) * 32, dh1);
                        o.z = dot32_row(WT + N::OFF_W1T + (4 * k4 + 2) * 32, dh1);
                        o.w = dot32_row(WT + N::OFF_W1T + (4 * k4 + 3) * 32, dh1);
                        *reinterpret_cast<float4*>(in_row + 4 * k4) = o;
                    }
                    if (TWO) {
#pragma unroll 4
                        for (int k4 = 0; k4 < N::F / 4; ++k4) {
                            float4 o;
                            o.x = dot32_row(WT + N::OFF_V1T + (kBlob + 4 * k4 + 0) * 32, dhc);
                            o.y = dot32_row(WT + N::OFF_V1T + (kBlob + 4 * k4 + 1) * 32, dhc);
                            o.z = dot32_row(WT + N::OFF_V1T + (kBlob + 4 * k4 + 2) * 32, dhc);
                            o.w = dot32_row(WT + N::OFF_V1T + (kBlob + 4 * k4 + 3) * 32, dhc);
                            *reinterpret_cast<float4*>(CF + lane * kCfStride + 4 * k4) = o;
                        }
                    }
                }
            }
            __syncwarp();      // every lane is done with its H1 / HC row: the TAB alias may be rewritten
            write_taps<TWO>(P.sc, TAB + lane * N::N_TAPS, px, py, pz);
            __syncwarp();
            // P10 (lane = channel): scatter plane gradients; ray gradients through the bilinear taps
            float dn_keep[3] = {0.f, 0.f, 0.f};   // d loss / d normalised plane coords of this lane's point
            for (int p = 0; p < np; ++p) {
                float dn[3] = {0.f, 0.f, 0.f};
#pragma unroll
                for (int grp = 0; grp < (TWO ? 2 : 1); ++grp) {
#pragma unroll
                    for (int l = 0; l < 2; ++l) {
                        const float gch = (grp == 0) ? IN[p * kInStride + l * 32 + lane] : CF[p * kCfStride + l * 32 + lane];
#pragma unroll
                        for (int a = 0; a < 3; ++a) {
                            const int t = grp * 6 + a * 2 + l;
                            const MnePlaneDesc& Pl = P.sc.planes[t];
                            const int4 tp = TAB[p * N::N_TAPS + t];
                            const int64_t off = (int64_t)lane * Pl.stride_c + tp.x;
                            const int64_t dx = (tp.w & 1) ? Pl.stride_w : 0, dy = (tp.w & 2) ? Pl.stride_h : 0;
                            const float fx = __int_as_float(tp.y), fy = __int_as_float(tp.z);
                            const float wx1 = (tp.w & 1) ? fx : 0.0f, wy1 = (tp.w & 2) ? fy : 0.0f;
                            const float wx0 = 1.0f - fx, wy0 = 1.0f - fy;
                            if (PARAM_GRAD && Pl.grad) {
                                float* gb = Pl.grad + off;
                                atomicAdd(gb, gch * (wy0 * wx0));
                                if (tp.w & 1) atomicAdd(gb + dx, gch * (wy0 * wx1));
                                if (tp.w & 2) atomicAdd(gb + dy, gch * (wy1 * wx0));
                                if ((tp.w & 3) == 3) atomicAdd(gb + dy + dx, gch * (wy1 * wx1));
                            }
                            if (RAY_GRAD && (tp.w & 12)) {
                                const float* vb = Pl.data + off;
                                const float t00 = __ldg(vb), t01 = __ldg(vb + dx), t10 = __ldg(vb + dy), t11 = __ldg(vb + dy + dx);
                                const int ax = (a == 2) ? 1 : 0, ay = (a == 0) ? 1 : 2;
                                // d value / d fx and / d fy, chained to the normalised coordinate: u = (g+1)/2*(size-1)
                                if (tp.w & 4) dn[ax] += gch * ((t01 - t00) * wy0 + (t11 - t10) * wy1) * (0.5f * (float)(Pl.W - 1));
                                if (tp.w & 8) dn[ay] += gch * ((t10 - t00) * wx0 + (t11 - t01) * wx1) * (0.5f * (float)(Pl.H - 1));
                            }
                        }
                    }
                }
                if (RAY_GRAD) {
                    dn[0] = warp_sum(dn[0]); dn[1] = warp_sum(dn[1]); dn[2] = warp_sum(dn[2]);
                    if (lane == p) { dn_keep[0] = dn[0]; dn_keep[1] = dn[1]; dn_keep[2] = dn[2]; }
                }
            }
            if (RAY_GRAD && active) {
#pragma unroll
                for (int a = 0; a < 3; ++a) {
                    const float gp = dn_keep[a] * (2.0f / P.sc.bound_ext[a]) + dpt[a];
                    atomicAdd(P.g_rays_o + ray * 3 + a, gp);
                    atomicAdd(P.g_rays_d + ray * 3 + a, gp * zq);
                }
            }
            __syncwarp();
        }
    }

    if (PARAM_GRAD) {
        __syncthreads();
        if (P.sc.g_sdf0)
            for (int i = threadIdx.x; i < N::IN_SDF * 32; i += blockDim.x)
                atomicAdd(P.sc.g_sdf0 + (i & 31) * N::IN_SDF + (i >> 5), DW[N::OFF_W1T + i]);
        if (P.sc.g_sdf1)
            for (int i = threadIdx.x; i < 32 * 16; i += blockDim.x) {
                const int j = i >> 4, c = i & 15;
                atomicAdd(P.sc.g_sdf1 + ((c == 15) ? 0 : c + 1) * 32 + j, DW[N::OFF_W2T + i]);
            }
        if (P.sc.g_col0)
            for (int i = threadIdx.x; i < N::IN_COL * 32; i += blockDim.x)
                atomicAdd(P.sc.g_col0 + (i & 31) * N::IN_COL + (i >> 5), DW[N::OFF_V1T + i]);
        if (P.sc.g_col1)
            for (int i = threadIdx.x; i < 32 * 4; i += blockDim.x)
                if ((i & 3) < 3) atomicAdd(P.sc.g_col1 + (i & 3) * 32 + (i >> 2), DW[N::OFF_V2T + i]);
    }
}

int check_scene(const MneSceneDesc* sc, const char* who, bool need_grads);
int choose_rays_per_group(int n_rays, int n_samples, int total_warps);
int sm_count();
int impl_version();
bool scene_is_vectorizable(const MneSceneDesc* sc);
int render_backward_v2(const MneSceneDesc*, const MneRenderCfg*, int32_t, int32_t, const float*, const float*, const float*,
                       const float*, const float*, const float*, const double*, const float*, const float*, const float*,
                       const float*, void*, float*, float*, bool, float*, cudaStream_t);

template <bool TWO, bool RAY_GRAD, bool PARAM_GRAD>
static int launch_bwd(const BwdParams& P, cudaStream_t stream) {
    const size_t smem = bwd_smem_bytes<TWO>();
    static std::atomic<bool> configured[64] = {};      // per device; set outside any stream capture by the first (warm-up) call
    int dev = 0;
    MNE_CHECK_CUDA(cudaGetDevice(&dev));
    if (dev >= 64 || !configured[dev]) {
        MNE_CHECK_CUDA(cudaFuncSetAttribute(render_bwd_kernel<TWO, RAY_GRAD, PARAM_GRAD>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        if (dev < 64) configured[dev] = true;
    }
    const int n_groups = (P.n_rays + P.rays_per_group - 1) / P.rays_per_group;
    const int blocks = max(1, min(n_groups, sm_count()));
    render_bwd_kernel<TWO, RAY_GRAD, PARAM_GRAD><<<blocks, bwd_warps<TWO>() * 32, smem, stream>>>(P);
    MNE_CHECK_CUDA(cudaGetLastError());
    return 0;
}

template <bool TWO>
static int dispatch_bwd(const BwdParams& P, bool ray_grad, bool param_grad, cudaStream_t stream) {
    if (ray_grad && param_grad) return launch_bwd<TWO, true, true>(P, stream);
    if (ray_grad) return launch_bwd<TWO, true, false>(P, stream);
    return launch_bwd<TWO, false, true>(P, stream);
}

}  // namespace mne

extern "C" size_t mne_backward_scratch_bytes(int32_t n_rays, int32_t n_samples) {
    if (n_rays <= 0 || n_samples <= 0) return 0;
    // d loss / d raw per sample, the list of samples where it is non-zero, and the length of that list
    const size_t n = (size_t)n_rays * (size_t)n_samples;
    return n * sizeof(float4) + ((n + 3) & ~(size_t)3) * sizeof(int) + 16;
}

extern "C" int mne_render_backward(const MneSceneDesc* scene, const MneRenderCfg* rcfg, int32_t n_rays, int32_t n_samples,
                                   const float* rays_o, const float* rays_d, const float* z_vals, const float* raw,
                                   const float* target_rgb, const float* target_d, const double* acc,
                                   const float* loss_grads, const float* g_rgb, const float* g_depth, const float* g_raw,
                                   void* scratch, float* g_rays_o, float* g_rays_d, float* act, mne_stream_t stream) {
    using namespace mne;
    MNE_CHECK_ARG(scene && rcfg, "mne_render_backward: null scene/rcfg");
    MNE_CHECK_ARG(((uintptr_t)act & 31) == 0, "mne_render_backward: act must be 32-byte aligned");
    if (int e = check_scene(scene, "mne_render_backward", false)) return e;
    MNE_CHECK_ARG(n_rays >= 0 && n_samples > 0, "mne_render_backward: bad sizes n_rays=%d n_samples=%d", n_rays, n_samples);
    if (n_rays == 0) return 0;
    MNE_CHECK_ARG(rays_o && rays_d && z_vals && raw && scratch, "mne_render_backward: rays_o, rays_d, z_vals, raw and scratch are required");
    MNE_CHECK_ARG(((uintptr_t)scratch & 15) == 0 && ((uintptr_t)raw & 15) == 0, "mne_render_backward: raw and scratch must be 16-byte aligned");
    MNE_CHECK_ARG(((uintptr_t)g_raw & 15) == 0, "mne_render_backward: g_raw must be 16-byte aligned");
    MNE_CHECK_ARG((int64_t)n_rays * n_samples < (int64_t)1 << 31, "mne_render_backward: n_rays * n_samples must be < 2^31");
    MNE_CHECK_ARG(!loss_grads || (acc && target_rgb && target_d), "mne_render_backward: loss_grads needs acc, target_rgb and target_d");
    MNE_CHECK_ARG((g_rays_o == nullptr) == (g_rays_d == nullptr), "mne_render_backward: pass both g_rays_o and g_rays_d or neither");
    bool param_grad = scene->g_sdf0 || scene->g_sdf1 || scene->g_col0 || scene->g_col1;
    if (scene->encoder == 1) param_grad = param_grad || scene->hash[0].grad || (scene->two_grid && scene->hash[1].grad);
    else for (int i = 0; i < (scene->two_grid ? 12 : 6); ++i) param_grad = param_grad || scene->planes[i].grad;
    const bool ray_grad = g_rays_o != nullptr;
    if (!param_grad && !ray_grad) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    if (ray_grad) {
        MNE_CHECK_CUDA(cudaMemsetAsync(g_rays_o, 0, sizeof(float) * 3 * (size_t)n_rays, st));
        MNE_CHECK_CUDA(cudaMemsetAsync(g_rays_d, 0, sizeof(float) * 3 * (size_t)n_rays, st));
    }
    MNE_CHECK_ARG(!scene->dirty || scene->dirty_base, "mne_render_backward: dirty without dirty_base");
    // (two grids with ray AND parameter gradients fit the tensor-pipe kernel's shared memory only without the recompute weights)
    const bool strided = scene->encoder == 0 && !scene_is_vectorizable(scene);
    if (impl_version() == 2 && ((act && !strided) || !(scene->two_grid && ray_grad && param_grad)))
        return render_backward_v2(scene, rcfg, n_rays, n_samples, rays_o, rays_d, z_vals, raw, target_rgb, target_d, acc,
                                  loss_grads, g_rgb, g_depth, g_raw, scratch, g_rays_o, g_rays_d, param_grad, act, st);
    MNE_CHECK_ARG(!scene->dirty, "mne_render_backward: MneSceneDesc.dirty is supported by impl 2 without combined two-grid ray + parameter gradients only");
    BwdParams P{};
    P.sc = *scene; P.rc = *rcfg; P.n_rays = n_rays; P.n_samples = n_samples;
    P.rays_per_group = choose_rays_per_group(n_rays, n_samples, sm_count() * (scene->two_grid ? bwd_warps<true>() : bwd_warps<false>()));
    P.rays_o = rays_o; P.rays_d = rays_d; P.z_vals = z_vals; P.raw = raw; P.target_rgb = target_rgb; P.target_d = target_d;
    P.acc = acc; P.loss_grads = loss_grads; P.g_rgb = g_rgb; P.g_depth = g_depth; P.g_raw = g_raw;
    P.draw = reinterpret_cast<float4*>(scratch); P.g_rays_o = g_rays_o; P.g_rays_d = g_rays_d;
    return scene->two_grid ? dispatch_bwd<true>(P, ray_grad, param_grad, st) : dispatch_bwd<false>(P, ray_grad, param_grad, st);
}

namespace mne {
int render_backward_points(const MneSceneDesc*, int32_t, int32_t, const float*, const float*, const float*, void*, float*, float*,
                           bool, float*, cudaStream_t);
int decoder_wgrad(const MneSceneDesc*, int32_t, int32_t, void*, const float*, int32_t, cudaStream_t);
}  // namespace mne

extern "C" int mne_render_backward_points(const MneSceneDesc* scene, int32_t n_rays, int32_t n_samples, const float* rays_o,
                                          const float* rays_d, const float* z_vals, void* scratch, float* g_rays_o,
                                          float* g_rays_d, float* act, mne_stream_t stream) {
    using namespace mne;
    MNE_CHECK_ARG(impl_version() == 2, "mne_render_backward_points: needs the tensor-pipe kernels (option impl = 2)");
    MNE_CHECK_ARG(scene, "mne_render_backward_points: null scene");
    if (int e = check_scene(scene, "mne_render_backward_points", false)) return e;
    MNE_CHECK_ARG(n_rays >= 0 && n_samples > 0, "mne_render_backward_points: bad sizes n_rays=%d n_samples=%d", n_rays, n_samples);
    if (n_rays == 0) return 0;
    MNE_CHECK_ARG(rays_o && rays_d && z_vals && scratch, "mne_render_backward_points: rays_o, rays_d, z_vals and scratch are required");
    MNE_CHECK_ARG(((uintptr_t)scratch & 15) == 0 && ((uintptr_t)act & 31) == 0, "mne_render_backward_points: scratch must be 16-byte, act 32-byte aligned");
    MNE_CHECK_ARG((int64_t)n_rays * n_samples < (int64_t)1 << 31, "mne_render_backward_points: n_rays * n_samples must be < 2^31");
    MNE_CHECK_ARG((g_rays_o == nullptr) == (g_rays_d == nullptr), "mne_render_backward_points: pass both g_rays_o and g_rays_d or neither");
    bool param_grad = scene->g_sdf0 || scene->g_sdf1 || scene->g_col0 || scene->g_col1;
    if (scene->encoder == 1) param_grad = param_grad || scene->hash[0].grad || (scene->two_grid && scene->hash[1].grad);
    else for (int i = 0; i < (scene->two_grid ? 12 : 6); ++i) param_grad = param_grad || scene->planes[i].grad;
    const bool ray_grad = g_rays_o != nullptr;
    if (!param_grad && !ray_grad) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    if (ray_grad) {
        MNE_CHECK_CUDA(cudaMemsetAsync(g_rays_o, 0, sizeof(float) * 3 * (size_t)n_rays, st));
        MNE_CHECK_CUDA(cudaMemsetAsync(g_rays_d, 0, sizeof(float) * 3 * (size_t)n_rays, st));
    }
    MNE_CHECK_ARG(!scene->dirty || scene->dirty_base, "mne_render_backward_points: dirty without dirty_base");
    MNE_CHECK_ARG(act || !(scene->two_grid && ray_grad && param_grad),
                  "mne_render_backward_points: two grids with ray AND parameter gradients need the forward's record (act)");
    return render_backward_points(scene, n_rays, n_samples, rays_o, rays_d, z_vals, scratch, g_rays_o, g_rays_d, param_grad, act, st);
}

extern "C" int mne_decoder_wgrad(const MneSceneDesc* scene, int32_t n_rays, int32_t n_samples, void* scratch, const float* act,
                                 int32_t max_ctas, mne_stream_t stream) {
    using namespace mne;
    MNE_CHECK_ARG(scene && scratch && act, "mne_decoder_wgrad: null argument");
    if (int e = check_scene(scene, "mne_decoder_wgrad", false)) return e;
    MNE_CHECK_ARG(n_rays >= 0 && n_samples > 0 && (int64_t)n_rays * n_samples < (int64_t)1 << 31, "mne_decoder_wgrad: bad sizes");
    MNE_CHECK_ARG(((uintptr_t)scratch & 15) == 0 && ((uintptr_t)act & 31) == 0, "mne_decoder_wgrad: scratch must be 16-byte, act 32-byte aligned");
    MNE_CHECK_ARG(scene->encoder == 1 || scene_is_vectorizable(scene), "mne_decoder_wgrad: needs channels-last planes or the hash grid");
    if (n_rays == 0 || !(scene->g_sdf0 || scene->g_sdf1 || scene->g_col0 || scene->g_col1)) return 0;
    return decoder_wgrad(scene, n_rays, n_samples, scratch, act, max_ctas, (cudaStream_t)stream);
}

// ---------------------------------------------------------------------------------------------------------
// Backward for callers that may not allocate.
//
// mne_render_backward needs a caller-supplied workspace of 16 bytes per sample.  Some callers cannot provide
// one: code that runs where cudaMalloc is off limits (inside a region the host application has fenced off
// from the allocator, in a real-time loop that must not take the driver's allocation lock, or in a plugin
// whose host only hands it input and output buffers).  For them the workspace below lives in the library
// image itself: every GPU that loads the library reserves MNE_STATIC_SCRATCH_SAMPLES * 20 bytes (2.5 MiB)
// until the process exits, and nothing is allocated at call time.
//
// In host code the name `g_static_draw` is only a handle for the symbol, not a device address, so the device
// address is looked up once per device with cudaGetSymbolAddress (with that device current) and cached.
// ---------------------------------------------------------------------------------------------------------
namespace mne {

constexpr int kStaticScratchSamples = MNE_STATIC_SCRATCH_SAMPLES;
static __device__ __align__(256) float4 g_static_draw[kStaticScratchSamples + kStaticScratchSamples / 4 + 1];   // = mne_backward_scratch_bytes

static int static_scratch_for_current_device(float4** out) {
    constexpr int kMaxDevices = 64;
    static float4* cached[kMaxDevices] = {};
    static std::mutex lock;
    int dev = 0;
    MNE_CHECK_CUDA(cudaGetDevice(&dev));
    MNE_CHECK_ARG(dev >= 0 && dev < kMaxDevices, "mne_render_backward_static: device index %d not supported", dev);
    std::lock_guard<std::mutex> guard(lock);
    if (!cached[dev]) {
        float4* ptr = nullptr;
        MNE_CHECK_CUDA(cudaGetSymbolAddress((void**)&ptr, g_static_draw));
        MNE_CHECK_ARG(ptr != nullptr, "mne_render_backward_static: static scratch has no device address");
        cached[dev] = ptr;
    }
    *out = cached[dev];
    return 0;
}

}  // namespace mne

extern "C" int mne_render_backward_static(const MneSceneDesc* scene, const MneRenderCfg* rcfg, int32_t n_rays,
                                          int32_t n_samples, const float* rays_o, const float* rays_d, const float* z_vals,
                                          const float* raw, const float* target_rgb, const float* target_d,
                                          const double* acc, const float* loss_grads, const float* g_rgb,
                                          const float* g_depth, const float* g_raw, float* g_rays_o, float* g_rays_d,
                                          mne_stream_t stream) {
    using namespace mne;
    MNE_CHECK_ARG(n_rays >= 0 && n_samples > 0, "mne_render_backward_static: bad sizes n_rays=%d n_samples=%d", n_rays, n_samples);
    MNE_CHECK_ARG((int64_t)n_rays * n_samples <= (int64_t)kStaticScratchSamples,
                  "mne_render_backward_static: %lld samples exceed the built-in workspace of %d samples "
                  "(MNE_STATIC_SCRATCH_SAMPLES); use mne_render_backward with your own scratch buffer",
                  (long long)n_rays * n_samples, kStaticScratchSamples);
    float4* scratch = nullptr;
    if (int e = static_scratch_for_current_device(&scratch)) return e;
    // Every sample slot of the workspace that the kernels read is written earlier in the same call (the compositing-
    // backward kernel fills d_raw for all n_rays * n_samples entries and the list of non-zero ones; the list length is
    // reset by a memset node on the stream), so leftovers of a previous call are never read.
    return mne_render_backward(scene, rcfg, n_rays, n_samples, rays_o, rays_d, z_vals, raw, target_rgb, target_d, acc,
                               loss_grads, g_rgb, g_depth, g_raw, scratch, g_rays_o, g_rays_d, nullptr, stream);
}
