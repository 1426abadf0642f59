// Per-ray work of the path, one warp per ray: raw2outputs / sdf2weights (reference model/scene_rep.py:183-230), the loss
// sums of JointEncoding.forward (:569-590), and its backward: d loss / d raw for every sample plus the compaction of the
// samples where that is non-zero.  Used by the stand-alone ray kernels (rays_fwd_kernel, rays_bwd_kernel) and by the tail
// of the point kernel, where the warp that writes the last sample of a ray finishes the ray on the spot.
#pragma once
#include "render_common.cuh"

namespace mne {

struct RayFwdParams {
    MneRenderCfg rc;
    int n_rays, n_samples;
    const float* z_vals;
    const float* raw;
    const float* target_rgb;
    const float* target_d;
    float* weights;
    float* rgb;
    float* depth;
    float* disp;
    float* acc_map;
    float* depth_var;
    int32_t* inds;
    double* acc;
};

struct RayBwdParams {
    MneRenderCfg rc;
    int n_rays, n_samples;
    const float* z_vals;
    const float* raw;
    const float* target_rgb;
    const float* target_d;
    const double* acc;
    const float* loss_grads;
    const float* g_rgb;
    const float* g_depth;
    const float* g_raw;
    float4* draw;
    int* list;                 // indices of the samples with a non-zero draw, [*n_active]
    int* n_active;             // zeroed before the launch
};

// loss sums, in this order (a ray's own sums are float; they are accumulated over rays in double)
enum RaySum { RS_RGB = 0, RS_DEPTH, RS_CO_FS, RS_CO_SDF, RS_E_FS, RS_E_CENTER, RS_E_TAIL, RS_COUNT };

// factors of d loss / d (per-sample residuals): loss weights (the upstream gradient) over the mask counts
struct RayGradScales {
    float rgb = 0.f, depth = 0.f, co_fs = 0.f, co_sdf = 0.f, e_fs = 0.f, e_center = 0.f, e_tail = 0.f;
};

__device__ __forceinline__ RayGradScales ray_grad_scales(const MneRenderCfg& rc, int S, const double* acc, const float* lg) {
    RayGradScales k;
    if (!lg) return k;
    const double Nn = rc.norm_rays, NS = rc.norm_rays * (double)S;
    auto inv = [](double c) { return c > 0.0 ? 1.0 / c : 0.0; };
    k.rgb = (float)(2.0 * lg[L_RGB] / (3.0 * Nn));
    k.depth = (float)(2.0 * lg[L_DEPTH] * inv(acc[A_N_VALID]));
    if (!rc.first_pass) {
        const float n_fs = (float)acc[A_N_CO_FRONT], n_sdf = (float)acc[A_N_CO_NEAR];
        const float fs_w = 1.0f - n_fs / (n_fs + n_sdf), sdf_w = 1.0f - n_sdf / (n_fs + n_sdf);
        k.co_fs = (n_fs + n_sdf > 0.f) ? (float)(2.0 * lg[L_CO_FS] * fs_w / NS) : 0.f;
        k.co_sdf = (n_fs + n_sdf > 0.f) ? (float)(2.0 * lg[L_CO_SDF] * sdf_w / NS) * rc.sc_trunc : 0.f;
        k.e_fs = (float)(2.0 * lg[L_E_FS] * inv(acc[A_N_E_FRONT]));
        k.e_center = (float)(2.0 * lg[L_E_CENTER] * inv(acc[A_N_E_CENTER])) * rc.model_trunc;
        k.e_tail = (float)(2.0 * lg[L_E_TAIL] * inv(acc[A_N_E_TAIL])) * rc.model_trunc;
    }
    return k;
}

// forward of one ray: rendered maps, weights and (with_loss) the ray's loss sums, warp-reduced: sums[] is valid in every lane
__device__ __forceinline__ RayComposite ray_forward_one(const RayFwdParams& P, int ray, int lane, bool with_loss, float (&sums)[RS_COUNT]) {
    const int S = P.n_samples;
    const size_t rb = (size_t)ray * S;
    const RayComposite c = composite_ray(P.raw, P.z_vals, rb, S, P.rc, lane);
    const float d = P.target_d ? P.target_d[ray] : 0.0f;
    float var = 0.f;
#pragma unroll
    for (int k = 0; k < RS_COUNT; ++k) sums[k] = 0.f;
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
            if (m & M_CO_FRONT) sums[RS_CO_FS] += fs;
            if (m & M_CO_NEAR) sums[RS_CO_SDF] += rc_co * rc_co;
            if (m & M_E_FRONT) sums[RS_E_FS] += fs;
            if (m & M_E_CENTER) sums[RS_E_CENTER] += rc_e * rc_e;
            if (m & M_E_TAIL) sums[RS_E_TAIL] += rc_e * rc_e;
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
            const float e0 = rr - P.target_rgb[ray * 3 + 0], e1 = gg - P.target_rgb[ray * 3 + 1], e2 = bb - P.target_rgb[ray * 3 + 2];
            sums[RS_RGB] = (e0 * e0 + e1 * e1) + e2 * e2;
            if (depth_valid(d, P.rc)) sums[RS_DEPTH] = (c.depth - d) * (c.depth - d);
        }
    }
    if (with_loss) {
#pragma unroll
        for (int k = 0; k < RS_COUNT; ++k) sums[k] = warp_sum(sums[k]);
    }
    return c;
}

// add a block's loss sums to the accumulators (call with all threads of the block).  warp_sums: [warps][RS_COUNT] doubles
// in shared memory, one row per warp (written by that warp's lane 0); `red` = RS_COUNT doubles of shared memory.
__device__ __forceinline__ void flush_loss_sums(const double* warp_sums, int n_warps, const MneRenderCfg& rc, double* acc, double* red) {
    __syncthreads();
    if (threadIdx.x < RS_COUNT) {
        double s = 0.0;
        for (int w = 0; w < n_warps; ++w) s += warp_sums[w * RS_COUNT + threadIdx.x];
        red[threadIdx.x] = s;
    }
    if (threadIdx.x < 7 && red[threadIdx.x] != 0.0) {
        const int slot_first[2] = {A_S_RGB0, A_S_DEPTH0};
        const int slot[7] = {A_S_RGB, A_S_DEPTH, A_S_CO_FS, A_S_CO_SDF, A_S_E_FS, A_S_E_CENTER, A_S_E_TAIL};
        if (rc.first_pass) { if (threadIdx.x < 2) atomicAdd(acc + slot_first[threadIdx.x], red[threadIdx.x]); }
        else atomicAdd(acc + slot[threadIdx.x], red[threadIdx.x]);
    }
}

// backward of one ray: d loss / d raw of every sample into P.draw, the non-zero ones appended to P.list.
// c = the ray's composite (recomputed by the caller, or handed over by ray_forward_one).
__device__ __forceinline__ void ray_backward_one(const RayBwdParams& P, int ray, int lane, const RayGradScales& K, const RayComposite& c) {
    const int S = P.n_samples;
    const float* lg = P.loss_grads;
    const size_t rb = (size_t)ray * S;
    const float d = P.target_d ? P.target_d[ray] : 0.0f;
    float rr[3] = {c.rgb[0], c.rgb[1], c.rgb[2]};
    if (P.rc.white_bkgd) { rr[0] += 1.f - c.acc; rr[1] += 1.f - c.acc; rr[2] += 1.f - c.acc; }
    float G[3] = {0.f, 0.f, 0.f}, Gd = 0.f;
    if (lg && P.target_rgb) {
#pragma unroll
        for (int k = 0; k < 3; ++k) G[k] = K.rgb * (rr[k] - P.target_rgb[ray * 3 + k]);
        if (depth_valid(d, P.rc)) Gd = K.depth * (c.depth - d);
    }
    if (P.g_rgb) { G[0] += P.g_rgb[ray * 3 + 0]; G[1] += P.g_rgb[ray * 3 + 1]; G[2] += P.g_rgb[ray * 3 + 2]; }
    if (P.g_depth) Gd += P.g_depth[ray];
    for (int s0 = 0; s0 < S; s0 += 32) {
        const int s = s0 + lane;
        bool active = false;
        if (s < S) {
            const float z = __ldcg(P.z_vals + rb + s);
            const float4 v = __ldcg(reinterpret_cast<const float4*>(P.raw) + rb + s);
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
                if (m & M_CO_FRONT) o.w += K.co_fs * (v.w - 1.0f);
                if (m & M_CO_NEAR) o.w += K.co_sdf * (__fadd_rn(z, __fmul_rn(v.w, P.rc.sc_trunc)) - d);
                if (m & M_E_FRONT) o.w += K.e_fs * (v.w - 1.0f);
                const float re = __fadd_rn(z, __fmul_rn(v.w, P.rc.model_trunc)) - d;
                if (m & M_E_CENTER) o.w += K.e_center * re;
                if (m & M_E_TAIL) o.w += K.e_tail * re;
            }
            if (P.g_raw) {
                const float4 e = __ldg(reinterpret_cast<const float4*>(P.g_raw) + rb + s);
                o.x += e.x; o.y += e.y; o.z += e.z; o.w += e.w;
            }
            P.draw[rb + s] = o;
            active = o.x != 0.0f || o.y != 0.0f || o.z != 0.0f || o.w != 0.0f;
        }
        const unsigned b = __ballot_sync(0xffffffffu, active);
        if (b) {
            int base = 0;
            if (lane == 0) base = atomicAdd(P.n_active, __popc(b));
            base = __shfl_sync(0xffffffffu, base, 0);
            if (active) P.list[base + __popc(b & ((1u << lane) - 1u))] = (int)(rb + s);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// The same per-ray work with the ray held in registers (up to 32 * NC samples): z and raw are read from L2 ONCE, every
// pass after that is register arithmetic and shuffles, and the compaction costs one atomic per ray instead of one per 32
// samples.  The arithmetic and its order are those of composite_ray / ray_forward_one / ray_backward_one (the results are
// bit-identical); what changes is the latency: one memory round trip + one atomic instead of a chain of six or seven.
// TRAIN: also d loss / d raw (upstream = the loss weights only) and the list of samples where it is non-zero.
// ---------------------------------------------------------------------------------------------
template <int NC, bool TRAIN>
__device__ __forceinline__ void ray_tail_regs(const RayFwdParams& P, int ray, int lane, bool with_loss, float (&sums)[RS_COUNT],
                                              const RayGradScales& K, const float* loss_grads, float4* draw, int* list, int* n_active) {
    const int S = P.n_samples;
    const size_t rb = (size_t)ray * S;
    const MneRenderCfg& rc = P.rc;
    float z[NC];
    float4 v[NC];
#pragma unroll
    for (int j = 0; j < NC; ++j) {
        const int s = lane + 32 * j;
        z[j] = 0.f; v[j] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (s < S) { z[j] = __ldcg(P.z_vals + rb + s); v[j] = __ldcg(reinterpret_cast<const float4*>(P.raw) + rb + s); }
    }
    // ---- first sign change of the sdf along the ray (0 if none), sdf2weights :195-197
    int first = 0x7fffffff;
#pragma unroll
    for (int j = 0; j < NC; ++j) {
        const int s = lane + 32 * j;
        const float nxt_lane = __shfl_down_sync(0xffffffffu, v[j].w, 1);
        const float nxt_chunk = (j + 1 < NC) ? __shfl_sync(0xffffffffu, v[(j + 1 < NC) ? j + 1 : j].w, 0) : 0.f;
        const float b = (lane == 31) ? nxt_chunk : nxt_lane;
        if (s + 1 < S && __fmul_rn(b, v[j].w) < 0.0f) first = min(first, s);
    }
    first = warp_min(first);
    const int ind = (first == 0x7fffffff) ? 0 : first;
    float z_ind = 0.f;
#pragma unroll
    for (int j = 0; j < NC; ++j) {
        const float cand = __shfl_sync(0xffffffffu, z[j], ind & 31);
        if ((ind >> 5) == j) z_ind = cand;
    }
    const float thresh = __fadd_rn(z_ind, rc.sc_trunc);
    // ---- weights: u = sigmoid(s/tr) sigmoid(-s/tr) inside the window, normalised
    float u[NC], sp[NC], c0[NC], c1[NC], c2[NC];
    float sw = 0.f, sr = 0.f, sg = 0.f, sb = 0.f, sz = 0.f;
#pragma unroll
    for (int j = 0; j < NC; ++j) {
        const int s = lane + 32 * j;
        u[j] = 0.f; sp[j] = 0.f; c0[j] = c1[j] = c2[j] = 0.f;
        if (s < S && z[j] < thresh) {
            u[j] = sdf_bell(v[j].w, rc.trunc, &sp[j]);
            c0[j] = sigmoidf(v[j].x); c1[j] = sigmoidf(v[j].y); c2[j] = sigmoidf(v[j].z);
            sw += u[j]; sr += u[j] * c0[j]; sg += u[j] * c1[j]; sb += u[j] * c2[j]; sz += u[j] * z[j];
        }
    }
    sw = warp_sum(sw); sr = warp_sum(sr); sg = warp_sum(sg); sb = warp_sum(sb); sz = warp_sum(sz);
    const float inv_norm = 1.0f / (sw + 1e-8f);
    const float depth = sz * inv_norm, acc_w = sw * inv_norm;
    float rr[3] = {sr * inv_norm, sg * inv_norm, sb * inv_norm};
    const float d = P.target_d ? P.target_d[ray] : 0.0f;
    // ---- per-sample outputs and loss sums
    float var = 0.f;
#pragma unroll
    for (int k = 0; k < RS_COUNT; ++k) sums[k] = 0.f;
#pragma unroll
    for (int j = 0; j < NC; ++j) {
        const int s = lane + 32 * j;
        if (s < S) {
            float w = 0.f;
            if (z[j] < thresh) w = u[j] * inv_norm;
            if (P.weights) P.weights[rb + s] = w;
            var += w * (z[j] - depth) * (z[j] - depth);
            if (with_loss && !rc.first_pass) {
                const unsigned m = sample_masks(z[j], d, rc);
                const float sdf = v[j].w;
                const float fs = (sdf - 1.0f) * (sdf - 1.0f);
                const float rc_co = __fadd_rn(z[j], __fmul_rn(sdf, rc.sc_trunc)) - d;
                const float rc_e = __fadd_rn(z[j], __fmul_rn(sdf, rc.model_trunc)) - d;
                if (m & M_CO_FRONT) sums[RS_CO_FS] += fs;
                if (m & M_CO_NEAR) sums[RS_CO_SDF] += rc_co * rc_co;
                if (m & M_E_FRONT) sums[RS_E_FS] += fs;
                if (m & M_E_CENTER) sums[RS_E_CENTER] += rc_e * rc_e;
                if (m & M_E_TAIL) sums[RS_E_TAIL] += rc_e * rc_e;
            }
        }
    }
    var = warp_sum(var);
    if (rc.white_bkgd) { rr[0] += 1.f - acc_w; rr[1] += 1.f - acc_w; rr[2] += 1.f - acc_w; }
    if (lane == 0) {
        if (P.rgb) { P.rgb[ray * 3 + 0] = rr[0]; P.rgb[ray * 3 + 1] = rr[1]; P.rgb[ray * 3 + 2] = rr[2]; }
        if (P.depth) P.depth[ray] = depth;
        if (P.disp) P.disp[ray] = 1.0f / fmaxf(1e-10f, depth / acc_w);
        if (P.acc_map) P.acc_map[ray] = acc_w;
        if (P.depth_var) P.depth_var[ray] = var;
        if (P.inds) P.inds[ray] = ind;
        if (with_loss) {
            const float e0 = rr[0] - P.target_rgb[ray * 3 + 0], e1 = rr[1] - P.target_rgb[ray * 3 + 1], e2 = rr[2] - P.target_rgb[ray * 3 + 2];
            sums[RS_RGB] = (e0 * e0 + e1 * e1) + e2 * e2;
            if (depth_valid(d, rc)) sums[RS_DEPTH] = (depth - d) * (depth - d);
        }
    }
    if (with_loss) {
#pragma unroll
        for (int k = 0; k < RS_COUNT; ++k) sums[k] = warp_sum(sums[k]);
    }
    if (!TRAIN) return;
    // ---- d loss / d raw (ray_backward_one with the loss weights as the only upstream gradient) and compaction
    float G[3] = {0.f, 0.f, 0.f}, Gd = 0.f;
    if (loss_grads && P.target_rgb) {
#pragma unroll
        for (int k = 0; k < 3; ++k) G[k] = K.rgb * (rr[k] - P.target_rgb[ray * 3 + k]);
        if (depth_valid(d, rc)) Gd = K.depth * (depth - d);
    }
    float4 o[NC];
    unsigned ball[NC];
    int total = 0;
#pragma unroll
    for (int j = 0; j < NC; ++j) {
        const int s = lane + 32 * j;
        o[j] = make_float4(0.f, 0.f, 0.f, 0.f);
        bool active = false;
        if (s < S) {
            if (z[j] < thresh) {
                const float w = u[j] * inv_norm;
                const float dLdu = inv_norm * (G[0] * (c0[j] - rr[0]) + G[1] * (c1[j] - rr[1]) + G[2] * (c2[j] - rr[2]) + Gd * (z[j] - depth));
                o[j].w = dLdu * u[j] * (1.0f - 2.0f * sp[j]) / rc.trunc;
                o[j].x = G[0] * w * c0[j] * (1.0f - c0[j]);
                o[j].y = G[1] * w * c1[j] * (1.0f - c1[j]);
                o[j].z = G[2] * w * c2[j] * (1.0f - c2[j]);
            }
            if (loss_grads && P.target_d && !rc.first_pass) {
                const unsigned m = sample_masks(z[j], d, rc);
                const float sdf = v[j].w;
                if (m & M_CO_FRONT) o[j].w += K.co_fs * (sdf - 1.0f);
                if (m & M_CO_NEAR) o[j].w += K.co_sdf * (__fadd_rn(z[j], __fmul_rn(sdf, rc.sc_trunc)) - d);
                if (m & M_E_FRONT) o[j].w += K.e_fs * (sdf - 1.0f);
                const float re = __fadd_rn(z[j], __fmul_rn(sdf, rc.model_trunc)) - d;
                if (m & M_E_CENTER) o[j].w += K.e_center * re;
                if (m & M_E_TAIL) o[j].w += K.e_tail * re;
            }
            draw[rb + s] = o[j];
            active = o[j].x != 0.0f || o[j].y != 0.0f || o[j].z != 0.0f || o[j].w != 0.0f;
        }
        ball[j] = __ballot_sync(0xffffffffu, active);
        total += __popc(ball[j]);
    }
    if (total) {
        int base = 0;
        if (lane == 0) base = atomicAdd(n_active, total);
        base = __shfl_sync(0xffffffffu, base, 0);
#pragma unroll
        for (int j = 0; j < NC; ++j) {
            if (ball[j] & (1u << lane)) list[base + __popc(ball[j] & ((1u << lane) - 1u))] = (int)(rb + lane + 32 * j);
            base += __popc(ball[j]);
        }
    }
}

}  // namespace mne
