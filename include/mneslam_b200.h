/*
 * mneslam_b200 -- C ABI of the B200-native JointEncoding ray-batch path.
 *
 * The reference (dtc111111/MNESLAM) has no FFI layer: its boundary for this path is the Python class
 * model/scene_rep.py::JointEncoding.  This library is what a native binding behind that class binds:
 * every entry point names the reference code it replaces.  Plain pointers and sizes only; no torch
 * types.  All pointers are DEVICE pointers unless a comment says "host".  All calls are asynchronous on
 * `stream` (a cudaStream_t passed as void*), never synchronise the device, never throw and never allocate
 * device memory (workspaces are passed in by the caller).  Return value: 0 = ok, nonzero = error; the
 * message is available from mne_last_error() (thread-local).
 *
 * Tensor layouts (row-major, float32 unless noted):
 *   rays_o, rays_d [N,3]   z_vals [N,S]   raw [N,S,4] = (rgb logits x3, sdf)   target_rgb [N,3]
 *   target_d [N]           noise [N,S]    rgb [N,3]   depth/disp/acc/depth_var [N]   inds [N] int32
 */
#ifndef MNESLAM_B200_H
#define MNESLAM_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MNE_MAX_PLANES 12
#define MNE_ACC_SLOTS 16
#define MNE_N_LOSSES 8

typedef void* mne_stream_t; /* cudaStream_t */

/* One feature plane, logical shape [1, C, H, W] (reference model/scene_rep.py:104-109), addressed through
 * element strides so that both torch.channels_last (stride_c == 1: the fast, coalesced layout) and plain
 * NCHW tensors (e.g. a checkpoint written by the reference, mneslam_mp.py:294-315) work.  `grad` has the
 * same shape/strides and is ACCUMULATED into (+=); NULL = no gradient for this plane. */
typedef struct MnePlaneDesc {
    const float* data;
    float* grad;
    int32_t H, W;
    int64_t stride_c, stride_h, stride_w;
} MnePlaneDesc;

/* Multi-resolution hash grid (the encoder the reference's configs name, model/encodings.py:31-46;
 * tiny-cuda-nn "HashGrid": n_levels x n_features tables concatenated, level l at table + offset[l]). */
typedef struct MneHashDesc {
    const float* table;      /* fp32 parameters [entries][2] (the optimiser's master copy) */
    float* grad;             /* fp32 gradient accumulator of the same shape (+=), or NULL */
    const void* table_h;     /* optional fp16 copy of `table` ([entries] half2, kept in step by the caller): when non-NULL the
                              * kernels READ the features from it with one 32-bit load per entry -- tiny-cuda-nn's default
                              * parameter precision; parity gate 1e-2 instead of 1e-3 (BASELINE.json).  Gradients stay fp32. */
    int32_t n_levels;        /* 16 */
    int32_t n_features;      /* 2 */
    int32_t base_resolution; /* 16 */
    float per_level_scale;
    uint32_t offsets[33];    /* entry offset of each level, offsets[n_levels] = total entries */
    /* per-level grid scale (2^(l log2 pls) * base - 1) and resolution (ceil(scale) + 1), computed once on the host
     * so that device and oracle use bit-identical values (an exp2f of the device could differ in the last bit) */
    float level_scale[32];
    uint32_t level_res[32];
} MneHashDesc;

/* Scene state read at call time (the reference reads model.all_planes / decoder weights at call time too:
 * SURVEY.md 8(b) "Attributes read/written from outside"). */
typedef struct MneSceneDesc {
    int32_t encoder;  /* 0 = tri-plane (live in the reference), 1 = hash grid: 16 levels x 2 features, one table
                       * (grid.oneGrid) or an sdf and a colour table (two_grid; its backward needs the `act` buffer);
                       * forward, table / decoder gradients and ray gradients */
    int32_t two_grid; /* 1 when grid.oneGrid is False: colour planes / colour table present */
    int32_t n_bins;   /* OneBlob bins per dimension; must be 16 */
    int32_t c_dim;    /* plane channels; must be 32 */
    /* tri-plane: index = group*6 + axis*2 + level; group 0 = sdf planes, 1 = colour planes;
     * axis 0 = xy [Ny,Nx], 1 = xz [Nz,Nx], 2 = yz [Nz,Ny]; level 0 = coarse, 1 = fine */
    MnePlaneDesc planes[MNE_MAX_PLANES];
    MneHashDesc hash[2]; /* [0] sdf table, [1] colour table (two_grid) */
    /* plane normalisation box = JointEncoding.bound (model/scene_rep.py:80-83), lo and (hi - lo) in fp32 */
    float bound_lo[3], bound_ext[3];
    /* OneBlob normalisation box = JointEncoding.bounding_box, float64 (model/scene_rep.py:292) */
    double bbox_lo[3], bbox_ext[3];
    /* decoder: the four bias-free nn.Linear weights, row-major [out, in] (model/decoder.py:39-55,92-108):
     * sdf0 [32, F+48], sdf1 [16, 32], col0 [32, 48 (+F) + 15], col1 [3, 32]; F = 64 (planes) or 32 (hash) */
    const float* w_sdf0;
    const float* w_sdf1;
    const float* w_col0;
    const float* w_col1;
    /* gradient accumulators of the same shapes (+=), or NULL */
    float* g_sdf0;
    float* g_sdf1;
    float* g_col0;
    float* g_col1;
    /* optional (multi-GPU optimiser, no counterpart in the reference): when dirty is non-NULL, mne_render_backward
     * also sets dirty[(g - dirty_base) / 32] = 1 for every 32-float texel of the plane / hash-table gradient it adds
     * into (g = the address it adds at; all plane / table gradients must then live in the one buffer dirty_base points
     * into).  Decoder-weight gradients are NOT flagged (they are dense; mne_dirty_push flags them wholesale).
     * Supported by the vectorised-plane and hash paths of impl 2; any other path rejects a non-NULL dirty. */
    uint8_t* dirty;
    const float* dirty_base;
} MneSceneDesc;

/* Constants of the sampling stage (reference model/scene_rep.py:361-381).  The linspace tables are built
 * by the host exactly as the reference builds them (torch.linspace on the CPU) so sample positions are
 * bit-identical by construction. */
typedef struct MneSampleCfg {
    int32_t n_uniform;            /* training.n_samples_d (target_d given) or training.n_samples */
    int32_t n_range;              /* training.n_range_d, 0 when target_d is NULL */
    const float* uniform;         /* [n_uniform] linspace(near, far) */
    const float* range_offsets;   /* [n_range]   linspace(-range_d, range_d) */
    const float* range_fallback;  /* [n_range]   linspace(near, far), used where target_d <= 0 */
    int32_t perturb;              /* training.perturb > 0 */
} MneSampleCfg;

/* Constants of compositing and of the losses.  Floats are the float32 roundings of the Python doubles the
 * reference combines with float32 tensors. */
typedef struct MneRenderCfg {
    float trunc;        /* training.trunc: sdf2weights sigmoid scale (model/scene_rep.py:193) */
    float sc_trunc;     /* data.sc_factor * training.trunc: mask window (:200) and Co-SLAM truncation (:586) */
    float model_trunc;  /* model.truncation: ESLAM losses (:489-507) */
    float center_trunc; /* 0.4 * model.truncation (:495-496) */
    float depth_trunc;  /* cam.depth_trunc (:570) */
    int32_t white_bkgd; /* training.white_bkgd (:227) */
    int32_t first_pass; /* 1: coarse pass of importance sampling -- only rgb0/depth0 losses (:579-581) */
    double norm_rays;   /* number of rays the means run over; = n_rays unless the batch is sharded over GPUs */
} MneRenderCfg;

/* Loss accumulators, MNE_ACC_SLOTS doubles on the device (one NCCL all-reduce when rays are sharded):
 * [0] n_valid_depth [1] n_co_front [2] n_co_near [3] n_e_front [4] n_e_center [5] n_e_tail   (sample stage)
 * [6] s_rgb [7] s_depth [8] s_co_fs [9] s_co_sdf [10] s_e_fs [11] s_e_center [12] s_e_tail
 * [13] s_rgb0 [14] s_depth0 [15] spare                                                        (forward)  */

/* ---- library ---- */
int mne_version(void);
const char* mne_last_error(void);
/* Process-wide switches.  "impl": 2 (default) = tensor-pipe kernels (16-point tiles, mma TF32 with split
 * operands), 1 = the first-generation fp32 CUDA-core kernels, kept as an on-device cross-check.
 * "adam_spare_sms" (default 0, at most 64): SMs the persistent optimiser grids leave free, so that small kernels on
 * another stream can run under them.  "hash_level0_smem" (default 0): stage level 0 of the hash table(s) in shared memory
 * (measured slower on B200).  "peer_barrier_polls_log2" (default 29, 10..31): mne_peer_barrier gives up after 2^v polls. */
int mne_set_option(const char* key, int value);
/* Current value of an option of mne_set_option (so that a caller can put it back). */
int mne_get_option(const char* key, int* value);

/* ---- ray assembly: replaces the gather of KeyFrameDatabase.sample_global_rays (model/keyframe.py:91-103) and the
 * assembly in mapping_optimize (mp_slam/mapper.py:135-153).  rays7: [*, 7] rows (camera-frame direction, rgb, depth);
 * idx [n_rays] (or NULL = rows 0..n_rays-1): the rows to take; kf_ids [n_rays] (or NULL = idx / per_kf): the pose
 * of each ray, -1 = the last one (the current frame); poses [n_kf, 4, 4] camera-to-world, row-major.  Writes
 * rays_o / rays_d / target_rgb [n_rays, 3] and target_d [n_rays]. */
int mne_assemble_rays(const float* rays7, const int64_t* idx, int64_t per_kf, const int64_t* kf_ids, const float* poses,
                      int32_t n_kf, int32_t n_rays, float* rays_o, float* rays_d, float* target_rgb, float* target_d,
                      mne_stream_t stream);

/* The draw itself on the device, fused with the assembly: replaces KeyFrameDatabase.sample_global_rays (reference
 * model/keyframe.py:91-103, random.sample = WITHOUT replacement over the n_stored = n_kf * per_kf stored rays) plus the draw
 * of max(sample // n_kf, min_pixels_cur) pixels of the current frame with pose id -1 (mp_slam/mapper.py:135-153).  key: one
 * int64 in device memory (refresh it per iteration, e.g. torch.randint -- under CUDA-graph replay too); the n_global keyframe
 * rays and the n_cur current-frame rays (current7 [n_current, 7], may be NULL when n_cur = 0) are distinct rows, element i of
 * each draw being a keyed bijection of the row range applied to i.  Outputs as mne_assemble_rays, [n_global + n_cur] rays,
 * keyframe rays first; rows_out (may be NULL) [n_global + n_cur] receives the drawn row numbers.  A draw larger than its
 * population is an error, as with random.sample. */
int mne_draw_rays(const int64_t* key, const float* rays7, int64_t n_stored, int64_t per_kf, int32_t n_global,
                  const float* current7, int64_t n_current, int32_t n_cur, const float* poses, int32_t n_kf, float* rays_o,
                  float* rays_d, float* target_rgb, float* target_d, int64_t* rows_out, mne_stream_t stream);

/* ---- camera pose: (rotation parameters, translation) -> rays, and ray gradients -> pose update.  rot_rep 0 = axis-angle
 * [3] (at_to_transform_matrix), 1 = real-first quaternion [4], not assumed unit (qt_to_transform_matrix; reference
 * optimization/utils.py:161-210).  mne_pose_rays: rays_o = t, rays_d = sum_k dirs_k R[:, k] (mp_slam/mapper.py:392-393).
 * mne_pose_step replaces the autograd chain ray gradients -> c2w -> pose parameters and the optimizer.step() of
 * get_pose_param_optim (mneslam_mp.py:577-584): reduces d loss / d rays over the rays, differentiates R analytically
 * (forward-mode dual numbers on the expression that builds it), applies torch.optim.Adam's update (no weight decay) with
 * one learning rate for the rotation and one for the translation.  exp_avg / exp_avg_sq [np + 3] (rotation first) and
 * *step_dev start at zero; grad_out (optional, [np + 3]) receives the pose gradient. */
int mne_pose_rays(int32_t rot_rep, const float* rot, const float* trans, const float* dirs_cam, int32_t n_rays,
                  float* rays_o, float* rays_d, mne_stream_t stream);
int mne_pose_step(int32_t rot_rep, float* rot, float* trans, const float* dirs_cam, const float* g_rays_o,
                  const float* g_rays_d, int32_t n_rays, float* exp_avg, float* exp_avg_sq, int64_t* step_dev,
                  float lr_rot, float lr_trans, float beta1, float beta2, float eps, float* grad_out, mne_stream_t stream);

/* ---- sampling: replaces the z_vals block of JointEncoding.render_rays (model/scene_rep.py:361-381) plus
 * the mask counts of get_masks (model/utils.py:117-145), sdf_losses (model/scene_rep.py:489-499) and
 * valid_depth_mask (:570).  target_d NULL = uniform sampling.  noise NULL = no perturbation.  acc may be
 * NULL (no counts); otherwise slots 0..5 are ADDED to (zero it first).  */
int mne_sample_rays(const MneSampleCfg* cfg, const MneRenderCfg* rcfg, int32_t n_rays, const float* target_d,
                    const float* noise, float* z_vals, double* acc, mne_stream_t stream);

/* ---- importance sampling: sample_pdf (model/utils.py:65-104) on the coarse pass's weights and the re-sort of
 * render_rays (model/scene_rep.py:393-397).  z_vals / weights [N,S] from the coarse pass; u = uniform numbers,
 * [N,n_importance] when u_per_ray != 0, else one shared row [n_importance] (the deterministic linspace the
 * reference uses when training.perturb == 0).  Outputs: z_out [N, S + n_importance] sorted; optional inds
 * [N,n_importance] int32 (searchsorted bin indices), z_std [N], and, when acc and target_d are given, the mask
 * counts of the FINAL sample positions added into acc slots 0..5. */
int mne_sample_importance(const MneRenderCfg* rcfg, int32_t n_rays, int32_t n_samples, int32_t n_importance,
                          const float* z_vals, const float* weights, const float* u, int32_t u_per_ray,
                          const float* target_d, float* z_out, int32_t* inds, float* z_std, double* acc,
                          mne_stream_t stream);

/* ---- forward: replaces run_network + raw2outputs/sdf2weights (model/scene_rep.py:183-230,303-317,384-386)
 * and the loss sums of JointEncoding.forward (:569-590).  Optional outputs may be NULL.  target_rgb /
 * target_d / acc NULL = no loss sums.  act (optional, [N*S, mne_act_floats(scene)] floats, 32-byte aligned): one record
 * per sample kept for mne_render_backward -- what autograd would save: the decoders' hidden activations (80 floats), the
 * encoder's features (64, or 32 for the hash grid; twice that with two grids), the ReLU masks, and room for the 80
 * per-sample decoder gradients the backward writes there.  With it the backward neither recomputes the forward nor
 * gathers the features again (impl 2, channels-last planes or hash grid; ignored otherwise). */
#define MNE_ACT_FLOATS_MAX 296
int mne_act_floats(const MneSceneDesc* scene);
int mne_render_forward(const MneSceneDesc* scene, const MneRenderCfg* rcfg, int32_t n_rays, int32_t n_samples,
                       const float* rays_o, const float* rays_d, const float* z_vals, const float* target_rgb,
                       const float* target_d, float* raw, float* weights, float* rgb, float* depth, float* disp,
                       float* acc_map, float* depth_var, int32_t* inds, double* acc, float* act, mne_stream_t stream);

/* ---- the same render as ONE launch: mne_render_forward (run_network + raw2outputs / sdf2weights + the loss sums of
 * JointEncoding.forward, model/scene_rep.py:183-230,303-317,384-386,569-590) with the per-ray work done inside the point
 * kernel by the warp that stores the last sample of a ray.  ray_counters: [n_rays] int32, all ZERO on entry; the kernel
 * leaves them zero (calls that share a buffer must not overlap on the GPU).  ray_counters == NULL: the point kernel is
 * followed by ONE ray kernel that holds each ray in registers (compositing, loss sums and, in training, the backward's
 * head): two launches, no counters -- the faster choice for small batches, where finished rays would otherwise queue up
 * behind the tiles of the warp that finished them.  All other arguments as mne_render_forward.
 * Training (loss_grads != NULL; needs scratch, acc, target_rgb, target_d; the mask counts in acc[0..5] must be final, i.e.
 * already reduced over the ranks when the batch is sharded): the kernel also does the head of the backward pass --
 * d loss / d raw of every sample for the upstream gradient loss_grads[8] (the loss weights of get_loss_from_ret,
 * mneslam_mp.py:350-372) and the list of the samples where it is non-zero, into scratch (mne_backward_scratch_bytes) --
 * so that mne_render_backward_points can follow directly: the whole of loss.backward() (mp_slam/mapper.py:159) up to the
 * point kernels costs no launch of its own. */
int mne_render_forward_fused(const MneSceneDesc* scene, const MneRenderCfg* rcfg, int32_t n_rays, int32_t n_samples,
                             const float* rays_o, const float* rays_d, const float* z_vals, const float* target_rgb,
                             const float* target_d, float* raw, float* weights, float* rgb, float* depth, float* disp,
                             float* acc_map, float* depth_var, int32_t* inds, double* acc, float* act,
                             int32_t* ray_counters, const float* loss_grads, void* scratch, mne_stream_t stream);

/* ---- compositing alone: JointEncoding.raw2outputs / sdf2weights (model/scene_rep.py:183-230) on raw [N,S,4] and
 * z_vals [N,S] supplied by the caller.  Every output may be NULL. */
int mne_composite(const MneRenderCfg* rcfg, int32_t n_rays, int32_t n_samples, const float* z_vals, const float* raw,
                  float* weights, float* rgb, float* depth, float* disp, float* acc_map, float* depth_var, int32_t* inds,
                  mne_stream_t stream);

/* ---- losses from the accumulators: compute_loss / get_sdf_loss / sdf_losses / mse2psnr
 * (model/utils.py:43-47,147-185; model/scene_rep.py:501-507).  losses[8] = rgb_loss, depth_loss,
 * co_sdf_loss, co_fs_loss, e_fs_loss, e_center_loss, e_tail_loss, psnr. */
int mne_loss_finalize(const MneRenderCfg* rcfg, int32_t n_samples, const double* acc, float* losses,
                      mne_stream_t stream);

/* ---- backward: replaces torch autograd through the whole path (loss.backward() at mp_slam/mapper.py:88,159).
 * Upstream gradients: loss_grads[8] (device; d total / d each entry of `losses`, psnr ignored) when the
 * losses are used, plus optional explicit g_rgb [N,3] / g_depth [N] / g_raw [N,S,4] when the caller
 * differentiates the rendered maps or the raw network output itself (loop-closure pose optimisation, mp_slam/mapper.py:388-410).
 * Accumulates (+=) into scene->planes[].grad / hash[].grad / g_* and, if non-NULL, WRITES g_rays_o /
 * g_rays_d [N,3].  `scratch` must hold mne_backward_scratch_bytes(n_rays, n_samples) bytes.  act: the buffer the
 * forward filled for the SAME scene, rays and z_vals (parameters unchanged in between), or NULL = recompute; the
 * backward writes the per-sample decoder gradients into it (its weight-gradient kernel reads them back). */
size_t mne_backward_scratch_bytes(int32_t n_rays, int32_t n_samples);
int mne_render_backward(const MneSceneDesc* scene, const MneRenderCfg* rcfg, int32_t n_rays, int32_t n_samples,
                        const float* rays_o, const float* rays_d, const float* z_vals, const float* raw,
                        const float* target_rgb, const float* target_d, const double* acc,
                        const float* loss_grads, const float* g_rgb, const float* g_depth, const float* g_raw,
                        void* scratch, float* g_rays_o, float* g_rays_d, float* act, mne_stream_t stream);

/* ---- the point stage of the backward alone, over a scratch buffer that already holds d loss / d raw and the sample list
 * (filled by mne_render_forward_fused in training mode): decoder backward and plane / table gradient scatter over the
 * listed samples (loss.backward() of mp_slam/mapper.py:159 from d raw downwards).  Accumulates (+=) into the gradient
 * buffers of `scene` and, if non-NULL, WRITES g_rays_o / g_rays_d.  act as in mne_render_backward.  With act, decoder
 * weight gradients are computed only if the scene carries their buffers (g_sdf0 ..); a caller may leave them NULL here and
 * call mne_decoder_wgrad afterwards, e.g. on another stream beside the planes' optimiser pass. */
int mne_render_backward_points(const MneSceneDesc* scene, int32_t n_rays, int32_t n_samples, const float* rays_o,
                               const float* rays_d, const float* z_vals, void* scratch, float* g_rays_o, float* g_rays_d,
                               float* act, mne_stream_t stream);
/* Decoder weight gradients (+= into scene->g_sdf0 / g_sdf1 / g_col0 / g_col1, each 16-byte aligned) from the records in
 * act that a mne_render_backward / mne_render_backward_points call with the same scratch completed: dW = X^T . dY as a
 * split-K product over the listed samples on the tensor pipe.  max_ctas > 0 caps the grid (to share the GPU with the
 * kernel it runs beside); 0 = one CTA per SM.  Reads only scratch and act (not the rays), so the caller may already
 * refill the ray buffers.  Needs channels-last planes or the hash grid. */
int mne_decoder_wgrad(const MneSceneDesc* scene, int32_t n_rays, int32_t n_samples, void* scratch, const float* act,
                      int32_t max_ctas, mne_stream_t stream);

/* ---- backward without a caller workspace.  This entry point has no counterpart in the reference: it exists
 * for callers that must not allocate device memory (a host that forbids cudaMalloc around the call, a real-time
 * loop that cannot take the allocator's lock, a plugin that is only handed input and output buffers).  It is
 * mne_render_backward with the 20-byte-per-sample workspace (d loss / d raw, 16 B, plus one list entry, 4 B) taken from
 * an array built into the library image (MNE_STATIC_SCRATCH_SAMPLES * 20 bytes = 2.5 MiB, held on every GPU that loads the
 * library until the process exits) instead of the `scratch` argument; inputs and outputs are still the caller's.  It calls
 * no cudaMalloc / cudaFree or other allocating routine.
 *   - It takes no `act` record: the decoder's forward is always recomputed inside the backward kernels.  The two-table
 *     hash grid (encoder 1 with two_grid), whose backward needs that record, is therefore rejected here.
 *   - n_rays * n_samples must not exceed MNE_STATIC_SCRATCH_SAMPLES; larger batches are rejected with an error.
 *   - The first call on each device looks up the array's device address (a one-time, synchronous-with-the-host
 *     query); make that first call outside of any CUDA stream capture.
 *   - All calls on a device share the one array.  Calls that can overlap on the GPU (different streams, or
 *     threads launching concurrently) race on it: serialise them on one stream.  Nothing in the array has to be
 *     cleared between calls -- each launch writes every slot it later reads -- and its contents after a call are
 *     unspecified. */
#define MNE_STATIC_SCRATCH_SAMPLES 131072
int mne_render_backward_static(const MneSceneDesc* scene, const MneRenderCfg* rcfg, int32_t n_rays, int32_t n_samples,
                               const float* rays_o, const float* rays_d, const float* z_vals, const float* raw,
                               const float* target_rgb, const float* target_d, const double* acc,
                               const float* loss_grads, const float* g_rgb, const float* g_depth, const float* g_raw,
                               float* g_rays_o, float* g_rays_d, mne_stream_t stream);

/* ---- point queries: query_color_sdf / run_network_flat / query_sdf(embed, return_geo)
 * (model/scene_rep.py:232-301,319-331).  pts [M,3].  Any output may be NULL:
 * raw [M,4], geo [M,15], feat [M,F] (the embed=True output). */
int mne_query_points(const MneSceneDesc* scene, int64_t n_pts, const float* pts, float* raw, float* geo,
                     float* feat, mne_stream_t stream);

/* ---- hash-grid inspection: entry indices (within their level) of the 8 cell corners of each point at every
 * level, for coordinates x01 [M,3] in [0,1]^3: indices [M, n_levels, 8] int32.  The table lookups of
 * tiny-cuda-nn's grid encoding (encodings/grid.h grid_index / coherent prime hash), exposed so that parity
 * tests can compare the one integer quantity of the encoder bit for bit. */
int mne_hash_indices(const MneHashDesc* hash, int64_t n_pts, const float* x01, int32_t* indices, mne_stream_t stream);

/* ---- optimiser: torch.optim.Adam.step() + zero_grad() of the mapping loop (mneslam_mp.py:459-469,
 * mp_slam/mapper.py:161-162), dense semantics, one pass: 16 B/param read + 16 B/param written.  All four
 * arrays hold n floats in the same (dense) storage order.  The step count t (>= 1, already incremented) is
 * read from `step_dev` when non-NULL (lets a CUDA graph replay advance it with mne_adam_tick), else
 * `step_host`.  zero_grad != 0 also clears `grad`. */
int mne_adam_tick(int64_t* step_dev, mne_stream_t stream);
int mne_adam_step(float* param, float* grad, float* exp_avg, float* exp_avg_sq, int64_t n, float lr, float beta1,
                  float beta2, float eps, float weight_decay, const int64_t* step_dev, int64_t step_host,
                  int32_t zero_grad, mne_stream_t stream);

/* Same update for several parameter groups stored back to back in four flat buffers (identical offsets in param /
 * grad / exp_avg / exp_avg_sq): one launch for the whole map.  segment_ends[i] = exclusive end of group i in
 * elements (ascending, multiples of 4); lr / eps / weight_decay are HOST arrays of n_segments entries (the
 * parameter groups of create_optimizer, mneslam_mp.py:459-467).
 * `touched` (device, one byte per 4 parameters, zero-initialised by the caller, or NULL): marks the 16-byte groups
 * that have ever received a non-zero gradient.  A group that never did has m = v = g = 0, so dense Adam leaves
 * it exactly unchanged; with the map the kernel skips its parameter / moment traffic (groups with weight decay
 * are always updated).  The resulting parameters are bit-identical to the dense update. */
#define MNE_ADAM_MAX_SEGMENTS 16
int mne_adam_step_segments(float* param, float* grad, float* exp_avg, float* exp_avg_sq, int32_t n_segments,
                           const int64_t* segment_ends, const float* lr, const float* eps, const float* weight_decay,
                           float beta1, float beta2, const int64_t* step_dev, int64_t step_host, int32_t zero_grad,
                           uint8_t* touched, mne_stream_t stream);
/* The same with the backward pass's texel flags (MneSceneDesc.dirty with dirty_base = grad): dirty holds one byte per 32
 * floats of [0, dirty_floats) -- the plane gradients; the rest of the buffer is read densely.  An unflagged texel's
 * gradient is zero by contract (everything that adds into it flags it, and this kernel clears what it read: zero_grad
 * must be set), so it is not read at all: ~80 % of the 129 MB gradient pass of a room0 iteration.  The flags are
 * cleared here for the next iteration.  Results are bit-identical to mne_adam_step_segments.  (Measured on B200: this pass
 * 111 -> 105 us and setting the flags costs the backward kernel 4.5 us -- a wash on one GPU, so mapping.FusedMapper leaves it
 * off there; with several GPUs the same flags drive mne_dirty_push_data and are on by default.) */
int mne_adam_step_segments_dirty(float* param, float* grad, float* exp_avg, float* exp_avg_sq, int32_t n_segments,
                                 const int64_t* segment_ends, const float* lr, const float* eps, const float* weight_decay,
                                 float beta1, float beta2, const int64_t* step_dev, int64_t step_host, int32_t zero_grad,
                                 uint8_t* touched, uint8_t* dirty, int64_t dirty_floats, mne_stream_t stream);

/* ---- data-parallel optimiser step over NVLink peer memory: ONE kernel instead of NCCL reduce-scatter + Adam +
 * all-gather (the reference has no multi-GPU training at all; this serves the ray-sharded mapping of BASELINE.json
 * configs[3]).  param_ptrs / grad_ptrs: HOST arrays of world_size device pointers to every rank's full flat
 * parameter / gradient buffer (n_total floats each, padding included), all mapped into this process (e.g. torch
 * symmetric memory).  Ownership is block-cyclic: chunk c (`chunk` floats, a multiple of 32) belongs to rank
 * c % world_size -- contiguous slices would hand one rank the densely observed planes and another the barely
 * touched ones.  exp_avg / exp_avg_sq / touched hold this rank's n_total / world_size elements packed in ownership
 * order; segment_ends are GLOBAL offsets.  The kernel sums the slice's gradient over all ranks (peer loads), clears what it read, applies Adam and
 * stores the new parameters into every rank's buffer (peer stores).  When multicast_param / multicast_grad (the
 * NVLink-switch multicast addresses of the same two buffers) are given, the reduction and the broadcast happen in the
 * switch: multimem.ld_reduce for the gradient, multimem.st for the parameters and for clearing the gradient.
 * dirty (device pointer on THIS rank, or NULL): [world_size][n_total / 32 / world_size] flag bytes, slot r holding
 * rank r's "non-zero this step" flags of the 32-float texels this rank owns (packed in ownership order), pushed
 * there by every rank's mne_grad_dirty_map before the barrier.  With it the owner fetches a peer's 16-byte group only
 * where that peer flagged the texel -- a mapping iteration touches about a fifth of the map, so most of the NVLink
 * traffic of a dense reduction disappears.  chunk must be a power of two >= 32.  The caller must place a cross-rank
 * barrier before the call (all gradients and flags complete) and
 * after it (all parameters written). */
#define MNE_MAX_PEERS 16
int mne_adam_step_peers(int32_t world_size, int32_t rank, float* const* param_ptrs, float* const* grad_ptrs,
                        int64_t n_total, int32_t chunk, float* exp_avg, float* exp_avg_sq, int32_t n_segments,
                        const int64_t* segment_ends, const float* lr, const float* eps, const float* weight_decay,
                        float beta1, float beta2, const int64_t* step_dev, int64_t step_host, uint8_t* touched,
                        float* multicast_param, float* multicast_grad, const uint8_t* dirty, const float* inbox,
                        mne_stream_t stream);
/* inbox (device pointer on THIS rank, or NULL; needs dirty): [world_size][n_total / world_size] floats, slot r holding the
 * gradient texels of this rank's slice that rank r flagged, pushed there by rank r's mne_grad_push before the barrier.  With
 * it the kernel reads the peers' contributions from local memory (its own contribution still from its own gradient buffer)
 * and nothing has to be fetched -- or cleared -- over NVLink. */
/* For every 32-float texel of this rank's gradient buffer: flag = 1 iff it holds a non-zero value; the flag is
 * stored into the texel OWNER's flag array (dirty_ptrs[owner], HOST array of world_size peer-mapped device pointers)
 * at [rank][owner-local texel index] -- the layout mne_adam_step_peers reads.  world_size = 1: a plain local map. */
int mne_grad_dirty_map(const float* grad, int64_t n_total, int32_t chunk, int32_t world_size, int32_t rank,
                       uint8_t* const* dirty_ptrs, mne_stream_t stream);
/* mne_grad_dirty_map that also MOVES the data: every non-zero texel owned by another rank is stored into that owner's
 * inbox (inbox_ptrs[owner], HOST array of world_size peer-mapped device pointers to [world_size][n_total / world_size]
 * floats; slot [rank][owner-local texel], the flags' layout) and cleared in `grad`; texels this rank owns stay where they
 * are.  Peer stores are fire-and-forget, so the exchange is bound by NVLink bandwidth instead of by the round trips of
 * peer loads issued from inside the optimiser kernel.  Only the first n_scan elements (a multiple of 32) are examined:
 * what lies beyond -- e.g. the decoder weights, exchanged by an all-reduce of their own -- is neither flagged nor moved. */
int mne_grad_push(float* grad, int64_t n_total, int64_t n_scan, int32_t chunk, int32_t world_size, int32_t rank,
                  uint8_t* const* dirty_ptrs, float* const* inbox_ptrs, mne_stream_t stream);
/* Same hand-over without reading the gradient buffer: local_map (n_total / 32 bytes) holds the flags the backward pass
 * set through MneSceneDesc.dirty; they are pushed to the owners in the same layout and local_map is cleared.  Texels
 * at or beyond element dense_from (the decoder weights at the tail of the flat buffer) are always flagged. */
int mne_dirty_push(uint8_t* local_map, int64_t n_total, int64_t dense_from, int32_t chunk, int32_t world_size, int32_t rank,
                   uint8_t* const* dirty_ptrs, mne_stream_t stream);
/* mne_dirty_push and mne_grad_push in one: the flags in local_map (set by the backward pass, cleared here) name the texels
 * of `grad` that may be non-zero; only those are read, stored into their owners' inboxes and cleared (texels this rank
 * owns stay in place), and the flags are handed to the owners.  Texels at or beyond n_scan are left alone. */
int mne_dirty_push_data(uint8_t* local_map, float* grad, int64_t n_total, int64_t n_scan, int32_t chunk, int32_t world_size,
                        int32_t rank, uint8_t* const* dirty_ptrs, float* const* inbox_ptrs, mne_stream_t stream);

/* ---- small all-reduce over peer memory (no counterpart in the reference; replaces NCCL all-reduces of the 16 loss
 * accumulators and of the decoder's 25 KB of weight gradients, which cost ~50 us each at eight ranks and serialise with one
 * another).  mne_peer_scatter stores src[0, n_floats) at float offset offset_floats of EVERY rank's exchange buffer
 * (dst_ptrs: HOST array of world_size peer-mapped device pointers; the caller chooses offset = region + rank * n_floats);
 * after a cross-rank barrier of the caller's, mne_peer_sum adds the world_size slots ([world_size][n] of float or double,
 * local memory) in rank order into out -- bit-identical on every rank.  A second barrier must precede the next scatter
 * into the same region. */
int mne_peer_scatter(const float* src, int32_t n_floats, float* const* dst_ptrs, int64_t offset_floats, int32_t world_size,
                     mne_stream_t stream);
int mne_peer_sum(const void* slots, int32_t n, int32_t world_size, void* out, int32_t is_double, mne_stream_t stream);

/* Cross-rank barrier over NVLink peer memory as one small kernel (replaces the barriers of the exchange above; a kernel per
 * barrier that only stores to the peers and polls local memory).  pad_ptrs: HOST array of world_size peer-mapped device
 * pointers, entry r = rank r's signal pad of MNE_PEER_PAD_WORDS(n_channels) uint32 words, zero-initialised by the caller and
 * identical in layout on every rank: [n_channels][MNE_MAX_PEERS] arrival words, [n_channels] epochs, 1 error word.  Every
 * rank calls it the same number of times per channel; different channels may be in flight on different streams.  Memory
 * written by earlier work of the calling stream (to any GPU) is visible to the peers' work after their barrier returns.  The
 * wait is bounded (2^29 polls, about two minutes; mne_set_option("peer_barrier_polls_log2", v)): a rank that never arrives
 * raises the error word (last word of the pad) instead of hanging the device. */
#define MNE_PEER_PAD_WORDS(n_channels) ((n_channels) * (MNE_MAX_PEERS + 1) + 1)
int mne_peer_barrier(uint32_t* const* pad_ptrs, int32_t world_size, int32_t rank, int32_t channel, int32_t n_channels,
                     mne_stream_t stream);

/* ---- measurement aid (no counterpart in the reference): a kernel that ONLY gathers.  offsets [n_offsets] uint32 = float
 * offsets into table; mode 0: each offset names a 128-byte texel (a multiple of 32 floats), fetched by eight lanes with one
 * 128-bit load each -- the tri-plane taps of the forward pass, 24 per sample; mode 1: each offset names an 8-byte entry (a
 * multiple of 2 floats), one 64-bit load per lane -- the hash grid's corners.  bench.py times it on the batch's own tap
 * addresses: the gather rate this GPU sustains for that access pattern is the ceiling the encoder kernels are held against
 * (beside the HBM copy bandwidth).  sink: one float of device memory (never written in practice). */
int mne_bench_gather(const float* table, const uint32_t* offsets, int64_t n_offsets, int32_t mode, float* sink,
                     mne_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* MNESLAM_B200_H */
