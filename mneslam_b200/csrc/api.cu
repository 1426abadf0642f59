// Library plumbing: error reporting, argument validation, launch heuristics.
#include <stdarg.h>
#include <string.h>

#include "common.cuh"

namespace mne {

static thread_local char g_error[512] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_error, sizeof(g_error), fmt, ap);
    va_end(ap);
}

int sm_count() {
    static thread_local int cached_dev = -1, cached = 148;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 148;
    if (dev != cached_dev) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && n > 0) cached = n;
        cached_dev = dev;
    }
    return cached;
}

// Rays per warp-group: a group's (ray, sample) points are flattened and processed 32 at a time, so the
// lane efficiency is (r*S) / (32*ceil(r*S/32)); but larger groups mean fewer of them to spread over the
// warps.  Pick the r that minimises (rounds of groups per warp) x (chunks per group).
int choose_rays_per_group(int n_rays, int n_samples, int total_warps) {
    int best_r = 1;
    long best_cost = -1;
    for (int r = 1; r <= 8; ++r) {
        if (r > 1 && (long)r * n_samples > 1024) break;
        const long groups = (n_rays + r - 1) / r;
        const long rounds = (groups + total_warps - 1) / total_warps;
        const long chunks = ((long)r * n_samples + 31) / 32;
        const long cost = rounds * chunks;
        if (best_cost < 0 || cost < best_cost) { best_cost = cost; best_r = r; }
    }
    return best_r;
}

int impl_version();

int check_scene(const MneSceneDesc* sc, const char* who, bool) {
    MNE_CHECK_ARG(sc->encoder == 0 || sc->encoder == 1, "%s: encoder %d unknown (0 = tri-plane, 1 = hash grid)", who, sc->encoder);
    if (sc->encoder == 1) {
        MNE_CHECK_ARG(impl_version() == 2, "%s: the hash-grid encoder needs the tensor-pipe kernels (option impl = 2)", who);
        for (int g = 0; g < (sc->two_grid ? 2 : 1); ++g) {       // [0] sdf table, [1] colour table (grid.oneGrid: False)
            const MneHashDesc& h = sc->hash[g];
            MNE_CHECK_ARG(h.table && h.n_levels == 16 && h.n_features == 2, "%s: hash grid %d must have a table, 16 levels, 2 features", who, g);
            MNE_CHECK_ARG((((uintptr_t)h.table | (uintptr_t)h.grad) & 7) == 0, "%s: hash table %d must be 8-byte aligned", who, g);
            for (int l = 0; l < 16; ++l)
                MNE_CHECK_ARG(h.offsets[l + 1] > h.offsets[l] && h.level_res[l] >= 2 && h.level_scale[l] > 0.f,
                              "%s: level %d of hash grid %d is empty or has no scale / resolution", who, l, g);
        }
        MNE_CHECK_ARG(sc->n_bins == kBins, "%s: n_bins = %d, this library is built for %d", who, sc->n_bins, kBins);
        MNE_CHECK_ARG(sc->w_sdf0 && sc->w_sdf1 && sc->w_col0 && sc->w_col1, "%s: decoder weights missing", who);
        for (int a = 0; a < 3; ++a) MNE_CHECK_ARG(sc->bbox_ext[a] > 0.0, "%s: empty bound on axis %d", who, a);
        return 0;
    }
    MNE_CHECK_ARG(sc->n_bins == kBins, "%s: n_bins = %d, this library is built for %d", who, sc->n_bins, kBins);
    MNE_CHECK_ARG(sc->c_dim == kCDim, "%s: c_dim = %d, this library is built for %d", who, sc->c_dim, kCDim);
    MNE_CHECK_ARG(sc->w_sdf0 && sc->w_sdf1 && sc->w_col0 && sc->w_col1, "%s: decoder weights missing", who);
    const int n = sc->two_grid ? 12 : 6;
    for (int i = 0; i < n; ++i) {
        const MnePlaneDesc& p = sc->planes[i];
        MNE_CHECK_ARG(p.data, "%s: plane %d has no data", who, i);
        MNE_CHECK_ARG(p.H >= 1 && p.W >= 1, "%s: plane %d has size %dx%d", who, i, p.H, p.W);
        const int64_t extent = (int64_t)(p.H - 1) * p.stride_h + (int64_t)(p.W - 1) * p.stride_w;
        MNE_CHECK_ARG(p.stride_h > 0 && p.stride_w > 0 && p.stride_c > 0 && extent < ((int64_t)1 << 31),
                      "%s: plane %d strides unsupported (c=%lld h=%lld w=%lld)", who, i, (long long)p.stride_c,
                      (long long)p.stride_h, (long long)p.stride_w);
    }
    for (int a = 0; a < 3; ++a) {
        MNE_CHECK_ARG(sc->bound_ext[a] > 0.f && sc->bbox_ext[a] > 0.0, "%s: empty bound on axis %d", who, a);
    }
    return 0;
}

static int g_impl = 2;
int impl_version() { return g_impl; }
static int g_adam_spare_sms = 0;
int adam_spare_sms() { return g_adam_spare_sms; }
static int g_hash_level0_smem = 0;
int hash_level0_smem() { return g_hash_level0_smem; }
static int g_peer_barrier_polls_log2 = 29;
int peer_barrier_polls_log2() { return g_peer_barrier_polls_log2; }

}  // namespace mne

extern "C" int mne_set_option(const char* key, int value) {
    using namespace mne;
    MNE_CHECK_ARG(key, "mne_set_option: null key");
    if (strcmp(key, "impl") == 0) {
        MNE_CHECK_ARG(value == 1 || value == 2, "mne_set_option: impl must be 1 (fp32 CUDA-core kernels) or 2 (tensor-pipe kernels)");
        g_impl = value;
        return 0;
    }
    if (strcmp(key, "adam_spare_sms") == 0) {
        // the optimiser kernels are persistent grids that fill every SM; leaving a few SMs out lets small kernels of another
        // stream (the next iteration's ray assembly and sampling, mapping.PipelinedMapper) run under them
        MNE_CHECK_ARG(value >= 0 && value <= 64, "mne_set_option: adam_spare_sms must be in [0, 64]");
        g_adam_spare_sms = value;
        return 0;
    }
    if (strcmp(key, "hash_level0_smem") == 0) {
        // hash grid forward: stage level 0 of the table(s) (17^3 entries, 39 KB) in shared memory.  Off by default: measured slower
        // on B200 (the shared memory comes out of the L1 that serves the other coarse levels)
        MNE_CHECK_ARG(value == 0 || value == 1, "mne_set_option: hash_level0_smem must be 0 or 1");
        g_hash_level0_smem = value;
        return 0;
    }
    if (strcmp(key, "peer_barrier_polls_log2") == 0) {
        // mne_peer_barrier gives up (and raises the pad's error word) after 2^value polls of a peer's arrival word
        MNE_CHECK_ARG(value >= 10 && value <= 31, "mne_set_option: peer_barrier_polls_log2 must be in [10, 31]");
        g_peer_barrier_polls_log2 = value;
        return 0;
    }
    set_error("mne_set_option: unknown option '%s'", key);
    return 1;
}

extern "C" int mne_get_option(const char* key, int* value) {
    using namespace mne;
    MNE_CHECK_ARG(key && value, "mne_get_option: null argument");
    if (strcmp(key, "impl") == 0) { *value = g_impl; return 0; }
    if (strcmp(key, "adam_spare_sms") == 0) { *value = g_adam_spare_sms; return 0; }
    if (strcmp(key, "hash_level0_smem") == 0) { *value = g_hash_level0_smem; return 0; }
    if (strcmp(key, "peer_barrier_polls_log2") == 0) { *value = g_peer_barrier_polls_log2; return 0; }
    set_error("mne_get_option: unknown option '%s'", key);
    return 1;
}

extern "C" int mne_version(void) { return 100; }
extern "C" const char* mne_last_error(void) { return mne::g_error; }
