// icp_api.cu — context management and the C-ABI of include/icp_b200.h.
//
// Everything the reference keeps in CPU::ICP / GPU::ICP (src/cpu.hh:45-114, src/GPU/gpu.hh:41-104)
// lives in an icp_ctx as device-resident buffers; the per-iteration loop (src/cpu.cc:55-79) is
// enqueued on one stream with no host round trip: kernels test a device-side `done` flag.
// There is no CPU fallback anywhere in this file.
#include <cub/device/device_radix_sort.cuh>
#include <dlfcn.h>

#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/icp_b200.h"
#include "icp_common.cuh"

using namespace icpb;

namespace {

thread_local std::string g_err;

#define CK(call)                                                                                    \
    do {                                                                                            \
        cudaError_t e__ = (call);                                                                   \
        if (e__ != cudaSuccess) {                                                                   \
            g_err = std::string(#call) + ": " + cudaGetErrorString(e__);                            \
            return ICP_ERR_CUDA;                                                                    \
        }                                                                                           \
    } while (0)

template <class T>
int dalloc(T** p, size_t n)
{
    if (*p) { cudaFree(*p); *p = nullptr; }
    CK(cudaMalloc((void**)p, (n ? n : 1) * sizeof(T)));
    return ICP_OK;
}
template <class T>
void dfree(T*& p)
{
    if (p) cudaFree(p);
    p = nullptr;
}

// ---- minimal NCCL surface, resolved at run time from the NCCL the launcher already uses ---------
typedef struct { char internal[128]; } nccl_uid;
typedef void* nccl_comm;
struct NcclApi {
    void* lib = nullptr;
    int (*GetUniqueId)(nccl_uid*) = nullptr;
    int (*CommInitRank)(nccl_comm*, int, nccl_uid, int) = nullptr;
    int (*AllReduce)(const void*, void*, size_t, int, int, nccl_comm, cudaStream_t) = nullptr;
    int (*CommDestroy)(nccl_comm) = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
};
constexpr int NCCL_FLOAT64 = 8, NCCL_SUM = 0;

int load_nccl(NcclApi& api, const char* path)
{
    if (api.lib) return ICP_OK;
    const char* p = (path && *path) ? path : "libnccl.so.2";
    api.lib = dlopen(p, RTLD_NOW | RTLD_GLOBAL);
    if (!api.lib) { g_err = std::string("dlopen NCCL failed: ") + dlerror(); return ICP_ERR_COMM; }
    api.GetUniqueId = (int (*)(nccl_uid*))dlsym(api.lib, "ncclGetUniqueId");
    api.CommInitRank = (int (*)(nccl_comm*, int, nccl_uid, int))dlsym(api.lib, "ncclCommInitRank");
    api.AllReduce = (int (*)(const void*, void*, size_t, int, int, nccl_comm, cudaStream_t))dlsym(api.lib, "ncclAllReduce");
    api.CommDestroy = (int (*)(nccl_comm))dlsym(api.lib, "ncclCommDestroy");
    api.GetErrorString = (const char* (*)(int))dlsym(api.lib, "ncclGetErrorString");
    if (!api.GetUniqueId || !api.CommInitRank || !api.AllReduce || !api.CommDestroy) {
        g_err = "NCCL symbols missing";
        return ICP_ERR_COMM;
    }
    return ICP_OK;
}

NcclApi g_nccl;

}  // namespace


namespace icpb {
// lets the other translation units of the library report through icp_last_error()
void set_last_error(const std::string& msg) { g_err = msg; }
}  // namespace icpb

constexpr int TRACE_WORDS = 16 + 5 * 32;   // block 0's phase stamps + (SM id, entry, barrier, solve, exit) of the first 32 blocks

struct icp_ctx {
    int device = 0, sm_count = 148;
    cudaStream_t st = nullptr;
    ModelDev m;
    SceneDev s;
    NNScratch sc;
    double* d_stage = nullptr; size_t stage_cap = 0;      // AoS staging (doubles)
    double *d_partials = nullptr, *d_mom = nullptr, *d_xf = nullptr, *d_err_sum = nullptr, *d_out6 = nullptr;
    double* d_err_hist = nullptr; int err_hist_cap = 0;
    int* d_ctl = nullptr;
    int n_red_blocks = 1;
    uint32_t *d_keys = nullptr, *d_keys_alt = nullptr; int32_t* d_vals = nullptr; void* d_cub = nullptr;
    size_t cub_bytes = 0, sort_cap = 0;
    int32_t* d_vals_alt = nullptr;         // sorted values of the model's Morton sort (tree build)
    // how K1 searches: ICP_NN=brute|tree|auto / icp_set_nn_mode.  auto (default): brute force, except inside icp_run on clouds
    // too small for the brute-force filter to fill one wave of the GPU, which take the exact tree search (same indices)
    int nn_mode = ICP_NN_AUTO;
    bool auto_tree_now = false;            // auto mode: icp_run decided for the tree on the current clouds
    TreeDev tree;
    bool tree_built[2] = {false, false};   // [0] binary form, [1] wide form: built on first use, for the model currently set
    int tree_form = 0;                     // ICP_TREE: 0 auto, 1 wide (warp per point, 32-ary), 2 binary (thread per point)
    double h_Rm = 0.0;                     // geo[3] of the current model (the tree needs 1e-12 <= Rm <= 1e12)
    int32_t* d_idx_out = nullptr; size_t idx_out_cap = 0;
    bool have_model = false, have_scene = false, have_idx = false;
    icp_stats stats{};
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    std::vector<cudaEvent_t> ev_filter;
    nccl_comm comm = nullptr; int rank = 0, world = 1;
    int exact_mode = 0;                    // ICP_EXACT: 0 auto, 1 always inside nn_refine, 2 always the nn_exact kernel
    bool model0_nan = false;               // model point 0 has a NaN coordinate: the reference's first-minimum scan never leaves it
    size_t cap_nm = 0, cap_np = 0;         // allocated capacities (points): buffers only ever grow
    // fused alignment kernel (cooperative) and the peer-memory exchange it carries when world > 1
    int align_blocks_max = 148;
    PeerMailbox* d_mailbox = nullptr;
    PeerComm peer;
    bool peer_ready = false, force_nccl = false;
    bool in_group = false;                 // member of an icp_group (one process, several devices): mailboxes reached by peer access
    double group_np = 0.0;                 // global scene size of the group (what alignement_check compares)
    int align_mode = 0;                    // 0 auto, 1 fused, 2 split (ICP_ALIGN)
    bool is_lane = false;                  // a context of icp_batch_run: shares the GPU with other lanes
    unsigned long long* d_trace = nullptr; // ICP_PROFILE=1: phase time stamps of the last fused alignment kernel
    // early stop of the enqueue loop: the done flag of every chunk of iterations is copied to pinned memory
    int* h_progress = nullptr;             // mapped pinned word the loop's closing kernel writes: (iterations closed << 1) | converged
    int* d_progress = nullptr;             // its device address
    int* progress_arg = nullptr;           // what nn_pass hands to the filter kernel (set by icp_run only)
    int lookahead = 3;                     // ICP_LOOKAHEAD: iterations icp_run keeps enqueued ahead of the device (0 = all at once)
    bool profile = false;                  // ICP_PROFILE=1: events between all kernels of icp_run (adds ~us per iteration)
};

namespace {

int ensure_stage(icp_ctx* c, size_t doubles)
{
    if (c->stage_cap >= doubles) return ICP_OK;
    dfree(c->d_stage);
    CK(cudaMalloc((void**)&c->d_stage, doubles * sizeof(double)));
    c->stage_cap = doubles;
    return ICP_OK;
}

// Radix-sort scratch (keys, values, cub temporary) for n items; grows only.
int ensure_sort(icp_ctx* c, size_t n)
{
    if (c->sort_cap >= n) return ICP_OK;
    int r;
    if ((r = dalloc(&c->d_keys, n)) || (r = dalloc(&c->d_keys_alt, n)) || (r = dalloc(&c->d_vals, n)) ||
        (r = dalloc(&c->d_vals_alt, n)))
        return r;
    const size_t bytes = tree_sort_bytes((int)n);
    if (c->d_cub) cudaFree(c->d_cub);
    c->d_cub = nullptr;
    CK(cudaMalloc(&c->d_cub, bytes ? bytes : 1));
    c->cub_bytes = bytes;
    c->sort_cap = n;
    return ICP_OK;
}

int env_nn_mode()
{
    const char* e = getenv("ICP_NN");
    if (e && !strcmp(e, "tree")) return ICP_NN_TREE;
    if (e && !strcmp(e, "brute")) return ICP_NN_BRUTE;
    return ICP_NN_AUTO;
}

// auto mode's rule for icp_run: the brute-force filter's grid would be less than one full wave (2 CTAs per SM), i.e. the
// clouds are too small for that kernel to approach its roofline whatever is done (cow: 48 CTAs of 296; up to ~16 384 x
// 16 384 points), and an iteration is three kernel latencies.  There the tree search wins by 25 % resident and end to end
// including the build (cow 7 iterations 0.34 -> 0.25 ms, profiles/r02f_tree_wl_cow.json); above, brute force stays.
bool small_for_brute(const icp_ctx* c)
{
    const long long total = (c->s.np_pad / NN_TILE) * (c->m.nm_pad / NN_G);
    return total / (NN_TM / NN_G) < (long long)c->sm_count * NN_MIN_CTAS;
}

// Rm = 0: every model point coincides with c0; the surrogate is then the same for all of them and FP64 decides.
bool tree_active(const icp_ctx* c)
{
    const bool wanted = c->nn_mode == ICP_NN_TREE || (c->nn_mode == ICP_NN_AUTO && c->auto_tree_now);
    return wanted && (c->h_Rm == 0.0 || (c->h_Rm >= 1e-12 && c->h_Rm <= 1e12));
}

// Which form searches the current scene: ICP_TREE forces one; otherwise a warp per point while the scene is small.
bool tree_use_wide(const icp_ctx* c) { return c->tree_form == 1 || (c->tree_form == 0 && c->s.np <= 131072); }

// Builds the model's bounding-box hierarchy (the form the current scene needs) on first use after icp_set_model.
int ensure_tree(icp_ctx* c)
{
    if (!tree_active(c)) return ICP_OK;
    const bool wide = tree_use_wide(c);
    if (c->tree_built[wide]) return ICP_OK;
    int r;
    const int64_t leaves = (c->m.nm + TREE_LEAF - 1) / TREE_LEAF;
    int nl = 1;
    while (nl < leaves) nl *= 2;
    if (!wide && c->tree.cap_leaves < (size_t)nl) {
        if ((r = dalloc(&c->tree.nodes, (size_t)nl * 2 * 6)) || (r = dalloc(&c->tree.pts, (size_t)nl * TREE_LEAF))) return r;
        c->tree.cap_leaves = (size_t)nl;
    }
    c->tree.n_leaves = nl;
    const size_t wfloats = wtree_plan(c->m.nm, c->tree.w);
    if (wide && c->tree.cap_wboxes < wfloats) {
        if ((r = dalloc(&c->tree.wboxes, wfloats))) return r;
        c->tree.cap_wboxes = wfloats;
    }
    if (wide && c->tree.cap_wpts < (size_t)c->tree.w.count[0] * 32) {
        if ((r = dalloc(&c->tree.wpts, (size_t)c->tree.w.count[0] * 32))) return r;
        c->tree.cap_wpts = (size_t)c->tree.w.count[0] * 32;
    }
    if ((r = ensure_sort(c, (size_t)c->m.nm))) return r;
    void* kd = nullptr;
    if (c->m.nm >= kd_global_min_points()) {
        const size_t need = kd_scratch_bytes((int)c->m.nm, nullptr);
        if (c->tree.cap_kd < need) {
            if (c->tree.kd_scratch) cudaFree(c->tree.kd_scratch);
            c->tree.kd_scratch = nullptr;
            c->tree.cap_kd = 0;
            CK(cudaMalloc(&c->tree.kd_scratch, need));
            c->tree.cap_kd = need;
        }
        kd = c->tree.kd_scratch;
    }
    const int e = tree_build(c->m, c->tree, wide, c->d_keys, c->d_keys_alt, c->d_vals, c->d_vals_alt, c->d_cub, c->cub_bytes, kd, c->st);
    if (e != 0) { g_err = std::string("tree_build: ") + cudaGetErrorString((cudaError_t)e); return ICP_ERR_CUDA; }
    c->tree_built[wide] = true;
    return ICP_OK;
}

// Stream-K decomposition of the filter (see nn_filter_kernel): the grid is a whole number of waves of
// NN_MIN_CTAS CTAs per SM, each CTA owning an equal contiguous range of chunk-units.  One wave while a CTA's range
// stays below ~64 K model points x one scene tile; beyond that, more waves of such ranges, so that the
// hardware's dynamic block scheduling evens out SM-to-SM speed differences on long runs (measured on 2^20
// points: a single wave of 28 000-chunk ranges is 1.9 % slower than 28 waves).
// Small clouds: a CTA normally gets at least one staged model tile (NN_TM / NN_G chunks), but never so few CTAs that
// SMs stay empty while there are units to hand out (round 2, profiles/r02a_minunits.md: cow's filter 35 -> 25 us with one
// unit per CTA; with more units than SMs the staged-tile floor wins again).
// n_splits = the most partial results a scene tile can receive.
void choose_splits(icp_ctx* c)
{
    const long long n_tiles = c->s.np_pad / NN_TILE;
    const long long cpt = c->m.nm_pad / NN_G;
    const long long total = n_tiles * cpt;
    const long long wave = (long long)c->sm_count * NN_MIN_CTAS;
    const long long per_cta = 65536 / NN_G;             // ~64 K model points x 1 scene tile per CTA and wave
    long long waves = (total + per_cta * wave - 1) / (per_cta * wave);
    if (waves < 1) waves = 1;
    long long n_cta = wave * waves;
    long long cap = total / (NN_TM / NN_G);             // one staged tile per CTA ...
    const long long one_per_sm = total < c->sm_count ? total : c->sm_count;
    if (cap < one_per_sm) cap = one_per_sm;             // ... unless that leaves SMs without work
    if (n_cta > cap) n_cta = cap;
    if (n_cta < 1) n_cta = 1;
    c->sc.n_cta = (int)n_cta;
    c->sc.n_splits = (int)((n_cta + n_tiles - 1) / n_tiles + 1);
}

int ensure_scratch(icp_ctx* c)
{
    choose_splits(c);
    const size_t need = (size_t)c->sc.n_splits * c->s.np_pad;
    if (c->sc.cap_points < need) {
        int r;
        if ((r = dalloc(&c->sc.best, need))) return r;
        if ((r = dalloc(&c->sc.runner, need))) return r;
        if ((r = dalloc(&c->sc.chunk, need))) return r;
        c->sc.cap_points = need;
    }
    return ICP_OK;
}

int allreduce(icp_ctx* c, double* buf, size_t n)
{
    if (!c->comm) return ICP_OK;
    int r = g_nccl.AllReduce(buf, buf, n, NCCL_FLOAT64, NCCL_SUM, c->comm, c->st);
    if (r != 0) {
        g_err = std::string("ncclAllReduce: ") + (g_nccl.GetErrorString ? g_nccl.GetErrorString(r) : "error");
        return ICP_ERR_COMM;
    }
    return ICP_OK;
}

// An undecided scene point is rescanned over the chunk ranges of the slots that can hold its nearest neighbour (typically
// one or two).  While such a range is short the block of nn_refine that met the point does it on the spot (no nn_exact
// launch: 3 kernels per iteration); long ranges are queued for nn_exact, which spreads the points over the whole GPU.
constexpr long long INLINE_EXACT_MAX_RANGE = 16384;     // model points per slot range
int inline_exact(const icp_ctx* c)
{
    if (c->exact_mode) return c->exact_mode == 1;     // ICP_EXACT=inline|kernel: test hook, both forms on any input
    const long long total = (c->s.np_pad / NN_TILE) * (c->m.nm_pad / NN_G);
    const long long per_cta = (total + c->sc.n_cta - 1) / (c->sc.n_cta > 0 ? c->sc.n_cta : 1);
    return per_cta * NN_G <= INLINE_EXACT_MAX_RANGE;
}

// One closest-matrix pass (K1) on the stream.  use_prev: idx[] holds valid matches of the same model (tree mode
// starts every search from them).
void nn_pass(icp_ctx* c, int honor_done, int use_prev, cudaEvent_t e0, cudaEvent_t e1, cudaEvent_t e2 = nullptr,
             cudaEvent_t e3 = nullptr)
{
    if (e0) cudaEventRecord(e0, c->st);
    if (tree_active(c) && c->tree_built[tree_use_wide(c)]) {
        // measured (profiles/r01h_tree.md): a warp per point wins while the grid is a few waves (cow 0.014 vs 0.052 ms,
        // horse 128 vs 285 us per search: the slowest thread no longer sets the kernel time); a thread per point
        // wins once there are enough points to fill the machine either way (2^18: 0.19 vs 0.47 ms, 2^22: 2x)
        if (tree_use_wide(c))
            launch_nn_wtree(c->m, c->s, c->tree, use_prev, c->d_ctl, honor_done, c->st, c->progress_arg);
        else
            launch_nn_tree(c->m, c->s, c->tree, use_prev, c->d_ctl, honor_done, c->st, c->progress_arg);
        c->stats.kernels_launched += 1;
    } else {
        const int inl = inline_exact(c);
        launch_nn_filter(c->m, c->s, c->sc, c->d_ctl, honor_done, c->st, c->progress_arg);
        if (e1) cudaEventRecord(e1, c->st);
        launch_nn_refine(c->m, c->s, c->sc, c->d_ctl, honor_done, inl, c->st);
        if (e2) cudaEventRecord(e2, c->st);
        e1 = e2 = nullptr;
        if (!inl)
            launch_nn_exact(c->m, c->s, c->sc, c->d_ctl, honor_done, c->sm_count, c->st, c->profile ? c->d_trace : nullptr);
        c->stats.kernels_launched += inl ? 2 : 3;
    }
    // Model point 0 is NaN: d(0) is NaN for every scene point and minCoeff's strict `<` never replaces it
    // (src/cpu.cc:21-22), so the reference matches everything to index 0; the kernels above ignore NaN candidates.
    if (c->model0_nan) {
        cudaMemsetAsync(c->s.idx, 0, (size_t)c->s.np * sizeof(int32_t), c->st);
        launch_gather_matches(c->s, c->m, c->st);
    }
    if (e1) cudaEventRecord(e1, c->st);
    if (e2) cudaEventRecord(e2, c->st);
    if (e3) cudaEventRecord(e3, c->st);
}

// Alignment step (K2 -> K3 -> K4) of one iteration.  Two device-only forms, both able to carry the multi-GPU
// exchange INSIDE the kernels once peer mailboxes are attached:
//   fused  one cooperative launch with a grid barrier (align_fused_kernel)
//   split  moments_kernel (last block: reduce, exchange, Horn solve) + transform_kernel (last block: reduce,
//          exchange, close the iteration): 2 launches, no grid barrier, no co-residency requirement
// ICP_ALIGN=fused|split forces one; the default picks per context (see align_step).
// A communicator without mailboxes (or ICP_COMM=nccl) takes the NCCL path: separate kernels + ncclAllReduce.
bool nccl_path(const icp_ctx* c) { return c->comm && (!c->peer_ready || c->force_nccl); }
// the in-kernel exchange: peer mailboxes attached (CUDA IPC between processes, or peer access inside an icp_group)
const PeerComm* peers_of(const icp_ctx* c) { return (c->peer_ready && !nccl_path(c) && (c->comm || c->in_group)) ? &c->peer : nullptr; }

int fused_blocks(const icp_ctx* c)
{
    int64_t b = (c->s.np + RED_THREADS - 1) / RED_THREADS;
    if (b > c->align_blocks_max) b = c->align_blocks_max;
    return b < 1 ? 1 : (int)b;
}

// Spins on the loop's progress word until `need` iterations are closed; `enqueued` = iterations on the stream.
// Returns 1 = stop enqueueing (converged, or the device stopped closing iterations), 0 = go on, < 0 = icp_status.
int wait_progress(icp_ctx* c, int need, int enqueued)
{
    volatile int* p = c->h_progress;
    for (unsigned spin = 1;; ++spin) {
        const int v = *p;
        if (v & 1) return 1;
        if ((v >> 1) >= need) return 0;
        if ((spin & 0x3ffu) == 0) {
            const cudaError_t q = cudaStreamQuery(c->st);
            if (q == cudaSuccess) {     // drained (the word lags one iteration: it is written when the next one starts)
                int ctl[2];
                CK(cudaMemcpy(ctl, c->d_ctl, sizeof(ctl), cudaMemcpyDeviceToHost));
                // all enqueued iterations closed and not converged: the host was slow, go on; otherwise the loop
                // converged or an error flag in ctl turned the kernels into no-ops
                return (!ctl[CTL_DONE] && ctl[CTL_ITERS] == enqueued) ? 0 : 1;
            }
            if (q != cudaErrorNotReady) { g_err = std::string("icp_run: ") + cudaGetErrorString(q); return ICP_ERR_CUDA; }
        }
#if defined(__x86_64__)
        __builtin_ia32_pause();
#endif
    }
}

// Enqueues one alignment step (not the NCCL path).
// Returns the number of kernels launched (> 0) or a negative icp_status.
int align_step(icp_ctx* c, int honor_done, int in_place, int finish, double n_global, double threshold)
{
    const PeerComm* peers = peers_of(c);
    int launched;
    // auto: the cooperative kernel where it was measured faster (one GPU, nothing else on it: 11.17 vs 11.27 ms per
    // horse loop), the split pair with peers (its second grid barrier costs more than a launch: 11.65 vs 11.61 ms
    // at N = 2) and in batch lanes (no co-residency requirement between concurrent contexts).
    // (Measured again in round 2 with the 32-byte model records: the fused kernel wins at every size on one GPU -
    // horse 18.9 vs 25.1 us, 2^18 points 23.2 vs 35.9 us, 2^22 points 158 vs 181 us, profiles/r02b_prof_tail_*.log.)
    const bool fused = c->align_mode == 1 || (c->align_mode == 0 && !c->comm && !c->in_group && !c->is_lane);
    if (fused) {
        const int e = launch_align_fused(c->m, c->s, c->d_partials, fused_blocks(c), c->d_mom, c->d_xf, c->d_err_sum,
                                         c->d_ctl, honor_done, in_place, finish, n_global, threshold, c->d_err_hist, peers,
                                         c->profile ? c->d_trace : nullptr, c->st);
        if (e != 0) { g_err = std::string("align_fused launch: ") + cudaGetErrorString((cudaError_t)e); return ICP_ERR_CUDA; }
        launched = 1;
    } else {
        launch_moments(c->m, c->s, c->d_partials, c->n_red_blocks, c->d_mom, c->d_ctl, honor_done, c->d_xf, c->st, peers);
        launch_transform(c->m, c->s, c->d_xf, in_place, c->d_partials, c->n_red_blocks, c->d_err_sum, c->d_ctl, honor_done,
                         finish, n_global, threshold, c->d_err_hist, c->st, peers);
        launched = 2;
    }
    return launched;
}

int reset_ctl(icp_ctx* c)
{
    CK(cudaMemsetAsync(c->d_ctl, 0, CTL_N * sizeof(int), c->st));
    return ICP_OK;
}

int set_identity_xf(icp_ctx* c)
{
    double xf[XF_N] = {0};
    xf[0] = 1.0; xf[1] = 1.0; xf[5] = 1.0; xf[9] = 1.0;
    CK(cudaMemcpyAsync(c->d_xf, xf, sizeof(xf), cudaMemcpyHostToDevice, c->st));
    CK(cudaStreamSynchronize(c->st));
    return ICP_OK;
}

int pull_amb_stats(icp_ctx* c)
{
    int ctl[CTL_N];
    CK(cudaMemcpyAsync(ctl, c->d_ctl, sizeof(ctl), cudaMemcpyDeviceToHost, c->st));
    CK(cudaStreamSynchronize(c->st));
    c->stats.rerank_full += ctl[CTL_AMB_TOTAL];
    c->stats.rerank_chunk += c->s.np - ctl[CTL_AMB_TOTAL];
    if (ctl[CTL_BOUNDS]) { g_err = "bounds check failed at csrc line " + std::to_string(ctl[CTL_BOUNDS]); return ICP_ERR_CUDA; }
    return ICP_OK;
}

}  // namespace

// =================================================================================================
extern "C" {

const char* icp_last_error(void) { return g_err.c_str(); }
const char* icp_version(void) { return "icp_b200 0.1 (sm_100a)"; }

static int create_body(icp_ctx* c, int device)
{
    CK(cudaSetDevice(device));
    c->device = device;
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, device));
    c->sm_count = prop.multiProcessorCount;
    CK(cudaStreamCreateWithFlags(&c->st, cudaStreamNonBlocking));
    CK(cudaEventCreate(&c->ev0));
    CK(cudaEventCreate(&c->ev1));
    int r;
    if ((r = dalloc(&c->d_mom, MOM_N)) || (r = dalloc(&c->d_xf, XF_N)) || (r = dalloc(&c->d_err_sum, 1)) ||
        (r = dalloc(&c->d_out6, 8)) || (r = dalloc(&c->d_ctl, CTL_N)) || (r = dalloc(&c->m.geo, 16)) ||
        (r = dalloc(&c->d_partials, (size_t)c->sm_count * 4 * 18)) || (r = dalloc(&c->d_trace, TRACE_WORDS)))
        return r;
    CK(cudaMemset(c->d_trace, 0, TRACE_WORDS * sizeof(unsigned long long)));
    c->align_blocks_max = align_fused_max_blocks(c->sm_count);
    if (const char* e = getenv("ICP_COMM")) c->force_nccl = !strcmp(e, "nccl");
    if (const char* e = getenv("ICP_ALIGN")) c->align_mode = !strcmp(e, "fused") ? 1 : (!strcmp(e, "split") ? 2 : 0);
    if (const char* e = getenv("ICP_PROFILE")) c->profile = atoi(e) != 0;
    if (const char* e = getenv("ICP_LOOKAHEAD")) c->lookahead = atoi(e);
    if (const char* e = getenv("ICP_TREE")) c->tree_form = !strcmp(e, "wide") ? 1 : (!strcmp(e, "binary") ? 2 : 0);
    c->nn_mode = env_nn_mode();
    if (const char* e = getenv("ICP_EXACT")) c->exact_mode = !strcmp(e, "inline") ? 1 : (!strcmp(e, "kernel") ? 2 : 0);
    CK(cudaHostAlloc((void**)&c->h_progress, sizeof(int), cudaHostAllocMapped));
    CK(cudaHostGetDevicePointer((void**)&c->d_progress, c->h_progress, 0));
    CK(cudaMemset(c->d_ctl, 0, CTL_N * sizeof(int)));
    return set_identity_xf(c);
}

int icp_create(icp_ctx** out, int device)
{
    if (!out) return ICP_ERR_INVALID;
    *out = nullptr;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0 || device < 0 || device >= ndev) {
        g_err = "no usable CUDA device (this library has no CPU fallback)";
        cudaGetLastError();
        return ICP_ERR_NO_DEVICE;
    }
    icp_ctx* c = new icp_ctx();
    c->device = device;
    const int r = create_body(c, device);
    if (r) {                                  // frees whatever was created so far; g_err keeps the cause
        const std::string why = g_err;
        icp_destroy(c);
        g_err = why;
        return r;
    }
    *out = c;
    return ICP_OK;
}

void icp_destroy(icp_ctx* c)
{
    if (!c) return;
    cudaSetDevice(c->device);
    if (c->st) cudaStreamSynchronize(c->st);
    if (c->comm && g_nccl.CommDestroy) g_nccl.CommDestroy(c->comm);
    for (int r = 0; r < c->peer.world; ++r)
        if (c->peer_ready && !c->in_group && r != c->peer.rank && c->peer.box[r]) cudaIpcCloseMemHandle(c->peer.box[r]);
    dfree(c->d_mailbox); dfree(c->d_trace);
    if (c->h_progress) cudaFreeHost(c->h_progress);
    dfree(c->m.m64); dfree(c->m.mhat); dfree(c->m.geo);
    dfree(c->s.p64x); dfree(c->s.p64y); dfree(c->s.p64z); dfree(c->s.q64x); dfree(c->s.q64y); dfree(c->s.q64z);
    dfree(c->s.phx); dfree(c->s.phy); dfree(c->s.phz); dfree(c->s.perm); dfree(c->s.idx);
    dfree(c->s.y64x); dfree(c->s.y64y); dfree(c->s.y64z);
    dfree(c->sc.best); dfree(c->sc.runner); dfree(c->sc.chunk); dfree(c->sc.amb_list); dfree(c->sc.amb_thr);
    dfree(c->d_stage); dfree(c->d_partials); dfree(c->d_mom); dfree(c->d_xf); dfree(c->d_err_sum); dfree(c->d_out6);
    dfree(c->d_err_hist); dfree(c->d_ctl); dfree(c->d_keys); dfree(c->d_keys_alt); dfree(c->d_vals);
    dfree(c->d_idx_out); dfree(c->d_vals_alt); dfree(c->tree.nodes); dfree(c->tree.pts); dfree(c->tree.wboxes); dfree(c->tree.wpts);
    if (c->tree.kd_scratch) cudaFree(c->tree.kd_scratch);
    if (c->d_cub) cudaFree(c->d_cub);
    for (auto e : c->ev_filter) cudaEventDestroy(e);
    if (c->ev0) cudaEventDestroy(c->ev0);
    if (c->ev1) cudaEventDestroy(c->ev1);
    if (c->st) cudaStreamDestroy(c->st);
    delete c;
}

int icp_set_model(icp_ctx* c, const double* m_xyz, int64_t nm)
{
    if (!c || !m_xyz || nm < 1 || nm > 0x7fffff00LL) { g_err = "icp_set_model: bad arguments"; return ICP_ERR_INVALID; }
    CK(cudaSetDevice(c->device));
    int r;
    if ((r = ensure_stage(c, (size_t)nm * 3))) return r;
    CK(cudaMemcpyAsync(c->d_stage, m_xyz, (size_t)nm * 3 * sizeof(double), cudaMemcpyHostToDevice, c->st));
    c->m.nm = nm;
    c->m.nm_pad = (nm + NN_TM - 1) / NN_TM * NN_TM;
    if (c->cap_nm < (size_t)c->m.nm_pad) {       // grow only: cudaFree/cudaMalloc synchronise the whole device
        c->cap_nm = 0;
        c->have_model = false;
        if ((r = dalloc(&c->m.m64, (size_t)c->m.nm_pad)) || (r = dalloc(&c->m.mhat, (size_t)c->m.nm_pad))) return r;
        c->cap_nm = (size_t)c->m.nm_pad;
    }
    launch_bbox(c->d_stage, nm, c->m.geo + 4, c->d_partials, c->sm_count * 2, c->d_ctl, c->st);
    launch_model_geo(c->m.geo, c->st);
    launch_model_build(c->d_stage, c->m, c->st);
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(&c->h_Rm, c->m.geo + 3, sizeof(double), cudaMemcpyDeviceToHost, c->st));
    CK(cudaStreamSynchronize(c->st));
    if (!(c->h_Rm == c->h_Rm) || c->h_Rm > 1.7e308) {
        // an infinite coordinate (or no finite point at all) leaves no finite bounding box: the FP32 shadow would be all
        // NaN and every search would degenerate into np x nm FP64 evaluations.  NaN coordinates alone are tolerated
        // (the bounding box ignores them; such a point is never a nearest neighbour, see nn_pass for index 0).
        c->have_model = false;
        g_err = "icp_set_model: the model has infinite coordinates (or no finite point)";
        return ICP_ERR_INVALID;
    }
    c->tree_built[0] = c->tree_built[1] = false;
    c->model0_nan = m_xyz[0] != m_xyz[0] || m_xyz[1] != m_xyz[1] || m_xyz[2] != m_xyz[2];
    c->have_model = true;
    c->have_idx = false;
    c->stats.nm = nm;
    if (c->have_scene) return icp_reset_scene(c);   // FP32 shadow depends on the model's c0
    return ICP_OK;
}

int icp_set_scene(icp_ctx* c, const double* p_xyz, int64_t np)
{
    if (!c || !p_xyz || np < 1 || np > 0x7fffff00LL) { g_err = "icp_set_scene: bad arguments"; return ICP_ERR_INVALID; }
    if (!c->have_model) { g_err = "icp_set_scene: set the model first"; return ICP_ERR_INVALID; }
    CK(cudaSetDevice(c->device));
    int r;
    if ((r = ensure_stage(c, (size_t)np * 3 + 8))) return r;
    CK(cudaMemcpyAsync(c->d_stage, p_xyz, (size_t)np * 3 * sizeof(double), cudaMemcpyHostToDevice, c->st));
    const int64_t np_pad = (np + NN_TILE - 1) / NN_TILE * NN_TILE;
    c->s.np = np; c->s.np_pad = np_pad;
    if (c->cap_np < (size_t)np_pad) {            // grow only
        c->cap_np = 0;                           // stays 0 (and the scene unset) if an allocation below fails
        c->have_scene = false;
        if ((r = dalloc(&c->s.p64x, (size_t)np_pad)) || (r = dalloc(&c->s.p64y, (size_t)np_pad)) ||
            (r = dalloc(&c->s.p64z, (size_t)np_pad)) || (r = dalloc(&c->s.q64x, (size_t)np_pad)) ||
            (r = dalloc(&c->s.q64y, (size_t)np_pad)) || (r = dalloc(&c->s.q64z, (size_t)np_pad)) ||
            (r = dalloc(&c->s.phx, (size_t)np_pad)) || (r = dalloc(&c->s.phy, (size_t)np_pad)) ||
            (r = dalloc(&c->s.phz, (size_t)np_pad)) || (r = dalloc(&c->s.perm, (size_t)np_pad)) ||
            (r = dalloc(&c->s.idx, (size_t)np_pad)) || (r = dalloc(&c->s.y64x, (size_t)np_pad)) ||
            (r = dalloc(&c->s.y64y, (size_t)np_pad)) || (r = dalloc(&c->s.y64z, (size_t)np_pad)) ||
            (r = dalloc(&c->sc.amb_list, (size_t)np_pad)) ||
            (r = dalloc(&c->sc.amb_thr, (size_t)np_pad)))
            return r;
        c->cap_np = (size_t)np_pad;
    }
    // Morton sort of the scene (once per scene; a similarity transform keeps tiles compact)
    if ((r = ensure_sort(c, (size_t)np))) return r;
    double* bbox = c->d_stage + (size_t)np * 3;
    launch_bbox(c->d_stage, np, bbox, c->d_partials, c->sm_count * 2, c->d_ctl, c->st);
    launch_scene_keys(c->d_stage, np, bbox, c->d_keys, c->d_vals, c->st);
    size_t bytes = c->cub_bytes;
    CK(cub::DeviceRadixSort::SortPairs(c->d_cub, bytes, c->d_keys, c->d_keys_alt, c->d_vals, c->s.perm, (int)np, 0, 30,
                                       c->st));
    launch_scene_build(c->d_stage, c->m, c->s, c->st);
    CK(cudaGetLastError());
    c->n_red_blocks = reduce_blocks_for(np, c->sm_count);
    if ((r = ensure_scratch(c))) return r;
    if ((r = set_identity_xf(c))) return r;
    c->have_scene = true;
    c->have_idx = false;
    c->stats.np = np;
    return ICP_OK;
}

int icp_reset_scene(icp_ctx* c)
{
    if (!c || !c->have_scene) { g_err = "icp_reset_scene: no scene"; return ICP_ERR_INVALID; }
    CK(cudaSetDevice(c->device));
    launch_scene_restore(c->m, c->s, c->st);
    CK(cudaGetLastError());
    int r;
    if ((r = ensure_scratch(c))) return r;
    c->have_idx = false;
    return set_identity_xf(c);
}

int icp_set_nn_mode(icp_ctx* c, int mode)
{
    if (!c || (mode != ICP_NN_BRUTE && mode != ICP_NN_TREE && mode != ICP_NN_AUTO)) { g_err = "icp_set_nn_mode: bad arguments"; return ICP_ERR_INVALID; }
    c->nn_mode = mode;
    c->auto_tree_now = false;
    return ICP_OK;
}

int icp_get_nn_mode(icp_ctx* c)
{
    if (!c) return ICP_ERR_INVALID;
    return tree_active(c) || (c->nn_mode == ICP_NN_TREE && !c->have_model) ? ICP_NN_TREE : ICP_NN_BRUTE;
}

int icp_get_scene(icp_ctx* c, double* out)
{
    if (!c || !out || !c->have_scene) { g_err = "icp_get_scene: no scene"; return ICP_ERR_INVALID; }
    CK(cudaSetDevice(c->device));
    int r;
    if ((r = ensure_stage(c, (size_t)c->s.np * 3))) return r;
    launch_scene_export(c->s, c->d_stage, c->st);
    CK(cudaMemcpyAsync(out, c->d_stage, (size_t)c->s.np * 3 * sizeof(double), cudaMemcpyDeviceToHost, c->st));
    CK(cudaStreamSynchronize(c->st));
    return ICP_OK;
}

int icp_get_indices(icp_ctx* c, int32_t* idx_out)
{
    if (!c || !idx_out || !c->have_idx) { g_err = "icp_get_indices: no correspondences yet"; return ICP_ERR_INVALID; }
    CK(cudaSetDevice(c->device));
    int r;
    if (c->idx_out_cap < (size_t)c->s.np) {
        if ((r = dalloc(&c->d_idx_out, (size_t)c->s.np))) return r;
        c->idx_out_cap = (size_t)c->s.np;
    }
    launch_indices_export(c->s, c->m, c->d_idx_out, nullptr, c->st);
    CK(cudaMemcpyAsync(idx_out, c->d_idx_out, (size_t)c->s.np * sizeof(int32_t), cudaMemcpyDeviceToHost, c->st));
    CK(cudaStreamSynchronize(c->st));
    return ICP_OK;
}

int icp_closest(icp_ctx* c, int32_t* idx_out, double* Y_out)
{
    if (!c || !c->have_model || !c->have_scene) { g_err = "icp_closest: model and scene required"; return ICP_ERR_INVALID; }
    CK(cudaSetDevice(c->device));
    int r;
    if ((r = reset_ctl(c))) return r;
    c->auto_tree_now = false;             // auto mode: a single search does not pay for building the hierarchy
    if ((r = ensure_tree(c))) return r;
    c->stats.kernels_launched = 0;
    CK(cudaEventRecord(c->ev0, c->st));
    nn_pass(c, 0, c->have_idx ? 1 : 0, nullptr, nullptr);
    CK(cudaEventRecord(c->ev1, c->st));
    CK(cudaGetLastError());
    c->have_idx = true;
    c->stats.nn_calls += 1;
    c->stats.nn_pairs += c->s.np * c->m.nm;
    if (Y_out) {
        if ((r = ensure_stage(c, (size_t)c->s.np * 3))) return r;
    }
    if (idx_out && c->idx_out_cap < (size_t)c->s.np) {
        if ((r = dalloc(&c->d_idx_out, (size_t)c->s.np))) return r;
        c->idx_out_cap = (size_t)c->s.np;
    }
    if (idx_out || Y_out) {
        launch_indices_export(c->s, c->m, idx_out ? c->d_idx_out : nullptr, Y_out ? c->d_stage : nullptr, c->st);
        if (idx_out)
            CK(cudaMemcpyAsync(idx_out, c->d_idx_out, (size_t)c->s.np * sizeof(int32_t), cudaMemcpyDeviceToHost, c->st));
        if (Y_out)
            CK(cudaMemcpyAsync(Y_out, c->d_stage, (size_t)c->s.np * 3 * sizeof(double), cudaMemcpyDeviceToHost, c->st));
    }
    if ((r = pull_amb_stats(c))) return r;
    cudaEventElapsedTime(&c->stats.ms_total, c->ev0, c->ev1);
    c->stats.ms_filter = c->stats.ms_total;
    return ICP_OK;
}

static int need_idx(icp_ctx* c, const char* who)
{
    if (!c || !c->have_model || !c->have_scene || !c->have_idx) {
        g_err = std::string(who) + ": needs a model, a scene and correspondences (icp_closest first)";
        return ICP_ERR_INVALID;
    }
    return ICP_OK;
}

int icp_centroid(icp_ctx* c, double mu_p[3], double mu_y[3])
{
    int r;
    if ((r = need_idx(c, "icp_centroid"))) return r;
    CK(cudaSetDevice(c->device));
    launch_moments(c->m, c->s, c->d_partials, c->n_red_blocks, c->d_mom, c->d_ctl, 0, nullptr, c->st, peers_of(c));
    if (!peers_of(c) && (r = allreduce(c, c->d_mom, MOM_N))) return r;
    launch_centroid_export(c->m, c->d_mom, c->d_out6, c->st);
    double h[6];
    CK(cudaMemcpyAsync(h, c->d_out6, sizeof(h), cudaMemcpyDeviceToHost, c->st));
    CK(cudaStreamSynchronize(c->st));
    if (mu_p) memcpy(mu_p, h, 24);
    if (mu_y) memcpy(mu_y, h + 3, 24);
    return ICP_OK;
}

int icp_get_transform(icp_ctx* c, double* s, double R[9], double t[3])
{
    if (!c) return ICP_ERR_INVALID;
    CK(cudaSetDevice(c->device));
    double xf[XF_N];
    CK(cudaMemcpyAsync(xf, c->d_xf, sizeof(xf), cudaMemcpyDeviceToHost, c->st));
    CK(cudaStreamSynchronize(c->st));
    if (s) *s = xf[0];
    if (R) memcpy(R, xf + 1, 72);
    if (t) memcpy(t, xf + 10, 24);
    return ICP_OK;
}

int icp_find_alignment(icp_ctx* c, double* s, double R[9], double t[3], double* err_align)
{
    int r;
    if ((r = need_idx(c, "icp_find_alignment"))) return r;
    CK(cudaSetDevice(c->device));
    int launched = 3;
    if (!nccl_path(c)) {
        if ((launched = align_step(c, 0, 0, 0, 1.0, 0.0)) < 0) return launched;
    } else {
        launch_moments(c->m, c->s, c->d_partials, c->n_red_blocks, c->d_mom, c->d_ctl, 0, nullptr, c->st);
        if ((r = allreduce(c, c->d_mom, MOM_N))) return r;
        launch_solve(c->m, c->d_mom, c->d_xf, c->d_ctl, 0, c->st);
        launch_transform(c->m, c->s, c->d_xf, 0, c->d_partials, c->n_red_blocks, c->d_err_sum, c->d_ctl, 0, 0, 1.0, 0.0,
                         nullptr, c->st);
        if ((r = allreduce(c, c->d_err_sum, 1))) return r;
    }
    CK(cudaGetLastError());
    double e;
    CK(cudaMemcpyAsync(&e, c->d_err_sum, sizeof(e), cudaMemcpyDeviceToHost, c->st));
    if ((r = icp_get_transform(c, s, R, t))) return r;
    if (err_align) *err_align = e;
    c->stats.kernels_launched = launched;
    return ICP_OK;
}

int icp_err_compute(icp_ctx* c, int in_place, double* err)
{
    int r;
    if ((r = need_idx(c, "icp_err_compute"))) return r;
    CK(cudaSetDevice(c->device));
    launch_transform(c->m, c->s, c->d_xf, in_place ? 1 : 0, c->d_partials, c->n_red_blocks, c->d_err_sum, c->d_ctl, 0,
                     0, 1.0, 0.0, nullptr, c->st, peers_of(c));
    if (!peers_of(c) && (r = allreduce(c, c->d_err_sum, 1))) return r;
    CK(cudaGetLastError());
    double e;
    CK(cudaMemcpyAsync(&e, c->d_err_sum, sizeof(e), cudaMemcpyDeviceToHost, c->st));
    CK(cudaStreamSynchronize(c->st));
    if (err) *err = e;
    c->stats.kernels_launched = 1;
    return ICP_OK;
}

int icp_run(icp_ctx* c, int max_iter, double threshold, int check_sizes, int* iters_done, double* err_hist)
{
    if (!c || !c->have_model || !c->have_scene) { g_err = "icp_run: model and scene required"; return ICP_ERR_INVALID; }
    CK(cudaSetDevice(c->device));
    int r;
    // alignement_check, src/cpu.cc:42-53 (with a communicator the GLOBAL scene size is what counts)
    double np_glob = (double)c->s.np;
    if (c->in_group) np_glob = c->group_np;
    if (c->comm) {
        CK(cudaMemcpyAsync(c->d_err_sum, &np_glob, sizeof(double), cudaMemcpyHostToDevice, c->st));
        if ((r = allreduce(c, c->d_err_sum, 1))) return r;
        CK(cudaMemcpyAsync(&np_glob, c->d_err_sum, sizeof(double), cudaMemcpyDeviceToHost, c->st));
        CK(cudaStreamSynchronize(c->st));
    }
    if (check_sizes) {
        if ((int64_t)np_glob != c->m.nm) { g_err = "Point sets need to have the same number of points."; return ICP_ERR_SIZE_MISMATCH; }
        if (np_glob < 4) { g_err = "Need at least 4 point pairs"; return ICP_ERR_TOO_FEW; }
    }
    if (iters_done) *iters_done = 0;
    if (max_iter <= 0) return ICP_OK;
    if (c->err_hist_cap < max_iter) {
        if ((r = dalloc(&c->d_err_hist, (size_t)max_iter))) return r;
        c->err_hist_cap = max_iter;
    }
    const int EV = c->profile ? 5 : 2;     // events per iteration
    while ((int)c->ev_filter.size() < EV * max_iter) {
        cudaEvent_t e;
        CK(cudaEventCreate(&e));
        c->ev_filter.push_back(e);
    }
    if ((r = reset_ctl(c))) return r;
    c->auto_tree_now = c->nn_mode == ICP_NN_AUTO && max_iter >= 2 && small_for_brute(c);
    if ((r = ensure_tree(c))) return r;
    c->stats.kernels_launched = 0;
    CK(cudaEventRecord(c->ev0, c->st));
    // The host runs ahead of the device and cannot know when the loop converges, so every kernel returns at entry
    // once the device-side done flag is set; but even a no-op launch costs ~2.3 us, which for a small cloud that
    // converges early (cow: 7 of 30 iterations) was 45 % of the whole call.  So the filter kernel
    // stores, at entry, (iterations closed, converged) into one word of mapped pinned memory, and the host keeps at
    // most `lookahead` iterations enqueued beyond the last one it saw closed: the device never waits for the host,
    // nothing is added to the stream, and after convergence at most `lookahead` no-op iterations remain.  Not on the NCCL
    // path: there every rank must issue the same collectives.  With peer mailboxes it is safe: the exchange
    // number lives on the device (PeerMailbox::seq_local) and kernels after convergence do not exchange.
    const int LA = (c->lookahead > 0 && !nccl_path(c)) ? (c->lookahead < 2 ? 2 : c->lookahead) : 0;   // >= 2: see nn_filter_kernel
    *(volatile int*)c->h_progress = 0;
    c->progress_arg = LA ? c->d_progress : nullptr;
    for (int it = 0; it < max_iter; ++it) {
        if (LA && it >= LA) {
            const int w = wait_progress(c, it + 1 - LA, it);
            if (w < 0) return w;
            if (w) break;
        }
        cudaEvent_t* ev = &c->ev_filter[(size_t)EV * it];
        nn_pass(c, 1, (it > 0 || c->have_idx) ? 1 : 0, ev[0], ev[1], c->profile ? ev[2] : nullptr, c->profile ? ev[3] : nullptr);
        if (!nccl_path(c)) {
            int launched = align_step(c, 1, 1, 1, np_glob, threshold);
            if (launched < 0) return launched;
            if (c->profile) cudaEventRecord(ev[4], c->st);
            c->stats.kernels_launched += launched;
        } else {
            launch_moments(c->m, c->s, c->d_partials, c->n_red_blocks, c->d_mom, c->d_ctl, 1, nullptr, c->st);
            if ((r = allreduce(c, c->d_mom, MOM_N))) return r;
            launch_solve(c->m, c->d_mom, c->d_xf, c->d_ctl, 1, c->st);
            launch_transform(c->m, c->s, c->d_xf, 1, c->d_partials, c->n_red_blocks, c->d_err_sum, c->d_ctl, 1, 0, np_glob,
                             threshold, c->d_err_hist, c->st);
            if ((r = allreduce(c, c->d_err_sum, 1))) return r;
            launch_finish(c->d_err_sum, np_glob, threshold, c->d_err_hist, c->d_ctl, c->st);
            c->stats.kernels_launched += 4;
        }
    }
    c->progress_arg = nullptr;
    CK(cudaEventRecord(c->ev1, c->st));
    CK(cudaGetLastError());
    int ctl[CTL_N];
    CK(cudaMemcpyAsync(ctl, c->d_ctl, sizeof(ctl), cudaMemcpyDeviceToHost, c->st));
    CK(cudaStreamSynchronize(c->st));
    if (ctl[CTL_COMM_TIMEOUT]) {
        g_err = "peer exchange timed out (a rank did not arrive); new_p holds the last completed iteration, the communicator is unusable";
        return ICP_ERR_COMM;
    }
    if (ctl[CTL_BOUNDS]) { g_err = "bounds check failed at csrc line " + std::to_string(ctl[CTL_BOUNDS]); return ICP_ERR_CUDA; }
    const int iters = ctl[CTL_ITERS];
    if (iters_done) *iters_done = iters;
    if (err_hist && iters > 0)
        CK(cudaMemcpy(err_hist, c->d_err_hist, (size_t)iters * sizeof(double), cudaMemcpyDeviceToHost));
    c->have_idx = iters > 0;
    c->stats.nn_calls += iters;
    c->stats.nn_pairs += (int64_t)iters * c->s.np * c->m.nm;
    c->stats.rerank_full += ctl[CTL_AMB_TOTAL];
    c->stats.rerank_chunk += (int64_t)iters * c->s.np - ctl[CTL_AMB_TOTAL];
    cudaEventElapsedTime(&c->stats.ms_total, c->ev0, c->ev1);
    float f = 0.f, f_ref = 0.f, f_ex = 0.f, f_al = 0.f;
    for (int it = 0; it < iters; ++it) {
        cudaEvent_t* ev = &c->ev_filter[(size_t)EV * it];
        float ms = 0.f;
        cudaEventElapsedTime(&ms, ev[0], ev[1]);
        f += ms;
        if (c->profile && !nccl_path(c)) {
            cudaEventElapsedTime(&ms, ev[1], ev[2]); f_ref += ms;
            cudaEventElapsedTime(&ms, ev[2], ev[3]); f_ex += ms;
            cudaEventElapsedTime(&ms, ev[3], ev[4]); f_al += ms;
        }
    }
    c->stats.ms_filter = f;
    c->stats.ms_refine = f_ref;
    c->stats.ms_exact = f_ex;
    c->stats.ms_align = f_al;
    return ICP_OK;
}

int icp_get_trace(icp_ctx* c, double out_us[10])
{
    if (!c || !out_us) return ICP_ERR_INVALID;
    CK(cudaSetDevice(c->device));
    unsigned long long t[16];
    CK(cudaMemcpyAsync(t, c->d_trace, sizeof(t), cudaMemcpyDeviceToHost, c->st));
    CK(cudaStreamSynchronize(c->st));
    for (int i = 0; i < 8; ++i) out_us[i] = t[i] >= t[0] ? (double)(t[i] - t[0]) * 1e-3 : 0.0;
    out_us[0] = t[0] >= t[8] ? (double)(t[0] - t[8]) * 1e-3 : -1.0;     // gap: end of K1 -> entry of block 0
    if (t[9] >= t[0]) out_us[7] = (double)(t[9] - t[0]) * 1e-3;         // exit of the LAST block instead of block 0
    out_us[8] = t[10] >= t[0] ? (double)(t[10] - t[0]) * 1e-3 : 0.0;
    out_us[9] = t[11] >= t[0] ? (double)(t[11] - t[0]) * 1e-3 : 0.0;
    return ICP_OK;
}

int icp_get_block_trace(icp_ctx* c, double out[160])
{
    if (!c || !out) return ICP_ERR_INVALID;
    CK(cudaSetDevice(c->device));
    unsigned long long t[TRACE_WORDS];
    CK(cudaMemcpyAsync(t, c->d_trace, sizeof(t), cudaMemcpyDeviceToHost, c->st));
    CK(cudaStreamSynchronize(c->st));
    for (int b = 0; b < 32; ++b) {
        out[5 * b] = (double)t[16 + 5 * b];
        for (int i = 1; i < 5; ++i) out[5 * b + i] = t[16 + 5 * b + i] >= t[0] && t[16 + 5 * b + i] ? (double)(t[16 + 5 * b + i] - t[0]) * 1e-3 : -1.0;
    }
    return ICP_OK;
}

int icp_get_stats(icp_ctx* c, icp_stats* out)
{
    if (!c || !out) return ICP_ERR_INVALID;
    c->stats.n_splits = c->sc.n_splits;
    c->stats.tile_points = NN_TILE;
    c->stats.chunk_points = NN_G;
    *out = c->stats;
    return ICP_OK;
}

// ---- stateless one-shot forms (src/GPU/gpu.hh:110-116) -------------------------------------------
// "Stateless" as seen by the caller (host matrices in, host results out, nothing to keep), like the
// reference's wrappers which allocate and free device memory on every call (src/GPU/compute.cu:154-245).
// Creating a context costs ~3 ms (stream, events, a dozen cudaMalloc), thirty times the work of a cow-sized
// call, so each host thread keeps ONE hidden context for these entry points and reuses its buffers.
// icp_release_cached() frees the calling thread's hidden context.

static thread_local icp_ctx* t_oneshot = nullptr;

// Hidden contexts (the one-shot wrappers' and icp_batch_run's lanes) outlive a call: they follow the environment
// as it is when they are used, not as it was when they were created.
static void apply_env_nn(icp_ctx* c)
{
    c->nn_mode = env_nn_mode();
    const char* e;
    e = getenv("ICP_EXACT");
    c->exact_mode = (e && !strcmp(e, "inline")) ? 1 : ((e && !strcmp(e, "kernel")) ? 2 : 0);
}

static int oneshot_ctx(icp_ctx** out)
{
    if (!t_oneshot) {
        int r = icp_create(&t_oneshot, 0);
        if (r) { t_oneshot = nullptr; return r; }
    }
    apply_env_nn(t_oneshot);
    *out = t_oneshot;
    return ICP_OK;
}

static void lane_pool_clear();

void icp_release_cached(void)
{
    if (t_oneshot) icp_destroy(t_oneshot);
    t_oneshot = nullptr;
    lane_pool_clear();
}

int icp_closest_matrix_host(const double* m_xyz, int64_t nm, const double* p_xyz, int64_t np, int32_t* idx_out,
                            double* Y_out)
{
    icp_ctx* c = nullptr;
    int r = oneshot_ctx(&c);
    if (r) return r;
    if (!(r = icp_set_model(c, m_xyz, nm)) && !(r = icp_set_scene(c, p_xyz, np))) r = icp_closest(c, idx_out, Y_out);
    return r;
}

int icp_nearest_index_host(const double* m_xyz, int64_t nm, const double pi[3], int32_t* idx_out)
{
    return icp_closest_matrix_host(m_xyz, nm, pi, 1, idx_out, nullptr);
}

// Pairs p_j with Y_j directly: the "model" is Y and the correspondence is the identity.
static int pair_ctx(icp_ctx** out, const double* Y_xyz, const double* p_xyz, int64_t n)
{
    icp_ctx* c = nullptr;
    int r = oneshot_ctx(&c);
    if (r) return r;
    if ((r = icp_set_model(c, Y_xyz, n)) || (r = icp_set_scene(c, p_xyz, n))) return r;
    // sorted position i holds original point perm[i], whose partner is Y[perm[i]]
    if (cudaMemcpyAsync(c->s.idx, c->s.perm, (size_t)n * sizeof(int32_t), cudaMemcpyDeviceToDevice, c->st) != cudaSuccess) {
        g_err = "pair_ctx: copy failed";
        return ICP_ERR_CUDA;
    }
    launch_gather_matches(c->s, c->m, c->st);
    c->have_idx = true;
    *out = c;
    return ICP_OK;
}

int icp_err_compute_host(const double* Y_xyz, double* p_xyz, int64_t np, int in_place, const double sr[9],
                         const double t[3], double* err)
{
    if (!Y_xyz || !p_xyz || !sr || !t || np < 1) { g_err = "icp_err_compute_host: bad arguments"; return ICP_ERR_INVALID; }
    icp_ctx* c = nullptr;
    int r = pair_ctx(&c, Y_xyz, p_xyz, np);
    if (r) return r;
    double xf[XF_N] = {0};
    xf[0] = 1.0;
    memcpy(xf + 1, sr, 72);
    memcpy(xf + 10, t, 24);
    if (cudaMemcpyAsync(c->d_xf, xf, sizeof(xf), cudaMemcpyHostToDevice, c->st) != cudaSuccess ||
        cudaStreamSynchronize(c->st) != cudaSuccess) {
        g_err = "icp_err_compute_host: upload failed";
        return ICP_ERR_CUDA;
    }
    r = icp_err_compute(c, in_place, err);
    if (!r && in_place) r = icp_get_scene(c, p_xyz);
    return r;
}

int icp_find_alignment_host(const double* Y_xyz, const double* p_xyz, int64_t np, double* s, double R[9], double t[3],
                            double* err_align)
{
    if (!Y_xyz || !p_xyz || np < 1) { g_err = "icp_find_alignment_host: bad arguments"; return ICP_ERR_INVALID; }
    icp_ctx* c = nullptr;
    int r = pair_ctx(&c, Y_xyz, p_xyz, np);
    if (r) return r;
    return icp_find_alignment(c, s, R, t, err_align);
}

int icp_subtract_col_host(const double* M_xyz, int64_t n, const double m[3], double* out_xyz)
{
    if (!M_xyz || !m || !out_xyz || n < 1) { g_err = "icp_subtract_col_host: bad arguments"; return ICP_ERR_INVALID; }
    icp_ctx* c = nullptr;
    int r = oneshot_ctx(&c);
    if (r) return r;
    CK(cudaSetDevice(c->device));
    if ((r = ensure_stage(c, (size_t)n * 6))) return r;
    cudaError_t e = cudaMemcpyAsync(c->d_stage, M_xyz, (size_t)n * 24, cudaMemcpyHostToDevice, c->st);
    launch_subtract_col(c->d_stage, n, m[0], m[1], m[2], c->d_stage + (size_t)n * 3, c->st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(out_xyz, c->d_stage + (size_t)n * 3, (size_t)n * 24, cudaMemcpyDeviceToHost, c->st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->st);
    if (e != cudaSuccess) { g_err = std::string("icp_subtract_col_host: ") + cudaGetErrorString(e); r = ICP_ERR_CUDA; }
    return r;
}

int icp_y_p_norm_host(const double* y_xyz, const double* p_xyz, int64_t n, double* d_caps, double* sp)
{
    if (!y_xyz || !p_xyz || n < 1) { g_err = "icp_y_p_norm_host: bad arguments"; return ICP_ERR_INVALID; }
    icp_ctx* c = nullptr;
    int r = pair_ctx(&c, y_xyz, p_xyz, n);
    if (r) return r;
    launch_moments(c->m, c->s, c->d_partials, c->n_red_blocks, c->d_mom, c->d_ctl, 0, nullptr, c->st);
    double mom[MOM_N], geo[4];
    cudaError_t e = cudaMemcpyAsync(mom, c->d_mom, sizeof(mom), cudaMemcpyDeviceToHost, c->st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(geo, c->m.geo, sizeof(geo), cudaMemcpyDeviceToHost, c->st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->st);
    if (e != cudaSuccess) { g_err = std::string("icp_y_p_norm_host: ") + cudaGetErrorString(e); return ICP_ERR_CUDA; }
    // sums were taken about c0: |v|^2 = |a|^2 + 2 c0.a + |c0|^2
    const double c2 = geo[0] * geo[0] + geo[1] * geo[1] + geo[2] * geo[2];
    if (sp) *sp = mom[16] + 2.0 * (geo[0] * mom[1] + geo[1] * mom[2] + geo[2] * mom[3]) + (double)n * c2;
    if (d_caps) *d_caps = mom[17] + 2.0 * (geo[0] * mom[4] + geo[1] * mom[5] + geo[2] * mom[6]) + (double)n * c2;
    return ICP_OK;
}

// ---- batched registration (BASELINE config 5: many independent cloud pairs) ----------------------
// Pairs are independent, so there is nothing to exchange: `lanes` host threads each own a context (own
// stream) on `device` and pull pairs from a shared counter.  While one lane is in the latency-bound
// tail of an iteration (refine / exact / alignment) or copying over PCIe, the other lanes' filter
// kernels keep the SMs busy.  Across GPUs the launcher deals pairs round-robin, one process per GPU.
// Pool of idle lane contexts per device (icp_release_cached() empties it).
static std::atomic<long long> g_batch_kernels{0};   // own kernels launched by the last icp_batch_run of this process
static std::mutex g_pool_mu;
static std::vector<std::pair<int, icp_ctx*>> g_pool;

static icp_ctx* lane_pool_take(int device)
{
    std::lock_guard<std::mutex> lk(g_pool_mu);
    for (size_t i = 0; i < g_pool.size(); ++i)
        if (g_pool[i].first == device) {
            icp_ctx* c = g_pool[i].second;
            g_pool.erase(g_pool.begin() + (long)i);
            return c;
        }
    return nullptr;
}

static void lane_pool_give(int device, icp_ctx* c)
{
    std::lock_guard<std::mutex> lk(g_pool_mu);
    g_pool.emplace_back(device, c);
}

static int icp_batch_run_device(int device, int lanes, int64_t n_pairs, const double* const* m_xyz, const int64_t* nm,
                                const double* const* p_xyz, const int64_t* np, int max_iter, double threshold, int check_sizes,
                                int32_t* iters_out, double* err_out, double* const* new_p_out, double* srt_out,
                                long long* launched_out)
{
    if (n_pairs < 0 || (n_pairs > 0 && (!m_xyz || !nm || !p_xyz || !np))) { g_err = "icp_batch_run: bad arguments"; return ICP_ERR_INVALID; }
    if (lanes < 1) lanes = 1;
    if (lanes > 8) lanes = 8;
    if ((int64_t)lanes > n_pairs) lanes = n_pairs > 0 ? (int)n_pairs : 1;
    std::atomic<int64_t> next{0};
    std::atomic<int> status{ICP_OK};
    std::atomic<long long> launched{0};
    std::mutex mu;
    std::string first_err;
    auto worker = [&]() {
        icp_ctx* c = lane_pool_take(device);       // contexts survive between calls: no per-call cudaMalloc storm
        int r = c ? ICP_OK : icp_create(&c, device);
        if (c) { c->is_lane = true; apply_env_nn(c); }
        std::vector<double> hist((size_t)(max_iter > 0 ? max_iter : 1));
        while (!r && status.load() == ICP_OK) {
            const int64_t k = next.fetch_add(1);
            if (k >= n_pairs) break;
            int it = 0;
            if ((r = icp_set_model(c, m_xyz[k], nm[k])) || (r = icp_set_scene(c, p_xyz[k], np[k])) ||
                (r = icp_run(c, max_iter, threshold, check_sizes, &it, hist.data())))
                break;
            // the loop's kernels (counted by icp_run) + the 7 preparation kernels of icp_set_model / icp_set_scene
            // (2 x bbox, model_geo, model_build, scene_keys, scene_build; + scene_export below); cub's sort is not ours
            launched.fetch_add(c->stats.kernels_launched + 6 + ((new_p_out && new_p_out[k]) ? 1 : 0));
            if (iters_out) iters_out[k] = it;
            if (err_out) err_out[k] = it > 0 ? hist[(size_t)it - 1] : 0.0;
            if (new_p_out && new_p_out[k] && (r = icp_get_scene(c, new_p_out[k]))) break;
            if (srt_out && (r = icp_get_transform(c, srt_out + 13 * k, srt_out + 13 * k + 1, srt_out + 13 * k + 10))) break;
        }
        if (r) {
            std::lock_guard<std::mutex> lk(mu);
            if (status.load() == ICP_OK) { status.store(r); first_err = g_err; }
        }
        if (c) lane_pool_give(device, c);
    };
    std::vector<std::thread> th;
    for (int i = 0; i < lanes; ++i) th.emplace_back(worker);
    for (auto& t : th) t.join();
    if (status.load() != ICP_OK) g_err = first_err;
    if (launched_out) *launched_out = launched.load();
    return status.load();
}

int icp_batch_run(int device, int lanes, int64_t n_pairs, const double* const* m_xyz, const int64_t* nm,
                  const double* const* p_xyz, const int64_t* np, int max_iter, double threshold, int check_sizes,
                  int32_t* iters_out, double* err_out, double* const* new_p_out, double* srt_out)
{
    long long launched = 0;
    const int r = icp_batch_run_device(device, lanes, n_pairs, m_xyz, nm, p_xyz, np, max_iter, threshold, check_sizes, iters_out,
                                       err_out, new_p_out, srt_out, &launched);
    g_batch_kernels.store(launched);
    return r;
}


long long icp_batch_kernels_launched(void) { return g_batch_kernels.load(); }

static void lane_pool_clear()
{
    std::lock_guard<std::mutex> lk(g_pool_mu);
    for (auto& e : g_pool) icp_destroy(e.second);
    g_pool.clear();
}

// ---- multi-GPU ----------------------------------------------------------------------------------

int icp_comm_unique_id(const char* nccl_library_path, void* id128)
{
    if (!id128) return ICP_ERR_INVALID;
    int r = load_nccl(g_nccl, nccl_library_path);
    if (r) return r;
    nccl_uid id;
    if (g_nccl.GetUniqueId(&id) != 0) { g_err = "ncclGetUniqueId failed"; return ICP_ERR_COMM; }
    memcpy(id128, &id, 128);
    return ICP_OK;
}

int icp_comm_init(icp_ctx* c, const char* nccl_library_path, const void* id128, int rank, int world)
{
    if (!c || !id128 || world < 1 || rank < 0 || rank >= world) { g_err = "icp_comm_init: bad arguments"; return ICP_ERR_INVALID; }
    int r = load_nccl(g_nccl, nccl_library_path);
    if (r) return r;
    CK(cudaSetDevice(c->device));
    nccl_uid id;
    memcpy(&id, id128, 128);
    int e = g_nccl.CommInitRank(&c->comm, world, id, rank);
    if (e != 0) {
        g_err = std::string("ncclCommInitRank: ") + (g_nccl.GetErrorString ? g_nccl.GetErrorString(e) : "error");
        c->comm = nullptr;
        return ICP_ERR_COMM;
    }
    c->rank = rank;
    c->world = world;
    return ICP_OK;
}

int icp_comm_peer_handle(icp_ctx* c, void* handle64)
{
    if (!c || !handle64) { g_err = "icp_comm_peer_handle: bad arguments"; return ICP_ERR_INVALID; }
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "CUDA IPC handle size");
    CK(cudaSetDevice(c->device));
    if (!c->d_mailbox) {
        CK(cudaMalloc((void**)&c->d_mailbox, sizeof(PeerMailbox)));
        CK(cudaMemset(c->d_mailbox, 0, sizeof(PeerMailbox)));
        CK(cudaDeviceSynchronize());
    }
    cudaIpcMemHandle_t h;
    CK(cudaIpcGetMemHandle(&h, c->d_mailbox));
    memcpy(handle64, &h, 64);
    return ICP_OK;
}

int icp_comm_peer_attach(icp_ctx* c, const void* handles, int rank, int world)
{
    if (!c || !handles || world < 1 || world > PEER_MAX_WORLD || rank < 0 || rank >= world || !c->d_mailbox) {
        g_err = "icp_comm_peer_attach: bad arguments (export the local handle first; world <= 16)";
        return ICP_ERR_INVALID;
    }
    CK(cudaSetDevice(c->device));
    for (int r = 0; r < PEER_MAX_WORLD; ++r) c->peer.box[r] = nullptr;
    for (int r = 0; r < world; ++r) {
        if (r == rank) { c->peer.box[r] = c->d_mailbox; continue; }
        cudaIpcMemHandle_t h;
        memcpy(&h, (const char*)handles + (size_t)r * 64, 64);
        void* p = nullptr;
        CK(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
        c->peer.box[r] = (PeerMailbox*)p;
    }
    c->peer.rank = rank;
    c->peer.world = world;
    c->peer_ready = true;
    return ICP_OK;
}

int icp_comm_world(icp_ctx* c, int* rank, int* world)
{
    if (!c) return ICP_ERR_INVALID;
    if (rank) *rank = c->rank;
    if (world) *world = c->world;
    return ICP_OK;
}


// ---- one process, several devices (icp_create(ctx, devices, ndev)) --------------------------------
// An icp_group owns one context per device.  The model is replicated, the scene is cut into contiguous blocks (member
// i holds block i), and the per-iteration exchange is the same in-kernel mailbox protocol the multi-process path uses —
// the mailboxes are simply reached through peer access (cudaDeviceEnablePeerAccess) instead of CUDA IPC; no NCCL, no
// launcher.  Every collective call runs its members concurrently on one host thread each (icp_run keeps its own
// look-ahead loop per device).  Loop being sharded: src/cpu.cc:55-79.

}  // extern "C" (the group's internals are C++)

struct icp_group {
    std::vector<icp_ctx*> ctx;
    std::vector<int64_t> lo;          // member i holds scene points [lo[i], lo[i+1])
    int64_t np = 0;
};

namespace {

// Runs f(i, member i) for every member concurrently; returns the first failing status (its message in g_err).
template <class F>
int group_parallel(icp_group* g, F f)
{
    const int n = (int)g->ctx.size();
    if (n == 1) return f(0, g->ctx[0]);
    std::vector<int> rc((size_t)n, ICP_OK);
    std::vector<std::string> msg((size_t)n);
    std::vector<std::thread> th;
    for (int i = 0; i < n; ++i)
        th.emplace_back([&, i]() {
            rc[(size_t)i] = f(i, g->ctx[(size_t)i]);
            if (rc[(size_t)i]) msg[(size_t)i] = g_err;       // g_err is thread-local
        });
    for (auto& t : th) t.join();
    for (int i = 0; i < n; ++i)
        if (rc[(size_t)i]) { g_err = "device " + std::to_string(g->ctx[(size_t)i]->device) + ": " + msg[(size_t)i]; return rc[(size_t)i]; }
    return ICP_OK;
}

}  // namespace

extern "C" {

int icp_group_create(icp_group** out, const int* devices, int ndev)
{
    if (!out || !devices || ndev < 1 || ndev > PEER_MAX_WORLD) { g_err = "icp_group_create: bad arguments (1..16 devices)"; return ICP_ERR_INVALID; }
    *out = nullptr;
    for (int i = 0; i < ndev; ++i)
        for (int j = 0; j < i; ++j)
            if (devices[i] == devices[j]) { g_err = "icp_group_create: a device is listed twice"; return ICP_ERR_INVALID; }
    icp_group* g = new icp_group();
    int r = ICP_OK;
    for (int i = 0; i < ndev && !r; ++i) {
        icp_ctx* c = nullptr;
        r = icp_create(&c, devices[i]);
        if (!r) g->ctx.push_back(c);
    }
    if (!r && ndev > 1) {
        for (int i = 0; i < ndev && !r; ++i) {
            icp_ctx* c = g->ctx[(size_t)i];
            if (cudaSetDevice(c->device) != cudaSuccess) { g_err = "icp_group_create: cudaSetDevice failed"; r = ICP_ERR_CUDA; break; }
            for (int j = 0; j < ndev && !r; ++j) {
                if (j == i) continue;
                int can = 0;
                cudaDeviceCanAccessPeer(&can, c->device, devices[j]);
                if (!can) { g_err = "icp_group_create: device " + std::to_string(c->device) + " cannot access device " + std::to_string(devices[j]) + " (no peer path)"; r = ICP_ERR_COMM; break; }
                const cudaError_t e = cudaDeviceEnablePeerAccess(devices[j], 0);
                if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) { g_err = std::string("cudaDeviceEnablePeerAccess: ") + cudaGetErrorString(e); r = ICP_ERR_CUDA; break; }
                cudaGetLastError();
            }
            if (r) break;
            if (cudaMalloc((void**)&c->d_mailbox, sizeof(PeerMailbox)) != cudaSuccess ||
                cudaMemset(c->d_mailbox, 0, sizeof(PeerMailbox)) != cudaSuccess || cudaDeviceSynchronize() != cudaSuccess) {
                g_err = "icp_group_create: mailbox allocation failed";
                r = ICP_ERR_CUDA;
            }
        }
        for (int i = 0; i < ndev && !r; ++i) {
            icp_ctx* c = g->ctx[(size_t)i];
            for (int k = 0; k < PEER_MAX_WORLD; ++k) c->peer.box[k] = k < ndev ? g->ctx[(size_t)k]->d_mailbox : nullptr;
            c->peer.rank = i;
            c->peer.world = ndev;
            c->rank = i;
            c->world = ndev;
            c->peer_ready = true;
            c->in_group = true;
        }
    }
    if (r) {
        const std::string why = g_err;
        icp_group_destroy(g);
        g_err = why;
        return r;
    }
    g->lo.assign((size_t)ndev + 1, 0);
    *out = g;
    return ICP_OK;
}

void icp_group_destroy(icp_group* g)
{
    if (!g) return;
    for (icp_ctx* c : g->ctx) icp_destroy(c);
    delete g;
}

int icp_group_size(icp_group* g) { return g ? (int)g->ctx.size() : 0; }
icp_ctx* icp_group_member(icp_group* g, int i) { return (g && i >= 0 && i < (int)g->ctx.size()) ? g->ctx[(size_t)i] : nullptr; }

int icp_group_set_model(icp_group* g, const double* m_xyz, int64_t nm)
{
    if (!g) return ICP_ERR_INVALID;
    return group_parallel(g, [&](int, icp_ctx* c) { return icp_set_model(c, m_xyz, nm); });
}

int icp_group_set_scene(icp_group* g, const double* p_xyz, int64_t np)
{
    const int n = g ? (int)g->ctx.size() : 0;
    if (!g || !p_xyz || np < n) { g_err = "icp_group_set_scene: bad arguments (at least one scene point per device)"; return ICP_ERR_INVALID; }
    const int64_t base = np / n, rem = np % n;
    for (int i = 0; i <= n; ++i) g->lo[(size_t)i] = i * base + (i < rem ? i : rem);
    g->np = np;
    return group_parallel(g, [&](int i, icp_ctx* c) {
        c->group_np = (double)np;
        return icp_set_scene(c, p_xyz + 3 * g->lo[(size_t)i], g->lo[(size_t)i + 1] - g->lo[(size_t)i]);
    });
}

int icp_group_reset_scene(icp_group* g)
{
    if (!g) return ICP_ERR_INVALID;
    return group_parallel(g, [&](int, icp_ctx* c) { return icp_reset_scene(c); });
}

int icp_group_set_nn_mode(icp_group* g, int mode)
{
    if (!g) return ICP_ERR_INVALID;
    for (icp_ctx* c : g->ctx) { const int r = icp_set_nn_mode(c, mode); if (r) return r; }
    return ICP_OK;
}

int icp_group_closest(icp_group* g, int32_t* idx_out, double* Y_out)
{
    if (!g) return ICP_ERR_INVALID;
    return group_parallel(g, [&](int i, icp_ctx* c) {
        return icp_closest(c, idx_out ? idx_out + g->lo[(size_t)i] : nullptr, Y_out ? Y_out + 3 * g->lo[(size_t)i] : nullptr);
    });
}

int icp_group_run(icp_group* g, int max_iter, double threshold, int check_sizes, int* iters_done, double* err_hist)
{
    if (!g) return ICP_ERR_INVALID;
    const int n = (int)g->ctx.size();
    std::vector<int> it((size_t)n, 0);
    const int r = group_parallel(g, [&](int i, icp_ctx* c) {
        return icp_run(c, max_iter, threshold, check_sizes, &it[(size_t)i], i == 0 ? err_hist : nullptr);
    });
    if (r) return r;
    for (int i = 1; i < n; ++i)
        if (it[(size_t)i] != it[0]) { g_err = "icp_group_run: the devices disagree on the iteration count"; return ICP_ERR_COMM; }
    if (iters_done) *iters_done = it[0];
    return ICP_OK;
}

int icp_group_find_alignment(icp_group* g, double* s, double R[9], double t[3], double* err_align)
{
    if (!g) return ICP_ERR_INVALID;
    return group_parallel(g, [&](int i, icp_ctx* c) {
        return i == 0 ? icp_find_alignment(c, s, R, t, err_align) : icp_find_alignment(c, nullptr, nullptr, nullptr, nullptr);
    });
}

int icp_group_err_compute(icp_group* g, int in_place, double* err)
{
    if (!g) return ICP_ERR_INVALID;
    return group_parallel(g, [&](int i, icp_ctx* c) { return icp_err_compute(c, in_place, i == 0 ? err : nullptr); });
}

int icp_group_centroid(icp_group* g, double mu_p[3], double mu_y[3])
{
    if (!g) return ICP_ERR_INVALID;
    return group_parallel(g, [&](int i, icp_ctx* c) { return icp_centroid(c, i == 0 ? mu_p : nullptr, i == 0 ? mu_y : nullptr); });
}

int icp_group_get_scene(icp_group* g, double* p_xyz_out)
{
    if (!g || !p_xyz_out) return ICP_ERR_INVALID;
    return group_parallel(g, [&](int i, icp_ctx* c) { return icp_get_scene(c, p_xyz_out + 3 * g->lo[(size_t)i]); });
}

int icp_group_get_indices(icp_group* g, int32_t* idx_out)
{
    if (!g || !idx_out) return ICP_ERR_INVALID;
    return group_parallel(g, [&](int i, icp_ctx* c) { return icp_get_indices(c, idx_out + g->lo[(size_t)i]); });
}

int icp_group_get_transform(icp_group* g, double* s, double R[9], double t[3])
{
    if (!g) return ICP_ERR_INVALID;
    return icp_get_transform(g->ctx[0], s, R, t);
}

// Whole-group view: sizes and pair counts summed, times = the slowest device, launches summed.
int icp_group_get_stats(icp_group* g, icp_stats* out)
{
    if (!g || !out) return ICP_ERR_INVALID;
    icp_stats a{};
    for (size_t i = 0; i < g->ctx.size(); ++i) {
        icp_stats s1;
        icp_get_stats(g->ctx[i], &s1);
        if (i == 0) a = s1;
        else {
            a.np += s1.np; a.nn_pairs += s1.nn_pairs; a.rerank_full += s1.rerank_full; a.rerank_chunk += s1.rerank_chunk;
            a.kernels_launched += s1.kernels_launched;
            a.ms_filter = s1.ms_filter > a.ms_filter ? s1.ms_filter : a.ms_filter;
            a.ms_total = s1.ms_total > a.ms_total ? s1.ms_total : a.ms_total;
        }
    }
    *out = a;
    return ICP_OK;
}

// icp_batch_run over several devices: `lanes` contexts per device, all pulling pairs from one shared counter.
int icp_batch_run_multi(const int* devices, int ndev, int lanes, int64_t n_pairs, const double* const* m_xyz, const int64_t* nm,
                        const double* const* p_xyz, const int64_t* np, int max_iter, double threshold, int check_sizes,
                        int32_t* iters_out, double* err_out, double* const* new_p_out, double* srt_out)
{
    if (!devices || ndev < 1) { g_err = "icp_batch_run_multi: bad arguments"; return ICP_ERR_INVALID; }
    if (ndev == 1)
        return icp_batch_run(devices[0], lanes, n_pairs, m_xyz, nm, p_xyz, np, max_iter, threshold, check_sizes, iters_out,
                             err_out, new_p_out, srt_out);
    // pair k goes to device k % ndev (any deal is correct: pairs are independent); the devices run concurrently
    std::vector<int> rc((size_t)ndev, ICP_OK);
    std::vector<std::string> msg((size_t)ndev);
    std::vector<long long> launched((size_t)ndev, 0);
    std::vector<std::thread> th;
    for (int d = 0; d < ndev; ++d)
        th.emplace_back([&, d]() {
            std::vector<const double*> mm, pp;
            std::vector<double*> oo;
            std::vector<int64_t> nmm, npp, ids;
            for (int64_t k = d; k < n_pairs; k += ndev) {
                ids.push_back(k); mm.push_back(m_xyz[k]); pp.push_back(p_xyz[k]); nmm.push_back(nm[k]); npp.push_back(np[k]);
                oo.push_back(new_p_out ? new_p_out[k] : nullptr);
            }
            const int64_t n = (int64_t)ids.size();
            std::vector<int32_t> it((size_t)(n ? n : 1));
            std::vector<double> er((size_t)(n ? n : 1)), srt((size_t)(n ? n : 1) * 13);
            rc[(size_t)d] = icp_batch_run_device(devices[d], lanes, n, mm.data(), nmm.data(), pp.data(), npp.data(), max_iter, threshold,
                                                 check_sizes, it.data(), er.data(), new_p_out ? oo.data() : nullptr, srt.data(),
                                                 &launched[(size_t)d]);
            if (rc[(size_t)d]) msg[(size_t)d] = g_err;
            for (int64_t j = 0; j < n && !rc[(size_t)d]; ++j) {
                if (iters_out) iters_out[ids[(size_t)j]] = it[(size_t)j];
                if (err_out) err_out[ids[(size_t)j]] = er[(size_t)j];
                if (srt_out) memcpy(srt_out + 13 * ids[(size_t)j], srt.data() + 13 * j, 13 * sizeof(double));
            }
        });
    for (auto& t : th) t.join();
    long long total = 0;
    for (int d = 0; d < ndev; ++d) {
        if (rc[(size_t)d]) { g_err = "device " + std::to_string(devices[d]) + ": " + msg[(size_t)d]; return rc[(size_t)d]; }
        total += launched[(size_t)d];
    }
    g_batch_kernels.store(total);
    return ICP_OK;
}

}  // extern "C"
