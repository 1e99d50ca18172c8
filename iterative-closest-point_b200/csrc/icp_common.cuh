// icp_common.cuh — shared declarations of the B200-native ICP hot path (sm_100a only).
//
// Device data layout (everything resident in HBM for the life of a context):
//   model   m64       : FP64 master, one 32-byte record (x,y,z,0) per point, original order (indices returned are positions here)
//           mhat      : float4 (x,y,z,0) = fl32(m - c0), c0 = centre of the model's bounding box;
//                       padded to a multiple of NN_TM with far-away sentinels
//   scene   p64x/y/z  : FP64 SoA master of new_p, Morton-sorted (perm[i] = original position)
//           phx/y/z   : FP32 shadow fl32(new_p - c0), padded to a multiple of NN_TILE
//           idx, y64* : int32 correspondence per sorted scene point and the matched model point itself (FP64 SoA)
//   scratch part_*    : per (model split, scene point) filter results; block partials of reductions
//   state   xf[]      : s, R[9], t[3], mu_p[3], mu_y[3]; ctl[] : done flag, iteration counter, ...
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace icpb {

// ---- nearest-neighbour filter geometry ---------------------------------------------------------
// The ICP_NN_* macros exist for A/B builds only (`make variants`, tools/filter_ab.sh: one library per variant,
// selected at run time with ICP_B200_LIB); the product is the defaults.  Round 2 measured the variants the static SASS
// study of round 1 favoured and then their neighbours (profiles/r02a_ab_summary.txt, r02b_, r02c_): 16 scene points per
// thread at 2 CTAs per SM (255 registers, no spills), 128-point chunks and the inner loop unrolled 8x replaced round 1's
// 8 / 4 / 64 / 2 (filter kernel on horse 0.795 -> 0.852 of the FP32 roofline, on 2^18 points 0.858 -> 0.914).
#ifndef ICP_NN_R
#define ICP_NN_R 16
#endif
#ifndef ICP_NN_G
#define ICP_NN_G 128
#endif
#ifndef ICP_NN_Q_UNROLL
#define ICP_NN_Q_UNROLL 8
#endif
#ifndef ICP_NN_MIN_CTAS
#define ICP_NN_MIN_CTAS 2
#endif
constexpr int NN_THREADS = 128;                 // threads per CTA
constexpr int NN_R       = ICP_NN_R;            // scene points register-blocked per thread (16)
constexpr int NN_TILE    = NN_THREADS * NN_R;   // scene points per CTA (2048)
constexpr int NN_G       = ICP_NN_G;            // model points per tracking chunk (128)
constexpr int NN_TM      = 512;                 // model points per shared-memory stage
constexpr int NN_Q_UNROLL = ICP_NN_Q_UNROLL;    // LDS.128 quadruples per trip of the filter's inner loop (8)
constexpr int NN_MIN_CTAS = ICP_NN_MIN_CTAS;    // resident filter CTAs per SM the launch bounds ask for (2)
static_assert(NN_TM % NN_G == 0 && NN_G % 32 == 0 && NN_G <= 256, "chunk size");
static_assert((NN_G / 4) % NN_Q_UNROLL == 0, "inner-loop unroll must divide the chunk");
constexpr float NN_SENTINEL = 1.0e18f;          // coordinate of padded model points (never wins)

// ---- reductions --------------------------------------------------------------------------------
constexpr int RED_THREADS = 256;
constexpr int MOM_N = 18;   // n, sum a(3), sum b(3), sum a b^T(9), sum |a|^2, sum |b|^2   (a = p-c0, b = y-c0)
constexpr int XF_N  = 20;   // s, R[9], t[3], mu_p[3], mu_y[3], spare

// ctl[] slots (int32, device)
enum { CTL_DONE = 0, CTL_ITERS = 1, CTL_AMB_COUNT = 2, CTL_TICKET_MOM = 3, CTL_TICKET_ERR = 4,
       CTL_AMB_TOTAL = 5, CTL_CHUNK_TOTAL = 6 /* unused: derived on the host */, CTL_TICKET_NN = 7, CTL_COMM_TIMEOUT = 8,
       CTL_TICKET_BBOX = 9, CTL_BOUNDS = 10 /* checked builds (ICP_BOUNDS_CHECK): source line of the first bad gather index */,
       CTL_N = 12 };

// ---- peer-memory exchange (multi-GPU, one process per GPU) ---------------------------------------
// Every rank owns one mailbox in its HBM, mapped into all peers through CUDA IPC.  The exchange number q is kept
// on the device (seq_local, advanced by the kernel that exchanges), so it is the same on every rank by
// construction.  Exchange q posts this rank's doubles into slot [q & 1][rank] of EVERY rank's mailbox (plain
// stores over NVLink), then publishes q in the matching flag with a system-scope release; the reader
// spins on its own flags (acquire) and adds the slots in rank order, so every rank obtains the
// bit-identical sum without a separate collective launch.
constexpr int PEER_MAX_WORLD = 16;
constexpr int PEER_SLOT = 32;                       // doubles per (parity, source rank)
struct PeerMailbox {
    double data[2][PEER_MAX_WORLD][PEER_SLOT];
    unsigned long long flag[2][PEER_MAX_WORLD];
    unsigned long long seq_local;                   // exchanges completed by the OWNER of this mailbox (peers never touch it)
};
struct PeerComm {
    PeerMailbox* box[PEER_MAX_WORLD];               // box[r] = rank r's mailbox as mapped in this process
    int rank = 0, world = 0;                        // world == 0: no peers attached (single GPU)
};

// FP64 model master: one 32-byte record per point (x, y, z, 0).  Every FP64 use of the model is a GATHER through a
// correspondence index (K1's re-rank, K2, K4, the exports), so one point = one 32-byte DRAM sector = one LDG.256;
// the SoA layout of round 1 cost three sectors for 24 useful bytes (ncu, 2^22 points: K2 read 113 B per point for 52
// algorithmic, profiles/r02a_syn4m_k2k4_split_metrics.csv).
struct __align__(32) Pt64 { double x, y, z, w; };

struct ModelDev {
    int64_t nm = 0, nm_pad = 0;
    Pt64* m64 = nullptr;
    float4* mhat = nullptr;
    double* geo = nullptr;       // device: c0[3], Rm (upper bound of |m - c0|), bbox min[3], max[3]
};

struct SceneDev {
    int64_t np = 0, np_pad = 0;
    double *p64x = nullptr, *p64y = nullptr, *p64z = nullptr;   // new_p, sorted order
    double *q64x = nullptr, *q64y = nullptr, *q64z = nullptr;   // backup of the initial scene, sorted order
    float *phx = nullptr, *phy = nullptr, *phz = nullptr;
    int32_t* perm = nullptr;     // sorted position -> original position
    int32_t* idx = nullptr;      // correspondence (sorted order)
    // Y = the matched model point of every scene point (FP64 SoA, sorted order), written by whichever kernel decides the
    // match (it has just read that model record for its FP64 evaluation).  K2 and K4 then STREAM p and Y instead of
    // gathering m64[idx[j]]: a random 32-byte gather costs a 64-byte DRAM burst and a dependent idx -> m64 chain per point
    // (ncu, 2^22 points: K2 read 78 B per point for 60 expected, profiles/r02b_syn4m_k2k4_split_metrics.csv).
    double *y64x = nullptr, *y64y = nullptr, *y64z = nullptr;
};

struct MatchOut {                // where a search kernel records its decision for scene point j
    int32_t* idx;
    double *yx, *yy, *yz;
};
inline MatchOut match_out(const SceneDev& s) { return MatchOut{s.idx, s.y64x, s.y64y, s.y64z}; }

struct NNScratch {
    int n_cta = 0;               // CTAs of the filter's single wave (stream-K decomposition)
    int n_splits = 0;            // max partial results ("slots") any scene tile receives
    float* best = nullptr;       // [n_splits][np_pad]
    float* runner = nullptr;
    uint32_t* chunk = nullptr;
    int32_t* amb_list = nullptr; // [np]   scene points the chunk test could not decide ...
    float* amb_thr = nullptr;    // [np]   ... with their FP32 acceptance threshold best + tau
    size_t cap_points = 0;       // allocated n_splits * np_pad
};

// Bounding-box hierarchy over the Morton-sorted model (tree_kernels.cu; opt-in exact search).
constexpr int TREE_LEAF = 8;     // model points per leaf
struct WideLevels {              // 32-ary form: level 0 = boxes of the 32-point leaves, level l = boxes of level l-1's blocks
    int T = 0;                   // number of box levels; the top one is a single block of <= 32 boxes
    int count[8] = {};           // boxes at each level
    long long off[8] = {};       // float offset of each level's blocks (192 floats per block of 32 boxes) in wboxes
};
struct TreeDev {
    int n_leaves = 0;            // power of two >= ceil(nm / TREE_LEAF); 0 = not built
    float* nodes = nullptr;      // [2 * n_leaves][6]  lo.xyz, hi.xyz; heap numbering, root = 1, leaves n_leaves..2n-1
    float4* pts = nullptr;       // [n_leaves * TREE_LEAF]  (m^.x, m^.y, m^.z, original index), Morton order
    size_t cap_leaves = 0;
    WideLevels w;
    float* wboxes = nullptr;     // all levels of the wide form
    float4* wpts = nullptr;      // [count[0] * 32] sorted points, padded with the last one
    size_t cap_wboxes = 0, cap_wpts = 0;
    void* kd_scratch = nullptr;  // global k-d levels of large models (tree_build)
    size_t cap_kd = 0;
};

// ---- launchers (defined in nn_kernels.cu / align_kernels.cu) ------------------------------------
// All launch on `st`, none synchronises.  `ctl` is the device control block; kernels return at entry
// when ctl[CTL_DONE] != 0 and `honor_done` is set (used by the device-side loop).
void launch_nn_filter(const ModelDev& m, const SceneDev& s, const NNScratch& sc, const int* ctl, int honor_done,
                      cudaStream_t st, int* progress = nullptr);
// inline_exact: the refine kernel rescans the model itself for the points it cannot decide (no nn_exact launch needed)
void launch_nn_refine(const ModelDev& m, const SceneDev& s, const NNScratch& sc, int* ctl, int honor_done,
                      int inline_exact, cudaStream_t st);
void launch_nn_exact(const ModelDev& m, const SceneDev& s, const NNScratch& sc, int* ctl, int honor_done,
                     int sm_count, cudaStream_t st, unsigned long long* trace = nullptr);
// tree search (tree_kernels.cu).  tree_build builds ONE form (wide or binary) and returns a cudaError_t; use_prev: idx[] holds the previous matches.
size_t tree_sort_bytes(int n);
int tree_build(const ModelDev& m, TreeDev& t, bool wide, uint32_t* keys, uint32_t* keys_alt,
               int32_t* vals, int32_t* perm, void* cub_tmp, size_t cub_bytes, void* kd_scratch /* nullable */, cudaStream_t st);
size_t kd_scratch_bytes(int nm, size_t* cub_bytes_out);
int kd_global_min_points();
size_t wtree_plan(int64_t nm, WideLevels& w);
void launch_nn_wtree(const ModelDev& m, const SceneDev& s, const TreeDev& t, int use_prev, int* ctl, int honor_done,
                     cudaStream_t st, int* progress = nullptr);
void launch_nn_tree(const ModelDev& m, const SceneDev& s, const TreeDev& t, int use_prev, int* ctl, int honor_done,
                    cudaStream_t st, int* progress = nullptr);

// xf_fused != nullptr: the last block also runs the Horn solve (K3) into xf_fused (single-GPU path).
// peer != nullptr (multi-GPU): the LAST block exchanges the sums with all peers inside the kernel before it solves / closes the iteration; every rank ends up with bit-identical values.
void launch_moments(const ModelDev& m, const SceneDev& s, double* partials, int n_blocks, double* mom, int* ctl,
                    int honor_done, double* xf_fused, cudaStream_t st, const PeerComm* peer = nullptr);
void launch_solve(const ModelDev& m, const double* mom, double* xf, const int* ctl, int honor_done, cudaStream_t st);
// in_place: p <- sR p + t (FP64 master and FP32 shadow).  finish: the last block also closes the iteration
// (err_hist[it] = 2*sum/n, done flag, iteration counter) — single-GPU loop only.
void launch_transform(const ModelDev& m, const SceneDev& s, const double* xf, int in_place, double* partials,
                      int n_blocks, double* err_sum, int* ctl, int honor_done, int finish, double n_global,
                      double threshold, double* err_hist, cudaStream_t st, const PeerComm* peer = nullptr);
// K2 + K3 + K4 in ONE cooperative launch (moments -> grid barrier -> Horn solve in every block -> transform
// + residual -> last block closes the iteration).  `peer` (world > 1): the 18 moment sums and the residual
// are exchanged through peer mailboxes inside the kernel; seq = sequence number of the first exchange.
// Returns a cudaError_t (cooperative launch can fail).
int launch_align_fused(const ModelDev& m, const SceneDev& s, double* partials, int n_blocks, double* mom, double* xf,
                       double* err_sum, int* ctl, int honor_done, int in_place, int finish, double n_global,
                       double threshold, double* err_hist, const PeerComm* peer, unsigned long long* trace,
                       cudaStream_t st);
int align_fused_max_blocks(int sm_count);
void launch_finish(const double* err_sum, double n_global, double threshold, double* err_hist, int* ctl,
                   cudaStream_t st);

// preparation
void launch_bbox(const double* xyz_aos, int64_t n, double* out_min3_max3, double* partials /* >= 6 * max_blocks */,
                 int max_blocks, int* ctl, cudaStream_t st);
void launch_model_geo(double* geo, cudaStream_t st);   // geo[4..9] bbox -> geo[0..3] = c0, Rm
void launch_model_build(const double* xyz_aos, const ModelDev& m, cudaStream_t st);
void launch_scene_keys(const double* xyz_aos, int64_t n, const double* bbox, uint32_t* keys, int32_t* vals,
                       cudaStream_t st);
void launch_scene_build(const double* xyz_aos, const ModelDev& m, const SceneDev& s, cudaStream_t st);
void launch_scene_restore(const ModelDev& m, const SceneDev& s, cudaStream_t st);
void launch_scene_export(const SceneDev& s, double* xyz_aos_out, cudaStream_t st);   // unsort + interleave
void launch_indices_export(const SceneDev& s, const ModelDev& m, int32_t* idx_out, double* Y_aos_out,
                           cudaStream_t st);
void launch_centroid_export(const ModelDev& m, const double* mom, double* out6, cudaStream_t st);
// idx[] was set from outside the search kernels (pairing by permutation, all-zero matches): rebuild Y = m64[idx]
void launch_gather_matches(const SceneDev& s, const ModelDev& m, cudaStream_t st);
void launch_subtract_col(const double* M_aos, int64_t n, double mx, double my, double mz, double* out_aos,
                         cudaStream_t st);

int reduce_blocks_for(int64_t n, int sm_count);

// ---- small device helpers -----------------------------------------------------------------------
#ifdef __CUDACC__
// one model point of the FP64 master: a single 256-bit read-only load (SASS LDG.E.ENL2.256.CONSTANT)
__device__ __forceinline__ void ld_pt64(const Pt64* __restrict__ a, int k, double& x, double& y, double& z)
{
    double w __attribute__((unused));
    asm("ld.global.nc.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(x), "=d"(y), "=d"(z), "=d"(w) : "l"(a + k));
}
// records the match of scene point j: the index and the model point itself (one 32-byte read, normally an L1 / L2 hit: the
// deciding kernel evaluated that record in FP64 a moment ago)
__device__ __forceinline__ void store_match(const MatchOut& mo, const Pt64* __restrict__ m64, int j, int k)
{
    double x, y, z;
    ld_pt64(m64, k, x, y, z);
    mo.idx[j] = k;
    mo.yx[j] = x;
    mo.yy[j] = y;
    mo.yz[j] = z;
}
// A scene point with a NaN or infinite coordinate has only NaN / +inf distances; Eigen's minCoeff visitor starts at
// column 0 and only moves on a strict `<` (src/cpu.cc:21-22), so the reference matches it to index 0.  Every search
// kernel answers that directly instead of scanning the model for it (after one such iteration the reference's
// transform is NaN and EVERY scene point is of this kind).
__device__ __forceinline__ bool scene_point_finite(double x, double y, double z)
{
    return isfinite(x) && isfinite(y) && isfinite(z);
}
// Checked builds (make checked: -DICP_BOUNDS_CHECK, the stand-in for compute-sanitizer, which is closed on this pool):
// every gather index is range-checked; the first offender's source line lands in ctl[CTL_BOUNDS] and the index is
// clamped, so a test run under the checked library fails loudly instead of faulting.
#ifdef ICP_BOUNDS_CHECK
#define ICP_CHK_IDX(k, n, ctl)                                                              \
    do {                                                                                    \
        if ((unsigned)(k) >= (unsigned)(n)) { atomicCAS((int*)&(ctl)[CTL_BOUNDS], 0, __LINE__); (k) = 0; } \
    } while (0)
#else
#define ICP_CHK_IDX(k, n, ctl) do { } while (0)
#endif
// Loop progress for the host (icp_run keeps only a few iterations enqueued ahead of the device): `progress`
// points into mapped pinned host memory, one word = (iterations closed << 1) | converged.  Called at the entry of
// the FIRST kernel of an iteration, so that the PCIe write drains under that kernel instead of at a kernel end.
__device__ __forceinline__ void publish_progress(int* progress, const int* ctl)
{
    if (progress != nullptr && blockIdx.x == 0 && threadIdx.x == 0)
        *reinterpret_cast<volatile int*>(progress) = (ctl[CTL_ITERS] << 1) | (ctl[CTL_DONE] != 0);
}

__device__ __forceinline__ double warp_sum(double v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
#endif

}  // namespace icpb
