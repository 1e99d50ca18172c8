/*
 * icp_b200.h — C-ABI of the B200-native ICP hot path (libicp_b200.so).
 *
 * Drop-in boundary for the per-iteration path of yassram/iterative-closest-point:
 *   closest_matrix -> centroid / cross-covariance -> Horn alignment (s, R, t) -> transform + residual,
 * iterated by find_corresponding.  Every entry point names the reference interface it replaces
 * (file:line relative to the reference repository).  Plain pointers and sizes only; no C++/torch
 * types.  All functions return ICP_OK (0) or a negative icp_status; nothing here ever calls exit().
 *
 * Conventions
 *   - Clouds cross this boundary as `n` points of 3 doubles, x0 y0 z0 x1 y1 z1 ...  This is exactly the
 *     in-memory image of the reference's column-major 3 x n Eigen::MatrixXd (src/cpu.hh:9-12), so a
 *     maintainer passes `matrix.data()` and `matrix.cols()` unchanged.
 *   - 3x3 matrices (R, sr) are row-major double[9]; vectors are double[3].
 *   - The caller owns every host buffer; the context owns all device memory.  One context per host
 *     thread; calls are synchronous with respect to their outputs.
 *   - There is no CPU fallback: every call that computes needs a CUDA device and fails with
 *     ICP_ERR_CUDA / ICP_ERR_NO_DEVICE otherwise.
 *   - Correspondence indices are bit-exact with the reference's FP64 argmin
 *     (sqrt(dx^2+dy^2+dz^2), first minimum wins, src/cpu.cc:14-22): an FP32 brute-force filter is
 *     followed by an FP64 re-rank that evaluates the reference's expression on every candidate
 *     the filter could not rule out with a rigorous error bound.
 */
#ifndef ICP_B200_H
#define ICP_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum icp_status {
    ICP_OK = 0,
    ICP_ERR_INVALID = -1,      /* bad argument / call order (e.g. no model set) */
    ICP_ERR_NO_DEVICE = -2,    /* no usable CUDA device: there is no CPU fallback */
    ICP_ERR_CUDA = -3,         /* a CUDA runtime call failed; see icp_last_error() */
    ICP_ERR_SIZE_MISMATCH = -4,/* np != nm   — reference: exit(-1), src/cpu.cc:44-47 */
    ICP_ERR_TOO_FEW = -5,      /* np < 4     — reference: exit(-1), src/cpu.cc:49-52 */
    ICP_ERR_COMM = -6          /* NCCL unavailable, a collective failed, or a peer did not arrive at an in-kernel exchange
                                * within 20 s.  After the latter icp_run stops at once: new_p (icp_get_scene) holds the scene
                                * as the last COMPLETED iteration left it, the transform is that iteration's, and the
                                * communicator (mailbox sequence numbers) is no longer usable: destroy the contexts. */
} icp_status;

typedef struct icp_ctx icp_ctx;

/* Counters of the last icp_closest / icp_run on this context (device work only). */
typedef struct icp_stats {
    int64_t np, nm;             /* points held by this context (scene shard, model) */
    int64_t nn_calls;           /* closest-matrix passes executed */
    int64_t nn_pairs;           /* distance evaluations = sum over passes of np*nm */
    int64_t rerank_full;        /* scene points sent to the full FP64 scan (filter ambiguous) */
    int64_t rerank_chunk;       /* scene points resolved by the FP64 re-rank of one model chunk */
    int32_t n_splits;           /* model splits of the 2-D decomposition */
    int32_t tile_points;        /* scene points per CTA */
    int32_t chunk_points;       /* model points per tracking chunk */
    int32_t kernels_launched;   /* kernels launched by the last icp_run / operator call */
    float   ms_filter;          /* CUDA-event time of the FP32 filter kernel, summed over the last run */
    float   ms_total;           /* CUDA-event time of the last icp_run (device loop only) */
    float   ms_refine;          /* ICP_PROFILE=1 only: event time of nn_refine / nn_exact / the fused alignment */
    float   ms_exact;           /*   kernel summed over the last run (the extra events cost ~1 us per iteration) */
    float   ms_align;
    float   reserved_;
} icp_stats;

/* ---- lifetime ------------------------------------------------------------------------------- */

/* Creates a context on CUDA device `device` (streams, scratch).  Replaces nothing in the reference
 * (its wrappers allocate and free per call, src/GPU/compute.cu:154-245,348-379). */
int icp_create(icp_ctx** ctx, int device);
void icp_destroy(icp_ctx* ctx);
const char* icp_last_error(void);
const char* icp_version(void);

/* ---- state: the ICP object (reference: CPU::ICP ctor src/cpu.hh:48-60, GPU::ICP src/GPU/gpu.hh:44-57)
 * icp_set_model  <- `m` (3 x nm).  Builds the FP64 master (one 32-byte record per point) and the FP32 shadow;
 *                   uploaded once (reference re-uploads every call: Matrix::toGpu, src/GPU/compute.cu:29-41).
 *                   Non-finite input: NaN coordinates are handled as the reference's first-minimum scan handles them
 *                   (a NaN model point is never chosen, unless it is point 0, which is then never left; a scene point
 *                   with a NaN / infinite coordinate matches index 0).  A model with an INFINITE coordinate (no finite
 *                   bounding box) is refused with ICP_ERR_INVALID.
 * icp_set_scene  <- `p` / `new_p` (3 x np).  Resets s = 1, R = I, t = 0 like the ctor.
 * icp_get_scene  -> `new_p` (src/cpu.hh:97).
 * icp_reset_scene   restores new_p to the cloud given to the last icp_set_scene without a host copy. */
int icp_set_model(icp_ctx* ctx, const double* m_xyz, int64_t nm);
int icp_set_scene(icp_ctx* ctx, const double* p_xyz, int64_t np);
int icp_get_scene(icp_ctx* ctx, double* p_xyz_out);
int icp_reset_scene(icp_ctx* ctx);

/* ---- how icp_closest / icp_run search (results are identical) ---------------
 * ICP_NN_AUTO (default)   brute force - except inside icp_run (>= 2 iterations) on clouds too small for the brute-force
 *                         kernel to fill one wave of the GPU (up to ~16 384 x 16 384 points; cow), which take the tree.
 * ICP_NN_BRUTE            exact brute force: every scene point meets every model point (north_star).
 * ICP_NN_TREE             exact search through a bounding-box hierarchy over the Morton-sorted model,
 *                         built on the device at the first search after icp_set_model; same first-minimum
 *                         rule on the reference's FP64 expression (src/cpu.cc:16-22), bit-identical
 *                         indices.  Also selected by ICP_NN=tree in the environment at icp_create.
 *                         Models whose extent is outside [1e-12, 1e12] keep the brute-force path
 *                         (icp_get_nn_mode reports what is in effect). */
enum { ICP_NN_BRUTE = 0, ICP_NN_TREE = 1, ICP_NN_AUTO = 2 };
int icp_set_nn_mode(icp_ctx* ctx, int mode);
int icp_get_nn_mode(icp_ctx* ctx);

/* ---- operators -------------------------------------------------------------------------------
 * icp_closest          CPU::closest_matrix(closest_matrix_params)   src/cpu.cc:5-27
 *                      compute_Y_w_opti(m, p, Y)                     src/GPU/gpu.hh:110, compute.cu:154-245
 *                      idx_out[j] (nullable) = argmin index, Y_out (nullable) = m[:, idx[j]].
 * icp_centroid         row means of new_p and Y                      src/cpu.cc:113-114 (bench.cc:146-157)
 * icp_find_alignment   CPU::ICP::find_alignment(Y)                   src/cpu.cc:105-175, src/GPU/gpu.cc:95-151
 *                      on the current correspondences; updates the context's s, R, t and returns
 *                      err_compute_alignment (src/cpu.cc:93-103) in *err_align.
 * icp_err_compute      in_place = 1: CPU::err_compute                 src/cpu.cc:29-40  (new_p <- sR new_p + t)
 *                      in_place = 0: CPU::err_compute_alignment       src/cpu.cc:93-103
 *                      mirrors compute_err_w's flag                   src/GPU/gpu.hh:112-113
 * icp_get_transform    the context's current s, R, t                  src/cpu.hh:100-103 */
int icp_closest(icp_ctx* ctx, int32_t* idx_out, double* Y_xyz_out);
int icp_centroid(icp_ctx* ctx, double mu_p[3], double mu_y[3]);
int icp_find_alignment(icp_ctx* ctx, double* s, double R[9], double t[3], double* err_align);
int icp_err_compute(icp_ctx* ctx, int in_place, double* err);
int icp_get_transform(icp_ctx* ctx, double* s, double R[9], double t[3]);

/* ---- the loop --------------------------------------------------------------------------------
 * icp_run              CPU::ICP::find_corresponding()                 src/cpu.cc:55-79
 *                      GPU::ICP::find_corresponding_opti()            src/GPU/gpu.cc:52-83
 * Runs up to max_iter iterations entirely on the device; stops after the first iteration whose
 * error value (src/cpu.cc:65-73: (err_alignment + err_compute) / np) is < threshold
 * (reference: 1e-5, src/cpu.hh:113).  err_hist (nullable, max_iter doubles) receives the value the
 * reference prints each iteration.  check_sizes != 0 applies alignement_check (src/cpu.cc:42-53)
 * and returns ICP_ERR_SIZE_MISMATCH / ICP_ERR_TOO_FEW where the reference exits.
 * The convergence test runs on the device; the host only throttles how far ahead it enqueues
 * (ICP_LOOKAHEAD iterations, default 3, read at icp_create; 0 = enqueue all max_iter at once). */
int icp_run(icp_ctx* ctx, int max_iter, double threshold, int check_sizes, int* iters_done, double* err_hist);

/* Correspondence indices of the last closest pass (any of icp_closest / icp_run). */
int icp_get_indices(icp_ctx* ctx, int32_t* idx_out);
int icp_get_stats(icp_ctx* ctx, icp_stats* out);
/* Diagnostic (ICP_PROFILE=1 in the environment at icp_create): microseconds, relative to kernel entry, at which
 * block 0 of the last fused alignment kernel passed: 1 moments loop done, 2 block partials written, 3 grid
 * barrier passed, 4 partials reduced (+ peer exchange), 5 Horn solve done, 6 transform loop done, 7 exit of
 * the LAST block; slot 0 = microseconds between the end of the preceding nearest-neighbour kernel and entry (large
 * models only: nn_exact stamps it); slots 8, 9 split step 4: partials re-read, block sum done. */
int icp_get_trace(icp_ctx* ctx, double out_us[10]);
/* Same diagnostic, per block: for each of the first 32 blocks of that kernel (SM id, then microseconds after block 0's
 * entry of: own entry, grid barrier passed, Horn solve done, exit); -1 = not stamped. */
int icp_get_block_trace(icp_ctx* ctx, double out[160]);

/* ---- stateless one-shot forms (host pointers in, host results out) — back the reference-shaped
 * wrappers of src/GPU/gpu.hh:110-116.  Stateless for the caller; internally each host thread keeps one hidden
 * context on device 0 whose buffers are reused from call to call (creating one costs ~3 ms).
 * icp_closest_matrix_host   compute_Y_w_opti / CPU::closest_matrix
 * icp_nearest_index_host    compute_distance_w_naive(m, pi) -> int     src/GPU/compute.cu:279-308
 * icp_find_alignment_host   ICP::find_alignment(Y) on explicit pairs    src/cpu.cc:105-175, src/GPU/gpu.cc:95-151
 * icp_err_compute_host      compute_err_w(Y, p, in_place, sr, t)        src/GPU/compute.cu:348-379
 * icp_subtract_col_host     substract_col_w(M, m)                       src/GPU/compute.cu:400-416
 * icp_y_p_norm_host         y_p_norm_w(y, p, n, d_caps, sp)             src/GPU/compute.cu:442-469 */
int icp_closest_matrix_host(const double* m_xyz, int64_t nm, const double* p_xyz, int64_t np,
                            int32_t* idx_out, double* Y_xyz_out);
void icp_release_cached(void);   /* frees the calling thread's hidden context and icp_batch_run's idle lane contexts; optional */
int icp_nearest_index_host(const double* m_xyz, int64_t nm, const double pi[3], int32_t* idx_out);
int icp_find_alignment_host(const double* Y_xyz, const double* p_xyz, int64_t np, double* s, double R[9],
                            double t[3], double* err_align);
int icp_err_compute_host(const double* Y_xyz, double* p_xyz, int64_t np, int in_place,
                         const double sr[9], const double t[3], double* err);
int icp_subtract_col_host(const double* M_xyz, int64_t n, const double m[3], double* out_xyz);
int icp_y_p_norm_host(const double* y_xyz, const double* p_xyz, int64_t n, double* d_caps, double* sp);

/* ---- batched registration (new: BASELINE.json configs[4], many independent pairs) ----------------------
 * Runs CPU::ICP::find_corresponding (src/cpu.cc:55-79) on n_pairs independent (model, scene) pairs on one
 * device.  `lanes` (1..8) contexts work concurrently, each on its own stream, so that one pair's latency-bound
 * alignment steps and PCIe copies hide behind another pair's nearest-neighbour kernel.  Outputs (all nullable):
 * iters_out[k], err_out[k] (last printed error value), new_p_out[k] (np[k] x 3 doubles), srt_out[13 k ..] =
 * s, R[9] row-major, t[3] of the last iteration.  Across GPUs: one process per GPU, pairs dealt round-robin by
 * the launcher; nothing is exchanged. */
int icp_batch_run(int device, int lanes, int64_t n_pairs, const double* const* m_xyz, const int64_t* nm,
                  const double* const* p_xyz, const int64_t* np, int max_iter, double threshold, int check_sizes,
                  int32_t* iters_out, double* err_out, double* const* new_p_out, double* srt_out);
/* Kernels of this library launched by the last icp_batch_run of the process (loop kernels as counted by icp_run + the
 * preparation kernels of each pair; the radix sort of the scene is cub's and not counted). */
long long icp_batch_kernels_launched(void);

/* ---- multi-GPU (new; the reference is single-GPU) ----------------------------------------------
 * One context per GPU / rank holds a contiguous block of the scene and a replica of the model.
 * The per-iteration exchange is an all-reduce(sum) of 18 doubles of moments and of 1 double of
 * residual, issued in-stream through NCCL.  icp_comm_unique_id fills a 128-byte NCCL unique id
 * (rank 0 creates it, the launcher broadcasts it); icp_comm_init joins the communicator.
 * nccl_library_path may be NULL ("libnccl.so.2" is resolved by the dynamic loader).
 * With a communicator attached, `np` in the error value and in the means is the GLOBAL scene size. */
int icp_comm_unique_id(const char* nccl_library_path, void* id128);
int icp_comm_init(icp_ctx* ctx, const char* nccl_library_path, const void* id128, int rank, int world);
int icp_comm_world(icp_ctx* ctx, int* rank, int* world);
/* Peer-memory exchange (preferred over the NCCL all-reduce once attached): every rank exports the CUDA IPC
 * handle (64 bytes) of a small mailbox in its HBM, the launcher all-gathers the handles, and every rank
 * maps its peers' mailboxes.  From then on the alignment kernel itself posts its 18 + 1 partial sums into
 * all peers' mailboxes over NVLink and adds what it received in rank order: no collective launch per
 * iteration.  One process per GPU; icp_comm_init must have been called (the size check still uses NCCL).
 * ICP_COMM=nccl in the environment keeps the NCCL path for comparison. */
int icp_comm_peer_handle(icp_ctx* ctx, void* handle64);
int icp_comm_peer_attach(icp_ctx* ctx, const void* handles /* world x 64 bytes, rank order */, int rank, int world);

/* ---- one process, several GPUs: the multi-device form of icp_create (icp_create(ctx, devices, ndev)) ----
 * An icp_group owns one context per listed device.  icp_group_set_model replicates the model, icp_group_set_scene cuts
 * the scene into `ndev` contiguous blocks (device i holds block i); icp_group_run is find_corresponding
 * (src/cpu.cc:55-79) over the sharded scene: per iteration the 18 moment sums and the residual are exchanged INSIDE the
 * alignment kernels through mailboxes in peer memory (cudaDeviceEnablePeerAccess over NVLink; no NCCL, no launcher, no
 * torch), every device solves for the identical (s, R, t), and results come back in the caller's point order.  Indices,
 * iteration count and error values are those of the single-device loop.  Devices must be distinct and peer-accessible
 * (ICP_ERR_COMM otherwise); with one device the group is a plain context.  The calls below are collective over the
 * group's devices (each device is driven by its own host thread for the duration of the call) and otherwise mirror
 * their icp_* namesakes.  icp_group_member gives the per-device context (for icp_get_stats / icp_get_trace only). */
typedef struct icp_group icp_group;
int icp_group_create(icp_group** grp, const int* devices, int ndev);
void icp_group_destroy(icp_group* grp);
int icp_group_size(icp_group* grp);
icp_ctx* icp_group_member(icp_group* grp, int i);
int icp_group_set_model(icp_group* grp, const double* m_xyz, int64_t nm);
int icp_group_set_scene(icp_group* grp, const double* p_xyz, int64_t np);
int icp_group_reset_scene(icp_group* grp);
int icp_group_set_nn_mode(icp_group* grp, int mode);
int icp_group_closest(icp_group* grp, int32_t* idx_out, double* Y_out);
int icp_group_centroid(icp_group* grp, double mu_p[3], double mu_y[3]);
int icp_group_find_alignment(icp_group* grp, double* s, double R[9], double t[3], double* err_align);
int icp_group_err_compute(icp_group* grp, int in_place, double* err);
int icp_group_run(icp_group* grp, int max_iter, double threshold, int check_sizes, int* iters_done, double* err_hist);
int icp_group_get_scene(icp_group* grp, double* p_xyz_out);
int icp_group_get_indices(icp_group* grp, int32_t* idx_out);
int icp_group_get_transform(icp_group* grp, double* s, double R[9], double t[3]);
int icp_group_get_stats(icp_group* grp, icp_stats* out);   /* sizes, pairs, launches summed; times of the slowest device */
/* icp_batch_run over several devices of this process: pair k is registered on device k % ndev, `lanes` contexts per device. */
int icp_batch_run_multi(const int* devices, int ndev, int lanes, int64_t n_pairs, const double* const* m_xyz,
                        const int64_t* nm, const double* const* p_xyz, const int64_t* np, int max_iter, double threshold,
                        int check_sizes, int32_t* iters_out, double* err_out, double* const* new_p_out, double* srt_out);

/* ---- no-allocation entry point -------------------------------------------------------------------------------------
 * No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * icp_find_alignment_noalloc = the fused moments + Horn solve (K2 + K3; the s, R, t of icp_find_alignment_host, without
 * its residual) on PAIRED clouds that already live in device memory: d_y_xyz[j] <-> d_p_xyz[j], n x 3 FP64 each, on
 * the CURRENT device.  d_out20 (device, 20 doubles) receives s, R[9] row-major, t[3], mu_p[3], mu_y[3], 0.  The work is
 * enqueued on `cuda_stream` (a cudaStream_t; NULL = the default stream) and the call returns without synchronising.
 * It calls no cudaMalloc / cudaFree / allocating routine and needs no icp_ctx: the kernel's scratch (block partials and a
 * ticket) is ONE file-scope device array of 80 768 bytes that every GPU loading the library holds until the process
 * exits.  Limits and sharing: n <= 4 194 304 point pairs (ICP_ERR_INVALID beyond, with a message); all calls on a device
 * share the array, so calls that overlap on the GPU (different streams, or several host threads) RACE - serialise them.
 * The very first call on a device resolves the array's address (cudaGetSymbolAddress): make it outside a stream capture. */
int icp_find_alignment_noalloc(const double* d_y_xyz, const double* d_p_xyz, int64_t n, double* d_out20, void* cuda_stream);

#ifdef __cplusplus
}
#endif
#endif /* ICP_B200_H */
