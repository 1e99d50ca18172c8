// align_kernels.cu — K2 (fused moments), K3 (Horn quaternion solve), K4 (fused transform + residual)
// for sm_100a.  All FP64, HBM-bound linear passes; they replace, without porting,
//   host means + substract_col + host 3xn.nx3 product + y_p_norm   (src/GPU/gpu.cc:98-104,142; compute.cu:381-440)
//   host Eigen::EigenSolver on Horn's N                             (src/cpu.cc:121-152, src/GPU/gpu.cc:106-140)
//   compute_err (both uses)                                         (src/GPU/compute.cu:315-379; src/cpu.cc:29-40,93-103)
//
// K2 reads each scene point once (24 B) with its matched model point (24 B, streamed from SceneDev::y64*, where the
// search kernel that decided the match recorded it; algorithmically: correspondence 4 B + gathered model point 24 B)
// and accumulates 18 sums about the pivot c0 (centre of the model's bounding box):
//   n, sum a, sum b, sum a b^T, sum |a|^2, sum |b|^2      with a = p - c0, b = y - c0.
// Centring is algebraic (S = sum a b^T - n mu_a mu_b^T, ...), so the centred copies p', y' that the
// reference materialises (src/cpu.cc:116-117) never exist.  Reduction is a fixed two-level tree
// (block partials -> last block), so results are run-to-run deterministic.
#include <cooperative_groups.h>

#include "icp_common.cuh"

namespace cg = cooperative_groups;

namespace icpb {

namespace {

// block-wide sum of N doubles held per thread; result valid in thread 0
template <int N>
__device__ __forceinline__ void block_sum(double (&v)[N], double* smem /* N * warps */)
{
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = blockDim.x >> 5;
#pragma unroll
    for (int i = 0; i < N; ++i) v[i] = warp_sum(v[i]);
    if (lane == 0)
#pragma unroll
        for (int i = 0; i < N; ++i) smem[i * nw + w] = v[i];
    __syncthreads();
    if (w == 0) {
#pragma unroll
        for (int i = 0; i < N; ++i) {
            double x = lane < nw ? smem[i * nw + lane] : 0.0;
            v[i] = warp_sum(x);
        }
    }
    __syncthreads();
}

__device__ void horn_solve(const double* __restrict__ geo, const double* __restrict__ mom, double* __restrict__ xf);

// ------------------------------------------------------------------------------------------------
// Peer-memory all-reduce(sum) of n <= PEER_SLOT doubles, executed by the FIRST WARP of one block.
// See PeerComm in icp_common.cuh.  `stage` is shared memory, PEER_MAX_WORLD x PEER_SLOT doubles.
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p)
{
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v)
{
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
#ifndef ICP_PEER_TIMEOUT_NS
#define ICP_PEER_TIMEOUT_NS 20000000000ull
#endif
constexpr unsigned long long PEER_TIMEOUT_NS = ICP_PEER_TIMEOUT_NS;    // 20 s
__device__ __forceinline__ unsigned long long globaltimer_ns()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// Returns false when a peer did not arrive within 20 s: CTL_COMM_TIMEOUT and CTL_DONE are set (every later kernel of
// the loop returns at entry), nothing is consumed and the exchange number does not advance, so new_p stays as the
// last COMPLETED iteration left it; the caller must not use `out`.
__device__ bool peer_allreduce(const PeerComm& pc, const double* vals, int n, double* out, double* stage, int* ctl)
{
    const int lane = threadIdx.x;                    // caller: threadIdx.x < 32
    // The exchange number lives in this rank's own mailbox and advances only when an exchange really happens, so
    // the ranks stay in step however many (no-op) launches each host has queued after convergence.
    PeerMailbox* self = pc.box[pc.rank];
    const unsigned long long seq = ld_acquire_sys(&self->seq_local) + 1ull;
    __syncwarp();
    const int par = (int)(seq & 1ull);
    bool arrived = true;
    if (lane < pc.world) {
        PeerMailbox* dst = pc.box[lane];             // post to rank `lane` (own mailbox included)
        for (int i = 0; i < n; ++i) dst->data[par][pc.rank][i] = vals[i];
        // the data stores above are this thread's own: the system-scope RELEASE of the flag orders them before it for
        // any reader that acquires the flag (round 1 had a full __threadfence_system() here as well: one fence too many)
        st_release_sys(&dst->flag[par][pc.rank], seq);
        // wait for rank `lane`'s post in my own mailbox
        PeerMailbox* mine = pc.box[pc.rank];
        const unsigned long long t0 = globaltimer_ns();
        while (ld_acquire_sys(&mine->flag[par][lane]) != seq) {
            if (globaltimer_ns() - t0 > PEER_TIMEOUT_NS) { arrived = false; break; }   // a peer died
        }
        if (arrived)
            for (int i = 0; i < n; ++i) stage[lane * PEER_SLOT + i] = __ldcv(&mine->data[par][lane][i]);
    }
    __syncwarp();
    if (__any_sync(0xffffffffu, !arrived)) {
        if (lane == 0) { ctl[CTL_COMM_TIMEOUT] = 1; ctl[CTL_DONE] = 1; __threadfence(); }
        __syncwarp();
        return false;
    }
    if (lane < n) {
        double acc = 0.0;
        for (int r = 0; r < pc.world; ++r) acc += stage[r * PEER_SLOT + lane];     // rank order: identical everywhere
        out[lane] = acc;
    }
    if (lane == 0) { self->seq_local = seq; __threadfence(); }
    __syncwarp();
    return true;
}

// ------------------------------------------------------------------------------------------------
// K2: moments.
// The streaming loop keeps TWO scene points in flight per thread (both index loads, then both gathers, then the sums) and
// the kernel is bounded to 64 registers (4 CTAs per SM): on multi-million-point clouds the pass is HBM-latency bound and
// needs ~6.5 MB in flight (round 1: one point per thread at 2 CTAs per SM = 37-40 % of the HBM copy peak on 2^22 points).
// The Horn solve rides on the last block (one thread, once per launch); its spills under the register bound do not matter.
__device__ __forceinline__ void moments_accumulate(double (&acc)[17], double ax, double ay, double az, double bx, double by,
                                                   double bz)
{
    acc[0] += ax; acc[1] += ay; acc[2] += az;
    acc[3] += bx; acc[4] += by; acc[5] += bz;
    acc[6] += ax * bx; acc[7] += ax * by; acc[8] += ax * bz;
    acc[9] += ay * bx; acc[10] += ay * by; acc[11] += ay * bz;
    acc[12] += az * bx; acc[13] += az * by; acc[14] += az * bz;
    acc[15] += ax * ax + ay * ay + az * az;
    acc[16] += bx * bx + by * by + bz * bz;
}

// Y = the matched model points, written by the search kernels (SceneDev::y64*): p and Y are streamed, nothing is gathered.
struct MatchIn {
    const double *yx, *yy, *yz;
};

__device__ __forceinline__ void moments_points(const MatchIn y, const double* __restrict__ p64x,
                                               const double* __restrict__ p64y, const double* __restrict__ p64z,
                                               int np, int n_blocks, double c0x, double c0y, double c0z, double (&acc)[17])
{
    const int stride = n_blocks * RED_THREADS;
    int j = blockIdx.x * RED_THREADS + threadIdx.x;
    for (; j + stride < np; j += 2 * stride) {          // two points in flight per thread
        const double a0x = p64x[j], a0y = p64y[j], a0z = p64z[j];
        const double a1x = p64x[j + stride], a1y = p64y[j + stride], a1z = p64z[j + stride];
        const double b0x = y.yx[j], b0y = y.yy[j], b0z = y.yz[j];
        const double b1x = y.yx[j + stride], b1y = y.yy[j + stride], b1z = y.yz[j + stride];
        moments_accumulate(acc, a0x - c0x, a0y - c0y, a0z - c0z, b0x - c0x, b0y - c0y, b0z - c0z);
        moments_accumulate(acc, a1x - c0x, a1y - c0y, a1z - c0z, b1x - c0x, b1y - c0y, b1z - c0z);
    }
    if (j < np)
        moments_accumulate(acc, p64x[j] - c0x, p64y[j] - c0y, p64z[j] - c0z, y.yx[j] - c0x, y.yy[j] - c0y, y.yz[j] - c0z);
}

__global__ void __launch_bounds__(RED_THREADS, 4)
moments_kernel(const MatchIn y,
               const double* __restrict__ geo, const double* __restrict__ p64x, const double* __restrict__ p64y,
               const double* __restrict__ p64z, int np, double* __restrict__ partials,
               double* __restrict__ mom, int* __restrict__ ctl, int honor_done, double* __restrict__ xf_fused,
               PeerComm pc)
{
    if (honor_done && ctl[CTL_DONE]) return;
    __shared__ double red[17 * (RED_THREADS / 32)];
    __shared__ double mom_s[MOM_N];
    __shared__ double stage[PEER_MAX_WORLD * PEER_SLOT];
    __shared__ int is_last;
    const double c0x = geo[0], c0y = geo[1], c0z = geo[2];
    double acc[17];
#pragma unroll
    for (int i = 0; i < 17; ++i) acc[i] = 0.0;
    moments_points(y, p64x, p64y, p64z, np, (int)gridDim.x, c0x, c0y, c0z, acc);
    block_sum<17>(acc, red);
    if (threadIdx.x == 0) {
#pragma unroll
        for (int i = 0; i < 17; ++i) partials[(size_t)blockIdx.x * 17 + i] = acc[i];
        __threadfence();
        const int t = atomicAdd(&ctl[CTL_TICKET_MOM], 1);
        is_last = (t == (int)gridDim.x - 1);
    }
    __syncthreads();
    if (is_last) {
        __threadfence();
        double tot[17];
#pragma unroll
        for (int i = 0; i < 17; ++i) tot[i] = 0.0;
        for (int b = threadIdx.x; b < (int)gridDim.x; b += RED_THREADS)
#pragma unroll
            for (int i = 0; i < 17; ++i) tot[i] += __ldcg(&partials[(size_t)b * 17 + i]);
        block_sum<17>(tot, red);
        if (threadIdx.x == 0) {
            mom_s[0] = (double)np;
#pragma unroll
            for (int i = 0; i < 17; ++i) mom_s[1 + i] = tot[i];
            ctl[CTL_TICKET_MOM] = 0;
        }
        __syncthreads();
        if (threadIdx.x < 32) {
            // multi-GPU: the 18 sums of all ranks meet here, inside the kernel (peer mailboxes over NVLink)
            bool ok = true;
            if (pc.world > 0) ok = peer_allreduce(pc, mom_s, MOM_N, mom_s, stage, ctl);
            if (ok) {
                if (threadIdx.x < MOM_N) mom[threadIdx.x] = mom_s[threadIdx.x];
                __syncwarp();
                if (threadIdx.x == 0 && xf_fused) horn_solve(geo, mom_s, xf_fused);   // K3 rides on K2's last block
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------
// K3: Horn's closed form.  One warp; the 4x4 symmetric eigenproblem is solved with cyclic Jacobi by
// lane 0 (a few hundred FP64 flops) and the eigenvector of the LARGEST eigenvalue is taken — the
// rule the reference intends (src/cpu.cc:131; its max_element_index has a bug, see oracle).
__device__ void jacobi4_dev(double A[4][4], double V[4][4])
{
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j) V[i][j] = (i == j) ? 1.0 : 0.0;
    for (int sweep = 0; sweep < 32; ++sweep) {
        double off = 0.0;
        for (int i = 0; i < 4; ++i)
            for (int j = i + 1; j < 4; ++j) off += A[i][j] * A[i][j];
        if (off == 0.0) break;
        for (int p = 0; p < 3; ++p)
            for (int q = p + 1; q < 4; ++q) {
                const double apq = A[p][q];
                if (apq == 0.0) continue;
                const double theta = (A[q][q] - A[p][p]) / (2.0 * apq);
                const double t = (theta >= 0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
                const double c = 1.0 / sqrt(t * t + 1.0), s = t * c;
                for (int k = 0; k < 4; ++k) {
                    const double akp = A[k][p], akq = A[k][q];
                    A[k][p] = c * akp - s * akq;
                    A[k][q] = s * akp + c * akq;
                }
                for (int k = 0; k < 4; ++k) {
                    const double apk = A[p][k], aqk = A[q][k];
                    A[p][k] = c * apk - s * aqk;
                    A[q][k] = s * apk + c * aqk;
                }
                for (int k = 0; k < 4; ++k) {
                    const double vkp = V[k][p], vkq = V[k][q];
                    V[k][p] = c * vkp - s * vkq;
                    V[k][q] = s * vkp + c * vkq;
                }
            }
    }
}

// Fast path for the largest eigenpair of the symmetric, trace-free 4x4 matrix N:
//   characteristic polynomial  l^4 + c2 l^2 + c1 l + c0   (c2 = -tr(N^2)/2, c1 = -tr(N^3)/3, c0 = det N),
//   Newton from an upper bound of the spectrum (monotone convergence to the largest root because
//   all roots are real), eigenvector = best-conditioned column of adj(N - l I).
// Returns false when the result fails the residual test |N q - l q| <= 1e-11 |N|_F (close or
// repeated top eigenvalues): the caller then falls back to Jacobi.
__device__ bool top_eigenpair_fast(const double N[4][4], double upper, double q[4])
{
    double fro2 = 0.0;
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j) fro2 += N[i][j] * N[i][j];
    if (!(fro2 > 0.0)) return false;
    const double fro = sqrt(fro2);
    double N2[4][4];
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j) {
            double a = 0.0;
            for (int k = 0; k < 4; ++k) a += N[i][k] * N[k][j];
            N2[i][j] = a;
        }
    double tr3 = 0.0;
    for (int i = 0; i < 4; ++i)
        for (int k = 0; k < 4; ++k) tr3 += N2[i][k] * N[k][i];
    // det N by cofactor expansion over 2x2 minors of rows (0,1) x rows (2,3)
    const double s0 = N[0][0] * N[1][1] - N[1][0] * N[0][1], s1 = N[0][0] * N[1][2] - N[1][0] * N[0][2];
    const double s2 = N[0][0] * N[1][3] - N[1][0] * N[0][3], s3 = N[0][1] * N[1][2] - N[1][1] * N[0][2];
    const double s4 = N[0][1] * N[1][3] - N[1][1] * N[0][3], s5 = N[0][2] * N[1][3] - N[1][2] * N[0][3];
    const double t5 = N[2][2] * N[3][3] - N[3][2] * N[2][3], t4 = N[2][1] * N[3][3] - N[3][1] * N[2][3];
    const double t3 = N[2][1] * N[3][2] - N[3][1] * N[2][2], t2 = N[2][0] * N[3][3] - N[3][0] * N[2][3];
    const double t1 = N[2][0] * N[3][2] - N[3][0] * N[2][2], t0 = N[2][0] * N[3][1] - N[3][0] * N[2][1];
    const double c0 = s0 * t5 - s1 * t4 + s2 * t3 + s3 * t2 - s4 * t1 + s5 * t0;
    const double c2 = -0.5 * fro2, c1 = -tr3 / 3.0;
    double l = fmin(fro, upper > 0.0 ? upper : fro) * (1.0 + 1e-12);
    bool conv = false;
    for (int it = 0; it < 64; ++it) {
        const double l2 = l * l;
        const double P = (l2 + c2) * l2 + c1 * l + c0;
        const double dP = (4.0 * l2 + 2.0 * c2) * l + c1;
        if (!(dP > 0.0)) break;
        const double step = P / dP;
        l -= step;
        if (fabs(step) <= 1e-15 * fabs(l)) { conv = true; break; }
    }
    if (!conv) return false;
    // adj(A), A = N - l I (symmetric).  adj = c q q^T: take the column with the largest diagonal entry.
    double A[4][4];
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j) A[i][j] = N[i][j] - (i == j ? l : 0.0);
    double adj[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            // cofactor C_ij = (-1)^(i+j) det(A without row i, col j); adj = C^T = C (symmetric)
            const int r0 = i == 0 ? 1 : 0, r1 = i <= 1 ? 2 : 1, r2 = i <= 2 ? 3 : 2;
            const int k0 = j == 0 ? 1 : 0, k1 = j <= 1 ? 2 : 1, k2 = j <= 2 ? 3 : 2;
            const double d = A[r0][k0] * (A[r1][k1] * A[r2][k2] - A[r1][k2] * A[r2][k1]) -
                             A[r0][k1] * (A[r1][k0] * A[r2][k2] - A[r1][k2] * A[r2][k0]) +
                             A[r0][k2] * (A[r1][k0] * A[r2][k1] - A[r1][k1] * A[r2][k0]);
            adj[i][j] = ((i + j) & 1) ? -d : d;
        }
    int col = 0;
    double bestd = fabs(adj[0][0]);
#pragma unroll
    for (int i = 1; i < 4; ++i)
        if (fabs(adj[i][i]) > bestd) { bestd = fabs(adj[i][i]); col = i; }
    double v[4] = {col == 0 ? adj[0][0] : col == 1 ? adj[0][1] : col == 2 ? adj[0][2] : adj[0][3],
                   col == 0 ? adj[1][0] : col == 1 ? adj[1][1] : col == 2 ? adj[1][2] : adj[1][3],
                   col == 0 ? adj[2][0] : col == 1 ? adj[2][1] : col == 2 ? adj[2][2] : adj[2][3],
                   col == 0 ? adj[3][0] : col == 1 ? adj[3][1] : col == 2 ? adj[3][2] : adj[3][3]};
    const double nn = v[0] * v[0] + v[1] * v[1] + v[2] * v[2] + v[3] * v[3];
    if (!(nn > 0.0)) return false;
    const double inv = 1.0 / sqrt(nn);
    double res2 = 0.0;
    for (int i = 0; i < 4; ++i) q[i] = v[i] * inv;
    for (int i = 0; i < 4; ++i) {
        const double ri = N[i][0] * q[0] + N[i][1] * q[1] + N[i][2] * q[2] + N[i][3] * q[3] - l * q[i];
        res2 += ri * ri;
    }
    return res2 <= 1e-22 * fro2;
}

// s, R, t from the 18 moment sums (src/cpu.cc:113-167).  Executed by ONE thread.
__device__ void horn_solve(const double* __restrict__ geo, const double* __restrict__ mom, double* __restrict__ xf)
{
    const double n = mom[0];
    const double mua[3] = {mom[1] / n, mom[2] / n, mom[3] / n};
    const double mub[3] = {mom[4] / n, mom[5] / n, mom[6] / n};
    double S[3][3];
    for (int a = 0; a < 3; ++a)
        for (int b = 0; b < 3; ++b) S[a][b] = mom[7 + a * 3 + b] - n * mua[a] * mub[b];   // src/cpu.cc:119
    // Horn's N, src/cpu.cc:121-126
    double N[4][4] = {
        {S[0][0] + S[1][1] + S[2][2], S[1][2] - S[2][1], S[2][0] - S[0][2], S[0][1] - S[1][0]},
        {S[1][2] - S[2][1], S[0][0] - S[2][2] - S[1][1], S[0][1] + S[1][0], S[0][2] + S[2][0]},
        {S[2][0] - S[0][2], S[1][0] + S[0][1], S[1][1] - S[2][2] - S[0][0], S[1][2] + S[2][1]},
        {S[0][1] - S[1][0], S[2][0] + S[0][2], S[2][1] + S[1][2], S[2][2] - S[1][1] - S[0][0]}};
    const double d_caps = mom[17] - n * (mub[0] * mub[0] + mub[1] * mub[1] + mub[2] * mub[2]);
    const double sp = mom[16] - n * (mua[0] * mua[0] + mua[1] * mua[1] + mua[2] * mua[2]);
    double q[4];
    if (!top_eigenpair_fast(N, 0.5 * (d_caps + sp), q)) {
        double V[4][4];
        jacobi4_dev(N, V);
        int kmax = 0;
        for (int i = 1; i < 4; ++i)
            if (N[i][i] > N[kmax][kmax]) kmax = i;
        for (int i = 0; i < 4; ++i) q[i] = V[i][kmax];
    }
    const double q0 = q[0], q1 = q[1], q2 = q[2], q3 = q[3];
    // R = (q_bar^T q_caps)[1:4,1:4], src/cpu.cc:138-152
    const double qb[4][4] = {{q0, -q1, -q2, -q3}, {q1, q0, q3, -q2}, {q2, -q3, q0, q1}, {q3, q2, -q1, q0}};
    const double qc[4][4] = {{q0, -q1, -q2, -q3}, {q1, q0, -q3, q2}, {q2, q3, q0, -q1}, {q3, -q2, q1, q0}};
    double R[9];
#pragma unroll
    for (int i = 1; i < 4; ++i)
#pragma unroll
        for (int j = 1; j < 4; ++j) {
            double acc = 0.0;
#pragma unroll
            for (int k = 0; k < 4; ++k) acc += qb[k][i] * qc[k][j];
            R[(i - 1) * 3 + (j - 1)] = acc;
        }
    // scale and translation, src/cpu.cc:154-167
    const double s = sqrt(d_caps / sp);
    const double mup[3] = {mua[0] + geo[0], mua[1] + geo[1], mua[2] + geo[2]};
    const double muy[3] = {mub[0] + geo[0], mub[1] + geo[1], mub[2] + geo[2]};
    xf[0] = s;
    for (int i = 0; i < 9; ++i) xf[1 + i] = R[i];
    for (int a = 0; a < 3; ++a)
        xf[10 + a] = muy[a] - ((s * R[a * 3] * mup[0] + s * R[a * 3 + 1] * mup[1]) + s * R[a * 3 + 2] * mup[2]);
    for (int a = 0; a < 3; ++a) { xf[13 + a] = mup[a]; xf[16 + a] = muy[a]; }
}

__global__ void solve_kernel(const double* __restrict__ geo, const double* __restrict__ mom, double* __restrict__ xf,
                             const int* __restrict__ ctl, int honor_done)
{
    if (honor_done && ctl[CTL_DONE]) return;
    if (threadIdx.x == 0) horn_solve(geo, mom, xf);
}

// The per-point body of K4, shared by transform_kernel and align_fused_kernel.  Two scene points are in flight per
// thread (j and j + stride: both index loads, both gathers, then the arithmetic; the per-thread summation order is the
// same as with one point per trip).  Warps stay converged over the last (ragged) 32 points so that the warp holding
// point np-1 can refresh the FP32 shadow's padding [np, np_pad) cooperatively (one thread doing it alone costs ~20 us
// on horse: 667 dependent stores).
struct XfPoint {
    double x, y, z, mx, my, mz;
    int j;
    bool act;
};

__device__ __forceinline__ XfPoint xf_load(const MatchIn y, const double* __restrict__ p64x,
                                           const double* __restrict__ p64y, const double* __restrict__ p64z, int np, int j0)
{
    XfPoint q;
    q.act = j0 < np;
    q.j = q.act ? j0 : np - 1;         // inactive lanes of the last warp re-read point np-1, which lives in this same warp
    q.x = p64x[q.j]; q.y = p64y[q.j]; q.z = p64z[q.j];
    q.mx = y.yx[q.j]; q.my = y.yy[q.j]; q.mz = y.yz[q.j];
    return q;
}

__device__ __forceinline__ void xf_apply(const XfPoint& q, int j0, double* __restrict__ p64x, double* __restrict__ p64y,
                                         double* __restrict__ p64z, float* __restrict__ phx, float* __restrict__ phy,
                                         float* __restrict__ phz, int np, int np_pad, const double (&sr)[9], double tx,
                                         double ty, double tz, double c0x, double c0y, double c0z, int in_place, double& acc)
{
    const double x = q.x, y = q.y, z = q.z;
    const double qx = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(sr[0], x), __dmul_rn(sr[1], y)), __dmul_rn(sr[2], z)), tx);
    const double qy = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(sr[3], x), __dmul_rn(sr[4], y)), __dmul_rn(sr[5], z)), ty);
    const double qz = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(sr[6], x), __dmul_rn(sr[7], y)), __dmul_rn(sr[8], z)), tz);
    const double ex = __dsub_rn(q.mx, qx), ey = __dsub_rn(q.my, qy), ez = __dsub_rn(q.mz, qz);
    if (q.act) acc += __dadd_rn(__dadd_rn(__dmul_rn(ex, ex), __dmul_rn(ey, ey)), __dmul_rn(ez, ez));
    if (in_place) {                                   // uniform
        float fx = (float)(qx - c0x), fy = (float)(qy - c0y), fz = (float)(qz - c0z);
        if (q.act) {
            p64x[q.j] = qx; p64y[q.j] = qy; p64z[q.j] = qz;
            phx[q.j] = fx; phy[q.j] = fy; phz[q.j] = fz;
        }
        const unsigned has_last = __ballot_sync(0xffffffffu, j0 == np - 1);
        if (has_last) {                               // warp-uniform: refresh the padding = last real point
            const int src = __ffs(has_last) - 1;
            fx = __shfl_sync(0xffffffffu, fx, src);
            fy = __shfl_sync(0xffffffffu, fy, src);
            fz = __shfl_sync(0xffffffffu, fz, src);
            for (int q2 = np + (int)(threadIdx.x & 31); q2 < np_pad; q2 += 32) { phx[q2] = fx; phy[q2] = fy; phz[q2] = fz; }
        }
    }
}

__device__ __forceinline__ void transform_points(const MatchIn y, double* __restrict__ p64x,
                                                 double* __restrict__ p64y, double* __restrict__ p64z,
                                                 float* __restrict__ phx, float* __restrict__ phy,
                                                 float* __restrict__ phz, int np,
                                                 int np_pad, const double (&sr)[9], double tx, double ty, double tz,
                                                 double c0x, double c0y, double c0z, int in_place, int n_blocks,
                                                 double& acc)
{
    const int np_round = (np + 31) & ~31;
    const int stride = n_blocks * RED_THREADS;
    for (int jA = blockIdx.x * RED_THREADS + threadIdx.x; jA < np_round; jA += 2 * stride) {
        const int jB = jA + stride;
        const bool hasB = jB < np_round;              // warp-uniform (32 consecutive, 32-aligned indices per warp)
        const XfPoint A = xf_load(y, p64x, p64y, p64z, np, jA);
        XfPoint B = A;
        if (hasB) B = xf_load(y, p64x, p64y, p64z, np, jB);
        // every lane has READ its points before any lane of this warp writes (only the warp that holds point np-1 has
        // inactive lanes, and they re-read that very point)
        __syncwarp();
        xf_apply(A, jA, p64x, p64y, p64z, phx, phy, phz, np, np_pad, sr, tx, ty, tz, c0x, c0y, c0z, in_place, acc);
        if (hasB) xf_apply(B, jB, p64x, p64y, p64z, phx, phy, phz, np, np_pad, sr, tx, ty, tz, c0x, c0y, c0z, in_place, acc);
    }
}

// ------------------------------------------------------------------------------------------------
// K4: fused transform + residual.  Per point q = sR p + t (evaluated as the reference does:
// sr = s*R element-wise, then (sr0 x + sr1 y + sr2 z) + t, no FMA contraction), e = |y - q|^2.
// in_place = 1 (CPU::err_compute, src/cpu.cc:29-40): q overwrites new_p (FP64 master + FP32 shadow).
// in_place = 0 (CPU::err_compute_alignment, src/cpu.cc:93-103): residual only.
// finish = 1: the last block also closes the ICP iteration (src/cpu.cc:67-77): the reference adds the
// two residual sums, which are the same expression evaluated twice, so err = 2*sum/np.
__global__ void __launch_bounds__(RED_THREADS)
transform_kernel(const MatchIn y,
                 const double* __restrict__ geo, double* __restrict__ p64x, double* __restrict__ p64y,
                 double* __restrict__ p64z, float* __restrict__ phx, float* __restrict__ phy, float* __restrict__ phz,
                 int np, int np_pad, const double* __restrict__ xf, int in_place,
                 double* __restrict__ partials, double* __restrict__ err_sum, int* __restrict__ ctl, int honor_done,
                 int finish, double n_global, double threshold, double* __restrict__ err_hist, PeerComm pc)
{
    if (honor_done && ctl[CTL_DONE]) return;
    __shared__ double red[RED_THREADS / 32];
    __shared__ double tot_s[1];
    __shared__ double stage[PEER_MAX_WORLD * PEER_SLOT];
    __shared__ int is_last;
    const double s = xf[0];
    double sr[9];
#pragma unroll
    for (int i = 0; i < 9; ++i) sr[i] = __dmul_rn(s, xf[1 + i]);
    const double tx = xf[10], ty = xf[11], tz = xf[12];
    const double c0x = geo[0], c0y = geo[1], c0z = geo[2];
    double acc[1] = {0.0};
    transform_points(y, p64x, p64y, p64z, phx, phy, phz, np, np_pad, sr, tx, ty, tz, c0x, c0y, c0z,
                     in_place, (int)gridDim.x, acc[0]);
    block_sum<1>(acc, red);
    if (threadIdx.x == 0) {
        partials[blockIdx.x] = acc[0];
        __threadfence();
        const int t = atomicAdd(&ctl[CTL_TICKET_ERR], 1);
        is_last = (t == (int)gridDim.x - 1);
    }
    __syncthreads();
    if (is_last) {
        __threadfence();
        double tot[1] = {0.0};
        for (int b = threadIdx.x; b < (int)gridDim.x; b += RED_THREADS) tot[0] += partials[b];
        block_sum<1>(tot, red);
        if (threadIdx.x == 0) tot_s[0] = tot[0];
        __syncthreads();
        if (threadIdx.x < 32) {
            bool ok = true;
            if (pc.world > 0) ok = peer_allreduce(pc, tot_s, 1, tot_s, stage, ctl);
            if (threadIdx.x == 0 && !ok) ctl[CTL_TICKET_ERR] = 0;
            if (threadIdx.x == 0 && ok) {
                const double total = tot_s[0];
                err_sum[0] = total;
                ctl[CTL_TICKET_ERR] = 0;
                if (finish) {
                    const int it = ctl[CTL_ITERS];
                    const double err = (total + total) / n_global;
                    err_hist[it] = err;
                    ctl[CTL_ITERS] = it + 1;
                    if (err < threshold) ctl[CTL_DONE] = 1;
                }
            }
        }
    }
}

// Multi-GPU: closes the iteration after the residual sums have been all-reduced.
__global__ void finish_kernel(const double* __restrict__ err_sum, double n_global, double threshold,
                              double* __restrict__ err_hist, int* __restrict__ ctl)
{
    if (ctl[CTL_DONE]) return;
    const int it = ctl[CTL_ITERS];
    const double err = (err_sum[0] + err_sum[0]) / n_global;
    err_hist[it] = err;
    ctl[CTL_ITERS] = it + 1;
    if (err < threshold) ctl[CTL_DONE] = 1;
}

// ------------------------------------------------------------------------------------------------
// K2 + K3 + K4 fused (cooperative launch, all blocks co-resident):
//   phase M  17 moment sums of this block's points                      -> partials[block][17]
//   grid barrier
//   phase S  every block adds the partials in the same fixed order (bit-identical totals) and its
//            thread 0 runs the Horn solve into shared memory: no second barrier, no broadcast.
//            peers attached (pc.world > 0): block 0 alone reduces, exchanges the 18 sums with its peers through NVLink
//            mailboxes, publishes the global sums, second grid barrier, then every block solves.
//   phase T  q = sR p + t, residual; in_place writes new_p and its FP32 shadow  -> partials_err[block]
//   last block (ticket): total residual (+ peer exchange), closes the iteration when `finish`.
// 3 CTAs per SM (85 registers; the Horn solve spills a little, once per block): measured against 2 and 4 on 2^22 points,
// K2+K4 159 / 150 / 156 us (profiles/r02c_ab_summary.txt)
#ifndef ICP_FUSED_MIN_CTAS
#define ICP_FUSED_MIN_CTAS 3
#endif
__global__ void __launch_bounds__(RED_THREADS, ICP_FUSED_MIN_CTAS)
align_fused_kernel(const MatchIn y,
                   const double* __restrict__ geo, double* __restrict__ p64x, double* __restrict__ p64y,
                   double* __restrict__ p64z, float* __restrict__ phx, float* __restrict__ phy, float* __restrict__ phz,
                   int np, int np_pad, double* __restrict__ partials,
                   double* __restrict__ mom, double* __restrict__ xf, double* __restrict__ err_sum,
                   int* __restrict__ ctl, int honor_done, int in_place, int finish, double n_global, double threshold,
                   double* __restrict__ err_hist, PeerComm pc, unsigned long long* __restrict__ trace)
{
    if (honor_done && ctl[CTL_DONE]) return;                 // uniform over the grid: nobody reaches a barrier
    // diagnostic (ICP_PROFILE=1): block 0 stamps %globaltimer at its phase boundaries
#define TRACE(slot) do { if (trace && blockIdx.x == 0 && threadIdx.x == 0) trace[slot] = globaltimer_ns(); } while (0)
    // ... and the first 32 blocks stamp (SM id, entry, barrier passed, solve done, exit) at trace[16 + 5 b ...]
#define BTRACE(i) do { if (trace && blockIdx.x < 32 && threadIdx.x == 0) trace[16 + 5 * blockIdx.x + (i)] = globaltimer_ns(); } while (0)
    TRACE(0);
    if (trace && blockIdx.x < 32 && threadIdx.x == 0) {
        unsigned smid;
        asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
        trace[16 + 5 * blockIdx.x] = smid;
    }
    BTRACE(1);
    cg::grid_group grid = cg::this_grid();
    __shared__ double red[17 * (RED_THREADS / 32)];
    __shared__ double mom_s[MOM_N];
    __shared__ double xf_s[XF_N];
    __shared__ double stage[PEER_MAX_WORLD * PEER_SLOT];
    __shared__ int is_last;
    const double c0x = geo[0], c0y = geo[1], c0z = geo[2];
    const int nb = (int)gridDim.x;
    double* partials_err = partials + (size_t)nb * 17;

    // ---- phase M ------------------------------------------------------------------------------
    double acc[17];
#pragma unroll
    for (int i = 0; i < 17; ++i) acc[i] = 0.0;
    moments_points(y, p64x, p64y, p64z, np, nb, c0x, c0y, c0z, acc);
    TRACE(1);
    block_sum<17>(acc, red);
    if (threadIdx.x == 0)
#pragma unroll
        for (int i = 0; i < 17; ++i) partials[(size_t)i * nb + blockIdx.x] = acc[i];   // [17][nb]: coalesced re-reads
    TRACE(2);
    grid.sync();
    TRACE(3);
    BTRACE(2);

    // ---- phase S ------------------------------------------------------------------------------
    if (pc.world == 0 || blockIdx.x == 0) {
        double tot[17];
#pragma unroll
        for (int i = 0; i < 17; ++i) tot[i] = 0.0;
        for (int b = threadIdx.x; b < nb; b += RED_THREADS)
#pragma unroll
            for (int i = 0; i < 17; ++i) tot[i] += __ldcg(&partials[(size_t)i * nb + b]);
        TRACE(10);
        block_sum<17>(tot, red);
        TRACE(11);
        if (threadIdx.x == 0) {
            mom_s[0] = (double)np;
#pragma unroll
            for (int i = 0; i < 17; ++i) mom_s[1 + i] = tot[i];
        }
        __syncthreads();
    }
    if (pc.world > 0) {
        if (blockIdx.x == 0) {
            if (threadIdx.x < 32) {
                const bool ok = peer_allreduce(pc, mom_s, MOM_N, mom_s, stage, ctl);
                if (ok && threadIdx.x < MOM_N) mom[threadIdx.x] = mom_s[threadIdx.x];
                __threadfence();
            }
        }
        grid.sync();
        if (*(volatile int*)&ctl[CTL_COMM_TIMEOUT]) return;      // a peer died: uniform over the grid, new_p untouched
        if (blockIdx.x != 0 && threadIdx.x < MOM_N) mom_s[threadIdx.x] = __ldcg(&mom[threadIdx.x]);
        __syncthreads();
    }
    TRACE(4);
    if (threadIdx.x == 0) {
        horn_solve(geo, mom_s, xf_s);
        if (blockIdx.x == 0) {
            for (int i = 0; i < XF_N; ++i) xf[i] = xf_s[i];
            if (pc.world == 0)
                for (int i = 0; i < MOM_N; ++i) mom[i] = mom_s[i];
        }
    }
    __syncthreads();
    TRACE(5);
    BTRACE(3);

    // ---- phase T ------------------------------------------------------------------------------
    const double s = xf_s[0];
    double sr[9];
#pragma unroll
    for (int i = 0; i < 9; ++i) sr[i] = __dmul_rn(s, xf_s[1 + i]);
    const double tx = xf_s[10], ty = xf_s[11], tz = xf_s[12];
    double e[1] = {0.0};
    transform_points(y, p64x, p64y, p64z, phx, phy, phz, np, np_pad, sr, tx, ty, tz, c0x, c0y, c0z,
                     in_place, nb, e[0]);
    TRACE(6);
    block_sum<1>(e, red);
    if (threadIdx.x == 0) {
        partials_err[blockIdx.x] = e[0];
        __threadfence();
        const int t = atomicAdd(&ctl[CTL_TICKET_ERR], 1);
        is_last = (t == nb - 1);
    }
    __syncthreads();
    if (is_last) {
        __threadfence();
        double tot[1] = {0.0};
        for (int b = threadIdx.x; b < nb; b += RED_THREADS) tot[0] += __ldcg(&partials_err[b]);
        block_sum<1>(tot, red);
        if (threadIdx.x == 0) mom_s[0] = tot[0];
        __syncthreads();
        if (threadIdx.x < 32) {
            bool ok = true;
            if (pc.world > 0) ok = peer_allreduce(pc, mom_s, 1, mom_s, stage, ctl);
            if (threadIdx.x == 0 && !ok) ctl[CTL_TICKET_ERR] = 0;
            if (threadIdx.x == 0 && ok) {
                const double total = mom_s[0];
                err_sum[0] = total;
                ctl[CTL_TICKET_ERR] = 0;
                if (finish) {   // src/cpu.cc:67-77: the two residual sums are the same expression
                    const int it = ctl[CTL_ITERS];
                    const double err = (total + total) / n_global;
                    err_hist[it] = err;
                    ctl[CTL_ITERS] = it + 1;
                    if (err < threshold) ctl[CTL_DONE] = 1;
                }
                if (trace) trace[9] = globaltimer_ns();      // the LAST block is done
            }
        }
    }
    TRACE(7);
    BTRACE(4);
#undef BTRACE
#undef TRACE
}

// ------------------------------------------------------------------------------------------------
// Preparation kernels (run once per icp_set_model / icp_set_scene, not per iteration).

// Bounding box of an AoS FP64 cloud: out = min[3], max[3].  Block partials -> the last block (ticket) folds them.
constexpr int BBOX_THREADS = 512;
__global__ void __launch_bounds__(BBOX_THREADS)
bbox_kernel(const double* __restrict__ xyz, int64_t n, double* __restrict__ out, double* __restrict__ partials,
            int* __restrict__ ticket)
{
    __shared__ double smin[3][BBOX_THREADS / 32], smax[3][BBOX_THREADS / 32];
    __shared__ int is_last;
    double mn[3] = {1e300, 1e300, 1e300}, mx[3] = {-1e300, -1e300, -1e300};
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    auto block_minmax = [&]() {     // result valid in thread 0
#pragma unroll
        for (int c = 0; c < 3; ++c) {
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                mn[c] = fmin(mn[c], __shfl_xor_sync(0xffffffffu, mn[c], o));
                mx[c] = fmax(mx[c], __shfl_xor_sync(0xffffffffu, mx[c], o));
            }
            if (lane == 0) { smin[c][w] = mn[c]; smax[c][w] = mx[c]; }
        }
        __syncthreads();
        if (w == 0) {
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                double a = lane < BBOX_THREADS / 32 ? smin[c][lane] : 1e300;
                double b = lane < BBOX_THREADS / 32 ? smax[c][lane] : -1e300;
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) {
                    a = fmin(a, __shfl_xor_sync(0xffffffffu, a, o));
                    b = fmax(b, __shfl_xor_sync(0xffffffffu, b, o));
                }
                mn[c] = a; mx[c] = b;
            }
        }
        __syncthreads();
    };
    for (int64_t i = (int64_t)blockIdx.x * BBOX_THREADS + threadIdx.x; i < n; i += (int64_t)gridDim.x * BBOX_THREADS)
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            const double v = xyz[3 * i + c];
            mn[c] = fmin(mn[c], v);
            mx[c] = fmax(mx[c], v);
        }
    block_minmax();
    if (threadIdx.x == 0) {
#pragma unroll
        for (int c = 0; c < 3; ++c) { partials[blockIdx.x * 6 + c] = mn[c]; partials[blockIdx.x * 6 + 3 + c] = mx[c]; }
        __threadfence();
        is_last = atomicAdd(ticket, 1) == (int)gridDim.x - 1;
    }
    __syncthreads();
    if (is_last) {
        __threadfence();
#pragma unroll
        for (int c = 0; c < 3; ++c) { mn[c] = 1e300; mx[c] = -1e300; }
        for (int b = threadIdx.x; b < (int)gridDim.x; b += BBOX_THREADS)
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                mn[c] = fmin(mn[c], __ldcg(&partials[b * 6 + c]));
                mx[c] = fmax(mx[c], __ldcg(&partials[b * 6 + 3 + c]));
            }
        block_minmax();
        if (threadIdx.x == 0) {
#pragma unroll
            for (int c = 0; c < 3; ++c) { out[c] = mn[c]; out[3 + c] = mx[c]; }
            *ticket = 0;
        }
    }
}

// geo[4..9] = bbox  ->  geo[0..2] = c0 (bbox centre), geo[3] = Rm >= max |m - c0|
__global__ void model_geo_kernel(double* geo)
{
    double h2 = 0.0;
    for (int c = 0; c < 3; ++c) {
        const double lo = geo[4 + c], hi = geo[7 + c];
        geo[c] = 0.5 * (lo + hi);
        const double h = fmax(hi - geo[c], geo[c] - lo);
        h2 += h * h;
    }
    geo[3] = sqrt(h2) * (1.0 + 1e-9);
}

// AoS FP64 model -> FP64 master (one 32-byte record per point) + FP32 float4 shadow (c0-relative), sentinel padding.
__global__ void model_build_kernel(const double* __restrict__ xyz, int64_t nm, int64_t nm_pad,
                                   const double* __restrict__ geo, Pt64* __restrict__ m64, float4* __restrict__ mhat)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nm_pad) return;
    if (i < nm) {
        const double x = xyz[3 * i], y = xyz[3 * i + 1], z = xyz[3 * i + 2];
        m64[i] = Pt64{x, y, z, 0.0};
        mhat[i] = make_float4((float)(x - geo[0]), (float)(y - geo[1]), (float)(z - geo[2]), 0.f);
    } else {
        m64[i] = Pt64{0.0, 0.0, 0.0, 0.0};
        mhat[i] = make_float4(NN_SENTINEL, NN_SENTINEL, NN_SENTINEL, 0.f);
    }
}

__device__ __forceinline__ uint32_t spread10(uint32_t v)
{
    v &= 0x3ffu;
    v = (v | (v << 16)) & 0x030000ffu;
    v = (v | (v << 8)) & 0x0300f00fu;
    v = (v | (v << 4)) & 0x030c30c3u;
    v = (v | (v << 2)) & 0x09249249u;
    return v;
}

// 30-bit Morton keys over the scene's own bounding box.
__global__ void scene_keys_kernel(const double* __restrict__ xyz, int64_t n, const double* __restrict__ bbox,
                                  uint32_t* __restrict__ keys, int32_t* __restrict__ vals)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t q[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        const double lo = bbox[c], ext = bbox[3 + c] - lo;
        double f = ext > 0.0 ? (xyz[3 * i + c] - lo) / ext : 0.0;
        f = fmin(fmax(f, 0.0), 1.0);
        q[c] = (uint32_t)fmin(f * 1024.0, 1023.0);
    }
    keys[i] = (spread10(q[0]) << 2) | (spread10(q[1]) << 1) | spread10(q[2]);
    vals[i] = (int32_t)i;
}

// Gather the scene into sorted SoA order: FP64 master, backup, FP32 shadow (+ padding = last point).
__global__ void scene_build_kernel(const double* __restrict__ xyz, int64_t np, int64_t np_pad,
                                   const int32_t* __restrict__ perm, const double* __restrict__ geo,
                                   double* __restrict__ p64x, double* __restrict__ p64y, double* __restrict__ p64z,
                                   double* __restrict__ q64x, double* __restrict__ q64y, double* __restrict__ q64z,
                                   float* __restrict__ phx, float* __restrict__ phy, float* __restrict__ phz)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= np_pad) return;
    const int64_t src = perm[i < np ? i : np - 1];
    const double x = xyz[3 * src], y = xyz[3 * src + 1], z = xyz[3 * src + 2];
    if (i < np) {
        p64x[i] = x; p64y[i] = y; p64z[i] = z;
        q64x[i] = x; q64y[i] = y; q64z[i] = z;
    }
    phx[i] = (float)(x - geo[0]); phy[i] = (float)(y - geo[1]); phz[i] = (float)(z - geo[2]);
}

__global__ void scene_restore_kernel(int64_t np, int64_t np_pad, const double* __restrict__ geo,
                                     const double* __restrict__ q64x, const double* __restrict__ q64y,
                                     const double* __restrict__ q64z, double* __restrict__ p64x,
                                     double* __restrict__ p64y, double* __restrict__ p64z, float* __restrict__ phx,
                                     float* __restrict__ phy, float* __restrict__ phz)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= np_pad) return;
    const int64_t src = i < np ? i : np - 1;
    const double x = q64x[src], y = q64y[src], z = q64z[src];
    if (i < np) { p64x[i] = x; p64y[i] = y; p64z[i] = z; }
    phx[i] = (float)(x - geo[0]); phy[i] = (float)(y - geo[1]); phz[i] = (float)(z - geo[2]);
}

__global__ void scene_export_kernel(int64_t np, const int32_t* __restrict__ perm, const double* __restrict__ p64x,
                                    const double* __restrict__ p64y, const double* __restrict__ p64z,
                                    double* __restrict__ out)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= np) return;
    const int64_t d = perm[i];
    out[3 * d] = p64x[i]; out[3 * d + 1] = p64y[i]; out[3 * d + 2] = p64z[i];
}

__global__ void indices_export_kernel(int64_t np, const int32_t* __restrict__ perm, const int32_t* __restrict__ idx,
                                      const Pt64* __restrict__ m64, int nm, int32_t* __restrict__ idx_out,
                                      double* __restrict__ Y_out)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= np) return;
    const int64_t d = perm[i];
    const int k = idx[i];
    if (idx_out) idx_out[d] = k;
    if (Y_out && (unsigned)k < (unsigned)nm) {
        double mx, my, mz;
        ld_pt64(m64, k, mx, my, mz);
        Y_out[3 * d] = mx; Y_out[3 * d + 1] = my; Y_out[3 * d + 2] = mz;
    }
}

// Y = m64[idx] for every scene point: only where idx was set from outside the search kernels
__global__ void gather_matches_kernel(int64_t np, int nm, const int32_t* __restrict__ idx, const Pt64* __restrict__ m64,
                                      double* __restrict__ yx, double* __restrict__ yy, double* __restrict__ yz)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= np) return;
    int k = idx[i];
    if ((unsigned)k >= (unsigned)nm) k = 0;
    double x, y, z;
    ld_pt64(m64, k, x, y, z);
    yx[i] = x; yy[i] = y; yz[i] = z;
}

// mu_p, mu_y from the moments (src/cpu.cc:113-114)
__global__ void centroid_export_kernel(const double* __restrict__ geo, const double* __restrict__ mom,
                                       double* __restrict__ out6)
{
    const double n = mom[0];
    for (int c = 0; c < 3; ++c) {
        out6[c] = mom[1 + c] / n + geo[c];
        out6[3 + c] = mom[4 + c] / n + geo[c];
    }
}

// substract_col (src/GPU/compute.cu:381-398): out[:,i] = M[:,i] - m
__global__ void subtract_col_kernel(const double* __restrict__ M, int64_t n, double mx, double my, double mz,
                                    double* __restrict__ out)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    out[3 * i] = M[3 * i] - mx; out[3 * i + 1] = M[3 * i + 1] - my; out[3 * i + 2] = M[3 * i + 2] - mz;
}

// ------------------------------------------------------------------------------------------------
// No-allocation variant of K2 + K3 (icp_find_alignment_noalloc, see the end of this file).  No counterpart in the
// reference; added for callers that cannot call cudaMalloc.
// Same maths as moments_kernel + horn_solve on PAIRED clouds given as AoS FP64 device arrays (p_j <-> y_j); the block
// partials and the ticket live in a file-scope device array instead of a context's cudaMalloc'ed scratch.
constexpr int NOALLOC_MAX_BLOCKS = 592;                       // 4 CTAs x 148 SMs
constexpr long long NOALLOC_MAX_POINTS = 1ll << 22;           // stated maximum of the entry point: 4 194 304 pairs
constexpr int NOALLOC_SCRATCH_DOUBLES = NOALLOC_MAX_BLOCKS * 17 + 32;      // partials [blocks][17] + a 256-byte tail (ticket)
// 80 768 bytes, held by every GPU that loads the library until the process exits.  In HOST code this name is a host-side
// stand-in, not a device address: the address comes from cudaGetSymbolAddress (noalloc_scratch_ptr below).
static __device__ __align__(256) double g_noalloc_scratch[NOALLOC_SCRATCH_DOUBLES];

__global__ void __launch_bounds__(RED_THREADS, 4)
moments_pairs_noalloc_kernel(const double* __restrict__ p_aos, const double* __restrict__ y_aos, int n,
                             double* __restrict__ scratch, double* __restrict__ out_xf)
{
    __shared__ double red[17 * (RED_THREADS / 32)];
    __shared__ double mom_s[MOM_N];
    __shared__ double geo_s[4];
    __shared__ int is_last;
    double* partials = scratch;
    int* ticket = reinterpret_cast<int*>(scratch + NOALLOC_MAX_BLOCKS * 17);
    // pivot of the sums: the first y point (any point near the clouds keeps the algebraic centring harmless)
    const double c0x = y_aos[0], c0y = y_aos[1], c0z = y_aos[2];
    double acc[17];
#pragma unroll
    for (int i = 0; i < 17; ++i) acc[i] = 0.0;
    for (int j = blockIdx.x * RED_THREADS + threadIdx.x; j < n; j += gridDim.x * RED_THREADS)
        moments_accumulate(acc, p_aos[3 * (size_t)j] - c0x, p_aos[3 * (size_t)j + 1] - c0y, p_aos[3 * (size_t)j + 2] - c0z,
                           y_aos[3 * (size_t)j] - c0x, y_aos[3 * (size_t)j + 1] - c0y, y_aos[3 * (size_t)j + 2] - c0z);
    block_sum<17>(acc, red);
    if (threadIdx.x == 0) {
#pragma unroll
        for (int i = 0; i < 17; ++i) partials[(size_t)blockIdx.x * 17 + i] = acc[i];
        __threadfence();
        is_last = atomicAdd(ticket, 1) == (int)gridDim.x - 1;
    }
    __syncthreads();
    if (is_last) {
        __threadfence();
        double tot[17];
#pragma unroll
        for (int i = 0; i < 17; ++i) tot[i] = 0.0;
        for (int b = threadIdx.x; b < (int)gridDim.x; b += RED_THREADS)
#pragma unroll
            for (int i = 0; i < 17; ++i) tot[i] += __ldcg(&partials[(size_t)b * 17 + i]);
        block_sum<17>(tot, red);
        if (threadIdx.x == 0) {
            mom_s[0] = (double)n;
#pragma unroll
            for (int i = 0; i < 17; ++i) mom_s[1 + i] = tot[i];
            geo_s[0] = c0x; geo_s[1] = c0y; geo_s[2] = c0z; geo_s[3] = 0.0;
            horn_solve(geo_s, mom_s, out_xf);          // s, R[9], t[3], mu_p[3], mu_y[3] (src/cpu.cc:113-167)
            out_xf[19] = 0.0;
            *ticket = 0;                               // the next call starts from zero (it clears the ticket again anyway)
        }
    }
}

}  // namespace

// ------------------------------------------------------------------------------------------------
// launchers
static inline int grid1d(int64_t n, int threads) { return (int)((n + threads - 1) / threads); }

int reduce_blocks_for(int64_t n, int sm_count)
{
    int64_t b = (n + RED_THREADS - 1) / RED_THREADS;
    const int64_t cap = (int64_t)sm_count * 4;
    if (b > cap) b = cap;
    if (b < 1) b = 1;
    return (int)b;
}

void launch_moments(const ModelDev& m, const SceneDev& s, double* partials, int n_blocks, double* mom, int* ctl,
                    int honor_done, double* xf_fused, cudaStream_t st, const PeerComm* peer)
{
    PeerComm pc;
    if (peer) pc = *peer;
    moments_kernel<<<n_blocks, RED_THREADS, 0, st>>>(MatchIn{s.y64x, s.y64y, s.y64z}, m.geo, s.p64x, s.p64y, s.p64z,
                                                     (int)s.np, partials, mom, ctl, honor_done, xf_fused, pc);
}

void launch_solve(const ModelDev& m, const double* mom, double* xf, const int* ctl, int honor_done, cudaStream_t st)
{
    solve_kernel<<<1, 32, 0, st>>>(m.geo, mom, xf, ctl, honor_done);
}

void launch_transform(const ModelDev& m, const SceneDev& s, const double* xf, int in_place, double* partials,
                      int n_blocks, double* err_sum, int* ctl, int honor_done, int finish, double n_global,
                      double threshold, double* err_hist, cudaStream_t st, const PeerComm* peer)
{
    PeerComm pc;
    if (peer) pc = *peer;
    transform_kernel<<<n_blocks, RED_THREADS, 0, st>>>(MatchIn{s.y64x, s.y64y, s.y64z}, m.geo, s.p64x, s.p64y, s.p64z, s.phx,
                                                       s.phy, s.phz, (int)s.np, (int)s.np_pad, xf, in_place,
                                                       partials, err_sum, ctl, honor_done, finish, n_global,
                                                       threshold, err_hist, pc);
}

int align_fused_max_blocks(int sm_count)
{
    int per_sm = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, align_fused_kernel, RED_THREADS, 0) != cudaSuccess || per_sm < 1)
        per_sm = 1;
    if (per_sm > 4) per_sm = 4;
    return per_sm * sm_count;
}

int launch_align_fused(const ModelDev& m, const SceneDev& s, double* partials, int n_blocks, double* mom, double* xf,
                       double* err_sum, int* ctl, int honor_done, int in_place, int finish, double n_global,
                       double threshold, double* err_hist, const PeerComm* peer, unsigned long long* trace,
                       cudaStream_t st)
{
    PeerComm pc;
    if (peer) pc = *peer;
    int np = (int)s.np, np_pad = (int)s.np_pad;
    MatchIn a0{s.y64x, s.y64y, s.y64z};
    const double* a3 = m.geo;
    double *b0 = s.p64x, *b1 = s.p64y, *b2 = s.p64z;
    float *f0 = s.phx, *f1 = s.phy, *f2 = s.phz;
    void* args[] = {&a0, &a3, &b0, &b1, &b2, &f0, &f1, &f2, &np, &np_pad, &partials, &mom, &xf, &err_sum,
                    &ctl, &honor_done, &in_place, &finish, &n_global, &threshold, &err_hist, &pc, &trace};
    return (int)cudaLaunchCooperativeKernel((const void*)align_fused_kernel, dim3(n_blocks), dim3(RED_THREADS), args, 0, st);
}

void launch_finish(const double* err_sum, double n_global, double threshold, double* err_hist, int* ctl,
                   cudaStream_t st)
{
    finish_kernel<<<1, 1, 0, st>>>(err_sum, n_global, threshold, err_hist, ctl);
}

void launch_bbox(const double* xyz_aos, int64_t n, double* out_min3_max3, double* partials, int max_blocks, int* ctl,
                 cudaStream_t st)
{
    int64_t blocks = (n + BBOX_THREADS * 4 - 1) / (BBOX_THREADS * 4);
    if (blocks > max_blocks) blocks = max_blocks;
    if (blocks < 1) blocks = 1;
    bbox_kernel<<<(int)blocks, BBOX_THREADS, 0, st>>>(xyz_aos, n, out_min3_max3, partials, ctl + CTL_TICKET_BBOX);
}

void launch_model_geo(double* geo, cudaStream_t st) { model_geo_kernel<<<1, 1, 0, st>>>(geo); }

void launch_model_build(const double* xyz_aos, const ModelDev& m, cudaStream_t st)
{
    model_build_kernel<<<grid1d(m.nm_pad, 256), 256, 0, st>>>(xyz_aos, m.nm, m.nm_pad, m.geo, m.m64, m.mhat);
}

void launch_scene_keys(const double* xyz_aos, int64_t n, const double* bbox, uint32_t* keys, int32_t* vals,
                       cudaStream_t st)
{
    scene_keys_kernel<<<grid1d(n, 256), 256, 0, st>>>(xyz_aos, n, bbox, keys, vals);
}

void launch_scene_build(const double* xyz_aos, const ModelDev& m, const SceneDev& s, cudaStream_t st)
{
    scene_build_kernel<<<grid1d(s.np_pad, 256), 256, 0, st>>>(xyz_aos, s.np, s.np_pad, s.perm, m.geo, s.p64x, s.p64y,
                                                              s.p64z, s.q64x, s.q64y, s.q64z, s.phx, s.phy, s.phz);
}

void launch_scene_restore(const ModelDev& m, const SceneDev& s, cudaStream_t st)
{
    scene_restore_kernel<<<grid1d(s.np_pad, 256), 256, 0, st>>>(s.np, s.np_pad, m.geo, s.q64x, s.q64y, s.q64z, s.p64x,
                                                                s.p64y, s.p64z, s.phx, s.phy, s.phz);
}

void launch_scene_export(const SceneDev& s, double* xyz_aos_out, cudaStream_t st)
{
    scene_export_kernel<<<grid1d(s.np, 256), 256, 0, st>>>(s.np, s.perm, s.p64x, s.p64y, s.p64z, xyz_aos_out);
}

void launch_indices_export(const SceneDev& s, const ModelDev& m, int32_t* idx_out, double* Y_aos_out, cudaStream_t st)
{
    indices_export_kernel<<<grid1d(s.np, 256), 256, 0, st>>>(s.np, s.perm, s.idx, m.m64, (int)m.nm, idx_out, Y_aos_out);
}

void launch_gather_matches(const SceneDev& s, const ModelDev& m, cudaStream_t st)
{
    gather_matches_kernel<<<grid1d(s.np, 256), 256, 0, st>>>(s.np, (int)m.nm, s.idx, m.m64, s.y64x, s.y64y, s.y64z);
}

void launch_centroid_export(const ModelDev& m, const double* mom, double* out6, cudaStream_t st)
{
    centroid_export_kernel<<<1, 1, 0, st>>>(m.geo, mom, out6);
}

void launch_subtract_col(const double* M_aos, int64_t n, double mx, double my, double mz, double* out_aos,
                         cudaStream_t st)
{
    subtract_col_kernel<<<grid1d(n, 256), 256, 0, st>>>(M_aos, n, mx, my, mz, out_aos);
}

}  // namespace icpb

// ------------------------------------------------------------------------------------------------
// icp_find_alignment_noalloc (include/icp_b200.h): K2 + K3 on paired device clouds without any allocation.
#include <mutex>
#include <string>

#include "../../include/icp_b200.h"

namespace icpb {
void set_last_error(const std::string& msg);      // icp_api.cu: what icp_last_error() returns on the calling thread
}

namespace {
// Device address of g_noalloc_scratch on the CURRENT device; resolved once per device, with that device current.
int noalloc_scratch_ptr(double** out)
{
    constexpr int MAX_DEV = 64;
    static std::mutex mu;
    static double* ptr[MAX_DEV] = {};
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= MAX_DEV) {
        icpb::set_last_error(std::string("icp_find_alignment_noalloc: no current CUDA device (") + cudaGetErrorString(e) + ")");
        cudaGetLastError();
        return ICP_ERR_NO_DEVICE;
    }
    std::lock_guard<std::mutex> lk(mu);
    if (!ptr[dev]) {
        double* p = nullptr;
        e = cudaGetSymbolAddress((void**)&p, icpb::g_noalloc_scratch);
        if (e != cudaSuccess || !p) {
            icpb::set_last_error(std::string("icp_find_alignment_noalloc: cudaGetSymbolAddress failed: ") + cudaGetErrorString(e));
            cudaGetLastError();
            return ICP_ERR_CUDA;
        }
        ptr[dev] = p;
    }
    *out = ptr[dev];
    return ICP_OK;
}
}  // namespace

extern "C" int icp_find_alignment_noalloc(const double* d_y_xyz, const double* d_p_xyz, int64_t n, double* d_out20,
                                          void* cuda_stream)
{
    if (!d_y_xyz || !d_p_xyz || !d_out20 || n < 1) {
        icpb::set_last_error("icp_find_alignment_noalloc: bad arguments (device pointers y, p, out and n >= 1 required)");
        return ICP_ERR_INVALID;
    }
    if (n > icpb::NOALLOC_MAX_POINTS) {
        icpb::set_last_error("icp_find_alignment_noalloc: n = " + std::to_string((long long)n) + " exceeds the maximum of " +
                             std::to_string(icpb::NOALLOC_MAX_POINTS) + " point pairs of the no-allocation entry point (use icp_find_alignment)");
        return ICP_ERR_INVALID;
    }
    double* scratch = nullptr;
    const int r = noalloc_scratch_ptr(&scratch);
    if (r) return r;
    cudaStream_t st = (cudaStream_t)cuda_stream;
    // the ticket must start at zero whatever an earlier (possibly aborted) call left in the shared array
    cudaError_t e = cudaMemsetAsync(scratch + icpb::NOALLOC_MAX_BLOCKS * 17, 0, 256, st);
    if (e == cudaSuccess) {
        int64_t blocks = (n + icpb::RED_THREADS - 1) / icpb::RED_THREADS;
        if (blocks > icpb::NOALLOC_MAX_BLOCKS) blocks = icpb::NOALLOC_MAX_BLOCKS;
        icpb::moments_pairs_noalloc_kernel<<<(int)blocks, icpb::RED_THREADS, 0, st>>>(d_p_xyz, d_y_xyz, (int)n, scratch, d_out20);
        e = cudaGetLastError();
    }
    if (e != cudaSuccess) {
        icpb::set_last_error(std::string("icp_find_alignment_noalloc: ") + cudaGetErrorString(e));
        return ICP_ERR_CUDA;
    }
    return ICP_OK;
}
