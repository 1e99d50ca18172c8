// tree_kernels.cu — EXACT nearest neighbour through a bounding-box hierarchy over the Morton-sorted model
// ("exact spatial acceleration behind the same icp_closest").  Opt-in (icp_set_nn_mode /
// ICP_NN=tree); the brute-force filter of nn_kernels.cu stays the default and the headline of bench.py.
//
// Same contract as the brute-force path: idx[j] = the FIRST minimum over k of the reference's FP64 value
// v_k = sqrt(pow(dx,2)+pow(dy,2)+pow(dz,2)) (src/cpu.cc:16-22), bit-exact.  How exactness is kept:
//   * the tree is searched with an FP32 surrogate d~2_k = fl32(|p^ - m^_k|^2) of the centred FP32 shadows
//     (p^ = fl32(p - c0), m^ = fl32(m - c0));
//   * with u = 2^-24:  sqrt(d~2) = D^(1 +- 3u),  |D - D^| <= E = u(1+3u)(|p^| + Rm)   (D = true distance), and the
//     reference's v_k = D_k(1 +- 4*2^-53).  If k0 minimises the surrogate and k* is ANY index attaining the
//     reference's minimum, then  D^_k* <= D^_k0 + 2E(1+eps)  and therefore
//         d~2_k*  <=  thr2 := ( sqrt(best2)(1+3u) + 2E )^2 (1+16u)         (best2 = d~2_k0)
//     thr2 is evaluated with every operation rounded UP and is monotone in best2, so the running thr2 of a
//     search in progress is always >= the final one;
//   * every model point met with d~2 <= thr2 is evaluated in FP64 with the reference's own expression and
//     merged by (v, lowest index) — order independent, hence the traversal order is free;
//   * a node is skipped only when a LOWER bound of its box distance (all operations rounded DOWN) exceeds
//     thr2, which (1+16u) keeps above d~2 of every point inside the box.
// The ORDER of the model points (Morton sort, then median splits: wtree_local_kd_kernel / kd_global_levels) only decides
// how many boxes a search opens; exactness never depends on it.
// Model scales outside [1e-12, 1e12] are refused at build time (FP32 squares would leave the normal range);
// icp_api.cu then keeps the brute-force path, whose bound has no such restriction.
#include <cub/device/device_radix_sort.cuh>

#include <cstdlib>

#include "icp_common.cuh"

namespace icpb {
namespace {

constexpr int TREE_THREADS = 128;
constexpr int TREE_STACK = 40;

__device__ __forceinline__ double ref_dist2(double px, double py, double pz, double mx, double my, double mz)
{
    const double dx = __dsub_rn(px, mx), dy = __dsub_rn(py, my), dz = __dsub_rn(pz, mz);
    return __dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(dz, dz));
}
__device__ __forceinline__ double ref_dist2_pt(double px, double py, double pz, const Pt64* __restrict__ m64, int k)
{
    double mx, my, mz;
    ld_pt64(m64, k, mx, my, mz);
    return ref_dist2(px, py, pz, mx, my, mz);
}

__device__ __forceinline__ uint32_t tree_spread10(uint32_t v)
{
    v &= 0x3ffu;
    v = (v | (v << 16)) & 0x030000ffu;
    v = (v | (v << 8)) & 0x0300f00fu;
    v = (v | (v << 4)) & 0x030c30c3u;
    v = (v | (v << 2)) & 0x09249249u;
    return v;
}

// 30-bit Morton keys of the centred FP32 model over its bounding box (geo[4..9] = bbox, geo[0..2] = c0)
__global__ void tree_keys_kernel(const float4* __restrict__ mhat, int nm, const double* __restrict__ geo,
                                 uint32_t* __restrict__ keys, int32_t* __restrict__ vals)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nm) return;
    const float4 m = mhat[i];
    const float v[3] = {m.x, m.y, m.z};
    uint32_t q[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        const float lo = (float)(geo[4 + c] - geo[c]), ext = (float)(geo[7 + c] - geo[4 + c]);
        float f = ext > 0.0f ? (v[c] - lo) / ext : 0.0f;
        f = fminf(fmaxf(f, 0.0f), 1.0f);
        q[c] = (uint32_t)fminf(f * 1024.0f, 1023.0f);
    }
    keys[i] = (tree_spread10(q[0]) << 2) | (tree_spread10(q[1]) << 1) | tree_spread10(q[2]);
    vals[i] = i;
}

// sorted leaf points: (m^.x, m^.y, m^.z, original index); positions >= nm repeat the last valid point
__global__ void tree_points_kernel(const float4* __restrict__ mhat, const int32_t* __restrict__ perm, int nm,
                                   int n_slots, float4* __restrict__ pts)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_slots) return;
    const int k = perm[i < nm ? i : nm - 1];
    const float4 m = mhat[k];
    pts[i] = make_float4(m.x, m.y, m.z, __int_as_float(k));
}

// node = 6 floats (lo.xyz, hi.xyz); heap numbering: root 1, children 2i / 2i+1, leaves NL .. 2NL-1
// Leaves past the last model point get the EMPTY box (lo = +inf, hi = -inf): its distance is +inf, above every
// finite threshold, and it is neutral under the min/max of tree_level_kernel.
__global__ void tree_leaf_boxes_kernel(const float4* __restrict__ pts, int n_leaves, int n_real, float* __restrict__ nodes)
{
    const int l = blockIdx.x * blockDim.x + threadIdx.x;
    if (l >= n_leaves) return;
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
#pragma unroll
    for (int i = 0; i < TREE_LEAF && l < n_real; ++i) {
        const float4 q = pts[(size_t)l * TREE_LEAF + i];
        lo[0] = fminf(lo[0], q.x); hi[0] = fmaxf(hi[0], q.x);
        lo[1] = fminf(lo[1], q.y); hi[1] = fmaxf(hi[1], q.y);
        lo[2] = fminf(lo[2], q.z); hi[2] = fmaxf(hi[2], q.z);
    }
    float* o = nodes + (size_t)(n_leaves + l) * 6;
    o[0] = lo[0]; o[1] = lo[1]; o[2] = lo[2]; o[3] = hi[0]; o[4] = hi[1]; o[5] = hi[2];
}

__global__ void tree_level_kernel(float* __restrict__ nodes, int first, int count)
{
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= count) return;
    const size_t i = (size_t)first + t;
    const float* a = nodes + 2 * i * 6;
    const float* b = a + 6;
    float* o = nodes + i * 6;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        o[c] = fminf(a[c], b[c]);
        o[3 + c] = fmaxf(a[3 + c], b[3 + c]);
    }
}

// lower bound of the squared distance from p to the box, every operation rounded down
__device__ __forceinline__ float box_dist2_rd(float px, float py, float pz, float lx, float ly, float lz, float hx,
                                              float hy, float hz)
{
    const float ex = fmaxf(fmaxf(__fsub_rd(lx, px), __fsub_rd(px, hx)), 0.0f);
    const float ey = fmaxf(fmaxf(__fsub_rd(ly, py), __fsub_rd(py, hy)), 0.0f);
    const float ez = fmaxf(fmaxf(__fsub_rd(lz, pz), __fsub_rd(pz, hz)), 0.0f);
    return __fmaf_rd(ez, ez, __fmaf_rd(ey, ey, __fmul_rd(ex, ex)));
}

struct TreeQuery {
    float px, py, pz;        // p^
    double qx, qy, qz;       // p (FP64 master)
    float twoE;              // 2E, rounded up
    float best2, thr2;       // running surrogate minimum and acceptance threshold
    double v;                // running reference minimum (v, k), lexicographic
    int k;
    const Pt64* m64;

    __device__ __forceinline__ void set_best(float b2)
    {
        best2 = b2;
        const float r = __fmul_ru(__fsqrt_ru(b2), 1.0000002f);            // (1 + 3u) < 1.0000002
        const float a = __fadd_ru(r, twoE);
        thr2 = __fmul_ru(__fmul_ru(a, a), 1.000001f);                     // (1 + 16u) < 1.000001
    }
    __device__ __forceinline__ void offer(float d2, int kk)
    {
        if (d2 <= thr2) {
            if (d2 < best2) set_best(d2);
            const double vv = __dsqrt_rn(ref_dist2_pt(qx, qy, qz, m64, kk));
            if (vv < v || (vv == v && kk < k)) { v = vv; k = kk; }
        }
    }
};

__device__ __forceinline__ float surrogate_direct(float px, float py, float pz, float mx, float my, float mz)
{
    const float dx = __fsub_rn(px, mx), dy = __fsub_rn(py, my), dz = __fsub_rn(pz, mz);
    return __fmaf_rn(dz, dz, __fmaf_rn(dy, dy, __fmul_rn(dx, dx)));
}

// One thread per scene point, near child first, far child on a small local stack; a descend-loop nested in a
// pop-loop (in a complete binary tree all lanes of a warp reach their leaves together).  This plain form is the
// fastest of those measured A/B on one box (profiles/r01h_tree.md): a flat one-node-per-trip loop +15 %; per-warp
// pools of scene points with idle lanes refilled on the fly +10 % to 6x (ncu shows 10.3 of 32 lanes active, but the
// kernel is bound by the latency of dependent loads and fewer, longer-lived warps hide less of it); the leaf
// scanned branch-free with a compact candidate loop and a square-root-free candidate merge +4..11 % (48 instead of
// 40 registers: 40 instead of 48 resident warps per SM).
__global__ void __launch_bounds__(TREE_THREADS)
nn_tree_kernel(const float* __restrict__ nodes, const float4* __restrict__ pts, int n_leaves /* power of two */,
               const float4* __restrict__ mhat, const Pt64* __restrict__ m64, const double* __restrict__ geo, const double* __restrict__ p64x,
               const double* __restrict__ p64y, const double* __restrict__ p64z, const float* __restrict__ phx,
               const float* __restrict__ phy, const float* __restrict__ phz, int np, int nm,
               MatchOut mo, int use_prev, int* __restrict__ ctl, int honor_done, int* __restrict__ progress)
{
    publish_progress(progress, ctl);
    if (honor_done && ctl[CTL_DONE]) return;
    const int j = blockIdx.x * TREE_THREADS + threadIdx.x;
    if (j >= np) return;

    TreeQuery q;
    q.px = phx[j]; q.py = phy[j]; q.pz = phz[j];
    q.qx = p64x[j]; q.qy = p64y[j]; q.qz = p64z[j];
    q.m64 = m64;
    if (!scene_point_finite(q.qx, q.qy, q.qz)) { store_match(mo, m64, j, 0); return; }
    {
        const float Rm = __fmul_ru(__double2float_ru(geo[3]), 1.000001f);
        const float H = __fsqrt_ru(__fmaf_ru(q.pz, q.pz, __fmaf_ru(q.py, q.py, __fmul_ru(q.px, q.px))));
        // 2E = 2 u (1+3u) (|p^| + Rm), plus an absolute floor against underflow
        q.twoE = __fadd_ru(__fmul_ru(1.1920932e-7f /* > 2^-23 (1+3u) */, __fadd_ru(H, Rm)), 1.0e-30f);
    }
    q.best2 = q.thr2 = 3.4028235e38f;     // FLT_MAX, not +inf: the empty boxes of padded leaves (+inf) stay above it
    q.v = __longlong_as_double(0x7ff0000000000000LL);
    q.k = 0x7fffffff;
    if (use_prev) {      // last iteration's match: a near-optimal bound before the first node is opened
        const int kp = mo.idx[j];
        if (kp >= 0 && kp < nm) {
            const float4 m = mhat[kp];
            q.offer(surrogate_direct(q.px, q.py, q.pz, m.x, m.y, m.z), kp);
        }
    }

    int st_node[TREE_STACK];
    float st_d[TREE_STACK];
    int sp = 0;
    int node = 1;
    float nd = 0.0f;
    const float4* nodes4 = reinterpret_cast<const float4*>(nodes);
    for (;;) {
        if (nd <= q.thr2) {
            while (node < n_leaves) {                                 // internal: open both children, near one first
                const float4 a = nodes4[(size_t)node * 3], b = nodes4[(size_t)node * 3 + 1],
                             c = nodes4[(size_t)node * 3 + 2];        // children 2*node, 2*node+1 = 12 floats
                const float d0 = box_dist2_rd(q.px, q.py, q.pz, a.x, a.y, a.z, a.w, b.x, b.y);
                const float d1 = box_dist2_rd(q.px, q.py, q.pz, b.z, b.w, c.x, c.y, c.z, c.w);
                const bool first0 = d0 <= d1;
                const float dn = first0 ? d0 : d1, df = first0 ? d1 : d0;
                const int near = 2 * node + (first0 ? 0 : 1), far = 2 * node + (first0 ? 1 : 0);
                if (df <= q.thr2) {      // one push per level at most: sp <= log2(n_leaves) < TREE_STACK
                    st_node[sp] = far; st_d[sp] = df; ++sp; }
                node = near;
                nd = dn;
                if (!(dn <= q.thr2)) { node = -1; break; }
            }
            if (node >= n_leaves) {
                const float4* lp = pts + (size_t)(node - n_leaves) * TREE_LEAF;
#pragma unroll
                for (int i = 0; i < TREE_LEAF; ++i) {
                    const float4 m = lp[i];
                    q.offer(surrogate_direct(q.px, q.py, q.pz, m.x, m.y, m.z), __float_as_int(m.w));
                }
            }
        }
        if (sp == 0) break;
        --sp;
        node = st_node[sp];
        nd = st_d[sp];
    }
    if (q.k == 0x7fffffff) {
        // nothing was accepted: the FP32 surrogate overflowed (a scene point > 1e19 away from the model).  Plain
        // FP64 scan in index order with the reference's expression, first minimum.
        double bv = __longlong_as_double(0x7ff0000000000000LL);
        q.k = 0;
        for (int kk = 0; kk < nm; ++kk) {
            const double vv = __dsqrt_rn(ref_dist2_pt(q.qx, q.qy, q.qz, m64, kk));
            if (vv < bv) { bv = vv; q.k = kk; }
        }
    }
    store_match(mo, m64, j, q.k);
}


// ------------------------------------------------------------------------------------------------
// Wide form: 32-ary hierarchy, ONE WARP PER SCENE POINT.  Leaf = 32 consecutive Morton-sorted model points
// (one per lane); a node = 32 child boxes stored as six 32-float rows (lo.x[32] .. hi.z[32]) so that the warp
// reads them with six coalesced 128-byte loads and every lane measures one child.  The whole warp follows
// one path (no divergence), the tree is 3 levels deep for horse and 4 for 2^22 points, and the per-level
// cursor (node, remaining-children mask) lives in the registers of lane `level` — there is no stack in memory.
// Same bound, same acceptance rule and same FP64 decision as the thread-per-point kernel above.
// Measured and rejected here too: 32 registers / 8 CTAs per SM (+8 %: the kernel is issue bound, not occupancy
// bound), a square-root-free candidate merge with the scene point re-read per candidate (+4 %).
constexpr int WT_WARPS = 8;     // scene points per CTA

// Leaf quality.  32 consecutive points of a Morton order make poor leaves on a surface (the curve jumps): on horse a
// search had 10.2 leaves inside its final radius.  Re-ordering every block of KD_B consecutive Morton-sorted points by
// median splits along the widest axis (a k-d order, 6 levels: 2048 -> 32 points) brings that to 4.7 leaves and 4.8 -> 3.9
// level-0 blocks (CPU emulation on horse, profiles/r01i_tree_order.md) while the coarse levels keep the Morton
// structure.  One CTA per block, everything in shared memory: per level a bounding box per segment (warp reduction +
// shared atomics on order-preserving integer keys), then one bitonic sort of ALL segments at once by the coordinate
// of their own widest axis.  Padding (+inf) sorts to the end of its segment at every level, so the valid points stay a
// contiguous prefix.
constexpr int KD_B = 2048;
constexpr int KD_T = KD_B / 2;
constexpr int KD_MIN_POINTS = 8192;     // below: the search is launch / latency bound and the ~0.1 ms of this kernel does not pay (cow)

__device__ __forceinline__ int kd_enc(float f)          // order-preserving float -> int
{
    const int i = __float_as_int(f);
    return i >= 0 ? i : i ^ 0x7fffffff;
}
__device__ __forceinline__ float kd_dec(int i) { return __int_as_float(i >= 0 ? i : i ^ 0x7fffffff); }
__device__ __forceinline__ float kd_comp(const float4& v, int ax) { return ax == 0 ? v.x : (ax == 1 ? v.y : v.z); }

__global__ void __launch_bounds__(KD_T) wtree_local_kd_kernel(float4* __restrict__ pts, int nm)
{
    __shared__ float4 P[KD_B];
    __shared__ int smin[KD_B / 64][3], smax[KD_B / 64][3];
    __shared__ int axis[KD_B / 64];
    const int tid = threadIdx.x, lane = tid & 31;
    const int base = blockIdx.x * KD_B;
    const int valid = nm - base < KD_B ? nm - base : KD_B;
    for (int e = tid; e < KD_B; e += KD_T)
        P[e] = e < valid ? pts[base + e] : make_float4(INFINITY, INFINITY, INFINITY, __int_as_float(-1));
    __syncthreads();
    for (int S = KD_B; S >= 64; S >>= 1) {
        const int nseg = KD_B / S;
        if (tid < nseg * 3) { smin[tid / 3][tid % 3] = 0x7fffffff; smax[tid / 3][tid % 3] = (int)0x80000000; }
        __syncthreads();
        for (int e = tid; e < KD_B; e += KD_T) {       // a warp's 32 elements lie in one segment (S >= 64)
            const float4 v = P[e];
            float lo[3] = {v.x, v.y, v.z}, hi[3] = {v.x, v.y, v.z};
            if (e >= valid) { lo[0] = lo[1] = lo[2] = INFINITY; hi[0] = hi[1] = hi[2] = -INFINITY; }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1)
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    lo[c] = fminf(lo[c], __shfl_xor_sync(0xffffffffu, lo[c], o));
                    hi[c] = fmaxf(hi[c], __shfl_xor_sync(0xffffffffu, hi[c], o));
                }
            if (lane == 0) {
                const int sg = e / S;
#pragma unroll
                for (int c = 0; c < 3; ++c) { atomicMin(&smin[sg][c], kd_enc(lo[c])); atomicMax(&smax[sg][c], kd_enc(hi[c])); }
            }
        }
        __syncthreads();
        if (tid < nseg) {
            float best = -1.0f;
            int ax = 0;
            for (int c = 0; c < 3; ++c) {
                const float ext = kd_dec(smax[tid][c]) - kd_dec(smin[tid][c]);     // NaN / -inf for an all-padding segment
                if (ext > best) { best = ext; ax = c; }
            }
            axis[tid] = ax;
        }
        __syncthreads();
        for (int k = 2; k <= S; k <<= 1)
            for (int j = k >> 1; j > 0; j >>= 1) {
                // a warp's 32 pairs span 64 consecutive elements: stages with j < 32 exchange inside the warp's own span
                // and need no block-wide barrier (201 of the 251 stages)
                if (j >= 32) __syncthreads(); else __syncwarp();
                const int i = ((tid & ~(j - 1)) << 1) | (tid & (j - 1));           // bit j clear; partner has it set
                const int ip = i | j;
                const bool asc = (k == S) || ((i & k) == 0);                       // last merge: every segment ascending
                const int ax = axis[i / S];
                const float4 a = P[i], b = P[ip];
                if ((kd_comp(a, ax) > kd_comp(b, ax)) == asc) { P[i] = b; P[ip] = a; }
            }
        __syncthreads();
    }
    for (int e = tid; e < valid; e += KD_T) pts[base + e] = P[e];
}

// The same median splits ABOVE the 2048-point blocks, for large models (CPU emulation, 2^17 uniform points, binary form:
// 35 boxes inside the final radius with local splits only, 22 with 4 more levels, 20 with a k-d order from the top).
// A level = bounding box per segment (block reduction + atomics on order-preserving keys), widest axis, one radix
// sort of (segment, coordinate) keys, one gather.  Segments are equal-sized and aligned, so the segment of an element
// is its index / S and nothing but the order changes.
constexpr int KD_GLOBAL_LEVELS = 4;
constexpr int KD_GLOBAL_MIN_POINTS = 131072;

__global__ void kd_fill_kernel(int* __restrict__ smin, int* __restrict__ smax, int n)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) { smin[i] = 0x7fffffff; smax[i] = (int)0x80000000; }
}

__global__ void __launch_bounds__(256) kd_seg_bbox_kernel(const float4* __restrict__ pts, int nm, int S,
                                                          int* __restrict__ smin, int* __restrict__ smax)
{
    __shared__ float red[6][8];
    const int i = blockIdx.x * 256 + threadIdx.x;           // S is a multiple of 256: one segment per block
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    if (i < nm) {
        const float4 v = pts[i];
        lo[0] = hi[0] = v.x; lo[1] = hi[1] = v.y; lo[2] = hi[2] = v.z;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            lo[c] = fminf(lo[c], __shfl_xor_sync(0xffffffffu, lo[c], o));
            hi[c] = fmaxf(hi[c], __shfl_xor_sync(0xffffffffu, hi[c], o));
        }
    if ((threadIdx.x & 31) == 0)
#pragma unroll
        for (int c = 0; c < 3; ++c) { red[c][threadIdx.x >> 5] = lo[c]; red[3 + c][threadIdx.x >> 5] = hi[c]; }
    __syncthreads();
    if (threadIdx.x < 3) {
        float a = INFINITY, b = -INFINITY;
        for (int w = 0; w < 8; ++w) { a = fminf(a, red[threadIdx.x][w]); b = fmaxf(b, red[3 + threadIdx.x][w]); }
        const int sg = (blockIdx.x * 256) / S;
        atomicMin(&smin[sg * 3 + threadIdx.x], kd_enc(a));
        atomicMax(&smax[sg * 3 + threadIdx.x], kd_enc(b));
    }
}

__global__ void kd_seg_axis_kernel(const int* __restrict__ smin, const int* __restrict__ smax, int nseg, int* __restrict__ axis)
{
    const int sg = blockIdx.x * blockDim.x + threadIdx.x;
    if (sg >= nseg) return;
    float best = -1.0f;
    int ax = 0;
    for (int c = 0; c < 3; ++c) {
        const float ext = kd_dec(smax[sg * 3 + c]) - kd_dec(smin[sg * 3 + c]);
        if (ext > best) { best = ext; ax = c; }
    }
    axis[sg] = ax;
}

__global__ void kd_keys_kernel(const float4* __restrict__ pts, int nm, int S, const int* __restrict__ axis,
                               unsigned long long* __restrict__ keys, int32_t* __restrict__ vals)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nm) return;
    const int sg = i / S;
    unsigned u = __float_as_uint(kd_comp(pts[i], axis[sg]));
    u ^= (u >> 31) ? 0xffffffffu : 0x80000000u;             // order-preserving float -> unsigned
    keys[i] = ((unsigned long long)sg << 32) | u;
    vals[i] = i;
}

__global__ void kd_gather_kernel(const float4* __restrict__ src, const int32_t* __restrict__ vals, int nm, float4* __restrict__ dst)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < nm) dst[i] = src[vals[i]];
}

__global__ void wtree_leaf_boxes_kernel(const float4* __restrict__ pts, int n_leaves, int n_padded,
                                        float* __restrict__ boxes0)
{
    const int l = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (l >= n_padded) return;
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    if (l < n_leaves) {
        const float4 q = pts[(size_t)l * 32 + lane];
        lo[0] = hi[0] = q.x; lo[1] = hi[1] = q.y; lo[2] = hi[2] = q.z;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1)
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                lo[c] = fminf(lo[c], __shfl_xor_sync(0xffffffffu, lo[c], o));
                hi[c] = fmaxf(hi[c], __shfl_xor_sync(0xffffffffu, hi[c], o));
            }
    }
    if (lane == 0) {
        float* blk = boxes0 + (size_t)(l >> 5) * 192 + (l & 31);
#pragma unroll
        for (int c = 0; c < 3; ++c) { blk[c * 32] = lo[c]; blk[(3 + c) * 32] = hi[c]; }
    }
}

// box i of the upper level = union of the 32 child boxes in block i of the lower level
__global__ void wtree_level_kernel(const float* __restrict__ lower, int n_upper, int n_padded, float* __restrict__ upper)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_padded) return;
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    if (i < n_upper) {
        const float* blk = lower + (size_t)i * 192;
        for (int ch = 0; ch < 32; ++ch)
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                lo[c] = fminf(lo[c], blk[c * 32 + ch]);
                hi[c] = fmaxf(hi[c], blk[(3 + c) * 32 + ch]);
            }
    }
    float* o = upper + (size_t)(i >> 5) * 192 + (i & 31);
#pragma unroll
    for (int c = 0; c < 3; ++c) { o[c * 32] = lo[c]; o[(3 + c) * 32] = hi[c]; }
}

__global__ void __launch_bounds__(WT_WARPS * 32)
nn_wtree_kernel(const float* __restrict__ boxes, WideLevels lv, const float4* __restrict__ pts,
                const float4* __restrict__ mhat, const Pt64* __restrict__ m64, const double* __restrict__ geo, const double* __restrict__ p64x,
                const double* __restrict__ p64y, const double* __restrict__ p64z, const float* __restrict__ phx,
                const float* __restrict__ phy, const float* __restrict__ phz, int np, int nm,
                MatchOut mo, int use_prev, int* __restrict__ ctl, int honor_done, int* __restrict__ progress)
{
    publish_progress(progress, ctl);
    if (honor_done && ctl[CTL_DONE]) return;
    const int lane = threadIdx.x & 31;
    const int j = blockIdx.x * WT_WARPS + (threadIdx.x >> 5);
    if (j >= np) return;                                    // warp-uniform
    constexpr unsigned FULL = 0xffffffffu;

    TreeQuery q;                                            // warp-uniform except (v, k), which are per lane
    q.px = phx[j]; q.py = phy[j]; q.pz = phz[j];
    q.qx = p64x[j]; q.qy = p64y[j]; q.qz = p64z[j];
    q.m64 = m64;
    if (!scene_point_finite(q.qx, q.qy, q.qz)) {            // warp-uniform
        if (lane == 0) store_match(mo, m64, j, 0);
        return;
    }
    {
        const float Rm = __fmul_ru(__double2float_ru(geo[3]), 1.000001f);
        const float H = __fsqrt_ru(__fmaf_ru(q.pz, q.pz, __fmaf_ru(q.py, q.py, __fmul_ru(q.px, q.px))));
        q.twoE = __fadd_ru(__fmul_ru(1.1920932e-7f, __fadd_ru(H, Rm)), 1.0e-30f);
    }
    q.best2 = q.thr2 = 3.4028235e38f;
    q.v = __longlong_as_double(0x7ff0000000000000LL);
    q.k = 0x7fffffff;
    if (use_prev) {
        const int kp = mo.idx[j];
        if (kp >= 0 && kp < nm) {
            const float4 m = mhat[kp];
            const float d2 = surrogate_direct(q.px, q.py, q.pz, m.x, m.y, m.z);
            // Only a FINITE surrogate may tighten the threshold: an overflowed one (scene point > 1e19 away) would set
            // thr2 = +inf, which opens the EMPTY boxes of padded leaves (+inf) - reads past the sorted points (found by
            // the non-finite-input test of round 2) - and would keep last iteration's match without a search.  Such a
            // point falls through to the FP64 scan below, like on the first iteration.  NaN (model point 0 is NaN): same.
            if (d2 <= 3.4028235e38f) {
                q.set_best(d2);
                if (lane == 0) {
                    q.v = __dsqrt_rn(ref_dist2_pt(q.qx, q.qy, q.qz, m64, kp));
                    q.k = kp;
                }
            }
        }
    }

    // cursor of level l in lane l: block index at that level, children not yet opened
    const long long my_off = lane < lv.T ? lv.off[lane] : 0;
    int my_node = 0;
    unsigned my_mask = FULL;
    int lvl = lv.T - 1;
    for (;;) {
        const int b = __shfl_sync(FULL, my_node, lvl);
        const unsigned msk = __shfl_sync(FULL, my_mask, lvl);
        const float* blk = boxes + __shfl_sync(FULL, my_off, lvl) + (size_t)b * 192 + lane;
        const float d = box_dist2_rd(q.px, q.py, q.pz, blk[0], blk[32], blk[64], blk[96], blk[128], blk[160]);
        const unsigned alive = __ballot_sync(FULL, d <= q.thr2) & msk;
        if (alive == 0u) {
            if (++lvl >= lv.T) break;
            continue;
        }
        if (lvl > 0) {
            const unsigned key = ((alive >> lane) & 1u) ? ((__float_as_uint(d) & ~31u) | (unsigned)lane) : 0xffffffffu;
            const int c = (int)(__reduce_min_sync(FULL, key) & 31u);        // nearest child first
            if (lane == lvl) my_mask = alive & ~(1u << c);                   // thr2 only shrinks: the others stay out
            --lvl;
            if (lane == lvl) { my_node = b * 32 + c; my_mask = FULL; }
            continue;
        }
        // lvl == 0: the children are leaves.  All of them are handled here, nearest first, with the box distances
        // kept in registers; a leaf drops out as soon as the shrinking threshold passes its box.
        unsigned rem = alive;
        while (rem != 0u) {
            const unsigned key = ((rem >> lane) & 1u) ? ((__float_as_uint(d) & ~31u) | (unsigned)lane) : 0xffffffffu;
            const int c = (int)(__reduce_min_sync(FULL, key) & 31u);
            const float4 m = pts[((size_t)b * 32 + c) * 32 + lane];
            const float d2 = surrogate_direct(q.px, q.py, q.pz, m.x, m.y, m.z);
            const float dmin = __uint_as_float(__reduce_min_sync(FULL, __float_as_uint(d2)));   // d2 >= 0: bits are monotone
            if (dmin < q.best2) q.set_best(dmin);
            if (d2 <= q.thr2) {
                const int kk = __float_as_int(m.w);
                const double vv = __dsqrt_rn(ref_dist2_pt(q.qx, q.qy, q.qz, m64, kk));
                if (vv < q.v || (vv == q.v && kk < q.k)) { q.v = vv; q.k = kk; }
            }
            rem &= ~(1u << c) & __ballot_sync(FULL, d <= q.thr2);
        }
        if (++lvl >= lv.T) break;
    }
    double v = q.v;
    int k = q.k;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const double ov = __shfl_xor_sync(FULL, v, o);
        const int ok = __shfl_xor_sync(FULL, k, o);
        if (ov < v || (ov == v && ok < k)) { v = ov; k = ok; }
    }
    k = __shfl_sync(FULL, k, 0);  // one decision for the warp (with NaN values the butterfly is not a total order)
    if (k == 0x7fffffff) {       // FP32 surrogate overflowed (scene point > 1e19 away): FP64 scan, first minimum
        v = __longlong_as_double(0x7ff0000000000000LL);
        for (int kk = lane; kk < nm; kk += 32) {
            const double vv = __dsqrt_rn(ref_dist2_pt(q.qx, q.qy, q.qz, m64, kk));
            if (vv < v) { v = vv; k = kk; }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const double ov = __shfl_xor_sync(FULL, v, o);
            const int ok = __shfl_xor_sync(FULL, k, o);
            if (ov < v || (ov == v && ok < k)) { v = ov; k = ok; }
        }
        if (k == 0x7fffffff) k = 0;
    }
    if (lane == 0) store_match(mo, m64, j, k);
}

}  // namespace

// Scratch of the global k-d levels for nm points: two key arrays, two value arrays, one point array, segment tables,
// the radix sort's temporary.
size_t kd_scratch_bytes(int nm, size_t* cub_bytes_out)
{
    size_t cub_bytes = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, cub_bytes, (unsigned long long*)nullptr, (unsigned long long*)nullptr,
                                    (int32_t*)nullptr, (int32_t*)nullptr, nm, 0, 64, (cudaStream_t)0);
    if (cub_bytes_out) *cub_bytes_out = cub_bytes;
    const size_t n = ((size_t)nm + 255) / 256 * 256;
    return n * (2 * 8 + 2 * 4 + 16) + 3 * 4 * 7 * (n / KD_B + 2) + cub_bytes + 1024;
}

// pts[0, nm): Morton order in, k-d order of the top KD_GLOBAL_LEVELS levels out (in place; `scratch` as sized above)
static cudaError_t kd_global_levels(float4* pts, int nm, void* scratch, cudaStream_t st)
{
    size_t cub_bytes = 0;
    kd_scratch_bytes(nm, &cub_bytes);
    const size_t n = ((size_t)nm + 255) / 256 * 256;
    char* p = static_cast<char*>(scratch);
    unsigned long long* keys = reinterpret_cast<unsigned long long*>(p); p += n * 8;
    unsigned long long* keys_alt = reinterpret_cast<unsigned long long*>(p); p += n * 8;
    float4* alt = reinterpret_cast<float4*>(p); p += n * 16;
    int32_t* vals = reinterpret_cast<int32_t*>(p); p += n * 4;
    int32_t* vals_alt = reinterpret_cast<int32_t*>(p); p += n * 4;
    const size_t seg_cap = 7 * (n / KD_B + 2);
    int* smin = reinterpret_cast<int*>(p); p += seg_cap * 4;
    int* smax = reinterpret_cast<int*>(p); p += seg_cap * 4;
    int* axis = reinterpret_cast<int*>(p); p += seg_cap * 4;
    p += (256 - (reinterpret_cast<uintptr_t>(p) & 255)) & 255;      // the sort's temporary wants cudaMalloc-like alignment
    void* cub_tmp = p;
    float4 *src = pts, *dst = alt;
    for (int g = 0; g < KD_GLOBAL_LEVELS; ++g) {
        const int S = KD_B << (KD_GLOBAL_LEVELS - g);           // 32768, 16384, 8192, 4096
        const int nseg = (nm + S - 1) / S;
        int seg_bits = 1;
        while ((1 << seg_bits) < nseg) ++seg_bits;
        kd_fill_kernel<<<(nseg * 3 + 255) / 256, 256, 0, st>>>(smin, smax, nseg * 3);
        kd_seg_bbox_kernel<<<(nm + 255) / 256, 256, 0, st>>>(src, nm, S, smin, smax);
        kd_seg_axis_kernel<<<(nseg + 255) / 256, 256, 0, st>>>(smin, smax, nseg, axis);
        kd_keys_kernel<<<(nm + 255) / 256, 256, 0, st>>>(src, nm, S, axis, keys, vals);
        size_t bytes = cub_bytes;
        const cudaError_t e = cub::DeviceRadixSort::SortPairs(cub_tmp, bytes, keys, keys_alt, vals, vals_alt, nm, 0,
                                                              32 + seg_bits, st);
        if (e != cudaSuccess) return e;
        kd_gather_kernel<<<(nm + 255) / 256, 256, 0, st>>>(src, vals_alt, nm, dst);
        float4* t = src; src = dst; dst = t;
    }
    if (src != pts) return cudaMemcpyAsync(pts, src, (size_t)nm * sizeof(float4), cudaMemcpyDeviceToDevice, st);
    return cudaSuccess;
}

// Builds the hierarchy for the model currently in `m` (mhat / geo valid); keys/vals/perm/cub: sort scratch for nm items.
int tree_build(const ModelDev& m, TreeDev& t, bool wide, uint32_t* keys, uint32_t* keys_alt,
               int32_t* vals, int32_t* perm, void* cub_tmp, size_t cub_bytes, void* kd_scratch, cudaStream_t st)
{
    const int nm = (int)m.nm;
    tree_keys_kernel<<<(nm + 255) / 256, 256, 0, st>>>(m.mhat, nm, m.geo, keys, vals);
    cudaError_t e = cub::DeviceRadixSort::SortPairs(cub_tmp, cub_bytes, keys, keys_alt, vals, perm, nm, 0, 30, st);
    if (e != cudaSuccess) return (int)e;
    if (!wide) {
    const int n_slots = t.n_leaves * TREE_LEAF;
    tree_points_kernel<<<(n_slots + 255) / 256, 256, 0, st>>>(m.mhat, perm, nm, n_slots, t.pts);
    if (kd_scratch && (e = kd_global_levels(t.pts, nm, kd_scratch, st)) != cudaSuccess) return (int)e;
    if (nm >= KD_MIN_POINTS) wtree_local_kd_kernel<<<(nm + KD_B - 1) / KD_B, KD_T, 0, st>>>(t.pts, nm);
    tree_leaf_boxes_kernel<<<(t.n_leaves + 255) / 256, 256, 0, st>>>(t.pts, t.n_leaves, (nm + TREE_LEAF - 1) / TREE_LEAF,
                                                                     t.nodes);
    for (int first = t.n_leaves / 2; first >= 1; first /= 2)
        tree_level_kernel<<<(first + 255) / 256, 256, 0, st>>>(t.nodes, first, first);
    return (int)cudaGetLastError();
    }
    // wide form (same sorted order)
    const int n0 = t.w.count[0];
    tree_points_kernel<<<(n0 * 32 + 255) / 256, 256, 0, st>>>(m.mhat, perm, nm, n0 * 32, t.wpts);
    if (kd_scratch && (e = kd_global_levels(t.wpts, nm, kd_scratch, st)) != cudaSuccess) return (int)e;
    if (nm >= KD_MIN_POINTS) wtree_local_kd_kernel<<<(nm + KD_B - 1) / KD_B, KD_T, 0, st>>>(t.wpts, nm);
    {
        const int padded = (n0 + 31) / 32 * 32;
        wtree_leaf_boxes_kernel<<<(padded * 32 + 255) / 256, 256, 0, st>>>(t.wpts, n0, padded, t.wboxes + t.w.off[0]);
    }
    for (int l = 1; l < t.w.T; ++l) {
        const int padded = (t.w.count[l] + 31) / 32 * 32;
        wtree_level_kernel<<<(padded + 127) / 128, 128, 0, st>>>(t.wboxes + t.w.off[l - 1], t.w.count[l], padded,
                                                                 t.wboxes + t.w.off[l]);
    }
    return (int)cudaGetLastError();
}

// Level geometry of the wide tree for nm points; returns the number of floats the boxes need.
size_t wtree_plan(int64_t nm, WideLevels& w)
{
    int64_t n = (nm + 31) / 32;          // leaves
    size_t off = 0;
    w.T = 0;
    for (;;) {
        const int64_t blocks = (n + 31) / 32;
        w.count[w.T] = (int)n;
        w.off[w.T] = (long long)off;
        off += (size_t)blocks * 192;
        ++w.T;
        if (blocks == 1) break;
        n = blocks;
    }
    return off;
}

void launch_nn_wtree(const ModelDev& m, const SceneDev& s, const TreeDev& t, int use_prev, int* ctl, int honor_done,
                     cudaStream_t st, int* progress)
{
    const int grid = (int)((s.np + WT_WARPS - 1) / WT_WARPS);
    nn_wtree_kernel<<<grid, WT_WARPS * 32, 0, st>>>(t.wboxes, t.w, t.wpts, m.mhat, m.m64, m.geo, s.p64x,
                                                    s.p64y, s.p64z, s.phx, s.phy, s.phz, (int)s.np, (int)m.nm, match_out(s),
                                                    use_prev, ctl, honor_done, progress);
}

// Global k-d levels only from 131 072 model points up.  Round 2 measured them at 16 384 too (profiles/r02a_kdmin.md): the
// resident loop gains 10 % (horse 2.79 -> 2.47 ms) but the costlier build loses 3-20 % end to end (horse_tr2 1.46 -> 1.79 ms),
// and a drop-in is called one-shot; the switch was removed.
int kd_global_min_points() { return KD_GLOBAL_MIN_POINTS; }

size_t tree_sort_bytes(int n)
{
    size_t bytes = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, bytes, (uint32_t*)nullptr, (uint32_t*)nullptr, (int32_t*)nullptr,
                                    (int32_t*)nullptr, n, 0, 30, (cudaStream_t)0);
    return bytes;
}

void launch_nn_tree(const ModelDev& m, const SceneDev& s, const TreeDev& t, int use_prev, int* ctl, int honor_done,
                    cudaStream_t st, int* progress)
{
    const int grid = (int)((s.np + TREE_THREADS - 1) / TREE_THREADS);
    nn_tree_kernel<<<grid, TREE_THREADS, 0, st>>>(t.nodes, t.pts, t.n_leaves, m.mhat, m.m64, m.geo,
                                                  s.p64x, s.p64y, s.p64z, s.phx, s.phy, s.phz, (int)s.np, (int)m.nm,
                                                  match_out(s), use_prev, ctl, honor_done, progress);
}

}  // namespace icpb
