// nn_kernels.cu — K1: exact brute-force nearest-neighbour correspondence for sm_100a.
//
// Replaces CPU::closest_matrix (reference src/cpu.cc:5-27) and the GPU pair compute_distance + find_Y
// (src/GPU/compute.cu:94-150) — neither is ported.  The reference materialises an np x nm FP64
// distance matrix in HBM and scans it with one thread per scene point; here nothing is
// materialised: model tiles stream through shared memory and the running minimum stays in
// registers.  Three kernels (geometry constants in icp_common.cuh; the numbers below are the defaults):
//
//  nn_filter  FP32, CUDA cores, issue bound.  Work = one linear sequence of (scene tile of 2048 points, 128-point model
//             chunk) units cut into equal contiguous ranges, one per CTA ("stream-K", 2 CTAs per SM, whole waves).
//             Model tiles (512 points) arrive by TMA bulk copy (cp.async.bulk + mbarrier) as float4,
//             are re-expressed relative to the CTA's own origin and staged as SoA float4 planes
//             X|Y|Z|W (W = |m|^2).  Each thread keeps 16 scene points in registers (coordinates
//             pre-scaled by -2) and evaluates, per LDS.128 of four consecutive model points,
//             d' = |m|^2 - 2 p.m   for 16 x 4 pairs with 96 FFMA2 (fma.rn.f32x2) and 32 FMNMX3 (3-input min).
//             Only the VALUE of the minimum is tracked per pair; the index is recovered lazily: per model chunk the
//             chunk minimum is merged into (best, best chunk id, runner-up over all other chunks), one result per
//             (scene tile, CTA) slot.
//  nn_refine  8 lanes per scene point: merges the tile's slots, derives a rigorous bound tau on
//             the FP32 evaluation error, and if the runner-up chunk is provably farther than tau
//             re-ranks the points of the best chunk in FP64 with the reference's own expression
//             sqrt((dx*dx+dy*dy)+dz*dz), lowest index on ties (shuffle argmin).  Otherwise the point is undecided:
//             the chunk ranges of the slots whose minimum is within the threshold are rescanned (pruned_rescan) -
//             by the block itself while such a range is short, else the point is queued for
//  nn_exact   one block per queued point, the same pruned rescan (large clouds only).
//  Every decision is recorded as (index, the matched model point) - store_match - so that K2 / K4 gather nothing.
//
// Error bound (DESIGN.md §K1 has the derivation).  With u = 2^-24, P = |p_loc|, r = FP32 distance of
// the filter's winner, M = P + r (largest |m_loc| a relevant candidate can have):
//   E   = 6.5 u (M^2 + 2 P M)            rounding of W and of the three chained FMAs
//   eta = 1.01 u (|p^| + Rm + 2P + r)    displacement caused by rounding inputs to FP32 (twice)
//   tau = 2E + 4 eta r + 4 eta^2 + 2^-40 r^2
// Any model point whose true FP64 distance is <= the winner's has d' <= best + tau.
#include "icp_common.cuh"

namespace icpb {

namespace {

__device__ __forceinline__ uint32_t smem_u32(const void* p)
{
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst_smem)),
                 "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity)
{
    uint32_t done;
    do {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.b32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(smem_u32(bar)), "r"(parity)
            : "memory");
    } while (!done);
}
__device__ __forceinline__ float min3(float a, float b, float c)
{
    float r;
    asm("min.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
    return r;
}

constexpr int SOA_PLANE = NN_TM;                       // floats per plane
constexpr int SOA_BUF = 4 * SOA_PLANE;                 // floats per buffer (X|Y|Z|W)
constexpr size_t FILTER_SMEM = NN_TM * sizeof(float4) + 2 * SOA_BUF * sizeof(float);

// ------------------------------------------------------------------------------------------------
// Work decomposition ("stream-K"): the np_pad/NN_TILE scene tiles x cpt model chunks (NN_G points each) form one
// linear sequence of T = n_tiles * cpt chunk-units, scene tile major.  The grid is whole waves of CTAs (NN_MIN_CTAS per SM)
// and CTA b owns the contiguous range [b*T/n, (b+1)*T/n): every CTA gets the same work to within one chunk,
// whatever the cloud sizes, so all SMs finish together.  A range may end inside a scene tile and continue in
// the next one; each (CTA, scene tile) segment writes its own partial result into slot = b - first CTA of that
// tile, and nn_refine merges a tile's slots (slot_range() gives their number from the same arithmetic).
__device__ __forceinline__ int cta_of_unit(long long x, long long T, int n)      // the CTA whose range holds unit x
{
    return (int)(((x + 1) * n - 1) / T);
}

// Chunks [c_lo, c_hi) of scene tile `tile` that its slot `slot` covers: the same integers nn_filter_kernel walks.
__device__ __forceinline__ void slot_chunks(int tile, int slot, long long T, int n_cta, int cpt, int& c_lo, int& c_hi)
{
    const long long x0 = (long long)tile * cpt;
    const long long b = (long long)cta_of_unit(x0, T, n_cta) + slot;
    const long long begin = b * T / n_cta, end = (b + 1) * T / n_cta;
    c_lo = (int)((begin > x0 ? begin : x0) - x0);
    c_hi = (int)((end < x0 + cpt ? end : x0 + cpt) - x0);
}

__global__ void __launch_bounds__(NN_THREADS, NN_MIN_CTAS)
nn_filter_kernel(const float4* __restrict__ mhat, const float* __restrict__ phx, const float* __restrict__ phy,
                 const float* __restrict__ phz, int np_pad, int n_tiles, int cpt,
                 float* __restrict__ out_best, float* __restrict__ out_runner, uint32_t* __restrict__ out_chunk,
                 const int* __restrict__ ctl, int honor_done, int* __restrict__ progress)
{
    publish_progress(progress, ctl);
    if (honor_done && ctl[CTL_DONE]) return;

    extern __shared__ __align__(128) unsigned char smem[];
    float4* raw = reinterpret_cast<float4*>(smem);
    float* soa = reinterpret_cast<float*>(smem + NN_TM * sizeof(float4));
    __shared__ __align__(8) uint64_t mbar;

    const int tid = threadIdx.x;
    constexpr int CPM = NN_TM / NN_G;                       // chunks per staged model tile (8)
    // this CTA's range, reduced to 32-bit loop state: (tile, c_lo, remaining)
    int tile, c_lo, remaining;
    {
        const long long T = (long long)n_tiles * cpt;
        const long long begin = (long long)blockIdx.x * T / (long long)gridDim.x;
        const long long end = (long long)(blockIdx.x + 1) * T / (long long)gridDim.x;
        tile = (int)(begin / cpt);
        c_lo = (int)(begin - (long long)tile * cpt);
        remaining = (int)(end - begin);
    }

    if (tid == 0) {
        mbar_init(&mbar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    uint32_t parity = 0;

    while (remaining > 0) {
        // ---- one (CTA, scene tile) segment: model chunks [c_lo, c_hi) of scene tile `tile` ------------
        const int c_hi = min(cpt, c_lo + remaining);
        const int slot = (int)blockIdx.x - cta_of_unit((long long)tile * cpt, (long long)n_tiles * cpt, (int)gridDim.x);
        const int mt0 = c_lo / CPM, mt1 = (c_hi + CPM - 1) / CPM;
        const int pbase = tile * NN_TILE;

        if (tid == 0) {
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            mbar_expect_tx(&mbar, NN_TM * sizeof(float4));
            bulk_g2s(raw, mhat + (size_t)mt0 * NN_TM, NN_TM * sizeof(float4), &mbar);
        }

        // CTA-local origin: the middle point of the (Morton-sorted) scene tile.
        const float cx = phx[pbase + NN_TILE / 2], cy = phy[pbase + NN_TILE / 2], cz = phz[pbase + NN_TILE / 2];

        float2 ax[NN_R], ay[NN_R], az[NN_R];
        float best[NN_R], runner[NN_R];
        uint32_t bchunk[NN_R];
#pragma unroll
        for (int r = 0; r < NN_R; ++r) {
            const int j = pbase + r * NN_THREADS + tid;
            const float x = -2.0f * __fsub_rn(phx[j], cx);
            const float y = -2.0f * __fsub_rn(phy[j], cy);
            const float z = -2.0f * __fsub_rn(phz[j], cz);
            ax[r] = make_float2(x, x);
            ay[r] = make_float2(y, y);
            az[r] = make_float2(z, z);
            best[r] = __int_as_float(0x7f800000);
            runner[r] = __int_as_float(0x7f800000);
            bchunk[r] = 0u;
        }

        for (int mt = mt0; mt < mt1; ++mt) {
            float* buf = soa + ((mt - mt0) & 1) * SOA_BUF;
            // ---- stage: raw float4 (TMA) -> CTA-local SoA planes --------------------------------------
            mbar_wait(&mbar, parity);
            parity ^= 1u;
#pragma unroll
            for (int k = 0; k < NN_TM / NN_THREADS; ++k) {
                const int i = k * NN_THREADS + tid;
                const float4 v = raw[i];
                const float x = __fsub_rn(v.x, cx), y = __fsub_rn(v.y, cy), z = __fsub_rn(v.z, cz);
                buf[i] = x;
                buf[SOA_PLANE + i] = y;
                buf[2 * SOA_PLANE + i] = z;
                buf[3 * SOA_PLANE + i] = __fmaf_rn(z, z, __fmaf_rn(y, y, __fmul_rn(x, x)));
            }
            __syncthreads();   // planes visible; everyone is done with `raw` and with the other buffer
            if (tid == 0 && mt + 1 < mt1) {
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                mbar_expect_tx(&mbar, NN_TM * sizeof(float4));
                bulk_g2s(raw, mhat + (size_t)(mt + 1) * NN_TM, NN_TM * sizeof(float4), &mbar);
            }

            // ---- compute: NN_R scene points x (up to) 512 model points per thread -----------------------
            const float4* X4 = reinterpret_cast<const float4*>(buf);
            const float4* Y4 = reinterpret_cast<const float4*>(buf + SOA_PLANE);
            const float4* Z4 = reinterpret_cast<const float4*>(buf + 2 * SOA_PLANE);
            const float4* W4 = reinterpret_cast<const float4*>(buf + 3 * SOA_PLANE);
            const int cb = max(c_lo - mt * CPM, 0), ce = min(c_hi - mt * CPM, CPM);   // the segment's share of this tile
#pragma unroll 1
            for (int c = cb; c < ce; ++c) {
                float tmin[NN_R];
#pragma unroll
                for (int r = 0; r < NN_R; ++r) tmin[r] = __int_as_float(0x7f800000);
#pragma unroll NN_Q_UNROLL
                for (int q = 0; q < NN_G / 4; ++q) {
                    const int o = c * (NN_G / 4) + q;
                    const float4 X = X4[o], Y = Y4[o], Z = Z4[o], W = W4[o];
                    const float2 x01 = make_float2(X.x, X.y), x23 = make_float2(X.z, X.w);
                    const float2 y01 = make_float2(Y.x, Y.y), y23 = make_float2(Y.z, Y.w);
                    const float2 z01 = make_float2(Z.x, Z.y), z23 = make_float2(Z.z, Z.w);
                    const float2 w01 = make_float2(W.x, W.y), w23 = make_float2(W.z, W.w);
#pragma unroll
                    for (int r = 0; r < NN_R; ++r) {
                        float2 t01 = __ffma2_rn(az[r], z01, w01);
                        float2 t23 = __ffma2_rn(az[r], z23, w23);
                        t01 = __ffma2_rn(ay[r], y01, t01);
                        t23 = __ffma2_rn(ay[r], y23, t23);
                        t01 = __ffma2_rn(ax[r], x01, t01);
                        t23 = __ffma2_rn(ax[r], x23, t23);
                        tmin[r] = min3(tmin[r], t01.x, t01.y);
                        tmin[r] = min3(tmin[r], t23.x, t23.y);
                    }
                }
                const uint32_t cid = (uint32_t)mt * CPM + (uint32_t)c;
#pragma unroll
                for (int r = 0; r < NN_R; ++r) {
                    const float tm = tmin[r];
                    const bool lt = tm < best[r];
                    runner[r] = fminf(runner[r], fmaxf(best[r], tm));
                    bchunk[r] = lt ? cid : bchunk[r];
                    best[r] = fminf(best[r], tm);
                }
            }
        }

        const size_t obase = (size_t)slot * np_pad + pbase;
#pragma unroll
        for (int r = 0; r < NN_R; ++r) {
            const size_t j = obase + r * NN_THREADS + tid;
            out_best[j] = best[r];
            out_runner[j] = runner[r];
            out_chunk[j] = bchunk[r];
        }
        remaining -= c_hi - c_lo;
        c_lo = 0;
        ++tile;
        __syncthreads();       // the shared buffers are reused by the next segment
    }
}

// ------------------------------------------------------------------------------------------------
// The reference's distance expression, FP64, no contraction (src/cpu.cc:16-19 with pow(x,2) == x*x):
// the squared sum; the sqrt is applied lazily (see Best64).
__device__ __forceinline__ double dist2_64(double px, double py, double pz, double mx, double my, double mz)
{
    const double dx = __dsub_rn(px, mx), dy = __dsub_rn(py, my), dz = __dsub_rn(pz, mz);
    return __dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(dz, dz));
}

// Running first-minimum of v = sqrt(s) over ascending indices, exactly as Eigen's minCoeff sees the
// reference's d(k) = sqrt(...) (src/cpu.cc:19-22), but evaluating the sqrt only when s improves:
// sqrt is monotone, so s >= s_best can never give a strictly smaller v.
constexpr int NO_INDEX = 0x7fffffff;
struct Best64 {
    double s, v;
    int k;
    __device__ __forceinline__ void init() { s = v = __longlong_as_double(0x7ff0000000000000LL); k = NO_INDEX; }
    __device__ __forceinline__ void feed(double s_new, int k_new)   // k_new > k of everything fed before
    {
        if (s_new < s) {
            const double v_new = __dsqrt_rn(s_new);
            if (v_new < v) { v = v_new; k = k_new; }
            s = s_new;   // equal sqrt: the earlier index stays, the smaller s keeps filtering
        }
    }
};
// Nothing was ever accepted: every distance of the row is +inf or NaN (a non-finite scene point, or one so far away
// that the FP64 squares overflow).  Eigen's minCoeff visitor starts at column 0 and only moves on a strict `<`
// (src/cpu.cc:21-22), so the reference's answer is index 0.
__device__ __forceinline__ int first_min_or_zero(int k) { return k == NO_INDEX ? 0 : k; }

// lexicographic (value, index) minimum over `width` adjacent lanes: lowest index wins ties
template <int WIDTH>
__device__ __forceinline__ void group_argmin(double& v, int& k, unsigned mask = 0xffffffffu)
{
#pragma unroll
    for (int o = WIDTH / 2; o > 0; o >>= 1) {
        const double ov = __shfl_xor_sync(mask, v, o);
        const int ok = __shfl_xor_sync(mask, k, o);
        if (ov < v || (ov == v && ok < k)) { v = ov; k = ok; }
    }
}

// The filter's FP32 surrogate d' = |m_loc|^2 - 2 p_loc.m_loc, re-evaluated with the SAME operation
// sequence as nn_filter_kernel (fsub, mul+fma+fma for W, three chained FMAs; FFMA2 is an IEEE FMA per
// lane), so the value is bit-identical to what the filter minimised.  (cx,cy,cz) = the CTA origin of the
// scene point's tile, (ax,ay,az) = -2 p_loc.
__device__ __forceinline__ float surrogate32(const float4 mh, float cx, float cy, float cz, float ax, float ay, float az)
{
    const float x = __fsub_rn(mh.x, cx), y = __fsub_rn(mh.y, cy), z = __fsub_rn(mh.z, cz);
    const float w = __fmaf_rn(z, z, __fmaf_rn(y, y, __fmul_rn(x, x)));
    return __fmaf_rn(ax, x, __fmaf_rn(ay, y, __fmaf_rn(az, z, w)));
}

// ---- block-cooperative exact scan (shared by nn_refine's in-kernel path and nn_exact) ---------------------------
// A block of SCAN_THREADS threads rescans model points [k_begin, k_end) for P undecided scene points at once: FP32
// surrogate against each point's acceptance threshold; every survivor (a superset of the reference's possible argmins)
// gets the reference's FP64 expression; first minimum wins.  A model point is loaded once for the P points.
constexpr int SCAN_THREADS = 256;
constexpr int EX_U = 4;          // model points in flight per thread
struct ScanPoint { float cx, cy, cz, ax, ay, az, thr; };

__device__ __forceinline__ ScanPoint scan_point(const float* __restrict__ phx, const float* __restrict__ phy,
                                                const float* __restrict__ phz, int j, float thr)
{
    ScanPoint q;
    const int cj = (j / NN_TILE) * NN_TILE + NN_TILE / 2;
    q.cx = phx[cj]; q.cy = phy[cj]; q.cz = phz[cj];
    q.ax = -2.0f * __fsub_rn(phx[j], q.cx);
    q.ay = -2.0f * __fsub_rn(phy[j], q.cy);
    q.az = -2.0f * __fsub_rn(phz[j], q.cz);
    q.thr = thr;
    return q;
}

template <int P>
__device__ __forceinline__ void scan_model(const float4* __restrict__ mhat, const Pt64* __restrict__ m64, int nm_pad,
                                           int k_begin, int k_end, const ScanPoint (&q)[P], const double (*sp)[3],
                                           Best64 (&bt)[P])
{
    for (int kb = k_begin + (int)threadIdx.x; kb < k_end; kb += SCAN_THREADS * EX_U) {
        float4 mh[EX_U];
#pragma unroll
        for (int u = 0; u < EX_U; ++u) mh[u] = mhat[min(kb + u * SCAN_THREADS, nm_pad - 1)];
#pragma unroll
        for (int u = 0; u < EX_U; ++u) {
            const int kk = kb + u * SCAN_THREADS;
#pragma unroll
            for (int p = 0; p < P; ++p) {
                const float d = surrogate32(mh[u], q[p].cx, q[p].cy, q[p].cz, q[p].ax, q[p].ay, q[p].az);
                if (!(d > q[p].thr) && kk < k_end) {
                    double mx, my, mz;
                    ld_pt64(m64, kk, mx, my, mz);
                    bt[p].feed(dist2_64(sp[p][0], sp[p][1], sp[p][2], mx, my, mz), kk);
                }
            }
        }
    }
}

// (value, index) minimum of the block's P running minima; the result is valid in thread 0.  Ends with a barrier, so
// sv / sk may be reused right after.
template <int P>
__device__ __forceinline__ void block_argmin(const Best64 (&bt)[P], double (*sv)[SCAN_THREADS / 32],
                                             int (*sk)[SCAN_THREADS / 32], double (&v_out)[P], int (&k_out)[P])
{
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
    for (int p = 0; p < P; ++p) {
        double v = bt[p].v;
        int k = bt[p].k;
        group_argmin<32>(v, k);
        if (lane == 0) { sv[p][w] = v; sk[p][w] = k; }
    }
    __syncthreads();
    if (w == 0) {
#pragma unroll
        for (int p = 0; p < P; ++p) {
            double v = lane < SCAN_THREADS / 32 ? sv[p][lane] : __longlong_as_double(0x7ff0000000000000LL);
            int k = lane < SCAN_THREADS / 32 ? sk[p][lane] : NO_INDEX;
            group_argmin<32>(v, k);
            v_out[p] = v;
            k_out[p] = k;
        }
    }
    __syncthreads();
}

constexpr int REFINE_THREADS = SCAN_THREADS;
#ifndef ICP_RF
#define ICP_RF 8
#endif
constexpr int RF = ICP_RF;    // lanes cooperating on one scene point (8; build-time variant like ICP_NN_*)
static_assert(RF == 4 || RF == 8 || RF == 16, "RF");
static_assert(NN_G / RF <= 32, "nn_refine keeps one candidate bit per (lane, step) of the best chunk in a 32-bit mask");
constexpr int RF_PPW = 32 / RF;                              // scene points per warp
constexpr int RF_PPB = (REFINE_THREADS / 32) * RF_PPW;       // scene points per block and trip (32)

// Rescan for ONE undecided scene point j by the whole block.  Every model point that can be the reference's argmin has
// surrogate d' <= thr, and the filter left, per slot of j's scene tile, the exact minimum of that very surrogate over
// the slot's chunk range: a slot whose minimum is above thr holds no candidate and is skipped.  On horse that leaves
// one or two of 15 slots (~3 500 of 48 485 model points), on 2^22 points one or two of 66 (round 2; round 1 rescanned
// the whole model).  NaN / +inf thresholds skip nothing.  Result (first minimum, or NO_INDEX) valid in thread 0.
__device__ __forceinline__ int pruned_rescan(const float4* __restrict__ mhat, const Pt64* __restrict__ m64, int nm, int nm_pad,
                                             const float* __restrict__ phx, const float* __restrict__ phy,
                                             const float* __restrict__ phz, const double* __restrict__ p64x,
                                             const double* __restrict__ p64y, const double* __restrict__ p64z,
                                             const float* __restrict__ part_best, int np_pad, int n_tiles, int cpt, int n_cta,
                                             int j, float thr, double (*sp)[3], double (*sv)[SCAN_THREADS / 32],
                                             int (*sk)[SCAN_THREADS / 32])
{
    ScanPoint q[1];
    q[0] = scan_point(phx, phy, phz, j, thr);
    if (threadIdx.x == 0) { sp[0][0] = p64x[j]; sp[0][1] = p64y[j]; sp[0][2] = p64z[j]; }
    __syncthreads();
    Best64 bt[1];
    bt[0].init();
    const int tile = j / NN_TILE;
    const long long T = (long long)n_tiles * cpt, x0 = (long long)tile * cpt;
    const int n_slots = cta_of_unit(x0 + cpt - 1, T, n_cta) - cta_of_unit(x0, T, n_cta) + 1;
    for (int s = 0; s < n_slots; ++s) {                       // ascending slots = ascending model indices
        const float sb = part_best[(size_t)s * np_pad + j];   // block-uniform
        if (sb > thr) continue;
        int c_lo, c_hi;
        slot_chunks(tile, s, T, n_cta, cpt, c_lo, c_hi);
        scan_model<1>(mhat, m64, nm_pad, c_lo * NN_G, min(c_hi * NN_G, nm), q, sp, bt);
    }
    double v[1];
    int k[1];
    block_argmin<1>(bt, sv, sk, v, k);                        // ends with a barrier: sp / sv / sk are free again
    return k[0];
}

// nn_refine: merges the model splits, derives the bound tau and the FP32 acceptance threshold
// thr = best + tau (rounded up).  Every model point that can be the reference's argmin has d' <= thr.
//  - runner > thr ("clear"): all such points lie in the best chunk.  Its NN_G points are re-scored with the
//    FP32 surrogate; only those with d' <= thr (typically one) are evaluated in FP64 with the
//    reference's own expression, first minimum wins.
//  - otherwise the point is undecided and the model is rescanned against its threshold (pruned_rescan):
//      inline_exact = 1 (a slot's chunk range is short: up to 16 384 model points): right here, by the block that met
//        the point (round 1 queued these for a separate kernel: on horse that launch served ~85 points per pass and
//        cost 20 us of the 370 us iteration);
//      inline_exact = 0 (long ranges, worth spreading over the whole GPU): the point is queued with its threshold for
//        nn_exact.
__global__ void __launch_bounds__(REFINE_THREADS, 4)
nn_refine_kernel(const Pt64* __restrict__ m64, const float4* __restrict__ mhat, int nm, int nm_pad,
                 const double* __restrict__ geo,
                 const double* __restrict__ p64x, const double* __restrict__ p64y, const double* __restrict__ p64z,
                 const float* __restrict__ phx, const float* __restrict__ phy, const float* __restrict__ phz, int np,
                 int np_pad, int n_tiles, int cpt, int n_cta, const float* __restrict__ part_best,
                 const float* __restrict__ part_runner,
                 const uint32_t* __restrict__ part_chunk, MatchOut mo, int32_t* __restrict__ amb_list,
                 float* __restrict__ amb_thr, int* __restrict__ ctl, int honor_done, int inline_exact)
{
    if (honor_done && ctl[CTL_DONE]) return;
    __shared__ int s_cnt;
    __shared__ int s_j[RF_PPB];
    __shared__ float s_thr[RF_PPB];
    __shared__ double sp[1][3];
    __shared__ double sv[1][SCAN_THREADS / 32];
    __shared__ int sk[1][SCAN_THREADS / 32];
    const int lane = threadIdx.x & 31, sub = lane & (RF - 1), grp = lane / RF, w = threadIdx.x >> 5;
    const float Rm = __double2float_ru(geo[3]);
    const float INF = __int_as_float(0x7f800000);
    const unsigned gmask = ((1u << RF) - 1u) << (grp * RF);
    int undecided = 0;                                 // thread 0: points this block rescanned itself

    for (int base = blockIdx.x * RF_PPB; base < np; base += gridDim.x * RF_PPB) {      // block-uniform trip count
        if (inline_exact) {
            if (threadIdx.x == 0) s_cnt = 0;
            __syncthreads();
        }
        const int jb = base + w * RF_PPW;
        const bool valid = jb + grp < np;
        const int j = valid ? jb + grp : np - 1;
        // everything the point needs, requested up front so the latencies overlap
        const int cj = (j / NN_TILE) * NN_TILE + NN_TILE / 2;
        const float hx = phx[j], hy = phy[j], hz = phz[j];
        const float cx = phx[cj], cy = phy[cj], cz = phz[cj];
        const double px = p64x[j], py = p64y[j], pz = p64z[j];
        // ---- merge the partial results of this scene tile (RF lanes stride over its slots) -----------
        const long long T = (long long)n_tiles * cpt, x0 = (long long)(j / NN_TILE) * cpt;
        const int n_slots = cta_of_unit(x0 + cpt - 1, T, n_cta) - cta_of_unit(x0, T, n_cta) + 1;
        float b = INF, rn = INF;
        uint32_t ch = 0;
        for (int s = sub; s < n_slots; s += RF) {
            const size_t o = (size_t)s * np_pad + j;
            const float sb = part_best[o], sr = part_runner[o];
            const uint32_t sc = part_chunk[o];
            if (sb < b) { rn = fminf(fminf(rn, b), sr); b = sb; ch = sc; }
            else        { rn = fminf(rn, fminf(sr, sb)); }     // sb == b: tie -> runner = best -> ambiguous
        }
        float best = b;
#pragma unroll
        for (int o = RF / 2; o > 0; o >>= 1) best = fminf(best, __shfl_xor_sync(0xffffffffu, best, o));
        const unsigned eq = (__ballot_sync(0xffffffffu, b == best) >> (grp * RF)) & ((1u << RF) - 1u);
        const int src = __ffs(eq) - 1;
        float runner = (sub == src) ? rn : b;
        if (__popc(eq) > 1 && best < INF) runner = best;
#pragma unroll
        for (int o = RF / 2; o > 0; o >>= 1) runner = fminf(runner, __shfl_xor_sync(0xffffffffu, runner, o));
        const uint32_t chunk = __shfl_sync(0xffffffffu, ch, grp * RF + max(src, 0));

        // ---- rigorous error bound, evaluated in FP32 with every operation rounded UP (all terms are
        //      non-negative, so the result is >= the real-valued formula of DESIGN.md §K1) --------------
        const float lx = __fsub_rn(hx, cx), ly = __fsub_rn(hy, cy), lz = __fsub_rn(hz, cz);   // p_loc, as in the filter
        const float P2 = __fmaf_ru(lz, lz, __fmaf_ru(ly, ly, __fmul_ru(lx, lx)));
        const float P = __fsqrt_ru(P2);
        const float H = __fsqrt_ru(__fmaf_ru(hz, hz, __fmaf_ru(hy, hy, __fmul_ru(hx, hx))));
        const float U65 = 3.8743021e-7f;                              // > 6.5 * 2^-24
        const float U101 = 6.0201e-8f;                                 // > 1.01 * 2^-24
        const float D0 = fmaxf(__fadd_ru(best, P2), 0.0f);
        float r = __fsqrt_ru(D0);
        float M = __fadd_ru(P, r);
        float E = __fmul_ru(U65, __fmaf_ru(M, M, __fmul_ru(__fmul_ru(2.0f, P), M)));
        r = __fmul_ru(__fsqrt_ru(__fadd_ru(D0, E)), 1.000001f);
        M = __fadd_ru(P, r);
        E = __fmul_ru(U65, __fmaf_ru(M, M, __fmul_ru(__fmul_ru(2.0f, P), M)));
        const float eta = __fmul_ru(U101, __fadd_ru(__fadd_ru(H, Rm), __fadd_ru(__fmul_ru(2.0f, P), r)));
        float tau = __fmul_ru(2.0f, E);
        tau = __fmaf_ru(__fmul_ru(4.0f, eta), r, tau);
        tau = __fmaf_ru(__fmul_ru(4.0f, eta), eta, tau);
        tau = __fmaf_ru(__fmul_ru(9.1e-13f, r), r, tau);
        tau = fmaxf(tau, 1.0e-37f);                                   // never let the margin underflow to zero
        const float thr = __fadd_ru(best, tau);                       // NaN / +inf: everything is a candidate
        const bool clear = runner > thr && thr < INF;                 // false for NaN/inf oddities too

        if (!scene_point_finite(px, py, pz)) {        // uniform per group: the reference's answer is index 0
            if (sub == 0 && valid) store_match(mo, m64, j, 0);
        } else if (clear) {
            // ---- best chunk: FP32 surrogate against thr, FP64 only for the survivors ---------------
            const float ax = -2.0f * lx, ay = -2.0f * ly, az = -2.0f * lz;
            const int k0 = (int)chunk * NN_G;
            unsigned cand = 0u;
#pragma unroll
            for (int i = 0; i < NN_G / RF; ++i) {    // mhat is padded to a multiple of NN_TM >= k0 + NN_G
                const float d = surrogate32(mhat[k0 + i * RF + sub], cx, cy, cz, ax, ay, az);
                if (!(d > thr) && k0 + i * RF + sub < nm) cand |= 1u << i;
            }
            Best64 bt;
            bt.init();
            while (cand) {                           // ascending i = ascending model index
                const int i = __ffs(cand) - 1;
                cand &= cand - 1u;
                int kk = k0 + i * RF + sub;
                ICP_CHK_IDX(kk, nm, ctl);
                double mx, my, mz;
                ld_pt64(m64, kk, mx, my, mz);
                bt.feed(dist2_64(px, py, pz, mx, my, mz), kk);
            }
            double v = bt.v;
            int k = bt.k;
            group_argmin<RF>(v, k, gmask);             // `clear` is uniform per group, not per warp
            if (sub == 0 && valid) store_match(mo, m64, j, first_min_or_zero(k));
        } else if (sub == 0 && valid) {
            if (inline_exact) {
                const int pos = atomicAdd(&s_cnt, 1);
                s_j[pos] = j;
                s_thr[pos] = thr;
            } else {
                const int pos = atomicAdd(&ctl[CTL_AMB_COUNT], 1);
                amb_list[pos] = j;
                amb_thr[pos] = thr;
            }
        }

        if (inline_exact) {
            // ---- the block rescans the model for the points of this trip it could not decide ----
            __syncthreads();
            const int nq = s_cnt;                    // block-uniform
            for (int a = 0; a < nq; ++a) {
                const int jq = s_j[a];
                const int k = pruned_rescan(mhat, m64, nm, nm_pad, phx, phy, phz, p64x, p64y, p64z, part_best, np_pad, n_tiles,
                                            cpt, n_cta, jq, s_thr[a], sp, sv, sk);
                if (threadIdx.x == 0) store_match(mo, m64, jq, first_min_or_zero(k));
            }
            if (threadIdx.x == 0) undecided += nq;
            __syncthreads();                         // s_j / s_cnt are rewritten by the next trip
        }
    }
    if (inline_exact && threadIdx.x == 0 && undecided > 0) atomicAdd(&ctl[CTL_AMB_TOTAL], undecided);
}

// nn_exact (long slot ranges only, see nn_refine): the queued scene points, one per block at a time (grid-stride over
// the queue), each rescanned with pruned_rescan.  Round 1 rescanned the whole model for batches of four queued points
// cut into model segments with a ticket merge; with the slot pruning a point needs 1-2 ranges of <= 64 K model points,
// and the segments, tickets and merge scratch are gone.
constexpr int EXACT_THREADS = SCAN_THREADS;

__global__ void __launch_bounds__(EXACT_THREADS)
nn_exact_kernel(const Pt64* __restrict__ m64, const float4* __restrict__ mhat, int nm, int nm_pad,
                const double* __restrict__ p64x,
                const double* __restrict__ p64y, const double* __restrict__ p64z, const float* __restrict__ phx,
                const float* __restrict__ phy, const float* __restrict__ phz, const float* __restrict__ part_best,
                int np_pad, int n_tiles, int cpt, int n_cta, const int32_t* __restrict__ amb_list,
                const float* __restrict__ amb_thr, MatchOut mo, int* __restrict__ ctl,
                int honor_done, unsigned long long* __restrict__ trace)
{
    if (honor_done && ctl[CTL_DONE]) return;
    __shared__ double sp[1][3];
    __shared__ double sv[1][EXACT_THREADS / 32];
    __shared__ int sk[1][EXACT_THREADS / 32];
    const int count = ctl[CTL_AMB_COUNT];
    for (int a = blockIdx.x; a < count; a += gridDim.x) {
        const int j = amb_list[a];
        const int k = pruned_rescan(mhat, m64, nm, nm_pad, phx, phy, phz, p64x, p64y, p64z, part_best, np_pad, n_tiles, cpt,
                                    n_cta, j, amb_thr[a], sp, sv, sk);
        if (threadIdx.x == 0) store_match(mo, m64, j, first_min_or_zero(k));
    }
    if (threadIdx.x == 0) {
        __threadfence();
        if (atomicAdd(&ctl[CTL_TICKET_NN], 1) == (int)gridDim.x - 1) {
            ctl[CTL_AMB_TOTAL] += count;
            ctl[CTL_AMB_COUNT] = 0;
            ctl[CTL_TICKET_NN] = 0;
            if (trace) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(trace[8]));   // diagnostic: last block of K1 done
        }
    }
}

}  // namespace

void launch_nn_filter(const ModelDev& m, const SceneDev& s, const NNScratch& sc, const int* ctl, int honor_done,
                      cudaStream_t st, int* progress)
{
    static bool attr_set = false;
    if (!attr_set) {
        cudaFuncSetAttribute(nn_filter_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)FILTER_SMEM);
        attr_set = true;
    }
    nn_filter_kernel<<<sc.n_cta, NN_THREADS, FILTER_SMEM, st>>>(m.mhat, s.phx, s.phy, s.phz, (int)s.np_pad,
                                                                (int)(s.np_pad / NN_TILE), (int)(m.nm_pad / NN_G), sc.best,
                                                                sc.runner, sc.chunk, ctl, honor_done, progress);
}

void launch_nn_refine(const ModelDev& m, const SceneDev& s, const NNScratch& sc, int* ctl, int honor_done,
                      int inline_exact, cudaStream_t st)
{
    int64_t grid = (s.np + RF_PPB - 1) / RF_PPB;
    if (grid > 148 * 16) grid = 148 * 16;
    if (grid < 1) grid = 1;
    nn_refine_kernel<<<(int)grid, REFINE_THREADS, 0, st>>>(m.m64, m.mhat, (int)m.nm, (int)m.nm_pad, m.geo, s.p64x, s.p64y,
                                                           s.p64z, s.phx, s.phy, s.phz, (int)s.np, (int)s.np_pad,
                                                           (int)(s.np_pad / NN_TILE), (int)(m.nm_pad / NN_G), sc.n_cta,
                                                           sc.best, sc.runner, sc.chunk, match_out(s), sc.amb_list,
                                                           sc.amb_thr, ctl, honor_done, inline_exact);
}

void launch_nn_exact(const ModelDev& m, const SceneDev& s, const NNScratch& sc, int* ctl, int honor_done, int sm_count,
                     cudaStream_t st, unsigned long long* trace)
{
    const int grid = sm_count * 4;
    nn_exact_kernel<<<grid, EXACT_THREADS, 0, st>>>(m.m64, m.mhat, (int)m.nm, (int)m.nm_pad, s.p64x,
                                                    s.p64y, s.p64z, s.phx, s.phy, s.phz, sc.best, (int)s.np_pad,
                                                    (int)(s.np_pad / NN_TILE), (int)(m.nm_pad / NN_G), sc.n_cta,
                                                    sc.amb_list, sc.amb_thr, match_out(s), ctl, honor_done, trace);
}

}  // namespace icpb
