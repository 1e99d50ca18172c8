// nn_kernels.cu — K1: exact brute-force nearest-neighbour correspondence for sm_100a.
//
// Replaces CPU::closest_matrix (reference src/cpu.cc:5-27) and the GPU pair compute_distance + find_Y
// (src/GPU/compute.cu:94-150) — neither is ported.  The reference materialises an np x nm FP64
// distance matrix in HBM and scans it with one thread per scene point; here nothing is
// materialised: model tiles stream through shared memory and the running minimum stays in
// registers.  Three kernels:
//
//  nn_filter  FP32, CUDA cores, FMA-pipe bound.  Work item = (scene tile of 1024 points, model split).
//             Model tiles (512 points) arrive by TMA bulk copy (cp.async.bulk + mbarrier) as float4,
//             are re-expressed relative to the CTA's own origin and staged as SoA float4 planes
//             X|Y|Z|W (W = |m|^2).  Each thread keeps 8 scene points in registers (coordinates
//             pre-scaled by -2 and duplicated into f32x2 pairs) and evaluates, per LDS.128 of four
//             consecutive model points,   d' = |m|^2 - 2 p.m   for 8 x 4 pairs with 48 FFMA2
//             (fma.rn.f32x2) and 16 FMNMX3 (3-input min): 3 FMA-pipe lane-cycles and 2.1 issue
//             slots per pair.  Only the VALUE of the minimum is tracked per pair; the index is
//             recovered lazily: per 64-point model chunk the chunk minimum is merged into (best,
//             best chunk id, runner-up over all other chunks).
//  nn_refine  one warp per scene point: merges the model splits, derives a rigorous bound tau on
//             the FP32 evaluation error, and if the runner-up chunk is provably farther than tau
//             re-ranks the 64 points of the best chunk in FP64 with the reference's own expression
//             sqrt((dx*dx+dy*dy)+dz*dz), lowest index on ties (warp-shuffle argmin).  Otherwise the
//             point is queued for
//  nn_exact   a full FP64 scan of the model for the (rare) ambiguous points.
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
__global__ void __launch_bounds__(NN_THREADS, 4)
nn_filter_kernel(const float4* __restrict__ mhat, const float* __restrict__ phx, const float* __restrict__ phy,
                 const float* __restrict__ phz, int np_pad, int n_splits, int tiles_per_split, int nm_tiles,
                 float* __restrict__ out_best, float* __restrict__ out_runner, uint32_t* __restrict__ out_chunk,
                 const int* __restrict__ ctl, int honor_done)
{
    if (honor_done && ctl[CTL_DONE]) return;

    extern __shared__ __align__(128) unsigned char smem[];
    float4* raw = reinterpret_cast<float4*>(smem);
    float* soa = reinterpret_cast<float*>(smem + NN_TM * sizeof(float4));
    __shared__ __align__(8) uint64_t mbar;

    const int tid = threadIdx.x;
    const int tile = blockIdx.x / n_splits;
    const int split = blockIdx.x - tile * n_splits;
    const int mt0 = split * tiles_per_split;
    const int mt1 = min(mt0 + tiles_per_split, nm_tiles);
    const int pbase = tile * NN_TILE;

    if (tid == 0) {
        mbar_init(&mbar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (tid == 0 && mt0 < mt1) {
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

    uint32_t parity = 0;
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

        // ---- compute: 8 scene points x 512 model points per thread --------------------------------
        const float4* X4 = reinterpret_cast<const float4*>(buf);
        const float4* Y4 = reinterpret_cast<const float4*>(buf + SOA_PLANE);
        const float4* Z4 = reinterpret_cast<const float4*>(buf + 2 * SOA_PLANE);
        const float4* W4 = reinterpret_cast<const float4*>(buf + 3 * SOA_PLANE);
#pragma unroll 1
        for (int c = 0; c < NN_TM / NN_G; ++c) {
            float tmin[NN_R];
#pragma unroll
            for (int r = 0; r < NN_R; ++r) tmin[r] = __int_as_float(0x7f800000);
#pragma unroll 2
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
            const uint32_t cid = (uint32_t)mt * (NN_TM / NN_G) + (uint32_t)c;
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

    const size_t obase = (size_t)split * np_pad + pbase;
#pragma unroll
    for (int r = 0; r < NN_R; ++r) {
        const size_t j = obase + r * NN_THREADS + tid;
        out_best[j] = best[r];
        out_runner[j] = runner[r];
        out_chunk[j] = bchunk[r];
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
struct Best64 {
    double s, v;
    int k;
    __device__ __forceinline__ void init() { s = v = __longlong_as_double(0x7ff0000000000000LL); k = 0x7fffffff; }
    __device__ __forceinline__ void feed(double s_new, int k_new)   // k_new > k of everything fed before
    {
        if (s_new < s) {
            const double v_new = __dsqrt_rn(s_new);
            if (v_new < v) { v = v_new; k = k_new; }
            s = s_new;   // equal sqrt: the earlier index stays, the smaller s keeps filtering
        }
    }
};

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

constexpr int REFINE_THREADS = 256;
constexpr int RF = 8;    // lanes cooperating on one scene point

__global__ void __launch_bounds__(REFINE_THREADS)
nn_refine_kernel(const double* __restrict__ m64x, const double* __restrict__ m64y, const double* __restrict__ m64z,
                 int nm, const double* __restrict__ geo, const double* __restrict__ p64x,
                 const double* __restrict__ p64y, const double* __restrict__ p64z, const float* __restrict__ phx,
                 const float* __restrict__ phy, const float* __restrict__ phz, int np, int np_pad, int n_splits,
                 const float* __restrict__ part_best, const float* __restrict__ part_runner,
                 const uint32_t* __restrict__ part_chunk, int32_t* __restrict__ idx, int32_t* __restrict__ amb_list,
                 int* __restrict__ ctl, int honor_done)
{
    if (honor_done && ctl[CTL_DONE]) return;
    const int lane = threadIdx.x & 31, sub = lane & (RF - 1), grp = lane / RF;
    const int warp = (blockIdx.x * REFINE_THREADS + threadIdx.x) >> 5;
    const int n_warps = (gridDim.x * REFINE_THREADS) >> 5;
    constexpr int PPW = 32 / RF;                       // scene points per warp
    const double Rm = geo[3];
    const float INF = __int_as_float(0x7f800000);

    for (int jb = warp * PPW; jb < np; jb += n_warps * PPW) {
        const bool valid = jb + grp < np;
        const int j = valid ? jb + grp : np - 1;
        // ---- merge the model splits (RF lanes stride over the splits) ------------------------------
        float b = INF, rn = INF;
        uint32_t ch = 0;
        for (int s = sub; s < n_splits; s += RF) {
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
        const uint32_t chunk = __shfl_sync(0xffffffffu, ch, grp * RF + src);

        // ---- rigorous FP32 error bound ------------------------------------------------------------
        const int cj = (j / NN_TILE) * NN_TILE + NN_TILE / 2;
        const float hx = phx[j], hy = phy[j], hz = phz[j];
        const float lx = __fsub_rn(hx, phx[cj]), ly = __fsub_rn(hy, phy[cj]), lz = __fsub_rn(hz, phz[cj]);
        const double P2 = (double)lx * lx + (double)ly * ly + (double)lz * lz;
        const double P = sqrt(P2);
        const double H = sqrt((double)hx * hx + (double)hy * hy + (double)hz * hz);
        const double u = 5.9604644775390625e-8;                       // 2^-24
        const double D0 = fmax((double)best + P2, 0.0);
        double r = sqrt(D0);
        double M = P + r;
        double E = 6.5 * u * (M * M + 2.0 * P * M);
        r = sqrt(D0 + E) * (1.0 + 1e-6);
        M = P + r;
        E = 6.5 * u * (M * M + 2.0 * P * M);
        const double eta = 1.01 * u * (H + Rm + 2.0 * P + r);
        const double tau = 2.0 * E + 4.0 * eta * r + 4.0 * eta * eta + 9.1e-13 * r * r;
        const bool clear = (double)runner > (double)best + tau;       // false for NaN/inf oddities too

        if (clear) {
            // ---- FP64 re-rank of the best chunk with the reference's expression -------------------
            const double px = p64x[j], py = p64y[j], pz = p64z[j];
            const int k0 = (int)chunk * NN_G;
            double s2[NN_G / RF];
#pragma unroll
            for (int i = 0; i < NN_G / RF; ++i) {
                const int kk = min(k0 + i * RF + sub, nm - 1);
                s2[i] = dist2_64(px, py, pz, m64x[kk], m64y[kk], m64z[kk]);
            }
            Best64 bt;
            bt.init();
#pragma unroll
            for (int i = 0; i < NN_G / RF; ++i) {
                const int kk = k0 + i * RF + sub;
                if (kk < nm) bt.feed(s2[i], kk);
            }
            double v = bt.v;
            int k = bt.k;
            group_argmin<RF>(v, k, ((1u << RF) - 1u) << (grp * RF));   // `clear` is uniform per group, not per warp
            if (sub == 0 && valid) idx[j] = k;
        } else if (sub == 0 && valid) {
            const int pos = atomicAdd(&ctl[CTL_AMB_COUNT], 1);
            amb_list[pos] = j;
        }
    }
}

// Full FP64 scan for the scene points the filter could not decide.  One CTA per queued point; the
// last CTA to finish folds the pass's counters into the totals.
constexpr int EXACT_THREADS = 256;

__global__ void __launch_bounds__(EXACT_THREADS)
nn_exact_kernel(const double* __restrict__ m64x, const double* __restrict__ m64y, const double* __restrict__ m64z,
                int nm, const double* __restrict__ p64x, const double* __restrict__ p64y,
                const double* __restrict__ p64z, const int32_t* __restrict__ amb_list, int32_t* __restrict__ idx,
                int* __restrict__ ctl, int np, int honor_done)
{
    if (honor_done && ctl[CTL_DONE]) return;
    __shared__ double sv[EXACT_THREADS / 32];
    __shared__ int sk[EXACT_THREADS / 32];
    const int count = ctl[CTL_AMB_COUNT];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (int a = blockIdx.x; a < count; a += gridDim.x) {
        const int j = amb_list[a];
        const double px = p64x[j], py = p64y[j], pz = p64z[j];
        Best64 bt;
        bt.init();
        int kk = threadIdx.x;
        for (; kk + 3 * EXACT_THREADS < nm; kk += 4 * EXACT_THREADS) {    // 4 independent evaluations in flight
            const double s0 = dist2_64(px, py, pz, m64x[kk], m64y[kk], m64z[kk]);
            const double s1 = dist2_64(px, py, pz, m64x[kk + EXACT_THREADS], m64y[kk + EXACT_THREADS], m64z[kk + EXACT_THREADS]);
            const double s2 = dist2_64(px, py, pz, m64x[kk + 2 * EXACT_THREADS], m64y[kk + 2 * EXACT_THREADS], m64z[kk + 2 * EXACT_THREADS]);
            const double s3 = dist2_64(px, py, pz, m64x[kk + 3 * EXACT_THREADS], m64y[kk + 3 * EXACT_THREADS], m64z[kk + 3 * EXACT_THREADS]);
            bt.feed(s0, kk); bt.feed(s1, kk + EXACT_THREADS); bt.feed(s2, kk + 2 * EXACT_THREADS); bt.feed(s3, kk + 3 * EXACT_THREADS);
        }
        for (; kk < nm; kk += EXACT_THREADS) bt.feed(dist2_64(px, py, pz, m64x[kk], m64y[kk], m64z[kk]), kk);
        double v = bt.v;
        int k = bt.k;
        group_argmin<32>(v, k);
        if (lane == 0) { sv[w] = v; sk[w] = k; }
        __syncthreads();
        if (w == 0) {
            v = lane < EXACT_THREADS / 32 ? sv[lane] : __longlong_as_double(0x7ff0000000000000LL);
            k = lane < EXACT_THREADS / 32 ? sk[lane] : 0x7fffffff;
            group_argmin<32>(v, k);
            if (lane == 0) idx[j] = k;
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        __threadfence();
        if (atomicAdd(&ctl[CTL_TICKET_NN], 1) == (int)gridDim.x - 1) {
            ctl[CTL_AMB_TOTAL] += count;
            ctl[CTL_CHUNK_TOTAL] += np - count;
            ctl[CTL_AMB_COUNT] = 0;
            ctl[CTL_TICKET_NN] = 0;
        }
    }
}

}  // namespace

void launch_nn_filter(const ModelDev& m, const SceneDev& s, const NNScratch& sc, const int* ctl, int honor_done,
                      cudaStream_t st)
{
    static bool attr_set = false;
    if (!attr_set) {
        cudaFuncSetAttribute(nn_filter_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)FILTER_SMEM);
        attr_set = true;
    }
    const int n_tiles = (int)(s.np_pad / NN_TILE);
    const int nm_tiles = (int)(m.nm_pad / NN_TM);
    const int grid = n_tiles * sc.n_splits;
    nn_filter_kernel<<<grid, NN_THREADS, FILTER_SMEM, st>>>(m.mhat, s.phx, s.phy, s.phz, (int)s.np_pad, sc.n_splits,
                                                            sc.tiles_per_split, nm_tiles, sc.best, sc.runner,
                                                            sc.chunk, ctl, honor_done);
}

void launch_nn_refine(const ModelDev& m, const SceneDev& s, const NNScratch& sc, int* ctl, int honor_done,
                      cudaStream_t st)
{
    const int64_t warps = (s.np + 32 / RF - 1) / (32 / RF);
    int grid = (int)((warps * 32 + REFINE_THREADS - 1) / REFINE_THREADS);
    if (grid > 148 * 16) grid = 148 * 16;
    if (grid < 1) grid = 1;
    nn_refine_kernel<<<grid, REFINE_THREADS, 0, st>>>(m.m64x, m.m64y, m.m64z, (int)m.nm, m.geo, s.p64x, s.p64y, s.p64z,
                                                      s.phx, s.phy, s.phz, (int)s.np, (int)s.np_pad, sc.n_splits,
                                                      sc.best, sc.runner, sc.chunk, s.idx, sc.amb_list, ctl,
                                                      honor_done);
}

void launch_nn_exact(const ModelDev& m, const SceneDev& s, const NNScratch& sc, int* ctl, int honor_done, int sm_count,
                     cudaStream_t st)
{
    int grid = sm_count * 4;
    if ((int64_t)grid > s.np) grid = (int)s.np;
    if (grid < 1) grid = 1;
    nn_exact_kernel<<<grid, EXACT_THREADS, 0, st>>>(m.m64x, m.m64y, m.m64z, (int)m.nm, s.p64x, s.p64y, s.p64z,
                                                    sc.amb_list, s.idx, ctl, (int)s.np, honor_done);
}

}  // namespace icpb
