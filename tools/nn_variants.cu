// nn_variants.cu — throughput study of the nearest-neighbour filter inner loop on sm_100a.
// Standalone (not part of the product): same tiling as csrc/nn_kernels.cu, plain smem staging, and
// compile-time variants of the FMA / min instruction mix.  Prints T pairs/s per variant.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o nn_variants nn_variants.cu
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s line %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)

constexpr int TM = 512, G = 64;
__device__ __forceinline__ float min3(float a, float b, float c) { float r; asm("min.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c)); return r; }

// FMA: 0 = FFMA2, 1 = scalar FFMA.   MIN: 0 = FMNMX3, 1 = 2x FMNMX, 2 = FMNMX3 on every 4th step only,
// 3 = tree: one FMNMX3 + one FMNMX3 (independent then merged)
template <int THREADS, int R, int FMA, int MIN, int MINB>
__global__ void __launch_bounds__(THREADS, MINB)
variant_kernel(const float4* __restrict__ mhat, const float* __restrict__ px, const float* __restrict__ py,
               const float* __restrict__ pz, int n_splits, int tiles_per_split, int nm_tiles, int np_pad,
               float* __restrict__ out_best, unsigned* __restrict__ out_chunk)
{
    extern __shared__ __align__(16) float soa[];   // 2 x 4 x TM
    const int tid = threadIdx.x;
    const int tile = blockIdx.x / n_splits, split = blockIdx.x - tile * n_splits;
    const int mt0 = split * tiles_per_split, mt1 = min(mt0 + tiles_per_split, nm_tiles);
    const int pbase = tile * THREADS * R;
    const float cx = px[pbase], cy = py[pbase], cz = pz[pbase];
    float ax[R], ay[R], az[R], best[R]; unsigned bch[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const int j = pbase + r * THREADS + tid;
        ax[r] = -2.f * (px[j] - cx); ay[r] = -2.f * (py[j] - cy); az[r] = -2.f * (pz[j] - cz);
        best[r] = __int_as_float(0x7f800000); bch[r] = 0;
    }
    for (int mt = mt0; mt < mt1; ++mt) {
        float* buf = soa + ((mt - mt0) & 1) * 4 * TM;
        for (int i = tid; i < TM; i += THREADS) {
            const float4 v = mhat[(size_t)mt * TM + i];
            const float x = v.x - cx, y = v.y - cy, z = v.z - cz;
            buf[i] = x; buf[TM + i] = y; buf[2 * TM + i] = z; buf[3 * TM + i] = fmaf(z, z, fmaf(y, y, x * x));
        }
        __syncthreads();
        const float4* X4 = reinterpret_cast<const float4*>(buf);
        const float4* Y4 = X4 + TM / 4; const float4* Z4 = Y4 + TM / 4; const float4* W4 = Z4 + TM / 4;
#pragma unroll 1
        for (int c = 0; c < TM / G; ++c) {
            float tmin[R]; float2 tsum[R];
#pragma unroll
            for (int r = 0; r < R; ++r) { tmin[r] = __int_as_float(0x7f800000); tsum[r] = make_float2(0.f, 0.f); }
#pragma unroll 2
            for (int q = 0; q < G / 4; ++q) {
                const int o = c * (G / 4) + q;
                const float4 X = X4[o], Y = Y4[o], Z = Z4[o], W = W4[o];
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    float d0, d1, d2, d3;
                    if (FMA == 0) {
                        const float2 a_x = make_float2(ax[r], ax[r]), a_y = make_float2(ay[r], ay[r]), a_z = make_float2(az[r], az[r]);
                        float2 t01 = __ffma2_rn(a_z, make_float2(Z.x, Z.y), make_float2(W.x, W.y));
                        float2 t23 = __ffma2_rn(a_z, make_float2(Z.z, Z.w), make_float2(W.z, W.w));
                        t01 = __ffma2_rn(a_y, make_float2(Y.x, Y.y), t01);
                        t23 = __ffma2_rn(a_y, make_float2(Y.z, Y.w), t23);
                        t01 = __ffma2_rn(a_x, make_float2(X.x, X.y), t01);
                        t23 = __ffma2_rn(a_x, make_float2(X.z, X.w), t23);
                        d0 = t01.x; d1 = t01.y; d2 = t23.x; d3 = t23.y;
                    } else {
                        d0 = fmaf(ax[r], X.x, fmaf(ay[r], Y.x, fmaf(az[r], Z.x, W.x)));
                        d1 = fmaf(ax[r], X.y, fmaf(ay[r], Y.y, fmaf(az[r], Z.y, W.y)));
                        d2 = fmaf(ax[r], X.z, fmaf(ay[r], Y.z, fmaf(az[r], Z.z, W.z)));
                        d3 = fmaf(ax[r], X.w, fmaf(ay[r], Y.w, fmaf(az[r], Z.w, W.w)));
                    }
                    if (MIN == 0) { tmin[r] = min3(tmin[r], d0, d1); tmin[r] = min3(tmin[r], d2, d3); }
                    else if (MIN == 1) { tmin[r] = fminf(tmin[r], d0); tmin[r] = fminf(tmin[r], d1); tmin[r] = fminf(tmin[r], d2); tmin[r] = fminf(tmin[r], d3); }
                    else if (MIN == 2) { if ((q & 3) == 0) { tmin[r] = min3(tmin[r], d0, d1); tmin[r] = min3(tmin[r], d2, d3); } else { tmin[r] = fminf(tmin[r], d0 + d1 + d2 + d3); } }
                    else if (MIN == 3) { const float u = min3(d0, d1, d2); tmin[r] = min3(tmin[r], u, d3); }
                    else if (MIN == 4) { tmin[r] += d0; tsum[r] = __ffma2_rn(make_float2(d1, d2), make_float2(1.0009765625f, 1.0009765625f), tsum[r]); tmin[r] += d3; }   // no min at all: 4 extra FMA-pipe lane-ops / 4 pairs
                    else if (MIN == 5) { tmin[r] = min3(tmin[r], d0, d1); tsum[r] = __ffma2_rn(make_float2(d2, d3), make_float2(1.0009765625f, 1.0009765625f), tsum[r]); }
                    else if (MIN == 6) { tsum[r] = __ffma2_rn(make_float2(d0, d1), make_float2(1.0009765625f, 1.0009765625f), tsum[r]); tsum[r] = __ffma2_rn(make_float2(d2, d3), make_float2(1.0009765625f, 1.0009765625f), tsum[r]); }
                }
            }
            const unsigned cid = (unsigned)mt * (TM / G) + c;
#pragma unroll
            for (int r = 0; r < R; ++r) { if (MIN >= 4) tmin[r] += tsum[r].x + tsum[r].y; if (tmin[r] < best[r]) { best[r] = tmin[r]; bch[r] = cid; } }
        }
    }
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const size_t j = (size_t)split * np_pad + pbase + r * THREADS + tid;
        out_best[j] = best[r]; out_chunk[j] = bch[r];
    }
}

template <int THREADS, int R, int FMA, int MIN, int MINB>
void run(const char* name, int n, int sms, const float4* mhat, const float* px, const float* py, const float* pz,
         float* ob, unsigned* oc)
{
    const int tile = THREADS * R, n_tiles = n / tile, nm_tiles = n / TM;
    // splits: aim at >= 4 waves-worth of CTAs
    int S = 1;
    while ((long)n_tiles * S < (long)sms * MINB * 2 && S < nm_tiles) S *= 2;
    const int tps = (nm_tiles + S - 1) / S;
    const size_t smem = 2 * 4 * TM * sizeof(float);
    auto k = variant_kernel<THREADS, R, FMA, MIN, MINB>;
    CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaFuncAttributes at; CK(cudaFuncGetAttributes(&at, k));
    int occ = 0; CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k, THREADS, smem));
    k<<<n_tiles * S, THREADS, smem>>>(mhat, px, py, pz, S, tps, nm_tiles, n, ob, oc);
    CK(cudaDeviceSynchronize());
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    float best_ms = 1e9f;
    for (int rep = 0; rep < 3; ++rep) {
        CK(cudaEventRecord(e0));
        k<<<n_tiles * S, THREADS, smem>>>(mhat, px, py, pz, S, tps, nm_tiles, n, ob, oc);
        CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); if (ms < best_ms) best_ms = ms;
    }
    const double pairs = (double)n * n;
    printf("%-44s regs=%3d occ=%d grid=%5d  %8.3f ms  %6.2f T pairs/s  (%5.1f%% of 8-flop roofline)\n", name,
           at.numRegs, occ, n_tiles * S, best_ms, pairs / best_ms / 1e9, pairs / best_ms / 1e9 / 9.306 * 100);
}

int main(int argc, char** argv)
{
    const int lg = argc > 1 ? atoi(argv[1]) : 18;
    const int n = 1 << lg;
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
    const int sms = p.multiProcessorCount;
    std::vector<float4> hm(n); std::vector<float> hx(n), hy(n), hz(n);
    srand(1);
    auto rnd = []() { return 2.f * rand() / RAND_MAX - 1.f; };
    for (int i = 0; i < n; ++i) { hm[i] = make_float4(rnd(), rnd(), rnd(), 0.f); hx[i] = rnd(); hy[i] = rnd(); hz[i] = rnd(); }
    float4* mhat; float *px, *py, *pz, *ob; unsigned* oc;
    CK(cudaMalloc(&mhat, n * sizeof(float4))); CK(cudaMalloc(&px, n * 4)); CK(cudaMalloc(&py, n * 4)); CK(cudaMalloc(&pz, n * 4));
    CK(cudaMalloc(&ob, (size_t)n * 64 * 4)); CK(cudaMalloc(&oc, (size_t)n * 64 * 4));
    CK(cudaMemcpy(mhat, hm.data(), n * sizeof(float4), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(px, hx.data(), n * 4, cudaMemcpyHostToDevice)); CK(cudaMemcpy(py, hy.data(), n * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(pz, hz.data(), n * 4, cudaMemcpyHostToDevice));
    printf("n = 2^%d, %d SMs\n", lg, sms);
#define RUN(T, R, F, M, B) run<T, R, F, M, B>("T=" #T " R=" #R " FMA=" #F " MIN=" #M " minB=" #B, n, sms, mhat, px, py, pz, ob, oc)
    RUN(128, 8, 0, 0, 4);   // product configuration
    RUN(128, 8, 0, 6, 4);   // no min: 4 FFMA2 per 2 pairs
    RUN(128, 8, 0, 5, 4);   // half the mins
    RUN(128, 8, 0, 4, 4);
    RUN(128, 8, 0, 1, 4);
    RUN(128, 8, 0, 2, 4);
    RUN(128, 8, 0, 3, 4);
    RUN(128, 8, 1, 0, 4);
    RUN(128, 8, 1, 1, 4);
    RUN(128, 8, 1, 2, 4);
    RUN(128, 4, 0, 0, 8);
    RUN(128, 4, 1, 0, 8);
    RUN(128, 6, 0, 0, 4);
    RUN(128, 12, 0, 0, 2);
    RUN(128, 12, 1, 0, 2);
    RUN(128, 16, 0, 0, 2);
    RUN(256, 8, 0, 0, 2);
    RUN(256, 4, 0, 0, 4);
    RUN(64, 8, 0, 0, 8);
    RUN(128, 8, 0, 0, 3);
    RUN(128, 8, 0, 0, 2);
    return 0;
}
