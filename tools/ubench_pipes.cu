// Pipe-throughput microbenchmark for sm_100a: measures FFMA, FFMA2 (fma.rn.f32x2), FMNMX, FMNMX3,
// the candidate nearest-neighbour inner-loop mixes and FP64 DFMA, in lane-ops / clk / SM.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ubench_pipes ubench_pipes.cu
// Not part of the product path; evidence for DESIGN.md's roofline discussion.
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)

typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float a, float b) { float2 v = make_float2(a, b); return *reinterpret_cast<u64*>(&v); }
__device__ __forceinline__ float2 upk(u64 v) { return *reinterpret_cast<float2*>(&v); }

constexpr int NCH = 16;    // independent chains per thread
constexpr int UNR = 8;

template <int MODE>
__global__ void __launch_bounds__(256) bench(float* out, int iters, float a, float b, long long* clk)
{
    float x[NCH]; u64 xx[NCH]; double dx[NCH];
#pragma unroll
    for (int i = 0; i < NCH; ++i) { x[i] = threadIdx.x * 1e-3f + i; xx[i] = pk(x[i], x[i] + 0.5f); dx[i] = x[i]; }
    u64 aa = pk(a, a * 1.0001f), bb = pk(b, b * 1.0001f);
    float y = b;
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < UNR; ++u) {
            if (MODE == 0) {         // FFMA
#pragma unroll
                for (int i = 0; i < NCH; ++i) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(x[i]) : "f"(a), "f"(b));
            } else if (MODE == 1) {  // FFMA2
#pragma unroll
                for (int i = 0; i < NCH; ++i) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(xx[i]) : "l"(aa), "l"(bb));
            } else if (MODE == 2) {  // FMNMX
#pragma unroll
                for (int i = 0; i < NCH; ++i) asm volatile("min.f32 %0, %0, %1;" : "+f"(x[i]) : "f"(y));
            } else if (MODE == 3) {  // FMNMX3
#pragma unroll
                for (int i = 0; i < NCH; ++i) asm volatile("min.f32 %0, %0, %1, %2;" : "+f"(x[i]) : "f"(y), "f"(a));
            } else if (MODE == 4) {  // scalar pair mix: 3 FFMA + 1 FMNMX  (NCH/2 pairs in flight)
#pragma unroll
                for (int i = 0; i < NCH / 2; ++i) {
                    float t;
                    asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(t) : "f"(x[2 * i]), "f"(a), "f"(b));
                    asm volatile("fma.rn.f32 %0, %1, %2, %0;" : "+f"(t) : "f"(x[2 * i]), "f"(b));
                    asm volatile("fma.rn.f32 %0, %1, %2, %0;" : "+f"(t) : "f"(x[2 * i]), "f"(y));
                    asm volatile("min.f32 %0, %0, %1;" : "+f"(x[2 * i + 1]) : "f"(t));
                }
            } else if (MODE == 5) {  // packed mix: 3 FFMA2 + 1 FMNMX3 per 2 pairs
#pragma unroll
                for (int i = 0; i < NCH / 2; ++i) {
                    u64 t;
                    asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(t) : "l"(xx[2 * i]), "l"(aa), "l"(bb));
                    asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(t) : "l"(xx[2 * i]), "l"(bb));
                    asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(t) : "l"(xx[2 * i]), "l"(aa));
                    float2 tt = upk(t);
                    asm volatile("min.f32 %0, %0, %1, %2;" : "+f"(x[2 * i + 1]) : "f"(tt.x), "f"(tt.y));
                }
            } else if (MODE == 6) {  // packed mix: 3 FFMA2 + 2 FMNMX per 2 pairs
#pragma unroll
                for (int i = 0; i < NCH / 2; ++i) {
                    u64 t;
                    asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(t) : "l"(xx[2 * i]), "l"(aa), "l"(bb));
                    asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(t) : "l"(xx[2 * i]), "l"(bb));
                    asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(t) : "l"(xx[2 * i]), "l"(aa));
                    float2 tt = upk(t);
                    asm volatile("min.f32 %0, %0, %1;" : "+f"(x[2 * i + 1]) : "f"(tt.x));
                    asm volatile("min.f32 %0, %0, %1;" : "+f"(x[2 * i]) : "f"(tt.y));
                }
            } else if (MODE == 7) {  // DFMA
#pragma unroll
                for (int i = 0; i < NCH; ++i) asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(dx[i]) : "d"((double)a), "d"((double)b));
            } else if (MODE == 8) {  // scalar mix with FMNMX3: 6 FFMA + 1 FMNMX3 per 2 pairs
#pragma unroll
                for (int i = 0; i < NCH / 2; ++i) {
                    float t, s;
                    asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(t) : "f"(x[2 * i]), "f"(a), "f"(b));
                    asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(s) : "f"(x[2 * i]), "f"(b), "f"(a));
                    asm volatile("fma.rn.f32 %0, %1, %2, %0;" : "+f"(t) : "f"(x[2 * i]), "f"(b));
                    asm volatile("fma.rn.f32 %0, %1, %2, %0;" : "+f"(s) : "f"(x[2 * i]), "f"(a));
                    asm volatile("fma.rn.f32 %0, %1, %2, %0;" : "+f"(t) : "f"(x[2 * i]), "f"(y));
                    asm volatile("fma.rn.f32 %0, %1, %2, %0;" : "+f"(s) : "f"(x[2 * i]), "f"(y));
                    asm volatile("min.f32 %0, %0, %1, %2;" : "+f"(x[2 * i + 1]) : "f"(t), "f"(s));
                }
            }
        }
        y += 1e-6f;
    }
    long long t1 = clock64();
    float acc = 0.f;
#pragma unroll
    for (int i = 0; i < NCH; ++i) { float2 t = upk(xx[i]); acc += x[i] + t.x + t.y + (float)dx[i]; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    if (threadIdx.x == 0 && blockIdx.x == 0) clk[0] = t1 - t0;
}

struct Case { const char* name; int mode; double ops_per_inner; const char* unit; };

template <int MODE>
void run(const Case& c, int sms, int ctas_per_sm, float* out, long long* clk)
{
    int iters = (MODE == 7) ? 2000 : 20000;
    int grid = sms * ctas_per_sm;
    bench<MODE><<<grid, 256>>>(out, 100, 1.0001f, 0.5f, clk);
    CK(cudaDeviceSynchronize());
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    CK(cudaEventRecord(e0));
    bench<MODE><<<grid, 256>>>(out, iters, 1.0001f, 0.5f, clk);
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
    long long cycles; CK(cudaMemcpy(&cycles, clk, sizeof(cycles), cudaMemcpyDeviceToHost));
    double inner = (double)iters * UNR;                      // inner bodies per thread
    double lane_ops = inner * c.ops_per_inner * 256.0 * ctas_per_sm;  // per SM
    double mhz = cycles / (ms * 1e3);
    printf("%-34s ctas/sm=%d  %.3f ms  clk(block0)=%lld (~%.0f MHz)  %s/clk/SM = %.1f   (%.2f T%s/s chip)\n",
           c.name, ctas_per_sm, ms, cycles, mhz, c.unit, lane_ops / cycles,
           lane_ops * sms / (ms * 1e-3) / 1e12, c.unit);
}

int main()
{
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
    int sms = p.multiProcessorCount;
    printf("device %s, %d SMs, clockRate %d kHz\n", p.name, sms, p.clockRate);
    float* out; long long* clk;
    CK(cudaMalloc(&out, sizeof(float) * sms * 8 * 256)); CK(cudaMalloc(&clk, 8));
    for (int cps : {2, 4}) {
        run<0>({"FFMA (fma lane-ops)", 0, NCH, "fma"}, sms, cps, out, clk);
        run<1>({"FFMA2 (fma lane-ops, 2/instr)", 1, NCH * 2, "fma"}, sms, cps, out, clk);
        run<2>({"FMNMX (min instr)", 2, NCH, "min"}, sms, cps, out, clk);
        run<3>({"FMNMX3 (min3 instr)", 3, NCH, "min3"}, sms, cps, out, clk);
        run<4>({"mix 3FFMA+1FMNMX (pairs)", 4, NCH / 2, "pair"}, sms, cps, out, clk);
        run<8>({"mix 6FFMA+1FMNMX3 (pairs)", 8, NCH, "pair"}, sms, cps, out, clk);
        run<5>({"mix 3FFMA2+1FMNMX3 (pairs)", 5, NCH, "pair"}, sms, cps, out, clk);
        run<6>({"mix 3FFMA2+2FMNMX (pairs)", 6, NCH, "pair"}, sms, cps, out, clk);
        run<7>({"DFMA (fp64 fma lane-ops)", 7, NCH, "dfma"}, sms, cps, out, clk);
    }
    return 0;
}
