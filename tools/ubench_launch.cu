// Launch-latency microbenchmark for sm_100a: how much does one more kernel / one grid barrier cost inside the
// ICP iteration?  (a) plain kernel back to back, (b) cooperative launch, (c) cooperative + grid.sync(),
// (d) plain launch + hand-rolled atomic grid barrier (single stream, co-resident grid), (e) programmatic
// dependent launch.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ubench_launch ubench_launch.cu
// Not part of the product path; evidence for DESIGN.md's kernel-count discussion.
#include <cooperative_groups.h>
#include <cstdio>
#include <cuda_runtime.h>
namespace cg = cooperative_groups;
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)

__global__ void k_plain(int* p) { if (threadIdx.x == 0 && blockIdx.x == 0) p[0] += 1; }
__global__ void k_coop(int* p, int syncs)
{
    cg::grid_group g = cg::this_grid();
    for (int i = 0; i < syncs; ++i) g.sync();
    if (threadIdx.x == 0 && blockIdx.x == 0) p[0] += 1;
}
__global__ void k_atomic_barrier(int* p, unsigned* bar, unsigned target)
{
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        atomicAdd(bar, 1u);
        while (*(volatile unsigned*)bar < target) {}
        __threadfence();
    }
    __syncthreads();
    if (threadIdx.x == 0 && blockIdx.x == 0) p[0] += 1;
}
__global__ void k_pdl(int* p)
{
    asm volatile("griddepcontrol.launch_dependents;");
    asm volatile("griddepcontrol.wait;" ::: "memory");
    if (threadIdx.x == 0 && blockIdx.x == 0) p[0] += 1;
}

int main()
{
    int* d; unsigned* bar;
    CK(cudaMalloc(&d, 4)); CK(cudaMalloc(&bar, 4)); CK(cudaMemset(d, 0, 4)); CK(cudaMemset(bar, 0, 4));
    cudaStream_t st; CK(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    const int N = 2000;
    for (int grid : {148, 296, 592}) {
        float ms;
        // (a)
        for (int rep = 0; rep < 2; ++rep) {
            CK(cudaEventRecord(e0, st));
            for (int i = 0; i < N; ++i) k_plain<<<grid, 256, 0, st>>>(d);
            CK(cudaEventRecord(e1, st)); CK(cudaEventSynchronize(e1));
        }
        CK(cudaEventElapsedTime(&ms, e0, e1));
        printf("grid %3d  plain back-to-back            %.2f us/launch\n", grid, ms * 1e3 / N);
        // (b), (c)
        for (int syncs : {0, 1, 3}) {
            void* args[] = {&d, &syncs};
            for (int rep = 0; rep < 2; ++rep) {
                CK(cudaEventRecord(e0, st));
                for (int i = 0; i < N; ++i) CK(cudaLaunchCooperativeKernel((const void*)k_coop, dim3(grid), dim3(256), args, 0, st));
                CK(cudaEventRecord(e1, st)); CK(cudaEventSynchronize(e1));
            }
            CK(cudaEventElapsedTime(&ms, e0, e1));
            printf("grid %3d  cooperative, %d grid.sync      %.2f us/launch\n", grid, syncs, ms * 1e3 / N);
        }
        // alternate plain / cooperative (as in the ICP loop)
        {
            int syncs = 1; void* args[] = {&d, &syncs};
            for (int rep = 0; rep < 2; ++rep) {
                CK(cudaEventRecord(e0, st));
                for (int i = 0; i < N; ++i) {
                    k_plain<<<grid, 256, 0, st>>>(d);
                    CK(cudaLaunchCooperativeKernel((const void*)k_coop, dim3(grid), dim3(256), args, 0, st));
                }
                CK(cudaEventRecord(e1, st)); CK(cudaEventSynchronize(e1));
            }
            CK(cudaEventElapsedTime(&ms, e0, e1));
            printf("grid %3d  plain + cooperative(1 sync)   %.2f us/pair\n", grid, ms * 1e3 / N);
        }
        // (d)
        if (grid <= 592) {
            CK(cudaMemset(bar, 0, 4));
            for (int rep = 0; rep < 2; ++rep) {
                CK(cudaMemsetAsync(bar, 0, 4, st));
                CK(cudaEventRecord(e0, st));
                for (int i = 0; i < N; ++i) k_atomic_barrier<<<grid, 256, 0, st>>>(d, bar, (unsigned)grid * (i + 1));
                CK(cudaEventRecord(e1, st)); CK(cudaEventSynchronize(e1));
            }
            CK(cudaEventElapsedTime(&ms, e0, e1));
            printf("grid %3d  plain + atomic grid barrier   %.2f us/launch\n", grid, ms * 1e3 / N);
        }
        // (e)
        {
            cudaLaunchConfig_t cfg = {};
            cfg.gridDim = dim3(grid); cfg.blockDim = dim3(256); cfg.stream = st;
            cudaLaunchAttribute at[1];
            at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
            at[0].val.programmaticStreamSerializationAllowed = 1;
            cfg.attrs = at; cfg.numAttrs = 1;
            for (int rep = 0; rep < 2; ++rep) {
                CK(cudaEventRecord(e0, st));
                for (int i = 0; i < N; ++i) CK(cudaLaunchKernelEx(&cfg, k_pdl, d));
                CK(cudaEventRecord(e1, st)); CK(cudaEventSynchronize(e1));
            }
            CK(cudaEventElapsedTime(&ms, e0, e1));
            printf("grid %3d  programmatic dependent launch %.2f us/launch\n", grid, ms * 1e3 / N);
        }
    }
    // CUDA graph of 4 plain kernels
    {
        cudaGraph_t g; cudaGraphExec_t ge;
        CK(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
        for (int i = 0; i < 4 * 30; ++i) k_plain<<<296, 256, 0, st>>>(d);
        CK(cudaStreamEndCapture(st, &g));
        CK(cudaGraphInstantiate(&ge, g, 0));
        float ms;
        for (int rep = 0; rep < 3; ++rep) {
            CK(cudaEventRecord(e0, st));
            for (int i = 0; i < 20; ++i) CK(cudaGraphLaunch(ge, st));
            CK(cudaEventRecord(e1, st)); CK(cudaEventSynchronize(e1));
        }
        CK(cudaEventElapsedTime(&ms, e0, e1));
        printf("graph of 120 plain kernels (grid 296)   %.2f us/kernel\n", ms * 1e3 / (20 * 120));
    }
    return 0;
}
