// Microbenchmark: cost of a grid-wide barrier on B200 for different grid shapes, and of a
// hand-rolled barrier (development aid for DESIGN.md's phase-latency numbers).
#include <cooperative_groups.h>
#include <cstdio>
#include <cuda_runtime.h>
namespace cg = cooperative_groups;

__global__ void k_cg(int iters, float* sink) {
    cg::grid_group g = cg::this_grid();
    float x = threadIdx.x;
    for (int i = 0; i < iters; ++i) {
        x = x * 1.0001f + 1.0f;
        g.sync();
    }
    if (x == 12345.f) *sink = x;
}

// sense-reversing barrier: one arrive counter, release by generation flag
__device__ __forceinline__ void my_barrier(unsigned* count, volatile unsigned* gen, unsigned nblocks) {
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned g = *gen;
        __threadfence();
        if (atomicAdd(count, 1u) == nblocks - 1) {
            *count = 0;
            __threadfence();
            *gen = g + 1;
        } else {
            while (*gen == g) { }
        }
        __threadfence();
    }
    __syncthreads();
}
__global__ void k_my(int iters, unsigned* count, unsigned* gen, float* sink) {
    float x = threadIdx.x;
    for (int i = 0; i < iters; ++i) {
        x = x * 1.0001f + 1.0f;
        my_barrier(count, gen, gridDim.x);
    }
    if (x == 12345.f) *sink = x;
}

// monotonic-counter barrier: arrive with a fire-and-forget release reduction (no round trip for
// a returned value, no reset, no generation flag), wait with acquire loads
__device__ __forceinline__ void mono_barrier(unsigned* ctr, unsigned& target, unsigned nblocks) {
    __syncthreads();
    if (threadIdx.x == 0) {
        target += nblocks;
        asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(ctr) : "memory");
        unsigned v;
        do {
            asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(ctr) : "memory");
        } while ((int)(v - target) < 0);
    }
    __syncthreads();
}
__global__ void k_mono(int iters, unsigned* ctr, float* sink) {
    float x = threadIdx.x;
    unsigned target = 0;
    for (int i = 0; i < iters; ++i) {
        x = x * 1.0001f + 1.0f;
        mono_barrier(ctr, target, gridDim.x);
    }
    if (x == 12345.f) *sink = x;
}
// the same with a store in front of each barrier (what a solver phase does), for both kinds
__global__ void k_mono_st(int iters, unsigned* ctr, float4* data, float* sink) {
    unsigned target = 0;
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    for (int i = 0; i < iters; ++i) {
        data[t] = make_float4(i, t, 0, 0);
        mono_barrier(ctr, target, gridDim.x);
    }
}
__global__ void k_cg_st(int iters, float4* data, float* sink) {
    cg::grid_group g = cg::this_grid();
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    for (int i = 0; i < iters; ++i) {
        data[t] = make_float4(i, t, 0, 0);
        g.sync();
    }
}

// a dependent gather chain between barriers: what one solver phase costs at minimum
__global__ void k_phase(int iters, const int* idx, float4* data, int n, float* sink) {
    cg::grid_group g = cg::this_grid();
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    float acc = 0;
    for (int i = 0; i < iters; ++i) {
        if (t < n) {
            int j = idx[(t + i * 977) % n];
            float4 v = data[j];
            v.x += 1.0f;
            data[j] = v;
            acc += v.x;
        }
        g.sync();
    }
    if (acc == 12345.f) *sink = acc;
}

int main() {
    int iters = 2000;
    float* sink; cudaMalloc(&sink, 4);
    unsigned *count, *gen; cudaMalloc(&count, 4); cudaMalloc(&gen, 4); cudaMemset(count, 0, 4); cudaMemset(gen, 0, 4);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    int shapes[][2] = {{148, 128}, {148, 256}, {148, 512}, {148, 1024}, {296, 256}, {444, 256}, {592, 256}, {74, 512}, {32, 256}, {8, 256}, {1, 256}};
    for (auto& s : shapes) {
        int nb = s[0], nt = s[1];
        void* args[] = {&iters, &sink};
        cudaLaunchCooperativeKernel((void*)k_cg, dim3(nb), dim3(nt), args, 0, 0);
        cudaDeviceSynchronize();
        cudaEventRecord(e0);
        cudaError_t err = cudaLaunchCooperativeKernel((void*)k_cg, dim3(nb), dim3(nt), args, 0, 0);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        void* args2[] = {&iters, &count, &gen, &sink};
        cudaEventRecord(e0);
        cudaError_t err2 = cudaLaunchCooperativeKernel((void*)k_my, dim3(nb), dim3(nt), args2, 0, 0);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms2; cudaEventElapsedTime(&ms2, e0, e1);
        cudaMemset(count, 0, 4);
        void* args3[] = {&iters, &count, &sink};
        cudaEventRecord(e0);
        cudaError_t err3 = cudaLaunchCooperativeKernel((void*)k_mono, dim3(nb), dim3(nt), args3, 0, 0);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms3; cudaEventElapsedTime(&ms3, e0, e1);
        cudaMemset(count, 0, 4);
        float4* scratch; cudaMalloc(&scratch, (size_t)nb * nt * 16);
        void* args4[] = {&iters, &count, &scratch, &sink};
        cudaEventRecord(e0);
        cudaLaunchCooperativeKernel((void*)k_mono_st, dim3(nb), dim3(nt), args4, 0, 0);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms4; cudaEventElapsedTime(&ms4, e0, e1);
        void* args5[] = {&iters, &scratch, &sink};
        cudaEventRecord(e0);
        cudaLaunchCooperativeKernel((void*)k_cg_st, dim3(nb), dim3(nt), args5, 0, 0);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms5; cudaEventElapsedTime(&ms5, e0, e1);
        cudaMemset(count, 0, 4);
        cudaFree(scratch);
        printf("grid %4d x %4d: cg::grid.sync %.3f us/sync (%s)   hand barrier %.3f us (%s)   monotonic red/acquire %.3f us (%s)   with a store before: cg %.3f, monotonic %.3f\n",
               nb, nt, ms * 1e3 / iters, cudaGetErrorString(err), ms2 * 1e3 / iters, cudaGetErrorString(err2),
               ms3 * 1e3 / iters, cudaGetErrorString(err3), ms5 * 1e3 / iters, ms4 * 1e3 / iters);
    }
    int n = 20000; int* idx; float4* data; cudaMalloc(&idx, n * 4); cudaMalloc(&data, 200000 * 16);
    int* h = new int[n]; for (int i = 0; i < n; ++i) h[i] = (int)((i * 2654435761u) % 200000u);
    cudaMemcpy(idx, h, n * 4, cudaMemcpyHostToDevice); cudaMemset(data, 0, 200000 * 16);
    for (int nb : {148, 296}) {
        void* args[] = {&iters, &idx, &data, &n, &sink};
        cudaLaunchCooperativeKernel((void*)k_phase, dim3(nb), dim3(256), args, 0, 0);
        cudaDeviceSynchronize();
        cudaEventRecord(e0);
        cudaLaunchCooperativeKernel((void*)k_phase, dim3(nb), dim3(256), args, 0, 0);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        printf("phase = idx load -> gather -> scatter -> grid.sync, grid %d x 256: %.3f us/phase\n", nb, ms * 1e3 / iters);
    }
    return 0;
}
