// Microbenchmark (development aid): latency of ONE warp running the position / velocity solve
// of a two-point manifold from registers, i.e. the arithmetic chain that every colour phase
// has on its critical path.   nvcc ... -I include scripts/solve_latency.cu
#include <cstdio>
#include <cuda_runtime.h>
#include "sfp_math.h"
using namespace sfp;

__global__ void k_pos(float* out, long long* cyc, int iters, float jitter) {
    BodyPos A, B;
    A.x = mk(0.0f, 0.0f); A.q = mk(1.0f, 0.0f); A.d = mk(0.001f, -0.002f); A.qp = A.q; A.im = 1.0f; A.ii = 6.0f;
    B.x = mk(0.02f * threadIdx.x * jitter, 0.49f); B.q = mk(0.99995f, 0.01f); B.d = mk(-0.001f, -0.003f); B.qp = B.q; B.im = 0.5f; B.ii = 3.0f;
    v2 n = mk(0.0f, 1.0f);
    v2 la0 = mk(-0.2f, 0.25f), lb0 = mk(-0.2f, -0.25f), la1 = mk(0.2f, 0.25f), lb1 = mk(0.2f, -0.25f);
    long long t0 = clock64();
    f4 lam;
    for (int i = 0; i < iters; ++i) {
        lam = solve_contact_manifold(&A, &B, n, 2u, la0, lb0, la1, lb1, 0.6f);
        B.x.y = B.x.y - 0.0005f;  // keep it penetrating
    }
    long long t1 = clock64();
    out[threadIdx.x] = A.x.x + B.x.y + A.q.y + B.q.x + lam.x + lam.y + lam.z;
    if (threadIdx.x == 0) cyc[0] = (t1 - t0) / iters;
}

__global__ void k_vel(float* out, long long* cyc, int iters) {
    BodyVel A, B;
    A.v = mk(0.1f, -0.2f); A.w = 0.3f; A.im = 1.0f; A.ii = 6.0f;
    B.v = mk(-0.1f, -0.4f); B.w = -0.2f; B.im = 0.5f; B.ii = 3.0f;
    v2 qa = mk(1.0f, 0.0f), qb = mk(0.99995f, 0.01f), n = mk(0.0f, 1.0f);
    v2 la0 = mk(-0.2f, 0.25f), lb0 = mk(-0.2f, -0.25f), la1 = mk(0.2f, 0.25f), lb1 = mk(0.2f, -0.25f);
    f4 lam; lam.x = 1e-5f; lam.y = 0.0f; lam.z = 2e-5f; lam.w = 0.0f;
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
        solve_manifold_velocity(&A, &B, qa, qb, n, 2u, la0, lb0, la1, lb1, lam, -0.3f, -0.2f, 0.6f, 0.4f, 0.1f,
                                1.0f / 480.0f, 0.04f);
        B.v.y = B.v.y - 0.02f;
    }
    long long t1 = clock64();
    out[threadIdx.x] = A.v.x + B.v.y + A.w + B.w;
    if (threadIdx.x == 0) cyc[1] = (t1 - t0) / iters;
}

int main() {
    float* out; long long* cyc;
    cudaMalloc(&out, 4096); cudaMallocManaged(&cyc, 64);
    for (int rep = 0; rep < 2; ++rep) {
        k_pos<<<1, 32>>>(out, cyc, 200, 0.0f);
        k_vel<<<1, 32>>>(out, cyc, 200);
        cudaDeviceSynchronize();
    }
    printf("one warp, uniform lanes: position solve %lld cycles, velocity solve %lld cycles per manifold\n", cyc[0], cyc[1]);
    return cudaGetLastError() != cudaSuccess;
}
