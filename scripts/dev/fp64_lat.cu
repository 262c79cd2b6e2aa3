// Developer microbenchmark: FP64 pipe latency / throughput on sm_100a (not part of the product).
#include <cstdio>
#include <cuda_runtime.h>
template <int ILP>
__global__ void k_dfma(double *out, int iters, double a, double b, long long *cyc)
{
    double v[ILP];
#pragma unroll
    for (int u = 0; u < ILP; u++) v[u] = threadIdx.x + u;
    long long t0 = clock64();
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int r = 0; r < 8; r++)
#pragma unroll
            for (int u = 0; u < ILP; u++) v[u] = fma(v[u], a, b);
    }
    long long t1 = clock64();
    double s = 0;
#pragma unroll
    for (int u = 0; u < ILP; u++) s += v[u];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
__global__ void k_rcp(double *out, int iters, double a, long long *cyc)
{
    double v = a + threadIdx.x;
    long long t0 = clock64();
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int r = 0; r < 8; r++) { double q; asm volatile("rcp.approx.ftz.f64 %0, %1;" : "=d"(q) : "d"(v)); v = q; }
    }
    long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = v;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
template <int ILP> void run(int threads, double *out, long long *cyc)
{
    const int iters = 2000;
    k_dfma<ILP><<<1, threads>>>(out, iters, 0.99999, 1e-9, cyc);
    cudaDeviceSynchronize();
    long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
    double per = (double)c / (iters * 8.0);
    printf("DFMA ILP=%d threads=%4d (warps/SMSP=%.1f): %.2f cycles per round of %d dep-chains -> %.2f cycles/warp-inst/SMSP\n", ILP, threads,
           threads / 128.0, per, ILP, per / (ILP * (threads / 128.0 < 1 ? 1 : threads / 128.0)));
}
int main()
{
    double *out; long long *cyc;
    cudaMalloc(&out, 1 << 20); cudaMalloc(&cyc, 8);
    run<1>(32, out, cyc); run<2>(32, out, cyc); run<4>(32, out, cyc); run<8>(32, out, cyc); run<16>(32, out, cyc);
    run<1>(128, out, cyc); run<4>(128, out, cyc); run<8>(128, out, cyc);
    run<1>(512, out, cyc); run<4>(512, out, cyc); run<1>(1024, out, cyc); run<2>(1024, out, cyc); run<4>(1024, out, cyc);
    k_rcp<<<1, 32>>>(out, 2000, 1.5, cyc); cudaDeviceSynchronize();
    long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
    printf("MUFU.RCP64H dependent latency: %.2f cycles\n", (double)c / 16000.0);
    return 0;
}
