// Developer microbenchmark: DFMA issue rate with distinct register operands (outer-product pattern) on sm_100a.
#include <cstdio>
#include <cuda_runtime.h>
template <int MODE>
__global__ void k_outer(double *out, const double *in, int iters, long long *cyc)
{
    double ci[6], ck[6], a[6][6];
#pragma unroll
    for (int u = 0; u < 6; u++) { ci[u] = in[threadIdx.x + 32 * u]; ck[u] = in[threadIdx.x + 7 * u + 1]; }
#pragma unroll
    for (int u = 0; u < 6; u++)
#pragma unroll
        for (int v = 0; v < 6; v++) a[u][v] = u + v;
    __syncthreads();
    long long t0 = clock64();
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int v = 0; v < 6; v++)
#pragma unroll
            for (int u = v; u < 6; u++) {
                if (MODE == 0) a[u][v] = fma(-ci[u], ck[v], a[u][v]);      // three distinct operands
                if (MODE == 1) a[u][v] = fma(a[u][v], ci[0], ck[0]);       // accumulator * const + const
                if (MODE == 2) a[u][v] = fma(ci[u], ci[u], a[u][v]);       // same register twice
            }
    }
    long long t1 = clock64();
    double s = 0;
#pragma unroll
    for (int u = 0; u < 6; u++)
#pragma unroll
        for (int v = 0; v <= u; v++) s += a[u][v];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
template <int MODE> void run(int threads, double *out, double *in, long long *cyc)
{
    const int iters = 2000;
    k_outer<MODE><<<1, threads>>>(out, in, iters, cyc);
    cudaDeviceSynchronize();
    long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
    const double wps = threads / 128.0 < 1 ? 1 : threads / 128.0;
    printf("mode %d threads=%4d: %.1f cycles per 21 DFMA -> %.2f cycles/warp-inst/SMSP\n", MODE, threads, (double)c / iters, (double)c / iters / 21.0 / wps);
}
int main()
{
    double *out, *in; long long *cyc;
    cudaMalloc(&out, 1 << 20); cudaMalloc(&in, 1 << 20); cudaMemset(in, 0, 1 << 20); cudaMalloc(&cyc, 8);
    for (int t : {32, 128, 256, 512, 1024}) { run<0>(t, out, in, cyc); run<1>(t, out, in, cyc); run<2>(t, out, in, cyc); }
    return 0;
}
