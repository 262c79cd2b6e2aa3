// Developer microbenchmark: standalone throughput of the accumulate inner loop (four components per reciprocal)
// with every warp of the SM inside it: the practical ceiling of the fused kernel's FP64 phase.
#include <cstdio>
#include <cuda_runtime.h>
#include "../../tamcmc-c-_b200/csrc/tg_eval.cuh"

template <bool FULL>
__global__ void __launch_bounds__(256, 2) k_acc(const double *xin, double *out, int n4, int reps)
{
    __shared__ TgComp L[224];
    __shared__ int full[56];
    const int tid = threadIdx.x;
    if (tid < 224) { TgComp q; q.nu = 2000.0 + 0.37 * tid; q.q = 4.0 / (1.0 + 0.01 * tid); q.h = 1.0 + tid; q.lo = FULL ? -10 : 5; q.len = 5000; q.ifc = 0; q.t2 = 0; L[tid] = q; }
    if (tid < 56) full[tid] = FULL ? 1 : 0;
    __syncthreads();
    double x[4], acc[4];
#pragma unroll
    for (int j = 0; j < 4; j++) { x[j] = xin[blockIdx.x * 1024 + j * 256 + tid]; acc[j] = 0.0; }
    for (int r = 0; r < reps; r++) tg_accumulate<4, false>(n4, L, full, x, acc, 0.0, tid);
    out[blockIdx.x * 256 + tid] = acc[0] + acc[1] + acc[2] + acc[3];
}

int main()
{
    const int nb = 148 * 2, reps = 200, n4 = 224;
    double *x, *o;
    cudaMalloc(&x, nb * 1024 * 8); cudaMalloc(&o, nb * 256 * 8);
    double *hx = new double[nb * 1024];
    for (int i = 0; i < nb * 1024; i++) hx[i] = 2000.0 + 1e-4 * i;
    cudaMemcpy(x, hx, nb * 1024 * 8, cudaMemcpyHostToDevice);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int full = 1; full >= 0; full--) {
        for (int it = 0; it < 3; it++) {
            cudaEventRecord(e0);
            if (full) k_acc<true><<<nb, 256>>>(x, o, n4, reps); else k_acc<false><<<nb, 256>>>(x, o, n4, reps);
            cudaEventRecord(e1); cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            const double compbins = (double)nb * 1024 * n4 * reps;
            const double fp64 = compbins * 6.25;                     // thread-level FP64 instructions (25 per 4 component-bins)
            if (it == 2) printf("%s path: %.3f ms  %.2f G comp-bins/s  FP64 thread-instr/s %.2f T  -> %.2f warp-instr/cycle/SM at 1.965 GHz (pipe peak 2.0)\n",
                                full ? "full-cover" : "windowed", ms, compbins / ms / 1e6, fp64 / ms / 1e9, fp64 / 32 / (ms * 1e-3) / 148 / 1.965e9);
        }
    }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
