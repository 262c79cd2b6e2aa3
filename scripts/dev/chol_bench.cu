// Developer microbenchmark of the in-CTA Cholesky (not part of the product).
#include <cstdio>
#include <vector>
#include <cmath>
#include "../../tamcmc-c-_b200/csrc/tg_adapt.cuh"
__global__ void __launch_bounds__(256) k(const double *A, double *Lout, int n, long long *cyc, int *fail)
{
    extern __shared__ double sm[];
    const int ld = n | 1;
    double *S = sm, *work = sm + (size_t)n * ld;
    for (int e = threadIdx.x; e < n * n; e += blockDim.x) { int i = e / n, k2 = e % n; S[i * ld + k2] = A[e]; }
    __syncthreads();
    long long t0 = clock64();
#ifdef USE_REG
    int f = tg_cholesky_reg96<true>(S, n, ld);
#else
    int f = tg_cholesky_inplace<true>(S, n, ld, work);
#endif
    __syncthreads();
    long long t1 = clock64();
    for (int e = threadIdx.x; e < n * n; e += blockDim.x) { int i = e / n, k2 = e % n; Lout[e] = (k2 <= i) ? S[i * ld + k2] : 0.0; }
    if (threadIdx.x == 0) { *cyc = t1 - t0; *fail = f; }
}
int main()
{
    const int n = 94;
    std::vector<double> B(n * n), A(n * n, 0.0), L(n * n);
    srand(1);
    for (auto &v : B) v = rand() / (double)RAND_MAX - 0.5;
    for (int i = 0; i < n; i++) for (int j = 0; j < n; j++) { double s = 0; for (int k2 = 0; k2 < n; k2++) s += B[i * n + k2] * B[j * n + k2]; A[i * n + j] = s + (i == j ? 1.0 : 0.0); }
    double *dA, *dL; long long *dc; int *df;
    cudaMalloc(&dA, n * n * 8); cudaMalloc(&dL, n * n * 8); cudaMalloc(&dc, 8); cudaMalloc(&df, 4);
    cudaMemcpy(dA, A.data(), n * n * 8, cudaMemcpyHostToDevice);
    size_t smem = ((size_t)n * (n | 1) + 6 * 96) * 8;
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    for (int rep = 0; rep < 3; rep++) {
        k<<<1, 256, smem>>>(dA, dL, n, dc, df);
        cudaError_t e = cudaDeviceSynchronize();
        long long c; int f; cudaMemcpy(&c, dc, 8, cudaMemcpyDeviceToHost); cudaMemcpy(&f, df, 4, cudaMemcpyDeviceToHost);
        printf("%s cholesky n=%d: %lld cycles (%.1f per column), fail=%d\n", cudaGetErrorString(e), n, c, (double)c / n, f);
    }
#ifdef TG_CHOL_DBG
    { long long st[8]; cudaMemcpyFromSymbol(st, tg_chol_dbg, sizeof(st)); for (int q = 0; q < 6; q++) printf("phase %d: %lld cycles (%.1f per column)\n", q, st[q + 1] - st[q], (st[q + 1] - st[q]) / (q == 5 ? 14.0 : 16.0)); }
#endif
    cudaMemcpy(L.data(), dL, n * n * 8, cudaMemcpyDeviceToHost);
    double err = 0;
    for (int i = 0; i < n; i++) for (int j = 0; j <= i; j++) { double s = 0; for (int k2 = 0; k2 <= j; k2++) s += L[i * n + k2] * L[j * n + k2]; err = fmax(err, fabs(s - A[i * n + j])); }
    printf("max |LL^T - A| = %.3e\n", err);
    return 0;
}
