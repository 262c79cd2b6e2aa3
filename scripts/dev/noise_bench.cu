// Developer microbenchmark: standalone throughput of the noise-only chain loop (two Harvey terms of classes 4 and 2
// + white noise + chi^2(2 dof) likelihood), NB = 8 and NB = 4, with every warp of the SM inside it.
#include <cstdio>
#include <cuda_runtime.h>
#include "../../tamcmc-c-_b200/csrc/tg_eval.cuh"

template <int NB>
__global__ void __launch_bounds__(TG_THREADS, 2) k_noise(TgEvalArgs a, int G, int reps, double *out)
{
    extern __shared__ __align__(16) unsigned char sm[];
    __shared__ int s_tot[TG_MAXG];
    const int tid = threadIdx.x;
    TgEvalSmem S;
    S.lists = (TgComp *)sm;
    S.nz = (double *)sm;
    S.sum = S.nz + G * TG_NOISE_STRIDE; S.pf = S.sum + G * TG_THREADS; S.pe = (int *)(S.pf + G * TG_THREADS);
    S.full = S.pe + G * TG_THREADS; S.cls = S.full; S.nh = S.cls + G * TG_MAXH;
    for (int e = tid; e < G * TG_NOISE_STRIDE; e += TG_THREADS) S.nz[e] = 0.0;
    __syncthreads();
    if (tid < G) {
        double *nz = S.nz + tid * TG_NOISE_STRIDE;
        tg_set_harvey(nz, 0, 11.0 + tid, 0.0496, 4.0, 9000.0);
        tg_set_harvey(nz, 1, 0.93, 0.00135, 2.0, 9000.0);
        nz[TG_N_N0] = 0.13; nz[TG_N_NH] = 2.0;
        S.cls[tid * TG_MAXH + 0] = 4; S.cls[tid * TG_MAXH + 1] = 2; S.nh[tid] = 2;
    }
    for (int g = 0; g < G; g++) { S.sum[g * TG_THREADS + tid] = 0.0; S.pf[g * TG_THREADS + tid] = 1.0; S.pe[g * TG_THREADS + tid] = 0; }
    __syncthreads();
    const int b0 = blockIdx.x * NB * TG_THREADS;
    for (int r = 0; r < reps; r++) {
        tg_chain_loop<NB, false, false>(a, S, s_tot, G, 0, b0, tid);
    }
    double s = 0;
    for (int g = 0; g < G; g++) s += S.sum[g * TG_THREADS + tid] + S.pf[g * TG_THREADS + tid] + S.pe[g * TG_THREADS + tid];
    out[blockIdx.x * TG_THREADS + tid] = s;
}

int main()
{
    const int nb = 148 * 2, reps = 200, G = 5, N = nb * 8 * TG_THREADS;
    double *x, *y, *o;
    cudaMalloc(&x, N * 8); cudaMalloc(&y, N * 8); cudaMalloc(&o, nb * TG_THREADS * 8);
    double *hx = new double[N];
    for (int i = 0; i < N; i++) hx[i] = 100.0 + 8000.0 * i / N;
    cudaMemcpy(x, hx, N * 8, cudaMemcpyHostToDevice);
    for (int i = 0; i < N; i++) hx[i] = 0.5 + (i % 7) * 0.1;
    cudaMemcpy(y, hx, N * 8, cudaMemcpyHostToDevice);
    TgEvalArgs a; memset(&a, 0, sizeof a);
    a.x = x; a.y = y; a.lnx = x; a.d.Nx = N; a.d.like_case = 0; a.list_cap = 4;
    const size_t smem = G * (TG_NOISE_STRIDE * 8 + TG_THREADS * 32 + 64 * 4) + 1024;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int nbv = 8; nbv >= 4; nbv -= 4)
        for (int it = 0; it < 3; it++) {
            cudaEventRecord(e0);
            if (nbv == 8) k_noise<8><<<nb, TG_THREADS, smem>>>(a, G, reps, o); else k_noise<4><<<nb, TG_THREADS, smem>>>(a, G, reps, o);
            cudaEventRecord(e1); cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            const double binchains = (double)nb * nbv * TG_THREADS * G * reps;
            if (it == 2) printf("NB=%d: %.3f ms  %.1f G bin-chains/s  -> %.1f cycles per bin-chain per SM-lane-group (1.965 GHz): %.2f bin-chains/cycle/SM\n", nbv, ms, binchains / ms / 1e6,
                                0.0, binchains / (ms * 1e-3) / 148 / 1.965e9);
        }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
