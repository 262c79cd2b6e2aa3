// tg_eval.cuh -- the fused hot kernel: model spectrum (truncated Lorentzian sums + Harvey-like
// noise) and chi^2(2 dof) / chi-square log-likelihood partial sums for chains x bins.
//
// Reference (under /root/reference/tamcmc/sources): build_lorentzian.cpp:61-102 (per-bin profile:
// d=x-nu; p=4*d*d/G^2; res += H*V/(1+p) [* asymmetry]), :453-459 (window accumulate);
// noise_models.cpp:17-41 (harvey_like); likelihoods.cpp:17-39; model_def.cpp:291-320.
//
// Layout: grid.x = bin tiles (TG_THREADS*BPT bins each), grid.y = chain groups.  A CTA loads its
// tile of x, y, ln x ONCE into registers (coalesced, bin = base + j*TG_THREADS + tid) and reuses it
// for every chain of its group; per chain it compacts the Lorentzian components whose truncation
// window intersects the tile into shared memory (order-preserving, so the accumulation order is
// fixed), accumulates them in registers, adds the noise background, forms y/m and ln m, and
// reduces them with warp shuffles + one shared-memory stage into a per-(chain,tile) fp64 pair.
// The model spectrum itself never goes to HBM (unless the parity/debug pointer is set).
#pragma once
#include "tg_internal.cuh"

struct TgEvalArgs {
    TgDesc        d;
    TgTables      t;
    const double *x, *y, *lnx, *sig;   // [Nx]; sig may be NULL
    double2      *partials;            // [B][ntiles]  (sum y/m , sum ln m)  or (sum chi2 , 0)
    double       *model;               // [B][Nx] or NULL
    int           B, ntiles, chains_per_block;
};

template <int BPT, bool ASYM>
__device__ __forceinline__ void tg_accumulate(int n, const double *s_nu, const double *s_h, const double *s_q,
                                              const double *s_ifc, const double *s_t2, const int *s_lo, const int *s_hi,
                                              const double (&x)[BPT], double (&acc)[BPT], double asym, int tid)
{
    constexpr int TILE = TG_THREADS * BPT;
    for (int k = 0; k < n; k++) {
        const double nu = s_nu[k], h = s_h[k], q = s_q[k];
        const int lo = s_lo[k], hi = s_hi[k];
        double ifc = 0, t2 = 0;
        if (ASYM) { ifc = s_ifc[k]; t2 = s_t2[k]; }
        if (lo <= 0 && hi >= TILE) {             // window covers the whole tile: no per-bin predicate
#pragma unroll
            for (int j = 0; j < BPT; j++) {
                const double dd = x[j] - nu;
                const double t = fma(dd * dd, q, 1.0);
                double r = tg_rcp(t);
                if (ASYM) { const double a = fma(asym, fma(x[j], ifc, -1.0), 1.0); r = r * fma(a, a, t2); }
                acc[j] = fma(h, r, acc[j]);
            }
        } else {
#pragma unroll
            for (int j = 0; j < BPT; j++) {
                const int li = j * TG_THREADS + tid;
                if (li >= lo && li < hi) {
                    const double dd = x[j] - nu;
                    const double t = fma(dd * dd, q, 1.0);
                    double r = tg_rcp(t);
                    if (ASYM) { const double a = fma(asym, fma(x[j], ifc, -1.0), 1.0); r = r * fma(a, a, t2); }
                    acc[j] = fma(h, r, acc[j]);
                }
            }
        }
    }
}

template <int BPT>
__global__ void __launch_bounds__(TG_THREADS) tg_k_eval(TgEvalArgs a)
{
    constexpr int TILE = TG_THREADS * BPT;
    __shared__ double s_nu[TG_THREADS], s_h[TG_THREADS], s_q[TG_THREADS], s_ifc[TG_THREADS], s_t2[TG_THREADS];
    __shared__ int    s_lo[TG_THREADS], s_hi[TG_THREADS];
    __shared__ int    s_wcnt[TG_THREADS / 32];
    __shared__ double s_red[2][TG_THREADS / 32];
    __shared__ double s_nz[TG_NOISE_STRIDE];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tile = blockIdx.x;
    const int base = tile * TILE;
    const int Nx = a.d.Nx;

    double x[BPT], y[BPT], lx[BPT];
#pragma unroll
    for (int j = 0; j < BPT; j++) {
        const int idx = base + j * TG_THREADS + tid;
        const int ii = idx < Nx ? idx : Nx - 1;
        x[j] = __ldg(a.x + ii); y[j] = __ldg(a.y + ii); lx[j] = __ldg(a.lnx + ii);
    }

    const int c0 = blockIdx.y * a.chains_per_block;
    const int c1 = min(a.B, c0 + a.chains_per_block);
    for (int c = c0; c < c1; c++) {
        const size_t cb = (size_t)c * a.d.ncomp_max;
        const int ncomp = a.t.ncomp[c];
        if (tid < TG_NOISE_STRIDE) s_nz[tid] = a.t.noise[(size_t)c * TG_NOISE_STRIDE + tid];
        __syncthreads();                          // s_nz ready; previous chain's smem lists are dead
        const double asym = s_nz[TG_N_ASYM];

        double acc[BPT];
        if (s_nz[TG_N_GAUSS] != 0.0) {            // models.cpp:1977-1978 / :2030-2031
            const double gA = s_nz[TG_N_GA], gnu = s_nz[TG_N_GNU], gs2 = s_nz[TG_N_GS2];
#pragma unroll
            for (int j = 0; j < BPT; j++) { const double dd = x[j] - gnu; acc[j] = gA * exp(-0.5 * (dd * dd) / gs2); }
        } else {
#pragma unroll
            for (int j = 0; j < BPT; j++) acc[j] = 0.0;
        }

        // ---- Lorentzian components, 256 candidates per pass ----
        for (int chunk = 0; chunk < ncomp; chunk += TG_THREADS) {
            const int k = chunk + tid;
            int lo = 0, hi = 0;
            bool ov = false;
            if (k < ncomp) {
                lo = a.t.lo[cb + k] - base; hi = a.t.hi[cb + k] - base;
                ov = (lo < TILE) && (hi > 0) && (hi > lo);
            }
            const unsigned bal = __ballot_sync(0xffffffffu, ov);
            if (lane == 0) s_wcnt[warp] = __popc(bal);
            __syncthreads();
            int off = 0, total = 0;
#pragma unroll
            for (int w = 0; w < TG_THREADS / 32; w++) { const int n = s_wcnt[w]; if (w < warp) off += n; total += n; }
            if (ov) {
                const int pos = off + __popc(bal & ((1u << lane) - 1u));
                s_nu[pos] = a.t.nu[cb + k]; s_h[pos] = a.t.h[cb + k]; s_q[pos] = a.t.q[cb + k];
                s_lo[pos] = lo; s_hi[pos] = hi;
                if (asym != 0.0) { s_ifc[pos] = a.t.ifc[cb + k]; s_t2[pos] = a.t.t2sq[cb + k]; }
            }
            __syncthreads();
            if (total > 0) {
                if (asym != 0.0) tg_accumulate<BPT, true>(total, s_nu, s_h, s_q, s_ifc, s_t2, s_lo, s_hi, x, acc, asym, tid);
                else             tg_accumulate<BPT, false>(total, s_nu, s_h, s_q, s_ifc, s_t2, s_lo, s_hi, x, acc, asym, tid);
            }
            __syncthreads();                      // lists are overwritten by the next pass
        }

        // ---- Harvey-like background: y += H/(1+(1e-3*tc*x)^p) ; y += N0  (noise_models.cpp:30-38) ----
        const int nH = (int)s_nz[TG_N_NH];
        for (int k = 0; k < nH; k++) {
            const double H = s_nz[4 * k], ak = s_nz[4 * k + 1], pk = s_nz[4 * k + 2], lna = s_nz[4 * k + 3];
            if (ak != 0.0) {
                const int ip = (pk == 0.0) ? 0 : (pk == 1.0) ? 1 : (pk == 2.0) ? 2 : (pk == 3.0) ? 3 : (pk == 4.0) ? 4 : -1;
#pragma unroll
                for (int j = 0; j < BPT; j++) {
                    const double u = ak * x[j];
                    double pw;
                    if (ip >= 0) {               // exact small-integer exponents: plain products
                        const double u2 = u * u;
                        pw = (ip == 0) ? 1.0 : (ip == 1) ? u : (ip == 2) ? u2 : (ip == 3) ? u2 * u : u2 * u2;
                    } else {                      // u^p = exp(p*(ln a + ln x)), ln x staged once for all chains
                        pw = exp(pk * (lna + lx[j]));
                    }
                    acc[j] = fma(H, tg_rcp(pw + 1.0), acc[j]);
                }
            }
        }
        const double N0 = s_nz[TG_N_N0];
        double s1 = 0.0, s2 = 0.0;
#pragma unroll
        for (int j = 0; j < BPT; j++) {
            const int idx = base + j * TG_THREADS + tid;
            const double m = acc[j] + N0;
            if (idx < Nx) {
                if (a.model) a.model[(size_t)c * Nx + idx] = m;
                if (a.d.like_case == 0) {        // likelihoods.cpp:23
                    s1 += y[j] * tg_rcp(m);
                    s2 += log(m);
                } else {                          // likelihoods.cpp:36
                    const double sg = __ldg(a.sig + idx);
                    const double dd = y[j] - m;
                    s1 += (dd * dd) * tg_rcp(sg * sg);
                }
            }
        }
        // ---- CTA reduction (fixed order: lanes by shuffle tree, then warps 0..7) ----
        s1 = tg_warp_sum(s1); s2 = tg_warp_sum(s2);
        if (lane == 0) { s_red[0][warp] = s1; s_red[1][warp] = s2; }
        __syncthreads();
        if (tid == 0) {
            double t1 = 0, t2 = 0;
#pragma unroll
            for (int w = 0; w < TG_THREADS / 32; w++) { t1 += s_red[0][w]; t2 += s_red[1][w]; }
            a.partials[(size_t)c * a.ntiles + tile] = make_double2(t1, t2);
        }
    }
}

// Sums the per-tile pairs of each chain in a fixed order (thread-strided serial sums, then a
// shared-memory tree) -> sums[c] = (sum y/m, sum ln m).  In the bin-sharded variant this is the
// buffer that is all-reduced across GPUs.
__global__ void __launch_bounds__(256) tg_k_reduce(const double2 *partials, int ntiles, double2 *sums)
{
    __shared__ double r1[256], r2[256];
    const int c = blockIdx.x, tid = threadIdx.x;
    double a1 = 0, a2 = 0;
    for (int t = tid; t < ntiles; t += 256) { const double2 v = partials[(size_t)c * ntiles + t]; a1 += v.x; a2 += v.y; }
    r1[tid] = a1; r2[tid] = a2;
    __syncthreads();
    for (int s = 128; s > 0; s >>= 1) {
        if (tid < s) { r1[tid] += r1[tid + s]; r2[tid] += r2[tid + s]; }
        __syncthreads();
    }
    if (tid == 0) sums[c] = make_double2(r1[0], r2[0]);
}

// logL = -p*(s1+s2)/T (likelihoods.cpp:25 + model_def.cpp:302) or -s1/T; logPost = logL + logPrior
// (model_def.cpp:365).  A chain whose window arithmetic failed gets logL = NaN (=> rejected).
__device__ __forceinline__ double tg_logl_from_sums(const TgDesc &d, double2 s, double T, int status)
{
    if (status & (TG_ST_EMPTY_WINDOW | TG_ST_NAN_WINDOW | TG_ST_MODE_OVERFLOW)) return CUDART_NAN;
    const double f = (d.like_case == 0) ? -((double)d.likelihood_p) * (s.x + s.y) : -s.x;
    return f / T;
}

__global__ void tg_k_finalize_eval(TgDesc d, const double2 *sums, const double *T, TgTables t, int B,
                                   double *logL, double *logPrior, double *logPost, int *status)
{
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= B) return;
    int st = t.status[c];
    const double L = tg_logl_from_sums(d, sums[c], T[c], st);
    if (L != L) st |= TG_ST_NAN_LOGL;
    logL[c] = L; logPrior[c] = t.logprior[c]; logPost[c] = L + t.logprior[c]; status[c] = st;
}

__global__ void tg_k_lnx(const double *x, double *lnx, int n)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) lnx[i] = log(x[i]);
}
