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
// reduces them into a per-(chain,tile) fp64 pair.  The model spectrum never goes to HBM (unless
// the parity/debug pointer is set).
//
// FP64-pipe economy (the kernel is bound by the FP64 pipe, not by HBM):
//  * four components share ONE reciprocal:  sum_k h_k/t_k = N/T with T = t0 t1 t2 t3 and
//    N = (h0 t1 + h1 t0) t2 t3 + (h2 t3 + h3 t2) t0 t1  -> 25 FP64 instructions per 4
//    component-bins instead of 4 x 9;
//  * the reciprocal is MUFU.RCP64H + one cubic correction (3 DFMA, |err| <= 1 ulp);
//  * sum_i ln m_i = ln prod_i m_i: a thread multiplies its BPT model values, the warp multiplies
//    the 32 mantissas and adds the exponents by shuffles, and ONE log per warp is evaluated;
//  * sum_i y_i/m_i over the 4 bins of a thread shares one reciprocal of the same product;
//  * integer Harvey exponents p in {1,2,3,4} use (a^p)*(x^p) with x^2 staged once per tile.
// All of these change rounding by a few ulp only (tests: 1e-12 relative against the oracle).
#pragma once
#include "tg_internal.cuh"

struct TgEvalArgs {
    TgDesc        d;
    TgTables      t;
    const double *x, *y, *lnx, *sig;   // [Nx]; sig may be NULL
    double2      *partials;            // [B][ntiles]  (sum y/m , sum ln m)  or (sum chi2 , 0)
    double       *model;               // [B][Nx] or NULL
    const int    *tile_order;          // [ntiles] launch order -> tile (heaviest first) or NULL
    int           B, ntiles, chains_per_block;
    int           nsub;                // sub-tiles (TG_THREADS*BPT bins each) per CTA tile
    int           list_cap;            // capacity (components) of one chain's list in shared memory, multiple of 4
};

struct __align__(16) TgComp {       // one Lorentzian m-component clipped to the tile, 48 bytes
    double nu, q;                   // centre, 4/Gamma^2
    double h; int lo, len;          // height, window start (tile-local), window length
    double ifc, t2;                 // asymmetry extras: 1/fc, (Gamma*asym/(2fc))^2
};
#define TG_MAXG 8                   // chains per CTA (chain group) at most == warps per CTA
// dynamic shared memory of tg_k_eval for a group of G chains and a list capacity of cap components:
// lists, noise tables, per-thread partial results (sum, mantissa product, exponent sum)
#define TG_EVAL_SMEM(G, cap) ((size_t)(G) * ((size_t)(cap) * sizeof(TgComp) + TG_NOISE_STRIDE * sizeof(double) + TG_THREADS * (2 * sizeof(double) + sizeof(int))))

// height of a component at tile-local bin li: h inside its window, 0 outside
__device__ __forceinline__ double tg_hsel(double h, int lo, int len, int li)
{
    return ((unsigned)(li - lo) < (unsigned)len) ? h : 0.0;
}

template <int BPT, bool ASYM>
__device__ __forceinline__ void tg_accumulate(int n4, const TgComp *L, const double (&x)[BPT], double (&acc)[BPT], double asym, int li0)
{
    constexpr int SUB = TG_THREADS * BPT;       // bins of one sub-tile; li0 = tile-local index of this thread's first bin
#pragma unroll 1
    for (int k = 0; k < n4; k += 4) {
        const double2 a0 = *(const double2 *)&L[k].nu, a1 = *(const double2 *)&L[k + 1].nu, a2 = *(const double2 *)&L[k + 2].nu, a3 = *(const double2 *)&L[k + 3].nu;
        const double h0 = L[k].h, h1 = L[k + 1].h, h2 = L[k + 2].h, h3 = L[k + 3].h;
        const int2 w0 = *(const int2 *)&L[k].lo, w1 = *(const int2 *)&L[k + 1].lo, w2 = *(const int2 *)&L[k + 2].lo, w3 = *(const int2 *)&L[k + 3].lo;
        const int s0 = li0 & ~(TG_THREADS - 1) , s1 = s0 + SUB;   // sub-tile range [s0, s1) (li0 = s0 + tid)
        const bool full = (w0.x <= s0 && w0.x + w0.y >= s1) && (w1.x <= s0 && w1.x + w1.y >= s1) &&
                          (w2.x <= s0 && w2.x + w2.y >= s1) && (w3.x <= s0 && w3.x + w3.y >= s1);
        if (full && !ASYM) {                     // CTA-uniform: every bin of the tile is inside all 4 windows
#pragma unroll
            for (int j = 0; j < BPT; j++) {
                const double d0 = x[j] - a0.x, d1 = x[j] - a1.x, d2 = x[j] - a2.x, d3 = x[j] - a3.x;
                const double t0 = fma(d0 * d0, a0.y, 1.0), t1 = fma(d1 * d1, a1.y, 1.0);
                const double t2 = fma(d2 * d2, a2.y, 1.0), t3 = fma(d3 * d3, a3.y, 1.0);
                const double t01 = t0 * t1, t23 = t2 * t3;
                const double n01 = fma(h0, t1, h1 * t0), n23 = fma(h2, t3, h3 * t2);
                const double T = t01 * t23;
                const double N = fma(n01, t23, n23 * t01);
                acc[j] = fma(N, tg_rcp_fast(T), acc[j]);
            }
        } else {
            double2 c0, c1, c2, c3;
            if (ASYM) { c0 = *(const double2 *)&L[k].ifc; c1 = *(const double2 *)&L[k + 1].ifc; c2 = *(const double2 *)&L[k + 2].ifc; c3 = *(const double2 *)&L[k + 3].ifc; }
#pragma unroll
            for (int j = 0; j < BPT; j++) {
                const int li = li0 + j * TG_THREADS;
                double g0 = tg_hsel(h0, w0.x, w0.y, li), g1 = tg_hsel(h1, w1.x, w1.y, li), g2 = tg_hsel(h2, w2.x, w2.y, li), g3 = tg_hsel(h3, w3.x, w3.y, li);
                if (ASYM) {                      // build_lorentzian.cpp:94-97: (1+asym(x/fc-1))^2 + (Gamma*asym/(2fc))^2
                    const double e0 = fma(asym, fma(x[j], c0.x, -1.0), 1.0), e1 = fma(asym, fma(x[j], c1.x, -1.0), 1.0);
                    const double e2 = fma(asym, fma(x[j], c2.x, -1.0), 1.0), e3 = fma(asym, fma(x[j], c3.x, -1.0), 1.0);
                    g0 *= fma(e0, e0, c0.y); g1 *= fma(e1, e1, c1.y); g2 *= fma(e2, e2, c2.y); g3 *= fma(e3, e3, c3.y);
                }
                const double d0 = x[j] - a0.x, d1 = x[j] - a1.x, d2 = x[j] - a2.x, d3 = x[j] - a3.x;
                const double t0 = fma(d0 * d0, a0.y, 1.0), t1 = fma(d1 * d1, a1.y, 1.0);
                const double t2 = fma(d2 * d2, a2.y, 1.0), t3 = fma(d3 * d3, a3.y, 1.0);
                const double t01 = t0 * t1, t23 = t2 * t3;
                const double n01 = fma(g0, t1, g1 * t0), n23 = fma(g2, t3, g3 * t2);
                const double T = t01 * t23;
                const double N = fma(n01, t23, n23 * t01);
                acc[j] = fma(N, tg_rcp_fast(T), acc[j]);
            }
        }
    }
}

// Branch-free exp for the Harvey power law (so that the BPT bins of a thread interleave on the
// FP64 pipe): z clamped to [-700, 322] (results 1e-304 .. 1e140, no denormals, no overflow of the
// paired denominators), NaN stays NaN.  k = rint(z log2 e) by the 1.5*2^52 trick, two-constant
// Cody-Waite reduction, degree-12 Taylor polynomial on |r| <= ln2/2 (truncation 1.7e-16), scaling
// by 2^k through the exponent field.  Relative error <= ~3e-16.
__device__ __forceinline__ double tg_exp_nb(double z)
{
    z = (z > 322.0) ? 322.0 : z;
    z = (z < -700.0) ? -700.0 : z;
    double kd = fma(z, 1.4426950408889634074, 6755399441055744.0);
    const int k = __double2loint(kd);
    kd -= 6755399441055744.0;
    double r = fma(kd, -6.93147180369123816490e-01, z);
    r = fma(kd, -1.90821492927058770002e-10, r);
    double p = 2.08767569878680989792e-09;            // 1/12!
    p = fma(p, r, 2.50521083854417187751e-08);         // 1/11!
    p = fma(p, r, 2.75573192239858906526e-07);         // 1/10!
    p = fma(p, r, 2.75573192239858906526e-06);         // 1/9!
    p = fma(p, r, 2.48015873015873015873e-05);         // 1/8!
    p = fma(p, r, 1.98412698412698412698e-04);         // 1/7!
    p = fma(p, r, 1.38888888888888888889e-03);         // 1/6!
    p = fma(p, r, 8.33333333333333333333e-03);         // 1/5!
    p = fma(p, r, 4.16666666666666666667e-02);         // 1/4!
    p = fma(p, r, 1.66666666666666666667e-01);         // 1/3!
    p = fma(p, r, 0.5);
    p = fma(p, r, 1.0);
    p = fma(p, r, 1.0);
    return p * __hiloint2double((k + 1023) << 20, 0);
}

// t[j] = 1 + (a x_j)^p for one Harvey term and the BPT bins of a thread.  The exponent class is
// chain-uniform, so the switch sits OUTSIDE the bin loop and every case is straight-line code.
template <int BPT>
__device__ __forceinline__ void tg_harvey_t(const double *q, const double (&x)[BPT], const double *lnx, const int (&ii)[BPT], double (&t)[BPT])
{
    const int cls = (int)q[TG_H_CLS];
    const double ap = q[TG_H_AP];
    switch (cls) {
    case 0:
#pragma unroll
        for (int j = 0; j < BPT; j++) t[j] = 2.0;                                   // (a x)^0 = 1
        break;
    case 1:
#pragma unroll
        for (int j = 0; j < BPT; j++) t[j] = fma(ap, x[j], 1.0);
        break;
    case 2:
#pragma unroll
        for (int j = 0; j < BPT; j++) t[j] = fma(ap * x[j], x[j], 1.0);
        break;
    case 3:
#pragma unroll
        for (int j = 0; j < BPT; j++) t[j] = fma((ap * x[j]) * x[j], x[j], 1.0);
        break;
    case 4:
#pragma unroll
        for (int j = 0; j < BPT; j++) { const double x2 = x[j] * x[j]; t[j] = fma(ap * x2, x2, 1.0); }
        break;
    default: {                                                                       // u^p = exp(p (ln a + ln x)); ln x staged once for all chains
        const double pk = q[TG_H_P], lna = q[TG_H_LNA];
#pragma unroll
        for (int j = 0; j < BPT; j++) t[j] = tg_exp_nb(pk * (lna + __ldg(lnx + ii[j]))) + 1.0;
        break; }
    }
}

template <int BPT>
__global__ void __launch_bounds__(TG_THREADS) tg_k_eval(TgEvalArgs a)
{
    constexpr int SUB = TG_THREADS * BPT;
    constexpr int NW = TG_THREADS / 32;
    extern __shared__ __align__(16) unsigned char tg_eval_smem[];
    __shared__ int s_cnt[TG_MAXG][NW];
    __shared__ int s_tot[TG_MAXG];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tile = a.tile_order ? __ldg(a.tile_order + blockIdx.x) : blockIdx.x;
    const int nsub = a.nsub;                      // sub-tiles of SUB bins per CTA tile
    const int TILE = SUB * nsub;
    const int base = tile * TILE;
    const int Nx = a.d.Nx;
    const int cap = a.list_cap;
    const int c0 = blockIdx.y * a.chains_per_block;
    const int G = min(a.B - c0, a.chains_per_block);
    TgComp *lists = (TgComp *)tg_eval_smem;                                      // [G][cap]
    double *s_nz  = (double *)(lists + (size_t)a.chains_per_block * cap);        // [G][TG_NOISE_STRIDE]
    double *s_sum = s_nz + (size_t)a.chains_per_block * TG_NOISE_STRIDE;         // [G][TG_THREADS]
    double *s_pf  = s_sum + (size_t)a.chains_per_block * TG_THREADS;             // [G][TG_THREADS]
    int    *s_pe  = (int *)(s_pf + (size_t)a.chains_per_block * TG_THREADS);     // [G][TG_THREADS]

    // ---- phase 1: per-chain lists of the components whose window meets this tile (order-preserving
    //      ballot compaction, all chains of the group at once), noise tables ----
    for (int e = tid; e < G * TG_NOISE_STRIDE; e += TG_THREADS) s_nz[e] = a.t.noise[(size_t)c0 * TG_NOISE_STRIDE + e];
    if (tid < TG_MAXG) s_tot[tid] = 0;
    for (int chunk = 0; chunk < a.d.ncomp_max; chunk += TG_THREADS) {
        const int k = chunk + tid;
        int lo[TG_MAXG], hi[TG_MAXG];
        unsigned bal[TG_MAXG];
        __syncthreads();                          // s_tot of the previous pass is final
#pragma unroll
        for (int g = 0; g < TG_MAXG; g++) {
            lo[g] = 0; hi[g] = 0;
            if (g < G && k < a.t.ncomp[c0 + g]) {
                const size_t e = (size_t)(c0 + g) * a.d.ncomp_max + k;
                lo[g] = a.t.lo[e] - base; hi[g] = a.t.hi[e] - base;
            }
        }
#pragma unroll
        for (int g = 0; g < TG_MAXG; g++) {
            const bool ov = (lo[g] < TILE) && (hi[g] > 0) && (hi[g] > lo[g]);
            bal[g] = __ballot_sync(0xffffffffu, ov);
            if (lane == 0 && g < G) s_cnt[g][warp] = __popc(bal[g]);
        }
        __syncthreads();
#pragma unroll
        for (int g = 0; g < TG_MAXG; g++) {
            if (g < G && ((bal[g] >> lane) & 1u)) {
                int off = s_tot[g];
#pragma unroll
                for (int w = 0; w < NW; w++) if (w < warp) off += s_cnt[g][w];
                const size_t e = (size_t)(c0 + g) * a.d.ncomp_max + k;
                TgComp q;
                q.nu = a.t.nu[e]; q.q = a.t.q[e]; q.h = a.t.h[e];
                q.lo = max(lo[g], -(1 << 29)); q.len = min(hi[g], 1 << 29) - q.lo;      // keep hi-lo inside int
                q.ifc = a.t.ifc[e]; q.t2 = a.t.t2sq[e];
                lists[(size_t)g * cap + off + __popc(bal[g] & ((1u << lane) - 1u))] = q;
            }
        }
        __syncthreads();                          // everyone has read s_tot / s_cnt of this pass
        if (tid < G) { int t = s_tot[tid]; for (int w = 0; w < NW; w++) t += s_cnt[tid][w]; s_tot[tid] = t; }
    }
    __syncthreads();
    if (tid < 4 * TG_MAXG) {                      // pad every list to a multiple of 4 with null components (t = 1, h = 0)
        const int g = tid >> 2, u = tid & 3;
        if (g < G) {
            const int total = s_tot[g], n4 = (total + 3) & ~3;
            if (total + u < n4) { TgComp q; q.nu = 0; q.q = 0; q.h = 0; q.lo = -(1 << 29); q.len = 1 << 30; q.ifc = 0; q.t2 = 0; lists[(size_t)g * cap + total + u] = q; }
        }
    }
    __syncthreads();

    // ---- phase 2: every chain of the group x every sub-tile, no CTA barrier inside (warps drift
    //      apart and hide each other's latencies) ----
    for (int g = 0; g < G; g++) {
        const int c = c0 + g;
        const double *nz = s_nz + (size_t)g * TG_NOISE_STRIDE;
        const TgComp *L = lists + (size_t)g * cap;
        const int n4 = (s_tot[g] + 3) & ~3;
        const double asym = nz[TG_N_ASYM], N0 = nz[TG_N_N0];
        const int nH = (int)nz[TG_N_NH];
        const bool gauss = nz[TG_N_GAUSS] != 0.0;
        double sum = 0.0;                         // sum y/m (or chi-square) + logs taken the slow way
        double pf = 1.0; int pe = 0;              // running product of model values: pf * 2^pe, pf in [1,2)
        for (int sub = 0; sub < nsub; sub++) {
            const int li0 = sub * SUB + tid;
            double x[BPT], acc[BPT];
            int ii[BPT];
            bool any = false;
#pragma unroll
            for (int j = 0; j < BPT; j++) {
                const int idx = base + li0 + j * TG_THREADS;
                ii[j] = idx < Nx ? idx : Nx - 1;
                any |= idx < Nx;
                x[j] = __ldg(a.x + ii[j]);
                acc[j] = 0.0;
            }
            if (base + sub * SUB >= Nx) break;    // whole sub-tile past the end (CTA-uniform)
            if (gauss) {                          // models.cpp:1977-1978 / :2030-2031
                const double gA = nz[TG_N_GA], gnu = nz[TG_N_GNU], gs2 = nz[TG_N_GS2];
#pragma unroll
                for (int j = 0; j < BPT; j++) { const double dd = x[j] - gnu; acc[j] = gA * exp(-0.5 * (dd * dd) / gs2); }
            }
            if (n4 > 0) {
                if (asym != 0.0) tg_accumulate<BPT, true>(n4, L, x, acc, asym, li0);
                else             tg_accumulate<BPT, false>(n4, L, x, acc, asym, li0);
            }
            // ---- Harvey-like background: y += H/(1+(1e-3*tc*x)^p) ; y += N0  (noise_models.cpp:30-38).
            //      The prologue lists only the active terms (tc != 0); consecutive pairs share a reciprocal.
            for (int k = 0; k < nH; k += 2) {
                double ta[BPT], tb[BPT];
                tg_harvey_t<BPT>(nz + TG_HS * k, x, a.lnx, ii, ta);
                const double Ha = nz[TG_HS * k + TG_H_H];
                if (k + 1 < nH) {
                    tg_harvey_t<BPT>(nz + TG_HS * (k + 1), x, a.lnx, ii, tb);
                    const double Hb = nz[TG_HS * (k + 1) + TG_H_H];
#pragma unroll
                    for (int j = 0; j < BPT; j++) acc[j] = fma(fma(Ha, tb[j], Hb * ta[j]), tg_rcp_fast(ta[j] * tb[j]), acc[j]);
                } else {
#pragma unroll
                    for (int j = 0; j < BPT; j++) acc[j] = fma(Ha, tg_rcp_fast(ta[j]), acc[j]);
                }
            }
            double m[BPT], y[BPT];
#pragma unroll
            for (int j = 0; j < BPT; j++) {
                const int idx = base + li0 + j * TG_THREADS;
                m[j] = acc[j] + N0;
                y[j] = __ldg(a.y + ii[j]);
                if (idx < Nx) { if (a.model) a.model[(size_t)c * Nx + idx] = m[j]; }
                else { m[j] = 1.0; y[j] = 0.0; }  // bins past the end: ln 1 = 0 and y/m = 0
            }
            if (a.d.like_case == 0) {             // likelihoods.cpp:23
#pragma unroll
                for (int j0 = 0; j0 < BPT; j0 += 4) {
                    double mm[4], yy[4];
#pragma unroll
                    for (int u = 0; u < 4; u++) {
                        const bool in = (j0 + u < BPT);
                        mm[u] = in ? m[in ? (j0 + u) : 0] : 1.0;
                        yy[u] = in ? y[in ? (j0 + u) : 0] : 0.0;
                    }
                    const double p01 = mm[0] * mm[1], p23 = mm[2] * mm[3], P = p01 * p23;
                    const int sgn = __double2hiint(mm[0]) | __double2hiint(mm[1]) | __double2hiint(mm[2]) | __double2hiint(mm[3]);
                    if (sgn >= 0 && P > 1e-250 && P < 1e250) {
                        const double n = fma(fma(yy[0], mm[1], yy[1] * mm[0]), p23, fma(yy[2], mm[3], yy[3] * mm[2]) * p01);
                        sum = fma(n, tg_rcp_fast(P), sum);
                        // fold P into the running product; the exponent is kept apart so it can never overflow
                        pf *= P;                  // pf in [1,2), P in (1e-250,1e250): no overflow
                        const int hi = __double2hiint(pf);
                        pe += ((hi >> 20) & 0x7ff) - 1023;
                        pf = __hiloint2double((hi & 0x000fffff) | 0x3ff00000, __double2loint(pf));
                    } else {                      // zero / negative / NaN / huge model values: the plain way
#pragma unroll
                        for (int u = 0; u < 4; u++) if (j0 + u < BPT && base + li0 + (j0 + u) * TG_THREADS < Nx) sum += yy[u] * tg_rcp(mm[u]) + log(mm[u]);
                    }
                }
            } else {                              // likelihoods.cpp:36
#pragma unroll
                for (int j = 0; j < BPT; j++) {
                    const int idx = base + li0 + j * TG_THREADS;
                    if (idx < Nx) {
                        const double sg = __ldg(a.sig + idx);
                        const double dd = y[j] - m[j];
                        sum += (dd * dd) * tg_rcp(sg * sg);
                    }
                }
            }
        }
        s_sum[g * TG_THREADS + tid] = sum; s_pf[g * TG_THREADS + tid] = pf; s_pe[g * TG_THREADS + tid] = pe;
    }
    __syncthreads();
    // ---- phase 3: warp g reduces chain g (fixed order: 8 strided entries per lane, shuffle tree),
    //      lane 0 takes ONE log of the mantissa product; all chains of the group in parallel ----
    if (warp < G) {
        double sum = 0.0, pf = 1.0;
        int pe = 0;
#pragma unroll
        for (int i = 0; i < NW; i++) {
            sum += s_sum[warp * TG_THREADS + i * 32 + lane];
            pf *= s_pf[warp * TG_THREADS + i * 32 + lane];           // 8 mantissas in [1,2): < 2^8
            pe += s_pe[warp * TG_THREADS + i * 32 + lane];
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {                           // 32 x 2^8 -> < 2^256
            sum += __shfl_down_sync(0xffffffffu, sum, o);
            pf *= __shfl_down_sync(0xffffffffu, pf, o);
            pe += __shfl_down_sync(0xffffffffu, pe, o);
        }
        if (lane == 0) {
            double t2 = 0.0;
            if (a.d.like_case == 0) t2 = fma((double)pe, 6.93147180369123816490e-01, fma((double)pe, 1.90821492927058770002e-10, log(pf)));
            a.partials[(size_t)(c0 + warp) * a.ntiles + tile] = make_double2(sum, t2);
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
