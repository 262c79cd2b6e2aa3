// tg_eval.cuh -- the fused hot kernel: model spectrum (truncated Lorentzian sums + Harvey-like
// noise) and chi^2(2 dof) / chi-square log-likelihood partial sums for chains x bins.
//
// Reference (under /root/reference/tamcmc/sources): build_lorentzian.cpp:61-102 (per-bin profile:
// d=x-nu; p=4*d*d/G^2; res += H*V/(1+p) [* asymmetry]), :453-459 (window accumulate);
// noise_models.cpp:17-41 (harvey_like); likelihoods.cpp:17-39; model_def.cpp:291-320.
//
// Layout: grid.x = work units (item x chain group), costliest first.  An item is a run of sub-tiles of 1024 bins
// (4 per thread, bin = b0 + j*256 + tid: coalesced): one sub-tile where truncation windows are expected,
// long runs where only the noise background exists.  Per sub-tile the CTA holds x, y in registers and
// reuses them for every chain of its group; where windows meet the sub-tile it first compacts, per chain,
// the components concerned into shared memory (order-preserving, so the accumulation order is fixed).
// Per chain: Lorentzians accumulated in registers, noise background added, y/m and the product of m
// folded into per-thread running results; one reduction per (chain, item) at the end.  The model
// spectrum never goes to HBM (unless the parity/debug pointer is set).
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
#include "tg_farfield.cuh"

#ifdef TG_PROFILE
#define TG_DBG_CTAS 8192
__device__ long long tg_dbg_cta[TG_DBG_CTAS][6];
__device__ long long tg_dbg_ff[TG_DBG_CTAS][4];   // far-field path: cycles in tg_build_lists, in tg_ff_setup, exact (quad, micro-tile) pairs, quads x 8
__device__ long long tg_dbg_tl[8];     // step timeline (globaltimer): propose start/end, finish entry/after wait/end, side start/end, previous finish end
__device__ __forceinline__ long long globaltimer_ns() { long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }
#endif

struct TgEvalArgs {
    TgDesc        d;
    TgTables      t;
    const double *x, *y, *lnx, *sig;   // [Nx]; sig may be NULL
    double2      *partials;            // [B][nitems]  (sum y/m + slow-path logs , ln of the product of m)  or (sum chi2 , 0)
    double       *model;               // [B][Nx] or NULL
    int           B, nitems, max_group;  // nitems: items (bin runs) = partial sums per chain; max_group: chains per unit at most
    int           list_cap;            // capacity (components) of one chain's list in shared memory, multiple of 4
    const double *mom;                 // [Nx / mom_bins][2][TG_MD] Chebyshev moments of the data over aligned tiles (tg_moment_tiles), or NULL
    int           mom_sub;             // sub-tiles of 1024 bins per moment tile
    int          *mom_fail;            // [Nx / mom_bins] set to 1 when a tile failed its check for some chain (scheduling hint for the host)
    int          *mom_note;            // pinned host word: set to 1 when a tile is flagged for the first time
    TgFFView      ff;                  // far-field path under the truncation windows (tg_farfield.cuh); ff.c == NULL: off
};

// ---- background-only stretches of the spectrum through Chebyshev moments of the data ------------------------------
// Outside every truncation window the model is the smooth background b(x) = N0 + sum_k H_k / (1 + (a_k x)^p_k), and over
// a tile of a few thousand bins both 1/b and ln b are polynomials of degree 8 in x to rounding error (the poles of the
// Harvey profiles are hundreds of tile widths away above a few hundred microHz).  With c, d the Chebyshev coefficients of 1/b and
// ln b on the tile, the likelihood terms of the tile are
//     sum_i y_i / b(x_i) = sum_k c_k YT_k ,   sum_i ln b(x_i) = sum_k d_k ST_k ,
// YT_k = sum_i y_i T_k(t_i), ST_k = sum_i T_k(t_i): moments of the DATA, computed once when the spectrum is loaded.  Per
// (tile, chain) that is 9 exact evaluations of b at the Chebyshev nodes instead of one per bin.  Nothing is assumed: five
// more exact evaluations at extrema of T_9 (where the interpolation error peaks) check both polynomials against b, and
// a tile whose check exceeds TG_MOM_TOL for a chain (low frequencies, exotic parameters, NaN, b <= 0) is evaluated bin by
// bin for that chain as before.
#define TG_MD      9                   // Chebyshev coefficients per function (degree 8)
#define TG_MCHK    5                   // check points
#define TG_MOM_TOL 2.0e-14
#define TG_MOM_HDR ((TG_MD + TG_MCHK) * TG_MD + TG_MD + TG_MCHK)   // doubles at the head of the moment buffer: the matrix that takes
                                       // the values at the nodes to the coefficients (TG_MD rows) and to the interpolant at the check
                                       // points (TG_MCHK rows), then the nodes and check points in [-1, 1]

struct __align__(16) TgComp {       // one Lorentzian m-component clipped to the tile, 48 bytes
    double nu, q;                   // centre, 4/Gamma^2
    double h; int lo, len;          // height, window start (tile-local), window length
    double ifc, t2;                 // asymmetry extras: 1/fc, (Gamma*asym/(2fc))^2
};
#ifndef TG_EVAL_MINBLOCKS
#define TG_EVAL_MINBLOCKS 2
#endif
#define TG_MAXG 8                   // chains per CTA (chain group) at most == warps per CTA
// dynamic shared memory of tg_k_eval for a group of G chains and a list capacity of cap components:
// lists, per-quad "covers the whole tile" flags, noise tables, per-thread partial results
// (sum, mantissa product, exponent sum)
#define TG_EVAL_SMEM(G, cap) ((size_t)(G) * ((size_t)(cap) * sizeof(TgComp) + (size_t)((cap) / 4 + 2 * TG_MAXH) * sizeof(int) + TG_NOISE_STRIDE * sizeof(double) + TG_THREADS * (2 * sizeof(double) + sizeof(int))))

// height of a component at tile-local bin li: h inside its window, 0 outside
__device__ __forceinline__ double tg_hsel(double h, int lo, int len, int li)
{
    return ((unsigned)(li - lo) < (unsigned)len) ? h : 0.0;
}

// acc[j] += sum over the list of h_k / (1 + q_k (x_j - nu_k)^2)  [* asymmetry factor], four components
// per reciprocal.  full[k/4] != 0: the four windows cover the whole tile (no per-bin predicate).
template <int BPT, bool ASYM>
__device__ __forceinline__ void tg_accumulate(int n4, const TgComp *L, const int *full, const double (&x)[BPT], double (&acc)[BPT], double asym, int tid, int li0 = 0)
{
#pragma unroll 1
    for (int k = 0; k < n4; k += 4) {
        const double2 a0 = *(const double2 *)&L[k].nu, a1 = *(const double2 *)&L[k + 1].nu, a2 = *(const double2 *)&L[k + 2].nu, a3 = *(const double2 *)&L[k + 3].nu;
        const double h0 = L[k].h, h1 = L[k + 1].h, h2 = L[k + 2].h, h3 = L[k + 3].h;
        if (!ASYM && full[k >> 2]) {             // CTA-uniform
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
            const int2 w0 = *(const int2 *)&L[k].lo, w1 = *(const int2 *)&L[k + 1].lo, w2 = *(const int2 *)&L[k + 2].lo, w3 = *(const int2 *)&L[k + 3].lo;
            double2 c0, c1, c2, c3;
            if (ASYM) { c0 = *(const double2 *)&L[k].ifc; c1 = *(const double2 *)&L[k + 1].ifc; c2 = *(const double2 *)&L[k + 2].ifc; c3 = *(const double2 *)&L[k + 3].ifc; }
#pragma unroll
            for (int j = 0; j < BPT; j++) {
                const int li = tid + j * TG_THREADS + li0;     // bin of the span the list was built for
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
__device__ __forceinline__ void tg_harvey_t(int cls, const double *q, const double (&x)[BPT], const double *lnx, const int (&ii)[BPT], double (&t)[BPT])
{
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

struct TgEvalSmem {                 // carved out of the dynamic shared memory
    TgComp *lists;                  // [G][cap]
    int    *full;                   // [G][cap/4]   1: the four windows of the quad cover the whole sub-tile
    int    *cls;                    // [G][TG_MAXH] Harvey exponent classes (ints, so the chain loop has no F2I)
    int    *nh;                     // [G] .. padded to TG_MAXH
    double *nz;                     // [G][TG_NOISE_STRIDE]
    double *sum, *pf;               // [G][TG_THREADS]  running per-thread results of the item
    int    *pe;                     // [G][TG_THREADS]
};

// Noise background, likelihood terms and the running results of chain g for the NB bins of a thread whose Lorentzian part is in
// acc (shared by every path of the fused kernel).  FF: nfold = 0xff when acc already holds the background and N0 (the far-field
// interpolant of the thread's micro-tile folded them in), 0 otherwise.
template <int NB, bool GEN, bool FF>
__device__ __forceinline__ void tg_bins_tail(const TgEvalArgs &a, const TgEvalSmem &S, int g, int c0, int b0, int tid, const double (&x)[NB], const double (&y)[NB],
                                             const int (&ii)[NB], double (&acc)[NB], int nfold)
{
    const int Nx = a.d.Nx;
    const double *nz = S.nz + (size_t)g * TG_NOISE_STRIDE;
    // ---- Harvey-like background: y += H/(1+(1e-3*tc*x)^p) ; y += N0  (noise_models.cpp:30-38).
    //      The prologue lists only the active terms (tc != 0); consecutive pairs share a reciprocal.
    // (far-field path: where the interpolant of the micro-tile already holds the background and N0, nothing is added)
    const int nH = (FF && nfold == 0xff) ? 0 : S.nh[g];
    for (int k = 0; k < nH; k += 2) {
        double ta[NB], tb[NB];
        tg_harvey_t<NB>(S.cls[g * TG_MAXH + k], nz + TG_HS * k, x, a.lnx, ii, ta);
        const double Ha = nz[TG_HS * k + TG_H_H];
        if (k + 1 < nH) {
            tg_harvey_t<NB>(S.cls[g * TG_MAXH + k + 1], nz + TG_HS * (k + 1), x, a.lnx, ii, tb);
            const double Hb = nz[TG_HS * (k + 1) + TG_H_H];
#pragma unroll
            for (int j = 0; j < NB; j++) acc[j] = fma(fma(Ha, tb[j], Hb * ta[j]), tg_rcp_fast(ta[j] * tb[j]), acc[j]);
        } else {
#pragma unroll
            for (int j = 0; j < NB; j++) acc[j] = fma(Ha, tg_rcp_fast(ta[j]), acc[j]);
        }
    }
    const double N0 = (FF && nfold == 0xff) ? 0.0 : nz[TG_N_N0];
    double m[NB];
#pragma unroll
    for (int j = 0; j < NB; j++) {
        m[j] = acc[j] + N0;
        if (GEN) {
            const int idx = b0 + j * TG_THREADS + tid;
            if (idx < Nx) { if (a.model) a.model[(size_t)(c0 + g) * Nx + idx] = m[j]; }
            else m[j] = 1.0;
        }
    }
    // running results of this thread for chain g over the whole item
    double sum = S.sum[g * TG_THREADS + tid];     // sum y/m (or chi-square) + logs taken the slow way
    double pf = S.pf[g * TG_THREADS + tid];       // product of the model values: pf * 2^pe, pf in [1,2)
    int pe = S.pe[g * TG_THREADS + tid];
    if (a.d.like_case == 0) {                 // likelihoods.cpp:23
#pragma unroll
        for (int j0 = 0; j0 < NB; j0 += 4) {
            double mm[4], yy[4];
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const bool in = (j0 + u < NB);
                mm[u] = in ? m[in ? (j0 + u) : 0] : 1.0;
                yy[u] = in ? y[in ? (j0 + u) : 0] : 0.0;
            }
            const double p01 = mm[0] * mm[1], p23 = mm[2] * mm[3], P = p01 * p23;
            // all four positive (sign bits clear) and the product a normal number far from the range ends
            const int sgn = __double2hiint(mm[0]) | __double2hiint(mm[1]) | __double2hiint(mm[2]) | __double2hiint(mm[3]);
            const int hiP = __double2hiint(P);
            const int eP = (hiP >> 20) & 0x7ff;
            if (sgn >= 0 && (unsigned)(eP - 200) < 1646u) {            // 2^-823 < P < 2^823
                const double n = fma(fma(yy[0], mm[1], yy[1] * mm[0]), p23, fma(yy[2], mm[3], yy[3] * mm[2]) * p01);
                sum = fma(n, tg_rcp_fast(P), sum);
                // fold P into the running product; the exponent is kept apart so it can never overflow
                pf *= __hiloint2double((hiP & 0x000fffff) | 0x3ff00000, __double2loint(P));      // [1,2) x [1,2)
                const int h2 = __double2hiint(pf);
                pe += (eP - 1023) + (((h2 >> 20) & 0x7ff) - 1023);
                pf = __hiloint2double((h2 & 0x000fffff) | 0x3ff00000, __double2loint(pf));
            } else {                          // zero / negative / NaN / huge model values: the plain way
#pragma unroll
                for (int u = 0; u < 4; u++) if (j0 + u < NB && (!GEN || b0 + (j0 + u) * TG_THREADS + tid < Nx)) sum += yy[u] * tg_rcp(mm[u]) + log(mm[u]);
            }
        }
    } else {                                  // likelihoods.cpp:36
#pragma unroll
        for (int j = 0; j < NB; j++) {
            const int idx = b0 + j * TG_THREADS + tid;
            if (!GEN || idx < Nx) {
                const double sg = __ldg(a.sig + idx);
                const double dd = y[j] - m[j];
                sum += (dd * dd) * tg_rcp(sg * sg);
            }
        }
    }
    S.sum[g * TG_THREADS + tid] = sum; S.pf[g * TG_THREADS + tid] = pf; S.pe[g * TG_THREADS + tid] = pe;
}

// Every chain of the group on NB*TG_THREADS consecutive bins starting at b0, held in registers; no CTA
// barrier inside, so warps drift apart and hide each other's latencies.  LISTS: the sub-tile meets
// truncation windows (lists in shared memory); otherwise only the noise background exists there and
// NB may be 8 (two sub-tiles at once: twice the instruction-level parallelism, half the per-chain
// overhead).  GEN = false is the interior fast path (all bins inside the spectrum, no model output, no
// Gaussian test model): no index clamps, no selects.
template <int NB, bool GEN, bool LISTS>
__device__ __forceinline__ void tg_chain_loop(const TgEvalArgs &a, const TgEvalSmem &S, const int *s_tot, int G, int c0, int b0, int tid, unsigned gmask = 0xffffffffu)
{
    const int Nx = a.d.Nx, cap = a.list_cap;
    double x[NB], y[NB];
    int ii[NB];
#pragma unroll
    for (int j = 0; j < NB; j++) {
        const int idx = b0 + j * TG_THREADS + tid;
        ii[j] = (!GEN || idx < Nx) ? idx : Nx - 1;
        x[j] = __ldg(a.x + ii[j]);
        y[j] = (!GEN || idx < Nx) ? __ldg(a.y + ii[j]) : 0.0;   // bins past the end: y = 0 and (below) m = 1
    }
    for (int g = 0; g < G; g++) {
        if (!((gmask >> g) & 1u)) continue;       // CTA-uniform: chains of the group this call is not for
        const double *nz = S.nz + (size_t)g * TG_NOISE_STRIDE;
        double acc[NB];
        if (GEN && nz[TG_N_GAUSS] != 0.0) {       // models.cpp:1977-1978 / :2030-2031 (Gaussian test models always take GEN)
            const double gA = nz[TG_N_GA], gnu = nz[TG_N_GNU], gs2 = nz[TG_N_GS2];
#pragma unroll
            for (int j = 0; j < NB; j++) { const double dd = x[j] - gnu; acc[j] = gA * exp(-0.5 * (dd * dd) / gs2); }
        } else {
#pragma unroll
            for (int j = 0; j < NB; j++) acc[j] = 0.0;
        }
        if constexpr (LISTS) {
            const int n4 = (s_tot[g] + 3) & ~3;
            const double asym = nz[TG_N_ASYM];
            if (n4 > 0) {
                if (asym != 0.0) tg_accumulate<NB, true>(n4, S.lists + (size_t)g * cap, S.full + g * (cap >> 2), x, acc, asym, tid);
                else             tg_accumulate<NB, false>(n4, S.lists + (size_t)g * cap, S.full + g * (cap >> 2), x, acc, asym, tid);
            }
        }
        tg_bins_tail<NB, GEN, false>(a, S, g, c0, b0, tid, x, y, ii, acc, 0);
    }
}

// npair consecutive pairs of sub-tiles starting at bin b0, all of them inside the range the far-field pre-pass prepared, for the
// G chains of the unit: per (pair, chain) the staged interpolant + exact records of the warp's micro-tile (tg_farfield.cuh), then
// the common tail.  What the pre-pass left travels from L2 one (pair, chain) ahead of the arithmetic, warp by warp: no CTA barrier.
__device__ __forceinline__ void tg_ff_run(const TgEvalArgs &a, const TgEvalSmem &S, int G, int c0, int b0, int npair, int tid, TgFFStage &st)
{
    const int lane = tid & 31, warp = tid >> 5;
    TgFFFetch q;
    tg_ff_fetch(a.ff, c0, b0, warp, lane, q);
    const int nstep = npair * G;
    int g = 0, pr = 0;
    double x[TG_FF_MT], y[TG_FF_MT];
    int ii[TG_FF_MT];
#ifdef TG_PROFILE
    long long pk[3] = {0, 0, 0};
#endif
    for (int s = 0; s < nstep; s++) {
        const int bp = b0 + pr * 2 * TG_THREADS * 4, bw = bp + warp * TG_FF_WBINS;
        if (g == 0) {
#pragma unroll
            for (int j = 0; j < TG_FF_MT; j++) { ii[j] = bw + 32 * j + lane; x[j] = __ldg(a.x + ii[j]); y[j] = __ldg(a.y + ii[j]); }
        }
#ifdef TG_PROFILE
        const long long pk0 = clock64();
#endif
        __syncwarp();                                 // the lanes are done with the previous step's data
        tg_ff_stage(st, warp, lane, q);
        __syncwarp();
        int gn = g + 1, pn = pr;
        if (gn == G) { gn = 0; pn++; }
        if (s + 1 < nstep) tg_ff_fetch(a.ff, c0 + gn, b0 + pn * 2 * TG_THREADS * 4, warp, lane, q);
        double acc[TG_FF_MT];
#ifdef TG_PROFILE
        const long long pk1 = clock64();
#endif
        const bool folded = tg_ff_microtile(st, a.t, a.d.ncomp_max, c0 + g, bw, warp, lane, x, acc);
#ifdef TG_PROFILE
        const long long pk2 = clock64();
#endif
        tg_bins_tail<TG_FF_MT, false, true>(a, S, g, c0, bp, tid, x, y, ii, acc, folded ? 0xff : 0);
#ifdef TG_PROFILE
        pk[0] += pk1 - pk0; pk[1] += pk2 - pk1; pk[2] += clock64() - pk2;
#endif
        g = gn; pr = pn;
    }
#ifdef TG_PROFILE
    if (lane == 0 && blockIdx.x < TG_DBG_CTAS) {      // [0] sum over warps of the cycles in the steps, [1] the slowest warp's, [2] sum over warps of the staging waits, [3] steps
        atomicAdd((unsigned long long *)&tg_dbg_ff[blockIdx.x][0], (unsigned long long)(pk[0] + pk[1] + pk[2]));
        atomicMax((unsigned long long *)&tg_dbg_ff[blockIdx.x][1], (unsigned long long)(pk[0] + pk[1] + pk[2]));
        atomicAdd((unsigned long long *)&tg_dbg_ff[blockIdx.x][2], (unsigned long long)pk[0]);
        if (tid == 0) tg_dbg_ff[blockIdx.x][3] += nstep;
    }
#endif
}

// Exact background at one abscissa (noise table of one chain in shared memory): N0 + sum_k H_k / (1 + (a_k x)^p_k)
__device__ __noinline__ double tg_background_at(const double *nz, const int *cls, int nH, double x)
{
    double acc = 0.0;
    for (int k = 0; k < nH; k++) {
        const double *q = nz + TG_HS * k;
        const double ap = q[TG_H_AP];
        double t;
        switch (cls[k]) {
        case 0: t = 2.0; break;
        case 1: t = fma(ap, x, 1.0); break;
        case 2: t = fma(ap * x, x, 1.0); break;
        case 3: t = fma((ap * x) * x, x, 1.0); break;
        case 4: { const double x2 = x * x; t = fma(ap * x2, x2, 1.0); break; }
        default: t = tg_exp_nb(q[TG_H_P] * (q[TG_H_LNA] + log(x))) + 1.0; break;
        }
        acc = fma(q[TG_H_H], tg_rcp(t), acc);
    }
    return acc + nz[TG_N_N0];
}

// n_mt background-only moment tiles starting at bin b0 (aligned) for the G chains of the unit: warp w takes the pairs
// (tile, chain) = w, w + 8, ...; lanes 0..8 evaluate the background at the nodes, lanes 9..13 at the check points.  A pair
// that passes adds its two sums to the running results of the chain (the slot of S.sum of its first lane, s_lw[g][half-warp]); a pair that
// fails sets bit g of s_fail[tile] and is evaluated bin by bin by the caller.  No barrier inside.
__device__ void tg_moment_tiles(const TgEvalArgs &a, const TgEvalSmem &S, const double (*s_WE)[TG_MD], const double *s_tn,
                                double (*s_lw)[2 * (TG_THREADS / 32)], unsigned *s_fail, unsigned *s_flag, int G, int b0, int n_mt)
{
    // Two (tile, chain) pairs per warp and round, one per half-warp (14 of its 16 lanes work: 9 nodes + 5 check points) -- the
    // rounds are chains of latencies (background, reciprocal and logarithm at a node, a 9-term product by shuffles, two reductions),
    // so two of them side by side cost one.
    static_assert(TG_MD + TG_MCHK <= 16, "nodes + check points of a pair fit a half-warp");
    constexpr int NW = TG_THREADS / 32;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, half = lane >> 4, hl = lane & 15;
    const unsigned hmask = 0xffffu << (16 * half);
    const int mtb = a.mom_sub * (TG_THREADS * 4);
    const int npairs = n_mt * G;
    for (int p0 = 2 * warp; p0 < npairs; p0 += 2 * NW) {
        const int p = p0 + half;
        const bool active = (p < npairs);
        const int pc = active ? p : p0;                          // an idle half-warp shadows its neighbour (it takes part in the shuffles, writes nothing)
        const int mt = pc / G, g = pc - mt * G;
        const int bb = b0 + mt * mtb;
        const double xa = __ldg(a.x + bb), xb = __ldg(a.x + bb + mtb - 1);
        const double xm = 0.5 * (xa + xb), xh = 0.5 * (xb - xa);
        const double *mo = a.mom + TG_MOM_HDR + (size_t)(bb / mtb) * 2 * TG_MD;
        double r = 0.0, l = 0.0, mom_s = 0.0, mom_y = 0.0;
        bool ok = (xh > 0.0), bval = true;
        if (hl < TG_MD) { mom_s = __ldg(mo + hl); mom_y = __ldg(mo + TG_MD + hl); }
        if (hl == 0) ok = ok && (mom_s == (double)mtb);         // ST_0 = number of bins; NaN marks a tile the host could not use
        if (hl < TG_MD + TG_MCHK) {
            const double b = tg_background_at(S.nz + (size_t)g * TG_NOISE_STRIDE, S.cls + g * TG_MAXH, S.nh[g], fma(xh, s_tn[hl], xm));
            bval = (b > 0.0) && (b < 1.0e300);                  // false for NaN
            ok = ok && bval;
            r = tg_rcp(b); l = log(b);
        }
        // a background that is not a positive finite number at the nodes (NaN parameters of a failed proposal, ...) is no
        // statement about the tile: it is evaluated bin by bin like any failure, but the tile is not flagged for the host
        const bool sane = ((__ballot_sync(0xffffffffu, bval) & hmask) == hmask);
        const int row = (hl < TG_MD + TG_MCHK) ? hl : 0;
        double cr = 0.0, cl = 0.0;                              // lanes 0..8 of the half: coefficients; 9..13: interpolants at the check points
#pragma unroll
        for (int j = 0; j < TG_MD; j++) {
            const double w = s_WE[row][j];
            cr = fma(w, __shfl_sync(0xffffffffu, r, j, 16), cr);
            cl = fma(w, __shfl_sync(0xffffffffu, l, j, 16), cl);
        }
        if (hl >= TG_MD && hl < TG_MD + TG_MCHK)
            ok = ok && (fabs(cr - r) <= TG_MOM_TOL * fabs(r)) && (fabs(cl - l) <= TG_MOM_TOL * fmax(1.0, fabs(l)));
        double ps = (hl < TG_MD) ? cr * mom_y : 0.0, pl = (hl < TG_MD) ? cl * mom_s : 0.0;
#pragma unroll
        for (int o = 8; o > 0; o >>= 1) { ps += __shfl_down_sync(0xffffffffu, ps, o, 16); pl += __shfl_down_sync(0xffffffffu, pl, o, 16); }
        const bool all_ok = ((__ballot_sync(0xffffffffu, ok) & hmask) == hmask);
        if (hl == 0 && active) {
            if (all_ok) { S.sum[g * TG_THREADS + tid] += ps; s_lw[g][2 * warp + half] += pl; }
            else { atomicOr(s_fail + mt, 1u << g); if (sane) atomicOr(s_flag + mt, 1u << g); }
        }
    }
}

// Per-chain lists of the components whose window meets the tile [b0, b0+SUB) (SUB = one or two sub-tiles): order-preserving
// ballot compaction for all chains of the group at once, padded to a multiple of four, plus the
// per-quad "covers the whole sub-tile" flags.  Called by the whole CTA (contains barriers).
__device__ void tg_build_lists(const TgEvalArgs &a, const TgEvalSmem &S, int (*s_cnt)[TG_THREADS / 32], int *s_tot, int G, int c0, int b0, const int SUB)
{
    constexpr int NW = TG_THREADS / 32;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, cap = a.list_cap;
    __syncthreads();                              // the previous sub-tile's lists are dead
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
                lo[g] = a.t.lo[e] - b0; hi[g] = a.t.hi[e] - b0;
            }
        }
#pragma unroll
        for (int g = 0; g < TG_MAXG; g++) {
            const bool ov = (lo[g] < SUB) && (hi[g] > 0) && (hi[g] > lo[g]);
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
                S.lists[(size_t)g * cap + off + __popc(bal[g] & ((1u << lane) - 1u))] = q;
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
            if (total + u < n4) { TgComp q; q.nu = 0; q.q = 0; q.h = 0; q.lo = -(1 << 29); q.len = 1 << 30; q.ifc = 0; q.t2 = 0; S.lists[(size_t)g * cap + total + u] = q; }
        }
    }
    __syncthreads();
    for (int e = tid; e < G * (cap >> 2); e += TG_THREADS) {
        const int g = e / (cap >> 2), qd = e - g * (cap >> 2);
        int f = 0;
        if (4 * qd < s_tot[g]) {
            f = 1;
#pragma unroll
            for (int u = 0; u < 4; u++) { const TgComp *q = S.lists + (size_t)g * cap + 4 * qd + u; if (!(q->lo <= 0 && q->lo + q->len >= SUB)) f = 0; }
        }
        S.full[e] = f;
    }
    __syncthreads();
}

// The fused kernel.  grid.x = work units = (item, chain group); an item is a run of sub-tiles of 1024 bins:
// ONE sub-tile where truncation windows are expected (few chains per unit), long runs in the noise-only
// parts of the spectrum (more chains per unit).  The unit table is a scheduling hint only: every sub-tile is classified again from the per-chain
// window unions of THIS launch, so any partition of the bins gives the same results.
// units [grid]: work units in launch order, (first bin, sub-tiles | chains << 16, first chain, item | star << TG_UNIT_STAR_SHIFT);
// argv [stars]: the argument record of every star the launch spans (one for a single handle).
#define TG_UNIT_STAR_SHIFT 21
__global__ void __launch_bounds__(TG_THREADS, TG_EVAL_MINBLOCKS) tg_k_eval(const TgEvalArgs *__restrict__ argv, const int4 *__restrict__ units)
{
    constexpr int SUB = TG_THREADS * 4;
    constexpr int NW = TG_THREADS / 32;
    extern __shared__ __align__(16) unsigned char tg_eval_smem[];
    __shared__ int s_cnt[TG_MAXG][NW];
    __shared__ int s_tot[TG_MAXG];
    __shared__ int s_w[2][TG_MAXG];
    __shared__ double s_WE[TG_MD + TG_MCHK][TG_MD];   // moment path: node values -> coefficients / check interpolants
    __shared__ double s_tn[TG_MD + TG_MCHK];
    __shared__ double s_lw[TG_MAXG][2 * NW];          // moment path: sum of ln b per chain and half-warp
    __shared__ unsigned s_fail[32], s_flag[32];       // moment path: chains that take the tile bin by bin / that flag it for the host

    __shared__ __align__(16) TgEvalArgs s_args;
    __shared__ TgFFStage s_ffstage;                  // far-field path: slot w of every array belongs to warp w
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int4 unit = __ldg(units + blockIdx.x);
    tg_load_args(&s_args, argv + ((unsigned)unit.w >> TG_UNIT_STAR_SHIFT));
    const TgEvalArgs &a = s_args;
    const int item_index = unit.w & ((1 << TG_UNIT_STAR_SHIFT) - 1);
    const int2 item = make_int2(unit.x, unit.y & 0xffff);   // (first bin, number of sub-tiles)
#ifdef TG_PROFILE
    const int dbg_id = blockIdx.x;
    if (tid == 0 && dbg_id < TG_DBG_CTAS) { tg_dbg_ff[dbg_id][0] = tg_dbg_ff[dbg_id][1] = tg_dbg_ff[dbg_id][2] = tg_dbg_ff[dbg_id][3] = 0; unsigned sm; asm volatile("mov.u32 %0, %smid;" : "=r"(sm)); tg_dbg_cta[dbg_id][0] = clock64(); tg_dbg_cta[dbg_id][3] = ((long long)sm << 32) | (unsigned)item.x; tg_dbg_cta[dbg_id][4] = globaltimer_ns(); }
#endif
    const int cap = a.list_cap, cpb = a.max_group, Nx = a.d.Nx;
    const int c0 = unit.z;
    const int G = unit.y >> 16;
    TgEvalSmem S;
    S.lists = (TgComp *)tg_eval_smem;
    S.nz    = (double *)(S.lists + (size_t)cpb * cap);
    S.sum   = S.nz + (size_t)cpb * TG_NOISE_STRIDE;
    S.pf    = S.sum + (size_t)cpb * TG_THREADS;
    S.pe    = (int *)(S.pf + (size_t)cpb * TG_THREADS);
    S.full  = S.pe + (size_t)cpb * TG_THREADS;
    S.cls   = S.full + (size_t)cpb * (cap >> 2);
    S.nh    = S.cls + (size_t)cpb * TG_MAXH;

    // ---- phase 0: noise tables, window unions, running results ----
    for (int g = 0; g < G; g++) { S.sum[g * TG_THREADS + tid] = 0.0; S.pf[g * TG_THREADS + tid] = 1.0; S.pe[g * TG_THREADS + tid] = 0; }
    const bool use_mom = (a.mom != nullptr) && (a.d.like_case == 0) && (a.model == nullptr) && (a.d.model_case > 1);
    if (tid < TG_MAXG * 2 * NW) (&s_lw[0][0])[tid] = 0.0;
    bool we_loaded = false;                       // the interpolation matrices are staged by the first moment tile of the unit (most units have none)
    const bool ff_on = (a.ff.c != nullptr) && (a.model == nullptr);
    tg_pdl_wait();                                // everything above (unit table, shared-memory set-up) may overlap the proposal kernel
    tg_pdl_trigger();
    for (int e = tid; e < G * TG_NOISE_STRIDE; e += TG_THREADS) S.nz[e] = a.t.noise[(size_t)c0 * TG_NOISE_STRIDE + e];
    if (tid < G) { s_w[0][tid] = a.t.wlo[c0 + tid]; s_w[1][tid] = a.t.whi[c0 + tid]; }
    __syncthreads();
    if (tid < TG_MAXG * TG_MAXH) {                // Harvey exponent classes / counts as ints
        const int g = tid / TG_MAXH, k = tid % TG_MAXH;
        if (g < G) {
            S.cls[g * TG_MAXH + k] = (int)S.nz[(size_t)g * TG_NOISE_STRIDE + TG_HS * k + TG_H_CLS];
            if (k == 0) S.nh[g] = (int)S.nz[(size_t)g * TG_NOISE_STRIDE + TG_N_NH];
        }
    }
    int wlo = 0x7fffffff, whi = 0;                // union of the group's windows
    for (int g = 0; g < G; g++) if (s_w[1][g] > s_w[0][g]) { wlo = min(wlo, s_w[0][g]); whi = max(whi, s_w[1][g]); }
    const bool gen_all = (a.model != nullptr) || (a.d.model_case <= 1);
    bool sym_all = true;                          // no chain of the unit has an asymmetric profile
    for (int g = 0; g < G; g++) sym_all = sym_all && (S.nz[(size_t)g * TG_NOISE_STRIDE + TG_N_ASYM] == 0.0);
    __syncthreads();
#ifdef TG_PROFILE
    if (tid == 0 && dbg_id < TG_DBG_CTAS) tg_dbg_cta[dbg_id][1] = clock64();
#endif

    // ---- phases 1+2 per sub-tile ----
    for (int sub = 0; sub < item.y; ) {
        const int b0 = item.x + sub * SUB;
        if (b0 >= Nx) break;
        const bool gen = gen_all || (b0 + SUB > Nx);
        if (use_mom) {                            // aligned background-only tiles ahead: through the moments of the data
            const int mtb = a.mom_sub * SUB;
            int n_mt = 0;
            if (b0 % mtb == 0)
                while (n_mt < 32 && sub + (n_mt + 1) * a.mom_sub <= item.y && b0 + (n_mt + 1) * mtb <= Nx && !(wlo < b0 + (n_mt + 1) * mtb && whi > b0 + n_mt * mtb)) n_mt++;
            if (n_mt > 0) {
                __syncthreads();                  // s_fail of the previous round has been read
                if (tid < 32) { s_fail[tid] = 0u; s_flag[tid] = 0u; }
                if (!we_loaded) {                 // constants of the spectrum, not of the step
                    if (tid < (TG_MD + TG_MCHK) * TG_MD) (&s_WE[0][0])[tid] = a.mom[tid];
                    if (tid < TG_MD + TG_MCHK) s_tn[tid] = a.mom[(TG_MD + TG_MCHK) * TG_MD + tid];
                    we_loaded = true;
                }
                __syncthreads();
                tg_moment_tiles(a, S, s_WE, s_tn, s_lw, s_fail, s_flag, G, b0, n_mt);
                __syncthreads();
                for (int t = 0; t < n_mt; t++) {
                    const unsigned fm = s_fail[t];
                    if (s_flag[t] && tid == 0 && atomicExch(a.mom_fail + b0 / mtb + t, 1) == 0) *(volatile int *)a.mom_note = 1;
                    if (fm) for (int s2 = 0; s2 < a.mom_sub; s2++) tg_chain_loop<4, false, false>(a, S, s_tot, G, c0, b0 + (t * a.mom_sub + s2) * SUB, tid, fm);
                }
                sub += n_mt * a.mom_sub;
                continue;
            }
        }
        if (!gen_all && sub + 1 < item.y && b0 + 2 * SUB <= Nx && wlo < b0 + 2 * SUB && whi > b0) {
            // truncation windows meet this pair of sub-tiles: 8 bins per thread against one list (half the list traffic and
            // loop overhead per bin, twice the independent work per warp behind every reciprocal)
            // inside the range the pre-pass covered (and no chain of the unit with an asymmetric profile): no lists at all, and
            // every following pair of the item that lies inside it too (whether windows meet it or not) in the same run
            if (ff_on && sym_all && (b0 >> TG_FF_SHIFT) >= a.ff.mt0 && (b0 >> TG_FF_SHIFT) + TG_FF_PAIR_MT <= a.ff.mt0 + a.ff.nmt) {
                int npair = 1;
                while (sub + 2 * (npair + 1) <= item.y && b0 + 2 * (npair + 1) * SUB <= Nx && (b0 >> TG_FF_SHIFT) + (npair + 1) * TG_FF_PAIR_MT <= a.ff.mt0 + a.ff.nmt) npair++;
                __syncthreads();                  // the staging buffers of an earlier run are dead
                tg_ff_run(a, S, G, c0, b0, npair, tid, s_ffstage);
                sub += 2 * npair;
            } else {
                tg_build_lists(a, S, s_cnt, s_tot, G, c0, b0, 2 * SUB);
                tg_chain_loop<8, false, true>(a, S, s_tot, G, c0, b0, tid);
                sub += 2;
            }
        } else if (wlo < b0 + SUB && whi > b0) {  // truncation windows meet this sub-tile
            tg_build_lists(a, S, s_cnt, s_tot, G, c0, b0, SUB);
            if (gen) tg_chain_loop<4, true, true>(a, S, s_tot, G, c0, b0, tid);
            else     tg_chain_loop<4, false, true>(a, S, s_tot, G, c0, b0, tid);
            sub += 1;
        } else if (!gen_all && sub + 1 < item.y && b0 + 2 * SUB <= Nx && !(wlo < b0 + 2 * SUB && whi > b0 + SUB)) {
            tg_chain_loop<8, false, false>(a, S, s_tot, G, c0, b0, tid);    // two noise-only sub-tiles at once
            sub += 2;
        } else {
            if (gen) tg_chain_loop<4, true, false>(a, S, s_tot, G, c0, b0, tid);
            else     tg_chain_loop<4, false, false>(a, S, s_tot, G, c0, b0, tid);
            sub += 1;
        }
    }
    __syncthreads();

    // ---- phase 3: warp g reduces chain g (fixed order: 8 strided entries per lane, shuffle tree),
    //      lane 0 takes ONE log of the mantissa product; all chains of the group in parallel ----
    if (warp < G) {
        double sum = 0.0, pf = 1.0;
        int pe = 0;
#pragma unroll
        for (int i = 0; i < NW; i++) {
            sum += S.sum[warp * TG_THREADS + i * 32 + lane];
            pf *= S.pf[warp * TG_THREADS + i * 32 + lane];           // 8 mantissas in [1,2): < 2^8
            pe += S.pe[warp * TG_THREADS + i * 32 + lane];
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {                           // 32 x 2^8 -> < 2^256
            sum += __shfl_down_sync(0xffffffffu, sum, o);
            pf *= __shfl_down_sync(0xffffffffu, pf, o);
            pe += __shfl_down_sync(0xffffffffu, pe, o);
        }
        if (lane == 0) {
            double t2 = 0.0;
            if (a.d.like_case == 0) {
                t2 = fma((double)pe, 6.93147180369123816490e-01, fma((double)pe, 1.90821492927058770002e-10, log(pf)));
#pragma unroll
                for (int w = 0; w < 2 * NW; w++) t2 += s_lw[warp][w];
            }
            a.partials[(size_t)(c0 + warp) * a.nitems + item_index] = make_double2(sum, t2);
        }
    }
#ifdef TG_PROFILE
    if (tid == 0 && dbg_id < TG_DBG_CTAS) { tg_dbg_cta[dbg_id][2] = clock64(); tg_dbg_cta[dbg_id][5] = globaltimer_ns(); }
#endif
}

// Sums the per-tile pairs of each chain in a fixed order (thread-strided serial sums, then a
// shared-memory tree) -> sums[c] = (sum y/m, sum ln m).  In the bin-sharded variant this is the
// buffer that is all-reduced across GPUs.
// Fixed-order sum of one pair per thread over a CTA of 256 (used by every kernel that totals the
// partials of a chain, so that they agree to the bit): xor-shuffle tree inside each warp, then the eight
// warp sums in order.  One barrier; s_part is [8] shared double2.
__device__ __forceinline__ double2 tg_block_sum2(double a1, double a2, double2 *s_part)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { a1 += __shfl_xor_sync(0xffffffffu, a1, o); a2 += __shfl_xor_sync(0xffffffffu, a2, o); }
    if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = make_double2(a1, a2);
    __syncthreads();
    double t1 = 0, t2 = 0;
#pragma unroll
    for (int w = 0; w < TG_THREADS / 32; w++) { const double2 v = s_part[w]; t1 += v.x; t2 += v.y; }
    return make_double2(t1, t2);
}

__global__ void __launch_bounds__(256) tg_k_reduce(const double2 *partials, int ntiles, double2 *sums)
{
    __shared__ double2 s_part[8];
    const int c = blockIdx.x, tid = threadIdx.x;
    double a1 = 0, a2 = 0;
    for (int t = tid; t < ntiles; t += 256) { const double2 v = partials[(size_t)c * ntiles + t]; a1 += v.x; a2 += v.y; }
    const double2 sm = tg_block_sum2(a1, a2, s_part);
    if (tid == 0) sums[c] = sm;
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
    int st = t.status[c] | t.pstatus[c];
    const double L = tg_logl_from_sums(d, sums[c], T[c], st);
    if (L != L) st |= TG_ST_NAN_LOGL;
    logL[c] = L; logPrior[c] = t.logprior[c]; logPost[c] = L + t.logprior[c]; status[c] = st;
}

// tg_k_reduce + tg_k_finalize_eval in one launch (the tg_eval call path): CTA c reduces the partials of
// chain c in the same fixed order as tg_k_reduce, thread 0 finishes the chain.
__global__ void __launch_bounds__(256) tg_k_reduce_finalize(TgDesc d, const double2 *partials, int ntiles, double2 *sums, const double *T, TgTables t,
                                                            double *logL, double *logPrior, double *logPost, int *status)
{
    __shared__ double2 s_part[8];
    const int c = blockIdx.x, tid = threadIdx.x;
    tg_pdl_wait();                                // (a no-op unless launched with the programmatic-serialization attribute)
    double a1 = 0, a2 = 0;
    for (int k = tid; k < ntiles; k += 256) { const double2 v = partials[(size_t)c * ntiles + k]; a1 += v.x; a2 += v.y; }
    const double2 sm = tg_block_sum2(a1, a2, s_part);
    if (tid == 0) {
        sums[c] = sm;
        int st = t.status[c] | t.pstatus[c];
        const double L = tg_logl_from_sums(d, sm, T[c], st);
        if (L != L) st |= TG_ST_NAN_LOGL;
        logL[c] = L; logPrior[c] = t.logprior[c]; logPost[c] = L + t.logprior[c]; status[c] = st;
    }
}

__global__ void tg_k_lnx(const double *x, double *lnx, int n)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) lnx[i] = log(x[i]);
}
