// tg_farfield.cuh -- far field of the truncated Lorentzian sums (the part of the fused kernel under the truncation windows).
//
// Reference arithmetic being reproduced: build_lorentzian.cpp:61-102 (profile h/(1+4(x-nu)^2/Gamma^2)), :365-460 (window
// [imin, imax) of every m-component).  Inside the windows ~30 components overlap on every bin, but all except one or two are
// many line widths away and vary slowly across a few hundred bins.  A micro-tile is TG_FF_BINS = 512 consecutive bins (half-width w, centre
// xc).  A component whose window covers the micro-tile and whose poles nu +- i Gamma/2 lie at least R w from xc is analytic
// inside the Bernstein ellipse of parameter rho >= R + sqrt(R^2 - 1) of the micro-tile: its interpolant at TG_FF_N = 16
// Chebyshev nodes is exact to rounding (<= 7e-15 of the component's own value at R = 5, 2e-15 at R = 6; R^2 = ff_r2 = 30.25).
//
//   pre-pass  tg_k_ff   one warp per (chain, micro-tile): classifies the chain's components (far / exact / not there), sums
//                       the far ones (+ the noise background where its poles are far too) at the 16 nodes, turns the node
//                       values into Chebyshev coefficients, checks the trailing coefficients, and leaves in HBM per
//                       (chain, micro-tile): 16 coefficients, a header, and the records (nu, 4/Gamma^2, height, window) of the
//                       components to be evaluated exactly there (near ones, and those whose window edge cuts the micro-tile);
//                       the coefficients converted to powers of the local variable (tg_ff_to_powers);
//   fused kernel        per bin: one Horner recurrence (15 FMA) + the exact records of its micro-tile, window predicate where a
//                       window edge cuts it -- the windows stay bin-exact -- instead of ~30 profile evaluations (6.25 each);
//                       a warp takes 256 bins (half a micro-tile) per step, eight per lane.
//
// Nothing is assumed about the parameters: the classification is a function of (nu, Gamma, window) and the grid alone; a
// micro-tile whose trailing coefficients are not small (|c14| + |c15| > TG_FF_TOL |c0|: NaN parameters, ...) or that has more
// exact components than a record holds is marked and evaluated by the fused kernel component by component from the chain's
// table (slow, exact, never taken in the tests' workloads).  Chains with an asymmetric profile, spectra whose abscissae do not
// increase, and the model-output (parity / getmodel) calls do not use the path at all.
#pragma once
#include "tg_internal.cuh"

#define TG_FF_N    16                  // Chebyshev nodes / coefficients per micro-tile
#define TG_FF_MT   8                   // micro-tiles per group of the pre-pass (one per warp of its CTAs) == bins per lane of the fused kernel there
#ifndef TG_FF_BINS
#define TG_FF_BINS 512                 // bins per micro-tile (256 or 512)
#endif
#define TG_FF_SHIFT (TG_FF_BINS == 512 ? 9 : 8)      // global micro-tile index = bin >> TG_FF_SHIFT
#define TG_FF_WBINS 256                // bins a warp of the fused kernel takes per step (8 per lane)
#define TG_FF_WPM  (TG_FF_BINS / TG_FF_WBINS)        // warps of the fused kernel per micro-tile
#define TG_FF_PAIR_MT (2048 / TG_FF_BINS)            // micro-tiles of a pair of sub-tiles
static_assert(TG_FF_BINS == 256 || TG_FF_BINS == 512, "micro-tiles of 256 or 512 bins");
#define TG_FF_E    32                  // exact components a micro-tile record holds
#define TG_FF_TOL  1.0e-12
#define TG_FF_HDR  (2 * TG_FF_N * TG_FF_N + TG_FF_N)   // doubles of the constant buffer: node values -> Chebyshev coefficients [k][i], the nodes,
                                                        // then the Chebyshev -> power-basis matrix M[k][j] (T_j = sum_k M[k][j] t^k: integers, exact)
#define TG_FF_FALLBACK (-1)

struct __align__(16) TgFFRec { double nu, q, h; int lo, len; };            // 32 bytes; window [lo, lo+len) in LOCAL bin indices
struct __align__(16) TgFFHdr { int count; int nfree; int flags; int nfar; };       // count: records, or TG_FF_FALLBACK; the first nfree of them cover the
                                                                            // whole micro-tile (no window predicate), the rest need it; flags bit 0: background + N0 folded in
struct __align__(16) TgFFGeo { double xc, hw, ihw, pad; };                 // centre, half-width, 1/half-width of a micro-tile (constants of the spectrum)

struct TgFFView {                      // what the fused kernel reads (part of its per-star argument record)
    const double  *c;                  // [B][stride][TG_FF_N]
    const TgFFHdr *hdr;                // [B][stride]
    const TgFFRec *rec;                // [B][stride][TG_FF_E]
    const TgFFGeo *geo;                // [Nx / TG_FF_BINS]  (global micro-tile index = bin >> TG_FF_SHIFT)
    int            stride;             // micro-tiles per chain the buffers hold
    int            mt0, nmt;           // micro-tiles [mt0, mt0 + nmt) were prepared by the pre-pass of this launch
    int            pad_;
};

struct TgFFArgs {                      // per-star argument record of the pre-pass
    TgDesc         d;
    TgTables       t;
    const double  *ffw;                // constant buffer (TG_FF_HDR doubles)
    double         r2, rs;             // (pole distance / half-width)^2 from which a component is far, and its square root
    double        *c; TgFFHdr *hdr; TgFFRec *rec; const TgFFGeo *geo;
    int            stride, mt0, nmt, B;
};

__global__ void tg_k_ffgeo(const double *x, int Nx, TgFFGeo *geo, int nmt)
{
    const int m = blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= nmt) return;
    TgFFGeo g; g.xc = 0.0; g.hw = 0.0; g.ihw = 0.0; g.pad = 0.0;
    if ((m + 1) * TG_FF_BINS <= Nx) {
        const double xa = x[m * TG_FF_BINS], xb = x[m * TG_FF_BINS + TG_FF_BINS - 1];
        g.xc = 0.5 * (xa + xb); g.hw = 0.5 * (xb - xa); g.ihw = (g.hw > 0.0) ? 1.0 / g.hw : 0.0;
    }
    geo[m] = g;
}

// Exact background at one abscissa from a noise table in shared memory (same arithmetic as the fused kernel's bin-by-bin path up
// to the shared reciprocals): N0 + sum_k H_k / (1 + (a_k x)^p_k)
__device__ __forceinline__ double tg_ff_background(const double *nz, const int *cls, int nH, double x)
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
        default: t = exp(q[TG_H_P] * (q[TG_H_LNA] + log(x))) + 1.0; break;
        }
        acc = fma(q[TG_H_H], tg_rcp(t), acc);
    }
    return acc + nz[TG_N_N0];
}

// Dynamic shared memory of the pre-pass for tables of ncm components: the chain's table (nu, q, h, lo, hi), the pair-level lists
// (pair-far, local) and one mid-far list per warp
#define TG_FF_SMEM(ncm) ((size_t)(ncm) * (3 * sizeof(double) + 2 * sizeof(int) + (2 + 8) * sizeof(unsigned short)) + 64)

// 16 partial products of the node-values -> coefficients matrix per half-warp, then the two halves: lane k (and k + 16) of the
// warp returns coefficient k of the values v held by lanes 0..15 (and, identically, 16..31)
__device__ __forceinline__ double tg_ff_transform(const double (*s_W)[TG_FF_N + 1], double v, int lane)
{
    const int i = lane & 15, n0 = (lane >> 4) * (TG_FF_N / 2);
    double ca = 0.0, cb = 0.0;
#pragma unroll
    for (int n = 0; n < TG_FF_N / 2; n += 2) {
        ca = fma(s_W[i][n0 + n], __shfl_sync(0xffffffffu, v, n0 + n), ca);
        cb = fma(s_W[i][n0 + n + 1], __shfl_sync(0xffffffffu, v, n0 + n + 1), cb);
    }
    const double c = ca + cb;
    return c + __shfl_xor_sync(0xffffffffu, c, 16);
}

// Chebyshev coefficients c (lane k and k + 16 hold c_k) -> coefficients of the same polynomial in powers of t, a_k = sum_j M[k][j] c_j
// (only j >= k of the parity of k: four terms per half-warp).  Why this is safe here and not in general: the interpolated functions
// are analytic in a Bernstein ellipse of parameter rho >= 10.9, so |c_j| <= ~rho^-j |c_0| and |a_k| ~ (2/rho)^k |c_0|: the power
// series is as well conditioned on [-1, 1] as the Chebyshev one, every term M[k][j] c_j is below |c_0|, and the rounding noise of the
// trailing c_j enters a as M (delta c) exactly (M holds integers), i.e. as the same small polynomial sum_j delta c_j T_j(t).  The
// fused kernel then takes one Horner recurrence (15 FMA per bin) instead of Clenshaw's (30).
__device__ __forceinline__ double tg_ff_to_powers(const double (*s_M)[TG_FF_N + 1], double c, int lane)
{
    const int k = lane & 15, h4 = (lane >> 4) * 4;
    double s = 0.0;
#pragma unroll
    for (int t = 3; t >= 0; t--) {                                      // smallest terms first
        const int j = k + 2 * (h4 + t);
        const double cj = __shfl_sync(0xffffffffu, c, j & 15);
        s = fma((j < TG_FF_N) ? s_M[k][j & 15] : 0.0, cj, s);
    }
    return s + __shfl_xor_sync(0xffffffffu, s, 16);
}

// The pre-pass.  grid = (CTAs per chain, chains of the launch); cmap[chain of the launch] = (star, chain of that star) or NULL
// (one star, identity); argv[star] = its argument record.  CTA r of a chain takes the pairs r*ppc .. r*ppc + ppc - 1 of the
// prepared range, one after the other (the chain's table is staged once), and of each pair the 8 micro-tiles, one per warp.  Two
// levels: the components whose poles are far from the whole PAIR (2048 bins) are summed at the 16 nodes of the pair by all
// threads and interpolated once (they are about half of those that overlap it); every warp then sums the components that are
// far from its micro-tile only, adds the pair's interpolant at its nodes (a polynomial of the same degree: reproduced exactly),
// and leaves coefficients, header and exact records of its micro-tile in HBM.
__global__ void __launch_bounds__(256, 4) tg_k_ff(const TgFFArgs *__restrict__ argv, const int2 *__restrict__ cmap, int ppc)
{
    extern __shared__ __align__(16) unsigned char tg_ff_smem[];
    __shared__ double s_W[TG_FF_N][TG_FF_N + 1];
    __shared__ double s_M[TG_FF_N][TG_FF_N + 1];
    __shared__ double s_tn[TG_FF_N];
    __shared__ double s_nz[TG_NOISE_STRIDE];
    __shared__ int s_cls[TG_MAXH];
    __shared__ int s_xl[8][TG_FF_E];                                    // exact list of every warp's micro-tile
    __shared__ double s_part[TG_FF_N][TG_FF_N + 1];                     // pair level: [slice][node]
    __shared__ double s_cp[TG_FF_N];                                    // pair level: coefficients
    __shared__ int s_cnt[2][8], s_tot[2], s_bgpair;
    __shared__ __align__(16) TgFFArgs s_args;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int chain_l = blockIdx.y;
    int2 cm = make_int2(0, chain_l);
    if (cmap) cm = __ldg(cmap + chain_l);
    tg_load_args(&s_args, argv + cm.x);
    const TgFFArgs &a = s_args;
    const int b = cm.y;
    if ((int)blockIdx.x * ppc * TG_FF_MT >= a.nmt) return;              // another star of the launch has the longer range
    const int ncm = a.d.ncomp_max;
    double *t_nu = (double *)tg_ff_smem, *t_q = t_nu + ncm, *t_h = t_q + ncm;
    int *t_lo = (int *)(t_h + ncm), *t_hi = t_lo + ncm;
    unsigned short *l_pf = (unsigned short *)(t_hi + ncm), *l_loc = l_pf + ncm, *l_mid = l_loc + ncm + (size_t)warp * ncm;
    s_W[tid >> 4][tid & 15] = __ldg(a.ffw + tid);
    s_M[tid >> 4][tid & 15] = __ldg(a.ffw + TG_FF_N * TG_FF_N + TG_FF_N + tid);
    if (tid < TG_FF_N) s_tn[tid] = __ldg(a.ffw + TG_FF_N * TG_FF_N + tid);
    tg_pdl_wait();                                                      // the tables of the proposed points are complete
    tg_pdl_trigger();
    // ---- the chain's noise table and component table -> shared memory (one round trip) ----
    if (tid < TG_NOISE_STRIDE) s_nz[tid] = a.t.noise[(size_t)b * TG_NOISE_STRIDE + tid];
    const int ncomp = min(a.t.ncomp[b], ncm);
    for (int k = tid; k < ncm; k += 256) {
        const size_t e = (size_t)b * ncm + k;
        t_lo[k] = a.t.lo[e]; t_hi[k] = a.t.hi[e]; t_nu[k] = a.t.nu[e]; t_q[k] = a.t.q[e]; t_h[k] = a.t.h[e];
    }
    __syncthreads();
    if (tid < TG_MAXH) s_cls[tid] = (int)s_nz[TG_HS * tid + TG_H_CLS];
    const int nH = (int)s_nz[TG_N_NH];
    const bool sym = (s_nz[TG_N_ASYM] == 0.0) && (s_nz[TG_N_GAUSS] == 0.0);
    const double r2 = a.r2, rs = a.rs;
    const int i = lane & 15, half = lane >> 4;
    const double tn_i = s_tn[i];

#pragma unroll 1
    for (int pq = 0; pq < ppc; pq++) {
        const int m_first = ((int)blockIdx.x * ppc + pq) * TG_FF_MT;    // within the prepared range
        if (m_first >= a.nmt) break;
        // geometry: the pair (first .. last whole micro-tile of it inside the range)
        const int m_last = min(m_first + TG_FF_MT, a.nmt) - 1;
        const TgFFGeo g0 = a.geo[a.mt0 + m_first], g1 = a.geo[a.mt0 + m_last];
        const double pxa = g0.xc - g0.hw, pxb = g1.xc + g1.hw, pxc = 0.5 * (pxa + pxb), phw = 0.5 * (pxb - pxa);
        const int P0 = (a.mt0 + m_first) * TG_FF_BINS, P1 = (a.mt0 + m_last + 1) * TG_FF_BINS;
        const bool geo_ok = (g0.hw > 0.0) && (g1.hw > 0.0) && (phw > 0.0);
        __syncthreads();                                                // the lists of the previous pair are dead
        if (tid < 2) s_tot[tid] = 0;
        if (tid == 32) {
            // is the background far from the whole pair?  Its poles lie on |x| = 1/a_k (x = 0: branch point of a non-integer power)
            const double pedge = pxc - rs * phw;
            bool bg = sym && geo_ok && (pedge > 0.0);
            for (int k = 0; k < nH; k++) bg = bg && (s_nz[TG_HS * k + TG_H_A] * pedge >= 1.0);
            s_bgpair = bg ? 1 : 0;
        }
        // ---- pair-level classification: pair-far / local (meets the pair, not pair-far) ----
        const double r2p = r2 * (phw * phw);
        for (int k0 = 0; k0 < ncomp; k0 += 256) {
            const int k = k0 + tid;
            bool pf = false, loc = false;
            if (k < ncomp) {
                const int lo = t_lo[k], hi = t_hi[k];
                const double dd = t_nu[k] - pxc, qq = t_q[k];
                const bool ov = (hi > lo) && (lo < P1) && (hi > P0);
                pf = ov && sym && geo_ok && (lo <= P0) && (hi >= P1) && (fma(qq * dd, dd, 1.0) >= r2p * qq);   // false for NaN
                loc = ov && !pf;
            }
            const unsigned bp = __ballot_sync(0xffffffffu, pf), bl = __ballot_sync(0xffffffffu, loc);
            if (lane == 0) { s_cnt[0][warp] = __popc(bp); s_cnt[1][warp] = __popc(bl); }
            __syncthreads();
            int op = s_tot[0], ol = s_tot[1];
            for (int w = 0; w < warp; w++) { op += s_cnt[0][w]; ol += s_cnt[1][w]; }
            const unsigned below = (1u << lane) - 1u;
            if (pf) l_pf[op + __popc(bp & below)] = (unsigned short)k;
            if (loc) l_loc[ol + __popc(bl & below)] = (unsigned short)k;
            __syncthreads();
            if (tid < 2) { int t = s_tot[tid]; for (int w = 0; w < 8; w++) t += s_cnt[tid][w]; s_tot[tid] = t; }
        }
        __syncthreads();
        const int npf = s_tot[0], nloc = s_tot[1];
        const bool bg_pair = (s_bgpair != 0);
        // ---- pair level: thread (slice, node) sums every 16th pair-far component at a node of the pair ----
        {
            const int sl = tid >> 4;
            const double xn = fma(phw, tn_i, pxc);
            double v = 0.0;
            for (int e = sl; e < npf; e += 16) {
                const int k = l_pf[e];
                const double d = xn - t_nu[k];
                v = fma(t_h[k], tg_rcp_fast(fma(d * d, t_q[k], 1.0)), v);
            }
            if (sl == 15 && bg_pair) v += tg_ff_background(s_nz, s_cls, nH, xn);
            s_part[sl][i] = v;
        }
        __syncthreads();
        if (warp == 0) {
            double v = 0.0;
#pragma unroll
            for (int sl = 0; sl < 16; sl++) v += s_part[sl][i];         // fixed order (both half-warps: the same sums)
            const double c = tg_ff_to_powers(s_M, tg_ff_transform(s_W, v, lane), lane);
            if (lane < TG_FF_N) s_cp[lane] = c;
        }
        __syncthreads();
        // ---- micro-tile level: warp = micro-tile ----
        const int m = m_first + warp;
        if (m < a.nmt) {
            const TgFFGeo gm = a.geo[a.mt0 + m];
            const double xc = gm.xc, hw = gm.hw;
            const int t0 = (a.mt0 + m) * TG_FF_BINS, t1 = t0 + TG_FF_BINS;
            const double r2w2 = r2 * (hw * hw);
            int nmid = 0, nex = 0;
            for (int e0 = 0; e0 < nloc; e0 += 32) {
                const int e = e0 + lane;
                bool mid = false, ex = false;
                int k = 0;
                if (e < nloc) {
                    k = l_loc[e];
                    const int lo = t_lo[k], hi = t_hi[k];
                    const bool ov = (lo < t1) && (hi > t0), cv = (lo <= t0) && (hi >= t1);
                    const double dd = t_nu[k] - xc, qq = t_q[k];
                    mid = ov && cv && sym && (hw > 0.0) && (fma(qq * dd, dd, 1.0) >= r2w2 * qq);      // far from the micro-tile; false for NaN
                    ex = ov && !mid;
                }
                const unsigned bm = __ballot_sync(0xffffffffu, mid), be = __ballot_sync(0xffffffffu, ex);
                const unsigned below = (1u << lane) - 1u;
                if (mid) l_mid[nmid + __popc(bm & below)] = (unsigned short)k;
                if (ex) { const int pos = nex + __popc(be & below); if (pos < TG_FF_E) s_xl[warp][pos] = k; }
                nmid += __popc(bm); nex += __popc(be);
            }
            __syncwarp();
            const double xn = fma(hw, tn_i, xc);
            double v0 = 0.0, v1 = 0.0;
            for (int e = 2 * half; e < nmid; e += 4) {                  // half-warp h: entries 2h, 2h+1, 2h+4, 2h+5, ...
                const int k0 = l_mid[e];
                const double d0 = xn - t_nu[k0];
                v0 = fma(t_h[k0], tg_rcp_fast(fma(d0 * d0, t_q[k0], 1.0)), v0);
                if (e + 1 < nmid) {
                    const int k1 = l_mid[e + 1];
                    const double d1 = xn - t_nu[k1];
                    v1 = fma(t_h[k1], tg_rcp_fast(fma(d1 * d1, t_q[k1], 1.0)), v1);
                }
            }
            double v = v0 + v1;
            // the background: with the pair's interpolant, else at this micro-tile's nodes if far from it, else left to the fused kernel
            bool bg_here = false;
            if (!bg_pair) {
                const double edge = xc - rs * hw;
                bg_here = sym && (edge > 0.0) && (hw > 0.0);
                for (int k = 0; k < nH; k++) bg_here = bg_here && (s_nz[TG_HS * k + TG_H_A] * edge >= 1.0);
            }
            if (half == 1) {
                if (bg_here) v += tg_ff_background(s_nz, s_cls, nH, xn);
            } else if (geo_ok) {                                        // the pair's interpolant at this node: Horner
                const double tt = (xn - pxc) * tg_rcp(phw);
                double pv = s_cp[TG_FF_N - 1];
#pragma unroll
                for (int k = TG_FF_N - 2; k >= 0; k--) pv = fma(tt, pv, s_cp[k]);
                v += pv;
            }
            v += __shfl_xor_sync(0xffffffffu, v, 16);
            const double c = tg_ff_transform(s_W, v, lane);
            const double tail = fabs(__shfl_sync(0xffffffffu, c, TG_FF_N - 1)) + fabs(__shfl_sync(0xffffffffu, c, TG_FF_N - 2));
            const double cz = __shfl_sync(0xffffffffu, c, 0);
            const bool good_tail = (tail <= TG_FF_TOL * fabs(cz));      // false for NaN; 0 <= 0 when nothing is far
            const bool good = good_tail && (hw > 0.0) && (nex <= TG_FF_E);
            const size_t slot = (size_t)b * a.stride + m;
            const double ap = tg_ff_to_powers(s_M, c, lane);
            if (lane < TG_FF_N) a.c[slot * TG_FF_N + lane] = good ? ap : 0.0;
            // exact records, one lane each: those whose window covers the whole micro-tile first (list order kept inside both groups),
            // so the fused kernel's loop over them carries no window predicate
            const bool have = good && lane < nex;
            TgFFRec rc; rc.nu = 0.0; rc.q = 0.0; rc.h = 0.0; rc.lo = 0; rc.len = 0;
            if (have) { const int k = s_xl[warp][lane]; rc.nu = t_nu[k]; rc.q = t_q[k]; rc.h = t_h[k]; rc.lo = t_lo[k]; rc.len = t_hi[k] - rc.lo; }
            const bool need = have && !((rc.lo <= t0) && (rc.lo + rc.len >= t1));
            const unsigned pm = __ballot_sync(0xffffffffu, need), fm = __ballot_sync(0xffffffffu, have && !need);
            const int nfree = __popc(fm);
            if (have) {
                const unsigned below = (1u << lane) - 1u;
                a.rec[slot * TG_FF_E + (need ? nfree + __popc(pm & below) : __popc(fm & below))] = rc;
            }
            if (lane == 0) {
                TgFFHdr hd; hd.count = good ? nex : TG_FF_FALLBACK; hd.nfree = nfree;
                hd.flags = ((good && (bg_pair || bg_here)) ? 1 : 0) | (good_tail ? 0 : 2) | (nex <= TG_FF_E ? 0 : 4) | (hw > 0.0 ? 0 : 8); hd.nfar = nmid + npf;
                a.hdr[slot] = hd;
            }
        }
    }
}

// ---- fused-kernel side --------------------------------------------------------------------------------------------------
// On a pair of sub-tiles (8 consecutive micro-tiles) WARP w of the CTA takes micro-tile w: lane l holds its bins l, l + 32, ...,
// l + 224.  Everything the pre-pass left for (chain, micro-tile) -- 16 coefficients, header, geometry, records -- is therefore
// private to one warp: it travels from L2 into the warp's slot of the staging buffer one (pair, chain) step ahead of the
// arithmetic (three 16-byte loads per lane) and is handed over with __syncwarp only -- no CTA barrier inside a run, the warps drift
// apart and hide each other's latencies.  The eight bins of a thread share one interpolant and one record list: every coefficient
// and every record is read once for eight independent chains of arithmetic.
struct __align__(16) TgFFStage {
    TgFFRec rec[TG_FF_MT][TG_FF_E];    // 8192 bytes   [warp][record]
    double  c[TG_FF_MT][TG_FF_N];      // 1024 bytes   [warp][k]: coefficient of t^k
    TgFFHdr hdr[TG_FF_MT];             //  128 bytes
    TgFFGeo geo[TG_FF_MT];             //  256 bytes
};
struct TgFFFetch { int4 r0, r1, r2; };
static_assert(sizeof(TgFFRec) == 32 && TG_FF_E == 32, "one record per lane: two 16-byte pieces");
static_assert(TG_FF_MT * 32 == TG_FF_WBINS, "eight bins per lane: 256 bins per warp and step (a whole micro-tile, or half of one)");

// what the pre-pass left for `chain` on the micro-tile of this warp in the pair starting at bin bp -> registers
__device__ __forceinline__ void tg_ff_fetch(const TgFFView &f, int chain, int bp, int warp, int lane, TgFFFetch &q)
{
    const int mg = (bp >> TG_FF_SHIFT) + warp / TG_FF_WPM;
    const size_t slot = (size_t)chain * f.stride + (mg - f.mt0);
    const int4 *rp = (const int4 *)(f.rec + slot * TG_FF_E) + 2 * lane;
    q.r0 = __ldg(rp); q.r1 = __ldg(rp + 1);                            // record `lane` (slots past the count are never read back)
    q.r2 = make_int4(0, 0, 0, 0);
    if (lane < 8)       q.r2 = __ldg((const int4 *)(f.c + slot * TG_FF_N) + lane);
    else if (lane == 8) q.r2 = __ldg((const int4 *)(f.hdr + slot));
    else if (lane < 11) q.r2 = __ldg((const int4 *)(f.geo + mg) + (lane - 9));
}
__device__ __forceinline__ void tg_ff_stage(TgFFStage &st, int warp, int lane, const TgFFFetch &q)
{
    int4 *rp = (int4 *)&st.rec[warp][lane];
    rp[0] = q.r0; rp[1] = q.r1;
    if (lane < 8)       ((int4 *)&st.c[warp][0])[lane] = q.r2;
    else if (lane == 8) *(int4 *)&st.hdr[warp] = q.r2;
    else if (lane < 11) ((int4 *)&st.geo[warp])[lane - 9] = q.r2;
}

// Far + exact parts of chain `chain` on this warp's micro-tile (first bin bw; bin j of the lane = bw + 32 j + lane) from the staged
// data: acc[j] = interpolant + exact records.  Returns whether the interpolant includes the background and N0.
template <int NB>
__device__ __forceinline__ bool tg_ff_microtile(const TgFFStage &st, const TgTables &t, int ncm, int chain, int bw, int warp, int lane, const double (&x)[NB], double (&acc)[NB])
{
    static_assert(NB == TG_FF_MT, "8 bins per thread");
    const int4 h4 = *(const int4 *)&st.hdr[warp];
    const int cnt = h4.x, nfree = h4.y;
    // Horner in the local variable t = (x - centre) / half-width: eight independent recurrences, the coefficients two per load
    {
        const double xc = st.geo[warp].xc, ihw = st.geo[warp].ihw;
        double tt[NB];
#pragma unroll
        for (int j = 0; j < NB; j++) tt[j] = (x[j] - xc) * ihw;
#pragma unroll
        for (int k = TG_FF_N - 2; k >= 0; k -= 2) {
            const double2 cc = *(const double2 *)&st.c[warp][k];                                        // (a_k, a_k+1)
#pragma unroll
            for (int j = 0; j < NB; j++) acc[j] = (k == TG_FF_N - 2) ? fma(tt[j], cc.y, cc.x) : fma(tt[j], fma(tt[j], acc[j], cc.y), cc.x);
        }
    }
    if (cnt >= 0) {
        // exact records: those whose window covers the whole micro-tile first (no predicate), then the ones cut by a window edge
#pragma unroll 1
        for (int e = 0; e < nfree; e++) {
            const double2 r0 = *(const double2 *)&st.rec[warp][e].nu;      // (nu, q)
            const double hh = st.rec[warp][e].h;
#pragma unroll
            for (int j = 0; j < NB; j++) { const double d = x[j] - r0.x; acc[j] = fma(hh, tg_rcp_fast(fma(d * d, r0.y, 1.0)), acc[j]); }
        }
#pragma unroll 1
        for (int e = nfree; e < cnt; e++) {
            const double2 r0 = *(const double2 *)&st.rec[warp][e].nu;
            const double hh = st.rec[warp][e].h;
            const int2 w = *(const int2 *)&st.rec[warp][e].lo;
            const int rel = bw + lane - w.x;
#pragma unroll
            for (int j = 0; j < NB; j++) {
                const double d = x[j] - r0.x;
                const double rr = tg_rcp_fast(fma(d * d, r0.y, 1.0));
                acc[j] = fma(((unsigned)(rel + 32 * j) < (unsigned)w.y) ? hh : 0.0, rr, acc[j]);
            }
        }
    } else {
        // marked micro-tile (failed check / too many exact components): every component of the chain, window predicate included
        const int nc = min(t.ncomp[chain], ncm);
#pragma unroll 1
        for (int k = 0; k < nc; k++) {
            const size_t e = (size_t)chain * ncm + k;
            const int lo = t.lo[e], len = t.hi[e] - lo;
            const double nu = t.nu[e], q = t.q[e], hh = t.h[e];
            const int rel = bw + lane - lo;
#pragma unroll
            for (int j = 0; j < NB; j++) {
                if ((unsigned)(rel + 32 * j) < (unsigned)len) { const double d = x[j] - nu; acc[j] = fma(hh, tg_rcp_fast(fma(d * d, q, 1.0)), acc[j]); }
            }
        }
    }
    return (h4.z & 1) != 0;
}
