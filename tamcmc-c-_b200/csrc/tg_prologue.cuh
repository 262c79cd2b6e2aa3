// tg_prologue.cuh -- per-chain scalar prologue: parameter vector -> Lorentzian component table,
// noise constants, log-prior.  One CTA per chain; one thread per mode, one thread per prior term.
//
// Reference (all under /root/reference/tamcmc/sources):
//   models.cpp  :15-167 (case 6) :169-322 (7) :325-478 (8) :480-637 (2) :803-946 (3) :952-1113 (12)
//               :1121-1290 (13) :1295-1493 (9) :1496-1674 (10) :1683-1829 (11) :1832-1963 (14)
//               :1968-1992 (1) :2021-2034 (0)
//   build_lorentzian.cpp :377-427 (window), :61-102 / :15-59 / :154-205 (component centres)
//   function_rot.cpp :20-106 (amplitude_ratio), interpol.cpp :13-55 (lin_interpol)
//   priors_calc.cpp :24-172, :175-280, :315-417 ; stats_dictionary.cpp :31-241 ;
//   derivatives_handler.cpp :342-408
#pragma once
#include "tg_internal.cuh"

#ifdef TG_PROFILE
__device__ long long tg_dbg_clk[32];
#define TG_STAMP(k) do { __syncthreads(); if (threadIdx.x == 0 && blockIdx.x == 0 && blockIdx.y == 0) tg_dbg_clk[k] = clock64(); } while (0)
#else
#define TG_STAMP(k) do { } while (0)
#endif

struct TgProArgs {
    TgDesc        d;
    TgTables      t;
    const double *params;   // [B][Np]
    const double *pri;      // [4*Np]
    const int    *sw;       // [Np]
};

#define TG_PI 3.141592653589793238462643383279502884

// ---- function_rot.cpp:20-106 ----------------------------------------------------------------
// amplitude_ratio(l, beta)[m+l] = (d^l_{m0}(beta))^2.  The reference builds the full Wigner matrix
// with pow() and factorials; the column it uses has the closed forms below (s = sin beta,
// c = cos beta), which agree with the reference's own output to 6e-16 absolute (checked against
// oracle/_ref for beta in [-30, 1234] degrees).
__device__ double tg_amp_ratio(int l, int m, double beta_deg)
{
    const double angle = 3.141592653589793238462643 * beta_deg / 180.;
    double s, c;
    sincos(angle, &s, &c);
    const int am = m < 0 ? -m : m;
    const double s2 = s * s, c2 = c * c;
    if (l == 1) return am == 1 ? 0.5 * s2 : c2;
    if (l == 2) {
        if (am == 2) return 0.375 * s2 * s2;
        if (am == 1) return 1.5 * s2 * c2;
        const double t = 3. * c2 - 1.; return 0.25 * t * t;
    }
    if (l == 3) {
        if (am == 3) return 0.3125 * s2 * s2 * s2;
        if (am == 2) return 1.875 * c2 * s2 * s2;
        if (am == 1) { const double t = 5. * c2 - 1.; return 0.1875 * s2 * t * t; }
        const double t = 5. * c2 * c - 3. * c; return 0.25 * t * t;
    }
    return 1.0;
}

// ---- interpol.cpp:13-55 -------------------------------------------------------------------
__device__ double tg_lin_interpol(const double *x, const double *y, int Nx, double x_int)
{
    int i = 0;
    double a = 0, b = 0;
    if (x_int >= x[0] && x_int <= x[Nx - 1]) {
        while ((x_int < x[i] || x_int > x[i + 1])) {
            i = i + 1;
            if (i >= Nx - 1) { i = Nx - 2; break; }
        }
        if (i == 0 && (x_int < x[i] || x_int > x[i + 1])) { i = i + 1; }
        a = (y[i + 1] - y[i]) / (x[i + 1] - x[i]);
        b = __dsub_rn(y[i], __dmul_rn(a, x[i]));
    }
    if (x_int < x[0]) {
        a = (y[1] - y[0]) / (x[1] - x[0]);
        b = __dsub_rn(y[0], __dmul_rn(a, x[0]));
    }
    if (x_int > x[Nx - 1]) {
        a = (y[Nx - 1] - y[Nx - 2]) / (x[Nx - 1] - x[Nx - 2]);
        b = __dsub_rn(y[Nx - 2], __dmul_rn(a, x[Nx - 2]));
    }
    return __dadd_rn(__dmul_rn(a, x_int), b);   // a*x_int+b without contraction, as the reference
}

// ---- build_lorentzian.cpp:377-427 truncation window (quirks Q1, Q2) -----------------------
// All arithmetic un-contracted (__dmul_rn/__dadd_rn): a one-bin disagreement with the reference
// would change the model by ~6e-4*H, so the floor/ceil arguments must round identically.
__device__ int tg_window(const TgDesc &d, double fc, double f_s, double g, int l, double c, long *imin_o, long *imax_o)
{
    double pmin = 0, pmax = 0;
    int hit = 0;
    const double step = __dsub_rn(d.x_second, d.x_first);
    const double dl = (double)l;
    if (g >= 1 && f_s >= 1) {
        hit = 1;
        if (l != 0) { double w = __dmul_rn(c, __dadd_rn(__dmul_rn(dl, f_s), g)); pmin = __dsub_rn(fc, w); pmax = __dadd_rn(fc, w); }
        else        { double w = __dmul_rn(__dmul_rn(c, g), 2.2);                pmin = __dsub_rn(fc, w); pmax = __dadd_rn(fc, w); }
    }
    if (g <= 1 && f_s >= 1) {
        hit = 1;
        if (l != 0) { double w = __dmul_rn(c, __dadd_rn(__dmul_rn(dl, f_s), 1.0)); pmin = __dsub_rn(fc, w); pmax = __dadd_rn(fc, w); }
        else        { double w = __dmul_rn(c, 2.2);                                pmin = __dsub_rn(fc, w); pmax = __dadd_rn(fc, w); }
    }
    if (g >= 1 && f_s <= 1) {
        hit = 1;
        if (l != 0) { double w = __dmul_rn(c, __dadd_rn(dl, g));        pmin = __dsub_rn(fc, w); pmax = __dadd_rn(fc, w); }
        else        { double w = __dmul_rn(__dmul_rn(c, 2.2), g);       pmin = __dsub_rn(fc, w); pmax = __dadd_rn(fc, w); }
    }
    if (g <= 1 && f_s <= 1) {
        hit = 1;
        if (l != 0) { double w = __dmul_rn(c, (double)(l + 1)); pmin = __dsub_rn(fc, w); pmax = __dadd_rn(fc, w); }
        else        { double w = __dmul_rn(c, 2.2);             pmin = __dsub_rn(fc, w); pmax = __dadd_rn(fc, w); }
    }
    if (!hit) return TG_ST_NAN_WINDOW;
    if (__dsub_rn(pmax, step) < d.x_first)  pmax = __dadd_rn(d.x_first, c);
    if (__dadd_rn(pmin, step) >= d.x_last)  pmin = __dsub_rn(d.x_last, c);
    double fmin = floor(__ddiv_rn(__dsub_rn(pmin, d.x_first), step));
    double fmax = ceil(__ddiv_rn(__dsub_rn(pmax, d.x_first), step));
    if (!(fmin == fmin) || !(fmax == fmax)) return TG_ST_NAN_WINDOW;
    fmin = fmin < -4.0e18 ? -4.0e18 : (fmin > 4.0e18 ? 4.0e18 : fmin);
    fmax = fmax < -4.0e18 ? -4.0e18 : (fmax > 4.0e18 ? 4.0e18 : fmax);
    long imin = (long)fmin, imax = (long)fmax;
    if (imin < 0) imin = 0;
    if (imax > d.Nx_global) imax = d.Nx_global;
    if (imax - imin <= 0) return TG_ST_EMPTY_WINDOW;
    *imin_o = imin; *imax_o = imax;
    return 0;
}

// ---- stats_dictionary.cpp:31-241 (double instead of the reference's x87 long double) -------
__device__ __forceinline__ double tg_ninf() { return -CUDART_INF; }
__device__ double tg_logP_uniform(double a, double b, double x) { return ((x <= b) && (x >= a)) ? -log(fabs(b - a)) : tg_ninf(); }
__device__ double tg_logP_uniform_abs(double a, double b, double x) { return ((fabs(x) <= b) && (fabs(x) >= a)) ? -log(fabs(b - a)) : tg_ninf(); }
__device__ double tg_logP_uniform_cos(double a, double b, double x)
{
    const double xr = TG_PI * x / 180.;
    return ((cos(xr) < b) && (cos(xr) > a)) ? -log(fabs(b - a)) : tg_ninf();
}
__device__ double tg_logP_gaussian(double mean, double sigma, double x)
{
    const double z = (x - mean) / sigma;
    return -log(sqrt(2 * TG_PI) * sigma) - 0.5 * (z * z);
}
__device__ double tg_logP_jeffrey(double hmin, double hmax, double h)
{
    if (h < hmax && h > 0) { double prior = 1. / (h + hmin); double norm = log((hmax + hmin) / hmin); return log(prior / norm); }
    return tg_ninf();
}
__device__ double tg_logP_jeffrey_abs(double hmin, double hmax, double h)
{
    if (fabs(h) < hmax) { double prior = 1. / (fabs(h) + hmin); double norm = log((hmax + hmin) / hmin); return log(prior / norm); }
    return tg_ninf();
}
__device__ double tg_logP_ug(double a, double b, double s, double x)
{
    double lp = CUDART_NAN;
    if (x < a) lp = tg_ninf();
    if ((x <= b) && (x >= a)) lp = 0;
    if (x > b) { double z = (x - b) / s; lp = -0.5 * (z * z); }
    return lp - log(fabs(b - a) + 0.5 * sqrt(2 * TG_PI) * s);
}
__device__ double tg_logP_gu(double a, double b, double s, double x)
{
    double lp = CUDART_NAN;
    if (x > b) lp = tg_ninf();
    if ((x <= b) && (x >= a)) lp = 0;
    if (x < a) { double z = (x - a) / s; lp = -0.5 * (z * z); }
    return lp - log(fabs(b - a) + 0.5 * sqrt(2 * TG_PI) * s);
}
__device__ double tg_logP_gug(double a, double b, double s1, double s2, double x)
{
    double lp = CUDART_NAN;
    if (x < a) { double z = (x - a) / s1; lp = -0.5 * (z * z); }
    if ((x <= b) && (x >= a)) lp = 0;
    if (x > b) { double z = (x - b) / s2; lp = -0.5 * (z * z); }
    return lp - log(fabs(b - a) + 0.5 * sqrt(2 * TG_PI) * (s1 + s2));
}
// one term of apply_generic_priors, priors_calc.cpp:315-417
__device__ double tg_generic_prior(int sw, const double *q, double x, int *bad)
{
    switch (sw) {
    case 0: case 11: return 0.0;
    case 1: return tg_logP_uniform(q[0], q[1], x);
    case 2: return tg_logP_gaussian(q[0], q[1], x);
    case 4: return tg_logP_jeffrey(q[0], q[1], x);
    case 5: return tg_logP_ug(q[0], q[1], q[2], x);
    case 6: return tg_logP_gu(q[0], q[1], q[2], x);
    case 7: return tg_logP_gug(q[0], q[1], q[2], q[3], x);
    case 8: return tg_logP_uniform_abs(q[0], q[1], x);
    case 9: return tg_logP_uniform_cos(q[0], q[1], x);
    case 10: return tg_logP_jeffrey_abs(q[0], q[1], x);
    default: *bad = 1; return 0.0;   // 3 (multivariate Gaussian) and unknown switches: reference exit()s
    }
}

// ---- per-chain scalars shared by the mode threads -------------------------------------------
struct TgProShared {
    double a1, eta, a3, asym, trunc_c, inclination, numax, qmax;
    double Vl[4];
    double ratios[4][7];
    double work;
    int    do_amp, status, bad_prior, wlo, whi;
    int    modeStart[5], compStart[5];   // local family: prefix over degrees
};

__device__ __forceinline__ double tg_height0(double p, double W, int do_amp) { return do_amp ? fabs(p / (TG_PI * W)) : fabs(p); }
__device__ __forceinline__ double tg_heightl(double p, double W, double V, int do_amp) { return do_amp ? fabs(p / (TG_PI * W)) * V : fabs(p * V); }

__device__ double tg_appwidth_v1(const double *w, double f, double numax)
{
    double lnGamma0 = w[1] * log(f / numax) + log(w[2]);
    double e = 2. * log(f / w[0]) / log(w[3] / numax);
    double lnLorentz = -log(w[4]) / (1. + e * e);
    return exp(lnGamma0 + lnLorentz);
}
__device__ double tg_appwidth_v2(const double *w, double f)
{
    double lnGamma0 = w[2] * log(f / w[0]) + log(w[3]);
    double e = 2. * log(f / w[1]) / log(w[4] / w[0]);
    double lnLorentz = -log(w[5]) / (1. + e * e);
    return exp(lnGamma0 + lnLorentz);
}

// Writes component m of one mode (one thread per component; the thread of m = -l also accounts for the
// mode's window in the work / status / union accumulators).  variant 0: a1etaa3, 1: a1l (f_s per degree, c_1m=0).
__device__ void tg_emit_comp(const TgProArgs &a, TgProShared &S, int b, int coff, int l, int m, double fc, double W,
                             double hm, double f_s, double f_s1, double f_s2, int variant, double &work_acc, int &st_acc, int &wlo_acc, int &whi_acc)
{
    const TgDesc &d = a.d;
    long imin = 0, imax = 0;
    const double fs_win = (variant == 1) ? (f_s1 + f_s2) / 2. : f_s;     // Q3
    int st = tg_window(d, fc, fs_win, W, l, S.trunc_c, &imin, &imax);
    int lo = 0, hi = 0;
    if (st == 0) {
        long a0 = imin - d.shard_lo, a1 = imax - d.shard_lo;
        a0 = a0 < 0 ? 0 : (a0 > d.Nx ? d.Nx : a0);
        a1 = a1 < 0 ? 0 : (a1 > d.Nx ? d.Nx : a1);
        lo = (int)a0; hi = (int)a1;
        if (m == -l) {
            work_acc += (double)(2 * l + 1) * (double)(hi - lo);       // integer-valued: exact, order-free
            if (hi > lo) { wlo_acc = min(wlo_acc, lo); whi_acc = max(whi_acc, hi); }
        }
    } else {
        st_acc |= st;
    }
    const double g2 = W * W;                 // pow(gamma_l,2)
    double q = 4.0 / g2;                     // 4*profile/g2 -> profile*(4/g2): <= 1 ulp from the reference
    if (q > S.qmax) q = S.qmax;              // keep 1+q*d^2 <= ~1e70: the product of four such terms stays finite (Gamma < ~1e-31 only)
    double ifc = 0.0, t2 = 0.0;              // asymmetry extras (only read by the kernel when asym != 0)
    if (S.asym != 0.0) { ifc = 1.0 / fc; t2 = 0.5 * W * S.asym / fc; }
    double nu = fc;
    if (l != 0) {
        // Q_lm = (l(l+1) - 3m^2) / ((2l-1)(2l+3)): small-integer quotients, folded at compile time with the
        // same single rounding as the reference's run-time division (build_lorentzian.cpp:77)
        const int am = m < 0 ? -m : m;
        const double Qlm = (l == 1) ? (am == 0 ? 2.0 / 5.0 : -1.0 / 5.0)
                         : (l == 2) ? (am == 0 ? 6.0 / 21.0 : (am == 1 ? 3.0 / 21.0 : -6.0 / 21.0))
                         :            (am == 0 ? 12.0 / 45.0 : (am == 1 ? 9.0 / 45.0 : (am == 2 ? 0.0 / 45.0 : -15.0 / 45.0)));
        double clm = 0, fs = f_s;
        if (variant == 1) {
            if (l == 1) { clm = 0; fs = f_s1; }
            if (l == 2) { clm = (double)((5 * m * m * m - 17 * m) / 3); fs = f_s2; }   // exact integers -2,4,0,-4,2
            if (l == 3) { clm = 0; fs = (f_s1 + f_s2) / 2.; }
        } else {
            if (l == 1) clm = m;
            if (l == 2) clm = (double)((5 * m * m * m - 17 * m) / 3);                  // exact integers -2,4,0,-4,2
            if (l == 3) clm = 0;
        }
        // fc*(1+eta*Q) + m*fs + clm*a3, un-contracted
        nu = __dadd_rn(__dadd_rn(__dmul_rn(fc, __dadd_rn(1., __dmul_rn(S.eta, Qlm))), __dmul_rn((double)m, fs)), __dmul_rn(clm, S.a3));
    }
    const size_t k = (size_t)b * d.ncomp_max + coff + (m + l);
    a.t.nu[k] = nu; a.t.h[k] = hm; a.t.q[k] = q;
    a.t.ifc[k] = ifc; a.t.t2sq[k] = t2 * t2;
    a.t.lo[k] = lo; a.t.hi[k] = hi;
}

// Body shared by the stand-alone prologue kernel (tg_eval path) and the sampler's propose kernel.
// sm_terms: [Np + 5*TG_MAXORD] doubles of shared memory for the prior terms.  Any blockDim.x >= 32.
// parts: TG_PRO_MODEL (component table, noise constants, window union: what the fused kernel reads) and/or
// TG_PRO_PRIOR (log-prior: only the accept decision reads it, so the sampler runs it in another CTA, off the
// path to the fused kernel).  staged: the caller already put the parameter vector into sm_terms[0..Np).
#define TG_PRO_MODEL 1
#define TG_PRO_PRIOR 2
__device__ void tg_prologue_body(const TgProArgs &a, const int b, double *sm_terms, const int parts = TG_PRO_MODEL | TG_PRO_PRIOR, const bool staged = false)
{
    __shared__ TgProShared S;
    __shared__ double s_red[256];
    const TgDesc &d = a.d;
    const int tid = threadIdx.x;
    // stage the parameter vector in shared memory: every later access (serial sections included) is an LDS
    double *sp = sm_terms, *terms = sm_terms + d.Np;
    {
        const double *pg = a.params + (size_t)b * d.Np;
        double *nzg = a.t.noise + (size_t)b * TG_NOISE_STRIDE;
        if (!staged) for (int i = tid; i < d.Np; i += blockDim.x) sp[i] = pg[i];
        if (parts & TG_PRO_MODEL) for (int k = tid; k < TG_NOISE_STRIDE; k += blockDim.x) nzg[k] = 0.0;
        if (tid == 0) { S.status = 0; S.bad_prior = 0; }
    }
    __syncthreads();
    const double *p = sp;
    const int mc = d.model_case;
    const int fam = (mc <= 1) ? 0 : ((mc == 11 || mc == 14) ? 2 : 1);
    const int Nmax = d.pl[0], lmax = d.pl[1], Nfl0 = d.pl[2], Nfl1 = d.pl[3], Nfl2 = d.pl[4], Nfl3 = d.pl[5];
    const int Nsplit = d.pl[6], Nwidth = d.pl[7], Nnoise = d.pl[8], Ninc = d.pl[9];
    const int Nf = Nfl0 + Nfl1 + Nfl2 + Nfl3;
    const int nV = d.pl[1];                        // lmax (global) or Nvis (local)
    const int iF = Nmax + nV, iS = iF + Nf, iW = iS + Nsplit, iN = iW + Nwidth, iI = iN + Nnoise;
    double *nz = a.t.noise + (size_t)b * TG_NOISE_STRIDE;

    TG_STAMP(4);
    // ---------------- A. chain scalars (thread 0) ----------------
    if ((parts & TG_PRO_MODEL) && tid == 0) {
        S.wlo = 0x7fffffff; S.whi = 0; S.work = 0.0; S.numax = 0; S.a1 = 0; S.inclination = 0;
        { const double range = d.x_last - d.x_first; S.qmax = 1e70 / (range * range); }
        S.Vl[0] = 1; S.Vl[1] = S.Vl[2] = S.Vl[3] = 0;
        if (fam == 0) {
            S.eta = S.a3 = S.asym = 0; S.trunc_c = 0; S.do_amp = 0;
            if (mc == 0) {           // model_Test_Gaussian, models.cpp:2021-2034 (no abs anywhere)
                nz[TG_N_GA] = p[0]; nz[TG_N_GNU] = p[2]; nz[TG_N_GS2] = p[1] * p[1]; nz[TG_N_N0] = p[3];
                nz[TG_N_NH] = 0; nz[TG_N_GAUSS] = 1;
            } else {                 // model_Harvey_Gaussian, :1968-1992 : noise = |params[3..6]|, Nharvey=1
                nz[TG_N_GA] = fabs(p[0]); nz[TG_N_GNU] = p[2]; nz[TG_N_GS2] = fabs(p[1]) * fabs(p[1]);
                nz[TG_N_N0] = fabs(p[6]); nz[TG_N_NH] = 0; nz[TG_N_GAUSS] = 1;
                if (fabs(p[4]) != 0.0) {   // harvey_like skips a term iff tc == 0 (noise_models.cpp:31, Q5)
                    tg_set_harvey(nz, 0, fabs(p[3]), 1e-3 * fabs(p[4]), fabs(p[5]), d.x_last); nz[TG_N_NH] = 1;
                }
            }
        } else {
            S.trunc_c = p[iI + Ninc];
            S.do_amp = (p[iI + Ninc + 1] != 0);
            S.eta = p[iS + 1]; S.a3 = p[iS + 2]; S.asym = p[iS + 5];
            if (mc == 2 || mc == 9 || mc == 10 || mc == 11) {          // sqrt(a1)cos i, sqrt(a1) sin i  (Q7)
                S.a1 = __dadd_rn(__dmul_rn(p[iS + 3], p[iS + 3]), __dmul_rn(p[iS + 4], p[iS + 4]));
                S.inclination = (atan(p[iS + 4] / p[iS + 3]) * 180.) / TG_PI;
            } else if (mc == 3 || mc == 12 || mc == 13) {
                S.a1 = fabs(p[iS]); S.inclination = p[iI];
            } else if (mc == 14) {
                S.a1 = fabs(p[iS]);
            } else {                                                   // a1l family 6,7,8
                S.inclination = p[iI];
            }
            if (fam == 1 && mc != 13) {
                if (lmax >= 1) S.Vl[1] = fabs(p[Nmax]);
                if (lmax >= 2) S.Vl[2] = fabs(p[Nmax + 1]);
                if (lmax >= 3) S.Vl[3] = fabs(p[Nmax + 2]);
            }
            if (mc == 9) {                                             // models.cpp:1372-1390
                double numax = 0, Htot = 0;
                for (int n = 0; n < Nmax; n++) {
                    numax = numax + p[n] * p[Nmax + lmax + n]; Htot = Htot + p[n];
                    if (lmax >= 1) { numax = numax + p[n] * S.Vl[1] * p[Nmax + lmax + Nfl0 + n]; Htot = Htot + p[n] * S.Vl[1]; }
                    if (lmax >= 2) { numax = numax + p[n] * S.Vl[2] * p[Nmax + lmax + Nfl0 + Nfl1 + n]; Htot = Htot + p[n] * S.Vl[2]; }
                    if (lmax >= 3) { numax = numax + p[n] * S.Vl[3] * p[Nmax + lmax + Nfl0 + Nfl1 + Nfl2 + n]; Htot = Htot + p[n] * S.Vl[3]; }
                }
                S.numax = numax / Htot;
            }
            // noise: |noise_params| ; Nharvey=(Nnoise-1)/3 (global) or 0 (local)
            const int nH = (fam == 1) ? (Nnoise - 1) / 3 : 0;
            if (nH > TG_MAXH) S.status |= TG_ST_MODE_OVERFLOW;
            int na = 0;                    // only the ACTIVE terms (tc != 0, noise_models.cpp:31) are listed, in order;
            for (int k = 0; k < nH && k < TG_MAXH; k++) if (fabs(p[iN + 3 * k + 1]) != 0.0) na++;   // their tables: threads 32.. below
            nz[TG_N_N0] = fabs(p[iN + Nnoise - 1]);
            nz[TG_N_NH] = na;
            nz[TG_N_ASYM] = S.asym;
            nz[TG_N_GAUSS] = 0;
            if (fam == 2) {
                const int nfl[4] = {Nfl0, Nfl1, Nfl2, Nfl3};
                int ms = 0, cs = 0;
                for (int l = 0; l < 4; l++) { S.modeStart[l] = ms; S.compStart[l] = cs; ms += nfl[l]; cs += nfl[l] * (2 * l + 1); }
                S.modeStart[4] = ms; S.compStart[4] = cs;
            }
        }
    }
    if ((parts & TG_PRO_MODEL) && fam == 1 && tid >= 32 && tid < 32 + TG_MAXH) {
        // one thread per Harvey term (log, powers: ~1 us each when thread 0 did them one after the other), next to thread 0's scalars
        const int nH = (Nnoise - 1) / 3, k = tid - 32;
        if (k < nH && fabs(p[iN + 3 * k + 1]) != 0.0) {
            int slot = 0;
            for (int j = 0; j < k; j++) if (fabs(p[iN + 3 * j + 1]) != 0.0) slot++;
            tg_set_harvey(nz, slot, fabs(p[iN + 3 * k]), 1e-3 * fabs(p[iN + 3 * k + 1]), fabs(p[iN + 3 * k + 2]), d.x_last);
        }
    }
    __syncthreads();

    TG_STAMP(5);
    // ---------------- B. inclination -> height ratios (15 threads) ----------------
    if ((parts & TG_PRO_MODEL) && fam != 0 && tid < 16) {
        // tid -> (l, m): l=0:{0} l=1:{1..3} l=2:{4..8} l=3:{9..15}
        const int l = tid == 0 ? 0 : (tid < 4 ? 1 : (tid < 9 ? 2 : 3));
        const int m = tid - l * l - l;
        double r = 1.0;
        if (l > 0) {
            if (mc == 12) {                       // free ratios, models.cpp:1010-1030
                const int am = m < 0 ? -m : m;
                const int off = (l == 1) ? 0 : (l == 2 ? 2 : 5);
                r = fabs(p[iI + off + am]);
            } else if (mc == 13 || mc == 14) {
                r = 0.0;                          // heights come from Hlm parameters
            } else {
                const int present = (fam == 1) ? (lmax >= l) : (d.pl[2 + l] >= 1);
                r = present ? tg_amp_ratio(l, m, S.inclination) : 0.0;
            }
        }
        S.ratios[l][m + l] = r;
    }
    __syncthreads();

    TG_STAMP(6);
    // ---------------- C. one thread per component (the mode-level quantities are recomputed by the 2l+1
    //                    threads of a mode: redundant arithmetic, but no serial loop over m) ----------------
    int nmodes = 0, ncomp = 0;
    if (fam == 1) { const int L1 = (lmax > 3 ? 3 : lmax) + 1; nmodes = Nmax * L1; ncomp = Nmax * L1 * L1; }
    if (fam == 2) { nmodes = Nf; ncomp = Nfl0 + 3 * Nfl1 + 5 * Nfl2 + 7 * Nfl3; }
    if (ncomp > d.ncomp_max) { if (tid == 0) S.status |= TG_ST_MODE_OVERFLOW; nmodes = 0; ncomp = 0; }
    double work_acc = 0.0;
    int st_acc = 0, wlo_acc = 0x7fffffff, whi_acc = 0;
    for (int j = tid; j < ((parts & TG_PRO_MODEL) ? ncomp : 0); j += blockDim.x) {
        if (fam == 1) {
            const int L1 = (lmax > 3 ? 3 : lmax) + 1, L2 = L1 * L1;
            const int n = j / L2, r = j - n * L2;
            const int l = (r >= 9) ? 3 : ((r >= 4) ? 2 : ((r >= 1) ? 1 : 0));
            const int m = r - l * l - l;
            const int foff = Nmax + lmax + (l >= 1 ? Nfl0 : 0) + (l >= 2 ? Nfl1 : 0) + (l >= 3 ? Nfl2 : 0);
            const double fl = p[foff + n];
            const double *fl0_all = p + Nmax + lmax, *Wl0_all = p + iW;
            double Wl;
            if (mc == 9)       Wl = tg_appwidth_v1(p + iW, fl, S.numax);
            else if (mc == 10) Wl = tg_appwidth_v2(p + iW, fl);
            else if (l == 0)   Wl = fabs(Wl0_all[n]);
            else               Wl = fabs(tg_lin_interpol(fl0_all, Wl0_all, Nfl0 < Nmax ? Nfl0 : Nmax, fl));
            const int coff = n * L2 + l * l;
            if (mc == 13) {                        // models.cpp:1185-1237
                double hm;
                if (l == 0) hm = tg_height0(p[n], Wl, S.do_amp);
                else { double v = p[iI + (l + 1) * n + (m < 0 ? -m : m)]; if (S.do_amp) v = v / (TG_PI * Wl); hm = fabs(v); }
                tg_emit_comp(a, S, b, coff, l, m, fl, Wl, hm, S.a1, 0, 0, 0, work_acc, st_acc, wlo_acc, whi_acc);
            } else {
                const double Hl = (l == 0) ? tg_height0(p[n], Wl, S.do_amp) : tg_heightl(p[n], Wl, S.Vl[l], S.do_amp);
                const double hm = Hl * S.ratios[l][m + l];
                if (mc == 6 || mc == 7 || mc == 8) {
                    double a11, a12;
                    if (mc == 6)      { a11 = fabs(p[iS]); a12 = fabs(p[iS + 6]); }
                    else if (mc == 7) { a11 = fabs(p[iS + 6 + n]); a12 = a11; }
                    else              { a11 = fabs(p[iS + 6 + n]); a12 = fabs(p[iS + 6 + Nmax + n]); }
                    tg_emit_comp(a, S, b, coff, l, m, fl, Wl, hm, 0, a11, a12, 1, work_acc, st_acc, wlo_acc, whi_acc);
                } else {
                    tg_emit_comp(a, S, b, coff, l, m, fl, Wl, hm, S.a1, 0, 0, 0, work_acc, st_acc, wlo_acc, whi_acc);
                }
            }
        } else {                                   // local family, models.cpp:1762-1810 / :1888-1943
            int l = 0;
            while (l < 3 && j >= S.compStart[l + 1]) l++;
            const int r = j - S.compStart[l];
            const int n = r / (2 * l + 1), m = r - n * (2 * l + 1) - l;
            const double fl = p[iF + S.modeStart[l] + n];
            const double Wl = fabs(p[iW + S.modeStart[l] + n]);
            const int coff = S.compStart[l] + n * (2 * l + 1);
            double hm;
            if (mc == 11) {
                hm = tg_height0(p[S.modeStart[l] + n], Wl, S.do_amp) * S.ratios[l][m + l];
            } else {
                if (l == 0) hm = tg_height0(p[n], Wl, S.do_amp);
                else { double v = p[S.modeStart[l] + (l + 1) * n + (m < 0 ? -m : m)]; if (S.do_amp) v = v / (TG_PI * Wl); hm = fabs(v); }
            }
            tg_emit_comp(a, S, b, coff, l, m, fl, Wl, hm, S.a1, 0, 0, 0, work_acc, st_acc, wlo_acc, whi_acc);
        }
    }

    {   // one shared-memory atomic per warp (the sums are integer-valued, hence exact in any order)
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            work_acc += __shfl_xor_sync(0xffffffffu, work_acc, o); st_acc |= __shfl_xor_sync(0xffffffffu, st_acc, o);
            wlo_acc = min(wlo_acc, __shfl_xor_sync(0xffffffffu, wlo_acc, o)); whi_acc = max(whi_acc, __shfl_xor_sync(0xffffffffu, whi_acc, o));
        }
        if ((tid & 31) == 0) {
            if (work_acc != 0.0) atomicAdd(&S.work, work_acc);
            if (st_acc) atomicOr(&S.status, st_acc);
            if (whi_acc > wlo_acc) { atomicMin(&S.wlo, wlo_acc); atomicMax(&S.whi, whi_acc); }
        }
    }
    TG_STAMP(7);
    const bool do_prior = (parts & TG_PRO_PRIOR) != 0;
    // ---------------- D. prior terms, one thread per term ----------------
    int bad = 0;
    double *ex = terms + d.Np;                     // extra terms: [0..MAXORD) d02, then 4 x MAXORD smoothness
    int *ssw = (int *)(ex + 5 * TG_MAXORD);        // priors_names_switch staged for the serial tail
    // the prior terms go to the threads in REVERSE order: the low threads are busy with the modes (C),
    // and there is no barrier between C and D
    const int rt = blockDim.x - 1 - tid;
    if (do_prior) for (int i = rt; i < d.Np; i += blockDim.x) { const int swi = a.sw[i]; ssw[i] = swi; terms[i] = tg_generic_prior(swi, a.pri + 4 * i, p[i], &bad); }
    if (bad) atomicOr(&S.bad_prior, 1);
    double Dnu = 0;
    if (do_prior && d.prior_case == 2) {
        // Dnu = SUM of the adaptive first derivative of fl0 (priors_calc.cpp:87-88)
        const double *f0 = p + Nmax + lmax;
        if (Nfl0 >= 2) {
            Dnu = f0[1] - f0[0];
            for (int i = 0; i < Nfl0 - 2; i++) Dnu += (f0[i + 2] - f0[i]) / 2.;   // order: deriv[0], mid..., then last
            // Eigen .sum() runs index order 0..N-1: deriv[N-1] is added last
            Dnu += f0[Nfl0 - 1] - f0[Nfl0 - 2];
        }
        if (Nfl0 == Nfl2)
            for (int i = rt; i < Nfl0 && i < TG_MAXORD; i += blockDim.x)
                ex[i] = tg_logP_gu(0, Dnu / 3., 0.015 * Dnu, p[Nmax + lmax + i] - p[Nmax + lmax + Nfl0 + Nfl1 + i]);
        if ((int)d.extra[0] == 1) {
            const int off[4] = {Nmax + lmax, Nmax + lmax + Nfl0, Nmax + lmax + Nfl0 + Nfl1, Nmax + lmax + Nfl0 + Nfl1 + Nfl2};
            const int nfl[4] = {Nfl0, Nfl1, Nfl2, Nfl3};
            for (int e = rt; e < 4 * 32; e += blockDim.x) {              // one thread per (degree, order); orders beyond 32: loop below
                const int l = e >> 5; for (int i = e & 31; i < nfl[l] && i < TG_MAXORD; i += 32) {
                const double *y = p + off[l];
                const int n = nfl[l];
                {
                    double der;
                    if (n < 3) der = CUDART_NAN;                                  // reference would exit()
                    else if (i == 0) der = y[2] - 2. * y[1] + y[0];
                    else if (i == n - 1) der = y[n - 1] - 2. * y[n - 2] + y[n - 3];
                    else der = y[i + 1] - 2. * y[i] + y[i - 1];
                    ex[TG_MAXORD * (1 + l) + i] = tg_logP_gaussian(0, d.extra[1], der);
                }
                }
            }
        }
    }
    __syncthreads();

    TG_STAMP(8);
    // ---------------- E. sum of all terms (fixed-order block tree; the reference sums sequentially in
    //                    long double -- same value to ~1e-15 relative), then thread 0 applies the hard limits ----
    if (do_prior) {
        double part = 0.0;
        for (int i = tid; i < d.Np; i += blockDim.x) part += terms[i];
        if (d.prior_case == 2) {
            if (Nfl0 == Nfl2) for (int i = tid; i < Nfl0 && i < TG_MAXORD; i += blockDim.x) part += ex[i];
            if ((int)d.extra[0] == 1) {
                const int nfl[4] = {Nfl0, Nfl1, Nfl2, Nfl3};
                for (int l = 0; l < 4; l++) for (int i = tid; i < nfl[l] && i < TG_MAXORD; i += blockDim.x) part += ex[TG_MAXORD * (1 + l) + i];
            }
        }
        part = tg_warp_sum(part);
        if ((tid & 31) == 0) s_red[tid >> 5] = part;
        __syncthreads();
    }
    if (tid == 0 && do_prior) {
        double f = 0.0;
        for (int w = 0; w < (int)(blockDim.x >> 5); w++) f += s_red[w];
        if (d.prior_case == 2) {            // priors_MS_Global, priors_calc.cpp:24-172
            const int impose = (int)d.extra[3];
            if (impose == 1) {
                const int bI = iI;
                f = f + tg_logP_uniform(0, 1. + 1e-10, p[bI] + 2 * p[bI + 1]);
                f = f + tg_logP_uniform(0, 1 + 1e-10, p[bI + 2] + 2 * p[bI + 3] + 2 * p[bI + 4]);
                f = f + tg_logP_uniform(0, 1 + 1e-10, p[bI + 5] + 2 * p[bI + 6] + 2 * p[bI + 7] + 2 * p[bI + 8]);
            } else if (impose == 2) S.bad_prior = 1;
            for (int i = Nmax; i <= Nmax + lmax; i++) if (p[i] < 0) f = tg_ninf();
            if (fabs(p[iS + 2] / p[iS]) >= d.extra[2]) f = tg_ninf();
            if (p[iS + 4] < 0) f = tg_ninf();
            const int w = iS + Nsplit + Nwidth;
            if (ssw[w + 3] != 0) if ((p[w + 3] < 0) || (p[w + 4] < 0) || (p[w + 5] < 0)) f = tg_ninf();
            if (ssw[w + 6] != 0) if ((p[w + 6] < 0) || (p[w + 7] < 0) || (p[w + 8] < 0)) f = tg_ninf();
            if ((ssw[iS + 9] != 0) && (p[w + 9] < 0)) f = tg_ninf();
        } else if (d.prior_case == 3) {     // priors_local, priors_calc.cpp:175-280
            if (p[iS] != 0) {
                if (fabs(p[iS + 2] / p[iS]) >= d.extra[2]) f = tg_ninf();
            } else if ((p[iS + 3] != 0) && (p[iS + 4] != 0)) {
                if (fabs(p[iS + 2] / (p[iS + 3] * p[iS + 3] + p[iS + 4] * p[iS + 4])) >= d.extra[2]) f = tg_ninf();
            }
            if ((ssw[iI] != 0) && (p[iI] < 0)) f = tg_ninf();
        }
        a.t.logprior[b] = f;
        a.t.pstatus[b] = S.bad_prior ? TG_ST_BAD_PRIOR : 0;
    }
    if (tid == 0 && (parts & TG_PRO_MODEL)) {
#ifdef TG_PROFILE
        if (blockIdx.x == 0 && blockIdx.y == 0) tg_dbg_clk[9] = clock64();
#endif
        a.t.ncomp[b] = ncomp;
        a.t.wlo[b] = S.wlo; a.t.whi[b] = S.whi;
        a.t.work[b] = S.work;
        a.t.status[b] = S.status;
    }
}

#define TG_PRO_SMEM(Np) ((size_t)(2 * (Np) + 5 * TG_MAXORD) * sizeof(double) + (size_t)(((Np) + 1) & ~1) * sizeof(int))

__global__ void __launch_bounds__(TG_PRO_THREADS) tg_k_prologue(TgProArgs a)
{
    extern __shared__ __align__(16) double tg_dyn_smem[];
    tg_prologue_body(a, blockIdx.x, tg_dyn_smem);
}
