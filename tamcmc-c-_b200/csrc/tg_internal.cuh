// tg_internal.cuh -- shared device/host declarations of libtamcmc_gpu (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <math.h>
#include <math_constants.h>
#include "../../include/tamcmc_gpu.h"

#define TG_THREADS      256      // threads per CTA of the fused model+likelihood kernel
#define TG_MAXH         8        // Harvey terms per chain the noise table holds
#define TG_PRO_THREADS  128      // threads per CTA of the per-chain prologue
#define TG_MAXORD       128      // max frequencies per degree (prior derivative scratch)

// Problem description passed BY VALUE to the kernels (plain data, < 4 KB).
struct TgDesc {
    int    model_case, like_case, prior_case;
    int    pl[11];
    int    Np, Nv;
    int    ncomp_max;          // Lorentzian components per chain the table holds
    int    nmodes_max;
    long   likelihood_p;       // (long) likelihood_params, model_def.cpp:300-301
    double extra[4];
    // grid geometry (global grid for the window arithmetic, local shard for evaluation)
    int    Nx;                 // local bins
    long   Nx_global;
    long   shard_lo;
    double x_first, x_second, x_last;
};

// Per-chain tables produced by the prologue and consumed by the fused kernel (SoA in HBM).
struct TgTables {
    double *nu, *h, *q;        // [B][ncomp_max]  centre, height (H*V or Hlm), 4/Gamma^2
    double *ifc, *t2sq;        // [B][ncomp_max]  asymmetry extras: 1/fc, (0.5*Gamma*asym/fc)^2
    int    *lo, *hi;           // [B][ncomp_max]  window [lo,hi) in LOCAL bin indices
    int    *ncomp;             // [B]
    int    *wlo, *whi;         // [B]  union [wlo,whi) of all truncation windows in LOCAL bins (empty: wlo >= whi)
    double *noise;             // [B][TG_NOISE_STRIDE] : per ACTIVE Harvey term (H, a=1e-3*tc, p, ln a, a^p clamped, exponent class), then N0, asym, gaussA, gaussNu, gaussS2, flags
    int    *status;            // [B]  TG_ST_* bits of the model part of the prologue
    int    *pstatus;           // [B]  TG_ST_BAD_PRIOR or 0, written by the prior part (it may run in another kernel)
    double *work;              // [B]  L = sum (2l+1)*(hi-lo)
    double *logprior;          // [B]
};
#define TG_HS       6                    // doubles per Harvey term
#define TG_H_H      0
#define TG_H_A      1
#define TG_H_P      2
#define TG_H_LNA    3
#define TG_H_AP     4                    // a^p for the integer classes (clamped), else a
#define TG_H_CLS    5                    // exponent class: 0..4 exact small integer, -1 general
#define TG_NOISE_STRIDE (TG_HS * TG_MAXH + 8)
#define TG_N_N0     (TG_HS * TG_MAXH + 0)
#define TG_N_ASYM   (TG_HS * TG_MAXH + 1)
#define TG_N_GA     (TG_HS * TG_MAXH + 2)
#define TG_N_GNU    (TG_HS * TG_MAXH + 3)
#define TG_N_GS2    (TG_HS * TG_MAXH + 4)
#define TG_N_NH     (TG_HS * TG_MAXH + 5)   // number of ACTIVE Harvey terms (as double)
#define TG_N_GAUSS  (TG_HS * TG_MAXH + 6)   // 0 none, 1 gaussian term present

// ---------------------------------------------------------------------------------------
// device helpers
// ---------------------------------------------------------------------------------------
// Programmatic dependent launch (sm_90+): a kernel launched with the programmatic-serialization attribute may become
// resident while its predecessor in the stream still runs; tg_pdl_wait() blocks until the predecessor grid has completed
// and its writes are visible (a no-op for a kernel launched without the attribute); tg_pdl_trigger() lets the successor
// become resident (every CTA must call it or exit).  All three step kernels trigger only AFTER their own wait, so at most
// one kernel runs ahead and nothing of a successor executes past its wait before the predecessor is complete.
__device__ __forceinline__ void tg_pdl_wait()    { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void tg_pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// Per-star kernel arguments live in DEVICE memory, one array per kernel indexed by star (a launch may span several stars:
// cfg5, several tempered ensembles of different spectra per GPU); a CTA copies the record of its star into shared memory.
// The arrays are written by the host long before the launch, never by a kernel, so they may be read before tg_pdl_wait().
template <class A> __device__ __forceinline__ void tg_load_args(A *dst, const A *src)
{
    static_assert(sizeof(A) % 8 == 0, "argument records are copied in 8-byte words");
    const unsigned long long *s = (const unsigned long long *)src;
    unsigned long long *d = (unsigned long long *)dst;
    for (int i = threadIdx.x; i < (int)(sizeof(A) / 8); i += blockDim.x) d[i] = __ldg(s + i);
    __syncthreads();
}

// Reciprocal: MUFU.RCP64H seed (2^-23 rel. error) + 2 Newton steps = full fp64 precision
// (<= 1 ulp; not IEEE-exact, which the 1e-10 parity budget does not need).  Special values:
// t=inf -> 0, t=0 -> inf, NaN -> NaN come out right because the seed is already exact there
// and the Newton correction is skipped when it would produce NaN from inf*0.
__device__ __forceinline__ double tg_rcp(double t)
{
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(t));
    double e = fma(-t, r, 1.0);
    double r1 = fma(r, e, r);
    e = fma(-t, r1, 1.0);
    double r2 = fma(r1, e, r1);
    // inf*0 = NaN in the residual when t is 0/inf/denormal-flushed: keep the seed there
    return (r2 == r2) ? r2 : ((t == t) ? r : t);
}

// Fast reciprocal for arguments known to be finite, normal and >= ~1e-290 in magnitude (products
// of 1+profile terms, products of positive model values): MUFU.RCP64H seed r (relative error
// e <= ~2^-20) and ONE cubic correction r(1+e+e^2) -> truncation error e^3 <= 2^-60, i.e. <= 1 ulp
// after rounding.  3 DFMA instead of the 4 of two Newton steps; no special-value handling
// (inf/0 give NaN, which the likelihood turns into a rejected proposal).
__device__ __forceinline__ double tg_rcp_fast(double t)
{
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(t));
    const double e = fma(-t, r, 1.0);
    const double p = fma(e, e, e);
    return fma(r, p, r);
}

// Harvey exponent class: 0..4 for exact small integers (plain products), -1 otherwise
__device__ __forceinline__ int tg_pclass(double p)
{
    return (p == 0.0) ? 0 : (p == 1.0) ? 1 : (p == 2.0) ? 2 : (p == 3.0) ? 3 : (p == 4.0) ? 4 : -1;
}
// a^ip, clamped so that a^ip * xmax^ip <= 1e140 (keeps products of two Harvey denominators finite;
// a clamped term contributes < 1e-140 H, the reference's own value there is smaller still)
__device__ __forceinline__ double tg_apow(double a, int ip, double xmax)
{
    if (ip < 1) return a;
    const double a2 = a * a, x2 = xmax * xmax;
    const double ap = (ip == 1) ? a : (ip == 2) ? a2 : (ip == 3) ? a2 * a : a2 * a2;
    const double xp = (ip == 1) ? xmax : (ip == 2) ? x2 : (ip == 3) ? x2 * xmax : x2 * x2;
    const double lim = 1e140 / fabs(xp);
    return (ap > lim) ? lim : ap;
}

// one active Harvey term of the noise table: y += H/(1+(a x)^p), a = 1e-3*tc (noise_models.cpp:32)
__device__ __forceinline__ void tg_set_harvey(double *nz, int slot, double H, double a, double p, double xmax)
{
    double *q = nz + TG_HS * slot;
    const int cls = tg_pclass(p);
    q[TG_H_H] = H; q[TG_H_A] = a; q[TG_H_P] = p; q[TG_H_LNA] = log(a);
    q[TG_H_AP] = tg_apow(a, cls, xmax); q[TG_H_CLS] = (double)cls;
}

// Philox4x32-10 (Salmon et al. 2011), counter-based: no shared RNG state between chains
// (the reference shares rand() and Box-Muller statics across OpenMP threads, quirk Q10).
__device__ __forceinline__ uint4 tg_philox(uint4 c, uint2 k)
{
#pragma unroll
    for (int r = 0; r < 10; r++) {
        const uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
        c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
        k.x += 0x9E3779B9u; k.y += 0xBB67AE85u;
    }
    return c;
}
// 53-bit uniform in (0,1)
__device__ __forceinline__ double tg_u01(uint32_t hi, uint32_t lo)
{
    const double v = (double)(hi >> 5) * 67108864.0 + (double)(lo >> 6);
    return (v + 0.5) * (1.0 / 9007199254740992.0);
}

__device__ __forceinline__ double tg_warp_sum(double v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    return v;
}
