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
    double *noise;             // [B][TG_NOISE_STRIDE] : per Harvey (H, a=1e-3*tc, p, ln a), then N0, asym, gaussA, gaussNu, gaussS2, flags
    int    *status;            // [B]  TG_ST_* bits
    double *work;              // [B]  L = sum (2l+1)*(hi-lo)
    double *logprior;          // [B]
};
#define TG_NOISE_STRIDE (4 * TG_MAXH + 8)
#define TG_N_N0     (4 * TG_MAXH + 0)
#define TG_N_ASYM   (4 * TG_MAXH + 1)
#define TG_N_GA     (4 * TG_MAXH + 2)
#define TG_N_GNU    (4 * TG_MAXH + 3)
#define TG_N_GS2    (4 * TG_MAXH + 4)
#define TG_N_NH     (4 * TG_MAXH + 5)   // number of Harvey slots (as double)
#define TG_N_GAUSS  (4 * TG_MAXH + 6)   // 0 none, 1 gaussian term present

// ---------------------------------------------------------------------------------------
// device helpers
// ---------------------------------------------------------------------------------------
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
