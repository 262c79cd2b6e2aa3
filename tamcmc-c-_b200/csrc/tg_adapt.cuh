// tg_adapt.cuh -- adaptive-proposal state and its speculative update (Robbins-Monro + Cholesky).
//
// Reference (under /root/reference/tamcmc/sources/MALA.cpp): update_proposal :292-315, p1/p2/p3 :135-176,
// new_prop_values :335-353 (llt of (Sigma + eps2 I) sigma), execute :608-737 (gamma = c0/(1+i), learning zones).
#pragma once
#include "tg_internal.cuh"

#ifdef TG_PROFILE
__device__ long long tg_dbg_spec[16];
#define TG_SSTAMP(k) do { __syncthreads(); if (threadIdx.x == 0 && blockIdx.x == 0) tg_dbg_spec[k] = clock64(); } while (0)
#else
#define TG_SSTAMP(k) do { } while (0)
#endif

#define TG_TAG_PROP 0x50524f50u
#define TG_TAG_MH   0x4d485f55u
#define TG_TAG_PT   0x50545f55u
#define TG_MAX_LEARN 16

struct TgMala {
    double c0, eps1, eps2, A1, target;
    int    n_learn;
    long   Nt_learn[TG_MAX_LEARN + 1];
    long   periods[TG_MAX_LEARN];
    long   dN_mixing;
    unsigned long long seed;
    int    flags;
    int    chain0, ens0;   // offsets of the chain / ensemble indices in the Philox counters (tg_problem.philox_chain0 / philox_ensemble0)
    int    pad_;
};

struct TgCtrl {            // lives in device memory so that a captured CUDA graph can be replayed
    long i;                // current iteration
    long i0;               // first iteration of the current record chunk
    unsigned int ticket;   // CTA completion counter of the finish kernel
    int pad;
};

struct TgState {
    int C, E, Ce, Nv, Np;
    double *vars_cur, *vars_prop;          // [C][Nv]
    double *params_cur, *params_prop;      // [C][Np]
    double *logL_cur, *logPrior_cur, *logPost_cur, *Pmove;   // [C]
    int    *moved;                         // [C]
    // Adaptive proposal.  (mu, Sigma, chol(Sigma+eps2 I)) live in THREE slots per chain: the live one
    // and the two candidates of the running step ("proposal rejected" / "proposal accepted"), which the
    // speculative kernel fills while the fused model+likelihood kernel runs; the finish kernel only
    // switches live[c].  Slot-major: mu [3][C][Nv], cov / chol [3][C][Nv][Nv] (row-major, lower part of chol).
    double *mu, *cov, *chol;
    double *sigma;                         // [C]
    int    *live;                          // [C]  live slot 0..2
    int    *chol_fail;                     // [3][C]
    int    *sticky_status;                 // [C]
    unsigned int *ens_ticket;              // [E]  CTA completion counters of the finish kernel, one per ensemble
    const int    *index_to_relax;          // [Nv]
    const double *Tcoefs;                  // [C]
    TgCtrl *ctrl;
    // Running sums of (sigma, mu, Sigma, vars) over the steps since the last drain: the *_mean blocks of the restore files
    // (outputs.cpp:1758-1818).  NULL unless the problem asks for them (TG_F_RESTORE_MEANS).
    double *acc_sigma, *acc_mu, *acc_cov, *acc_vars;   // [C], [C][Nv], [C][Nv][Nv], [C][Nv]
    long long *acc_count;                              // steps summed
};

__device__ __forceinline__ size_t tg_slot_vec(const TgState &s, int slot, int c) { return ((size_t)slot * s.C + c) * s.Nv; }
__device__ __forceinline__ size_t tg_slot_mat(const TgState &s, int slot, int c) { return ((size_t)slot * s.C + c) * s.Nv * s.Nv; }

// learning zone of iteration it?  (MALA.cpp:641-652)
__device__ __forceinline__ int tg_learning(const TgMala &m, long it)
{
    int learn = 0, which = 0;
    for (int l = 0; l < m.n_learn; l++) if (it >= m.Nt_learn[l] && it < m.Nt_learn[l + 1]) { learn = 1; which = l; }
    if (learn && (it % m.periods[which]) != 0) learn = 0;
    return learn;
}

// 1/sqrt(x) for the Cholesky pivots: MUFU.RSQ64H seed (relative error e0 < 2^-20) + one third-order step
// y = y0 (1 + e/2 + 3e^2/8), e = 1 - x y0^2 (residual error ~ e0^3: full fp64 precision; four dependent
// operations, the pivot sits on the critical path of every column).  x <= 0 / NaN give NaN or inf, which
// the caller flags as a failed factorisation.
__device__ __forceinline__ double tg_rsqrt_fast(double x)
{
    double y;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
    const double e = fma(-(x * y), y, 1.0);
    return fma(y * e, fma(e, 0.375, 0.5), y);
}

// Loads / stores of the matrix being factorised: SH = true -> explicit shared-memory instructions
// (the generic-pointer form costs an address-space check per access and, worse, made the compiler wrap
// every update in its own branch), SH = false -> plain global accesses.
template <bool SH> __device__ __forceinline__ double tg_mld(const double *p)
{
    if (SH) { double v; asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"((unsigned)__cvta_generic_to_shared(p))); return v; }
    return *p;
}
template <bool SH> __device__ __forceinline__ void tg_mst(double *p, double v)
{
    if (SH) asm volatile("st.shared.f64 [%0], %1;" :: "r"((unsigned)__cvta_generic_to_shared(p)), "d"(v) : "memory");
    else *p = v;
}

// In-CTA Cholesky (lower, right-looking) of the n x n matrix S (leading dimension ld; shared memory when
// SH, else global), 256 threads as a 16x16 grid, ONE barrier per column: while the rank-1 update of
// column j runs, every thread recomputes the next pivot from the double-buffered diagonal (its reciprocal
// square root is in flight during the update), so the freshly updated entries of column j+1 are scaled in
// the same pass.  The update walks the 16x16 tiles of the trailing triangle (its cost follows the
// shrinking block).  On exit the strictly lower part and the diagonal of S hold L.
// work: 4*n doubles of SHARED memory (two column buffers, two diagonal buffers).  Returns 1 if a pivot
// was <= 0 / NaN.
template <bool SH>
__device__ int tg_cholesky_inplace(double *S, int n, int ld, double *work)
{
    __shared__ int s_fail;
    const int tid = threadIdx.x, ty = tid >> 4, tx = tid & 15;
    double *col0 = work, *dg0 = work + 2 * n;     // column buffers col0 + (j&1)*n, diagonal buffers dg0 + (j&1)*n
    if (tid == 0) s_fail = 0;
    for (int i = tid; i < n; i += blockDim.x) dg0[i] = tg_mld<SH>(S + (size_t)i * ld + i);
    __syncthreads();
    {   // column 0
        const double pv = dg0[0];
        const double inv = tg_rsqrt_fast(pv);
        if (tid == 0) { if (!(pv > 0.0)) s_fail = 1; tg_mst<SH>(S, pv * inv); }
        for (int i = 1 + tid; i < n; i += blockDim.x) { const double v = tg_mld<SH>(S + (size_t)i * ld) * inv; tg_mst<SH>(S + (size_t)i * ld, v); col0[i] = v; }
    }
    __syncthreads();
    for (int j = 0; j + 1 < n; j++) {
        const double *cj = col0 + (j & 1) * n; double *cn = col0 + ((j + 1) & 1) * n;
        const double *dj = dg0 + (j & 1) * n;  double *dn = dg0 + ((j + 1) & 1) * n;
        // next pivot, recomputed by every thread with the same arithmetic as the diagonal update below
        const double cnext = cj[j + 1];
        const double pv = fma(-cnext, cnext, dj[j + 1]);
        const double inv = tg_rsqrt_fast(pv);
        if (tid == 0) { if (!(pv > 0.0)) s_fail = 1; tg_mst<SH>(S + (size_t)(j + 1) * ld + (j + 1), pv * inv); }
        for (int t = j + 1 + tid; t < n; t += blockDim.x) { const double c = cj[t]; dn[t] = fma(-c, c, dj[t]); }
        // strictly-lower trailing update; column j+1 is scaled on the fly
        const int r = n - j - 1;
        const double *cb = cj + j + 1;            // cb[ii] = L[j+1+ii][j]
        double *Sb = S + (size_t)(j + 1) * ld + (j + 1);   // trailing block origin
        const int nU = (r + 15) >> 4;             // 16x16 tiles of the trailing triangle, one element per thread and tile
        for (int U = 0; U < nU; U++) {
            const int ii = 16 * U + ty, iic = min(ii, r - 1);
            const double ci = cb[iic];
            double *rowp = Sb + (size_t)iic * ld;
            for (int V0 = 0; V0 <= U; V0 += 4) {          // four tiles at a time: all loads first, then the stores
                double old[4], ck[4];
#pragma unroll
                for (int q = 0; q < 4; q++) {
                    const int kkc = min(16 * (V0 + q) + tx, r - 1);
                    ck[q] = cb[kkc];
                    old[q] = tg_mld<SH>(rowp + kkc);       // clamped: a valid address, result discarded when off
                }
#pragma unroll
                for (int q = 0; q < 4; q++) {
                    const int kk = 16 * (V0 + q) + tx;
                    double val = fma(-ci, ck[q], old[q]);
                    if (kk == 0) val *= inv;
                    if ((ii < r) && (kk < ii)) { tg_mst<SH>(rowp + kk, val); if (kk == 0) cn[j + 1 + ii] = val; }
                }
            }
        }
        __syncthreads();
    }
    return s_fail;
}

// Register-resident Cholesky for n <= 96 (the usual case: ~100 free parameters).  256 threads as a
// 16x16 grid: thread (ty,tx) OWNS the strictly-lower elements a[u][v] = A(tx+16u, ty+16v), v <= u < 6, in
// registers for the whole factorisation (0 where the position is outside the strict lower triangle), so
// a column step is at most 12 shared-memory loads (the thread's row and column entries of L[:,j]) and 21
// UNPREDICATED FMAs: the column vector is zero above its first valid row, so finished columns and rows
// are updated by exactly 0.  The column loop is straight-line code (a taken-or-not branch costs a
// two-warp scheduler ~13 cycles, more than the arithmetic it would skip): six phases, one per 16-column
// block, each compiled for its own set of live blocks, two columns per trip so that the buffer parity is a
// compile-time constant.  Column j+1 belongs to half of ONE warp (ty == (j+1) mod 16): only that warp
// runs the pivot chain (1/sqrt from the double-buffered diagonal, in flight during its updates) and
// publishes the scaled column; the other seven go straight to the barrier.  One barrier per column.
// The registers keep the UNSCALED columns: L(i,k) = a * Li[k] on exit.
#ifdef TG_CHOL_DBG
__device__ long long tg_chol_dbg[8];
#define TG_CSTAMP(k) if (threadIdx.x == 0) tg_chol_dbg[k] = clock64()
#else
#define TG_CSTAMP(k)
#endif
struct TgChol96 {           // shared-memory side of tg_chol_reg96 (statically allocated: constant addresses)
    double col[2][96];      // column j of L, double-buffered by the parity of j; zero above its first valid row
    double dg[2][96];       // running diagonal, same parity scheme; dg[1] = diag(A) on entry
    double Ld[96], Li[96];  // diag(L) and its reciprocal
};

template <int VMIN, int PJ>
__device__ __forceinline__ void tg_chol96_col(double (&a)[6][6], TgChol96 &W, const int j, int &bad)
{
    constexpr int PN = PJ ^ 1;
    const int tid = threadIdx.x, ty = tid >> 4, tx = tid & 15;
    const int sty = (j + 1) & 15;                 // ty of the half-warp that owns column j+1 (block column VMIN)
    double ci[6], ck[6];
    const double cnext = W.col[PJ][j + 1], dnext = W.dg[PJ][j + 1];
#pragma unroll
    for (int u = VMIN; u < 6; u++) { ci[u] = W.col[PJ][tx + 16 * u]; ck[u] = W.col[PJ][ty + 16 * u]; }
    double dcur = 0.0, ccur = 0.0;
    if (tid < 96) { dcur = W.dg[PJ][tid]; ccur = W.col[PJ][tid]; }
    if ((tid >> 5) == (sty >> 1)) {               // the owning warp: pivot chain in flight during its updates
        const double pv = fma(-cnext, cnext, dnext);
        const double inv = tg_rsqrt_fast(pv);
#pragma unroll
        for (int v = VMIN; v < 6; v++)
#pragma unroll
            for (int u = v; u < 6; u++) a[u][v] = fma(-ci[u], ck[v], a[u][v]);
        const bool own = (ty == sty);
#pragma unroll
        for (int u = VMIN; u < 6; u++) {
            const int ii = tx + 16 * u;           // rows >= n hold zeros
            if (own && ii > j + 1) W.col[PN][ii] = a[u][VMIN] * inv;
        }
        if (own && tx == 0) { W.Ld[j + 1] = pv * inv; W.Li[j + 1] = inv; W.col[PN][j + 1] = 0.0; W.col[PN][max(j, 0)] = 0.0; }
        bad |= !(pv > 0.0);
    } else {
#pragma unroll
        for (int v = VMIN; v < 6; v++)
#pragma unroll
            for (int u = v; u < 6; u++) a[u][v] = fma(-ci[u], ck[v], a[u][v]);
    }
    if (tid < 96) W.dg[PN][tid] = fma(-ccur, ccur, dcur);
    __syncthreads();
}

template <int VMIN>
__device__ __forceinline__ void tg_chol96_phase(double (&a)[6][6], const int n, TgChol96 &W, int &bad)
{
    const int jend = min(16 * VMIN + 15, n - 1);  // columns j = 16 VMIN - 1 .. jend - 1 (odd first: parity 1)
#pragma unroll 1
    for (int j = 16 * VMIN - 1; j < jend; j += 2) {
        tg_chol96_col<VMIN, 1>(a, W, j, bad);     // j = -1: no update (col[1] = 0), publishes column 0
        if (j + 1 < jend) tg_chol96_col<VMIN, 0>(a, W, j + 1, bad);
    }
}

// a[][]: see above; W.col zero and W.dg[1] = diag(A) (0 beyond n) on entry.  Returns 1 if a pivot was
// <= 0 / NaN.  Contains barriers; all 256 threads must call.
__device__ int tg_chol_reg96(double (&a)[6][6], int n, TgChol96 &W)
{
    int bad = 0;
    TG_CSTAMP(0);
    tg_chol96_phase<0>(a, n, W, bad); TG_CSTAMP(1);
    if (n > 16) tg_chol96_phase<1>(a, n, W, bad); TG_CSTAMP(2);
    if (n > 32) tg_chol96_phase<2>(a, n, W, bad); TG_CSTAMP(3);
    if (n > 48) tg_chol96_phase<3>(a, n, W, bad); TG_CSTAMP(4);
    if (n > 64) tg_chol96_phase<4>(a, n, W, bad); TG_CSTAMP(5);
    if (n > 80) tg_chol96_phase<5>(a, n, W, bad); TG_CSTAMP(6);
    return __syncthreads_or(bad);
}

// wrapper around tg_chol_reg96 for a matrix held in memory (S shared when SH)
template <bool SH>
__device__ int tg_cholesky_reg96(double *S, int n, int ld)
{
    __shared__ TgChol96 W;
    const int tid = threadIdx.x, ty = tid >> 4, tx = tid & 15;
    double a[6][6];
#pragma unroll
    for (int u = 0; u < 6; u++)
#pragma unroll
        for (int v = 0; v <= u; v++) {
            const int ii = tx + 16 * u, kk = ty + 16 * v;
            a[u][v] = (ii < n && kk < ii) ? tg_mld<SH>(S + (size_t)ii * ld + kk) : 0.0;
        }
    for (int i = tid; i < 2 * 96; i += blockDim.x) { (&W.col[0][0])[i] = 0.0; (&W.dg[0][0])[i] = 0.0; }
    __syncthreads();
    for (int i = tid; i < n; i += blockDim.x) W.dg[1][i] = tg_mld<SH>(S + (size_t)i * ld + i);
    __syncthreads();
    const int fail = tg_chol_reg96(a, n, W);
#pragma unroll
    for (int u = 0; u < 6; u++)
#pragma unroll
        for (int v = 0; v <= u; v++) {
            const int ii = tx + 16 * u, kk = ty + 16 * v;
            if (ii < n && kk < ii) tg_mst<SH>(S + (size_t)ii * ld + kk, a[u][v] * W.Li[kk]);
        }
    for (int i = tid; i < n; i += blockDim.x) tg_mst<SH>(S + (size_t)i * ld + i, W.Ld[i]);
    __syncthreads();
    return fail;
}

// Packed lower triangle, column by column: column k starts at k*n - k(k-1)/2 and holds rows k..n-1.
// (Thread-per-row consumers read consecutive addresses for a fixed column.)
__device__ __forceinline__ int tg_colpack(int n, int k) { return k * n - (k * (k - 1)) / 2; }

// One Cholesky job: L = chol(cov + eps2*I) of (slot, chain) -> chol[slot][c]; staged in shared memory
// when it fits (dyn = shared-memory buffer of Nv*ld + 4*Nv doubles, or NULL: factorise in place in HBM).
__device__ void tg_factor_slot(const TgState &s, double eps2, int slot, int c, double *dyn, double *work)
{
    const int Nv = s.Nv, tid = threadIdx.x, nt = blockDim.x;
    const double *cv = s.cov + tg_slot_mat(s, slot, c);
    double *Lg = s.chol + tg_slot_mat(s, slot, c);
    const int ld = dyn ? (Nv | 1) : Nv;
    double *L = dyn ? dyn : Lg;
    for (int e = tid; e < Nv * Nv; e += nt) {
        const int i = e / Nv, k = e - i * Nv;
        if (k <= i) L[i * ld + k] = cv[e] + ((i == k) ? eps2 : 0.0);
    }
    __syncthreads();
    TG_SSTAMP(3);
    const int fail = dyn ? tg_cholesky_inplace<true>(L, Nv, ld, work) : tg_cholesky_inplace<false>(L, Nv, ld, work);
    TG_SSTAMP(4);
    // staged factor goes out PACKED by columns (tg_colpack): the propose kernel copies it flat into shared memory
    if (dyn) for (int k = tid >> 5; k < Nv; k += nt >> 5) for (int i = k + (tid & 31); i < Nv; i += 32) Lg[tg_colpack(Nv, k) + i - k] = L[i * ld + k];
    if (tid == 0) s.chol_fail[slot * s.C + c] = fail;
}

// Fixed-order block sum (warp shuffle tree, then the eight warp sums in order); all 256 threads call.
__device__ __forceinline__ double tg_block_sum256(double v, double *s_red)
{
    v = tg_warp_sum(v);
    __syncthreads();                              // s_red of the previous call is dead
    if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = v;
    __syncthreads();
    double t = 0.0;
#pragma unroll
    for (int w = 0; w < 8; w++) t += s_red[w];
    return t;
}

// Adaptation + factorisation for Nv <= 96, everything in registers (ownership as in tg_chol_reg96): the
// new Sigma is formed in the registers that the factorisation then consumes; Sigma is read from and
// written to HBM once (both halves are written from the one value, so it stays exactly symmetric; the
// update reads the upper half, which is contiguous along tx).  update = 0: factorise the live slot as it
// is (create / restore; reads the LOWER half as the reference's llt() does and mirrors it).
// Factor layout: packed by columns (tg_colpack).
__device__ void tg_spec96(const TgState &s, const TgMala &m, const int update, const int c, const int src, const int dst,
                          const double gamma, const double *xv)
{
    __shared__ TgChol96 W;
    __shared__ double s_d[96];
    __shared__ double s_red[8];
    const int tid = threadIdx.x, ty = tid >> 4, tx = tid & 15, n = s.Nv;
    const double *cov0 = s.cov + tg_slot_mat(s, src, c);
    double *cov = s.cov + tg_slot_mat(s, dst, c), *Lg = s.chol + tg_slot_mat(s, dst, c);
    double a[6][6], dgv = 0.0;
    TG_SSTAMP(0);
#pragma unroll
    for (int u = 0; u < 6; u++)
#pragma unroll
        for (int v = 0; v <= u; v++) {
            const int ii = tx + 16 * u, kk = ty + 16 * v;
            a[u][v] = (ii < n && kk < ii) ? (update ? cov0[(size_t)kk * n + ii] : cov0[(size_t)ii * n + kk]) : 0.0;
        }
    if (tid < n) dgv = cov0[(size_t)tid * n + tid];
    for (int i = tid; i < 2 * 96; i += 256) { (&W.col[0][0])[i] = 0.0; (&W.dg[0][0])[i] = 0.0; }
    double scale = 1.0;
    if (update) {
        // mu <- p3(mu + gamma (x - mu))   (MALA.cpp:292-315, :135-176)
        const double *mu0 = s.mu + tg_slot_vec(s, src, c);
        double *mu = s.mu + tg_slot_vec(s, dst, c);
        double mn = 0.0, xi = 0.0;
        if (tid < n) { xi = xv[tid]; const double m0 = mu0[tid]; mn = m0 + gamma * (xi - m0); }
        const double nrm3 = sqrt(tg_block_sum256(mn * mn, s_red));
        if (!(nrm3 <= m.A1)) mn = mn * (m.A1 / nrm3);
        if (tid < 96) s_d[tid] = (tid < n) ? xi - mn : 0.0;
        if (tid < n) mu[tid] = mn;
        __syncthreads();
        TG_SSTAMP(1);
        // Sigma <- p2(Sigma + gamma ((x-mu)(x-mu)^T - Sigma))
        double di[6], dk[6], part = 0.0;
#pragma unroll
        for (int u = 0; u < 6; u++) { di[u] = s_d[tx + 16 * u]; dk[u] = s_d[ty + 16 * u]; }
#pragma unroll
        for (int u = 0; u < 6; u++)
#pragma unroll
            for (int v = 0; v <= u; v++) {
                const int ii = tx + 16 * u, kk = ty + 16 * v;
                if (ii < n && kk < ii) { const double o = a[u][v]; const double cn = o + gamma * (di[u] * dk[v] - o); a[u][v] = cn; part = fma(cn, cn, part); }
            }
        part *= 2.0;                              // both halves
        if (tid < n) { const double dd = s_d[tid]; dgv = dgv + gamma * (dd * dd - dgv); part = fma(dgv, dgv, part); }
        const double nrm2 = sqrt(tg_block_sum256(part, s_red));
        if (!(nrm2 <= m.A1)) scale = m.A1 / nrm2;
        if (scale != 1.0) {
#pragma unroll
            for (int u = 0; u < 6; u++)
#pragma unroll
                for (int v = 0; v <= u; v++) a[u][v] *= scale;
            dgv *= scale;
        }
    }
    {                                             // write Sigma[dst] (update), or mirror the lower half (create / restore)
#pragma unroll
        for (int u = 0; u < 6; u++)
#pragma unroll
            for (int v = 0; v <= u; v++) {
                const int ii = tx + 16 * u, kk = ty + 16 * v;
                if (ii < n && kk < ii) { cov[(size_t)kk * n + ii] = a[u][v]; if (update) cov[(size_t)ii * n + kk] = a[u][v]; }
            }
        if (update && tid < n) cov[(size_t)tid * n + tid] = dgv;
    }
    __syncthreads();                              // W zeroed
    if (tid < n) W.dg[1][tid] = dgv + m.eps2;
    __syncthreads();
    TG_SSTAMP(2);
    const int fail = tg_chol_reg96(a, n, W);
    TG_SSTAMP(4);
#pragma unroll
    for (int v = 0; v < 6; v++) {
        const int kk = ty + 16 * v;
        if (kk < n) {
            const double li = W.Li[kk];
            const int cs = tg_colpack(n, kk) - kk;
#pragma unroll
            for (int u = v; u < 6; u++) { const int ii = tx + 16 * u; if (ii < n && kk < ii) Lg[cs + ii] = a[u][v] * li; }
        }
    }
    if (tid < n) Lg[tg_colpack(n, tid)] = W.Ld[tid];
    if (tid == 0) s.chol_fail[dst * s.C + c] = fail;
    TG_SSTAMP(5);
}

struct TgSpecArgs {
    TgState s;
    TgMala  m;
    int     smem_matrix;    // 1: the Nv x Nv matrix is staged in shared memory
    int     init;           // 1: factorise the LIVE slot of every chain (create / restore); 0: speculative step
};

// Speculative adaptation, one CTA per (chain, candidate).  cand = 0: "proposal rejected" candidate (x = vars_cur), 1: "accepted"
// (x = vars_prop).  Robbins-Monro update of (mu, Sigma) with the p3/p2 truncations (MALA.cpp:292-315,
// :135-176) from the live slot into candidate slot (live+1+cand)%3, then its Cholesky factor.  These CTAs ride
// in the grid of the propose kernel (the GPU is otherwise idle there: one CTA per chain), so the O(Nv^3)
// factorisation costs nothing on the step's critical path; the finish kernel selects the candidate that
// matches the Metropolis decision.  The acceptance only enters sigma (a scalar,
// applied as sqrt(sigma) at proposal time), so both candidates are exact.  chol layout: packed lower
// triangle when smem_matrix, else full row-major (factorised in place in HBM).
__device__ void tg_spec_body(const TgState &s, const TgMala &m, const int smem_matrix, const int init, const int c, const int cand,
                             double *tg_dyn_smem, const double *x_accept = nullptr)
{
    __shared__ double r1[256];
    __shared__ double s_scale;
    const int tid = threadIdx.x, nt = blockDim.x;
    const int Nv = s.Nv;
    double *dyn = smem_matrix ? tg_dyn_smem : nullptr;
    double *work = smem_matrix ? tg_dyn_smem + (size_t)Nv * (Nv | 1) : tg_dyn_smem;
    const int live = s.live[c];
    if (init) {
        if (cand != 0) return;
        if (Nv <= 96 && smem_matrix) tg_spec96(s, m, 0, c, live, live, 0.0, nullptr);
        else tg_factor_slot(s, m.eps2, live, c, dyn, work);
        return;
    }
    const long it = s.ctrl->i;
    if (!tg_learning(m, it)) return;
    const int slot = (live + 1 + cand) % 3;
    const double gamma = m.c0 / (1.0 + (double)it);
    // x of the candidate: the current point (rejected) or the proposed one (accepted; x_accept when the caller holds it)
    const double *xv = cand ? (x_accept ? x_accept : s.vars_prop + (size_t)c * Nv) : s.vars_cur + (size_t)c * Nv;
    if (Nv <= 96 && smem_matrix) { tg_spec96(s, m, 1, c, live, slot, gamma, xv); return; }
    TG_SSTAMP(0);
    const double *mu0 = s.mu + tg_slot_vec(s, live, c), *cov0 = s.cov + tg_slot_mat(s, live, c);
    double *mu = s.mu + tg_slot_vec(s, slot, c), *cov = s.cov + tg_slot_mat(s, slot, c);
    // mu <- p3(mu + gamma (x - mu))
    double part = 0;
    for (int v = tid; v < Nv; v += nt) { const double m = mu0[v] + gamma * (xv[v] - mu0[v]); mu[v] = m; part += m * m; }
    r1[tid] = part; __syncthreads();
    for (int h = 128; h > 0; h >>= 1) { if (tid < h) r1[tid] += r1[tid + h]; __syncthreads(); }
    if (tid == 0) { const double nrm = sqrt(r1[0]); s_scale = (nrm <= m.A1) ? 1.0 : m.A1 / nrm; }
    __syncthreads();
    if (s_scale != 1.0) for (int v = tid; v < Nv; v += nt) mu[v] = mu[v] * s_scale;
    __syncthreads();
    TG_SSTAMP(1);
    // Sigma <- p2(Sigma + gamma ((x-mu)(x-mu)^T - Sigma))
    part = 0;
    for (int e = tid; e < Nv * Nv; e += nt) {
        const int i = e / Nv, k = e - i * Nv;
        const double mat = (xv[i] - mu[i]) * (xv[k] - mu[k]);
        const double cn = cov0[e] + gamma * (mat - cov0[e]);
        cov[e] = cn; part += cn * cn;
    }
    __syncthreads();
    r1[tid] = part; __syncthreads();
    for (int h = 128; h > 0; h >>= 1) { if (tid < h) r1[tid] += r1[tid + h]; __syncthreads(); }
    if (tid == 0) { const double nrm = sqrt(r1[0]); s_scale = (nrm <= m.A1) ? 1.0 : m.A1 / nrm; }
    __syncthreads();
    if (s_scale != 1.0) for (int e = tid; e < Nv * Nv; e += nt) cov[e] = cov[e] * s_scale;
    __syncthreads();
    TG_SSTAMP(2);
    tg_factor_slot(s, m.eps2, slot, c, dyn, work);
    TG_SSTAMP(5);
}

// stand-alone launch: grid (C, 2) speculative step, or (C, 1) with init = 1 (create / restore)
__global__ void __launch_bounds__(256) tg_k_spec(TgSpecArgs a)
{
    extern __shared__ __align__(16) double tg_dyn_smem[];
    tg_spec_body(a.s, a.m, a.smem_matrix, a.init, blockIdx.x, blockIdx.y, tg_dyn_smem);
}
