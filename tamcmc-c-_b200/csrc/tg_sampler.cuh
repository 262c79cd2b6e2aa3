// tg_sampler.cuh -- device-resident adaptive Metropolis-Hastings + parallel tempering.
//
// Reference (under /root/reference/tamcmc/sources/MALA.cpp):
//   new_prop_values :335-353   x' = x + chol((Sigma+eps2 I)*sigma) z
//   update_position_MH :447-538, update_proposal :292-315, p1/p2/p3 :135-176,
//   parallel_tempering :381-445, execute :608-737 (gamma=c0/(1+i), learning zones, swap cadence)
// RNG: counter-based Philox4x32-10 keyed by (seed, purpose), counter (iteration, chain, index) --
// replaces rand()/Box-Muller with racy static state (random_JB.cpp, quirk Q10); trajectories are
// therefore only statistically comparable with the reference.
#pragma once
#include "tg_internal.cuh"
#include "tg_prologue.cuh"
#include "tg_adapt.cuh"
#include "tg_eval.cuh"

struct TgRecDev {          // device ring of per-step records (drained once per chunk)
    long    cap;           // steps the ring holds
    double *vars, *logL, *logPrior, *logPost, *Pmove;
    unsigned char *moved;
    int    *swap_attempt, *swap_chain0, *swap_done, *status;
    double *swap_Pswap;
};

// two standard normals from one Philox block
__device__ __forceinline__ void tg_normal_pair(uint4 r, double *z0, double *z1)
{
    const double u1 = tg_u01(r.x, r.y), u2 = tg_u01(r.z, r.w);
    const double rad = sqrt(-2.0 * log(u1));
    double s, c;
    sincospi(2.0 * u2, &s, &c);
    *z0 = rad * c; *z1 = rad * s;
}

struct TgProposeArgs {
    TgState   s;
    TgMala    m;
    TgProArgs pro;          // pro.params == s.params_prop
    int       packed;       // 1: chol is stored packed (tg_k_spec staged it in shared memory), 0: full row-major
    int       tri_smem;     // 1: the packed lower triangle of L (Nv(Nv+1)/2 doubles) fits the dynamic shared memory
    int       fused_spec;   // 1: grid (C, 3): rows 1 and 2 are the speculative adaptation CTAs (tg_adapt.cuh)
    int       smem_matrix;  // TgSpecArgs.smem_matrix of those CTAs
    int       role_base;    // role of grid row 0: 0 = proposal (main stream), 1 or 3 = side-stream grid (roles 1,2 adaptation, role 3 log-prior)
    int       parts;        // role 0: TG_PRO_MODEL, or TG_PRO_MODEL | TG_PRO_PRIOR when no role-3 CTA computes the prior
    int       nroles;       // grid rows this star uses (a launch that spans several stars has as many rows as its most demanding star)
    int       pad_;
};

// K1, role = blockIdx.y + role_base.  Role 0, one CTA per chain: draw z, form x' = x + sqrt(sigma)*L z with
// L = chol(Sigma+eps2 I) of the live slot (== chol((Sigma+eps2 I)*sigma) z, MALA.cpp:335-353, up to
// rounding), scatter into params, then run the model prologue + log-prior on the proposed parameter
// vector.  Roles 1 and 2 (fused_spec): the adapted proposal (Robbins-Monro + Cholesky) this chain will
// use next step if the proposal is rejected / accepted; role 2 recomputes x' itself (same arithmetic,
// same bits) so that the three CTAs of a chain are independent -- the host launches roles 1 and 2 as their
// own grid on a side stream, concurrent with role 0 and the fused kernel.
// grid.x = every chain of the launch; cmap[blockIdx.x] = (star, chain of that star), argv[star] = its argument record.
__global__ void __launch_bounds__(256) tg_k_propose(const TgProposeArgs *__restrict__ argv, const int2 *__restrict__ cmap)
{
    extern __shared__ __align__(16) double tg_dyn_smem[];
    __shared__ double s_z[512];
    __shared__ __align__(16) TgProposeArgs s_args;
    const int2 cm = __ldg(cmap + blockIdx.x);
    tg_load_args(&s_args, argv + cm.x);
    const TgProposeArgs &a = s_args;
    if ((int)blockIdx.y >= a.nroles) return;
    const TgState &s = a.s;
    const int c = cm.y, role = blockIdx.y + a.role_base, tid = threadIdx.x, nt = blockDim.x;
    const int Nv = s.Nv, Np = s.Np;
    const uint32_t cid = (uint32_t)(c + a.m.chain0);              // chain index of the Philox counters
    tg_pdl_wait();                                                  // role 0 may be resident before the previous finish kernel is done
    if (role == 0) tg_pdl_trigger();
#ifdef TG_PROFILE
    if (tid == 0 && c == 0 && blockIdx.y == 0) { if (role == 0) { tg_dbg_tl[7] = tg_dbg_tl[4]; tg_dbg_tl[0] = globaltimer_ns(); } else tg_dbg_tl[5] = globaltimer_ns(); }
#endif                                // the side-stream roles complete in full before the finish kernel starts
    const long it = s.ctrl->i;
    if (role == 1 || role == 2) {
        if (!tg_learning(a.m, it)) return;
        if (role == 1) { tg_spec_body(s, a.m, a.smem_matrix, 0, c, 0, tg_dyn_smem); return; }
    }
    const double *L = s.chol + tg_slot_mat(s, s.live[c], c);
    // role 0, fast path: the proposed parameter vector is assembled in shared memory and handed to the prologue there (no
    // round trip through global memory between the proposal and the prologue); its ingredients are requested now
    double pc0 = 0.0; int ir0 = 0;
    if (role == 0 && a.tri_smem) { if (tid < Np) pc0 = s.params_cur[(size_t)c * Np + tid]; if (tid < Nv) ir0 = s.index_to_relax[tid]; }
    TG_STAMP(0);
    const double sq = sqrt(s.sigma[c]);
    const uint2 key = make_uint2((uint32_t)a.m.seed ^ TG_TAG_PROP, (uint32_t)(a.m.seed >> 32));
    if (a.tri_smem) {
        // ---- fast path: z and the packed lower triangle of L in shared memory, one thread per row ----
        double *tri = tg_dyn_smem;                                  // packed by columns (tg_colpack)
        double *zz = tri + (size_t)Nv * (Nv + 1) / 2;
        bool async_copy = false;
        if (a.packed) {
            const int ntri = Nv * (Nv + 1) / 2;
            if ((((size_t)L) & 15) == 0) {
                // all of the triangle in flight at once (16-byte asynchronous copies through L2, no registers), under the normals
                const unsigned sb = (unsigned)__cvta_generic_to_shared(tri);
                for (int e = 2 * tid; e + 1 < ntri; e += 2 * nt)
                    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(sb + 8u * (unsigned)e), "l"(L + e) : "memory");
                asm volatile("cp.async.commit_group;" ::: "memory");
                if ((ntri & 1) && tid == 0) tri[ntri - 1] = __ldcg(L + ntri - 1);
                async_copy = true;
            } else {
#pragma unroll 8
                for (int e = tid; e < ntri; e += nt) tri[e] = __ldcg(L + e);
            }
        } else {
            for (int e = tid; e < Nv * Nv; e += nt) { const int i = e / Nv, k = e - i * Nv; if (k <= i) tri[tg_colpack(Nv, k) + i - k] = __ldcg(L + e); }
        }
        for (int v = tid; v < Nv; v += nt) {
            double z0, z1;
            tg_normal_pair(tg_philox(make_uint4((uint32_t)it, (uint32_t)(it >> 32), cid, (uint32_t)(v >> 1)), key), &z0, &z1);
            zz[v] = (v & 1) ? z1 : z0;
        }
        if (async_copy) asm volatile("cp.async.wait_group 0;" ::: "memory");
        __syncthreads();
        TG_STAMP(1);
        for (int i = tid; i < Nv; i += nt) {
            const double *el = tri + i;                             // L(i,k) = tri[colpack(k) - k + i]: conflict-free across rows
            double a0 = 0.0, a1 = 0.0;                              // fixed order: even / odd columns, then a0 + a1
            int k = 0;
            for (; k + 1 <= i; k += 2) {
                a0 = fma(el[0], zz[k], a0); a1 = fma(el[Nv - k - 1], zz[k + 1], a1);
                el += 2 * (Nv - k) - 3;                             // colpack(k+2)-(k+2) - (colpack(k)-k)
            }
            if (k <= i) a0 = fma(el[0], zz[k], a0);
            const double xp = s.vars_cur[(size_t)c * Nv + i] + sq * (a0 + a1);
            s_z[i] = xp;
            if (role == 0) s.vars_prop[(size_t)c * Nv + i] = xp;
        }
        if (role == 2) {                                            // fused_spec implies tri_smem and Nv <= 512
            __syncthreads();
            tg_spec_body(s, a.m, a.smem_matrix, 0, c, 1, tg_dyn_smem, s_z);
            return;
        }
        if (role == 3) {                                            // log-prior of the proposed point, off the path to the fused kernel
            __syncthreads();                                        // the triangle is dead: its memory becomes the parameter vector
            for (int i = tid; i < Np; i += nt) tg_dyn_smem[i] = s.params_cur[(size_t)c * Np + i];
            __syncthreads();
            for (int v = tid; v < Nv; v += nt) tg_dyn_smem[s.index_to_relax[v]] = s_z[v];
            tg_prologue_body(a.pro, c, tg_dyn_smem, TG_PRO_PRIOR, true);
            return;
        }
        // role 0: params_prop = params_cur with the free variables replaced (Model_def::update_params_with_vars), in shared memory
        TG_STAMP(2);
        __syncthreads();                                            // the triangle is dead
        for (int i = tid; i < Np; i += nt) tg_dyn_smem[i] = (i == tid) ? pc0 : s.params_cur[(size_t)c * Np + i];
        __syncthreads();
        for (int v = tid; v < Nv; v += nt) tg_dyn_smem[(v == tid) ? ir0 : s.index_to_relax[v]] = s_z[v];
        __syncthreads();
        for (int i = tid; i < Np; i += nt) s.params_prop[(size_t)c * Np + i] = tg_dyn_smem[i];
        TG_STAMP(3);
        tg_prologue_body(a.pro, c, tg_dyn_smem, a.parts, true);
#ifdef TG_PROFILE
        __syncthreads(); if (tid == 0 && c == 0) tg_dbg_tl[1] = globaltimer_ns();
#endif
        return;
    } else
    // general path (very large Nv): standard normals 512 per pass, rows from HBM
    for (int v0 = 0; v0 < Nv; v0 += 512) {
        __syncthreads();
        for (int v = v0 + tid; v < Nv && v < v0 + 512; v += nt) {
            double z0, z1;
            tg_normal_pair(tg_philox(make_uint4((uint32_t)it, (uint32_t)(it >> 32), cid, (uint32_t)(v >> 1)), key), &z0, &z1);
            s_z[v - v0] = (v & 1) ? z1 : z0;
        }
        __syncthreads();
        TG_STAMP(1);
        // partial products of rows with this block of z (lower triangle => k <= i): a warp takes four rows
        // at a time (their loads are in flight together), lanes stride the row (coalesced), fixed shuffle tree
        const int lane = tid & 31, nw = nt >> 5;
        for (int i0 = (tid >> 5); i0 < Nv; i0 += 4 * nw) {
            double acc[4] = {0.0, 0.0, 0.0, 0.0};
            int kend[4];
#pragma unroll
            for (int u = 0; u < 4; u++) { const int i = i0 + u * nw; kend[u] = (i < Nv) ? min(i, v0 + 511) : -1; }
            const int kmax = max(max(kend[0], kend[1]), max(kend[2], kend[3]));
            for (int k = v0 + lane; k <= kmax; k += 32) {
                const double z = s_z[k - v0];
#pragma unroll
                for (int u = 0; u < 4; u++) if (k <= kend[u]) acc[u] = fma(__ldcg(L + (size_t)(i0 + u * nw) * Nv + k), z, acc[u]);
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
                for (int u = 0; u < 4; u++) acc[u] += __shfl_down_sync(0xffffffffu, acc[u], o);
            }
            if (lane == 0) {
#pragma unroll
                for (int u = 0; u < 4; u++) {
                    const int i = i0 + u * nw;
                    if (i < Nv && i >= v0) {
                        const double prev = (v0 == 0) ? s.vars_cur[(size_t)c * Nv + i] : s.vars_prop[(size_t)c * Nv + i];
                        s.vars_prop[(size_t)c * Nv + i] = prev + sq * acc[u];
                    }
                }
            }
        }
    }
    TG_STAMP(2);
    // params_prop = params_cur with the free variables replaced (Model_def::update_params_with_vars)
    for (int i = tid; i < Np; i += nt) s.params_prop[(size_t)c * Np + i] = s.params_cur[(size_t)c * Np + i];
    __syncthreads();
    for (int v = tid; v < Nv; v += nt) s.params_prop[(size_t)c * Np + s.index_to_relax[v]] = s.vars_prop[(size_t)c * Nv + v];
    __syncthreads();
    TG_STAMP(3);
    tg_prologue_body(a.pro, c, tg_dyn_smem, a.parts);
}

struct TgFinishArgs {
    TgDesc    d;
    TgState   s;
    TgMala    m;
    TgTables  t;            // tables of the PROPOSED points (status, logprior)
    const double2 *partials; int ntiles;
    double2  *sums;         // [C]
    int       presummed;    // 1: sums already hold the (all-reduced) totals
    TgRecDev  rec;
    int       record;
};

__device__ __forceinline__ void tg_write_record(const TgFinishArgs &a, long slot, int c, int st, int tid, int nt)
{
    const TgState &s = a.s;
    if (slot < 0 || slot >= a.rec.cap) return;
    const size_t rc = (size_t)slot * s.C + c;
    if (a.rec.vars) for (int v = tid; v < s.Nv; v += nt) a.rec.vars[rc * s.Nv + v] = s.vars_cur[(size_t)c * s.Nv + v];
    if (tid == 0) {
        a.rec.logL[rc] = s.logL_cur[c]; a.rec.logPrior[rc] = s.logPrior_cur[c]; a.rec.logPost[rc] = s.logPost_cur[c];
        a.rec.Pmove[rc] = s.Pmove[c]; a.rec.moved[rc] = (unsigned char)s.moved[c];
        if (st >= 0) a.rec.status[rc] = st;
    }
}

// K3: one CTA per chain: likelihood reduction -> MH accept -> Robbins-Monro update -> record; the last CTA of an
// ensemble to finish performs that ensemble's tempering swap (ensembles in parallel), and the last ensemble of the star
// advances the star's iteration counter.  grid.x = every chain of the launch (cmap / argv as in tg_k_propose).
__global__ void __launch_bounds__(256) tg_k_finish(const TgFinishArgs *__restrict__ argv, const int2 *__restrict__ cmap)
{
    __shared__ double2 s_part[8];
    __shared__ int s_accept, s_last, s_swap;
    __shared__ double s_sw[3];
    __shared__ __align__(16) TgFinishArgs s_args;
    const int2 cm = __ldg(cmap + blockIdx.x);
    tg_load_args(&s_args, argv + cm.x);
    const TgFinishArgs &a = s_args;
    const TgState &s = a.s;
    const int c = cm.y, tid = threadIdx.x, nt = blockDim.x;
    const int Nv = s.Nv, Np = s.Np;
    // Every load that does not depend on the decision is issued up front: the kernel is a chain of
    // global-memory round trips, so they are kept to the unavoidable ones.
    // Under programmatic dependent launch this kernel becomes resident while the fused kernel drains: the proposal kernel and
    // the adaptation candidates are complete by then (the fused kernel triggers after its own wait; the candidates are a full
    // dependency), so everything but the partial sums is loaded before the wait.
#ifdef TG_PROFILE
    if (tid == 0 && blockIdx.x == 0) tg_dbg_tl[2] = globaltimer_ns();
#endif
    const long it = s.ctrl->i;
    const bool small = (Nv <= 256 && Np <= 256);   // one element of the proposed point per thread, held in registers
    double vprop = 0.0, pprop = 0.0;
    if (small) { if (tid < Nv) vprop = s.vars_prop[(size_t)c * Nv + tid]; if (tid < Np) pprop = s.params_prop[(size_t)c * Np + tid]; }
    int st = 0, live = 0, fail0 = 0, fail1 = 0;
    double T = 1.0, logPr = 0.0, logPo_cur = 0.0, sig = 0.0;
    if (tid == 0) {
        st = a.t.status[c] | s.sticky_status[c]; T = s.Tcoefs[c]; logPo_cur = s.logPost_cur[c];
        live = s.live[c]; sig = s.sigma[c];
    }
    const long slot = it - s.ctrl->i0;
    tg_pdl_wait();
    tg_pdl_trigger();
#ifdef TG_PROFILE
    if (tid == 0 && blockIdx.x == 0) tg_dbg_tl[3] = globaltimer_ns();
#endif
    if (tid == 0) {                                                 // products of the side-stream CTAs: only after the wait
        st |= a.t.pstatus[c]; logPr = a.t.logprior[c];
        fail0 = s.chol_fail[((live + 1) % 3) * s.C + c]; fail1 = s.chol_fail[((live + 2) % 3) * s.C + c];
    }
    double a1 = 0, a2 = 0;
    if (!a.presummed) for (int t = tid; t < a.ntiles; t += nt) { const double2 v = a.partials[(size_t)c * a.ntiles + t]; a1 += v.x; a2 += v.y; }

    // 1. logL of the proposed point
    const double2 sum = a.presummed ? a.sums[c] : tg_block_sum2(a1, a2, s_part);
    // 2. Metropolis-Hastings decision (MALA.cpp:475-535)
    if (tid == 0) {
        const double logL = tg_logl_from_sums(a.d, sum, T, st);
        const double logPo = logL + logPr;
        double r;
        if (logL != logL) { r = 0.0; st |= TG_ST_NAN_LOGL; }
        else if (logPo == -CUDART_INF) r = 0.0;
        else { r = exp(logPo - logPo_cur); r = (r < 1.0) ? r : 1.0; if (r != r) r = 0.0; }   // NaN ratio: reference is fatal (:504-506); reject here
        const uint2 key = make_uint2((uint32_t)a.m.seed ^ TG_TAG_MH, (uint32_t)(a.m.seed >> 32));
        const uint4 rn = tg_philox(make_uint4((uint32_t)it, (uint32_t)(it >> 32), (uint32_t)(c + a.m.chain0), 0u), key);
        const double u = tg_u01(rn.x, rn.y);
        const int acc = (u <= r);
        s_accept = acc;
        if (acc) { s.logL_cur[c] = logL; s.logPrior_cur[c] = logPr; s.logPost_cur[c] = logPo; }
        s.moved[c] = acc; s.Pmove[c] = r;
        if (a.record && slot >= 0 && slot < a.rec.cap) a.rec.status[(size_t)slot * s.C + c] = st;
        // 3. Robbins-Monro update (MALA.cpp:292-315): (mu, Sigma, chol) were prepared for both outcomes by the
        //    speculative CTAs; select the candidate slot of this outcome and update sigma with the p1 truncation.
        if (tg_learning(a.m, it)) {
            const double gamma = a.m.c0 / (1.0 + (double)it);
            s.live[c] = (live + 1 + acc) % 3;
            const int sticky = s.sticky_status[c];
            s.sticky_status[c] = (acc ? fail1 : fail0) ? (sticky | TG_ST_CHOL_FAIL) : (sticky & ~TG_ST_CHOL_FAIL);
            double sg = sig + gamma * (r - a.m.target);
            if (sg < a.m.eps1) sg = a.m.eps1;
            if (sg > a.m.A1) sg = a.m.A1;
            s.sigma[c] = sg;
        }
    }
    __syncthreads();
    if (s_accept) {
        if (small) { if (tid < Nv) s.vars_cur[(size_t)c * Nv + tid] = vprop; if (tid < Np) s.params_cur[(size_t)c * Np + tid] = pprop; }
        else {
            for (int v = tid; v < Nv; v += nt) s.vars_cur[(size_t)c * Nv + v] = s.vars_prop[(size_t)c * Nv + v];
            for (int i = tid; i < Np; i += nt) s.params_cur[(size_t)c * Np + i] = s.params_prop[(size_t)c * Np + i];
        }
    }
    __syncthreads();
    if (a.record) tg_write_record(a, slot, c, -1, tid, nt);
    // running sums of the proposal law after this step's update (Outputs::update_buffer_restore, outputs.cpp:1758-1818; the
    // reference sums sigma, mu and covarmat of every step and divides when it writes the restore files)
    if (s.acc_cov) {
        const int lv = s.live[c];                                    // written by thread 0 before the barriers above
        const size_t nn = (size_t)Nv * Nv;
        const double *cv = s.cov + tg_slot_mat(s, lv, c);
        double *ac = s.acc_cov + (size_t)c * nn;
        for (size_t e = tid; e < nn; e += nt) ac[e] += cv[e];
        const double *m = s.mu + tg_slot_vec(s, lv, c);
        for (int v = tid; v < Nv; v += nt) s.acc_mu[(size_t)c * Nv + v] += m[v];
        if (tid == 0) s.acc_sigma[c] += s.sigma[c];
    }

    // 4. last CTA of the ensemble: its tempering swap (MALA.cpp:381-445, :676-692)
    const int e = c / s.Ce;
    __threadfence();
    __syncthreads();
    if (tid == 0) { const unsigned t = atomicAdd(s.ens_ticket + e, 1u); s_last = (t == (unsigned)(s.Ce - 1)); }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    const int attempt = (a.m.dN_mixing > 0) && (it % a.m.dN_mixing == 0) && (it != 0) && (s.Ce >= 2);
    const bool rec_ok = a.record && slot >= 0 && slot < a.rec.cap;
    const size_t re = (size_t)(rec_ok ? slot : 0) * s.E + e;
    if (!attempt) {
        if (rec_ok && tid == 0) { a.rec.swap_attempt[re] = 0; a.rec.swap_chain0[re] = -1; a.rec.swap_Pswap[re] = 0.0; a.rec.swap_done[re] = 0; }
    } else {
        // The pair is a function of the counter alone, so every thread knows it and fetches its elements of the two rows in the
        // same round trip in which thread 0 fetches the likelihoods (values written by other CTAs: read through L2).  ONE thread
        // takes the decision and publishes it through shared memory: nobody re-reads a likelihood that thread 0 may already have
        // overwritten.
        const uint2 key = make_uint2((uint32_t)a.m.seed ^ TG_TAG_PT, (uint32_t)(a.m.seed >> 32));
        const uint4 rn = tg_philox(make_uint4((uint32_t)it, (uint32_t)(it >> 32), (uint32_t)(e + a.m.ens0), 0u), key);
        const int A = (int)(rn.z % (uint32_t)(s.Ce - 1));
        const int gA = e * s.Ce + A, gB = gA + 1;
        double va = 0, vb = 0, pa = 0, pb = 0;
        if (small) {
            if (tid < Nv) { va = __ldcg(s.vars_cur + (size_t)gA * Nv + tid); vb = __ldcg(s.vars_cur + (size_t)gB * Nv + tid); }
            if (tid < Np) { pa = __ldcg(s.params_cur + (size_t)gA * Np + tid); pb = __ldcg(s.params_cur + (size_t)gB * Np + tid); }
        }
        if (tid == 0) {
            const double TA = s.Tcoefs[gA], TB = s.Tcoefs[gB];
            const double LA = __ldcg(s.logL_cur + gA), LB = __ldcg(s.logL_cur + gB);
            const double LA_TB = LA * TA / TB, LB_TA = LB * TB / TA;
            double rT = exp(LA_TB + LB_TA - LA - LB); rT = (rT < 1.0) ? rT : 1.0;
            const double u = tg_u01(rn.x, rn.y);
            s_swap = (u <= rT); s_sw[0] = rT; s_sw[1] = LA_TB; s_sw[2] = LB_TA;
        }
        __syncthreads();
        const int doswap = s_swap;
        if (doswap) {
            if (tid == 0) {
                const double LA_TB = s_sw[1], LB_TA = s_sw[2];
                const double prA = __ldcg(s.logPrior_cur + gA), prB = __ldcg(s.logPrior_cur + gB);
                s.logL_cur[gA] = LB_TA; s.logPrior_cur[gA] = prB; s.logPost_cur[gA] = LB_TA + prB;
                s.logL_cur[gB] = LA_TB; s.logPrior_cur[gB] = prA;
                // Q8: the reference adds logPrior[ind_A] AFTER it was overwritten with B's prior (MALA.cpp:417,:428)
                s.logPost_cur[gB] = LA_TB + ((a.m.flags & TG_F_FIX_SWAP_BOOKKEEPING) ? prA : prB);
                const int mv = __ldcg(s.moved + gA); s.moved[gA] = __ldcg(s.moved + gB); s.moved[gB] = mv;
                const double pm = __ldcg(s.Pmove + gA); s.Pmove[gA] = __ldcg(s.Pmove + gB); s.Pmove[gB] = pm;
            }
            if (small) {
                if (tid < Nv) { s.vars_cur[(size_t)gA * Nv + tid] = vb; s.vars_cur[(size_t)gB * Nv + tid] = va; }
                if (tid < Np) { s.params_cur[(size_t)gA * Np + tid] = pb; s.params_cur[(size_t)gB * Np + tid] = pa; }
            } else {
                for (int v = tid; v < Nv; v += nt) { const double t = __ldcg(s.vars_cur + (size_t)gA * Nv + v); s.vars_cur[(size_t)gA * Nv + v] = __ldcg(s.vars_cur + (size_t)gB * Nv + v); s.vars_cur[(size_t)gB * Nv + v] = t; }
                for (int i = tid; i < Np; i += nt) { const double t = __ldcg(s.params_cur + (size_t)gA * Np + i); s.params_cur[(size_t)gA * Np + i] = __ldcg(s.params_cur + (size_t)gB * Np + i); s.params_cur[(size_t)gB * Np + i] = t; }
            }
            // the records of the two chains are rewritten by the threads that hold the new values (thread v wrote element v, thread 0 the scalars)
            if (a.record) { tg_write_record(a, slot, gA, -1, tid, nt); tg_write_record(a, slot, gB, -1, tid, nt); }
        }
        if (rec_ok && tid == 0) {
            a.rec.swap_attempt[re] = 1; a.rec.swap_chain0[re] = A;
            a.rec.swap_Pswap[re] = doswap ? s_sw[0] : 0.0; a.rec.swap_done[re] = doswap;
        }
    }
    // the variables of the ensemble's chains as the step leaves them (after the swap): running sums for vars_mean
    if (s.acc_vars) {
        __syncthreads();                                             // the swapped rows are written
        for (int q = tid; q < s.Ce * Nv; q += nt) { const size_t k = (size_t)e * s.Ce * Nv + q; s.acc_vars[k] += __ldcg(s.vars_cur + k); }
    }
    // 5. the last ensemble of the star advances the iteration
    if (tid == 0) {
        s.ens_ticket[e] = 0u;
        const unsigned t2 = atomicAdd(&s.ctrl->ticket, 1u);
        if (t2 == (unsigned)(s.E - 1)) { s.ctrl->ticket = 0u; s.ctrl->i = it + 1; if (s.acc_count) *s.acc_count += 1; }
    }
#ifdef TG_PROFILE
    if (tid == 0) tg_dbg_tl[4] = globaltimer_ns();
#endif
}

// ---- bin-sharded variant: exchange of the per-chain partial sums over NVLink peer memory ---------------------------------
// Every rank owns one buffer (cudaMalloc, opened by its peers through CUDA IPC): data[2][nranks][2C] doubles and
// stamp[2][nranks] 64-bit words.  Exchange number `seq` (the same on every rank: all ranks run the same sequence of steps):
// rank r writes its 2C sums into slot [seq & 1][r] of EVERY rank's buffer (plain stores through the peer pointers), fences, and
// releases stamp [seq & 1][r] = seq there; it then waits until its own nranks stamps of that parity read seq (acquire) and adds
// the nranks contributions in rank order -- the same order on every rank, so the totals are bit-identical everywhere, which the
// replicated accept / swap decisions need.  A rank cannot lap a peer by two exchanges (it needs the peer's stamp of the exchange
// in between), so two parities suffice.  One CTA; 2C doubles and nranks stamps cross NVLink per rank and step.
#define TG_XCHG_MAXR 8
struct TgXchg {
    int nranks, rank;
    double *peer_data[TG_XCHG_MAXR];               // [2][nranks][2C] of every rank (own buffer included)
    unsigned long long *peer_stamp[TG_XCHG_MAXR];  // [2][nranks]
    unsigned long long *seq;                       // this rank's exchange counter (device memory: graph replays advance it)
    int *err;                                      // set to 1 if a peer did not answer within ~2 s (the run is void then; no hang)
};
__global__ void __launch_bounds__(256) tg_k_xchg(TgXchg x, double2 *sums, int C)
{
    const int t = threadIdx.x, n = 2 * C, R = x.nranks;
    const unsigned long long seq = *x.seq + 1ull;
    const int par = (int)(seq & 1ull);
    const double *mine = (const double *)sums;
    for (int p = 0; p < R; p++) {
        double *dst = x.peer_data[p] + ((size_t)par * R + x.rank) * n;
        for (int e = t; e < n; e += blockDim.x) dst[e] = mine[e];
    }
    __threadfence_system();
    __syncthreads();
    if (t < R) {
        unsigned long long *st = x.peer_stamp[t] + (size_t)par * R + x.rank;
        asm volatile("st.release.sys.global.u64 [%0], %1;" :: "l"(st), "l"(seq) : "memory");
        const unsigned long long *my = x.peer_stamp[x.rank] + (size_t)par * R + t;
        unsigned long long v = 0ull, t0;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
        for (;;) {
            asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(my) : "memory");
            if (v == seq) break;
            unsigned long long t1;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
            if (t1 - t0 > 2000000000ull) { *x.err = 1; break; }
        }
    }
    __syncthreads();
    const double *all = x.peer_data[x.rank] + (size_t)par * R * n;
    for (int e = t; e < n; e += blockDim.x) {
        double s = 0.0;
        for (int r = 0; r < R; r++) s += __ldcv(all + (size_t)r * n + e);          // written by peers: not through a stale cache line
        ((double *)sums)[e] = s;
    }
    __syncthreads();
    if (t == 0) *x.seq = seq;
}

// Live (mu, Sigma) of every chain -> contiguous [C][Nv], [C][Nv][Nv] (restore files / tg_get_state)
__global__ void tg_k_gather_live(TgState s, double *mu_out, double *cov_out)
{
    const int c = blockIdx.x, live = s.live[c];
    const size_t nv = s.Nv, nn = nv * nv;
    if (mu_out) for (size_t e = threadIdx.x; e < nv; e += blockDim.x) mu_out[c * nv + e] = s.mu[tg_slot_vec(s, live, c) + e];
    if (cov_out) for (size_t e = threadIdx.x; e < nn; e += blockDim.x) cov_out[c * nn + e] = s.cov[tg_slot_mat(s, live, c) + e];
}

// Installs the evaluation of params_cur as the current state (Model_def ctor, model_def.cpp:139-143)
__global__ void tg_k_install_current(TgDesc d, TgState s, TgTables t, const double2 *sums)
{
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= s.C) return;
    const double L = tg_logl_from_sums(d, sums[c], s.Tcoefs[c], t.status[c] | t.pstatus[c]);
    s.logL_cur[c] = L; s.logPrior_cur[c] = t.logprior[c]; s.logPost_cur[c] = L + t.logprior[c];
    s.Pmove[c] = 0.0; s.moved[c] = 0;
}

// FP64 roofline denominator: 8 independent DFMA chains per thread, no memory traffic.
__global__ void __launch_bounds__(256) tg_k_dfma_peak(double *out, int iters, double a, double b)
{
    double v0 = threadIdx.x, v1 = v0 + 1, v2 = v0 + 2, v3 = v0 + 3, v4 = v0 + 4, v5 = v0 + 5, v6 = v0 + 6, v7 = v0 + 7;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int u = 0; u < 8; u++) {
            v0 = fma(v0, a, b); v1 = fma(v1, a, b); v2 = fma(v2, a, b); v3 = fma(v3, a, b);
            v4 = fma(v4, a, b); v5 = fma(v5, a, b); v6 = fma(v6, a, b); v7 = fma(v7, a, b);
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = v0 + v1 + v2 + v3 + v4 + v5 + v6 + v7;
}
