// tamcmc_gpu.cu -- C-ABI (include/tamcmc_gpu.h) over the sm_100a kernels.  Host side only
// allocates, uploads and launches; there is no CPU implementation of any numerical step here.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include <limits>
#include <cmath>
#include <algorithm>
#include "tg_internal.cuh"
#include "tg_prologue.cuh"
#include "tg_eval.cuh"
#include "tg_sampler.cuh"

static thread_local std::string g_err;
static int fail(int code, const std::string &msg) { g_err = msg; return code; }
#define CK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail(TG_E_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_)); } while (0)

struct EvalBuf {                 // tables + results for a batch of B parameter vectors
    int       cap = 0;
    TgTables  t{};
    // inputs and results of a batch are contiguous, laid out for the batch size B of the call:
    // in = params[B][Np] | T[B];  out = logL[B] | logPrior[B] | logPost[B] | status[B] (int) -- one copy each way
    double   *in = nullptr, *out = nullptr;
    double   *params = nullptr, *T = nullptr;
    double2  *partials = nullptr, *sums = nullptr;
    double   *logL = nullptr, *logPrior = nullptr, *logPost = nullptr;
    int      *status = nullptr;
    int       B = 0;               // batch size the pointers above are laid out for
    double   *pin_in = nullptr, *pin_out = nullptr;    // pinned host staging of the same layout
    void layout(int b, int Np) {
        B = b; params = in; T = in + (size_t)b * Np;
        logL = out; logPrior = out + b; logPost = out + 2 * (size_t)b; status = (int *)(out + 3 * (size_t)b);
    }
};

struct tg_handle {
    int        device = 0, n_sm = 0;
    cudaStream_t stream = nullptr;
    cudaStream_t side = nullptr;           // the adaptation candidates of a step run here, beside the proposal and the fused kernel
    cudaEvent_t  ev_fork = nullptr, ev_join = nullptr;
    bool         side_pending = false;     // a join is owed before the next finish kernel
    bool         pdl = true;                           // programmatic dependent launch between the step kernels (environment TG_PDL=0 turns it off)
    bool         no_split = false, no_side = false;   // developer switches (environment TG_NO_SPLIT_PRIOR / TG_NO_SIDE_STREAM)
    TgDesc     d{};
    TgMala     m{};
    TgState    s{};
    TgRecDev   rec{};
    double    *x = nullptr, *y = nullptr, *lnx = nullptr, *sig = nullptr;
    double    *mom = nullptr; int mom_sub = 4, mom_run = 16; int *mom_fail = nullptr; long n_mom_tiles = 0; bool mom_probed = false;
    int        hybrid_pct = 0, hybrid_group = 2;
    int       *mom_note = nullptr;                      // pinned, mapped: the kernel stores 1 when it flags a tile for the first time          // Chebyshev moments of the data over aligned tiles (tg_eval.cuh), NULL = off
    double    *pri = nullptr; int *sw = nullptr;
    int       *index_to_relax = nullptr; double *Tcoefs = nullptr;
    EvalBuf    ev;               // tg_eval batches
    EvalBuf    sp;               // sampler's proposed points (cap == C)
    int        nitems = 0, max_items = 0, light_sub = 8, heavy_sub = 2, chains_per_block = 0, heavy_group = 2, use_graph = 1, steps_per_graph = 16;
    int        smem_matrix = 0, tri_smem = 0; size_t propose_smem = 0, spec_smem = 0; int fused_spec = 0;
    double    *stage_mu = nullptr, *stage_cov = nullptr;   // tg_get_state staging ([C][Nv], [C][Nv][Nv])
    int        list_cap = 4, max_cpb = 1;
    struct UnitTable { int4 *dev = nullptr; int n = 0, B = -1, cap = 0; };
    UnitTable  u_sp, u_ev;                 // work units of the fused kernel in launch order: sampler (B = C), tg_eval batches
    std::vector<int4> items_host;          // bin partition: (first bin, sub-tiles, weight = windows crossing, heavy?)
    bool       no_order = false;
    cudaGraphExec_t graph = nullptr;
    cudaGraphExec_t ev_graph = nullptr; int ev_graph_B = 0;   // tg_eval: copy in, prologue, fused kernel, reduce+finalize, copy out
    // eval-kernel timing
    bool       timing = false;
    std::vector<cudaEvent_t> ev_pool; size_t ev_used = 0;
    long       n_eval_launch = 0, n_launch_total = 0; double ms_eval = 0;
    std::vector<void *> allocs;
    cudaEvent_t watch[2] = {nullptr, nullptr};
};

template <typename T> static int dalloc(tg_handle *h, T **p, size_t n)
{
    void *q = nullptr;
    CK(cudaMalloc(&q, std::max<size_t>(n, 1) * sizeof(T)));
    h->allocs.push_back(q);
    *p = (T *)q;
    return TG_OK;
}
#define DA(ptr, n) do { int r_ = dalloc(h, &(ptr), (n)); if (r_) return r_; } while (0)

extern "C" const char *tg_last_error(void) { return g_err.c_str(); }
extern "C" int tg_version(void) { return 100; }
#ifdef TG_PROFILE
extern "C" int tg_dbg_ctas(long long *out) { return cudaMemcpyFromSymbol(out, tg_dbg_cta, sizeof(long long) * TG_DBG_CTAS * 6) == cudaSuccess ? 0 : -2; }
extern "C" int tg_dbg_timeline(long long *out) { return cudaMemcpyFromSymbol(out, tg_dbg_tl, sizeof(long long) * 8) == cudaSuccess ? 0 : -2; }
extern "C" int tg_dbg_specs(long long *out) { return cudaMemcpyFromSymbol(out, tg_dbg_spec, sizeof(long long) * 16) == cudaSuccess ? 0 : -2; }
extern "C" int tg_dbg_clocks(long long *out) { return cudaMemcpyFromSymbol(out, tg_dbg_clk, sizeof(long long) * 32) == cudaSuccess ? 0 : -2; }
#endif

static void drop_graphs(tg_handle *h)
{
    if (h->graph) { cudaGraphExecDestroy(h->graph); h->graph = nullptr; }
    if (h->ev_graph) { cudaGraphExecDestroy(h->ev_graph); h->ev_graph = nullptr; }
}

static int alloc_evalbuf(tg_handle *h, EvalBuf &b, int B)
{
    if (B <= b.cap) return TG_OK;
    const size_t nc = (size_t)B * h->d.ncomp_max;
    b.cap = B;
    if (h->ev_graph) { cudaGraphExecDestroy(h->ev_graph); h->ev_graph = nullptr; }     // the buffers move
    DA(b.t.nu, nc); DA(b.t.h, nc); DA(b.t.q, nc); DA(b.t.ifc, nc); DA(b.t.t2sq, nc); DA(b.t.lo, nc); DA(b.t.hi, nc);
    DA(b.t.ncomp, B); DA(b.t.wlo, B); DA(b.t.whi, B); DA(b.t.noise, (size_t)B * TG_NOISE_STRIDE); DA(b.t.status, B); DA(b.t.pstatus, B); DA(b.t.work, B); DA(b.t.logprior, B);
    DA(b.in, (size_t)B * (h->d.Np + 1)); DA(b.out, (size_t)B * 4);
    DA(b.partials, (size_t)B * h->max_items); DA(b.sums, B);
    if (b.pin_in) cudaFreeHost(b.pin_in);
    if (b.pin_out) cudaFreeHost(b.pin_out);
    CK(cudaHostAlloc((void **)&b.pin_in, (size_t)B * (h->d.Np + 1) * 8, cudaHostAllocDefault));
    CK(cudaHostAlloc((void **)&b.pin_out, (size_t)B * 4 * 8, cudaHostAllocDefault));
    b.layout(B, h->d.Np);
    return TG_OK;
}

static int ncomp_for(const tg_problem *p, int *ncomp, int *nmodes)
{
    const int mc = p->model_case;
    const int *pl = p->plength;
    if (mc == 0 || mc == 1) { *ncomp = 0; *nmodes = 0; return TG_OK; }
    if (mc == 11 || mc == 14) { *nmodes = pl[2] + pl[3] + pl[4] + pl[5]; *ncomp = pl[2] + 3 * pl[3] + 5 * pl[4] + 7 * pl[5]; return TG_OK; }
    const int L1 = std::min(pl[1], 3) + 1;
    *nmodes = pl[0] * L1; *ncomp = pl[0] * L1 * L1;
    return TG_OK;
}

// Moment buffer of the background-only path (tg_eval.cuh, tg_moment_tiles): the interpolation matrices, then for every
// aligned tile of mt_bins bins ST_k = sum_i T_k(t_i) and YT_k = sum_i y_i T_k(t_i), k < TG_MD, t_i = (x_i - xm) / xh with
// xm, xh the centre and half-width of the tile (the kernel forms them from the same two abscissae).  Extended precision on
// the host; a tile whose abscissae do not increase across it gets ST_0 = NaN and is never used.
static void build_moments(const double *x, const double *y, long Nx, int mt_bins, std::vector<double> &out)
{
    const int n = TG_MD, nc = TG_MCHK;
    const long nt = Nx / mt_bins;
    out.assign((size_t)TG_MOM_HDR + (size_t)nt * 2 * n, 0.0);
    const long double pi = 3.141592653589793238462643383279502884L;
    long double tn[TG_MD + TG_MCHK];
    for (int j = 0; j < n; j++) tn[j] = cosl(pi * (j + 0.5L) / n);                 // roots of T_n
    { const int ext[TG_MCHK] = {0, 2, 4, 6, 9}; for (int c = 0; c < nc; c++) tn[n + c] = cosl(pi * ext[c] / n); }   // extrema of T_n: the node polynomial peaks there
    long double W[TG_MD][TG_MD];
    for (int k = 0; k < n; k++) for (int j = 0; j < n; j++) W[k][j] = (k == 0 ? 1.0L : 2.0L) / n * cosl(pi * k * (j + 0.5L) / n);
    for (int k = 0; k < n; k++) for (int j = 0; j < n; j++) out[(size_t)k * n + j] = (double)W[k][j];
    for (int c = 0; c < nc; c++) for (int j = 0; j < n; j++) {
        long double e = 0.0L;
        for (int k = 0; k < n; k++) e += W[k][j] * cosl(k * acosl(tn[n + c]));
        out[(size_t)(n + c) * n + j] = (double)e;
    }
    for (int j = 0; j < n + nc; j++) out[(size_t)(n + nc) * n + j] = (double)tn[j];
    for (long t = 0; t < nt; t++) {
        const double *xs = x + t * mt_bins, *ys = y + t * mt_bins;
        const double xa = xs[0], xb = xs[mt_bins - 1], xm = 0.5 * (xa + xb), xh = 0.5 * (xb - xa);
        long double S[TG_MD] = {0}, Y[TG_MD] = {0};
        bool good = (xh > 0.0) && std::isfinite(xm) && std::isfinite(xh);
        for (int i = 0; i < mt_bins && good; i++) {
            const long double tt = ((long double)xs[i] - xm) / xh;
            if (!(tt >= -1.0L - 1e-12L && tt <= 1.0L + 1e-12L) || !std::isfinite(ys[i])) { good = false; break; }
            long double T0 = 1.0L, T1 = tt;
            S[0] += 1.0L; Y[0] += ys[i]; S[1] += T1; Y[1] += ys[i] * T1;
            for (int k = 2; k < n; k++) { const long double T2 = 2.0L * tt * T1 - T0; S[k] += T2; Y[k] += ys[i] * T2; T0 = T1; T1 = T2; }
        }
        double *o = out.data() + TG_MOM_HDR + (size_t)t * 2 * n;
        for (int k = 0; k < n; k++) { o[k] = (double)S[k]; o[n + k] = (double)Y[k]; }
        if (!good) o[0] = std::numeric_limits<double>::quiet_NaN();
    }
}

// Launch with (pdl) or without the programmatic-serialization attribute: with it the kernel may become resident while its
// predecessor in the stream still runs and synchronises on it with tg_pdl_wait() (tg_internal.cuh).
template <class Arg>
static cudaError_t launch_k(void (*kern)(Arg), dim3 grid, dim3 block, size_t smem, cudaStream_t st, bool pdl, const Arg &arg)
{
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kern, arg);
}

static cudaEvent_t next_event(tg_handle *h)
{
    if (h->ev_used == h->ev_pool.size()) { cudaEvent_t e; cudaEventCreate(&e); h->ev_pool.push_back(e); }
    return h->ev_pool[h->ev_used++];
}

static int harvest_events(tg_handle *h)
{
    if (h->ev_used == 0) return TG_OK;
    CK(cudaStreamSynchronize(h->stream));
    for (size_t k = 0; k + 1 < h->ev_used; k += 2) {
        float ms = 0;
        CK(cudaEventElapsedTime(&ms, h->ev_pool[k], h->ev_pool[k + 1]));
        h->ms_eval += ms;
    }
    h->ev_used = 0;
    return TG_OK;
}

static int build_units(tg_handle *h, tg_handle::UnitTable &u, int B);

// prologue (optional) + fused kernel + per-chain reduction for B points
static int launch_eval_kernels(tg_handle *h, EvalBuf &b, const double *params_dev, int B, double *model, bool do_prologue, bool do_reduce, bool pdl = false)
{
    if (do_prologue) {
        TgProArgs pa{h->d, b.t, params_dev, h->pri, h->sw};
        tg_k_prologue<<<B, TG_PRO_THREADS, TG_PRO_SMEM(h->d.Np), h->stream>>>(pa);
        h->n_launch_total++;
    }
    tg_handle::UnitTable &u = (&b == &h->sp) ? h->u_sp : h->u_ev;
    if (B != u.B) { int r = build_units(h, u, B); if (r) return r; }     // never inside a graph capture: tg_run builds u_sp first
    const size_t esm = TG_EVAL_SMEM(h->chains_per_block, h->list_cap);
    TgEvalArgs ea{h->d, b.t, h->x, h->y, h->lnx, h->sig, b.partials, model, u.dev, B, h->nitems, h->chains_per_block, h->list_cap, h->mom, h->mom_sub, h->mom_fail, h->mom_note};
    dim3 grid(u.n);
    if (h->timing) {
        if (h->ev_used >= 8192) { int r = harvest_events(h); if (r) return r; }
        cudaEventRecord(next_event(h), h->stream);
    }
    CK(launch_k(tg_k_eval, grid, dim3(TG_THREADS), esm, h->stream, pdl && !h->timing, ea));
    if (h->timing) cudaEventRecord(next_event(h), h->stream);
    h->n_eval_launch++; h->n_launch_total++;
    if (do_reduce) { tg_k_reduce<<<B, 256, 0, h->stream>>>(b.partials, h->nitems, b.sums); h->n_launch_total++; }
    CK(cudaGetLastError());
    return TG_OK;
}

// Work items of the fused kernel (scheduling only -- the kernel classifies every sub-tile again from
// the window unions of the launch, so results never depend on this table): single sub-tiles of 1024
// bins where truncation windows are expected, heaviest first (they cost several times more than
// noise-only bins, and dispatching them first lets the light items fill the tail of the grid), then
// runs of light_sub sub-tiles over the noise-only parts.  Estimated from the B chains held in `b`.
static int build_items(tg_handle *h, EvalBuf &b, int B)
{
    const int SUB = TG_THREADS * 4, nt = (h->d.Nx + SUB - 1) / SUB, nc = h->d.ncomp_max;
    std::vector<int4> items;
    if (h->no_order) {
        for (int t = 0; t < nt; t++) items.push_back(make_int4(t * SUB, 1, 0, 1));
    } else {
        std::vector<int> wlo(B), whi(B), lo(std::max(nc, 1)), hi(std::max(nc, 1));
        int ncomp = 0;
        CK(cudaStreamSynchronize(h->stream));
        CK(cudaMemcpy(wlo.data(), b.t.wlo, (size_t)B * 4, cudaMemcpyDeviceToHost));
        CK(cudaMemcpy(whi.data(), b.t.whi, (size_t)B * 4, cudaMemcpyDeviceToHost));
        CK(cudaMemcpy(&ncomp, b.t.ncomp, sizeof(int), cudaMemcpyDeviceToHost));
        if (nc > 0) {
            CK(cudaMemcpy(lo.data(), b.t.lo, (size_t)nc * 4, cudaMemcpyDeviceToHost));
            CK(cudaMemcpy(hi.data(), b.t.hi, (size_t)nc * 4, cudaMemcpyDeviceToHost));
        }
        int ulo = h->d.Nx, uhi = 0;
        for (int c = 0; c < B; c++) if (whi[c] > wlo[c]) { ulo = std::min(ulo, wlo[c]); uhi = std::max(uhi, whi[c]); }
        int t0 = nt, t1 = nt;                       // heavy tiles [t0, t1), one tile of margin for drifting windows
        if (uhi > ulo) { t0 = std::max(0, ulo / SUB - 1); t1 = std::min(nt, (uhi + SUB - 1) / SUB + 1); }
        std::vector<int> diff(nt + 1, 0), w(nt, 0);
        for (int k = 0; k < ncomp && k < nc; k++) {
            if (hi[k] <= lo[k]) continue;
            const int a0 = std::max(0, lo[k] / SUB), a1 = std::min(nt - 1, (hi[k] - 1) / SUB);
            if (a1 >= a0) { diff[a0]++; diff[a1 + 1]--; }
        }
        for (int t = 0, c = 0; t < nt; t++) { c += diff[t]; w[t] = c; }
        const int L = std::max(1, h->light_sub);
        // background-only stretches: runs of aligned moment tiles the kernel evaluates through the moments of the data (kind 2, cheap),
        // bin-by-bin runs (kind 0) for misaligned ends and for the tiles earlier launches flagged (their polynomial check failed)
        const int ms = h->mom ? h->mom_sub : 0;
        std::vector<int> flag((size_t)std::max<long>(h->n_mom_tiles, 1), 0);
        if (ms > 0 && h->n_mom_tiles > 0) CK(cudaMemcpy(flag.data(), h->mom_fail, (size_t)h->n_mom_tiles * 4, cudaMemcpyDeviceToHost));
        auto usable = [&](int t, int tb) { return ms > 0 && t % ms == 0 && t + ms <= tb && (long)(t + ms) * SUB <= (long)h->d.Nx && !flag[t / ms]; };
        auto light_region = [&](int ta, int tb) {
            int t = ta;
            while (t < tb) {
                if (usable(t, tb)) {
                    int n = 1;
                    while (n < h->mom_run && usable(t + n * ms, tb)) n++;
                    items.push_back(make_int4(t * SUB, n * ms, 0, 2)); t += n * ms;
                } else {
                    int e = t + 1;
                    while (e < tb && e - t < L && !usable(e, tb)) e++;
                    items.push_back(make_int4(t * SUB, e - t, 0, 0)); t = e;
                }
            }
        };
        light_region(0, t0);
        const int Hs = std::max(1, h->heavy_sub);   // sub-tiles per item where windows are expected (2: the kernel takes them 8 bins per thread)
        for (int t = t0; t < t1; t += Hs) { int wm = 0; for (int q = t; q < std::min(t + Hs, t1); q++) wm = std::max(wm, w[q]); items.push_back(make_int4(t * SUB, std::min(Hs, t1 - t), wm, 1)); }
        light_region(t1, nt);
    }
    bool same = items.size() == h->items_host.size();
    for (size_t k = 0; same && k < items.size(); k++) same = items[k].x == h->items_host[k].x && items[k].y == h->items_host[k].y && items[k].z == h->items_host[k].z && items[k].w == h->items_host[k].w;
    if (same) return TG_OK;
    h->items_host = items;
    h->nitems = (int)items.size();
    h->u_sp.B = h->u_ev.B = -1;                     // units are rebuilt at the next launch
    drop_graphs(h);                                 // grid size changed
    return TG_OK;
}

// Units = items x chain groups: few chains per unit where windows are expected (an item there costs
// 5-10x a noise-only sub-tile), up to chains_per_block elsewhere; costliest first (LPT scheduling).
static int build_units(tg_handle *h, tg_handle::UnitTable &u, int B)
{
    std::vector<int4> units;
    std::vector<double> cost;
    const int gl = std::max(1, h->chains_per_block), gh = std::max(1, std::min(h->heavy_group, gl));
    // hybrid: the costliest hybrid_pct percent of the window-region items take hybrid_group chains per unit (set-up and
    // list building shared by the chains of a unit), the rest heavy_group (small units for the tail of the grid)
    std::vector<char> big(h->items_host.size(), 0);
    if (h->hybrid_pct > 0 && !h->no_order) {
        std::vector<int> hv;
        for (size_t k = 0; k < h->items_host.size(); k++) if (h->items_host[k].w == 1) hv.push_back((int)k);
        std::stable_sort(hv.begin(), hv.end(), [&](int a, int b2) { const int4 &p = h->items_host[a], &q = h->items_host[b2]; return p.y * (20.0 + 6.25 * p.z) > q.y * (20.0 + 6.25 * q.z); });
        const size_t nb = hv.size() * (size_t)std::min(100, h->hybrid_pct) / 100;
        for (size_t k = 0; k < nb; k++) big[hv[k]] = 1;
    }
    for (size_t k = 0; k < h->items_host.size(); k++) {
        const int4 it = h->items_host[k];
        const int gmax = (it.w == 1 && !h->no_order) ? (big[k] ? std::max(1, std::min(h->hybrid_group, gl)) : gh) : gl;
        const int groups = (B + gmax - 1) / gmax, g = (B + groups - 1) / groups;      // balanced
        for (int c0 = 0; c0 < B; c0 += g) {
            const int G = std::min(g, B - c0);
            units.push_back(make_int4(it.x, it.y | (G << 16), c0, (int)k));
            cost.push_back(it.w == 2 ? (double)G * (it.y / std::max(1, h->mom_sub)) * 8.0 + 15.0 : (double)G * it.y * (20.0 + 6.25 * it.z) + 15.0);
        }
    }
    if (!h->no_order) {
        std::vector<int> idx(units.size());
        for (size_t k = 0; k < idx.size(); k++) idx[k] = (int)k;
        std::stable_sort(idx.begin(), idx.end(), [&](int a, int b2) { return cost[a] > cost[b2]; });
        std::vector<int4> u2(units.size());
        for (size_t k = 0; k < idx.size(); k++) u2[k] = units[idx[k]];
        units.swap(u2);
    }
    if ((int)units.size() > u.cap) {
        void *q = nullptr;
        CK(cudaMalloc(&q, units.size() * sizeof(int4)));
        h->allocs.push_back(q);                     // the old (smaller) table is released with the handle
        u.dev = (int4 *)q; u.cap = (int)units.size();
    }
    CK(cudaStreamSynchronize(h->stream));
    CK(cudaMemcpy(u.dev, units.data(), units.size() * sizeof(int4), cudaMemcpyHostToDevice));
    u.n = (int)units.size(); u.B = B;
    return TG_OK;
}

// The fused kernel flagged background-only tiles whose polynomial check failed since the work items were built: rebuild the
// items around them (they cost a bin-by-bin evaluation; left inside a long run of cheap moment tiles they would be the tail
// of the launch).  Results never depend on the items.
static int refresh_items_if_flagged(tg_handle *h)
{
    if (!h->mom_note || !*(volatile int *)h->mom_note) return TG_OK;
    *(volatile int *)h->mom_note = 0;
    return build_items(h, h->sp, h->s.C);
}

extern "C" int tg_set_tuning(tg_handle *h, int chains_per_block, int heavy_group, int light_subtiles, int use_graph)
{
    if (!h) return fail(TG_E_BADARG, "null handle");
    if (chains_per_block > 0) { h->chains_per_block = std::min(chains_per_block, h->max_cpb); h->u_sp.B = h->u_ev.B = -1; }
    if (heavy_group > 0) { h->heavy_group = heavy_group; h->u_sp.B = h->u_ev.B = -1; }
    if (use_graph >= 0) h->use_graph = use_graph;
    drop_graphs(h);
    if (light_subtiles > 0 && light_subtiles != h->light_sub) {
        h->light_sub = std::min(light_subtiles, 4096);
        int r = build_items(h, h->sp, h->s.C); if (r) return r;
    }
    return TG_OK;
}

static int create_impl(const tg_problem *p, tg_handle *h);

extern "C" int tg_destroy(tg_handle *h);

extern "C" int tg_create(const tg_problem *p, tg_handle **out)
{
    if (!p || !out) return fail(TG_E_BADARG, "null argument");
    *out = nullptr;
    tg_handle *h = new tg_handle();
    const int r = create_impl(p, h);
    if (r != TG_OK) { const std::string keep = g_err; tg_destroy(h); g_err = keep; return r; }
    *out = h;
    return TG_OK;
}

static int create_impl(const tg_problem *p, tg_handle *h)
{
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return fail(TG_E_NODEVICE, "no CUDA device (libtamcmc_gpu has no CPU fallback)");
    if (p->device < 0 || p->device >= ndev) return fail(TG_E_BADARG, "bad device ordinal");
    if (p->Nx < 2 || p->Nx > 2000000000L || !p->x || !p->y) return fail(TG_E_BADARG, "bad data arrays");
    if (p->model_case == 4 || p->model_case == 5) return fail(TG_E_UNSUPPORTED, "model case 4/5: the reference itself exits (models.cpp:724, model_def.cpp:231-237)");
    if (p->model_case < 0 || p->model_case > 14) return fail(TG_E_BADARG, "unknown model_fct_name_switch");
    if (p->likelihood_case < 0 || p->likelihood_case > 1) return fail(TG_E_BADARG, "unknown likelihood_fct_name_switch");
    if (p->prior_case < 0 || p->prior_case > 3) return fail(TG_E_BADARG, "unknown prior_fct_name_switch");
    if (p->likelihood_case == 1 && !p->sigma_y) return fail(TG_E_BADARG, "chi_square needs sigma_y");
    int np = 0;
    for (int k = 0; k < 11; k++) { if (p->plength[k] < 0) return fail(TG_E_BADARG, "negative plength"); np += p->plength[k]; }
    if (p->model_case >= 2 && np != p->Np) return fail(TG_E_BADARG, "Np != sum(plength)");
    if (p->Np < 1 || !p->params0 || !p->relax || !p->priors || !p->priors_switch) return fail(TG_E_BADARG, "bad parameter arrays");
    if (p->model_case == 0 && p->Np < 4) return fail(TG_E_BADARG, "model_Test_Gaussian needs 4 parameters");
    if (p->model_case == 1 && p->Np < 7) return fail(TG_E_BADARG, "model_Harvey_Gaussian needs 7 parameters");
    if (p->model_case >= 2 && p->plength[6] < 6) return fail(TG_E_BADARG, "Nsplit < 6");
    if (p->model_case >= 2 && p->plength[8] < 1) return fail(TG_E_BADARG, "Nnoise < 1");
    for (int l = 0; l < 4; l++) if (p->plength[2 + l] > TG_MAXORD) return fail(TG_E_BADARG, "more than TG_MAXORD frequencies per degree");
    if (p->Nchains < 1 || !p->Tcoefs) return fail(TG_E_BADARG, "bad chain setup");
    const int Ce = p->chains_per_ensemble > 0 ? p->chains_per_ensemble : p->Nchains;
    if (p->Nchains % Ce != 0) return fail(TG_E_BADARG, "Nchains not a multiple of chains_per_ensemble");
    if (p->n_learn < 0 || p->n_learn > TG_MAX_LEARN || (p->n_learn > 0 && (!p->Nt_learn || !p->periods_learn))) return fail(TG_E_BADARG, "bad learning zones");
    for (int l = 0; l < p->n_learn; l++) if (p->periods_learn[l] < 1) return fail(TG_E_BADARG, "periods_learn must be >= 1");

    h->device = p->device;
    CK(cudaSetDevice(h->device));
    cudaDeviceProp prop; CK(cudaGetDeviceProperties(&prop, h->device));
    h->n_sm = prop.multiProcessorCount;
    CK(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
    CK(cudaStreamCreateWithFlags(&h->side, cudaStreamNonBlocking));
    h->no_split = getenv("TG_NO_SPLIT_PRIOR") != nullptr; h->no_side = getenv("TG_NO_SIDE_STREAM") != nullptr;
    if (const char *e = getenv("TG_PDL")) h->pdl = atoi(e) != 0;
    CK(cudaEventCreateWithFlags(&h->ev_fork, cudaEventDisableTiming));
    CK(cudaEventCreateWithFlags(&h->ev_join, cudaEventDisableTiming));

    TgDesc &d = h->d;
    d.model_case = p->model_case; d.like_case = p->likelihood_case; d.prior_case = p->prior_case;
    for (int k = 0; k < 11; k++) d.pl[k] = p->plength[k];
    d.Np = p->Np;
    d.likelihood_p = (long)p->likelihood_params;
    for (int k = 0; k < 4; k++) d.extra[k] = p->extra_priors[k];
    d.Nx = (int)p->Nx;
    if (p->Nx_global > 0) { d.Nx_global = p->Nx_global; d.shard_lo = p->shard_lo; d.x_first = p->x_first; d.x_second = p->x_second; d.x_last = p->x_last; }
    else { d.Nx_global = p->Nx; d.shard_lo = 0; d.x_first = p->x[0]; d.x_second = p->x[1]; d.x_last = p->x[p->Nx - 1]; }
    ncomp_for(p, &d.ncomp_max, &d.nmodes_max);
    std::vector<int> idx;
    for (int i = 0; i < p->Np; i++) if (p->relax[i] == 1) idx.push_back(i);
    d.Nv = (int)idx.size();
    if (d.Nv < 1) return fail(TG_E_BADARG, "no free parameter (relax==1)");

    // ---- data ----
    const size_t Nx = (size_t)p->Nx;
    DA(h->x, Nx); DA(h->y, Nx); DA(h->lnx, Nx);
    CK(cudaMemcpyAsync(h->x, p->x, Nx * 8, cudaMemcpyHostToDevice, h->stream));
    CK(cudaMemcpyAsync(h->y, p->y, Nx * 8, cudaMemcpyHostToDevice, h->stream));
    if (p->sigma_y) { DA(h->sig, Nx); CK(cudaMemcpyAsync(h->sig, p->sigma_y, Nx * 8, cudaMemcpyHostToDevice, h->stream)); }
    tg_k_lnx<<<(unsigned)((Nx + 255) / 256), 256, 0, h->stream>>>(h->x, h->lnx, (int)Nx);
    {   // moments of the data for the background-only stretches (chi22p likelihood only; TG_MOMENTS=0 turns the path off)
        const char *e = getenv("TG_MOMENTS"), *e2 = getenv("TG_MOM_SUB");
        if (e2) h->mom_sub = std::max(1, std::min(8, atoi(e2)));
        if (const char *e3 = getenv("TG_MOM_RUN")) h->mom_run = std::max(1, atoi(e3));
        if (!(e && atoi(e) == 0) && p->likelihood_case == 0 && (long)Nx >= (long)h->mom_sub * TG_THREADS * 4) {
            std::vector<double> mom;
            build_moments(p->x, p->y, (long)Nx, h->mom_sub * TG_THREADS * 4, mom);
            DA(h->mom, mom.size());
            CK(cudaMemcpy(h->mom, mom.data(), mom.size() * 8, cudaMemcpyHostToDevice));
            h->n_mom_tiles = (long)Nx / (h->mom_sub * TG_THREADS * 4);
            DA(h->mom_fail, (size_t)h->n_mom_tiles);
            CK(cudaMemset(h->mom_fail, 0, (size_t)h->n_mom_tiles * 4));
            CK(cudaHostAlloc((void **)&h->mom_note, sizeof(int), cudaHostAllocMapped));
            *h->mom_note = 0;
        }
    }
    DA(h->pri, (size_t)4 * p->Np); DA(h->sw, p->Np); DA(h->index_to_relax, d.Nv); DA(h->Tcoefs, p->Nchains);
    CK(cudaMemcpyAsync(h->pri, p->priors, (size_t)4 * p->Np * 8, cudaMemcpyHostToDevice, h->stream));
    CK(cudaMemcpyAsync(h->sw, p->priors_switch, (size_t)p->Np * 4, cudaMemcpyHostToDevice, h->stream));
    CK(cudaMemcpyAsync(h->index_to_relax, idx.data(), (size_t)d.Nv * 4, cudaMemcpyHostToDevice, h->stream));
    CK(cudaMemcpyAsync(h->Tcoefs, p->Tcoefs, (size_t)p->Nchains * 8, cudaMemcpyHostToDevice, h->stream));

    // ---- tiling: partial sums are sized for one item per 1024-bin sub-tile (the finest partition) ----
    h->max_items = (d.Nx + TG_THREADS * 4 - 1) / (TG_THREADS * 4);
    h->no_order = (p->flags & TG_F_NO_TILE_ORDER) != 0;
    for (int t = 0; t < h->max_items; t++) h->items_host.push_back(make_int4(t * TG_THREADS * 4, 1, 0, 1));   // until chain tables exist
    h->nitems = h->max_items;
    const int C = p->Nchains;
    { int r = alloc_evalbuf(h, h->sp, C); if (r) return r; }
    { int r = alloc_evalbuf(h, h->ev, C); if (r) return r; }
    // shared-memory lists of the fused kernel: one list of <= ncomp_max components per chain of a group
    {
        h->list_cap = std::max(4, (d.ncomp_max + 3) & ~3);
        int maxdyn = 0; CK(cudaDeviceGetAttribute(&maxdyn, cudaDevAttrMaxSharedMemoryPerBlockOptin, h->device));
        const size_t budget = std::min<size_t>((size_t)maxdyn - 2048, 100 * 1024);       // <= 100 KB: two CTAs per SM stay resident
        int m = TG_MAXG;
        while (m > 1 && TG_EVAL_SMEM(m, h->list_cap) > budget) m--;
        if (TG_EVAL_SMEM(m, h->list_cap) > (size_t)maxdyn - 2048) return fail(TG_E_BADARG, "too many Lorentzian components for the shared-memory list of the fused kernel");
        h->max_cpb = m;
        const size_t need = TG_EVAL_SMEM(m, h->list_cap);
        CK(cudaFuncSetAttribute(tg_k_eval, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)need));
        const int groups = (C + m - 1) / m;                                               // balanced chain groups of <= max_cpb
        h->chains_per_block = (C + groups - 1) / groups;
        // where truncation windows lie: items of two sub-tiles (taken 8 bins per thread); one chain per unit for a single
        // ensemble (more, smaller units: better tail of the grid), three when many chains fill the GPU anyway (measured with the
        // background-only tiles on the moment path: one chain per unit is best up to ~100 chains, 3.9 vs 4.0 us per chain at 80)
        h->heavy_sub = 2;
        h->heavy_group = (C >= 128) ? 3 : 1;
        // a spectrum so short that pairs of sub-tiles would not even fill the GPU once (S-local: 10k bins, 10 chains): single
        // sub-tiles, twice the units (25.7 vs 29.9 us per step there)
        const bool short_spectrum = (long)((d.Nx + TG_THREADS * 4 - 1) / (TG_THREADS * 4)) * C < 4L * 296;
        if (short_spectrum) h->heavy_sub = 1;
        // developer knobs for scheduling experiments (never change results)
        // the costliest window-region items take three chains per unit (set-up and list building shared), the rest one chain
        // (small units for the tail): 70 % of them at 10 chains (49.0 -> 48.1 us), all of them from 32 chains on (-7 %)
        h->hybrid_group = (C < 64) ? 3 : ((C < 128) ? 4 : 5);
        h->hybrid_pct = short_spectrum ? 0 : ((C < 16) ? 70 : ((C < 32) ? 80 : 100));      // short spectra: as many units as possible
        if (const char *e = getenv("TG_GRAPH_STEPS")) h->steps_per_graph = std::max(1, std::min(256, atoi(e)));
        if (const char *e = getenv("TG_HYBRID_PCT")) h->hybrid_pct = atoi(e);
        if (const char *e = getenv("TG_HYBRID_GROUP")) h->hybrid_group = std::max(1, atoi(e));
        if (const char *e = getenv("TG_HEAVY_SUB")) h->heavy_sub = std::max(1, atoi(e));
        if (const char *e = getenv("TG_HEAVY_GROUP")) h->heavy_group = std::max(1, atoi(e));
    }

    // ---- MALA configuration ----
    TgMala &m = h->m;
    m.c0 = p->c0; m.eps1 = p->epsilon1; m.eps2 = p->epsilon2; m.A1 = p->A1; m.target = p->target_acceptance;
    m.n_learn = p->n_learn;
    for (int l = 0; l < p->n_learn; l++) { m.Nt_learn[l] = p->Nt_learn[l]; m.periods[l] = p->periods_learn[l]; }
    if (p->n_learn > 0) m.Nt_learn[p->n_learn] = p->Nt_learn[p->n_learn];
    m.dN_mixing = p->dN_mixing; m.seed = p->seed; m.flags = p->flags;

    // ---- sampler state (MALA::MALA + init_proposal, Model_def ctor) ----
    TgState &s = h->s;
    s.C = C; s.Ce = Ce; s.E = C / Ce; s.Nv = d.Nv; s.Np = d.Np;
    const size_t Nv = d.Nv, Np = d.Np;
    DA(s.vars_cur, C * Nv); DA(s.vars_prop, C * Nv); DA(s.params_cur, C * Np);
    s.params_prop = h->sp.params;
    DA(s.logL_cur, C); DA(s.logPrior_cur, C); DA(s.logPost_cur, C); DA(s.Pmove, C); DA(s.moved, C);
    DA(s.mu, 3 * C * Nv); DA(s.sigma, C); DA(s.cov, 3 * C * Nv * Nv); DA(s.chol, 3 * C * Nv * Nv); DA(s.live, C); DA(s.chol_fail, 3 * C); DA(s.sticky_status, C);
    DA(h->stage_mu, C * Nv); DA(h->stage_cov, C * Nv * Nv);
    DA(s.ctrl, 1);
    s.index_to_relax = h->index_to_relax; s.Tcoefs = h->Tcoefs;
    {
        std::vector<double> vars(C * Nv), params(C * Np), sigma(C), cov(C * Nv * Nv, 0.0);
        for (int c = 0; c < C; c++) {
            for (size_t i = 0; i < Np; i++) params[c * Np + i] = p->params0[i];
            for (size_t v = 0; v < Nv; v++) vars[c * Nv + v] = p->params0[idx[v]];
            sigma[c] = 2.38 * 2.38 * std::pow(p->Tcoefs[c], 0.2) / (double)Nv;         // MALA.cpp:273
            for (size_t v = 0; v < Nv; v++) { const double e = p->init_err ? p->init_err[v] : 1.0; cov[(c * Nv + v) * Nv + v] = e * e; }  // :264-266
        }
        CK(cudaMemcpyAsync(s.vars_cur, vars.data(), vars.size() * 8, cudaMemcpyHostToDevice, h->stream));
        CK(cudaMemcpyAsync(s.mu, vars.data(), vars.size() * 8, cudaMemcpyHostToDevice, h->stream));           // mu.row(m)=vars, :284
        CK(cudaMemcpyAsync(s.params_cur, params.data(), params.size() * 8, cudaMemcpyHostToDevice, h->stream));
        CK(cudaMemcpyAsync(s.sigma, sigma.data(), sigma.size() * 8, cudaMemcpyHostToDevice, h->stream));
        CK(cudaMemcpyAsync(s.cov, cov.data(), cov.size() * 8, cudaMemcpyHostToDevice, h->stream));
        CK(cudaMemsetAsync(s.live, 0, C * 4, h->stream));
        CK(cudaMemsetAsync(s.chol_fail, 0, 3 * C * 4, h->stream));
        CK(cudaMemsetAsync(s.sticky_status, 0, C * 4, h->stream));
        CK(cudaMemsetAsync(s.ctrl, 0, sizeof(TgCtrl), h->stream));
        CK(cudaStreamSynchronize(h->stream));
    }
    // shared memory: the speculative kernel stages the Nv x Nv matrix when it fits; the propose kernel only
    // needs the prologue scratch
    {
        const size_t ld = (size_t)(d.Nv | 1);
        const size_t need_mat = ((size_t)d.Nv * ld + 4 * (size_t)d.Nv) * 8, need_pro = TG_PRO_SMEM(d.Np);
        int maxdyn = 0; CK(cudaDeviceGetAttribute(&maxdyn, cudaDevAttrMaxSharedMemoryPerBlockOptin, h->device));
        const size_t budget = (size_t)maxdyn > 12 * 1024 ? (size_t)maxdyn - 12 * 1024 : 0;   // static smem of the kernels
        h->smem_matrix = (need_mat <= budget) ? 1 : 0;
        h->spec_smem = h->smem_matrix ? need_mat : 4 * (size_t)d.Nv * 8;
        const size_t need_tri = ((size_t)d.Nv * (d.Nv + 1) / 2 + d.Nv) * 8;
        h->tri_smem = (need_tri <= budget) ? 1 : 0;
        h->propose_smem = std::max(need_pro, h->tri_smem ? need_tri : (size_t)0);
        // Nv <= 96: the adaptation works out of registers (tg_spec96, no dynamic shared memory) and its CTAs
        // ride in the propose grid; larger problems run it as a separate kernel after the propose kernel
        h->fused_spec = (h->tri_smem && h->smem_matrix && d.Nv <= 96) ? 1 : 0;
        CK(cudaFuncSetAttribute(tg_k_spec, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->spec_smem));
        CK(cudaFuncSetAttribute(tg_k_propose, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->propose_smem));   // static + dynamic may exceed 48 KB
        if (need_pro > 48 * 1024) {
            CK(cudaFuncSetAttribute(tg_k_prologue, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)need_pro));
        }
    }
    // record ring
    {
        const size_t per_step = (size_t)C * (Nv + 4) * 8 + (size_t)C * 5 + (size_t)s.E * 24;
        long cap = (long)std::min<size_t>(1024, std::max<size_t>(1, ((size_t)256 << 20) / per_step));
        TgRecDev &r = h->rec; r.cap = cap;
        DA(r.vars, (size_t)cap * C * Nv); DA(r.logL, (size_t)cap * C); DA(r.logPrior, (size_t)cap * C); DA(r.logPost, (size_t)cap * C);
        DA(r.Pmove, (size_t)cap * C); DA(r.moved, (size_t)cap * C); DA(r.status, (size_t)cap * C);
        DA(r.swap_attempt, (size_t)cap * s.E); DA(r.swap_chain0, (size_t)cap * s.E); DA(r.swap_done, (size_t)cap * s.E); DA(r.swap_Pswap, (size_t)cap * s.E);
    }
    // evaluate every chain once at the initial point (model_def.cpp:139-143)
    {
        int r = launch_eval_kernels(h, h->sp, s.params_cur, C, nullptr, true, true); if (r) return r;
        tg_k_install_current<<<(C + 127) / 128, 128, 0, h->stream>>>(h->d, h->s, h->sp.t, h->sp.sums);
        TgSpecArgs sa{h->s, h->m, h->smem_matrix, 1};
        tg_k_spec<<<dim3(C, 1), 256, h->spec_smem, h->stream>>>(sa);      // chol(Sigma + eps2 I) of the live slot
        CK(cudaGetLastError());
        CK(cudaStreamSynchronize(h->stream));
    }
    { int r = build_items(h, h->sp, C); if (r) return r; }
    if (h->mom_note) {   // once more with the real work items: flags the background-only tiles whose polynomial check fails at the start
        int r = launch_eval_kernels(h, h->sp, s.params_cur, C, nullptr, false, true); if (r) return r;
        CK(cudaStreamSynchronize(h->stream));
        r = refresh_items_if_flagged(h); if (r) return r;
    }
    return TG_OK;
}

extern "C" int tg_destroy(tg_handle *h)
{
    if (!h) return TG_OK;
    cudaSetDevice(h->device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    drop_graphs(h);
    for (auto e : h->ev_pool) cudaEventDestroy(e);
    for (int k = 0; k < 2; k++) if (h->watch[k]) cudaEventDestroy(h->watch[k]);
    for (void *q : h->allocs) cudaFree(q);
    for (EvalBuf *b : {&h->sp, &h->ev}) { if (b->pin_in) cudaFreeHost(b->pin_in); if (b->pin_out) cudaFreeHost(b->pin_out); }
    if (h->mom_note) cudaFreeHost(h->mom_note);
    if (h->ev_fork) cudaEventDestroy(h->ev_fork);
    if (h->ev_join) cudaEventDestroy(h->ev_join);
    if (h->side) cudaStreamDestroy(h->side);
    if (h->stream) cudaStreamDestroy(h->stream);
    delete h;
    return TG_OK;
}

extern "C" int tg_sizes(tg_handle *h, int *Nv, int *Np, int *C, int *E, long *Nx)
{
    if (!h) return fail(TG_E_BADARG, "null handle");
    if (Nv) *Nv = h->d.Nv; if (Np) *Np = h->d.Np; if (C) *C = h->s.C; if (E) *E = h->s.E; if (Nx) *Nx = h->d.Nx;
    return TG_OK;
}

// ------------------------------------------------------------------------------------------
// Batched generate_model (parity / getmodel / reference-facing entry)
// ------------------------------------------------------------------------------------------
extern "C" int tg_eval_upload(tg_handle *h, const double *params, int B, const double *T)
{
    if (!h || !params || !T || B < 1) return fail(TG_E_BADARG, "tg_eval_upload: bad argument");
    CK(cudaSetDevice(h->device));
    { int r = alloc_evalbuf(h, h->ev, B); if (r) return r; }
    EvalBuf &b = h->ev;
    if (b.B != B) { CK(cudaStreamSynchronize(h->stream)); b.layout(B, h->d.Np); }
    // one host->device copy from the pinned staging buffer (a pageable source would make the copy synchronous)
    memcpy(b.pin_in, params, (size_t)B * h->d.Np * 8);
    memcpy(b.pin_in + (size_t)B * h->d.Np, T, (size_t)B * 8);
    CK(cudaMemcpyAsync(b.in, b.pin_in, (size_t)B * (h->d.Np + 1) * 8, cudaMemcpyHostToDevice, h->stream));
    return TG_OK;
}

static int eval_resident(tg_handle *h, int B, double *model_dev)
{
    if (B < 1 || B != h->ev.B) return fail(TG_E_BADARG, "tg_eval_resident: B differs from the uploaded batch");
    int r = launch_eval_kernels(h, h->ev, h->ev.params, B, model_dev, true, false); if (r) return r;
    tg_k_reduce_finalize<<<B, 256, 0, h->stream>>>(h->d, h->ev.partials, h->nitems, h->ev.sums, h->ev.T, h->ev.t, h->ev.logL, h->ev.logPrior, h->ev.logPost, h->ev.status);
    h->n_launch_total++;
    CK(cudaGetLastError());
    return TG_OK;
}

extern "C" int tg_eval_resident(tg_handle *h, int B)
{
    if (!h) return fail(TG_E_BADARG, "null handle");
    CK(cudaSetDevice(h->device));
    { int r = refresh_items_if_flagged(h); if (r) return r; }
    return eval_resident(h, B, nullptr);
}

extern "C" int tg_eval_fetch(tg_handle *h, int B, double *logL, double *logPrior, double *logPost, int *status)
{
    if (!h || B < 1 || B != h->ev.B) return fail(TG_E_BADARG, "tg_eval_fetch: bad argument");
    CK(cudaSetDevice(h->device));
    EvalBuf &b = h->ev;
    CK(cudaMemcpyAsync(b.pin_out, b.out, (size_t)B * 28, cudaMemcpyDeviceToHost, h->stream));     // one device->host copy
    CK(cudaStreamSynchronize(h->stream));
    if (logL) memcpy(logL, b.pin_out, (size_t)B * 8);
    if (logPrior) memcpy(logPrior, b.pin_out + B, (size_t)B * 8);
    if (logPost) memcpy(logPost, b.pin_out + 2 * (size_t)B, (size_t)B * 8);
    if (status) memcpy(status, b.pin_out + 3 * (size_t)B, (size_t)B * 4);
    return TG_OK;
}

extern "C" int tg_eval(tg_handle *h, const double *params, int B, const double *T,
                       double *logL, double *logPrior, double *logPost, int *status, double *model)
{
    if (!h || !params || !T || B < 1) return fail(TG_E_BADARG, "tg_eval: bad argument");
    { int r = refresh_items_if_flagged(h); if (r) return r; }
    if (!model && h->use_graph && !h->timing) {
        // the whole call as one graph launch: H2D of the packed inputs, prologue, fused kernel, reduce+finalize, D2H of the packed results
        CK(cudaSetDevice(h->device));
        { int r = alloc_evalbuf(h, h->ev, B); if (r) return r; }
        EvalBuf &b = h->ev;
        if (b.B != B) { CK(cudaStreamSynchronize(h->stream)); b.layout(B, h->d.Np); }
        if (h->ev_graph && h->ev_graph_B != B) { cudaGraphExecDestroy(h->ev_graph); h->ev_graph = nullptr; }
        if (!h->ev_graph) {
            if (h->u_ev.B != B) { int r = build_units(h, h->u_ev, B); if (r) return r; }
            cudaGraph_t g;
            const long lt = h->n_launch_total, le = h->n_eval_launch;
            CK(cudaStreamBeginCapture(h->stream, cudaStreamCaptureModeThreadLocal));
            cudaMemcpyAsync(b.in, b.pin_in, (size_t)B * (h->d.Np + 1) * 8, cudaMemcpyHostToDevice, h->stream);
            int r = eval_resident(h, B, nullptr);
            cudaMemcpyAsync(b.pin_out, b.out, (size_t)B * 28, cudaMemcpyDeviceToHost, h->stream);
            cudaError_t ce = cudaStreamEndCapture(h->stream, &g);
            h->n_launch_total = lt; h->n_eval_launch = le;
            if (r) return r;
            CK(ce);
            CK(cudaGraphInstantiate(&h->ev_graph, g, 0));
            cudaGraphDestroy(g);
            h->ev_graph_B = B;
        }
        memcpy(b.pin_in, params, (size_t)B * h->d.Np * 8);
        memcpy(b.pin_in + (size_t)B * h->d.Np, T, (size_t)B * 8);
        CK(cudaGraphLaunch(h->ev_graph, h->stream));
        h->n_launch_total += 3; h->n_eval_launch += 1;
        CK(cudaStreamSynchronize(h->stream));
        if (logL) memcpy(logL, b.pin_out, (size_t)B * 8);
        if (logPrior) memcpy(logPrior, b.pin_out + B, (size_t)B * 8);
        if (logPost) memcpy(logPost, b.pin_out + 2 * (size_t)B, (size_t)B * 8);
        if (status) memcpy(status, b.pin_out + 3 * (size_t)B, (size_t)B * 4);
        return TG_OK;
    }
    int r = tg_eval_upload(h, params, B, T); if (r) return r;
    double *model_dev = nullptr;
    if (model) CK(cudaMalloc(&model_dev, (size_t)B * h->d.Nx * 8));
    r = eval_resident(h, B, model_dev);
    if (r == TG_OK) r = tg_eval_fetch(h, B, logL, logPrior, logPost, status);
    if (model_dev) {
        if (r == TG_OK && cudaMemcpy(model, model_dev, (size_t)B * h->d.Nx * 8, cudaMemcpyDeviceToHost) != cudaSuccess) r = fail(TG_E_CUDA, "model copy failed");
        cudaFree(model_dev);
    }
    return r;
}

extern "C" int tg_last_work(tg_handle *h, int B, double *L, int *K)
{
    if (!h || B < 1 || B > h->ev.cap) return fail(TG_E_BADARG, "tg_last_work: bad argument");
    CK(cudaSetDevice(h->device));
    CK(cudaStreamSynchronize(h->stream));
    if (L) CK(cudaMemcpy(L, h->ev.t.work, (size_t)B * 8, cudaMemcpyDeviceToHost));
    if (K) {
        std::vector<double> nz(TG_NOISE_STRIDE);
        CK(cudaMemcpy(nz.data(), h->ev.t.noise, TG_NOISE_STRIDE * 8, cudaMemcpyDeviceToHost));
        int k = 0;
        for (int j = 0; j < (int)nz[TG_N_NH]; j++) if (nz[TG_HS * j + TG_H_A] != 0.0) k++;
        *K = k;
    }
    return TG_OK;
}

// ------------------------------------------------------------------------------------------
// Sampler state
// ------------------------------------------------------------------------------------------
extern "C" int tg_get_state(tg_handle *h, double *vars, double *mu, double *sigma, double *covarmat,
                            double *logL, double *logPrior, double *logPost)
{
    if (!h) return fail(TG_E_BADARG, "null handle");
    CK(cudaSetDevice(h->device));
    CK(cudaStreamSynchronize(h->stream));
    const size_t C = h->s.C, Nv = h->s.Nv;
    if (vars) CK(cudaMemcpy(vars, h->s.vars_cur, C * Nv * 8, cudaMemcpyDeviceToHost));
    if (mu || covarmat) {
        tg_k_gather_live<<<(unsigned)C, 256, 0, h->stream>>>(h->s, mu ? h->stage_mu : nullptr, covarmat ? h->stage_cov : nullptr);
        CK(cudaGetLastError());
        CK(cudaStreamSynchronize(h->stream));
    }
    if (mu) CK(cudaMemcpy(mu, h->stage_mu, C * Nv * 8, cudaMemcpyDeviceToHost));
    if (sigma) CK(cudaMemcpy(sigma, h->s.sigma, C * 8, cudaMemcpyDeviceToHost));
    if (covarmat) CK(cudaMemcpy(covarmat, h->stage_cov, C * Nv * Nv * 8, cudaMemcpyDeviceToHost));
    if (logL) CK(cudaMemcpy(logL, h->s.logL_cur, C * 8, cudaMemcpyDeviceToHost));
    if (logPrior) CK(cudaMemcpy(logPrior, h->s.logPrior_cur, C * 8, cudaMemcpyDeviceToHost));
    if (logPost) CK(cudaMemcpy(logPost, h->s.logPost_cur, C * 8, cudaMemcpyDeviceToHost));
    return TG_OK;
}

__global__ void tg_k_scatter_vars(TgState s)
{
    const int c = blockIdx.x;
    for (int v = threadIdx.x; v < s.Nv; v += blockDim.x) s.params_cur[(size_t)c * s.Np + s.index_to_relax[v]] = s.vars_cur[(size_t)c * s.Nv + v];
}

extern "C" int tg_set_state(tg_handle *h, const double *vars, const double *mu, const double *sigma, const double *covarmat)
{
    if (!h) return fail(TG_E_BADARG, "null handle");
    CK(cudaSetDevice(h->device));
    const size_t C = h->s.C, Nv = h->s.Nv;
    if (mu || covarmat) {
        // bring the live (mu, Sigma) of every chain into slot 0 (contiguous [C][..] block), overwrite what was given
        tg_k_gather_live<<<(unsigned)C, 256, 0, h->stream>>>(h->s, h->stage_mu, h->stage_cov);
        CK(cudaMemcpyAsync(h->s.mu, h->stage_mu, C * Nv * 8, cudaMemcpyDeviceToDevice, h->stream));
        CK(cudaMemcpyAsync(h->s.cov, h->stage_cov, C * Nv * Nv * 8, cudaMemcpyDeviceToDevice, h->stream));
        CK(cudaMemsetAsync(h->s.live, 0, C * 4, h->stream));
        if (mu) CK(cudaMemcpyAsync(h->s.mu, mu, C * Nv * 8, cudaMemcpyHostToDevice, h->stream));
        if (covarmat) CK(cudaMemcpyAsync(h->s.cov, covarmat, C * Nv * Nv * 8, cudaMemcpyHostToDevice, h->stream));
        TgSpecArgs sa{h->s, h->m, h->smem_matrix, 1};
        tg_k_spec<<<dim3((unsigned)C, 1), 256, h->spec_smem, h->stream>>>(sa);
        CK(cudaGetLastError());
    }
    if (sigma) CK(cudaMemcpyAsync(h->s.sigma, sigma, C * 8, cudaMemcpyHostToDevice, h->stream));
    if (vars) {
        CK(cudaMemcpyAsync(h->s.vars_cur, vars, C * Nv * 8, cudaMemcpyHostToDevice, h->stream));
        tg_k_scatter_vars<<<(unsigned)C, 128, 0, h->stream>>>(h->s);
        int r = launch_eval_kernels(h, h->sp, h->s.params_cur, (int)C, nullptr, true, true); if (r) return r;
        tg_k_install_current<<<((int)C + 127) / 128, 128, 0, h->stream>>>(h->d, h->s, h->sp.t, h->sp.sums);
        CK(cudaGetLastError());
        CK(cudaStreamSynchronize(h->stream));
        { int r = build_items(h, h->sp, (int)C); if (r) return r; }
    }
    CK(cudaStreamSynchronize(h->stream));
    return TG_OK;
}

// ------------------------------------------------------------------------------------------
// MALA::execute on the device
// ------------------------------------------------------------------------------------------
static int launch_step(tg_handle *h, bool begin, bool end, int presummed)
{
    const int C = h->s.C;
    if (begin) {
        // The adaptation candidates (Robbins-Monro + Cholesky for both outcomes of the decision) depend on the
        // state the previous finish kernel left and on nothing of this step (the CTA of the "accepted" outcome
        // recomputes the proposed point itself), and only the finish kernel needs them: they run on a side stream
        // next to the proposal, the prologue and the fused kernel (a parallel branch once captured in a graph).
        const int learn = h->m.n_learn > 0, fused = learn && h->fused_spec;
        const int split = h->tri_smem && !h->no_split;        // the log-prior of the proposed point gets its own CTAs (role 3) on the side stream
        TgProposeArgs pa{h->s, h->m, TgProArgs{h->d, h->sp.t, h->sp.params, h->pri, h->sw}, h->smem_matrix, h->tri_smem, fused, h->smem_matrix, 0,
                         split ? TG_PRO_MODEL : (TG_PRO_MODEL | TG_PRO_PRIOR)};
        if (fused || split) {
            CK(cudaEventRecord(h->ev_fork, h->stream));
            CK(cudaStreamWaitEvent(h->side, h->ev_fork, 0));
            TgProposeArgs ps = pa; ps.role_base = fused ? 1 : 3;
            tg_k_propose<<<dim3(C, (fused ? 2 : 0) + (split ? 1 : 0)), 256, h->propose_smem, h->no_side ? h->stream : h->side>>>(ps);
            CK(cudaEventRecord(h->ev_join, h->side));
            h->side_pending = true;
            h->n_launch_total++;
        }
        CK(launch_k(tg_k_propose, dim3(C, 1), dim3(256), h->propose_smem, h->stream, h->pdl && !h->timing, pa));
        h->n_launch_total++;
        if (learn && !fused) {       // very large Nv: the candidates read the proposed point from HBM, so they follow the propose kernel
            CK(cudaEventRecord(h->ev_fork, h->stream));
            CK(cudaStreamWaitEvent(h->side, h->ev_fork, 0));
            TgSpecArgs sa{h->s, h->m, h->smem_matrix, 0};
            tg_k_spec<<<dim3(C, 2), 256, h->spec_smem, h->side>>>(sa);
            CK(cudaEventRecord(h->ev_join, h->side));
            h->side_pending = true;
            h->n_launch_total++;
        }
        int r = launch_eval_kernels(h, h->sp, h->sp.params, C, nullptr, false, presummed != 0, h->pdl); if (r) return r;
    }
    if (end) {
        if (h->side_pending) { CK(cudaStreamWaitEvent(h->stream, h->ev_join, 0)); h->side_pending = false; }
        TgFinishArgs fa{h->d, h->s, h->m, h->sp.t, h->sp.partials, h->nitems, h->sp.sums, presummed, h->rec, 1};
        CK(launch_k(tg_k_finish, dim3(C), dim3(256), 0, h->stream, h->pdl && !presummed && !h->timing, fa));
        h->n_launch_total++;
    }
    CK(cudaGetLastError());
    return TG_OK;
}

static int set_ctrl(tg_handle *h, long i, long i0)
{
    TgCtrl c{i, i0, 0u, 0};
    CK(cudaMemcpyAsync(h->s.ctrl, &c, sizeof(c), cudaMemcpyHostToDevice, h->stream));
    return TG_OK;
}

static int drain_records(tg_handle *h, tg_records *rec, long done, long n)
{
    if (!rec || n <= 0) return TG_OK;
    const size_t C = h->s.C, Nv = h->s.Nv, E = h->s.E;
    const TgRecDev &r = h->rec;
#define CP(dst, src, per, T) if (rec->dst) CK(cudaMemcpyAsync(rec->dst + (size_t)done * (per), r.src, (size_t)n * (per) * sizeof(T), cudaMemcpyDeviceToHost, h->stream))
    CP(vars, vars, C * Nv, double); CP(logL, logL, C, double); CP(logPrior, logPrior, C, double); CP(logPost, logPost, C, double);
    CP(Pmove, Pmove, C, double); CP(moved, moved, C, unsigned char); CP(status, status, C, int);
    CP(swap_attempt, swap_attempt, E, int); CP(swap_chain0, swap_chain0, E, int); CP(swap_done, swap_done, E, int); CP(swap_Pswap, swap_Pswap, E, double);
#undef CP
    CK(cudaStreamSynchronize(h->stream));
    return TG_OK;
}

extern "C" int tg_run(tg_handle *h, long i0, long nsteps, tg_records *rec)
{
    if (!h || nsteps < 0) return fail(TG_E_BADARG, "tg_run: bad argument");
    CK(cudaSetDevice(h->device));
    if (nsteps >= 64) { if (h->mom_note) *(volatile int *)h->mom_note = 0; int r = build_items(h, h->sp, h->s.C); if (r) return r; }   // follow the windows of the chains (scheduling only)
    else { int r = refresh_items_if_flagged(h); if (r) return r; }
    if (h->u_sp.B != h->s.C) { int r = build_units(h, h->u_sp, h->s.C); if (r) return r; }
    long done = 0;
    while (done < nsteps) {
        const long n = std::min(nsteps - done, h->rec.cap);
        { int r = set_ctrl(h, i0 + done, i0 + done); if (r) return r; }
        long k = 0;
        for (int pass = 0; pass < 2 && h->use_graph && !h->timing; pass++) {
            const int G = h->steps_per_graph;
            for (int attempt = 0; !h->graph && n >= G && attempt < 2; attempt++) {
                cudaGraph_t g = nullptr;
                const long lt = h->n_launch_total, le = h->n_eval_launch;
                CK(cudaStreamBeginCapture(h->stream, cudaStreamCaptureModeThreadLocal));
                int r = TG_OK;
                for (int q = 0; q < G && r == TG_OK; q++) r = launch_step(h, true, true, 0);
                cudaError_t ce = cudaStreamEndCapture(h->stream, &g);
                h->n_launch_total = lt; h->n_eval_launch = le; h->side_pending = false;
                if (ce == cudaSuccess && r == TG_OK) ce = cudaGraphInstantiate(&h->graph, g, 0);
                if (g) cudaGraphDestroy(g);
                if (ce == cudaSuccess && r == TG_OK) break;
                h->graph = nullptr;
                if (!h->pdl || attempt == 1) { if (r) return r; CK(ce); }
                cudaGetLastError();                  // a driver that cannot capture programmatic edges: same graph with full dependencies
                h->pdl = false;
            }
            bool again = false;
            if (h->graph) {
                for (; k + G <= n; k += G) {
                    CK(cudaGraphLaunch(h->graph, h->stream)); h->n_launch_total += ((h->m.n_learn > 0 || (h->tri_smem && !h->no_split)) ? 4L : 3L) * G; h->n_eval_launch += G;
                    if (h->mom_note && !h->mom_probed) {
                        // first steps of this handle: which background-only tiles fail their polynomial check is known now; without
                        // this look the rest of the call would run with those tiles inside long cheap items (2x slower launches)
                        h->mom_probed = true;
                        CK(cudaStreamSynchronize(h->stream));
                        if (*(volatile int *)h->mom_note) {
                            int r = refresh_items_if_flagged(h); if (r) return r;
                            if (h->u_sp.B != h->s.C) { r = build_units(h, h->u_sp, h->s.C); if (r) return r; }
                            if (!h->graph) { k += G; again = true; break; }
                        }
                    }
                }
            }
            if (!again) break;
        }
        for (; k < n; k++) { int r = launch_step(h, true, true, 0); if (r) return r; }
        { int r = drain_records(h, rec, done, n); if (r) return r; }
        done += n;
    }
    CK(cudaStreamSynchronize(h->stream));
    return TG_OK;
}

extern "C" int tg_step_begin(tg_handle *h, long i)
{
    if (!h) return fail(TG_E_BADARG, "null handle");
    CK(cudaSetDevice(h->device));
    { int r = refresh_items_if_flagged(h); if (r) return r; }
    if (h->u_sp.B != h->s.C) { int r = build_units(h, h->u_sp, h->s.C); if (r) return r; }
    { int r = set_ctrl(h, i, i); if (r) return r; }
    return launch_step(h, true, false, 1);
}
extern "C" int tg_step_end(tg_handle *h, long i)
{
    (void)i;
    if (!h) return fail(TG_E_BADARG, "null handle");
    CK(cudaSetDevice(h->device));
    return launch_step(h, false, true, 1);
}
extern "C" int tg_shard_install_current(tg_handle *h)
{
    if (!h) return fail(TG_E_BADARG, "null handle");
    CK(cudaSetDevice(h->device));
    tg_k_install_current<<<(h->s.C + 127) / 128, 128, 0, h->stream>>>(h->d, h->s, h->sp.t, h->sp.sums);
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize(h->stream));
    return TG_OK;
}
extern "C" int tg_partials_device_ptr(tg_handle *h, void **ptr, int *count)
{
    if (!h || !ptr || !count) return fail(TG_E_BADARG, "null argument");
    *ptr = (void *)h->sp.sums; *count = 2 * h->s.C;
    return TG_OK;
}
extern "C" int tg_sync(tg_handle *h)
{
    if (!h) return fail(TG_E_BADARG, "null handle");
    CK(cudaSetDevice(h->device));
    CK(cudaStreamSynchronize(h->stream));
    return TG_OK;
}

extern "C" int tg_kernel_time(tg_handle *h, int reset, long *launches, double *ms_eval, long *total_launches)
{
    if (!h) return fail(TG_E_BADARG, "null handle");
    CK(cudaSetDevice(h->device));
    { int r = harvest_events(h); if (r) return r; }
    if (launches) *launches = h->n_eval_launch;
    if (ms_eval) *ms_eval = h->ms_eval;
    if (total_launches) *total_launches = h->n_launch_total;
    if (reset == 1) { h->n_eval_launch = 0; h->ms_eval = 0; h->n_launch_total = 0; h->timing = true; }
    if (reset == 2) { h->n_eval_launch = 0; h->ms_eval = 0; h->n_launch_total = 0; h->timing = false; }
    return TG_OK;
}

extern "C" int tg_measure_fp64_peak(int device, double *tflops, double *ms_out)
{
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return fail(TG_E_NODEVICE, "no CUDA device");
    CK(cudaSetDevice(device));
    cudaDeviceProp prop; CK(cudaGetDeviceProperties(&prop, device));
    const int blocks = prop.multiProcessorCount * 8, iters = 4096;
    double *out = nullptr;
    CK(cudaMalloc(&out, (size_t)blocks * 256 * 8));
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e30f;
    for (int rep = 0; rep < 6; rep++) {
        cudaEventRecord(e0);
        tg_k_dfma_peak<<<blocks, 256>>>(out, iters, 0.999999, 1e-9);
        cudaEventRecord(e1);
        CK(cudaEventSynchronize(e1));
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (rep > 0 && ms < best) best = ms;
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(out);
    const double flops = (double)blocks * 256.0 * iters * 64.0 * 2.0;
    if (tflops) *tflops = flops / (best * 1e-3) / 1e12;
    if (ms_out) *ms_out = best;
    return TG_OK;
}

extern "C" int tg_event_record(tg_handle *h, int which)
{
    if (!h || which < 0 || which > 1) return fail(TG_E_BADARG, "tg_event_record: bad argument");
    CK(cudaSetDevice(h->device));
    if (!h->watch[which]) CK(cudaEventCreate(&h->watch[which]));
    CK(cudaEventRecord(h->watch[which], h->stream));
    return TG_OK;
}
extern "C" int tg_event_elapsed_ms(tg_handle *h, double *ms)
{
    if (!h || !ms || !h->watch[0] || !h->watch[1]) return fail(TG_E_BADARG, "tg_event_elapsed_ms: record both events first");
    CK(cudaSetDevice(h->device));
    CK(cudaEventSynchronize(h->watch[1]));
    float f = 0; CK(cudaEventElapsedTime(&f, h->watch[0], h->watch[1]));
    *ms = f;
    return TG_OK;
}
