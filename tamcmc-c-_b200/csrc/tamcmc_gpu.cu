// tamcmc_gpu.cu -- C-ABI (include/tamcmc_gpu.h) over the sm_100a kernels.  Host side only
// allocates, uploads and launches; there is no CPU implementation of any numerical step here.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include <set>
#include <mutex>
#include <limits>
#include <cmath>
#include <algorithm>
#include <dlfcn.h>
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
    double   *pin_in2 = nullptr, *pin_out2 = nullptr;  // second staging slot (tg_eval_async: the host fills one while the GPU works from the other)
    TgEvalArgs *d_args = nullptr;                      // the fused kernel reads its arguments from device memory (tg_load_args)
    TgEvalArgs  h_args;                                // what d_args holds
    bool        args_valid = false;
    // far-field pre-pass (tg_farfield.cuh): coefficients, headers and exact records per (vector, micro-tile); its argument record
    double     *ffc = nullptr; TgFFHdr *ffh = nullptr; TgFFRec *ffr = nullptr;
    TgFFArgs   *d_ffargs = nullptr; TgFFArgs h_ffargs; bool ffargs_valid = false;
    void layout(int b, int Np) {
        B = b; params = in; T = in + (size_t)b * Np;
        logL = out; logPrior = out + b; logPost = out + 2 * (size_t)b; status = (int *)(out + 3 * (size_t)b);
    }
};

struct tg_batch;
struct tg_handle {
    int        device = 0, n_sm = 0;
    cudaStream_t stream = nullptr;
    tg_batch    *self = nullptr;           // the sampler steps of a handle run as a batch of one star (tg_run == tg_batch_run of it)
    long         items_version = 0;        // bumped whenever the work items change (batches rebuild their unit tables)
    bool         no_split = false;         // developer switch (environment TG_NO_SPLIT_PRIOR)
    TgDesc     d{};
    TgMala     m{};
    TgState    s{};
    TgRecDev   rec{};
    double    *x = nullptr, *y = nullptr, *lnx = nullptr, *sig = nullptr;
    double    *mom = nullptr; int mom_sub = 4, mom_run = 16; int *mom_fail = nullptr; long n_mom_tiles = 0;
    double    *ffw = nullptr; double ff_r2 = 30.25;    // far-field path (tg_farfield.cuh): matrix + nodes, (pole distance / half-width)^2 from which a component is far
    TgFFGeo   *ffgeo = nullptr; int ff_stride = 0;    // geometry of the micro-tiles of the spectrum; micro-tiles per vector the buffers hold
    int        ff_mt0 = 0, ff_nmt = 0;                 // micro-tiles the pre-pass prepares: where truncation windows are expected (the heavy items)
    int        heavy_t0 = 0, heavy_t1 = 0;             // window region of the current partition, in sub-tiles
    int        hybrid_pct = 0, hybrid_group = 2;
    int       *mom_note = nullptr;                      // pinned, mapped: the kernel stores 1 when it flags a tile for the first time          // Chebyshev moments of the data over aligned tiles (tg_eval.cuh), NULL = off
    double    *pri = nullptr; int *sw = nullptr;
    int       *index_to_relax = nullptr; double *Tcoefs = nullptr;
    EvalBuf    ev;               // tg_eval batches
    EvalBuf    ev2;              // tg_eval_async, slot 1: its own tables and results on its own stream, so that the copy-in, prologue and pre-pass of one
    cudaStream_t stream2 = nullptr;   // submission run beside the fused kernel of the other
    EvalBuf    sp;               // sampler's proposed points (cap == C)
    int        nitems = 0, max_items = 0, light_sub = 8, heavy_sub = 2, chains_per_block = 0, heavy_group = 2, use_graph = 1, steps_per_graph = 16;
    int        smem_matrix = 0, tri_smem = 0; size_t propose_smem = 0, spec_smem = 0; int fused_spec = 0;
    double    *stage_mu = nullptr, *stage_cov = nullptr;   // tg_get_state staging ([C][Nv], [C][Nv][Nv])
    int        list_cap = 4, max_cpb = 1;
    struct UnitTable { int4 *dev = nullptr; int n = 0, B = -1, cap = 0; };
    UnitTable  u_sp, u_ev;                 // work units of the fused kernel in launch order: sampler (B = C), tg_eval batches
    std::vector<int4> items_host;          // bin partition: (first bin, sub-tiles, weight = windows crossing, heavy?)
    bool       no_order = false;
    cudaGraphExec_t ev_graph = nullptr; int ev_graph_B = 0;   // tg_eval: copy in, prologue, fused kernel, reduce+finalize, copy out
    cudaGraphExec_t ev_graph2 = nullptr;                       // the same through the second staging slot
    cudaEvent_t ev_done[2] = {nullptr, nullptr}; int slot_B[2] = {0, 0};   // tg_eval_async / tg_eval_wait
    // eval-kernel timing
    bool       timing = false;
    std::vector<cudaEvent_t> ev_pool; size_t ev_used = 0;
    long       n_eval_launch = 0, n_launch_total = 0; double ms_eval = 0;
    std::vector<void *> allocs;
    cudaEvent_t watch[2] = {nullptr, nullptr};
    void       *nccl_comm = nullptr;       // bin-sharded variant: ncclComm_t of the shard group (tg_shard_attach)
    // bin-sharded variant over peer memory (tg_shard_ipc_*): this rank's exchange buffer, the peers' buffers opened through CUDA IPC
    bool        xchg_on = false; TgXchg xchg{}; void *xbuf = nullptr; size_t xbuf_bytes = 0; void *xpeer[TG_XCHG_MAXR] = {}; int *xerr_host = nullptr;
};

// A launch group: the sampler steps of several stars (handles on the same GPU) as ONE propose / fused / finish launch per
// step -- cfg5 of BASELINE.json (several tempered ensembles of different spectra per GPU).  The stars stay independent
// (own spectrum, parameters, priors, Philox key, iteration counter, records); only the grids are shared: the unit table of
// the fused kernel spans all stars, costliest first.  A handle on its own is a batch of one.
struct tg_batch {
    std::vector<tg_handle *> hs;
    int          device = 0, Ctot = 0;
    cudaStream_t stream = nullptr;         // == hs[0]->stream
    cudaStream_t side = nullptr;           // the adaptation candidates of a step run here, beside the proposal and the fused kernel
    cudaEvent_t  ev_fork = nullptr, ev_join = nullptr;
    bool         side_pending = false;     // a join is owed before the next finish kernel
    bool         pdl = true;               // programmatic dependent launch between the step kernels (environment TG_PDL=0 turns it off)
    bool         no_side = false;          // developer switch (environment TG_NO_SIDE_STREAM)
    int          use_graph = 1, steps_per_graph = 16, presummed = 0;
    // per-star argument records of the three step kernels + the chain map, host copies and device arrays
    std::vector<TgEvalArgs>    h_ev;   TgEvalArgs    *d_ev = nullptr;
    std::vector<TgProposeArgs> h_pm, h_ps; TgProposeArgs *d_pm = nullptr, *d_ps = nullptr;
    std::vector<TgFinishArgs>  h_fin;  TgFinishArgs  *d_fin = nullptr;
    std::vector<TgFFArgs>      h_ff;   TgFFArgs      *d_ff = nullptr;   // far-field pre-pass
    int          ff_cpc = 0, ff_ppc = 1; size_t ff_smem = 0;
    int2        *d_cmap = nullptr;
    bool         args_valid = false;
    cudaEvent_t  ev_run[2] = {nullptr, nullptr};   // device-side stopwatch around the steps of the last tg_run / tg_batch_run call
    int          side_rows = 0; size_t propose_smem = 0, eval_smem = 0;
    tg_handle::UnitTable units;            // work units of the fused kernel over all stars, costliest first
    std::vector<long> units_version;       // items_version of every star the table was built from
    cudaGraphExec_t graph = nullptr;       // steps_per_graph steps
    cudaGraphExec_t graph_small[2] = {nullptr, nullptr};   // 4 steps and 1 step: the remainder of a call is replayed too, never launched kernel by kernel
    bool         probed = false;           // the look at the flagged background tiles after the first graph of a long run
    // eval-kernel timing (and, beside it, the far-field pre-pass)
    bool         timing = false;
    std::vector<cudaEvent_t> ev_pool; size_t ev_used = 0;
    std::vector<cudaEvent_t> ffev_pool; size_t ffev_used = 0; double ms_ff = 0;
    long         n_eval_launch = 0, n_launch_total = 0; double ms_eval = 0;
    std::vector<void *> allocs;
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

// cudaFuncAttributeMaxDynamicSharedMemorySize is a property of the function, not of a handle: only ever raised
static int raise_dyn_smem(const void *func, size_t bytes)
{
    static std::vector<std::pair<std::pair<const void *, int>, size_t>> seen;   // (function, device) -> bytes
    int dev = 0; CK(cudaGetDevice(&dev));
    const std::pair<const void *, int> key(func, dev);
    for (auto &e : seen) if (e.first == key) {
        if (bytes <= e.second) return TG_OK;
        CK(cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
        e.second = bytes; return TG_OK;
    }
    CK(cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    seen.push_back({key, bytes});
    return TG_OK;
}

// handles that exist (a launch group gives its stars their own unit shapes back when it is destroyed -- but only those that are
// still there: at interpreter exit a binding may destroy the stars first)
static std::mutex g_live_mu;
static std::set<const void *> g_live;
static bool handle_alive(const void *h) { std::lock_guard<std::mutex> lk(g_live_mu); return g_live.count(h) != 0; }
static long g_dbg_count[4] = {0, 0, 0, 0};      // developer counters: work-item partitions built, step graphs captured, probe syncs, argument uploads
extern "C" int tg_dbg_counters(long *out) { for (int k = 0; k < 4; k++) out[k] = g_dbg_count[k]; return 0; }
extern "C" const char *tg_last_error(void) { return g_err.c_str(); }
extern "C" int tg_version(void) { return 100; }
#ifdef TG_PROFILE
extern "C" int tg_dbg_ctas(long long *out) { return cudaMemcpyFromSymbol(out, tg_dbg_cta, sizeof(long long) * TG_DBG_CTAS * 6) == cudaSuccess ? 0 : -2; }
extern "C" int tg_dbg_ffstats(long long *out) { return cudaMemcpyFromSymbol(out, tg_dbg_ff, sizeof(long long) * TG_DBG_CTAS * 4) == cudaSuccess ? 0 : -2; }
extern "C" int tg_dbg_timeline(long long *out) { return cudaMemcpyFromSymbol(out, tg_dbg_tl, sizeof(long long) * 8) == cudaSuccess ? 0 : -2; }
extern "C" int tg_dbg_specs(long long *out) { return cudaMemcpyFromSymbol(out, tg_dbg_spec, sizeof(long long) * 16) == cudaSuccess ? 0 : -2; }
extern "C" int tg_dbg_clocks(long long *out) { return cudaMemcpyFromSymbol(out, tg_dbg_clk, sizeof(long long) * 32) == cudaSuccess ? 0 : -2; }
#endif

static void drop_graphs(tg_handle *h)
{
    if (h->ev_graph) { cudaGraphExecDestroy(h->ev_graph); h->ev_graph = nullptr; }
    if (h->ev_graph2) { cudaGraphExecDestroy(h->ev_graph2); h->ev_graph2 = nullptr; }
    h->items_version++;                                   // batches holding this star re-capture their step graph
}

static int alloc_evalbuf(tg_handle *h, EvalBuf &b, int B)
{
    if (B <= b.cap) return TG_OK;
    const size_t nc = (size_t)B * h->d.ncomp_max;
    b.cap = B;
    if (h->ev_graph) { cudaGraphExecDestroy(h->ev_graph); h->ev_graph = nullptr; }     // the buffers move
    if (h->ev_graph2) { cudaGraphExecDestroy(h->ev_graph2); h->ev_graph2 = nullptr; }
    DA(b.t.nu, nc); DA(b.t.h, nc); DA(b.t.q, nc); DA(b.t.ifc, nc); DA(b.t.t2sq, nc); DA(b.t.lo, nc); DA(b.t.hi, nc);
    DA(b.t.ncomp, B); DA(b.t.wlo, B); DA(b.t.whi, B); DA(b.t.noise, (size_t)B * TG_NOISE_STRIDE); DA(b.t.status, B); DA(b.t.pstatus, B); DA(b.t.work, B); DA(b.t.logprior, B);
    DA(b.in, (size_t)B * (h->d.Np + 1)); DA(b.out, (size_t)B * 4);
    DA(b.partials, (size_t)B * h->max_items); DA(b.sums, B);
    if (h->ffw) {
        const size_t n = (size_t)B * h->ff_stride;
        DA(b.ffc, n * TG_FF_N); DA(b.ffh, n); DA(b.ffr, n * TG_FF_E);
        b.ffargs_valid = false;
    }
    if (b.pin_in) cudaFreeHost(b.pin_in);
    if (b.pin_out) cudaFreeHost(b.pin_out);
    if (b.pin_in2) cudaFreeHost(b.pin_in2);
    if (b.pin_out2) cudaFreeHost(b.pin_out2);
    CK(cudaHostAlloc((void **)&b.pin_in, (size_t)B * (h->d.Np + 1) * 8, cudaHostAllocDefault));
    CK(cudaHostAlloc((void **)&b.pin_out, (size_t)B * 4 * 8, cudaHostAllocDefault));
    CK(cudaHostAlloc((void **)&b.pin_in2, (size_t)B * (h->d.Np + 1) * 8, cudaHostAllocDefault));
    CK(cudaHostAlloc((void **)&b.pin_out2, (size_t)B * 4 * 8, cudaHostAllocDefault));
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
template <class... Arg>
static cudaError_t launch_k(void (*kern)(Arg...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, bool pdl, Arg... arg)
{
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kern, arg...);
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

// Far-field pre-pass of the B points held in `b`: view for the fused kernel, argument record, launch geometry
static TgFFView ff_view(const tg_handle *h, const EvalBuf &b)
{
    TgFFView v; memset(&v, 0, sizeof(v));
    if (!h->ffw || !b.ffc || h->ff_nmt <= 0) return v;
    v.c = b.ffc; v.hdr = b.ffh; v.rec = b.ffr; v.geo = h->ffgeo; v.stride = h->ff_stride; v.mt0 = h->ff_mt0; v.nmt = h->ff_nmt;
    return v;
}
static TgFFArgs ff_args(const tg_handle *h, const EvalBuf &b, int B)
{
    TgFFArgs a; memset(&a, 0, sizeof(a));
    a.d = h->d; a.t = b.t; a.ffw = h->ffw; a.r2 = h->ff_r2; a.rs = std::sqrt(h->ff_r2); a.c = b.ffc; a.hdr = b.ffh; a.rec = b.ffr; a.geo = h->ffgeo;
    a.stride = h->ff_stride; a.mt0 = h->ff_mt0; a.nmt = (h->ffw && b.ffc) ? h->ff_nmt : 0; a.B = B;
    return a;
}
// launch geometry of the pre-pass: pairs of sub-tiles per CTA (the chain's table is staged once per CTA) so that the grid stays at
// a few waves of 4 CTAs per SM, and CTAs per chain
static int ff_pairs_per_cta(int nmt, int chains, int n_sm)
{
    static const int forced = getenv("TG_FF_PPC") ? atoi(getenv("TG_FF_PPC")) : 0;      // developer knob
    if (forced > 0) return forced;
    const long pairs = (long)((nmt + TG_FF_MT - 1) / TG_FF_MT) * chains;
    return (int)std::max<long>(1, std::min<long>(4, (pairs + (long)n_sm * 2) / ((long)n_sm * 4)));     // about one wave of 4 CTAs per SM, 4 pairs at most (measured: 2 at 10 chains, 3-4 at 80)
}
static int ff_ctas_per_chain(int nmt, int ppc) { return ((nmt + TG_FF_MT - 1) / TG_FF_MT + ppc - 1) / ppc; }
static size_t ff_smem(const tg_handle *h) { return TG_FF_SMEM(std::max(h->d.ncomp_max, 1)); }

// The argument record of the fused kernel for the points held in `b` -> device memory (uploaded only when it changed; never
// called with a change pending inside a stream capture: the callers bring it up to date before they begin one).
static int sync_eval_args(tg_handle *h, EvalBuf &b, double *model, int B)
{
    TgEvalArgs ea; memset(&ea, 0, sizeof(ea));
    ea.d = h->d; ea.t = b.t; ea.x = h->x; ea.y = h->y; ea.lnx = h->lnx; ea.sig = h->sig; ea.partials = b.partials; ea.model = model;
    ea.B = B; ea.nitems = h->nitems; ea.max_group = h->chains_per_block; ea.list_cap = h->list_cap;
    ea.mom = h->mom; ea.mom_sub = h->mom_sub; ea.mom_fail = h->mom_fail; ea.mom_note = h->mom_note;
    if (!model) ea.ff = ff_view(h, b);
    if (h->ffw) {
        const TgFFArgs fa = ff_args(h, b, B);
        if (!(b.d_ffargs && b.ffargs_valid && memcmp(&fa, &b.h_ffargs, sizeof(fa)) == 0)) {
            if (!b.d_ffargs) DA(b.d_ffargs, 1);
            memcpy(&b.h_ffargs, &fa, sizeof(fa));
            CK(cudaMemcpyAsync(b.d_ffargs, &b.h_ffargs, sizeof(fa), cudaMemcpyHostToDevice, h->stream));
            b.ffargs_valid = true;
        }
    }
    if (b.d_args && b.args_valid && memcmp(&ea, &b.h_args, sizeof(ea)) == 0) return TG_OK;
    if (!b.d_args) DA(b.d_args, 1);
    memcpy(&b.h_args, &ea, sizeof(ea));
    CK(cudaMemcpyAsync(b.d_args, &b.h_args, sizeof(ea), cudaMemcpyHostToDevice, h->stream));
    b.args_valid = true;
    return TG_OK;
}

// prologue (optional) + fused kernel + per-chain reduction for B points of ONE handle (tg_eval, create / restore)
static int launch_eval_kernels(tg_handle *h, EvalBuf &b, const double *params_dev, int B, double *model, bool do_prologue, bool do_reduce, bool pdl = false, cudaStream_t st = nullptr)
{
    if (!st) st = h->stream;
    if (do_prologue) {
        TgProArgs pa{h->d, b.t, params_dev, h->pri, h->sw};
        tg_k_prologue<<<B, TG_PRO_THREADS, TG_PRO_SMEM(h->d.Np), st>>>(pa);
        h->n_launch_total++;
    }
    tg_handle::UnitTable &u = (&b == &h->sp) ? h->u_sp : h->u_ev;
    if (B != u.B) { int r = build_units(h, u, B); if (r) return r; }     // never inside a graph capture
    { int r = sync_eval_args(h, b, model, B); if (r) return r; }
    const size_t esm = TG_EVAL_SMEM(h->chains_per_block, h->list_cap);
    dim3 grid(u.n);
    if (h->ffw && !model && h->ff_nmt > 0 && b.ffc) {   // far-field pre-pass of the B points (tg_farfield.cuh)
        const int ppc = ff_pairs_per_cta(h->ff_nmt, B, h->n_sm), cpc = ff_ctas_per_chain(h->ff_nmt, ppc);
        CK(launch_k(tg_k_ff, dim3(cpc, B), dim3(256), ff_smem(h), st, pdl && !h->timing, (const TgFFArgs *)b.d_ffargs, (const int2 *)nullptr, ppc));
        h->n_launch_total++;
    }
    if (h->timing) {
        if (h->ev_used >= 8192) { int r = harvest_events(h); if (r) return r; }
        cudaEventRecord(next_event(h), st);
    }
    CK(launch_k(tg_k_eval, grid, dim3(TG_THREADS), esm, st, pdl && !h->timing, (const TgEvalArgs *)b.d_args, (const int4 *)u.dev));
    if (h->timing) cudaEventRecord(next_event(h), st);
    h->n_eval_launch++; h->n_launch_total++;
    if (do_reduce) { tg_k_reduce<<<B, 256, 0, st>>>(b.partials, h->nitems, b.sums); h->n_launch_total++; }
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
        // hysteresis: the window region of the previous partition is kept while it still covers the windows and is not more than
        // eight sub-tiles too wide on either side (the chains' windows breathe by a tile or two between calls; a new partition
        // means new unit tables and new step graphs)
        if (uhi > ulo && h->heavy_t1 > h->heavy_t0 && h->heavy_t0 * SUB <= ulo && h->heavy_t1 * SUB >= uhi &&
            t0 - h->heavy_t0 <= 8 && h->heavy_t1 - t1 <= 8 && h->heavy_t1 <= nt) { t0 = h->heavy_t0; t1 = h->heavy_t1; }
        h->heavy_t0 = t0; h->heavy_t1 = t1;
        // the far-field pre-pass prepares the micro-tiles of the window region (whole pairs of sub-tiles inside the spectrum)
        { const int mt0 = t0 * (SUB / TG_FF_BINS), nmt = std::max(0, std::min(t1 * (SUB / TG_FF_BINS), h->ff_stride) - mt0);
          if (h->ffw && (mt0 != h->ff_mt0 || nmt != h->ff_nmt)) { h->ff_mt0 = mt0; h->ff_nmt = nmt; h->sp.args_valid = h->ev.args_valid = false; h->sp.ffargs_valid = h->ev.ffargs_valid = false; drop_graphs(h); } }
        std::vector<int> diff(nt + 1, 0), w(nt, 0);
        for (int k = 0; k < ncomp && k < nc; k++) {
            if (hi[k] <= lo[k]) continue;
            const int a0 = std::max(0, lo[k] / SUB), a1 = std::min(nt - 1, (hi[k] - 1) / SUB);
            if (a1 >= a0) { diff[a0]++; diff[a1 + 1]--; }
        }
        for (int t = 0, c = 0; t < nt; t++) { c += diff[t]; w[t] = c; }
        // bin-by-bin background runs: shorter when the window region is cheap (far-field path), or they become the tail of the launch
        const int L = h->ffw ? std::max(1, std::min(h->light_sub, 4)) : std::max(1, h->light_sub);
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
        // (item.z = window crossings, a cost estimate only.  On the far-field path the crossings hardly matter for the cost, and a
        // value that follows every breath of the windows would make a new partition -- new unit tables, new step graphs -- of
        // every long call: it stays 0 there.)
        for (int t = t0; t < t1; t += Hs) { int wm = 0; if (!h->ffw) for (int q = t; q < std::min(t + Hs, t1); q++) wm = std::max(wm, w[q]); items.push_back(make_int4(t * SUB, std::min(Hs, t1 - t), wm, 1)); }
        light_region(t1, nt);
    }
    bool same = items.size() == h->items_host.size();
    for (size_t k = 0; same && k < items.size(); k++) same = items[k].x == h->items_host[k].x && items[k].y == h->items_host[k].y && items[k].z == h->items_host[k].z && items[k].w == h->items_host[k].w;
    if (same) return TG_OK;
    h->items_host = items;
    g_dbg_count[0]++;
    h->nitems = (int)items.size();
    h->u_sp.B = h->u_ev.B = -1;                     // units are rebuilt at the next launch
    drop_graphs(h);                                 // grid size changed
    return TG_OK;
}

// Units = items x chain groups: few chains per unit where windows are expected (an item there costs
// 5-10x a noise-only sub-tile), up to chains_per_block elsewhere.  Appends the units of star `star` (B points) with their
// estimated costs; host side only.
static void make_units(tg_handle *h, int B, int star, std::vector<int4> &units, std::vector<double> &cost)
{
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
        int gmax = (it.w == 1 && !h->no_order) ? (big[k] ? std::max(1, std::min(h->hybrid_group, gl)) : gh) : gl;
        if (it.w == 0 && h->ffw && !h->no_order) gmax = std::min(gmax, 2);
        const int groups = (B + gmax - 1) / gmax, g = (B + groups - 1) / groups;      // balanced
        for (int c0 = 0; c0 < B; c0 += g) {
            const int G = std::min(g, B - c0);
            units.push_back(make_int4(it.x, it.y | (G << 16), c0, (int)k | (star << TG_UNIT_STAR_SHIFT)));
            const double per_sub = (it.w == 1 && h->ffw) ? 22.0 + 0.5 * it.z : 20.0 + 6.25 * it.z;      // far-field path: the window crossings hardly matter
            cost.push_back(it.w == 2 ? (double)G * (it.y / std::max(1, h->mom_sub)) * 8.0 + 15.0 : (double)G * it.y * per_sub + 15.0);
        }
    }
}

// costliest first (LPT scheduling) unless the tile order is forced, then to the device
static int upload_units(std::vector<void *> &allocs, cudaStream_t stream, tg_handle::UnitTable &u, std::vector<int4> &units, const std::vector<double> &cost, bool keep_order)
{
    if (!keep_order) {
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
        allocs.push_back(q);                        // the old (smaller) table is released with its owner
        u.dev = (int4 *)q; u.cap = (int)units.size();
    }
    CK(cudaStreamSynchronize(stream));
    CK(cudaMemcpy(u.dev, units.data(), units.size() * sizeof(int4), cudaMemcpyHostToDevice));
    u.n = (int)units.size();
    return TG_OK;
}

static int build_units(tg_handle *h, tg_handle::UnitTable &u, int B)
{
    std::vector<int4> units;
    std::vector<double> cost;
    make_units(h, B, 0, units, cost);
    int r = upload_units(h->allocs, h->stream, u, units, cost, h->no_order); if (r) return r;
    u.B = B;
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

// NCCL for the bin-sharded variant: the library already in the process (the one torch ships) or the first libnccl.so.2 found;
// only four entry points are bound (nccl.h: ncclGetUniqueId, ncclCommInitRank, ncclAllReduce, ncclCommDestroy)
struct tg_nccl_id { char internal[128]; };                       // == ncclUniqueId
typedef int (*tg_nccl_allreduce_t)(const void *, void *, size_t, int, int, void *, cudaStream_t);
typedef int (*tg_nccl_getid_t)(tg_nccl_id *);
typedef int (*tg_nccl_init_t)(void **, int, tg_nccl_id, int);
typedef int (*tg_nccl_destroy_t)(void *);
static tg_nccl_allreduce_t g_nccl_allreduce = nullptr;
static tg_nccl_getid_t g_nccl_getid = nullptr;
static tg_nccl_init_t g_nccl_init = nullptr;
static tg_nccl_destroy_t g_nccl_destroy = nullptr;
static int load_nccl()
{
    if (g_nccl_allreduce) return TG_OK;
    void *lib = nullptr;
    if (const char *e = getenv("TG_NCCL_LIB")) lib = dlopen(e, RTLD_NOW | RTLD_GLOBAL);
    if (!lib) lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);                  // already loaded (torch.distributed)
    if (!lib) lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!lib) lib = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!lib) return fail(TG_E_UNSUPPORTED, std::string("bin-sharded variant: no libnccl.so.2 (set TG_NCCL_LIB): ") + (dlerror() ? dlerror() : ""));
    g_nccl_getid = (tg_nccl_getid_t)dlsym(lib, "ncclGetUniqueId");
    g_nccl_init = (tg_nccl_init_t)dlsym(lib, "ncclCommInitRank");
    g_nccl_destroy = (tg_nccl_destroy_t)dlsym(lib, "ncclCommDestroy");
    tg_nccl_allreduce_t ar = (tg_nccl_allreduce_t)dlsym(lib, "ncclAllReduce");
    if (!g_nccl_getid || !g_nccl_init || !g_nccl_destroy || !ar) return fail(TG_E_UNSUPPORTED, "bin-sharded variant: libnccl lacks an entry point");
    g_nccl_allreduce = ar;
    return TG_OK;
}

static int shard_reduce_and_install(tg_handle *h)
{
    if (h->xchg_on) {
        tg_k_xchg<<<1, 256, 0, h->stream>>>(h->xchg, h->sp.sums, h->s.C);
        tg_k_install_current<<<(h->s.C + 127) / 128, 128, 0, h->stream>>>(h->d, h->s, h->sp.t, h->sp.sums);
        CK(cudaGetLastError());
        CK(cudaStreamSynchronize(h->stream));
        if (h->xerr_host && *(volatile int *)h->xerr_host) return fail(TG_E_CUDA, "bin-sharded exchange: a peer did not answer");
        return TG_OK;
    }
    const int rc = g_nccl_allreduce(h->sp.sums, h->sp.sums, (size_t)2 * h->s.C, /*ncclDouble*/ 8, /*ncclSum*/ 0, h->nccl_comm, h->stream);
    if (rc != 0) return fail(TG_E_CUDA, "ncclAllReduce failed (" + std::to_string(rc) + ")");
    tg_k_install_current<<<(h->s.C + 127) / 128, 128, 0, h->stream>>>(h->d, h->s, h->sp.t, h->sp.sums);
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize(h->stream));
    return TG_OK;
}
static int batch_refresh(tg_batch *b);
static void batch_drop_graph(tg_batch *b);

extern "C" int tg_set_tuning(tg_handle *h, int chains_per_block, int heavy_group, int light_subtiles, int use_graph)
{
    if (!h) return fail(TG_E_BADARG, "null handle");
    CK(cudaSetDevice(h->device));
    if (chains_per_block > 0) { h->chains_per_block = std::min(chains_per_block, h->max_cpb); h->u_sp.B = h->u_ev.B = -1; }
    if (heavy_group > 0) { h->heavy_group = heavy_group; h->u_sp.B = h->u_ev.B = -1; }
    if (use_graph >= 0) { h->use_graph = use_graph; if (h->self) { h->self->use_graph = use_graph; batch_drop_graph(h->self); } }
    drop_graphs(h);
    if (light_subtiles > 0 && light_subtiles != h->light_sub) {
        h->light_sub = std::min(light_subtiles, 4096);
        int r = build_items(h, h->sp, h->s.C); if (r) return r;
    }
    return h->self ? batch_refresh(h->self) : TG_OK;
}

static int create_impl(const tg_problem *p, tg_handle *h);
static int batch_create(tg_handle **hs, int n, tg_batch **out);
static int batch_destroy(tg_batch *b);
extern "C" int tg_event_record(tg_handle *h, int which);

extern "C" int tg_destroy(tg_handle *h);

extern "C" int tg_create(const tg_problem *p, tg_handle **out)
{
    if (!p || !out) return fail(TG_E_BADARG, "null argument");
    *out = nullptr;
    tg_handle *h = new tg_handle();
    const int r = create_impl(p, h);
    if (r != TG_OK) { const std::string keep = g_err; tg_destroy(h); g_err = keep; return r; }
    { std::lock_guard<std::mutex> lk(g_live_mu); g_live.insert(h); }
    *out = h;
    return TG_OK;
}

// Shape of the work units by the number of chains a LAUNCH carries (the chains of a handle; of every star of a launch group):
// scheduling only, results never depend on it.  Developer knobs override.
static void schedule_defaults(tg_handle *h, int C)
{
    const TgDesc &d = h->d;
    // where truncation windows lie: items of two sub-tiles (taken 8 bins per thread); one chain per unit for a single
    // ensemble (more, smaller units: better tail of the grid), three when many chains fill the GPU anyway (measured with the
    // background-only tiles on the moment path: one chain per unit is best up to ~100 chains, 3.9 vs 4.0 us per chain at 80)
    h->heavy_sub = 2;
    h->heavy_group = (C >= 128) ? 3 : 1;
    // a spectrum so short that pairs of sub-tiles would not even fill the GPU once (S-local: 10k bins, 10 chains): single
    // sub-tiles, twice the units (25.7 vs 29.9 us per step there)
    const bool short_spectrum = (long)((d.Nx + TG_THREADS * 4 - 1) / (TG_THREADS * 4)) * C < 4L * 296;
    if (short_spectrum) h->heavy_sub = 1;
    // the costliest window-region items take three chains per unit (set-up and list building shared), the rest one chain
    // (small units for the tail): 70 % of them at 10 chains (49.0 -> 48.1 us), all of them from 32 chains on (-7 %)
    h->hybrid_group = (C < 64) ? 3 : ((C < 128) ? 4 : 5);
    h->hybrid_pct = short_spectrum ? 0 : ((C < 16) ? 70 : ((C < 32) ? 80 : 100));      // short spectra: as many units as possible
    if (h->ffw && !short_spectrum) {
        // far-field path: a window-region unit is a run of pairs of sub-tiles x a few chains (per pair and chain one staged
        // interpolant + the exact records; no lists); measured on cfg3: 4 sub-tiles x 2 chains at 10 chains (28.8 us against
        // 30-36 for the neighbours), 8 sub-tiles x 3 chains at 80 chains (1.45 us per chain)
        h->heavy_sub = (C < 32) ? 4 : 8;
        h->hybrid_pct = 100;
        h->hybrid_group = (C < 32) ? 2 : 4;          // (8 x 4 against 8 x 3 at 8 stars x 10 chains: 246 against 254 us per step)
    }
    // moment tiles per work item: a run of 16 tiles x 5 chains is the longest unit of a launch (23 us: ten rounds of one
    // (tile, chain) per warp) -- with few chains, where the launch is one wave and a bit, shorter runs end it earlier
    // (live step at 10 chains 64.4 -> 62.7 us); with many chains the longer runs amortise their set-up (measured at 80 chains)
    h->mom_run = (C < 32) ? 4 : 16;
    if (const char *e = getenv("TG_MOM_RUN")) h->mom_run = std::max(1, atoi(e));
    if (const char *e = getenv("TG_HYBRID_PCT")) h->hybrid_pct = atoi(e);
    if (const char *e = getenv("TG_HYBRID_GROUP")) h->hybrid_group = std::max(1, atoi(e));
    if (const char *e = getenv("TG_HEAVY_SUB")) h->heavy_sub = std::max(1, atoi(e));
    if (const char *e = getenv("TG_HEAVY_GROUP")) h->heavy_group = std::max(1, atoi(e));
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
    h->no_split = getenv("TG_NO_SPLIT_PRIOR") != nullptr;

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
        // moment tiles per work item: a run of 16 tiles x 5 chains is the longest unit of a launch (23 us: ten rounds of one
        // (tile, chain) per warp) -- with few chains, where the launch is one wave and a bit, shorter runs end it earlier
        // (live step at 10 chains 64.4 -> 62.7 us); with many chains the longer runs amortise their set-up (measured at 80 chains)
        // (the run length itself: schedule_defaults)
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
    {   // far-field path under the truncation windows (TG_FARFIELD=0 turns it off): needs abscissae that increase
        const char *e = getenv("TG_FARFIELD");
        bool inc = true;
        for (size_t i = 1; i < Nx && inc; i++) inc = p->x[i] > p->x[i - 1];
        // (not for a spectrum so short that the launch is a fraction of a wave anyway -- S-local, 10 k bins x 10 chains: there the
        //  pre-pass is one more latency in the step, 31.9 against 27.2 us; TG_FARFIELD=1 forces it)
        const bool short0 = (long)((Nx + TG_THREADS * 4 - 1) / (TG_THREADS * 4)) * p->Nchains < 4L * 296 && !(e && atoi(e) == 1);
        if (!(e && atoi(e) == 0) && inc && p->model_case >= 2 && !(p->flags & TG_F_NO_TILE_ORDER) && !short0) {
            if (const char *e2 = getenv("TG_FF_R2")) h->ff_r2 = std::max(4.0, atof(e2));
            std::vector<double> w((size_t)TG_FF_HDR);
            const long double pi = 3.141592653589793238462643383279502884L;
            for (int k = 0; k < TG_FF_N; k++) for (int i = 0; i < TG_FF_N; i++)
                w[(size_t)k * TG_FF_N + i] = (double)((k == 0 ? 1.0L : 2.0L) / TG_FF_N * cosl(pi * k * (i + 0.5L) / TG_FF_N));
            for (int i = 0; i < TG_FF_N; i++) w[(size_t)TG_FF_N * TG_FF_N + i] = (double)cosl(pi * (i + 0.5L) / TG_FF_N);
            {   // T_j(t) = sum_k M[k][j] t^k by the three-term recurrence, in integers (|M| <= 92160 at degree 15: exact in double)
                long long T[TG_FF_N][TG_FF_N] = {};
                T[0][0] = 1; T[1][1] = 1;
                for (int j = 2; j < TG_FF_N; j++) for (int k = 0; k <= j; k++) T[j][k] = (k > 0 ? 2 * T[j - 1][k - 1] : 0) - T[j - 2][k];
                double *M = w.data() + (size_t)TG_FF_N * TG_FF_N + TG_FF_N;
                for (int k = 0; k < TG_FF_N; k++) for (int j = 0; j < TG_FF_N; j++) M[(size_t)k * TG_FF_N + j] = (double)T[j][k];
            }
            DA(h->ffw, w.size());
            CK(cudaMemcpy(h->ffw, w.data(), w.size() * 8, cudaMemcpyHostToDevice));
            h->ff_stride = (int)(Nx / TG_FF_BINS);                  // whole micro-tiles of the spectrum
            if (h->ff_stride < TG_FF_PAIR_MT * 4) { h->ffw = nullptr; h->ff_stride = 0; }
            else {
                DA(h->ffgeo, (size_t)h->ff_stride);
                tg_k_ffgeo<<<(h->ff_stride + 255) / 256, 256, 0, h->stream>>>(h->x, (int)Nx, h->ffgeo, h->ff_stride);
            }
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
        size_t budget = std::min<size_t>((size_t)maxdyn - 2048, 96 * 1024);              // <= 96 KB (+ 14 KB static): two CTAs per SM stay resident
        if (const char *e = getenv("TG_SMEM_BUDGET_KB")) budget = std::min<size_t>(budget, (size_t)std::max(16, atoi(e)) * 1024);   // developer knob (occupancy experiments)
        int m = TG_MAXG;
        while (m > 1 && TG_EVAL_SMEM(m, h->list_cap) > budget) m--;
        if (TG_EVAL_SMEM(m, h->list_cap) > (size_t)maxdyn - 2048) return fail(TG_E_BADARG, "too many Lorentzian components for the shared-memory list of the fused kernel");
        h->max_cpb = m;
        const size_t need = TG_EVAL_SMEM(m, h->list_cap);
        { int r = raise_dyn_smem((const void *)tg_k_eval, need); if (r) return r; }
        if (h->ffw && ff_smem(h) > 40 * 1024) { int r = raise_dyn_smem((const void *)tg_k_ff, ff_smem(h)); if (r) return r; }
        const int groups = (C + m - 1) / m;                                               // balanced chain groups of <= max_cpb
        h->chains_per_block = (C + groups - 1) / groups;
        schedule_defaults(h, C);
        if (const char *e = getenv("TG_GRAPH_STEPS")) h->steps_per_graph = std::max(1, std::min(256, atoi(e)));
    }

    // ---- MALA configuration ----
    TgMala &m = h->m;
    m.c0 = p->c0; m.eps1 = p->epsilon1; m.eps2 = p->epsilon2; m.A1 = p->A1; m.target = p->target_acceptance;
    m.n_learn = p->n_learn;
    for (int l = 0; l < p->n_learn; l++) { m.Nt_learn[l] = p->Nt_learn[l]; m.periods[l] = p->periods_learn[l]; }
    if (p->n_learn > 0) m.Nt_learn[p->n_learn] = p->Nt_learn[p->n_learn];
    m.dN_mixing = p->dN_mixing; m.seed = p->seed; m.flags = p->flags;
    m.chain0 = p->philox_chain0; m.ens0 = p->philox_ensemble0;

    // ---- sampler state (MALA::MALA + init_proposal, Model_def ctor) ----
    TgState &s = h->s;
    s.C = C; s.Ce = Ce; s.E = C / Ce; s.Nv = d.Nv; s.Np = d.Np;
    const size_t Nv = d.Nv, Np = d.Np;
    DA(s.vars_cur, C * Nv); DA(s.vars_prop, C * Nv); DA(s.params_cur, C * Np);
    s.params_prop = h->sp.params;
    DA(s.logL_cur, C); DA(s.logPrior_cur, C); DA(s.logPost_cur, C); DA(s.Pmove, C); DA(s.moved, C);
    DA(s.mu, 3 * C * Nv); DA(s.sigma, C); DA(s.cov, 3 * C * Nv * Nv); DA(s.chol, 3 * C * Nv * Nv); DA(s.live, C); DA(s.chol_fail, 3 * C); DA(s.sticky_status, C);
    DA(h->stage_mu, C * Nv); DA(h->stage_cov, C * Nv * Nv);
    DA(s.ctrl, 1); DA(s.ens_ticket, s.E);
    s.acc_sigma = s.acc_mu = s.acc_cov = s.acc_vars = nullptr; s.acc_count = nullptr;
    if (p->flags & TG_F_RESTORE_MEANS) {
        DA(s.acc_sigma, C); DA(s.acc_mu, C * Nv); DA(s.acc_cov, C * Nv * Nv); DA(s.acc_vars, C * Nv); DA(s.acc_count, 1);
        CK(cudaMemsetAsync(s.acc_sigma, 0, C * 8, h->stream)); CK(cudaMemsetAsync(s.acc_mu, 0, C * Nv * 8, h->stream));
        CK(cudaMemsetAsync(s.acc_cov, 0, C * Nv * Nv * 8, h->stream)); CK(cudaMemsetAsync(s.acc_vars, 0, C * Nv * 8, h->stream));
        CK(cudaMemsetAsync(s.acc_count, 0, 8, h->stream));
    }
    CK(cudaMemsetAsync(s.ens_ticket, 0, (size_t)s.E * 4, h->stream));
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
        { int r = raise_dyn_smem((const void *)tg_k_spec, h->spec_smem); if (r) return r; }
        { int r = raise_dyn_smem((const void *)tg_k_propose, h->propose_smem); if (r) return r; }   // static + dynamic may exceed 48 KB
        if (need_pro > 48 * 1024) { int r = raise_dyn_smem((const void *)tg_k_prologue, need_pro); if (r) return r; }
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
    // the handle's own launch group: unit table, argument records and the step graph are ready when tg_create returns
    { tg_handle *one[1] = {h}; int r = batch_create(one, 1, &h->self); if (r) return r; }
    return batch_refresh(h->self);
}

extern "C" int tg_destroy(tg_handle *h)
{
    if (!h) return TG_OK;
    { std::lock_guard<std::mutex> lk(g_live_mu); g_live.erase(h); }
    cudaSetDevice(h->device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    if (h->stream2) cudaStreamSynchronize(h->stream2);
    if (h->self) { batch_destroy(h->self); h->self = nullptr; }
    if (h->nccl_comm && g_nccl_destroy) { g_nccl_destroy(h->nccl_comm); h->nccl_comm = nullptr; }
    for (int r = 0; r < TG_XCHG_MAXR; r++) if (h->xpeer[r]) cudaIpcCloseMemHandle(h->xpeer[r]);
    if (h->xbuf) cudaFree(h->xbuf);
    if (h->xerr_host) cudaFreeHost(h->xerr_host);
    drop_graphs(h);
    for (auto e : h->ev_pool) cudaEventDestroy(e);
    for (int k = 0; k < 2; k++) if (h->watch[k]) cudaEventDestroy(h->watch[k]);
    for (void *q : h->allocs) cudaFree(q);
    for (EvalBuf *b : {&h->sp, &h->ev, &h->ev2}) { if (b->pin_in) cudaFreeHost(b->pin_in); if (b->pin_out) cudaFreeHost(b->pin_out); if (b->pin_in2) cudaFreeHost(b->pin_in2); if (b->pin_out2) cudaFreeHost(b->pin_out2); }
    for (int k = 0; k < 2; k++) if (h->ev_done[k]) cudaEventDestroy(h->ev_done[k]);
    if (h->mom_note) cudaFreeHost(h->mom_note);
    if (h->stream2) cudaStreamDestroy(h->stream2);
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

static int eval_resident(tg_handle *h, int B, double *model_dev, bool pdl = false, EvalBuf *buf = nullptr, cudaStream_t st = nullptr)
{
    EvalBuf &e = buf ? *buf : h->ev;
    if (!st) st = h->stream;
    if (B < 1 || B != e.B) return fail(TG_E_BADARG, "tg_eval_resident: B differs from the uploaded batch");
    int r = launch_eval_kernels(h, e, e.params, B, model_dev, true, false, pdl, st); if (r) return r;
    CK(launch_k(tg_k_reduce_finalize, dim3(B), dim3(256), 0, st, pdl && !h->timing, h->d, (const double2 *)e.partials, h->nitems, e.sums, (const double *)e.T, e.t,
                e.logL, e.logPrior, e.logPost, e.status));
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

// The whole batched generate_model call as ONE graph launch through staging slot `slot` (0 or 1): H2D of the packed inputs,
// prologue, far-field pre-pass, fused kernel, reduce+finalize, D2H of the packed results.  eval_submit returns once the work is
// enqueued; eval_collect blocks until that slot's results are in its pinned buffer.
static int eval_submit(tg_handle *h, const double *params, int B, const double *T, int slot)
{
    CK(cudaSetDevice(h->device));
    if (slot && !h->stream2) CK(cudaStreamCreateWithFlags(&h->stream2, cudaStreamNonBlocking));
    EvalBuf &b = slot ? h->ev2 : h->ev;
    cudaStream_t st = slot ? h->stream2 : h->stream;
    if (B > b.cap) { CK(cudaStreamSynchronize(h->stream)); if (h->stream2) CK(cudaStreamSynchronize(h->stream2)); }
    { int r = alloc_evalbuf(h, b, B); if (r) return r; }
    if (b.B != B) { CK(cudaStreamSynchronize(st)); b.layout(B, h->d.Np); }
    if (h->ev_graph_B != B) {
        if (h->ev_graph) { cudaGraphExecDestroy(h->ev_graph); h->ev_graph = nullptr; }
        if (h->ev_graph2) { cudaGraphExecDestroy(h->ev_graph2); h->ev_graph2 = nullptr; }
    }
    cudaGraphExec_t &ge = slot ? h->ev_graph2 : h->ev_graph;
    double *pin = b.pin_in, *pout = b.pin_out;
    if (!ge) {
        if (h->u_ev.B != B) {                                                  // the unit table is shared by the two slots: nothing of the other may be in flight
            CK(cudaStreamSynchronize(h->stream)); if (h->stream2) CK(cudaStreamSynchronize(h->stream2));
            int r = build_units(h, h->u_ev, B); if (r) return r;
        }
        { int r = sync_eval_args(h, b, nullptr, B); if (r) return r; }        // not inside the capture (uploads on the handle's first stream)
        CK(cudaStreamSynchronize(h->stream));
        cudaGraph_t g;
        const long lt = h->n_launch_total, le = h->n_eval_launch;
        CK(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
        cudaMemcpyAsync(b.in, pin, (size_t)B * (h->d.Np + 1) * 8, cudaMemcpyHostToDevice, st);
        int r = eval_resident(h, B, nullptr, getenv("TG_PDL") ? atoi(getenv("TG_PDL")) != 0 : true, &b, st);
        cudaMemcpyAsync(pout, b.out, (size_t)B * 28, cudaMemcpyDeviceToHost, st);
        cudaError_t ce = cudaStreamEndCapture(st, &g);
        h->n_launch_total = lt; h->n_eval_launch = le;
        if (r) return r;
        CK(ce);
        CK(cudaGraphInstantiate(&ge, g, 0));
        cudaGraphDestroy(g);
        h->ev_graph_B = B;
    }
    memcpy(pin, params, (size_t)B * h->d.Np * 8);
    memcpy(pin + (size_t)B * h->d.Np, T, (size_t)B * 8);
    CK(cudaGraphLaunch(ge, st));
    h->n_launch_total += 3 + ((h->ffw && h->ff_nmt > 0) ? 1 : 0); h->n_eval_launch += 1;
    if (!h->ev_done[slot]) CK(cudaEventCreateWithFlags(&h->ev_done[slot], cudaEventDisableTiming));
    CK(cudaEventRecord(h->ev_done[slot], st));
    h->slot_B[slot] = B;
    return TG_OK;
}
static int eval_collect(tg_handle *h, int slot, double *logL, double *logPrior, double *logPost, int *status)
{
    const int B = h->slot_B[slot];
    if (B < 1 || !h->ev_done[slot]) return fail(TG_E_BADARG, "tg_eval_wait: nothing was submitted through this slot");
    CK(cudaSetDevice(h->device));
    CK(cudaEventSynchronize(h->ev_done[slot]));
    const double *pout = slot ? h->ev2.pin_out : h->ev.pin_out;
    if (logL) memcpy(logL, pout, (size_t)B * 8);
    if (logPrior) memcpy(logPrior, pout + B, (size_t)B * 8);
    if (logPost) memcpy(logPost, pout + 2 * (size_t)B, (size_t)B * 8);
    if (status) memcpy(status, pout + 3 * (size_t)B, (size_t)B * 4);
    h->slot_B[slot] = 0;
    return TG_OK;
}

extern "C" int tg_eval_async(tg_handle *h, const double *params, int B, const double *T, int slot)
{
    if (!h || !params || !T || B < 1 || slot < 0 || slot > 1) return fail(TG_E_BADARG, "tg_eval_async: bad argument");
    if (h->slot_B[slot] != 0) return fail(TG_E_BADARG, "tg_eval_async: this slot still holds results that were not collected (tg_eval_wait)");
    if (h->slot_B[1 - slot] != 0 && h->slot_B[1 - slot] != B) return fail(TG_E_BADARG, "tg_eval_async: the two slots in flight must hold batches of the same size");
    if (h->slot_B[1 - slot] == 0) { int r = refresh_items_if_flagged(h); if (r) return r; }      // (scheduling only; never with work in flight)
    if (!h->use_graph || h->timing) return fail(TG_E_UNSUPPORTED, "tg_eval_async needs the graph path (tg_set_tuning use_graph = 1, no kernel timing)");
    return eval_submit(h, params, B, T, slot);
}
extern "C" int tg_eval_wait(tg_handle *h, int slot, double *logL, double *logPrior, double *logPost, int *status)
{
    if (!h || slot < 0 || slot > 1) return fail(TG_E_BADARG, "tg_eval_wait: bad argument");
    return eval_collect(h, slot, logL, logPrior, logPost, status);
}

extern "C" int tg_eval(tg_handle *h, const double *params, int B, const double *T,
                       double *logL, double *logPrior, double *logPost, int *status, double *model)
{
    if (!h || !params || !T || B < 1) return fail(TG_E_BADARG, "tg_eval: bad argument");
    { int r = refresh_items_if_flagged(h); if (r) return r; }
    if (!model && h->use_graph && !h->timing) {
        int r = eval_submit(h, params, B, T, 0); if (r) return r;
        return eval_collect(h, 0, logL, logPrior, logPost, status);
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

extern "C" int tg_get_state_means(tg_handle *h, double *vars_mean, double *mu_mean, double *sigma_mean, double *covarmat_mean, long *count, int reset)
{
    if (!h) return fail(TG_E_BADARG, "null handle");
    if (!h->s.acc_cov) return fail(TG_E_UNSUPPORTED, "tg_get_state_means: the handle was created without TG_F_RESTORE_MEANS");
    CK(cudaSetDevice(h->device));
    CK(cudaStreamSynchronize(h->stream));
    const size_t C = h->s.C, Nv = h->s.Nv;
    long long n = 0;
    CK(cudaMemcpy(&n, h->s.acc_count, 8, cudaMemcpyDeviceToHost));
    if (count) *count = (long)n;
    if (n == 0) {      // nothing summed yet: the current values (what the reference's first write after a restore holds)
        int r = tg_get_state(h, vars_mean, mu_mean, sigma_mean, covarmat_mean, nullptr, nullptr, nullptr); if (r) return r;
    } else {
        const double inv = 1.0 / (double)n;
        auto fetch = [&](double *dst, const double *src, size_t len) -> int {
            if (!dst) return TG_OK;
            CK(cudaMemcpy(dst, src, len * 8, cudaMemcpyDeviceToHost));
            for (size_t k = 0; k < len; k++) dst[k] *= inv;
            return TG_OK;
        };
        int r;
        if ((r = fetch(vars_mean, h->s.acc_vars, C * Nv)) || (r = fetch(mu_mean, h->s.acc_mu, C * Nv)) || (r = fetch(sigma_mean, h->s.acc_sigma, C)) ||
            (r = fetch(covarmat_mean, h->s.acc_cov, C * Nv * Nv))) return r;
    }
    if (reset) {
        CK(cudaMemsetAsync(h->s.acc_sigma, 0, C * 8, h->stream)); CK(cudaMemsetAsync(h->s.acc_mu, 0, C * Nv * 8, h->stream));
        CK(cudaMemsetAsync(h->s.acc_cov, 0, C * Nv * Nv * 8, h->stream)); CK(cudaMemsetAsync(h->s.acc_vars, 0, C * Nv * 8, h->stream));
        CK(cudaMemsetAsync(h->s.acc_count, 0, 8, h->stream));
        CK(cudaStreamSynchronize(h->stream));
    }
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
        if (h->nccl_comm || h->xchg_on) { r = shard_reduce_and_install(h); if (r) return r; }
        else tg_k_install_current<<<((int)C + 127) / 128, 128, 0, h->stream>>>(h->d, h->s, h->sp.t, h->sp.sums);
        CK(cudaGetLastError());
        CK(cudaStreamSynchronize(h->stream));
        { int r = build_items(h, h->sp, (int)C); if (r) return r; }
    }
    CK(cudaStreamSynchronize(h->stream));
    return h->self ? batch_refresh(h->self) : TG_OK;
}

// ------------------------------------------------------------------------------------------
// MALA::execute on the device: launch groups (a handle is a group of one star)
// ------------------------------------------------------------------------------------------
static void batch_drop_graph(tg_batch *b)
{
    if (b->graph) { cudaGraphExecDestroy(b->graph); b->graph = nullptr; }
    for (int k = 0; k < 2; k++) if (b->graph_small[k]) { cudaGraphExecDestroy(b->graph_small[k]); b->graph_small[k] = nullptr; }
}

static cudaEvent_t batch_next_event(tg_batch *b)
{
    if (b->ev_used == b->ev_pool.size()) { cudaEvent_t e; cudaEventCreate(&e); b->ev_pool.push_back(e); }
    return b->ev_pool[b->ev_used++];
}

static int batch_harvest_events(tg_batch *b)
{
    if (b->ev_used == 0 && b->ffev_used == 0) return TG_OK;
    CK(cudaStreamSynchronize(b->stream));
    for (size_t k = 0; k + 1 < b->ffev_used; k += 2) {
        float ms = 0;
        CK(cudaEventElapsedTime(&ms, b->ffev_pool[k], b->ffev_pool[k + 1]));
        b->ms_ff += ms;
    }
    b->ffev_used = 0;
    for (size_t k = 0; k + 1 < b->ev_used; k += 2) {
        float ms = 0;
        CK(cudaEventElapsedTime(&ms, b->ev_pool[k], b->ev_pool[k + 1]));
        b->ms_eval += ms;
    }
    b->ev_used = 0;
    return TG_OK;
}

template <class T> static int batch_dalloc(tg_batch *b, T **p, size_t n)
{
    void *q = nullptr;
    CK(cudaMalloc(&q, std::max<size_t>(n, 1) * sizeof(T)));
    b->allocs.push_back(q);
    *p = (T *)q;
    return TG_OK;
}

static int batch_destroy(tg_batch *b)
{
    if (!b) return TG_OK;
    cudaSetDevice(b->device);
    if (b->stream && !b->hs.empty() && handle_alive(b->hs[0])) cudaStreamSynchronize(b->stream);     // (the stream is the first star's)
    if (b->side) cudaStreamSynchronize(b->side);
    if (b->hs.size() > 1) for (tg_handle *h : b->hs) if (handle_alive(h) && h->self && h->self != b) { schedule_defaults(h, h->s.C); build_items(h, h->sp, h->s.C); }   // their own shape back
    batch_drop_graph(b);
    for (auto e : b->ev_pool) cudaEventDestroy(e);
    for (auto e : b->ffev_pool) cudaEventDestroy(e);
    for (void *q : b->allocs) cudaFree(q);
    for (int k = 0; k < 2; k++) if (b->ev_run[k]) cudaEventDestroy(b->ev_run[k]);
    if (b->ev_fork) cudaEventDestroy(b->ev_fork);
    if (b->ev_join) cudaEventDestroy(b->ev_join);
    if (b->side) cudaStreamDestroy(b->side);
    cudaGetLastError();                  // nothing of a teardown may surface in a later call
    delete b;
    return TG_OK;
}

static int batch_create(tg_handle **hs, int n, tg_batch **out)
{
    if (!hs || n < 1 || n >= (1 << (31 - TG_UNIT_STAR_SHIFT)) || !out) return fail(TG_E_BADARG, "tg_batch_create: bad argument");
    for (int k = 0; k < n; k++) {
        if (!hs[k]) return fail(TG_E_BADARG, "tg_batch_create: null handle");
        if (hs[k]->device != hs[0]->device) return fail(TG_E_BADARG, "tg_batch_create: the stars of a launch group live on one GPU");
        if (hs[k]->nccl_comm && n > 1) return fail(TG_E_BADARG, "tg_batch_create: a bin-sharded handle runs on its own");
        for (int j = 0; j < k; j++) if (hs[j] == hs[k]) return fail(TG_E_BADARG, "tg_batch_create: handle listed twice");
    }
    tg_batch *b = new tg_batch();
    b->hs.assign(hs, hs + n);
    b->device = hs[0]->device; b->stream = hs[0]->stream;
    b->use_graph = hs[0]->use_graph; b->steps_per_graph = hs[0]->steps_per_graph;
    b->units_version.assign(n, -1);
    int r = TG_OK;
    auto bail = [&](int code) { const std::string keep = g_err; batch_destroy(b); g_err = keep; return code; };
    if (cudaSetDevice(b->device) != cudaSuccess) return bail(fail(TG_E_CUDA, "cudaSetDevice"));
    if (cudaStreamCreateWithFlags(&b->side, cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreateWithFlags(&b->ev_fork, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&b->ev_join, cudaEventDisableTiming) != cudaSuccess) return bail(fail(TG_E_CUDA, "tg_batch_create: stream / events"));
    b->no_side = getenv("TG_NO_SIDE_STREAM") != nullptr;
    if (const char *e = getenv("TG_PDL")) b->pdl = atoi(e) != 0;
    std::vector<int2> cmap;
    for (int k = 0; k < n; k++) for (int c = 0; c < hs[k]->s.C; c++) cmap.push_back(make_int2(k, c));
    b->Ctot = (int)cmap.size();
    if (n > 1) {         // the units of every star are shaped for the chains of the whole launch
        for (int k = 0; k < n; k++) { schedule_defaults(hs[k], b->Ctot); int r2 = build_items(hs[k], hs[k]->sp, hs[k]->s.C); if (r2) return bail(r2); }
    }
    if ((r = batch_dalloc(b, &b->d_cmap, cmap.size())) || (r = batch_dalloc(b, &b->d_ev, n)) || (r = batch_dalloc(b, &b->d_pm, n)) ||
        (r = batch_dalloc(b, &b->d_ps, n)) || (r = batch_dalloc(b, &b->d_fin, n)) || (r = batch_dalloc(b, &b->d_ff, n))) return bail(r);
    for (int k = 0; k < n; k++) if (cudaStreamSynchronize(hs[k]->stream) != cudaSuccess) return bail(fail(TG_E_CUDA, "tg_batch_create: sync"));
    if (cudaMemcpy(b->d_cmap, cmap.data(), cmap.size() * sizeof(int2), cudaMemcpyHostToDevice) != cudaSuccess) return bail(fail(TG_E_CUDA, "tg_batch_create: chain map"));
    *out = b;
    return TG_OK;
}

// Argument records of the three step kernels for every star -> device arrays (uploaded when something changed)
static int batch_sync_args(tg_batch *b)
{
    const size_t n = b->hs.size();
    std::vector<TgEvalArgs> ev(n); std::vector<TgProposeArgs> pm(n), ps(n); std::vector<TgFinishArgs> fin(n); std::vector<TgFFArgs> ff(n);
    int cpc = 0, ppc = 1, nmt_max = 0; size_t fsm = 0;
    memset(ev.data(), 0, n * sizeof(TgEvalArgs)); memset(pm.data(), 0, n * sizeof(TgProposeArgs));
    memset(ps.data(), 0, n * sizeof(TgProposeArgs)); memset(fin.data(), 0, n * sizeof(TgFinishArgs)); memset(ff.data(), 0, n * sizeof(TgFFArgs));
    int rows = 0; size_t psm = 0, esm = 0;
    for (size_t k = 0; k < n; k++) {
        tg_handle *h = b->hs[k];
        TgEvalArgs &ea = ev[k];
        ea.d = h->d; ea.t = h->sp.t; ea.x = h->x; ea.y = h->y; ea.lnx = h->lnx; ea.sig = h->sig; ea.partials = h->sp.partials; ea.model = nullptr;
        ea.B = h->s.C; ea.nitems = h->nitems; ea.max_group = h->chains_per_block; ea.list_cap = h->list_cap;
        ea.mom = h->mom; ea.mom_sub = h->mom_sub; ea.mom_fail = h->mom_fail; ea.mom_note = h->mom_note; ea.ff = ff_view(h, h->sp);
        ff[k] = ff_args(h, h->sp, h->s.C);
        if (ff[k].nmt > 0) { nmt_max = std::max(nmt_max, ff[k].nmt); fsm = std::max(fsm, ff_smem(h)); }
        // The adaptation candidates (Robbins-Monro + Cholesky for both outcomes of the decision) depend on the state the previous
        // finish kernel left and on nothing of this step (the CTA of the "accepted" outcome recomputes the proposed point itself),
        // and only the finish kernel needs them: they run on a side stream next to the proposal, the prologue and the fused kernel.
        const int learn = h->m.n_learn > 0, fused = learn && h->fused_spec;
        const int split = h->tri_smem && !h->no_split;        // the log-prior of the proposed point gets its own CTAs (role 3) on the side stream
        TgProposeArgs &a = pm[k];
        a.s = h->s; a.m = h->m; a.pro.d = h->d; a.pro.t = h->sp.t; a.pro.params = h->sp.params; a.pro.pri = h->pri; a.pro.sw = h->sw;
        a.packed = h->smem_matrix; a.tri_smem = h->tri_smem; a.fused_spec = fused; a.smem_matrix = h->smem_matrix;
        a.role_base = 0; a.parts = split ? TG_PRO_MODEL : (TG_PRO_MODEL | TG_PRO_PRIOR); a.nroles = 1;
        ps[k] = a; ps[k].role_base = fused ? 1 : 3; ps[k].nroles = (fused ? 2 : 0) + (split ? 1 : 0);
        rows = std::max(rows, ps[k].nroles);
        psm = std::max(psm, h->propose_smem); esm = std::max(esm, TG_EVAL_SMEM(h->chains_per_block, h->list_cap));
        TgFinishArgs &f = fin[k];
        f.d = h->d; f.s = h->s; f.m = h->m; f.t = h->sp.t; f.partials = h->sp.partials; f.ntiles = h->nitems; f.sums = h->sp.sums;
        f.presummed = b->presummed; f.rec = h->rec; f.record = 1;
    }
    b->side_rows = rows; b->propose_smem = psm; b->eval_smem = esm;
    if (nmt_max > 0) { ppc = ff_pairs_per_cta(nmt_max, b->Ctot, b->hs[0]->n_sm); cpc = ff_ctas_per_chain(nmt_max, ppc); }
    if (cpc != b->ff_cpc || ppc != b->ff_ppc || fsm != b->ff_smem) { b->ff_cpc = cpc; b->ff_ppc = ppc; b->ff_smem = fsm; batch_drop_graph(b); }
    const bool same = b->args_valid && b->h_ev.size() == n && memcmp(ev.data(), b->h_ev.data(), n * sizeof(TgEvalArgs)) == 0 && memcmp(ff.data(), b->h_ff.data(), n * sizeof(TgFFArgs)) == 0 &&
                      memcmp(pm.data(), b->h_pm.data(), n * sizeof(TgProposeArgs)) == 0 && memcmp(ps.data(), b->h_ps.data(), n * sizeof(TgProposeArgs)) == 0 &&
                      memcmp(fin.data(), b->h_fin.data(), n * sizeof(TgFinishArgs)) == 0;
    if (same) return TG_OK;
    g_dbg_count[3]++;
    b->h_ev.swap(ev); b->h_pm.swap(pm); b->h_ps.swap(ps); b->h_fin.swap(fin); b->h_ff.swap(ff);
    CK(cudaStreamSynchronize(b->stream));
    CK(cudaMemcpy(b->d_ev, b->h_ev.data(), n * sizeof(TgEvalArgs), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(b->d_pm, b->h_pm.data(), n * sizeof(TgProposeArgs), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(b->d_ps, b->h_ps.data(), n * sizeof(TgProposeArgs), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(b->d_fin, b->h_fin.data(), n * sizeof(TgFinishArgs), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(b->d_ff, b->h_ff.data(), n * sizeof(TgFFArgs), cudaMemcpyHostToDevice));
    b->args_valid = true;
    return TG_OK;
}

static int batch_build_units(tg_batch *b)
{
    std::vector<int4> units; std::vector<double> cost;
    bool keep_order = false;
    for (size_t k = 0; k < b->hs.size(); k++) { make_units(b->hs[k], b->hs[k]->s.C, (int)k, units, cost); keep_order = keep_order || b->hs[k]->no_order; }
    int r = upload_units(b->allocs, b->stream, b->units, units, cost, keep_order); if (r) return r;
    for (size_t k = 0; k < b->hs.size(); k++) b->units_version[k] = b->hs[k]->items_version;
    return TG_OK;
}

// One MCMC step of every star of the group: begin = proposal + prologue + fused kernel, end = decision + swaps.
static int batch_launch_step(tg_batch *b, bool begin, bool end)
{
    const int Ct = b->Ctot;
    if (begin) {
        if (b->side_rows > 0) {
            CK(cudaEventRecord(b->ev_fork, b->stream));
            CK(cudaStreamWaitEvent(b->side, b->ev_fork, 0));
            tg_k_propose<<<dim3(Ct, b->side_rows), 256, b->propose_smem, b->no_side ? b->stream : b->side>>>(b->d_ps, b->d_cmap);
            CK(cudaEventRecord(b->ev_join, b->side));
            b->side_pending = true;
            b->n_launch_total++;
        }
        CK(launch_k(tg_k_propose, dim3(Ct, 1), dim3(256), b->propose_smem, b->stream, b->pdl && !b->timing, (const TgProposeArgs *)b->d_pm, (const int2 *)b->d_cmap));
        b->n_launch_total++;
        for (tg_handle *h : b->hs) if (h->m.n_learn > 0 && !h->fused_spec) {
            // very large Nv: the candidates read the proposed point from HBM, so they follow the propose kernel
            CK(cudaEventRecord(b->ev_fork, b->stream));
            CK(cudaStreamWaitEvent(b->side, b->ev_fork, 0));
            TgSpecArgs sa{h->s, h->m, h->smem_matrix, 0};
            tg_k_spec<<<dim3(h->s.C, 2), 256, h->spec_smem, b->side>>>(sa);
            CK(cudaEventRecord(b->ev_join, b->side));
            b->side_pending = true;
            b->n_launch_total++;
        }
        if (b->ff_cpc > 0) {      // far-field pre-pass of the proposed points of every star
            if (b->timing) {
                if (b->ffev_used + 2 > b->ffev_pool.size()) { cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1); b->ffev_pool.push_back(e0); b->ffev_pool.push_back(e1); }
                cudaEventRecord(b->ffev_pool[b->ffev_used], b->stream);
            }
            CK(launch_k(tg_k_ff, dim3((unsigned)b->ff_cpc, (unsigned)Ct), dim3(256), b->ff_smem, b->stream, b->pdl && !b->timing, (const TgFFArgs *)b->d_ff, (const int2 *)b->d_cmap, b->ff_ppc));
            b->n_launch_total++;
            if (b->timing) { cudaEventRecord(b->ffev_pool[b->ffev_used + 1], b->stream); b->ffev_used += 2; }
        }
        if (b->timing) {
            if (b->ev_used >= 8192) { int r = batch_harvest_events(b); if (r) return r; }
            cudaEventRecord(batch_next_event(b), b->stream);
        }
        CK(launch_k(tg_k_eval, dim3(b->units.n), dim3(TG_THREADS), b->eval_smem, b->stream, b->pdl && !b->timing, (const TgEvalArgs *)b->d_ev, (const int4 *)b->units.dev));
        if (b->timing) cudaEventRecord(batch_next_event(b), b->stream);
        b->n_eval_launch++; b->n_launch_total++;
        if (b->presummed) for (tg_handle *h : b->hs) {
            tg_k_reduce<<<h->s.C, 256, 0, b->stream>>>(h->sp.partials, h->nitems, h->sp.sums); b->n_launch_total++;
            if (h->xchg_on) {         // bin-sharded variant over peer memory: one kernel, 2*C doubles per rank over NVLink, part of the step's graph
                tg_k_xchg<<<1, 256, 0, b->stream>>>(h->xchg, h->sp.sums, h->s.C);
                b->n_launch_total++;
            } else
            if (h->nccl_comm) {       // bin-sharded variant: per-chain partial sums of all shards, part of the step's graph
                const int rc = g_nccl_allreduce(h->sp.sums, h->sp.sums, (size_t)2 * h->s.C, /*ncclDouble*/ 8, /*ncclSum*/ 0, h->nccl_comm, b->stream);
                if (rc != 0) return fail(TG_E_CUDA, "ncclAllReduce failed (" + std::to_string(rc) + ")");
                b->n_launch_total++;
            }
        }
    }
    if (end) {
        if (b->side_pending) { CK(cudaStreamWaitEvent(b->stream, b->ev_join, 0)); b->side_pending = false; }
        CK(launch_k(tg_k_finish, dim3(Ct), dim3(256), 0, b->stream, b->pdl && !b->presummed && !b->timing, (const TgFinishArgs *)b->d_fin, (const int2 *)b->d_cmap));
        b->n_launch_total++;
    }
    CK(cudaGetLastError());
    return TG_OK;
}

// Capture + instantiate the graphs of steps_per_graph, 4 and 1 steps (never inside a caller's timed loop: batch_refresh runs at create,
// after a restore, and at the END of a run when the work items moved)
static int batch_capture_one(tg_batch *b, int nsteps, cudaGraphExec_t *out)
{
    for (int attempt = 0; !*out && attempt < 2; attempt++) {
        cudaGraph_t g = nullptr;
        const long lt = b->n_launch_total, le = b->n_eval_launch;
        CK(cudaStreamBeginCapture(b->stream, cudaStreamCaptureModeThreadLocal));
        int r = TG_OK;
        for (int q = 0; q < nsteps && r == TG_OK; q++) r = batch_launch_step(b, true, true);
        cudaError_t ce = cudaStreamEndCapture(b->stream, &g);
        b->n_launch_total = lt; b->n_eval_launch = le; b->side_pending = false;
        if (ce == cudaSuccess && r == TG_OK) ce = cudaGraphInstantiate(out, g, 0);
        if (g) cudaGraphDestroy(g);
        if (ce == cudaSuccess && r == TG_OK) { cudaGraphUpload(*out, b->stream); g_dbg_count[1]++; break; }
        *out = nullptr;
        if (!b->pdl || attempt == 1) { if (r) return r; CK(ce); }
        cudaGetLastError();                  // a driver that cannot capture programmatic edges: the same graphs with full dependencies
        b->pdl = false;
        batch_drop_graph(b);                 // (the ones captured so far carry programmatic edges; batch_capture takes a second pass)
    }
    return TG_OK;
}
static const int TG_SMALL_GRAPH[2] = {4, 1};
static int batch_capture(tg_batch *b)
{
    for (int pass = 0; pass < 2; pass++) {   // a second pass if programmatic edges had to be given up half way
        const bool pdl0 = b->pdl;
        if (!b->graph) { int r = batch_capture_one(b, b->steps_per_graph, &b->graph); if (r) return r; }
        for (int k = 0; k < 2; k++) if (!b->graph_small[k] && TG_SMALL_GRAPH[k] < b->steps_per_graph) { int r = batch_capture_one(b, TG_SMALL_GRAPH[k], &b->graph_small[k]); if (r) return r; }
        if (b->pdl == pdl0) break;
    }
    return TG_OK;
}

// Brings the group up to date with its stars: work items around freshly flagged background tiles, unit table, argument
// records, step graph.  Cheap when nothing changed.
static int batch_refresh(tg_batch *b)
{
    for (tg_handle *h : b->hs) { int r = refresh_items_if_flagged(h); if (r) return r; }
    bool stale = false;
    for (size_t k = 0; k < b->hs.size(); k++) stale = stale || (b->units_version[k] != b->hs[k]->items_version);
    if (stale) { batch_drop_graph(b); int r = batch_build_units(b); if (r) return r; }
    { int r = batch_sync_args(b); if (r) return r; }
    if (b->use_graph && !b->timing && !b->graph) { int r = batch_capture(b); if (r) return r; }
    return TG_OK;
}

static int batch_set_ctrl(tg_batch *b, long i, long i0)
{
    for (tg_handle *h : b->hs) {
        TgCtrl c{i, i0, 0u, 0};
        CK(cudaMemcpyAsync(h->s.ctrl, &c, sizeof(c), cudaMemcpyHostToDevice, b->stream));
    }
    return TG_OK;
}

static int drain_records(tg_handle *h, cudaStream_t st, tg_records *rec, long done, long n)
{
    if (!rec || n <= 0) return TG_OK;
    const size_t C = h->s.C, Nv = h->s.Nv, E = h->s.E;
    const TgRecDev &r = h->rec;
#define CP(dst, src, per, T) if (rec->dst) CK(cudaMemcpyAsync(rec->dst + (size_t)done * (per), r.src, (size_t)n * (per) * sizeof(T), cudaMemcpyDeviceToHost, st))
    CP(vars, vars, C * Nv, double); CP(logL, logL, C, double); CP(logPrior, logPrior, C, double); CP(logPost, logPost, C, double);
    CP(Pmove, Pmove, C, double); CP(moved, moved, C, unsigned char); CP(status, status, C, int);
    CP(swap_attempt, swap_attempt, E, int); CP(swap_chain0, swap_chain0, E, int); CP(swap_done, swap_done, E, int); CP(swap_Pswap, swap_Pswap, E, double);
#undef CP
    return TG_OK;
}

static int batch_launches_per_step(const tg_batch *b)
{
    int n = 3 + (b->side_rows > 0 ? 1 : 0) + (b->ff_cpc > 0 ? 1 : 0);
    for (tg_handle *h : b->hs) { if (h->m.n_learn > 0 && !h->fused_spec) n++; if (b->presummed) n += 1 + ((h->nccl_comm || h->xchg_on) ? 1 : 0); }
    return n;
}

static int batch_run(tg_batch *b, long i0, long nsteps, tg_records **recs)
{
    CK(cudaSetDevice(b->device));
    { int r = batch_refresh(b); if (r) return r; }
    long cap = b->hs[0]->rec.cap;
    for (tg_handle *h : b->hs) cap = std::min(cap, h->rec.cap);
    const int G = b->steps_per_graph;
    long done = 0;
    for (int k = 0; k < 2; k++) if (!b->ev_run[k]) CK(cudaEventCreate(&b->ev_run[k]));
    CK(cudaEventRecord(b->ev_run[0], b->stream));          // the steps of this call on the device's clock: from here ...
    while (done < nsteps) {
        const long n = std::min(nsteps - done, cap);
        { int r = batch_set_ctrl(b, i0 + done, i0 + done); if (r) return r; }
        long k = 0;
        while (b->graph && b->use_graph && !b->timing && k + G <= n) {
            CK(cudaGraphLaunch(b->graph, b->stream)); b->n_launch_total += (long)batch_launches_per_step(b) * G; b->n_eval_launch += G;
            k += G;
            if (!b->probed && nsteps >= 2L * G) {
                // first steps of a long run: which background-only tiles fail their polynomial check at the proposed points is known
                // now; without this look the rest of the call would run with those tiles inside long cheap items
                b->probed = true;
                g_dbg_count[2]++;
                CK(cudaStreamSynchronize(b->stream));
                int r = batch_refresh(b); if (r) return r;
            }
        }
        for (int q = 0; q < 2; q++) {          // the remainder: graphs of 4 steps and of 1 step
            const int gs = TG_SMALL_GRAPH[q];
            while (b->graph_small[q] && b->use_graph && !b->timing && k + gs <= n) {
                CK(cudaGraphLaunch(b->graph_small[q], b->stream)); b->n_launch_total += (long)batch_launches_per_step(b) * gs; b->n_eval_launch += gs;
                k += gs;
            }
        }
        for (; k < n; k++) { int r = batch_launch_step(b, true, true); if (r) return r; }
        if (done + n >= nsteps) CK(cudaEventRecord(b->ev_run[1], b->stream));     // ... to the end of its last step (before the records are drained)
        for (size_t s2 = 0; s2 < b->hs.size(); s2++) { int r = drain_records(b->hs[s2], b->stream, recs ? recs[s2] : nullptr, done, n); if (r) return r; }
        CK(cudaStreamSynchronize(b->stream));
        done += n;
    }
    CK(cudaStreamSynchronize(b->stream));
    // maintenance for the NEXT call, outside any timed loop of the caller's steps: follow the windows of the chains (scheduling
    // only) after a long run, rebuild around flagged tiles, re-capture the graph if the grid changed
    if (nsteps >= 64) for (tg_handle *h : b->hs) { if (h->mom_note) *(volatile int *)h->mom_note = 0; int r = build_items(h, h->sp, h->s.C); if (r) return r; }
    return batch_refresh(b);
}

extern "C" int tg_run(tg_handle *h, long i0, long nsteps, tg_records *rec)
{
    if (!h || nsteps < 0 || !h->self) return fail(TG_E_BADARG, "tg_run: bad argument");
    tg_records *recs[1] = {rec};
    return batch_run(h->self, i0, nsteps, recs);
}

extern "C" int tg_batch_create(tg_handle **handles, int n, tg_batch **out)
{
    if (out) *out = nullptr;
    tg_batch *b = nullptr;
    int r = batch_create(handles, n, &b); if (r) return r;
    r = batch_refresh(b);
    if (r) { const std::string keep = g_err; batch_destroy(b); g_err = keep; return r; }
    *out = b;
    return TG_OK;
}
extern "C" int tg_batch_destroy(tg_batch *b) { return batch_destroy(b); }
extern "C" int tg_batch_run(tg_batch *b, long i0, long nsteps, tg_records **recs)
{
    if (!b || nsteps < 0) return fail(TG_E_BADARG, "tg_batch_run: bad argument");
    return batch_run(b, i0, nsteps, recs);
}
extern "C" int tg_batch_sizes(tg_batch *b, int *nstars, int *Ctot, int *units)
{
    if (!b) return fail(TG_E_BADARG, "null batch");
    if (nstars) *nstars = (int)b->hs.size(); if (Ctot) *Ctot = b->Ctot; if (units) *units = b->units.n;
    return TG_OK;
}
static int batch_kernel_time(tg_batch *b, int reset, long *launches, double *ms_eval, long *total_launches)
{
    CK(cudaSetDevice(b->device));
    { int r = batch_harvest_events(b); if (r) return r; }
    if (launches) *launches = b->n_eval_launch;
    if (ms_eval) *ms_eval = b->ms_eval;
    if (total_launches) *total_launches = b->n_launch_total;
    if (reset == 1) { b->n_eval_launch = 0; b->ms_eval = 0; b->ms_ff = 0; b->n_launch_total = 0; b->timing = true; }
    if (reset == 2) { b->n_eval_launch = 0; b->ms_eval = 0; b->n_launch_total = 0; b->timing = false; return batch_refresh(b); }      // (ms_ff stays readable)
    return TG_OK;
}
extern "C" int tg_batch_kernel_time(tg_batch *b, int reset, long *launches, double *ms_eval, long *total_launches)
{
    if (!b) return fail(TG_E_BADARG, "null batch");
    return batch_kernel_time(b, reset, launches, ms_eval, total_launches);
}
extern "C" int tg_batch_prepass_time(tg_batch *b, double *ms_ff)
{
    if (!b || !ms_ff) return fail(TG_E_BADARG, "tg_batch_prepass_time: bad argument");
    CK(cudaSetDevice(b->device));
    { int r = batch_harvest_events(b); if (r) return r; }
    *ms_ff = b->ms_ff;
    return TG_OK;
}
extern "C" int tg_prepass_time(tg_handle *h, double *ms_ff)
{
    if (!h || !h->self) return fail(TG_E_BADARG, "null handle");
    return tg_batch_prepass_time(h->self, ms_ff);
}
static int batch_last_run_ms(tg_batch *b, double *ms)
{
    if (!b || !ms || !b->ev_run[0] || !b->ev_run[1]) return fail(TG_E_BADARG, "tg_last_run_ms: no run yet");
    CK(cudaSetDevice(b->device));
    CK(cudaEventSynchronize(b->ev_run[1]));
    float f = 0; CK(cudaEventElapsedTime(&f, b->ev_run[0], b->ev_run[1]));
    *ms = f;
    return TG_OK;
}
extern "C" int tg_batch_last_run_ms(tg_batch *b, double *ms) { return batch_last_run_ms(b, ms); }
extern "C" int tg_last_run_ms(tg_handle *h, double *ms) { return batch_last_run_ms(h ? h->self : nullptr, ms); }

// device-side stopwatch on the group's stream: which = 0 start, 1 stop (then the elapsed time is read with tg_event_elapsed_ms of the first star)
extern "C" int tg_batch_event_record(tg_batch *b, int which)
{
    if (!b) return fail(TG_E_BADARG, "null batch");
    return tg_event_record(b->hs[0], which);
}

// Bin-sharded variant, host-driven form: the caller all-reduces the partial sums between the two halves of a step
extern "C" int tg_step_begin(tg_handle *h, long i)
{
    if (!h || !h->self) return fail(TG_E_BADARG, "null handle");
    CK(cudaSetDevice(h->device));
    tg_batch *b = h->self;
    if (!b->presummed) { b->presummed = 1; batch_drop_graph(b); }
    { int r = batch_refresh(b); if (r) return r; }
    { int r = batch_set_ctrl(b, i, i); if (r) return r; }
    return batch_launch_step(b, true, false);
}
extern "C" int tg_step_end(tg_handle *h, long i)
{
    (void)i;
    if (!h || !h->self) return fail(TG_E_BADARG, "null handle");
    CK(cudaSetDevice(h->device));
    return batch_launch_step(h->self, false, true);
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
// Bin-sharded variant inside the library: every rank holds a contiguous bin range of ONE spectrum (tg_problem.Nx_global, shard_lo, ...)
// and the same chain state and Philox streams; from tg_shard_attach on, every step all-reduces the 2*C partial sums
// (sum y/m, sum ln m: likelihoods.cpp:17-28) over the ranks with ONE ncclAllReduce enqueued on the handle's stream between the
// fused kernel and the decision kernel -- part of the captured step graph, no host synchronisation per step.
extern "C" int tg_shard_unique_id(char *id128)
{
    if (!id128) return fail(TG_E_BADARG, "null argument");
    { int r = load_nccl(); if (r) return r; }
    tg_nccl_id id; memset(&id, 0, sizeof(id));
    const int rc = g_nccl_getid(&id);
    if (rc != 0) return fail(TG_E_CUDA, "ncclGetUniqueId failed (" + std::to_string(rc) + ")");
    memcpy(id128, id.internal, 128);
    return TG_OK;
}
extern "C" int tg_shard_attach(tg_handle *h, const char *id128, int nranks, int rank)
{
    if (!h || !h->self || !id128 || nranks < 1 || rank < 0 || rank >= nranks) return fail(TG_E_BADARG, "tg_shard_attach: bad argument");
    if (h->nccl_comm) return fail(TG_E_BADARG, "tg_shard_attach: already attached");
    { int r = load_nccl(); if (r) return r; }
    CK(cudaSetDevice(h->device));
    CK(cudaStreamSynchronize(h->stream));
    tg_nccl_id id; memcpy(id.internal, id128, 128);
    void *comm = nullptr;
    const int rc = g_nccl_init(&comm, nranks, id, rank);
    if (rc != 0 || !comm) return fail(TG_E_CUDA, "ncclCommInitRank failed (" + std::to_string(rc) + ")");
    h->nccl_comm = comm;
    // tg_create evaluated the current points on this rank's bins only: the sums of all ranks make their logL
    { int r = shard_reduce_and_install(h); if (r) return r; }
    tg_batch *b = h->self;
    b->presummed = 1;
    batch_drop_graph(b);
    return batch_refresh(b);
}
extern "C" int tg_shard_detach(tg_handle *h)
{
    if (!h) return fail(TG_E_BADARG, "null handle");
    if (!h->nccl_comm) return TG_OK;
    CK(cudaSetDevice(h->device));
    CK(cudaStreamSynchronize(h->stream));
    if (h->self) batch_drop_graph(h->self);
    g_nccl_destroy(h->nccl_comm);
    h->nccl_comm = nullptr;
    return TG_OK;
}

// The same exchange without NCCL: the ranks' buffers are mapped into one another through CUDA IPC and ONE kernel per step
// (tg_k_xchg, tg_sampler.cuh) stores this rank's 2*C sums into every peer's buffer over NVLink, waits for the peers' stamps and adds
// the contributions in rank order.  tg_shard_ipc_export allocates this rank's buffer and returns its 64-byte IPC handle; every rank
// passes the handles of all ranks (rank order) to tg_shard_ipc_attach.  One process per GPU, all GPUs of one node, <= 8 ranks.
extern "C" int tg_shard_ipc_export(tg_handle *h, int nranks, char *handle64)
{
    if (!h || !handle64 || nranks < 1 || nranks > TG_XCHG_MAXR) return fail(TG_E_BADARG, "tg_shard_ipc_export: bad argument (<= 8 ranks)");
    if (h->xbuf) return fail(TG_E_BADARG, "tg_shard_ipc_export: already exported");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    CK(cudaSetDevice(h->device));
    const size_t n = 2 * (size_t)h->s.C;
    h->xbuf_bytes = (2 * (size_t)nranks * n) * 8 + 2 * (size_t)nranks * 8 + 64;      // data, stamps, counter + error word
    CK(cudaMalloc(&h->xbuf, h->xbuf_bytes));
    CK(cudaMemset(h->xbuf, 0, h->xbuf_bytes));
    CK(cudaDeviceSynchronize());
    cudaIpcMemHandle_t ih;
    CK(cudaIpcGetMemHandle(&ih, h->xbuf));
    memcpy(handle64, &ih, 64);
    return TG_OK;
}
extern "C" int tg_shard_ipc_attach(tg_handle *h, const char *handles, int nranks, int rank)
{
    if (!h || !h->self || !handles || !h->xbuf || nranks < 1 || nranks > TG_XCHG_MAXR || rank < 0 || rank >= nranks) return fail(TG_E_BADARG, "tg_shard_ipc_attach: bad argument");
    if (h->xchg_on || h->nccl_comm) return fail(TG_E_BADARG, "tg_shard_ipc_attach: already attached");
    CK(cudaSetDevice(h->device));
    CK(cudaStreamSynchronize(h->stream));
    const size_t n = 2 * (size_t)h->s.C, data_bytes = (2 * (size_t)nranks * n) * 8, stamp_bytes = 2 * (size_t)nranks * 8;
    TgXchg &x = h->xchg; memset(&x, 0, sizeof(x));
    x.nranks = nranks; x.rank = rank;
    for (int r = 0; r < nranks; r++) {
        void *p = h->xbuf;
        if (r != rank) {
            cudaIpcMemHandle_t ih; memcpy(&ih, handles + 64 * (size_t)r, 64);
            CK(cudaIpcOpenMemHandle(&p, ih, cudaIpcMemLazyEnablePeerAccess));
            h->xpeer[r] = p;
        }
        x.peer_data[r] = (double *)p;
        x.peer_stamp[r] = (unsigned long long *)((char *)p + data_bytes);
    }
    x.seq = (unsigned long long *)((char *)h->xbuf + data_bytes + stamp_bytes);
    CK(cudaHostAlloc((void **)&h->xerr_host, sizeof(int), cudaHostAllocMapped));
    *h->xerr_host = 0;
    x.err = h->xerr_host;
    h->xchg_on = true;
    { int r = shard_reduce_and_install(h); if (r) return r; }      // tg_create evaluated the current points on this rank's bins only
    tg_batch *b = h->self;
    b->presummed = 1;
    batch_drop_graph(b);
    return batch_refresh(b);
}
extern "C" int tg_shard_status(tg_handle *h)
{
    if (!h) return fail(TG_E_BADARG, "null handle");
    if (h->xerr_host && *(volatile int *)h->xerr_host) return fail(TG_E_CUDA, "bin-sharded exchange: a peer did not answer within 2 s; the run is void");
    return TG_OK;
}

extern "C" int tg_sync(tg_handle *h)
{
    if (!h) return fail(TG_E_BADARG, "null handle");
    CK(cudaSetDevice(h->device));
    CK(cudaStreamSynchronize(h->stream));
    if (h->stream2) CK(cudaStreamSynchronize(h->stream2));
    return TG_OK;
}

extern "C" int tg_kernel_time(tg_handle *h, int reset, long *launches, double *ms_eval, long *total_launches)
{
    if (!h) return fail(TG_E_BADARG, "null handle");
    CK(cudaSetDevice(h->device));
    { int r = harvest_events(h); if (r) return r; }
    long bl = 0, bt = 0; double bms = 0;
    if (h->self) { int r = batch_kernel_time(h->self, 0, &bl, &bms, &bt); if (r) return r; }
    if (launches) *launches = h->n_eval_launch + bl;
    if (ms_eval) *ms_eval = h->ms_eval + bms;
    if (total_launches) *total_launches = h->n_launch_total + bt;
    if (reset == 1) { h->n_eval_launch = 0; h->ms_eval = 0; h->n_launch_total = 0; h->timing = true; }
    if (reset == 2) { h->n_eval_launch = 0; h->ms_eval = 0; h->n_launch_total = 0; h->timing = false; }
    if (reset && h->self) return batch_kernel_time(h->self, reset, nullptr, nullptr, nullptr);
    return TG_OK;
}

extern "C" int tg_work_profile(tg_handle *h, long *bins_direct, long *moment_tiles, int *units)
{
    if (!h) return fail(TG_E_BADARG, "null handle");
    long bd = 0, mt = 0;
    for (const int4 &it : h->items_host) {
        if (it.w == 2) mt += it.y / std::max(1, h->mom_sub);
        else bd += std::min<long>((long)it.y * TG_THREADS * 4, (long)h->d.Nx - it.x);
    }
    if (bins_direct) *bins_direct = bd;
    if (moment_tiles) *moment_tiles = mt;
    if (units) *units = h->self ? h->self->units.n : 0;
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

// ------------------------------------------------------------------------------------------
// tg_eval_noalloc: batched generate_model for callers that may not allocate (include/tamcmc_gpu.h).  The scratch of the fused
// kernel -- its argument record, its work-unit table, the per-(chain, item) partial sums and the per-chain totals -- lives in
// ONE file-scope device array of fixed size instead of cudaMalloc'd buffers.  Every GPU that loads the library holds the
// array until the process exits; every call on a device shares it and finds the previous call's contents in it.
// ------------------------------------------------------------------------------------------
#define TG_NA_ARGS_BYTES 4096                                   // the argument record of tg_k_eval
static_assert(sizeof(TgEvalArgs) <= TG_NA_ARGS_BYTES / 2 && sizeof(TgFFArgs) <= TG_NA_ARGS_BYTES / 2, "argument records outgrew their slot of the static scratch");
#define TG_NA_OFF_UNITS  ((size_t)TG_NA_ARGS_BYTES)
#define TG_NA_OFF_PAIRS  (TG_NA_OFF_UNITS + (size_t)TG_NOALLOC_MAX_UNITS * sizeof(int4))
#define TG_NA_OFF_SUMS   (TG_NA_OFF_PAIRS + (size_t)TG_NOALLOC_MAX_PAIRS * sizeof(double2))
#define TG_NA_BYTES      (TG_NA_OFF_SUMS + (size_t)TG_NOALLOC_MAX_B * sizeof(double2))     // 4 KiB + 1 MiB + 4 MiB + 64 KiB
static __device__ __align__(256) unsigned char tg_na_scratch[TG_NA_BYTES];

// In host code `tg_na_scratch` is only the host-side stand-in of the symbol, not a device address: the address a kernel can
// use comes from cudaGetSymbolAddress, asked once per device with that device current.
static int na_scratch(int device, unsigned char **out)
{
    static unsigned char *ptr[64] = {};
    if (device < 0 || device >= 64) return fail(TG_E_BADARG, "tg_eval_noalloc: device ordinal out of range");
    if (!ptr[device]) {
        CK(cudaSetDevice(device));
        void *p = nullptr;
        CK(cudaGetSymbolAddress(&p, tg_na_scratch));
        if (!p) return fail(TG_E_CUDA, "tg_eval_noalloc: cudaGetSymbolAddress returned a null address");
        ptr[device] = (unsigned char *)p;
    }
    *out = ptr[device];
    return TG_OK;
}

extern "C" int tg_eval_noalloc(tg_handle *h, const double *params, int B, const double *T,
                               double *logL, double *logPrior, double *logPost, int *status)
{
    if (!h || !params || !T || B < 1) return fail(TG_E_BADARG, "tg_eval_noalloc: bad argument");
    CK(cudaSetDevice(h->device));
    EvalBuf &b = h->ev;
    if (B > b.cap) return fail(TG_E_BADARG, "tg_eval_noalloc: B = " + std::to_string(B) + " parameter vectors, but this handle's tables hold " + std::to_string(b.cap) +
                               " (tg_create sizes them for Nchains, a larger tg_eval call grows them; this entry point never allocates)");
    if (B > TG_NOALLOC_MAX_B) return fail(TG_E_BADARG, "tg_eval_noalloc: B = " + std::to_string(B) + " exceeds TG_NOALLOC_MAX_B = " + std::to_string(TG_NOALLOC_MAX_B));
    const int nitems = h->nitems;
    if ((long)B * nitems > (long)TG_NOALLOC_MAX_PAIRS)
        return fail(TG_E_BADARG, "tg_eval_noalloc: B x work items = " + std::to_string(B) + " x " + std::to_string(nitems) + " partial sums exceed the static scratch (TG_NOALLOC_MAX_PAIRS = " +
                                 std::to_string(TG_NOALLOC_MAX_PAIRS) + "); use tg_eval");
    std::vector<int4> units; std::vector<double> cost;              // host memory only
    make_units(h, B, 0, units, cost);
    if (units.size() > (size_t)TG_NOALLOC_MAX_UNITS)
        return fail(TG_E_BADARG, "tg_eval_noalloc: " + std::to_string(units.size()) + " work units exceed the static scratch (TG_NOALLOC_MAX_UNITS = " + std::to_string(TG_NOALLOC_MAX_UNITS) + "); use tg_eval");
    if (!h->no_order) {                                             // costliest first, like upload_units
        std::vector<int> idx(units.size());
        for (size_t k = 0; k < idx.size(); k++) idx[k] = (int)k;
        std::stable_sort(idx.begin(), idx.end(), [&](int a, int b2) { return cost[a] > cost[b2]; });
        std::vector<int4> u2(units.size());
        for (size_t k = 0; k < idx.size(); k++) u2[k] = units[idx[k]];
        units.swap(u2);
    }
    unsigned char *base = nullptr;
    { int r = na_scratch(h->device, &base); if (r) return r; }
    TgEvalArgs *d_args = (TgEvalArgs *)base;
    int4 *d_units = (int4 *)(base + TG_NA_OFF_UNITS);
    double2 *d_pairs = (double2 *)(base + TG_NA_OFF_PAIRS), *d_sums = (double2 *)(base + TG_NA_OFF_SUMS);
    if (b.B != B) { CK(cudaStreamSynchronize(h->stream)); b.layout(B, h->d.Np); }
    // inputs through the handle's pinned staging buffer (allocated by tg_create), one copy
    memcpy(b.pin_in, params, (size_t)B * h->d.Np * 8);
    memcpy(b.pin_in + (size_t)B * h->d.Np, T, (size_t)B * 8);
    CK(cudaMemcpyAsync(b.in, b.pin_in, (size_t)B * (h->d.Np + 1) * 8, cudaMemcpyHostToDevice, h->stream));
    TgEvalArgs ea; memset(&ea, 0, sizeof(ea));
    ea.d = h->d; ea.t = b.t; ea.x = h->x; ea.y = h->y; ea.lnx = h->lnx; ea.sig = h->sig; ea.partials = d_pairs; ea.model = nullptr;
    ea.B = B; ea.nitems = nitems; ea.max_group = h->chains_per_block; ea.list_cap = h->list_cap;
    ea.mom = h->mom; ea.mom_sub = h->mom_sub; ea.mom_fail = h->mom_fail; ea.mom_note = h->mom_note; ea.ff = ff_view(h, b);
    const TgFFArgs fa = ff_args(h, b, B);
    TgFFArgs *d_ffa = (TgFFArgs *)(base + TG_NA_ARGS_BYTES / 2);   // the pre-pass's record shares the slot of the argument record
    // the scratch keeps the last call's contents: the record and the table are overwritten, the partial sums start at zero
    // (pageable sources: the runtime stages them before the call returns, so the locals may die)
    CK(cudaMemcpyAsync(d_args, &ea, sizeof(ea), cudaMemcpyHostToDevice, h->stream));
    CK(cudaMemcpyAsync(d_ffa, &fa, sizeof(fa), cudaMemcpyHostToDevice, h->stream));
    CK(cudaMemcpyAsync(d_units, units.data(), units.size() * sizeof(int4), cudaMemcpyHostToDevice, h->stream));
    CK(cudaMemsetAsync(d_pairs, 0, (size_t)B * nitems * sizeof(double2), h->stream));
    CK(cudaMemsetAsync(d_sums, 0, (size_t)B * sizeof(double2), h->stream));
    TgProArgs pa{h->d, b.t, b.params, h->pri, h->sw};
    tg_k_prologue<<<B, TG_PRO_THREADS, TG_PRO_SMEM(h->d.Np), h->stream>>>(pa);
    if (fa.nmt > 0) {
        const int ppc = ff_pairs_per_cta(fa.nmt, B, h->n_sm), cpc = ff_ctas_per_chain(fa.nmt, ppc);
        tg_k_ff<<<dim3(cpc, B), 256, ff_smem(h), h->stream>>>((const TgFFArgs *)d_ffa, (const int2 *)nullptr, ppc);
    }
    tg_k_eval<<<(unsigned)units.size(), TG_THREADS, TG_EVAL_SMEM(h->chains_per_block, h->list_cap), h->stream>>>((const TgEvalArgs *)d_args, (const int4 *)d_units);
    tg_k_reduce_finalize<<<B, 256, 0, h->stream>>>(h->d, d_pairs, nitems, d_sums, b.T, b.t, b.logL, b.logPrior, b.logPost, b.status);
    CK(cudaGetLastError());
    h->n_launch_total += 3; h->n_eval_launch += 1;
    CK(cudaMemcpyAsync(b.pin_out, b.out, (size_t)B * 28, cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    if (logL) memcpy(logL, b.pin_out, (size_t)B * 8);
    if (logPrior) memcpy(logPrior, b.pin_out + B, (size_t)B * 8);
    if (logPost) memcpy(logPost, b.pin_out + 2 * (size_t)B, (size_t)B * 8);
    if (status) memcpy(status, b.pin_out + 3 * (size_t)B, (size_t)B * 4);
    return TG_OK;
}

// What the far-field pre-pass left for the B vectors of the last tg_eval batch (measurement support: the flops the two kernels
// actually execute under the truncation windows follow from these counts)
extern "C" int tg_farfield_profile(tg_handle *h, int B, long *out)
{
    if (!h || !out || B < 1) return fail(TG_E_BADARG, "tg_farfield_profile: bad argument");
    for (int k = 0; k < 6; k++) out[k] = 0;
    out[5] = TG_FF_BINS;
    if (!h->ffw || !h->ev.ffh || h->ff_nmt <= 0 || B > h->ev.cap) return TG_OK;      // path off: nothing prepared
    CK(cudaSetDevice(h->device));
    CK(cudaStreamSynchronize(h->stream));
    std::vector<TgFFHdr> hd((size_t)h->ff_nmt);
    for (int b = 0; b < B; b++) {
        CK(cudaMemcpy(hd.data(), h->ev.ffh + (size_t)b * h->ff_stride, hd.size() * sizeof(TgFFHdr), cudaMemcpyDeviceToHost));
        for (const TgFFHdr &q : hd) { out[0]++; if (q.count < 0) out[1]++; else out[2] += q.count; out[3] += q.nfar; out[4] += q.flags & 1; }
    }
    return TG_OK;
}

// developer check (not part of the ABI): summary of the far-field headers the last tg_eval batch left
extern "C" int tg_dbg_ff_summary(tg_handle *h, int B, long *out /* [6]: micro-tiles, fallbacks, records, far components, noise-folded, max records */)
{
    if (!h || !h->ffw || !h->ev.ffh || B < 1 || B > h->ev.cap) return fail(TG_E_BADARG, "tg_dbg_ff_summary: nothing to look at");
    CK(cudaSetDevice(h->device));
    CK(cudaStreamSynchronize(h->stream));
    std::vector<TgFFHdr> hd((size_t)B * h->ff_stride);
    CK(cudaMemcpy(hd.data(), h->ev.ffh, hd.size() * sizeof(TgFFHdr), cudaMemcpyDeviceToHost));
    for (int k = 0; k < 6; k++) out[k] = 0;
    for (int b = 0; b < B; b++) for (int m = 0; m < h->ff_nmt; m++) {
        const TgFFHdr &q = hd[(size_t)b * h->ff_stride + m];
        out[0]++; if (q.count < 0) out[1]++; else { out[2] += q.count; out[5] = std::max<long>(out[5], q.count); }
        out[3] += q.nfar; out[4] += q.flags & 1;
        if (q.count < 0) fprintf(stderr, "fallback: vector %d micro-tile %d flags %d nfar %d\n", b, m, q.flags, q.nfar);
    }
    return TG_OK;
}
