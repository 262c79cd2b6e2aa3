/*
 * tamcmc_gpu.h -- C-ABI of the B200-native TAMCMC hot path (libtamcmc_gpu.so).
 *
 * Plain pointers and sizes only; no torch / Eigen types.  Every entry point returns an int
 * status (0 = ok, <0 = fatal, see TG_E_*) and never calls exit(); the message of the last
 * failure on the calling thread is available from tg_last_error().  The reference has no
 * FFI: its operator API for this path is the Model_def registry, so each entry point cites
 * the reference interface (file:line under /root/reference/tamcmc) it stands in for.
 * INTEGRATION.md shows the binding a maintainer would add on the reference side.
 *
 * Threading contract (reference: one OpenMP thread per chain on shared Model_def objects,
 * sources/MALA.cpp:632-655): one host thread per handle, one handle per GPU, all work of a
 * handle on one CUDA stream; calls on the same handle are not re-entrant, different handles
 * are independent.
 */
#ifndef TAMCMC_GPU_H
#define TAMCMC_GPU_H

#ifdef __cplusplus
extern "C" {
#endif

#define TG_OK              0
#define TG_E_BADARG       -1   /* bad sizes / null pointers / bad case id (reference: exit(), model_def.cpp:266-285) */
#define TG_E_CUDA         -2   /* CUDA runtime error */
#define TG_E_UNSUPPORTED  -3   /* model case the reference itself kills with exit() (cases 4, 5) */
#define TG_E_NODEVICE     -4   /* no CUDA device: there is NO CPU fallback */

/* per-chain status bits (reference behaviour in brackets) */
#define TG_ST_EMPTY_WINDOW  1  /* imax-imin<=0 [exit(EXIT_FAILURE), build_lorentzian.cpp:428-443] -> logL=NaN, rejected */
#define TG_ST_NAN_WINDOW    2  /* Gamma or f_s NaN: window undefined [uninitialised pmin/pmax, Q1]  -> logL=NaN, rejected */
#define TG_ST_NAN_LOGL      4  /* logL is NaN [warning + reject, MALA.cpp:507-509] */
#define TG_ST_BAD_PRIOR     8  /* prior kind the reference exits on (multivariate Gaussian / unknown switch) */
#define TG_ST_CHOL_FAIL    16  /* (Sigma+eps2 I)*sigma not positive definite [Eigen llt yields NaN silently] */
#define TG_ST_MODE_OVERFLOW 32 /* more Lorentzian components than the device table holds */

typedef struct tg_handle tg_handle;

/* Everything Config::setup (sources/config.cpp:80-189) leaves behind for MALA / Model_def. */
typedef struct tg_problem {
    /* ---- Data{x,y,sigma_y,Nx}, headers/data.h:24-35 (after the freq_range crop) ---- */
    long          Nx;             /* bins held by THIS handle (a shard when bin-sharded)                */
    const double *x, *y;          /* [Nx]                                                                */
    const double *sigma_y;        /* [Nx] or NULL (only the chi_square likelihood reads it)              */
    /* bin-sharded variant: the window arithmetic needs the GLOBAL grid.       */
    long          Nx_global;      /* 0 => not sharded (== Nx)                                            */
    long          shard_lo;       /* global index of this shard's first bin                              */
    double        x_first, x_second, x_last; /* global x[0], x[1], x[Nx_global-1]; ignored if Nx_global==0 */
    /* ---- Modeling: registry switches, Config/default/{models,likelihoods,priors}_ctrl.list ---- */
    int           model_case;     /* model_fct_name_switch   0..14, model_def.cpp:215-265                */
    int           likelihood_case;/* likelihood_fct_name_switch 0 chi(2,2p) | 1 chi_square, :298-307      */
    int           prior_case;     /* prior_fct_name_switch   0..3, model_def.cpp:329-341                 */
    double        likelihood_params; /* p of chi(2,2p), truncated to long like model_def.cpp:300-301     */
    int           plength[11];    /* Input_Data.plength, io_ms_global.cpp:996-1001                       */
    int           Np;             /* == sum(plength)                                                      */
    const double *params0;        /* [Np]  Input_Data.inputs                                              */
    const int    *relax;          /* [Np]  Input_Data.relax (1 = free variable)                           */
    const double *priors;         /* [4*Np] Input_Data.priors, 4 x Np column-major: priors[4*i+r]         */
    const int    *priors_switch;  /* [Np]  Input_Data.priors_names_switch (primepriors_ctrl.list)         */
    double        extra_priors[4];/* Input_Data.extra_priors                                              */
    /* ---- MALA: Config/default/config_default.cfg !MALA section, MALA.cpp:53-112 ---- */
    int           Nchains;        /* total chains C held by this handle                                   */
    int           chains_per_ensemble; /* tempering swaps stay inside an ensemble; 0 => Nchains           */
    const double *Tcoefs;         /* [Nchains] temperatures (lambda^m, MALA.cpp:103)                      */
    double        c0, epsilon1, epsilon2, A1, target_acceptance;
    int           n_learn;        /* number of learning zones = periods_learn.size()                      */
    const long   *Nt_learn;       /* [n_learn+1]                                                          */
    const long   *periods_learn;  /* [n_learn]                                                            */
    long          dN_mixing;      /* swap attempt every dN_mixing steps (i != 0)                          */
    const double *init_err;       /* [Nv] initial std-dev per free variable (A*x+B, MALA.cpp:253-258); NULL => 1 */
    unsigned long long seed;      /* Philox key (reference: time(NULL), MALA.cpp:62)                      */
    int           device;         /* CUDA device ordinal                                                  */
    int           flags;          /* TG_F_* */
    /* Offsets added to the chain / ensemble indices in the Philox counters: ensemble e of a handle that batches many replicated
     * ensembles (chains_per_ensemble) draws the same numbers as a single-ensemble handle with philox_chain0 = e*chains_per_ensemble,
     * philox_ensemble0 = e and the same seed.  0 for ordinary use. */
    int           philox_chain0, philox_ensemble0;
} tg_problem;

#define TG_F_NO_TILE_ORDER        2 /* one work item per 1024-bin sub-tile in index order instead of heaviest-first + long noise-only items (scheduling only; results identical) */
#define TG_F_RESTORE_MEANS        4 /* keep running sums of (sigma, mu, covarmat, vars) of every step on the device: the *_mean blocks of the restore files (outputs.cpp:1758-1818), read by tg_get_state_means */
#define TG_F_FIX_SWAP_BOOKKEEPING 1 /* fix quirk Q8 (stale logPosterior[B] after a swap, MALA.cpp:417-428); default replicates */

/* Per-step records the reference hands to Outputs (MALA.cpp:696-707), for nsteps steps.
 * Any pointer may be NULL (that record is skipped).  All are HOST buffers. */
typedef struct tg_records {
    double *vars;        /* [nsteps][C][Nv]  Outputs::update_buffer_params                  */
    double *logL;        /* [nsteps][C]      tempered logLikelihood  (update_buffer_stat_criteria) */
    double *logPrior;    /* [nsteps][C]                                                    */
    double *logPost;     /* [nsteps][C]                                                    */
    double *Pmove;       /* [nsteps][C]      update_buffer_proposals                        */
    unsigned char *moved;/* [nsteps][C]                                                    */
    int    *swap_attempt;/* [nsteps][E]      tempted_mixing  (update_buffer_ptempering), E = ensembles */
    int    *swap_chain0; /* [nsteps][E]      chain_A (ensemble-local index) or -1           */
    double *swap_Pswap;  /* [nsteps][E]                                                    */
    int    *swap_done;   /* [nsteps][E]      swaped                                        */
    int    *status;      /* [nsteps][C]      TG_ST_* bits of the PROPOSED point             */
} tg_records;

const char *tg_last_error(void);
int tg_version(void);

/* Uploads the problem, allocates all device state, initialises the proposal like
 * MALA::MALA + init_proposal (MALA.cpp:53-112, :246-289) and evaluates every chain once at
 * params0 like the Model_def constructor (model_def.cpp:139-143). */
int tg_create(const tg_problem *p, tg_handle **out);
int tg_destroy(tg_handle *h);

/* Batched Model_def::generate_model (model_def.cpp:358-367 = call_model :210-289 +
 * call_likelihood :291-320 + call_prior :322-356) for B parameter vectors.
 * params [B][Np] and T [B] are HOST buffers copied in, results are copied out:
 * logL (already divided by T), logPrior, logPost [B]; status [B] (TG_ST_*), may be NULL;
 * model [B][Nx] may be NULL (parity / getmodel path, tools/getmodel.cpp:111). */
int tg_eval(tg_handle *h, const double *params, int B, const double *T,
            double *logL, double *logPrior, double *logPost, int *status, double *model);

/* The same call in two halves, through one of two slots (0, 1), so that a host-side sampler (the loop of MALA.cpp:632-655
 * with its proposal on the host) prepares the parameter vectors of step k+1 while the GPU evaluates step k: tg_eval_async copies
 * the inputs into the slot's pinned buffer, enqueues the whole call (one graph launch: copy in, kernels, copy out) and returns;
 * tg_eval_wait blocks until that slot's results have arrived and copies them out.  Each slot has its own device tables, staging
 * buffers and stream (slot 1's are made at its first use), so the copy-in, prologue and far-field pre-pass of one submission run
 * beside the fused kernel of the other; results are the bits of tg_eval.  Calls of a handle are still issued from one host thread;
 * at most one submission per slot may be outstanding, and two outstanding submissions must have the same B. */
int tg_eval_async(tg_handle *h, const double *params, int B, const double *T, int slot);
int tg_eval_wait(tg_handle *h, int slot, double *logL, double *logPrior, double *logPost, int *status);

/* tg_eval for callers that cannot allocate device memory (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.  Same arguments, meaning
 * and results as tg_eval without the model pointer (HOST buffers in and out), but the call performs no cudaMalloc / cudaFree /
 * cudaHostAlloc or any other allocating routine: the scratch of the fused kernel (argument record, work-unit table,
 * per-(vector, work item) partial sums, per-vector totals) lives in ONE file-scope __device__ array of
 * 4 KiB + TG_NOALLOC_MAX_UNITS*16 + TG_NOALLOC_MAX_PAIRS*16 + TG_NOALLOC_MAX_B*16 bytes (about 5.1 MiB) that every GPU loading
 * the library holds until the process exits; the per-vector tables and the pinned staging buffers are the ones tg_create (or an
 * earlier, larger tg_eval) allocated for the handle.  Limits, each rejected with TG_E_BADARG and a message naming it:
 * B <= the vectors the handle's tables hold (Nchains of tg_create unless a larger tg_eval grew them), B <= TG_NOALLOC_MAX_B,
 * B x work items of the spectrum <= TG_NOALLOC_MAX_PAIRS, work units of the launch <= TG_NOALLOC_MAX_UNITS.
 * The array is shared by every call on a device and keeps the last call's contents (the call overwrites the record and the table
 * and clears the sums it uses): calls that overlap on the GPU -- two handles on one device driven from two host threads -- race
 * on it; calls on one handle, or serialised by the caller, do not. */
#define TG_NOALLOC_MAX_UNITS  65536
#define TG_NOALLOC_MAX_PAIRS  262144
#define TG_NOALLOC_MAX_B      4096
int tg_eval_noalloc(tg_handle *h, const double *params, int B, const double *T,
                    double *logL, double *logPrior, double *logPost, int *status);

/* Same evaluation with the parameter vectors already resident in HBM (uploaded once by
 * tg_eval_upload); nothing crosses PCIe.  Results stay on the device until tg_eval_fetch. */
int tg_eval_upload(tg_handle *h, const double *params, int B, const double *T);
int tg_eval_resident(tg_handle *h, int B);
int tg_eval_fetch(tg_handle *h, int B, double *logL, double *logPrior, double *logPost, int *status);

/* Algorithmic work of the last evaluation: L[b] = sum over modes (2l+1)*(imax-imin)
 * (SURVEY 8(d)); K = number of active Harvey terms of chain 0. */
int tg_last_work(tg_handle *h, int B, double *L, int *K);

/* Sampler state <-> host: restore files (outputs.cpp:863-1025, MALA::restore_proposal
 * MALA.cpp:191-244, model_def.cpp:104-128).  vars [C][Nv], mu [C][Nv], sigma [C],
 * covarmat [C][Nv][Nv] (row-major per chain).  NULL pointers are skipped.
 * tg_set_state re-evaluates every chain at the new vars. */
int tg_get_state(tg_handle *h, double *vars, double *mu, double *sigma, double *covarmat,
                 double *logL, double *logPrior, double *logPost);
int tg_set_state(tg_handle *h, const double *vars, const double *mu, const double *sigma, const double *covarmat);
/* The averages the reference keeps for its restore files (Outputs::update_buffer_restore, outputs.cpp:1758-1818: sigmas_mean,
 * mus_mean, covarmats_mean, vars_mean = means over the steps since the files were last written; MALA::restore_proposal uses them
 * instead of the last values when do_restore_proposal_mean = 1, MALA.cpp:206-232).  Needs TG_F_RESTORE_MEANS in tg_problem.flags
 * (TG_E_UNSUPPORTED otherwise): the sums are accumulated on the device by the decision kernel of every step.  Returns the means
 * over the *count steps run since the last reset (count = 0: the current values); reset != 0 starts a new average.  NULL
 * pointers are skipped; layouts as tg_get_state. */
int tg_get_state_means(tg_handle *h, double *vars_mean, double *mu_mean, double *sigma_mean, double *covarmat_mean, long *count, int reset);

/* MALA::execute (MALA.cpp:608-737) for iterations i0 .. i0+nsteps-1, entirely on the device:
 * proposal (new_prop_values :335-353), generate_model, MH accept (update_position_MH :447-538),
 * Robbins-Monro update (update_proposal :292-315), tempering swap (parallel_tempering :381-445).
 * No per-step host transfer; the per-step records are drained into rec (may be NULL) at the end. */
int tg_run(tg_handle *h, long i0, long nsteps, tg_records *rec);

/* Launch group (configs[4] of BASELINE.json: several stars -- independent tempered ensembles with their own spectrum, parameters
 * and priors -- per GPU; the reference runs stars as an array of independent processes, scripts/slurm/job_array.sh:8,24 and
 * main.cpp:111-127).  The stars stay independent handles; tg_batch_run advances all of them by nsteps iterations with ONE
 * proposal, ONE fused model+likelihood and ONE decision launch per step (the work-unit table of the fused kernel spans every
 * star, costliest first).  Each star's trajectory is the same, bit for bit, as tg_run on its handle alone.  All handles on one
 * device; they must outlive the group; recs[k] (may be NULL, as may recs) receives the records of star k. */
typedef struct tg_batch tg_batch;
int tg_batch_create(tg_handle **handles, int n, tg_batch **out);
int tg_batch_run(tg_batch *b, long i0, long nsteps, tg_records **recs);
int tg_batch_destroy(tg_batch *b);
int tg_batch_sizes(tg_batch *b, int *nstars, int *Ctot, int *units);
int tg_batch_kernel_time(tg_batch *b, int reset, long *launches, double *ms_eval, long *total_launches);
int tg_batch_prepass_time(tg_batch *b, double *ms_ff);   /* total time of the far-field pre-pass launches timed together with tg_batch_kernel_time (reset = 1 clears it) */
int tg_batch_event_record(tg_batch *b, int which);   /* stopwatch on the group's stream; read with tg_event_elapsed_ms(handles[0]) */

/* Sizes implied by the problem. */
int tg_sizes(tg_handle *h, int *Nv, int *Np, int *C, int *E, long *Nx);

/* Bin-sharded variant: per-chain partial sums (sum y/m, sum ln m) of the proposed points live in
 * a device buffer of 2*C doubles; the caller all-reduces it (ncclAllReduce sum, fp64) between
 * tg_step_begin and tg_step_end.  tg_run == loop { tg_step_begin; tg_step_end }. */
int tg_step_begin(tg_handle *h, long i);
int tg_step_end(tg_handle *h, long i);
int tg_partials_device_ptr(tg_handle *h, void **ptr, int *count);
/* Bin-sharded variant only: tg_create / tg_set_state evaluate the CURRENT points on this shard's bins and leave their
 * partial sums in the same buffer; all-reduce it, then call this to install the logL / logPosterior of the current
 * points from the reduced sums (Model_def constructor, model_def.cpp:139-143). */
int tg_shard_install_current(tg_handle *h);
/* The same variant driven by the library: tg_shard_unique_id (one rank; ncclGetUniqueId) gives the 128 bytes every rank passes,
 * with its rank, to tg_shard_attach (ncclCommInitRank; all-reduces the start values like tg_shard_install_current).  From then on
 * tg_run / tg_set_state all-reduce the 2*C partial sums themselves: one ncclAllReduce (fp64 sum) per step on the handle's stream,
 * inside the captured step graph, no host synchronisation per step.  NCCL is the libnccl.so.2 already in the process (or
 * TG_NCCL_LIB).  No reference counterpart (the reference caps a spectrum at 1 000 000 rows, config.cpp:531); SURVEY 8(e)-ii. */
int tg_shard_unique_id(char *id128);
int tg_shard_attach(tg_handle *h, const char *id128, int nranks, int rank);
int tg_shard_detach(tg_handle *h);
/* The exchange of the bin-sharded variant over NVLink peer memory instead of NCCL (one process per GPU, one node, <= 8 ranks):
 * tg_shard_ipc_export allocates this rank's exchange buffer and returns its 64-byte CUDA IPC handle; every rank passes the handles
 * of ALL ranks (rank order, nranks x 64 bytes) to tg_shard_ipc_attach.  Per step ONE kernel stores the rank's 2*C partial sums
 * into every peer's buffer, waits for the peers' stamps and adds the contributions in rank order (bit-identical totals on every
 * rank); it is part of the captured step graph.  tg_shard_status reports a peer that stopped answering (the kernel gives up after
 * 2 s instead of hanging). */
int tg_shard_ipc_export(tg_handle *h, int nranks, char *handle64);
int tg_shard_ipc_attach(tg_handle *h, const char *handles, int nranks, int rank);
int tg_shard_status(tg_handle *h);
int tg_sync(tg_handle *h);

/* Kernel timing of the dominant kernel (CUDA events on the handle's stream), accumulated since
 * the last reset: number of launches and total milliseconds. */
int tg_kernel_time(tg_handle *h, int reset, long *launches, double *ms_eval, long *total_launches);
/* The far-field pre-pass (the kernel that prepares, per chain and 512-bin micro-tile under the truncation windows, the
 * interpolant of the far components and the records of the near ones for the fused kernel): total milliseconds of its launches
 * in the sampler steps timed together with tg_kernel_time. */
int tg_prepass_time(tg_handle *h, double *ms_ff);
/* Tuning of the fused kernel (never changes results; 0 / -1 = keep): chains per work unit at most, chains per
 * unit where truncation windows are expected, sub-tiles of 1024 bins per unit in the noise-only parts of the
 * spectrum, CUDA-graph replay of tg_run. */
int tg_set_tuning(tg_handle *h, int chains_per_block, int heavy_group, int light_subtiles, int use_graph);

/* How the fused kernel covers the spectrum for the sampler's chains at present (scheduling only, results never depend on it):
 * bins evaluated bin by bin, background-only tiles of mom_sub*1024 bins taken through the Chebyshev moments of the data
 * (14 exact evaluations of the background per tile and chain instead of one per bin), work units per launch. */
int tg_work_profile(tg_handle *h, long *bins_direct, long *moment_tiles, int *units);

/* What the far-field pre-pass left for the B vectors of the last tg_eval batch (measurement support, no reference counterpart):
 * out[0] micro-tiles prepared (B x micro-tiles of the window region), out[1] of them marked for the
 * component-by-component loop, out[2] exact records in all of them, out[3] far components interpolated (summed over the
 * micro-tiles), out[4] micro-tiles whose interpolant also holds the background and N0, out[5] bins per micro-tile.
 * out[0..4] are zero when the path is off.  out has six entries. */
int tg_farfield_profile(tg_handle *h, int B, long *out);

/* Device time of the steps of the last tg_run / tg_batch_run call: CUDA events recorded by the call itself on its stream, right
 * before the first launch and right after the last step (before the records are drained), so no host code of the caller lies
 * between them. */
int tg_last_run_ms(tg_handle *h, double *ms);
int tg_batch_last_run_ms(tg_batch *b, double *ms);

/* Device-side stopwatch on the handle's stream (CUDA events): which = 0 start, 1 stop. */
int tg_event_record(tg_handle *h, int which);
int tg_event_elapsed_ms(tg_handle *h, double *ms);

/* FP64 roofline denominator: DFMA-saturating microkernel, returns measured TFLOP/s (2 flop per DFMA). */
int tg_measure_fp64_peak(int device, double *tflops, double *ms);

#ifdef __cplusplus
}
#endif
#endif
