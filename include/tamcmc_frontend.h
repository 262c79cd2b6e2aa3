/* tamcmc_frontend.h -- the reference's input front end for the hot path.
 *
 * Plain C interface of the host code in csrc/tg_frontend.cpp (no CUDA, no Eigen, no Boost).  It reads the
 * reference's own files, unchanged, from a working directory laid out like the reference's repository:
 *   Config/default/config_default.cfg    Config::read_cfg_file            sources/config.cpp:809-1319
 *   Config/default/errors_default.cfg    Config::read_defautlerrors       sources/config.cpp:1611-1662
 *   Config/default/{models,priors,likelihoods,primepriors}_ctrl.list      Config::read_listfiles
 *   Config/config_presets.cfg            Config_presets                   sources/config_presets.cpp:22-470
 *   <cfg_models_dir>/<id>.model          read_MCMC_file_MS_Global / _local, build_init_MS_Global / _local
 *                                        sources/io_ms_global.cpp:25-1090, sources/io_local.cpp:26-1158
 *   <cfg_models_dir>/<id>.data           Config::read_data_ascii_Ncols    sources/config.cpp:509-620
 * and leaves behind exactly what Config::setup (config.cpp:80-189) + MALA::MALA / init_proposal
 * (MALA.cpp:53-112, :246-289) hand to Model_def and MALA: a tg_problem (include/tamcmc_gpu.h) plus the run
 * description (phase, samples, output and restore names).  tools/cpptamcmc_gpu.cpp is the `cpptamcmc execute`
 * program built on it.  Reference failures (exit()) become TG_E_BADARG with a message in tg_fe_last_error().
 */
#ifndef TAMCMC_FRONTEND_H
#define TAMCMC_FRONTEND_H
#include "tamcmc_gpu.h"
#ifdef __cplusplus
extern "C" {
#endif

typedef struct tg_frontend tg_frontend;

/* What main.cpp / Outputs / Diagnostics read from Config besides the problem itself. */
typedef struct tg_fe_run {
    const char *star_id;            /* table_ids[star][0]                                                  */
    const char *phase;              /* processing[phase]: "Burn-in" | "Learning" | "Acquire"               */
    const char *model_fct_name, *likelihood_fct_name, *prior_fct_name;
    const char *dir_out, *restore_dir, *diags_dir;       /* <cfg_out_dir>/<id>/{outputs,restore,diags}/     */
    const char *output_root_name;   /* "<id>_[<slice>_]<core_out>_"                                        */
    const char *restore_file_in, *restore_file_out;      /* "<id>_[<slice>_]restore_<core>_"               */
    const char *params_txt_fileout, *stat_txt_fileout, *parallel_tempering_txt_fileout, *acceptance_txt_fileout;
    const char *file_evidence_diags;
    long  Nsamples, Nbuffer;
    int   erase_old_files, do_restore_proposal, do_restore_variables, do_restore_last_index, do_restore_proposal_mean;
    int   get_params, get_statcriteria, get_parallel_tempering, evidence_diags, evidence_interpolation_factor;
    int   Nslices, Nvars, Ncons;
    double lambda_temp;
    int   file_format;              /* !Outputs file_format: 0 binary, 1 text, 2 debug (tg_out_desc.file_format)   */
} tg_fe_run;

const char *tg_fe_last_error(void);
/* Config::Config + Config_presets::Config_presets (main.cpp:88-93) from <workdir>/Config/... */
int  tg_fe_open(const char *workdir, tg_frontend **fe);
void tg_fe_close(tg_frontend *fe);
/* stars (table_ids rows), phases (processing entries), start/last_index_processing of config_presets.cfg */
int  tg_fe_counts(const tg_frontend *fe, int *n_stars, int *n_phases, int *first_phase, int *last_phase);
/* get_slices_range (main.cpp:379-441): number of '*' range lines of <id>.model */
int  tg_fe_nslices(tg_frontend *fe, int star, int *nslices);
/* Config_presets::apply_presets + Config::setup(slice) + the MALA constructor's temperatures and initial errors
 * for star / slice / phase (all 0-based).  Creates the output directory tree like generate_default_dirtree. */
int  tg_fe_setup(tg_frontend *fe, int star, int slice, int phase);
/* The problem of the last tg_fe_setup; every pointer stays owned by fe (valid until the next setup / close).
 * seed, device, flags, chains_per_ensemble and the bin-shard fields are left 0 for the caller. */
int  tg_fe_problem(const tg_frontend *fe, tg_problem *p);
int  tg_fe_run_info(const tg_frontend *fe, tg_fe_run *r);
/* names of parameter i (0 <= i < Np): Input_Data.inputs_names / priors_names */
const char *tg_fe_input_name(const tg_frontend *fe, int i);
const char *tg_fe_prior_name(const tg_frontend *fe, int i);
/* "key: values" lines (17 significant digits) of everything above, the format of oracle/ref_shim/ref_frontend_dump.cpp.
 * Returns the number of bytes needed (excluding the terminator); writes at most cap bytes. */
long tg_fe_dump(const tg_frontend *fe, char *buf, long cap);

/* ---- numeric diagnostics ----
 * Diagnostics::evidence_calc (diagnostics.cpp:924-956): beta = 1/T, L_beta[m] = mean over samples of the (tempered)
 * logL of chain m, both resampled to interpolation_factor*C points by quad_interpol (interpol.cpp:58-103), evidence =
 * mean of the resampled L_beta.  logL is [nsamples][C].  L_beta [C] may be NULL. */
int tg_evidence(const double *Tcoefs, int C, const double *logL, long nsamples, int interpolation_factor, double *L_beta, double *evidence);
/* Outputs::reject_rate (outputs.cpp:1824-1843): accepted moves of each chain over a block of nsteps moves (moved is
 * [nsteps][C]) divided by Nbuffer -- the reference divides by Nbuffer even when the block is shorter -- and the x-axis
 * value (Ncopy+0.5)*Nbuffer + Nsamples_init it writes beside it. */
int tg_acceptance(const unsigned char *moved, long nsteps, int C, long Nbuffer, long Ncopy, long Nsamples_init, double *rate, double *xaxis);

#ifdef __cplusplus
}
#endif
#endif
