/* tamcmc_outputs.h -- on-disk formats of the reference for the records of the hot path.
 *
 * Plain C interface of libtamcmc_host.so (host code only, no CUDA).  Writers follow the reference's
 * Outputs class byte for byte for the binary files and keyword for keyword for the ASCII ones
 * (all paths relative to /root/reference/tamcmc/sources):
 *   <params_root>_chain-<m>.bin + <params_root>.hdr   outputs.cpp:1231-1333  write_bin_params
 *   <stat_root>.bin + .hdr                            outputs.cpp:1472-1540  write_bin_stat_criteria
 *   <ptemp_root>.bin + .hdr                           outputs.cpp:1336-1404  write_bin_parallel_temp_params
 *   <restore_root>{1,2,3}.dat                         outputs.cpp:863-1025   write_buffer_restore
 *   the ASCII variants (file_format = text / debug)   outputs.cpp:510-613, :615-676, :791-860
 * and the reader of the restore triplet follows config.cpp:1321-1579 (Config::read_restore_files).
 * Consumers in the reference: tools/bin2txt_params.cpp, tools/read_stats.cpp, diagnostics.cpp:677-922.
 */
#ifndef TAMCMC_OUTPUTS_H
#define TAMCMC_OUTPUTS_H
#include "tamcmc_gpu.h"
#ifdef __cplusplus
extern "C" {
#endif

typedef struct tg_out tg_out;

typedef struct tg_out_desc {
    const char *params_root;    /* Outputs::params_txtbin_fileout; NULL: not written                          */
    const char *stat_root;      /* Outputs::stat_criteria_txtbin_fileout; NULL: not written                   */
    const char *ptemp_root;     /* Outputs::parallel_tempering_txtbin_fileout; NULL: not written              */
    const char *restore_root;   /* Outputs::restore_file_out ("<dir>/restore_"); NULL: not written            */
    long  Nsamples;             /* samples planned for the phase (header field only)                          */
    long  Nsamples_sofar;       /* samples already on disk (buf_restore.Nsamples_sofar; 0 for a new run)      */
    int   erase_old_files;      /* 1: new files and headers; 0: append to the data files, headers untouched (outputs.cpp:1254) */
    int   Nchains, Nvars, Ncons, Np, n_plength;
    const int    *relax;        /* [Np]                                                                       */
    const int    *plength;      /* [n_plength]                                                                */
    const char *const *variable_names;  /* [Nvars]                                                            */
    const char *const *constant_names;  /* [Ncons]; NULL or Ncons == 0: "None" / -1 as the reference writes   */
    const double *constant_values;      /* [Ncons]                                                            */
    const double *Tcoefs;       /* [Nchains]                                                                  */
    int   file_format;          /* Outputs::file_format (outputs.cpp:38-56): 0 "binary" (.bin + .hdr), 1 "text" (.txt with the header
                                 * inside, write_txt_params :510-613, write_txt_stat_criteria :791-860, write_txt_parallel_temp_params
                                 * :615-676), 2 "debug" (the binary files with extension .dbg and the text ones as .dbg.txt)          */
} tg_out_desc;

int tg_out_open(const tg_out_desc *d, tg_out **out);
/* Appends nsteps samples of a tg_records block (as filled by tg_run; rec->swap_* are taken from ensemble 0,
 * the reference has one ensemble) and rewrites the headers with the new Nsamples_done. */
int tg_out_append(tg_out *o, const tg_records *rec, long nsteps, int n_ensembles);
/* Restore triplet at `iteration` (= samples done so far): vars[C][Nv], mu[C][Nv], sigma[C], covarmat[C][Nv][Nv].
 * The *_mean blocks hold the mean of vars over the last appended block and the current mu/sigma/covarmat.
 * Values are written with 17 significant digits (the reference's stream default keeps 6: a restart from its
 * own files is not exact; its reader accepts any precision). */
int tg_out_write_restore(tg_out *o, long iteration, const double *vars, const double *mu, const double *sigma, const double *covarmat);
/* The same with the averages of the reference's restore files (buf_restore.*_mean, outputs.cpp:1758-1818: means of sigma, mu,
 * covarmat and vars over the steps since the files were last written -- tg_get_state_means gives them).  A NULL mean pointer
 * falls back to what tg_out_write_restore writes in that block. */
int tg_out_write_restore_means(tg_out *o, long iteration, const double *vars, const double *mu, const double *sigma, const double *covarmat,
                               const double *vars_mean, const double *mu_mean, const double *sigma_mean, const double *covarmat_mean);
int tg_out_close(tg_out *o);

/* Reader of the restore triplet.  First call with all arrays NULL to get the sizes.  do_restore_last_index == 0
 * forces iteration = 0 (config.cpp:1370-1378).  Returns TG_OK, TG_E_BADARG (missing file / keyword). */
int tg_restore_read(const char *restore_root, int do_restore_last_index, int *Nchains, int *Nvars, long *iteration,
                    double *vars, double *mu, double *sigma, double *covarmat);
/* The *_mean blocks of the same files (what MALA::restore_proposal installs when do_restore_proposal_mean = 1, MALA.cpp:206-232). */
int tg_restore_read_means(const char *restore_root, int do_restore_last_index, int *Nchains, int *Nvars, long *iteration,
                          double *vars_mean, double *mu_mean, double *sigma_mean, double *covarmat_mean);
const char *tg_out_last_error(void);

#ifdef __cplusplus
}
#endif
#endif
