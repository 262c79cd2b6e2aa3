// cpptamcmc_gpu -- the reference's `cpptamcmc` program (tamcmc/sources/main.cpp) on the CUDA hot path.
//
//   cpptamcmc_gpu version
//   cpptamcmc_gpu execute 0
//   cpptamcmc_gpu execute 1 <first star> <last star> [first slice] [last slice]        (1-based, as main.cpp:44-58)
//
// Run it from a directory laid out like the reference's repository: Config/default/{config_default.cfg,
// errors_default.cfg, *_ctrl.list} and Config/config_presets.cfg.  For every star of table_ids, every frequency slice
// of its .model file and every phase between start_index_processing and last_index_processing it does what
// main.cpp:104-176 does: apply the presets, read the .model/.data files, build Model_def + MALA, run the sampler and
// write the reference's output files (params / stat_criteria / parallel_tempering binaries with their headers, the
// restore triplet, the acceptance-rate and evidence text files).  All of the sampling runs on the GPU through
// libtamcmc_gpu.so (include/tamcmc_gpu.hpp); this file only orchestrates.  Not done here, on purpose: gnuplot
// diagrams, zip backups of the inputs, the `proposals` / `models` output files.
//
// Environment: TAMCMC_SEED (default: time, as the reference), TAMCMC_DEVICE (default 0).
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <ctime>
#include <string>
#include <vector>
#include <unistd.h>
#include "tamcmc_frontend.h"
#include "tamcmc_gpu.hpp"
#include "tamcmc_outputs.h"

using tamcmc_gpu::check;

static void fe_check(int rc) { if (rc != TG_OK) throw tamcmc_gpu::Error(rc, std::string("front end: ") + tg_fe_last_error()); }
static void out_check(int rc) { if (rc != TG_OK) throw tamcmc_gpu::Error(rc, std::string("outputs: ") + tg_out_last_error()); }

static int usage()
{
    printf("Usage:\n  cpptamcmc_gpu version\n  cpptamcmc_gpu execute 0\n  cpptamcmc_gpu execute 1 <first star> <last star> [first slice] [last slice]\n"
           "  (indices start at 1 and follow table_ids of Config/config_presets.cfg; without slice indices every slice of the .model file is processed)\n");
    return 2;
}

// one phase of one slice of one star: main.cpp:139-176 + MALA::execute (MALA.cpp:540-739)
static void run_phase(tg_frontend *fe, int star, int slice, int phase, unsigned long long seed, int device)
{
    fe_check(tg_fe_setup(fe, star, slice, phase));
    tg_problem prob; tg_fe_run run;
    fe_check(tg_fe_problem(fe, &prob));
    fe_check(tg_fe_run_info(fe, &run));
    prob.seed = seed; prob.device = device;
    prob.flags |= TG_F_RESTORE_MEANS;      // the restore files carry the averages over each buffer (Outputs::update_buffer_restore, outputs.cpp:1758-1818)
    printf(" - Model: %s   likelihood: %s   priors: %s\n", run.model_fct_name, run.likelihood_fct_name, run.prior_fct_name);
    printf(" - Ndata=%ld  Nparams=%d  Nvars=%d  Nchains=%d  Nsamples=%ld  c0=%g  dN_mixing=%ld\n", prob.Nx, prob.Np, run.Nvars, prob.Nchains, run.Nsamples, prob.c0, prob.dN_mixing);

    tamcmc_gpu::Model_def model(prob);
    const int C = model.Nmodels, Nv = model.Nvars, E = model.Nensembles;
    long i0 = 0;
    if (run.do_restore_variables || run.do_restore_proposal) {       // Config::read_restore_files + model_def.cpp:104-128 + MALA::restore_proposal
        const std::string root = std::string(run.restore_dir) + run.restore_file_in;
        int rc_ = 0, rv = 0; long it = 0;
        out_check(tg_restore_read(root.c_str(), run.do_restore_last_index, &rc_, &rv, &it, nullptr, nullptr, nullptr, nullptr));
        if (rc_ != C || rv != Nv) throw tamcmc_gpu::Error(TG_E_BADARG, "restore files describe " + std::to_string(rc_) + " chains x " + std::to_string(rv) + " variables, the configuration " + std::to_string(C) + " x " + std::to_string(Nv));
        std::vector<double> v((size_t)C * Nv), mu((size_t)C * Nv), sg((size_t)C), cov((size_t)C * Nv * Nv);
        out_check(tg_restore_read(root.c_str(), run.do_restore_last_index, &rc_, &rv, &it, v.data(), mu.data(), sg.data(), cov.data()));
        const bool rp = run.do_restore_proposal != 0;
        if (rp && run.do_restore_proposal_mean) {                      // MALA::restore_proposal, MALA.cpp:206-232: the averaged proposal law instead of the last one
            std::vector<double> vm((size_t)C * Nv);
            out_check(tg_restore_read_means(root.c_str(), run.do_restore_last_index, &rc_, &rv, &it, vm.data(), mu.data(), sg.data(), cov.data()));
        }
        model.set_state(run.do_restore_variables ? v.data() : nullptr, rp ? mu.data() : nullptr, rp ? sg.data() : nullptr, rp ? cov.data() : nullptr);
        if (run.do_restore_last_index) i0 = it;
        printf(" - Restored %s%s%s from %s*.dat (iteration %ld)\n", run.do_restore_variables ? "variables " : "", rp ? "proposal " : "", (rp && run.do_restore_proposal_mean) ? "(averaged over the last buffer) " : "", root.c_str(), it);
    }
    {   // the consistency print of the Model_def constructor (model_def.cpp:139-150)
        std::vector<double> l((size_t)C), p((size_t)C), q((size_t)C);
        model.get_state(nullptr, nullptr, nullptr, nullptr, l.data(), p.data(), q.data());
        printf("      - logLikelihood[0]=%g\n      - logPrior[0]=%g\n      - logPosterior[0]=%g\n", l[0], p[0], q[0]);
    }

    // ---- Outputs (outputs.cpp:27-150): file names as the reference builds them ----
    const std::string base = std::string(run.dir_out) + run.output_root_name;
    const std::string f_params = base + run.params_txt_fileout, f_stat = base + run.stat_txt_fileout, f_pt = base + run.parallel_tempering_txt_fileout;
    const std::string f_restore = std::string(run.restore_dir) + run.restore_file_out, f_acc = base + run.acceptance_txt_fileout + ".txt";
    const std::string f_evid = std::string(run.diags_dir) + run.output_root_name + run.file_evidence_diags + ".txt";
    std::vector<std::string> vn, cn; std::vector<double> cv;
    for (int i = 0; i < prob.Np; i++) { if (prob.relax[i] == 1) vn.push_back(tg_fe_input_name(fe, i)); else { cn.push_back(tg_fe_input_name(fe, i)); cv.push_back(prob.params0[i]); } }
    std::vector<const char *> vnp, cnp; for (auto &s : vn) vnp.push_back(s.c_str()); for (auto &s : cn) cnp.push_back(s.c_str());
    tg_out_desc d; memset(&d, 0, sizeof d);
    d.params_root = run.get_params ? f_params.c_str() : nullptr; d.stat_root = run.get_statcriteria ? f_stat.c_str() : nullptr;
    d.ptemp_root = run.get_parallel_tempering ? f_pt.c_str() : nullptr; d.restore_root = f_restore.c_str();
    d.Nsamples = run.Nsamples; d.Nsamples_sofar = i0; d.erase_old_files = run.erase_old_files; d.file_format = run.file_format;
    d.Nchains = C; d.Nvars = Nv; d.Ncons = (int)cn.size(); d.Np = prob.Np; d.n_plength = 11; d.relax = prob.relax; d.plength = prob.plength;
    d.variable_names = vnp.data(); d.constant_names = cnp.empty() ? nullptr : cnp.data(); d.constant_values = cv.data(); d.Tcoefs = prob.Tcoefs;
    tg_out *out = nullptr;
    out_check(tg_out_open(&d, &out));

    tamcmc_gpu::MALA mala(model);
    const long Nbuffer = run.Nbuffer > 0 ? run.Nbuffer : run.Nsamples;
    std::vector<long double> sumL((size_t)C, 0.0L);                    // running sums of the tempered logL of every chain, for the evidence
    long nL = 0;                                                       // samples in them (those of this process: a resumed run averages what it computed)
    std::vector<double> v((size_t)C * Nv), mu((size_t)C * Nv), sg((size_t)C), cov((size_t)C * Nv * Nv), rate((size_t)C), Lb((size_t)C);
    std::vector<double> vm((size_t)C * Nv), mum((size_t)C * Nv), sgm((size_t)C), covm((size_t)C * Nv * Nv);
    model.get_state_means(nullptr, nullptr, nullptr, nullptr, true);          // the averages start with this phase
    long done = i0, ncopy = 0;
    const clock_t t0 = clock(); time_t w0 = time(nullptr);
    while (done < run.Nsamples) {
        const long n = std::min(Nbuffer, run.Nsamples - done);
        tamcmc_gpu::Records rec(n, C, E, Nv);
        mala.execute(done, n, &rec);
        out_check(tg_out_append(out, &rec.view, n, E));
        model.get_state(v.data(), mu.data(), sg.data(), cov.data());
        model.get_state_means(vm.data(), mum.data(), sgm.data(), covm.data(), true);
        out_check(tg_out_write_restore_means(out, done + n, v.data(), mu.data(), sg.data(), cov.data(), vm.data(), mum.data(), sgm.data(), covm.data()));
        // acceptance rates of the block (Outputs::reject_rate + write_txt_acceptance, outputs.cpp:747-789, :1824-1843)
        double xaxis = 0;
        fe_check(tg_acceptance(rec.moved.data(), n, C, Nbuffer, ncopy, i0, rate.data(), &xaxis));
        {
            const bool fresh = (ncopy == 0 && run.erase_old_files);
            FILE *f = fopen(f_acc.c_str(), fresh ? "w" : "a");
            if (!f) throw tamcmc_gpu::Error(TG_E_BADARG, "cannot open " + f_acc);
            if (fresh) fprintf(f, "# This is an output file for the acceptance rate. \n# This file contains values for the acceptance_rate[0:Nchains-1] in function of the average sample position\n# Averaging is done over Nbuffer \n! Nchains= %d\n", C);
            fprintf(f, "%g ", xaxis); for (int m = 0; m < C; m++) fprintf(f, "%g ", rate[(size_t)m]); fprintf(f, "\n");
            fclose(f);
        }
        // evidence from every sample so far (Diagnostics::evidence_calc + write_evidence, diagnostics.cpp:218-224, :924-1001)
        for (long k = 0; k < n; k++) for (int m = 0; m < C; m++) sumL[(size_t)m] += (long double)rec.logL[(size_t)k * C + m];
        nL += n;
        if (run.evidence_diags) {
            double ev = 0;
            std::vector<double> meanL((size_t)C);
            for (int m = 0; m < C; m++) meanL[(size_t)m] = (double)(sumL[(size_t)m] / (long double)nL);
            fe_check(tg_evidence(prob.Tcoefs, C, meanL.data(), 1, run.evidence_interpolation_factor, Lb.data(), &ev));   // L_beta = the means; O(C) per buffer
            const bool fresh = (ncopy == 0);
            FILE *f = fopen(f_evid.c_str(), fresh ? "w" : "a");
            if (!f) throw tamcmc_gpu::Error(TG_E_BADARG, "cannot open " + f_evid);
            if (fresh) {
                fprintf(f, "# This is an output file for the evidence. Evidence is calculated after a quadratic interpolation of L_beta. \n# This file contains values for the L_beta[0:Nchains-1], the evidence calculated at each time the buffer was written \n"
                           "# col(1): Number of samples used to compute the evidence \n# col(2:2+Nchains): averaged probability <P(D|M,I)> over the samples of each chain \n# col(2+Nchains+1): Evidence P(M|D, I) computed by (1) interpolation and (2) averaging \n! beta=");
                for (int m = 0; m < C; m++) fprintf(f, "%g ", 1. / prob.Tcoefs[m]);
                fprintf(f, "\n! interpolation_factor=%d\n", run.evidence_interpolation_factor);
            }
            fprintf(f, "%ld ", nL - 1); for (int m = 0; m < C; m++) fprintf(f, "%20.10g ", Lb[(size_t)m]); fprintf(f, " %20.10g\n", ev);
            fclose(f);
        }
        done += n; ncopy++;
        printf("[%ld]  %ld samples written; acceptance of the coldest chain over the block: %.3f\n", done - 1, done - i0, rate[0] * (double)Nbuffer / (double)n);
    }
    out_check(tg_out_close(out));
    const double wall = difftime(time(nullptr), w0);
    printf(" ----------------------------\n    Calculation finished in: %g s (cpu %g s), %.0f chain-steps/s\n ----------------------------\n", wall, (double)(clock() - t0) / CLOCKS_PER_SEC,
           wall > 0 ? (double)(run.Nsamples - i0) * C / wall : 0.0);
}

int main(int argc, char **argv)
{
    if (argc == 2 && std::string(argv[1]) == "version") { printf("cpptamcmc_gpu (B200 hot path of TAMCMC) library version %d\n", tg_version()); return 0; }
    if (argc < 3 || std::string(argv[1]) != "execute") return usage();
    const bool execute = std::string(argv[2]) != "0";
    if (execute && argc < 5) return usage();
    if (argc > 7) return usage();
    int startV = -1, lastV = -1, startslice = -1, lastslice = -1;
    if (argc >= 5) { startV = atoi(argv[3]) - 1; lastV = atoi(argv[4]) - 1; }
    if (argc >= 6) startslice = atoi(argv[5]) - 1;
    if (argc >= 7) lastslice = atoi(argv[6]) - 1;
    char cwd[4096];
    if (!getcwd(cwd, sizeof cwd)) { perror("getcwd"); return 1; }
    printf(" --------- TAMCMC (B200 hot path) ----------\n- Current directory: %s\n- Loading the default configurations and Config/config_presets.cfg...\n", cwd);
    tg_frontend *fe = nullptr;
    try {
        fe_check(tg_fe_open(cwd, &fe));
        if (!execute) { printf("Detection of execute=0... The code will exit now\n"); tg_fe_close(fe); return 0; }
        int n_stars = 0, n_phases = 0, first_phase = 0, last_phase = 0;
        fe_check(tg_fe_counts(fe, &n_stars, &n_phases, &first_phase, &last_phase));
        if (last_phase > n_phases - 1) last_phase = n_phases - 1;
        if (lastV > n_stars - 1) lastV = n_stars - 1;
        const unsigned long long seed = getenv("TAMCMC_SEED") ? strtoull(getenv("TAMCMC_SEED"), nullptr, 10) : (unsigned long long)time(nullptr);
        const int device = getenv("TAMCMC_DEVICE") ? atoi(getenv("TAMCMC_DEVICE")) : 0;
        for (int i = startV; i <= lastV; i++) {
            if (i < 0) continue;
            int Nslices = 1;
            fe_check(tg_fe_nslices(fe, i, &Nslices));
            // main.cpp:116-126: without slice arguments every slice; the last slice index is exclusive there, kept so
            const int s0 = startslice < 0 ? 0 : startslice;
            const int s1 = (lastslice < 0 || (lastslice < startslice && lastslice > 0)) ? Nslices : lastslice;
            for (int s = s0; s < s1 && s < Nslices; s++)
                for (int ph = first_phase; ph <= last_phase; ph++) {
                    printf("---------------------------------------------------------------------------------------\n"
                           "                       Processing Object %d/%d   Frequency Slice %d/%d   Phase %d/%d\n", i + 1, n_stars, s + 1, Nslices, ph + 1, n_phases);
                    run_phase(fe, i, s, ph, seed + 1000003ull * (unsigned long long)(i * 64 + s) + (unsigned long long)ph, device);
                }
        }
        tg_fe_close(fe);
    } catch (const tamcmc_gpu::Error &e) {
        printf("%s\nThe program will exit now\n", e.what());
        if (fe) tg_fe_close(fe);
        return e.code == TG_E_NODEVICE ? 3 : 1;
    }
    return 0;
}
