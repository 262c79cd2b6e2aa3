"""Writers for the reference's input files: a ``Config/`` tree (config_default.cfg, errors_default.cfg, the four
``*_ctrl.list`` registries, config_presets.cfg) and ``<id>.model`` / ``<id>.data`` files, in the formats
``cpptamcmc`` reads (tamcmc/sources/config.cpp, config_presets.cpp, io_ms_global.cpp, io_local.cpp).

Harness only: the tests and the CLI smoke run need a working directory on machines that do not hold the reference's
repository.  The keyword sets and the numeric defaults are the reference's (Config/default/*); the files written here
carry no comment text of the reference."""
import os

import numpy as np

MODELS = ["Test_Gaussian", "model_Harvey_Gaussian", "model_MS_Global_a1etaa3_HarveyLike", "model_MS_Global_a1etaa3_HarveyLike_Classic",
          "model_MS_Global_a1etaa3_Harvey1985", "model_MS_Global_a1acta3_Harvey1985", "model_MS_Global_a1l_etaa3_HarveyLike",
          "model_MS_Global_a1n_etaa3_HarveyLike", "model_MS_Global_a1nl_etaa3_HarveyLike", "model_MS_Global_a1etaa3_AppWidth_HarveyLike_v1",
          "model_MS_Global_a1etaa3_AppWidth_HarveyLike_v2", "model_MS_local_basic", "model_MS_Global_a1etaa3_HarveyLike_Classic_v2",
          "model_MS_Global_a1etaa3_HarveyLike_Classic_v3", "model_MS_local_Hnlm"]
PRIOR_FCTS = ["priors_Test_Gaussian", "priors_Harvey_Gaussian", "io_MS_Global", "io_local"]
LIKELIHOODS = ["chi(2,2p)", "chi_square"]
PRIME_PRIORS = [(0, "None"), (0, "Fix"), (1, "Uniform"), (2, "Gaussian"), (3, "multivar_Gaussian"), (4, "Jeffreys"), (5, "UG"), (6, "GU"),
                (7, "GUG"), (8, "Uniform_abs"), (9, "Uniform_cos"), (10, "Jeffreys_abs"), (11, "Auto")]
# name, A, B : initial error of a free variable = A * value + B (errors_default.cfg)
ERRORS = [("Numax", 0.01, 20), ("Gsigma", 0.01, 10), ("Hmax", 0.1, 0), ("Frequency_l", 0.0, 0.07), ("Height_l", 0.02, 0.01), ("Amplitude_l", 0.01, 0.01),
          ("Width_l", 0.015, 0.005), ("Visibility_l1", 0, 0.1), ("Visibility_l2", 0, 0.05), ("Visibility_l3", 0, 0.05), ("Splitting_a1", 0.02, 0.),
          ("sqrt(splitting_a1).cosi", 0.1, 0.05), ("sqrt(Splitting_a1).sini", 0.1, 0.05), ("Asphericity_eta", 0.2, 0.000001), ("Splitting_a3", 0.15, 0.005),
          ("Asphericity_mag_b", 0.05, 0.05), ("Asphericity_mag_alfa", 0.0, 0.10), ("Lorentzian_asymetry", 0.02, 1.0), ("Inclination", 0.0, 5.0),
          ("Harvey-Noise_H", 0.03, 0.0), ("Harvey-Noise_tc", 0.015, 0.0), ("Harvey-Noise_p", 0.015, 0.0), ("White_Noise_N0", 0.015, 0.0002)]


def write_config_tree(workdir, models_dir, out_dir, ids, prior_fct="io_local", phases=("Burn-in", "Learning", "Acquire"), Nsamples=(200, 300, 400),
                      c0=(1.8, 1.7, 0.0), restore=(0, 1, 2), first_phase=0, last_phase=0, Nchains=10, Nbuffer=10000, Nt_learn=(500, 1500, 100000),
                      dN_mixing=1, lambda_temp=1.7):
    """Config/default/* and Config/config_presets.cfg under ``workdir``; ``ids`` are the table_ids (file stems in models_dir)."""
    d = os.path.join(workdir, "Config", "default")
    os.makedirs(d, exist_ok=True)
    os.makedirs(os.path.join(workdir, "tmp"), exist_ok=True)
    os.makedirs(out_dir, exist_ok=True)
    kv = lambda pairs: "".join("\t%s=%s;\n" % (k, v) for k, v in pairs)
    cfg = "# default configuration (written by tamcmc_b200.config_tree)\n!MALA:\n" + kv([
        ("target_acceptance", 0.234), ("c0", 10), ("epsilon1", 1e-12), ("epsilon2", 1e-12), ("A1", 1e14), ("Nt_learn", ", ".join(str(v) for v in Nt_learn)),
        ("periods_learn", ", ".join("1" for _ in Nt_learn[1:])), ("use_drift", 0), ("delta", 0), ("delta_x", 1e-10), ("Nchains", Nchains), ("dN_mixing", dN_mixing),
        ("lambda_temp", lambda_temp), ("proposal_type", "Random")])
    cfg += "!Modeling:\n" + kv([("prior_fct_name", prior_fct), ("model_fct_name", prior_fct), ("likelihood_fct_name", "chi(2,2p)"), ("likelihood_params", 1)])
    cfg += "!Data:\n" + kv([("verbose_data", 0), ("x_col", 0), ("y_col", 1), ("ysig_col", -1)])
    cfg += "!Outputs:\n" + kv([
        ("Nsamples", 150), ("Nbuffer", Nbuffer), ("erase_old_files", 1), ("file_format", "binary"), ("get_params", 1), ("get_statcriteria", 1), ("get_proposal_params", 0),
        ("get_parallel_tempering", 1), ("get_models", 0), ("output_root_name", ""), ("output_dir", ""), ("params_txt_fileout", "params"), ("proposal_txt_fileout", "proposals"),
        ("parallel_tempering_txt_fileout", "parallel_tempering"), ("model_txt_fileout", "models"), ("stat_txt_fileout", "stat_criteria"), ("acceptance_txt_fileout", "acceptance"),
        ("do_restore_last_index", 0), ("do_restore_proposal", 0), ("do_restore_variables", 0), ("do_restore_proposal_mean", 0), ("restore_dir", ""), ("restore_file_in", "restore_"),
        ("restore_file_out", "restore_"), ("do_backup_input_file", 0), ("do_backup_cfg_files", 0)])
    cfg += "!Diagnostics:\n" + kv([
        ("chains_diags", 1), ("evidence_diags", 1), ("pdfs_diags", 0), ("model_initial_diags", 0), ("model_buffer_diags", 0), ("model_final_diags", 0), ("output_root_name", ""),
        ("output_dir", ""), ("file_chains_diags", "acceptance"), ("file_evidence_diags", "evidence"), ("file_pdfs_diags", "pdfs_plots"), ("file_model_init_diags", "model_init"),
        ("file_model_buffer_diags", "model_buffer"), ("file_model_final_diags", "model_final"), ("show_original_data", 0), ("data_scoef1", 0.1), ("data_scoef2", 0.6), ("Nclasses", 200),
        ("evidence_interpolation_factor", 1000)]) + "/END;\n"
    open(os.path.join(d, "config_default.cfg"), "w").write(cfg)
    open(os.path.join(d, "errors_default.cfg"), "w").write("# name  A  B : initial error = A*value + B\n" + "".join("%s\t%g\t%g\n" % e for e in ERRORS))
    lst = lambda rows: "# case index, name\n" + "".join("%d %s\n" % r for r in rows)
    open(os.path.join(d, "models_ctrl.list"), "w").write(lst(list(enumerate(MODELS))))
    open(os.path.join(d, "priors_ctrl.list"), "w").write(lst(list(enumerate(PRIOR_FCTS))))
    open(os.path.join(d, "likelihoods_ctrl.list"), "w").write(lst(list(enumerate(LIKELIHOODS))))
    open(os.path.join(d, "primepriors_ctrl.list"), "w").write(lst(PRIME_PRIORS))
    core = {"Burn-in": "B", "Learning": "L", "Acquire": "A"}
    cores = [core[p] for p in phases]
    j = lambda v: " , ".join(str(a) for a in v)
    md = models_dir if models_dir.endswith("/") else models_dir + "/"
    pre = "# presets (written by tamcmc_b200.config_tree)\n   force_manual_config=0;\n   manual_config_file=;\n   cfg_models_dir=%s;\n   cfg_out_dir=%s;\n" % (md, out_dir.rstrip("/"))
    pre += "   processing = %s;\n   Nsamples = %s;\n   c0 = %s;\n   restore = %s;\n   core_out = %s;\n   core_in = %s;\n" % (
        j(phases), j(Nsamples), j(c0), j(restore), j(cores), j([cores[0]] + cores[:-1]))
    pre += "   start_index_processing=%d;\n   last_index_processing=%d;\n   table_ids=%d, 2;\n" % (first_phase, last_phase, len(ids)) + "".join("%s   1;\n" % i for i in ids) + "/END;\n"
    open(os.path.join(workdir, "Config", "config_presets.cfg"), "w").write(pre)


def write_data_file(path, x, y):
    """``.data`` file: '#' header, '!' labels, '*' units, then two columns (config.cpp:509-620)."""
    with open(path, "w") as f:
        f.write("# synthetic power spectrum\n!            frequency               power\n*            (microHz)     (ppm^2/microHz)\n")
        np.savetxt(f, np.column_stack([x, y]), fmt="%19.8f")


def write_model_file(path, modes, ranges, model_fullname, Dnu, C_l=0.0, noise=((0, 0, 1), (0, 0, 1), (0, 0, 1)), N0=1.0, kic="000000001",
                     a1=1.0, inclination=45.0, trunc_c=None, height_prior=("Jeffreys", 1.0, 1.0, 10000.0), extra_common=()):
    """``.model`` file of the io_MS_Global / io_local family (io_ms_global.cpp:25-306).

    modes: rows (l, nu, nu_min, nu_max, Gamma, H); ranges: list of (lo, hi) -- one for global models, one per slice for
    local ones; noise: three (H, tc, p) Harvey terms (all-zero terms are written as 0 0 1)."""
    modes = sorted(modes, key=lambda m: (m[0], m[1]))
    with open(path, "w") as f:
        f.write("#KIC =%s\n! %.6f\n!! %.6f\n" % (kic, Dnu, C_l))
        for lo, hi in ranges:
            f.write("* %.4f %.4f\n" % (lo, hi))
        f.write("# type / l / value / relax_f / relax_H / relax_W\n")
        for m in modes:
            f.write("p  %d  %.5f  1      1       1\n" % (m[0], m[1]))
        f.write("# hyper priors\n# extra parameters\n0.0\n")
        f.write("# eigen solutions: l / nu / window_min / window_max / Gamma / H\n")
        for m in modes:
            f.write("    %d   %.5f   %.5f   %.5f   %.5f   %.5f\n" % tuple(m))
        f.write("# noise parameters\n")
        for h in noise:
            f.write("   %.7f   %.7f   %.7f\n" % tuple(h))
        f.write("   %.7f\n" % N0)
        f.write("# noise information of the previous step: value / err_m / err_p\n")
        for h in noise:
            for v in h:
                f.write("   %.7f   %.7f   %.7f\n" % (v, 0.05 * v, 0.05 * v))
        f.write("   %.7f   %.7f   %.7f\n" % (N0, 0.05 * N0, 0.05 * N0))
        f.write("# common parameters\n")
        f.write("   model_fullname   %s\n   freq_smoothness   bool   1.0   2.0\n" % model_fullname)
        if trunc_c is not None:
            f.write("   trunc_c   Fix   %.3f\n" % trunc_c)
        f.write("   Splitting_a1   Uniform   %.6f   0.0   %.6f\n   Asphericity_eta   Fix_Auto   1.0\n   Splitting_a3   Fix   0.0\n   Asymetry   Fix   0.0\n" % (a1, max(3.0 * a1, 1.5)))
        f.write("   Inclination   Uniform   %.6f   0.0   90.0\n" % inclination)
        f.write("   Visibility_l1   Gaussian   1.5   1.5   0.15\n   Visibility_l2   Gaussian   0.53   0.53   0.05\n   Visibility_l3   Gaussian   0.08   0.08   0.02\n")
        f.write("   Height   %s   %s\n   Width   Fix_Auto   1.0\n" % (height_prior[0], "   ".join("%.6f" % v for v in height_prior[1:])))
        for line in extra_common:
            f.write("   %s\n" % line)
