// ref_frontend_dump -- test infrastructure.  Runs the UNMODIFIED reference front end (Config +
// Config_presets::apply_presets + Config::setup, i.e. main.cpp:88-151 without the sampler) from the
// current directory (which must hold Config/...) and prints what it hands to MALA / Model_def, with
// 17 significant digits, one "key: values" line per item.  tests/test_frontend_cpu.py compares the
// output of csrc/tg_frontend.cpp with it field by field.
//
//   ref_frontend_dump <star index (1-based)> <slice index (1-based)> <phase index (0-based)>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <unistd.h>
#include "config.h"
#include "config_presets.h"
#include "MALA.h"

#include <fstream>
// number of '*' range lines of the .model file (get_slices_range lives in main.cpp, which cannot be linked here)
static long count_slices(const std::string &f) { std::ifstream in(f.c_str()); std::string l; long n = 0; while (std::getline(in, l)) { size_t p = l.find_first_not_of(" \t"); if (p != std::string::npos && l[p] == '*') n++; } return n; }

static void pv(const char *k, const VectorXd &v) { printf("%s:", k); for (long i = 0; i < v.size(); i++) printf(" %.17g", v[i]); printf("\n"); }
static void pi(const char *k, const VectorXi &v) { printf("%s:", k); for (long i = 0; i < v.size(); i++) printf(" %d", v[i]); printf("\n"); }

int main(int argc, char **argv)
{
    if (argc < 4) { fprintf(stderr, "usage: ref_frontend_dump star slice phase\n"); return 2; }
    const int star = atoi(argv[1]) - 1, slice = atoi(argv[2]) - 1, phase = atoi(argv[3]);
    std::string cpath = getcwd(NULL, 0);
    // the reference prints its progress on stdout: keep it apart from the dump
    FILE *dump = fdopen(dup(1), "w");
    if (!freopen("/dev/null", "w", stdout)) return 3;
    Config config(cpath, cpath + "/Config/default/config_default.cfg", cpath + "/Config/default/errors_default.cfg",
                  cpath + "/Config/default/models_ctrl.list", cpath + "/Config/default/priors_ctrl.list",
                  cpath + "/Config/default/likelihoods_ctrl.list", cpath + "/Config/default/primepriors_ctrl.list");
    Config_presets master(cpath + "/Config/config_presets.cfg", &config);
    master.first_id_ind = star; master.last_id_ind = star;
    master.current_id_ind = star; master.current_slice_ind = slice; master.current_process_ind = phase;
    std::string modelfile = master.cfg_models_dir + master.table_ids[star].at(0) + ".model";
    const long nslices = count_slices(modelfile);
    master.Nslices = nslices;
    master.apply_presets(&config, 0);
    config.setup(slice);
    MALA mala(&config);
    fflush(stdout);
    stdout = dump;
    const Input_Data &in = config.modeling.inputs;
    printf("Nslices: %ld\n", nslices);
    printf("model_fct_name: %s\n", config.modeling.model_fct_name.c_str());
    printf("likelihood_fct_name: %s\n", config.modeling.likelihood_fct_name.c_str());
    printf("prior_fct_name: %s\n", config.modeling.prior_fct_name.c_str());
    printf("switches: %d %d %d\n", config.modeling.model_fct_name_switch, config.modeling.likelihood_fct_name_switch, config.modeling.prior_fct_name_switch);
    printf("likelihood_params: %.17g\n", config.modeling.likelihood_params);
    pv("inputs", in.inputs); pi("relax", in.relax); pi("plength", in.plength); pv("extra_priors", in.extra_priors);
    pi("priors_names_switch", in.priors_names_switch);
    for (long k = 0; k < in.priors.rows(); k++) { char key[32]; snprintf(key, sizeof key, "priors_row%ld", k); pv(key, in.priors.row(k)); }
    printf("inputs_names:"); for (size_t i = 0; i < in.inputs_names.size(); i++) printf(" |%s", in.inputs_names[i].c_str()); printf("\n");
    printf("priors_names:"); for (size_t i = 0; i < in.priors_names.size(); i++) printf(" |%s", in.priors_names[i].c_str()); printf("\n");
    const Data &d = config.data.data;
    printf("Nx: %ld\n", (long)d.Nx);
    double sx = 0, sy = 0; for (long i = 0; i < d.x.size(); i++) { sx += d.x[i]; sy += d.y[i]; }
    printf("data: %.17g %.17g %.17g %.17g %.17g %.17g\n", d.x[0], d.x[d.x.size() - 1], d.y[0], d.y[d.y.size() - 1], sx, sy);
    printf("mala: Nchains %ld dN_mixing %d lambda_temp %.17g c0 %.17g epsilon1 %.17Lg epsi2 %.17Lg A1 %.17Lg target %.17g\n",
           config.MALA.Nchains, config.MALA.dN_mixing, config.MALA.lambda_temp, config.MALA.c0, config.MALA.epsilon1, config.MALA.epsi2, config.MALA.A1, config.MALA.target_acceptance);
    printf("Nt_learn:"); for (size_t i = 0; i < config.MALA.Nt_learn.size(); i++) printf(" %d", config.MALA.Nt_learn[i]); printf("\n");
    printf("periods_learn:"); for (size_t i = 0; i < config.MALA.periods_learn.size(); i++) printf(" %d", config.MALA.periods_learn[i]); printf("\n");
    printf("outputs: Nsamples %ld Nbuffer %ld erase_old_files %d restore %d %d %d %d\n", config.outputs.Nsamples, config.outputs.Nbuffer, (int)config.outputs.erase_old_files,
           (int)config.outputs.do_restore_proposal, (int)config.outputs.do_restore_variables, (int)config.outputs.do_restore_last_index, (int)config.outputs.do_restore_proposal_mean);
    printf("output_root_name: %s\n", config.outputs.output_root_name.c_str());
    printf("dir_out: %s\n", config.outputs.dir_out.c_str());
    printf("restore_dir: %s\n", config.outputs.restore_dir.c_str());
    printf("restore_file_in: %s\n", config.outputs.restore_file_in.c_str());
    printf("restore_file_out: %s\n", config.outputs.restore_file_out.c_str());
    // MALA::MALA + init_proposal (MALA.cpp:53-112, :246-289): temperatures, sigma, diagonal of the initial covariance
    pv("Tcoefs", mala.Tcoefs); pv("sigma0", mala.sigma);
    pv("cov0_diag_chain0", mala.covarmat[0]->diagonal());
    pv("cov0_diag_chainlast", mala.covarmat[config.MALA.Nchains - 1]->diagonal());
    fflush(stdout);
    return 0;
}
