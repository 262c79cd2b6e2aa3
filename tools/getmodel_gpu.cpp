// getmodel_gpu -- the reference's stand-alone `getmodel` tool (tools/getmodel.cpp:25-186) on the GPU path:
//     getmodel_gpu <data file> <params file> <model name> [output file]
// data file   : ASCII columns x, y [, sigma_y] with optional '#' header, '!' labels and '*' units lines
//               (Config::read_data_ascii_Ncols, config.cpp:519-647)
// params file : '#' comments, then the plength line, then up to 5 lines of parameters (getmodel.cpp:75-112)
// model name  : a name of Config/default/models_ctrl.list (or its case number)
// output      : the data columns followed by one model column per parameter line, 12 significant digits
//               (getmodel.cpp:127-137); default name output_model.ascii.
// The model is evaluated by Model_def::call_model through the C-ABI (tg_eval with a model buffer): this is
// also the user-facing parity checker of the CUDA path against the reference's getmodel.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <fstream>
#include <sstream>
#include <string>
#include <vector>
#include "tamcmc_gpu.hpp"

static const char *MODEL_NAMES[15] = {
    "Test_Gaussian", "model_Harvey_Gaussian", "model_MS_Global_a1etaa3_HarveyLike", "model_MS_Global_a1etaa3_HarveyLike_Classic",
    "model_MS_Global_a1etaa3_Harvey1985", "model_MS_Global_a1acta3_Harvey1985", "model_MS_Global_a1l_etaa3_HarveyLike",
    "model_MS_Global_a1n_etaa3_HarveyLike", "model_MS_Global_a1nl_etaa3_HarveyLike", "model_MS_Global_a1etaa3_AppWidth_HarveyLike_v1",
    "model_MS_Global_a1etaa3_AppWidth_HarveyLike_v2", "model_MS_local_basic", "model_MS_Global_a1etaa3_HarveyLike_Classic_v2",
    "model_MS_Global_a1etaa3_HarveyLike_Classic_v3", "model_MS_local_Hnlm"};

static std::string trim(const std::string &s)
{
    const size_t a = s.find_first_not_of(" \t\r\n");
    return a == std::string::npos ? "" : s.substr(a, s.find_last_not_of(" \t\r\n") - a + 1);
}
static std::vector<double> numbers(const std::string &line)
{
    std::vector<double> v;
    std::istringstream ss(line);
    std::string tok;
    while (ss >> tok) { char *end = nullptr; const double d = strtod(tok.c_str(), &end); v.push_back(end == tok.c_str() ? std::nan("") : d); }
    return v;
}

int main(int argc, char **argv)
{
    if (argc < 4) {
        std::printf("usage: %s <data file> <params file> <model name> [output file]\n", argv[0]);
        return 2;
    }
    const std::string fdata = argv[1], fpar = argv[2], name = argv[3], fout = argc > 4 ? argv[4] : "output_model.ascii";
    int model_case = -1;
    for (int k = 0; k < 15; k++) if (name == MODEL_NAMES[k] || name == std::string("model_") + MODEL_NAMES[k]) model_case = k;
    if (model_case < 0) { char *e; const long v = strtol(name.c_str(), &e, 10); if (*e == 0 && v >= 0 && v < 15) model_case = (int)v; }
    if (model_case < 0) { std::printf("unknown model name %s (see Config/default/models_ctrl.list)\n", name.c_str()); return 2; }

    // ---- data ----
    std::ifstream fd(fdata.c_str());
    if (!fd.is_open()) { std::printf("Unable to open the file: %s\n", fdata.c_str()); return 2; }
    std::vector<std::vector<double>> rows;
    std::string line;
    while (std::getline(fd, line)) {
        line = trim(line);
        if (line.empty() || line[0] == '#' || line[0] == '!' || line[0] == '*') continue;
        rows.push_back(numbers(line));
    }
    if (rows.empty()) { std::printf("no data in %s\n", fdata.c_str()); return 2; }
    const size_t ncol = rows[0].size(), Nx = rows.size();
    std::vector<double> x(Nx), y(Nx, 0.0), sig;
    if (ncol >= 3) sig.resize(Nx);
    for (size_t i = 0; i < Nx; i++) { x[i] = rows[i][0]; if (ncol > 1 && rows[i].size() > 1) y[i] = rows[i][1]; if (ncol >= 3 && rows[i].size() > 2) sig[i] = rows[i][2]; }

    // ---- parameters ----
    std::ifstream fp(fpar.c_str());
    if (!fp.is_open()) { std::printf("Unable to open the file: %s\n", fpar.c_str()); return 2; }
    std::vector<int> plength;
    std::vector<std::vector<double>> params;
    while (std::getline(fp, line)) {
        line = trim(line);
        if (line.empty() || line[0] == '#') continue;
        const std::vector<double> v = numbers(line);
        if (plength.empty()) { for (double d : v) plength.push_back((int)d); continue; }
        if (params.size() < 5) params.push_back(v);
    }
    int Np = 0;
    for (int v : plength) Np += v;
    if (plength.empty() || plength.size() > 11 || params.empty()) { std::printf("params file: need a plength line (<= 11 entries) and at least one parameter line\n"); return 2; }
    for (const auto &p : params) if ((int)p.size() != Np) { std::printf("params file: a line has %zu values, plength sums to %d\n", p.size(), Np); return 2; }

    // ---- one handle, one chain; only call_model is used, so the sampler side gets neutral settings ----
    std::vector<int> relax((size_t)Np, 0), psw((size_t)Np, 0);
    relax[0] = 1;
    std::vector<double> pri((size_t)4 * Np, -9999.0);
    const double T[1] = {1.0};
    const long Nt_learn[2] = {0, 0}, periods[1] = {1};
    tg_problem P = tg_problem();
    P.Nx = (long)Nx; P.x = x.data(); P.y = y.data(); P.sigma_y = sig.empty() ? nullptr : sig.data();
    P.model_case = model_case; P.likelihood_case = 0; P.likelihood_params = 1.0;
    P.prior_case = model_case <= 1 ? model_case : ((model_case == 11 || model_case == 14) ? 3 : 2);
    for (size_t k = 0; k < plength.size(); k++) P.plength[k] = plength[k];
    P.Np = Np; P.params0 = params[0].data(); P.relax = relax.data(); P.priors = pri.data(); P.priors_switch = psw.data();
    P.Nchains = 1; P.Tcoefs = T; P.c0 = 1.0; P.epsilon1 = 1e-12; P.epsilon2 = 1e-12; P.A1 = 1e14; P.target_acceptance = 0.234;
    P.n_learn = 0; P.Nt_learn = Nt_learn; P.periods_learn = periods; P.dN_mixing = 0; P.seed = 1; P.device = 0;
    std::vector<std::vector<double>> models;
    try {
        tamcmc_gpu::Model_def md(P);
        for (const auto &p : params) models.push_back(md.call_model(p.data(), 1, T));
    } catch (const tamcmc_gpu::Error &e) {
        std::printf("%s\n", e.what());
        return e.code == TG_E_NODEVICE ? 3 : 1;
    }
    std::FILE *fo = std::fopen(fout.c_str(), "w");
    if (!fo) { std::printf("Unable to open the output file %s\n", fout.c_str()); return 2; }
    for (size_t i = 0; i < Nx; i++) {
        for (size_t c = 0; c < ncol; c++) std::fprintf(fo, "%.12g ", c < rows[i].size() ? rows[i][c] : std::nan(""));
        for (const auto &m : models) std::fprintf(fo, "%.12g ", m[i]);
        std::fprintf(fo, "\n");
    }
    std::fclose(fo);
    std::printf("Output model file successfully written\n");
    return 0;
}
