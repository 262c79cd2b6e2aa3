// tg_outputs.cpp -- writers / reader of the reference's on-disk formats (include/tamcmc_outputs.h).
// Host code only; built into libtamcmc_host.so with the system C++ compiler.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>
#include <string>
#include <vector>
#include "../../include/tamcmc_outputs.h"

static thread_local std::string g_err;
static int fail(int code, const std::string &msg) { g_err = msg; return code; }
extern "C" const char *tg_out_last_error(void) { return g_err.c_str(); }

struct tg_out {
    std::string params_root, stat_root, ptemp_root, restore_root;
    long Nsamples = 0, sofar = 0, written = 0;       // written: samples appended through this object
    int erase = 1, C = 0, Nv = 0, Ncons = 0, format = 0;
    std::vector<int> relax, plength;
    std::vector<std::string> vnames, cnames;
    std::vector<double> cvals, Tcoefs, vars_mean;
};

// Eigen's default IOFormat for a row of coefficients (what `stream << v.transpose()` prints): every
// coefficient formatted with the stream's precision, right-aligned to the widest one, separated by " ".
template <typename T> static std::string eigen_row(const T *v, long n, int precision)
{
    std::vector<std::string> s((size_t)n);
    size_t w = 0;
    for (long i = 0; i < n; i++) { std::ostringstream o; o.precision(precision); o << v[i]; s[(size_t)i] = o.str(); if (s[(size_t)i].size() > w) w = s[(size_t)i].size(); }
    std::string r;
    for (long i = 0; i < n; i++) { if (i) r += " "; r += std::string(w - s[(size_t)i].size(), ' ') + s[(size_t)i]; }
    return r;
}

extern "C" int tg_out_open(const tg_out_desc *d, tg_out **out)
{
    if (!d || !out || d->Nchains < 1 || d->Nvars < 1) return fail(TG_E_BADARG, "tg_out_open: bad argument");
    tg_out *o = new tg_out;
    if (d->params_root) o->params_root = d->params_root;
    if (d->stat_root) o->stat_root = d->stat_root;
    if (d->ptemp_root) o->ptemp_root = d->ptemp_root;
    if (d->restore_root) o->restore_root = d->restore_root;
    o->Nsamples = d->Nsamples; o->sofar = d->Nsamples_sofar; o->erase = d->erase_old_files;
    o->C = d->Nchains; o->Nv = d->Nvars; o->Ncons = d->Ncons;
    o->format = (d->file_format >= 0 && d->file_format <= 2) ? d->file_format : 0;
    if (d->relax) o->relax.assign(d->relax, d->relax + d->Np);
    if (d->plength) o->plength.assign(d->plength, d->plength + d->n_plength);
    for (int i = 0; i < d->Nvars; i++) o->vnames.push_back(d->variable_names ? d->variable_names[i] : "var" + std::to_string(i));
    if (d->constant_names && d->Ncons > 0) {
        for (int i = 0; i < d->Ncons; i++) o->cnames.push_back(d->constant_names[i]);
        if (d->constant_values) o->cvals.assign(d->constant_values, d->constant_values + d->Ncons);
    }
    if (d->Tcoefs) o->Tcoefs.assign(d->Tcoefs, d->Tcoefs + d->Nchains);
    o->vars_mean.assign((size_t)o->C * o->Nv, 0.0);
    *out = o;
    return TG_OK;
}

static std::string names_line(const std::vector<std::string> &n) { std::string s; for (const auto &x : n) s += x + "   "; return s; }

// The ASCII variants (file_format = text: extension txt; debug: dbg.txt beside the binary files): the header sits inside the file,
// written with the first block of a new file only; one line per sample in the stream's default format (6 significant digits,
// Eigen's column alignment for the vectors).
static int append_text(tg_out *o, const tg_records *rec, long n, int E)
{
    const int C = o->C, Nv = o->Nv;
    const std::string ext = (o->format == 1) ? "txt" : "dbg.txt";
    const bool fresh = (o->erase == 1 && o->written == 0);
    const bool header = (o->written == 0 && o->erase != 0);          // Ncopy == 0 && need_header
    const std::ios_base::openmode mode = fresh ? std::ios::trunc : std::ios::app;
    if (!o->params_root.empty() && rec->vars) {                       // outputs.cpp:510-613
        for (int c = 0; c < C; c++) {
            std::ofstream f((o->params_root + "_chain-" + std::to_string(c) + "." + ext).c_str(), mode);
            if (!f.is_open()) return fail(TG_E_BADARG, "cannot open " + o->params_root + "_chain-*." + ext);
            if (header) {
                f << "# This is an output file for the model parameters \n";
                f << "# This file contains values for vars[0:Nchains-1][ 0:Nvars-1]. Each matrix is in a different file, indexed by the chain number\n";
                f << "! Nsamples= " << o->Nsamples << "\n";
                f << "! Nchains= " << C << "\n";
                f << "! Nvars= " << Nv << "\n";
                f << "! Ncons= " << o->Ncons << "\n";
                f << "! chain= " << c << "\n";
                f << "! relax= " << eigen_row(o->relax.data(), (long)o->relax.size(), 6) << "\n";
                f << "! plength= " << eigen_row(o->plength.data(), (long)o->plength.size(), 6) << "\n";
                f << "! constant_names= " << (o->cnames.empty() ? std::string("None   ") : names_line(o->cnames)) << "\n";
                f << "! constant_values= " << (o->cnames.empty() ? std::string("-1") : eigen_row(o->cvals.data(), (long)o->cvals.size(), 6)) << "\n";
                f << "! variable_names=" << names_line(o->vnames) << "\n";
            }
            for (long i = 0; i < n; i++) f << eigen_row(rec->vars + ((size_t)i * C + c) * Nv, Nv, 6) << "\n";
        }
    }
    if (!o->stat_root.empty() && rec->logL && rec->logPrior && rec->logPost) {     // outputs.cpp:791-860
        std::ofstream f((o->stat_root + "." + ext).c_str(), mode);
        if (!f.is_open()) return fail(TG_E_BADARG, "cannot open " + o->stat_root + "." + ext);
        if (header) {
            f << "# This is an output file for the statistical information.\n";
            f << "# This file contains values for the logLikelihood (columns 0:Nchains-1), logPrior (columns Nchains:2*Nchains-1) and logPosterior (columns 2*Nchains:3*Nchains-1),  \n";
            f << "! Nchains= " << C << "\n";
            f << "! labels= ";
            const char *labels[3] = {"logLikelihood", "logPrior", "logPosteriors"};
            for (int k = 0; k < 3; k++) for (int i = 0; i < C; i++) f << labels[k] << "[" << i << "]   ";
            f << "\n";
        }
        for (long i = 0; i < n; i++)
            f << eigen_row(rec->logL + (size_t)i * C, C, 6) << "     " << eigen_row(rec->logPrior + (size_t)i * C, C, 6) << "     " << eigen_row(rec->logPost + (size_t)i * C, C, 6) << "\n";
    }
    if (!o->ptemp_root.empty() && rec->swap_attempt) {                // outputs.cpp:615-676
        std::ofstream f((o->ptemp_root + "." + ext).c_str(), mode);
        if (!f.is_open()) return fail(TG_E_BADARG, "cannot open " + o->ptemp_root + "." + ext);
        if (header) {
            f << "# This is an output file for the parameters of the parallel tempering.\n";
            f << "# This file contains values for \n";
            f << "# Correspondance between chain0=[0:Nchains-1] and temperature Tcoefs[chain] \n";
            f << "! Tcoefs = " << eigen_row(o->Tcoefs.data(), (long)o->Tcoefs.size(), 6) << "\n";
            f << "! labels= attempt_mixing    chain0    Pswitch    switched \n";
        }
        for (long i = 0; i < n; i++) {
            const size_t e = (size_t)i * E;
            f << (rec->swap_attempt[e] ? 1 : 0) << "   " << (rec->swap_chain0 ? rec->swap_chain0[e] : -1) << "   " << (rec->swap_Pswap ? rec->swap_Pswap[e] : 0.0)
              << "   " << ((rec->swap_done && rec->swap_done[e]) ? 1 : 0) << "\n";
        }
    }
    return TG_OK;
}

static int write_headers(const tg_out *o)
{
    const long done = o->written + o->sofar;          // Nbuffer*Ncopy + counts + Nsamples_sofar
    if (!o->params_root.empty()) {                    // outputs.cpp:1262-1305
        std::ofstream f((o->params_root + ".hdr").c_str());
        if (!f.is_open()) return fail(TG_E_BADARG, "cannot open " + o->params_root + ".hdr");
        f << "# This is the header file of the BINARY output file for the model parameters \n";
        f << "# This file contains values for vars[0:Nchains-1][ 0:Nvars-1]. Each matrix is in a different file, indexed by the chain number\n";
        f << "! Nsamples= " << o->Nsamples << "\n";
        f << "! Nchains= " << o->C << "\n";
        f << "! Nsamples_done=" << done << "\n";
        f << "! Nvars= " << o->Nv << "\n";
        f << "! Ncons= " << o->Ncons << "\n";
        f << "! relax= " << eigen_row(o->relax.data(), (long)o->relax.size(), 6) << "\n";
        f << "! plength= " << eigen_row(o->plength.data(), (long)o->plength.size(), 6) << "\n";
        f << "! constant_names= " << (o->cnames.empty() ? std::string("None   ") : names_line(o->cnames)) << "\n";
        f << "! constant_values= " << (o->cnames.empty() ? std::string("-1") : eigen_row(o->cvals.data(), (long)o->cvals.size(), 6)) << "\n";
        f << "! variable_names=" << names_line(o->vnames) << "\n";
    }
    if (!o->stat_root.empty()) {                      // outputs.cpp:1500-1516
        std::ofstream f((o->stat_root + ".hdr").c_str());
        if (!f.is_open()) return fail(TG_E_BADARG, "cannot open " + o->stat_root + ".hdr");
        f << "# This is the header of the BINARY output file for the statistical information.\n";
        f << "# This file contains values for the logLikelihood (columns 0:Nchains-1), logPrior (columns Nchains:2*Nchains-1) and logPosterior (columns 2*Nchains:3*Nchains-1),  \n";
        f << "! Nsamples_done=" << done << "\n";
        f << "! Nchains= " << o->C << "\n";
        f << "! labels= ";
        const char *labels[3] = {"logLikelihood", "logPrior", "logPosteriors"};
        for (int k = 0; k < 3; k++) for (int i = 0; i < o->C; i++) f << labels[k] << "[" << i << "]   ";
        f << "\n";
    }
    if (!o->ptemp_root.empty()) {                     // outputs.cpp:1363-1375
        std::ofstream f((o->ptemp_root + ".hdr").c_str());
        if (!f.is_open()) return fail(TG_E_BADARG, "cannot open " + o->ptemp_root + ".hdr");
        f << "# This is the header of the BINARY output file for the parameters of the parallel tempering.\n";
        f << "# This file contains values for \n";
        f << "# Correspondance between chain0=[0:Nchains-1] and temperature Tcoefs[chain] \n";
        f << "! Nsamples_done=" << done << "\n";
        f << "! Tcoefs = " << eigen_row(o->Tcoefs.data(), (long)o->Tcoefs.size(), 6) << "\n";
        f << "! labels= attempt_mixing    chain0    Pswitch    switched \n";
    }
    return TG_OK;
}

extern "C" int tg_out_append(tg_out *o, const tg_records *rec, long n, int E)
{
    if (!o || !rec || n < 0 || E < 1) return fail(TG_E_BADARG, "tg_out_append: bad argument");
    const int C = o->C, Nv = o->Nv;
    const bool fresh = (o->erase == 1 && o->written == 0);          // erase_old_files == 1 && Ncopy == 0: truncate
    const std::ios_base::openmode mode = std::ios::binary | (fresh ? std::ios::trunc : std::ios::app);
    if (o->format != 0) { int r = append_text(o, rec, n, E); if (r) return r; }
    const std::string bext = (o->format == 2) ? ".dbg" : ".bin";     // file_ext of the reference: bin, or dbg in debug mode
    const bool binary = (o->format != 1);
    if (binary && !o->params_root.empty() && rec->vars) {           // one file per chain, raw f64 [sample][Nv]
        std::vector<double> row((size_t)n * Nv);
        for (int c = 0; c < C; c++) {
            std::ofstream f((o->params_root + "_chain-" + std::to_string(c) + bext).c_str(), mode);
            if (!f.is_open()) return fail(TG_E_BADARG, "cannot open " + o->params_root + "_chain-*.bin");
            for (long i = 0; i < n; i++) memcpy(&row[(size_t)i * Nv], rec->vars + ((size_t)i * C + c) * Nv, (size_t)Nv * 8);
            f.write(reinterpret_cast<const char *>(row.data()), (std::streamsize)(row.size() * 8));
        }
    }
    if (rec->vars) {
        std::fill(o->vars_mean.begin(), o->vars_mean.end(), 0.0);
        for (long i = 0; i < n; i++) for (size_t k = 0; k < (size_t)C * Nv; k++) o->vars_mean[k] += rec->vars[(size_t)i * C * Nv + k];
        if (n > 0) for (auto &v : o->vars_mean) v /= (double)n;
    }
    if (binary && !o->stat_root.empty() && rec->logL && rec->logPrior && rec->logPost) {   // per sample: logL[C] | logPrior[C] | logPost[C]
        std::ofstream f((o->stat_root + bext).c_str(), mode);
        if (!f.is_open()) return fail(TG_E_BADARG, "cannot open " + o->stat_root + ".bin");
        for (long i = 0; i < n; i++) {
            f.write(reinterpret_cast<const char *>(rec->logL + (size_t)i * C), C * 8);
            f.write(reinterpret_cast<const char *>(rec->logPrior + (size_t)i * C), C * 8);
            f.write(reinterpret_cast<const char *>(rec->logPost + (size_t)i * C), C * 8);
        }
    }
    if (binary && !o->ptemp_root.empty() && rec->swap_attempt) {    // per sample: bool(1) int(4) double(8) bool(1), packed
        std::ofstream f((o->ptemp_root + bext).c_str(), mode);
        if (!f.is_open()) return fail(TG_E_BADARG, "cannot open " + o->ptemp_root + ".bin");
        for (long i = 0; i < n; i++) {
            const size_t e = (size_t)i * E;
            const unsigned char att = rec->swap_attempt[e] ? 1 : 0, sw = (rec->swap_done && rec->swap_done[e]) ? 1 : 0;
            const int c0 = rec->swap_chain0 ? rec->swap_chain0[e] : -1;
            const double ps = rec->swap_Pswap ? rec->swap_Pswap[e] : 0.0;
            f.write(reinterpret_cast<const char *>(&att), 1);
            f.write(reinterpret_cast<const char *>(&c0), 4);
            f.write(reinterpret_cast<const char *>(&ps), 8);
            f.write(reinterpret_cast<const char *>(&sw), 1);
        }
    }
    o->written += n;
    if (o->erase == 1 && binary) return write_headers(o);                     // appends leave the headers alone (outputs.cpp:1254)
    return TG_OK;
}

static void restore_preamble(std::ofstream &f, int number, const char *what, const tg_out *o, long iteration, const char *flag)
{
    f << "# This is an output file containing what is required to restore a run to its last saved position \n";
    f << "# File number: " << number << " \n";
    f << what;
    f << "# Use this if you wish to: \n";
    f << "#       (1) complete a finished job that requires more samples ==> set erase_old_file=0 and " << flag << "=1 \n";
    f << "#       (2) restart a finished job by ignoring old samples (e.g. ignoring a Burn-in) ==> set erase_old_file=1 and " << (number == 1 ? "do_restore_proposal" : flag) << "=1 \n";
    f << "#       (3) terminate an unfinished job which failed to finished (e.g. due to computer unexpected shutdown) ==> set erase_old_file=0 and " << flag << "=1 \n";
    f << "! Nchains= " << o->C << "\n";
    f << "! Nvars= " << o->Nv << "\n";
    f << "! iteration=" << iteration << "\n";
    f << "! variable_names=" << names_line(o->vnames) << "\n";
}

extern "C" int tg_out_write_restore(tg_out *o, long iteration, const double *vars, const double *mu, const double *sigma, const double *covarmat)
{
    return tg_out_write_restore_means(o, iteration, vars, mu, sigma, covarmat, nullptr, nullptr, nullptr, nullptr);
}

extern "C" int tg_out_write_restore_means(tg_out *o, long iteration, const double *vars, const double *mu, const double *sigma, const double *covarmat,
                                          const double *vars_mean, const double *mu_mean, const double *sigma_mean, const double *covarmat_mean)
{
    if (!o || o->restore_root.empty() || !vars || !mu || !sigma || !covarmat) return fail(TG_E_BADARG, "tg_out_write_restore: bad argument");
    const int C = o->C, Nv = o->Nv, P = 17;
    if (!vars_mean) vars_mean = o->written ? o->vars_mean.data() : vars;
    {
        std::ofstream f((o->restore_root + "1.dat").c_str());
        if (!f.is_open()) return fail(TG_E_BADARG, "cannot open " + o->restore_root + "1.dat");
        restore_preamble(f, 1, "# Contains the last values for the variables vars[0:Nchain-1]. vars_mean denotes averaged values of Nbuffer \n", o, iteration, "do_restore_[X]");
        f << "! vars= \n";
        for (int c = 0; c < C; c++) f << eigen_row(vars + (size_t)c * Nv, Nv, P) << "\n";
        f << "! vars_mean= \n";
        for (int c = 0; c < C; c++) f << eigen_row(vars_mean + (size_t)c * Nv, Nv, P) << "\n";
    }
    {
        std::ofstream f((o->restore_root + "2.dat").c_str());
        if (!f.is_open()) return fail(TG_E_BADARG, "cannot open " + o->restore_root + "2.dat");
        restore_preamble(f, 2, "# Contains the last values of (a) sigmas[0:Nchains-1] and (b) mus[0:Nchains-1, 0:Nvars-1].  sigmas_mean and mus_mean denotes averaged values of Nbuffer\n", o, iteration, "do_restore");
        for (int rep = 0; rep < 2; rep++) {
            const double *sg = (rep && sigma_mean) ? sigma_mean : sigma, *m = (rep && mu_mean) ? mu_mean : mu;
            f << (rep ? "! sigmas_mean= " : "! sigmas= ") << eigen_row(sg, C, P) << "\n";
            f << (rep ? "! mus_mean= \n" : "! mus= \n");
            for (int c = 0; c < C; c++) f << eigen_row(m + (size_t)c * Nv, Nv, P) << "\n";
        }
    }
    {
        std::ofstream f((o->restore_root + "3.dat").c_str());
        if (!f.is_open()) return fail(TG_E_BADARG, "cannot open " + o->restore_root + "3.dat");
        restore_preamble(f, 3, "# Contains the last value of the covariance matrix covarmats[0:Nchains-1, 0:Nvars-1, 0:Nvars-1]. covarmats_mean denotes the averaged values over Nbuffer\n", o, iteration, "do_restore");
        for (int rep = 0; rep < 2; rep++) {
            const double *cv = (rep && covarmat_mean) ? covarmat_mean : covarmat;
            f << (rep ? "! covarmats_mean= \n" : "! covarmats= \n");
            for (int c = 0; c < C; c++) {
                f << "*" << c << "\n";
                for (int i = 0; i < Nv; i++) f << eigen_row(cv + ((size_t)c * Nv + i) * Nv, Nv, P) << "\n";
            }
        }
    }
    return TG_OK;
}

extern "C" int tg_out_close(tg_out *o) { delete o; return TG_OK; }

// ---------------------------------------------------------------------------------------------
// Reader (config.cpp:1321-1579): lines starting with '!' are "keyword= value"; the value of vars / mus /
// covarmats follows on the next lines, one row per line, '*<chain>' opening every matrix.
// ---------------------------------------------------------------------------------------------
static std::string trim(const std::string &s)
{
    const size_t a = s.find_first_not_of(" \t\r\n");
    if (a == std::string::npos) return "";
    return s.substr(a, s.find_last_not_of(" \t\r\n") - a + 1);
}
static bool parse_row(const std::string &line, double *dst, int n)
{
    std::istringstream ss(line);
    for (int i = 0; i < n; i++) if (!(ss >> dst[i])) return false;
    return true;
}

static int restore_read_impl(const char *root, int last_index, int *Nchains, int *Nvars, long *iteration,
                             double *vars, double *mu, double *sigma, double *covarmat, bool mean);
extern "C" int tg_restore_read(const char *root, int last_index, int *Nchains, int *Nvars, long *iteration,
                               double *vars, double *mu, double *sigma, double *covarmat)
{
    return restore_read_impl(root, last_index, Nchains, Nvars, iteration, vars, mu, sigma, covarmat, false);
}
extern "C" int tg_restore_read_means(const char *root, int last_index, int *Nchains, int *Nvars, long *iteration,
                                     double *vars_mean, double *mu_mean, double *sigma_mean, double *covarmat_mean)
{
    return restore_read_impl(root, last_index, Nchains, Nvars, iteration, vars_mean, mu_mean, sigma_mean, covarmat_mean, true);
}
static int restore_read_impl(const char *root, int last_index, int *Nchains, int *Nvars, long *iteration,
                             double *vars, double *mu, double *sigma, double *covarmat, bool mean)
{
    if (!root) return fail(TG_E_BADARG, "tg_restore_read: bad argument");
    const std::string k_vars = mean ? "! vars_mean" : "! vars", k_mus = mean ? "! mus_mean" : "! mus", k_sig = mean ? "! sigmas_mean" : "! sigmas",
                      k_cov = mean ? "! covarmats_mean" : "! covarmats";
    int C = -1, Nv = -1;
    long it = 0;
    for (int number = 1; number <= 3; number++) {
        std::ifstream f((std::string(root) + std::to_string(number) + ".dat").c_str());
        if (!f.is_open()) return fail(TG_E_BADARG, std::string("cannot open ") + root + std::to_string(number) + ".dat");
        std::string line;
        int found = 0;
        while (std::getline(f, line)) {
            line = trim(line);
            if (line.empty() || line[0] != '!') continue;
            const size_t eq = line.find('=');
            if (eq == std::string::npos) continue;
            const std::string key = trim(line.substr(0, eq)), val = trim(line.substr(eq + 1));
            if (key == "! Nchains") C = atoi(val.c_str());
            else if (key == "! Nvars") Nv = atoi(val.c_str());
            else if (key == "! iteration") it = last_index ? atol(val.c_str()) : 0;
            else if (key == k_vars || key == k_mus) {
                if (C < 1 || Nv < 1) return fail(TG_E_BADARG, "restore file: sizes must precede the values");
                double *dst = (key == k_vars) ? vars : mu;
                std::vector<double> tmp((size_t)Nv);
                for (int c = 0; c < C; c++) {
                    if (!std::getline(f, line) || !parse_row(line, tmp.data(), Nv)) return fail(TG_E_BADARG, "restore file: short row in " + key);
                    if (dst) memcpy(dst + (size_t)c * Nv, tmp.data(), (size_t)Nv * 8);
                }
                found++;
            } else if (key == k_sig) {
                if (C < 1) return fail(TG_E_BADARG, "restore file: sizes must precede the values");
                std::vector<double> tmp((size_t)C);
                if (!parse_row(val, tmp.data(), C)) return fail(TG_E_BADARG, "restore file: short sigmas");
                if (sigma) memcpy(sigma, tmp.data(), (size_t)C * 8);
                found++;
            } else if (key == k_cov) {
                if (C < 1 || Nv < 1) return fail(TG_E_BADARG, "restore file: sizes must precede the values");
                std::vector<double> tmp((size_t)Nv);
                for (int c = 0; c < C; c++) {
                    if (!std::getline(f, line) || trim(line) != "*" + std::to_string(c)) return fail(TG_E_BADARG, "restore file: matrix indicator not found");
                    for (int i = 0; i < Nv; i++) {
                        if (!std::getline(f, line) || !parse_row(line, tmp.data(), Nv)) return fail(TG_E_BADARG, "restore file: short covarmat row");
                        if (covarmat) memcpy(covarmat + ((size_t)c * Nv + i) * Nv, tmp.data(), (size_t)Nv * 8);
                    }
                }
                found++;
            }
        }
        const int need = (number == 2) ? 2 : 1;
        if (found < need) return fail(TG_E_BADARG, "restore file " + std::to_string(number) + ": missing keywords");
    }
    if (Nchains) *Nchains = C;
    if (Nvars) *Nvars = Nv;
    if (iteration) *iteration = it;
    return TG_OK;
}
