// tg_frontend.cpp -- the reference's input front end, rewritten as plain host C++ (include/tamcmc_frontend.h).
//
// Reads config_default.cfg / errors_default.cfg / *_ctrl.list / config_presets.cfg / <id>.model / <id>.data and
// builds the parameter vector, relax flags, priors, plength and extra_priors of the model family exactly as the
// reference does (file:line citations are relative to /root/reference/tamcmc/sources).  The data model is a list
// of parameter blocks (heights, visibilities, frequencies, splitting, widths, noise, inclination) that are
// concatenated at the end, as in io_models.cpp; the keyword handling of the `.model` "common parameters" section is
// table-driven here instead of the reference's if-chains.  Checked field by field against the unmodified reference
// front end (oracle/_ref/ref_frontend_dump) in tests/test_frontend_cpu.py.
#include <array>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <map>
#include <sstream>
#include <string>
#include <sys/stat.h>
#include <vector>
#include "../../include/tamcmc_frontend.h"

namespace {

thread_local std::string g_err;
struct FeError { std::string msg; };
[[noreturn]] void die(const std::string &m) { throw FeError{m}; }

// ------------------------------------------------------------------ strings (string_handler.cpp)
const char *WS = " \t";
std::string trim(const std::string &s)                       // strtrim, string_handler.cpp:23-38
{
    const size_t b = s.find_first_not_of(WS);
    if (b == std::string::npos) return "";
    return s.substr(b, s.find_last_not_of(WS) - b + 1);
}
// strsplit (string_handler.cpp:40-56): tokens between any of `delims`, each trimmed; a delimiter at position 0 yields
// no token; the remainder is always appended (so "a=" gives {"a", ""}).
std::vector<std::string> split(const std::string &str, const std::string &delims)
{
    std::string s = trim(str);
    std::vector<std::string> out;
    size_t pos;
    while ((pos = s.find_first_of(delims)) != std::string::npos) {
        if (pos != 0) out.push_back(trim(s.substr(0, pos)));
        s.erase(0, pos + 1);
        s = trim(s);
    }
    out.push_back(s);
    return out;
}
double to_dbl(const std::string &s) { double v = 0; std::stringstream(trim(s)) >> v; return v; }
int    to_int(const std::string &s) { int v = 0; std::stringstream(trim(s)) >> v; return v; }
bool   to_bool(const std::string &s) { bool v = false; std::stringstream(trim(s)) >> v; return v; }
std::string first_char(const std::string &l) { return trim(l.substr(0, 1)); }
std::vector<double> to_dbls(const std::string &s, const std::string &d) { std::vector<double> v; for (auto &w : split(s, d)) { double x = 0; if (trim(w) != "") std::stringstream(w) >> x; v.push_back(x); } return v; }

// a text file as the reference walks it: getline + eof() tests
struct Lines {
    std::ifstream in; std::string cur; bool ok;
    explicit Lines(const std::string &f) : in(f.c_str()) { ok = in.is_open(); }
    bool next() { std::getline(in, cur); return true; }
    bool eof() const { return in.eof(); }
};

// ------------------------------------------------------------------ parameter blocks (io_models.cpp)
struct Block {                                               // Input_Data, data.h:37-47
    std::vector<std::string> names, pnames;
    std::vector<double> inputs;
    std::vector<int> relax;
    std::vector<std::array<double, 4>> priors;               // priors(k, pos) -> priors[pos][k]
    void init(int n)                                         // IO_models::initialise_param, io_models.cpp:232-266
    {
        names.assign((size_t)n, "Empty"); pnames.assign((size_t)n, "Fix");
        inputs.assign((size_t)n, 0.0); relax.assign((size_t)n, 0);
        priors.assign((size_t)n, std::array<double, 4>{{-9999, -9999, -9999, -9999}});
    }
    int size() const { return (int)inputs.size(); }
    // IO_models::fill_param, io_models.cpp:24-58.  pv holds at least i0+4 values unless the prior is "Fix".
    void fill(const std::string &name, const std::string &prior, double val, const std::vector<double> &pv, int pos, int i0)
    {
        if (pos < 0 || pos >= size()) die("front end: parameter slot out of range while filling '" + name + "'");
        names[(size_t)pos] = name; pnames[(size_t)pos] = prior; inputs[(size_t)pos] = val;
        if (prior == "Fix") { relax[(size_t)pos] = 0; priors[(size_t)pos] = {{-9999, -9999, -9999, -9999}}; }
        else {
            relax[(size_t)pos] = 1;
            for (int k = 0; k < 4; k++) priors[(size_t)pos][(size_t)k] = ((size_t)(k + i0) < pv.size()) ? pv[(size_t)(k + i0)] : -9999;
        }
    }
    // IO_models::fill_param_vect, io_models.cpp:60-77
    void fill_vect(const std::vector<double> &v, const std::vector<bool> &rl, const std::string &name, const std::string &prior,
                   const std::vector<double> &pv, int pos, int i0_if, int i0_else)
    {
        for (size_t i = 0; i < v.size(); i++) {
            if (rl[i]) fill(name, prior, v[i], pv, (int)i + pos, i0_if);
            else       fill(name, "Fix", v[i], pv, (int)i + pos, i0_else);
        }
    }
    std::vector<double> prior_col(int pos) const { return std::vector<double>(priors[(size_t)pos].begin(), priors[(size_t)pos].end()); }
    void put(const Block &b, int pos)                        // IO_models::add_param, io_models.cpp:99-120
    {
        for (int i = 0; i < b.size(); i++) {
            names[(size_t)(pos + i)] = b.names[(size_t)i]; pnames[(size_t)(pos + i)] = b.pnames[(size_t)i];
            inputs[(size_t)(pos + i)] = b.inputs[(size_t)i]; relax[(size_t)(pos + i)] = b.relax[(size_t)i]; priors[(size_t)(pos + i)] = b.priors[(size_t)i];
        }
    }
};

struct ModelFile {                                           // MCMC_files, data.h:73-93
    std::string ID; double Dnu = 0, numax = -9999, C_l = 0;
    std::vector<int> els; std::vector<double> freqs_ref; double freq_range[2] = {0, 0};
    std::vector<bool> relax_freq, relax_gamma, relax_H;
    std::vector<double> hyper_priors;
    std::vector<std::array<double, 6>> eigen;                // l, nu, window_min, window_max, Gamma, H
    std::vector<double> noise_params;                        // 10
    std::vector<std::array<double, 3>> noise_s2;             // 10 x (value, err_m, err_p)
    std::vector<std::string> common_names, common_priors;
    std::vector<std::vector<double>> common;                 // rows of 5: value + 4 prior values, -9999 where absent
};

// read_MCMC_file_MS_Global (io_ms_global.cpp:25-306) and read_MCMC_file_local (io_local.cpp:26-301): the same walk
// over the '#'-separated sections; slice < 0 is the global reader (a second '*' range line is an error), otherwise
// the range line number `slice` is kept and the others are skipped.
ModelFile read_model_file(const std::string &file, int slice)
{
    Lines f(file);
    if (!f.ok) die("Unable to open the file: " + file);
    ModelFile m;
    int out = 0, range_counter = 0; bool range_done = false;
    f.next();
    while (out < 3 && !f.eof()) {                            // identification, ranges, list of relaxed modes
        const std::string l = trim(f.cur), c0 = first_char(l), c1 = l.size() > 1 ? l.substr(1, 1) : "";
        if (c0 == "#" && c1 == "K") { auto w = split(l, "= \t"); if (w.size() > 1) m.ID = trim(w[1]); }
        if (c0 == "!" && c1 == "n") m.numax = to_dbl(split(l, " ").at(1));
        if (c0 == "!" && c1 != "!" && c1 != "n") m.Dnu = to_dbl(split(l, " ").at(1));
        if (c0 == "!" && c1 == "!") m.C_l = to_dbl(split(l, " ").at(1));
        if (c0 == "*") {
            auto w = split(l, " ");
            if (w.size() < 3) die("front end: malformed range line in " + file);
            if (slice < 0) {
                if (range_done) die("Multiple range detected. This is not allowed with io_ms_global models and priors");
                m.freq_range[0] = to_dbl(w[1]); m.freq_range[1] = to_dbl(w[2]); range_done = true;
            } else {
                if (range_counter == slice) { m.freq_range[0] = to_dbl(w[1]); m.freq_range[1] = to_dbl(w[2]); }
                range_counter++;
            }
        }
        if (c0 != "#" && c0 != "!" && c0 != "*") {
            auto w = split(l, " ");
            const std::string t = trim(w[0]);
            if (!(t == "p" || t == "g" || t == "co") || w.size() < 3) die("Problem with the keyword that identifies the type of mode (p, g or co) in " + file);
            m.els.push_back(to_int(w[1])); m.freqs_ref.push_back(to_dbl(w[2]));
            m.relax_freq.push_back(w.size() >= 4 ? to_bool(w[3]) : true);
            m.relax_H.push_back(w.size() >= 5 ? to_bool(w[4]) : true);      // column order in the file: relax_f, relax_H, relax_W
            m.relax_gamma.push_back(w.size() >= 6 ? to_bool(w[5]) : true);
        }
        if (c0 == "#") out++;
        f.next();
    }
    while (out < 4 && !f.eof()) {                            // hyper priors (the line read above is dropped, as in the reference)
        f.next();
        const std::string l = trim(f.cur);
        if (first_char(l) != "#") m.hyper_priors.push_back(to_dbl(l)); else out++;
    }
    f.next();
    while (out < 5 && !f.eof()) {                            // eigen solutions
        const std::string l = trim(f.cur);
        if (first_char(l) != "#") { auto v = to_dbls(l, " \t"); std::array<double, 6> r{{0, 0, 0, 0, 0, 0}}; for (size_t k = 0; k < 6 && k < v.size(); k++) r[k] = v[k]; m.eigen.push_back(r); }
        else out++;
        f.next();
    }
    std::vector<double> np;
    while (out < 6 && !f.eof()) {                            // noise parameters, any number per line
        const std::string l = trim(f.cur);
        if (first_char(l) != "#") { for (auto &w : split(l, " \t")) np.push_back(to_dbl(w)); } else out++;
        f.next();
    }
    if (np.size() > 10) die("front end: more than 10 noise parameters in " + file);
    m.noise_params.assign(10, -1.0);                         // fewer than 3 Harvey + white noise: right-aligned, -1 in front
    if (np.size() == 10) m.noise_params = np; else for (size_t k = 0; k < np.size(); k++) m.noise_params[10 - np.size() + k] = np[k];
    std::vector<std::array<double, 3>> s2;
    while (out < 7 && !f.eof()) {                            // noise information of the previous fitting step
        const std::string l = trim(f.cur);
        if (first_char(l) != "#") { auto v = to_dbls(l, " \t"); std::array<double, 3> r{{0, 0, 0}}; for (size_t k = 0; k < 3 && k < v.size(); k++) r[k] = v[k]; s2.push_back(r); }
        else out++;
        f.next();
    }
    if (s2.size() > 10) die("front end: more than 10 noise_s2 rows in " + file);
    m.noise_s2.assign(10, std::array<double, 3>{{-1, -1, -1}});
    if (s2.size() == 10) m.noise_s2 = s2; else for (size_t k = 0; k < s2.size(); k++) m.noise_s2[10 - s2.size() + k] = s2[k];
    while (out < 9 && !f.eof()) {                            // common parameters: name, prior name, value, prior values
        const std::string l = trim(f.cur);
        if (l != "") {
            if (first_char(l) != "#") {
                auto w = split(l, " \t");
                if (w.size() < 2) die("front end: malformed common-parameter line '" + l + "'");
                m.common_names.push_back(trim(w[0])); m.common_priors.push_back(trim(w[1]));
                std::vector<double> row(5, -9999.0);
                for (size_t k = 2; k < w.size() && k < 7; k++) { double x = 0; std::stringstream(w[k]) >> x; row[k - 2] = x; }
                m.common.push_back(row);
            } else out++;
        }
        f.next();
    }
    return m;
}

// ------------------------------------------------------------------ small numerics used while building the vector
const long double PI_L = 3.141592653589793238L;

// amplitude_ratio(l, beta) (function_rot.cpp:20-47): squared reduced rotation matrix elements d^l_{m0}(beta)^2, m = -l..l,
// through the same factorial sum as the reference (function_rot.cpp:49-93).
double factorial(int n) { double f = 1; for (int i = 2; i <= n; i++) f *= i; return f; }
std::vector<double> amplitude_ratio(int l, double beta_deg)
{
    const double beta = beta_deg * (double)PI_L / 180.0;
    std::vector<double> out((size_t)(2 * l + 1));
    for (int m1 = -l; m1 <= l; m1++) {
        const int m2 = 0;
        double s = 0;
        for (int k = 0; k <= 2 * l; k++) {
            const int a = l + m2 - k, b = k, c = m1 - m2 + k, d = l - m1 - k;
            if (a < 0 || c < 0 || d < 0) continue;
            const double num = std::sqrt(factorial(l + m1) * factorial(l - m1) * factorial(l + m2) * factorial(l - m2));
            const double term = ((m1 - m2 + k) % 2 == 0 ? 1.0 : -1.0) * num / (factorial(a) * factorial(b) * factorial(c) * factorial(d))
                              * std::pow(std::cos(beta / 2), 2 * l - 2 * k - m1 + m2) * std::pow(std::sin(beta / 2), 2 * k + m1 - m2);
            s += term;
        }
        out[(size_t)(m1 + l)] = s * s;
    }
    return out;
}
// harvey_like on the two ends of the analysed range (noise_models.cpp:17-41), used by set_noise_params_local
std::vector<double> harvey_like(const std::vector<double> &p, const double *x, int nx, int Nharvey)
{
    std::vector<double> y((size_t)nx, 0.0);
    for (int k = 0; k < Nharvey; k++) {
        const double H = p[(size_t)(3 * k)], tc = p[(size_t)(3 * k + 1)], pw = p[(size_t)(3 * k + 2)];
        if (tc != 0) for (int i = 0; i < nx; i++) y[(size_t)i] += H / (1.0 + std::pow(1e-3 * tc * x[i], pw));
    }
    for (int i = 0; i < nx; i++) y[(size_t)i] += p[(size_t)(3 * Nharvey)];
    return y;
}

struct InputData {                                           // what build_init_* returns
    Block all; std::string model_fullname; std::vector<int> plength; double extra_priors[4] = {0, 0, 0, 0};
};

// rows of the `.model` common section
struct Common {
    const ModelFile &m;
    explicit Common(const ModelFile &mf) : m(mf) {}
    size_t n() const { return m.common_names.size(); }
    bool is(size_t i, const char *a, const char *b = nullptr) const { return m.common_names[i] == a || (b && m.common_names[i] == b); }
    const std::string &prior(size_t i) const { return m.common_priors[i]; }
    double v(size_t i, int k) const { return m.common[i][(size_t)k]; }
    const std::vector<double> &row(size_t i) const { return m.common[i]; }
};
[[noreturn]] void bad_keyword(const std::string &name, const std::string &type)   // fatalerror_msg_io_*
{
    if (type == "Fix_Auto") die("Fix_Auto is not implemented for the parameter '" + name + "': choose the prior yourself");
    die("'" + name + "' should always be defined as '" + type + "'");
}

// The splitting block (a1, eta, a3, sqrt(a1)cos i, sqrt(a1)sin i, asymmetry [, a1 per l / n / (n,l)]) and the inclination
// block are filled from the same keywords by the global and the local builders (io_ms_global.cpp:704-857, io_local.cpp:829-928).
struct RotState { bool a1cosi = false, a1sini = false; };
void rotation_keywords(const Common &c, size_t i, Block &Snlm, Block &Inc, bool a11_eq_a12, bool avg_a1n, const int Nf_el[4],
                       double Dnu, double rho, RotState &st)
{
    const double G = 6.667e-8, Dnl = 0.75;
    const std::string &pr = c.prior(i);
    if (c.is(i, "splitting_a1", "Splitting_a1")) {
        if (pr == "Fix_Auto") bad_keyword("splitting_a1", "Fix_Auto");
        Snlm.fill("Splitting_a1", pr, c.v(i, 0), c.row(i), 0, 1);
        int extra = 0;
        if (!a11_eq_a12 && avg_a1n) Snlm.fill(Snlm.names[0], Snlm.pnames[0], Snlm.inputs[0], c.row(i), 6, 1);          // a1(l): same prior for a1(2)
        if (a11_eq_a12 && !avg_a1n) extra = Nf_el[1];                                                                      // a1(n)
        if (!a11_eq_a12 && !avg_a1n) extra = Nf_el[1] + Nf_el[2];                                                          // a1(n,l)
        if (!avg_a1n) {
            for (int k = 0; k < extra; k++) Snlm.fill(Snlm.names[0], Snlm.pnames[0], Snlm.inputs[0], c.row(i), 6 + k, 1);
            Snlm.fill("Empty", "Fix", 0, c.row(i), 0, 1);                                                                  // the single-a1 slot is emptied
        }
    }
    if (c.is(i, "asphericity_eta", "Asphericity_eta")) {
        Snlm.names[1] = "Asphericity_eta";
        if (pr == "Fix_Auto") {                              // fixed to the centrifugal distortion computed from a1 and the mean density
            Snlm.pnames[1] = "Fix"; Snlm.relax[1] = 0;
            if (c.v(i, 0) == 1) Snlm.inputs[1] = (Dnu > 0) ? (double)((4. / 3.) * PI_L * Dnl * std::pow(Snlm.inputs[0] * 1e-6, 2.) / (rho * G)) : 0.0;
            else Snlm.inputs[1] = 0;
        } else Snlm.fill("Asphericity_eta", pr, c.v(i, 0), c.row(i), 1, 1);
    }
    if (c.is(i, "splitting_a3", "Splitting_a3")) { if (pr == "Fix_Auto") bad_keyword("splitting_a3", "Fix_Auto"); Snlm.fill("Splitting_a3", pr, c.v(i, 0), c.row(i), 2, 1); }
    if (c.is(i, "asymetry", "Asymetry")) { if (pr == "Fix_Auto") bad_keyword("asymetry", "Fix_Auto"); Snlm.fill("Lorentzian_asymetry", pr, c.v(i, 0), c.row(i), 5, 1); }
    if (c.is(i, "inclination", "Inclination")) {
        if (pr == "Fix_Auto") bad_keyword("inclination", "Fix_Auto");
        Inc.fill("Inclination", pr, c.v(i, 0) >= 90 ? 89.99999 : c.v(i, 0), c.row(i), 0, 1);
    }
    for (int which = 0; which < 2; which++) {
        const char *kw = which == 0 ? "sqrt(splitting_a1).cosi" : "sqrt(splitting_a1).sini";
        if (!c.is(i, kw)) continue;
        if (pr == "Fix_Auto") bad_keyword(kw, "Fix_Auto");
        Snlm.fill(kw, pr, c.v(i, 0), c.row(i), 3 + which, 1);
        if (!a11_eq_a12 || !avg_a1n) die("models *_a1n* / *_a1l* are not available with sqrt(a1).cosi and sqrt(a1).sini");
        (which == 0 ? st.a1cosi : st.a1sini) = true;
    }
}
// (a1, inclination) -> (sqrt(a1) cos i, sqrt(a1) sin i): io_ms_global.cpp:869-902, io_local.cpp:935-960
void project_a1_inclination(Block &Snlm, Block &Inc)
{
    const double ci = (double)(std::sqrt(Snlm.inputs[0]) * std::cos(Inc.inputs[0] * PI_L / 180.)), si = (double)(std::sqrt(Snlm.inputs[0]) * std::sin(Inc.inputs[0] * PI_L / 180.));
    if (Inc.pnames[0] == "Fix" && Snlm.pnames[0] == "Fix") {
        Snlm.fill("sqrt(splitting_a1).cosi", "Fix", ci, Snlm.prior_col(0), 3, 0);
        Snlm.fill("sqrt(splitting_a1).sini", "Fix", si, Snlm.prior_col(0), 4, 0);
    } else {
        Snlm.priors[0][1] = std::sqrt(Snlm.priors[0][1]);    // upper bound of a1 -> upper bound of sqrt(a1)
        Snlm.fill("sqrt(splitting_a1).cosi", Snlm.pnames[0], ci, Snlm.prior_col(0), 3, 0);
        Snlm.fill("sqrt(splitting_a1).sini", Snlm.pnames[0], si, Snlm.prior_col(0), 4, 0);
    }
    if (Snlm.inputs[3] < 1e-2) Snlm.inputs[3] = 1e-2;
    if (Snlm.inputs[4] < 1e-2) Snlm.inputs[4] = 1e-2;
    Inc.fill("Empty", "Fix", 0, Inc.prior_col(0), 0, 1);
    Snlm.fill("Empty", "Fix", 0, Snlm.prior_col(0), 0, 1);
}
int snlm_size(bool a11_eq_a12, bool avg_a1n, const int Nf_el[4])      // io_ms_global.cpp:598-618
{
    if (a11_eq_a12 && avg_a1n) return 6;
    if (!a11_eq_a12 && avg_a1n) return 7;
    if (a11_eq_a12 && !avg_a1n) { if (Nf_el[1] != Nf_el[2]) die("When considering a11=a22 you must have as many l=1 as l=2"); return 6 + Nf_el[1]; }
    return 6 + Nf_el[1] + Nf_el[2];
}
double mean_density(double Dnu)
{
    const double Dnu_sun = 135.1, R_sun = 6.96342e5, M_sun = 1.98855e30;
    const double rho_sun = (double)(M_sun * 1e3 / (4 * PI_L * std::pow(R_sun * 1e5, 3) / 3));
    return std::pow(Dnu / Dnu_sun, 2.) * rho_sun;
}
// per degree: the modes of the eigen table with their relax flags looked up (within 1e-2) in the first table
struct Degree { std::vector<double> f, fmin, fmax, w, h; std::vector<bool> rf, rw, rh; };
Degree collect_degree(const ModelFile &m, int el)
{
    Degree d;
    std::vector<double> f_el; std::vector<bool> rf, rw, rh;
    for (size_t i = 0; i < m.els.size(); i++) if (m.els[i] == el) { f_el.push_back(m.freqs_ref[i]); rf.push_back(m.relax_freq[i]); rw.push_back(m.relax_gamma[i]); rh.push_back(m.relax_H[i]); }
    for (size_t r = 0; r < m.eigen.size(); r++) {
        if ((int)m.eigen[r][0] != el) continue;
        const double nu = m.eigen[r][1];
        int found = -1, nfound = 0;
        for (size_t k = 0; k < f_el.size(); k++) if (f_el[k] > nu - 1e-2 && f_el[k] < nu + 1e-2) { if (!nfound) found = (int)k; nfound++; }
        if (nfound != 1) { char b[160]; snprintf(b, sizeof b, "The uniqueness of the frequency %.6f is not respected between the relax list and the eigen list of the .model file", nu); die(b); }
        d.f.push_back(nu); d.fmin.push_back(m.eigen[r][2]); d.fmax.push_back(m.eigen[r][3]); d.w.push_back(m.eigen[r][4]); d.h.push_back(m.eigen[r][5]);
        d.rf.push_back(rf[(size_t)found]); d.rw.push_back(rw[(size_t)found]); d.rh.push_back(rh[(size_t)found]);
    }
    return d;
}
void finish_vector(InputData &out, const std::vector<const Block *> &blocks, const std::vector<int> &offsets, double trunc_c, bool do_amp)
{
    int total = 0; for (int v : out.plength) total += v;
    out.all.init(total);
    for (size_t b = 0; b < blocks.size(); b++) out.all.put(*blocks[b], offsets[b]);
    const int p0 = total - 2;
    out.all.fill("Truncation parameter", "Fix", trunc_c, {}, p0, 1);
    if (out.all.inputs[(size_t)p0] <= 0) out.all.inputs[(size_t)p0] = 10000.;      // no truncation
    out.all.fill("Switch for fit of Amplitudes or Heights", "Fix", do_amp ? 1.0 : 0.0, {}, p0 + 1, 1);
}

// ------------------------------------------------------------------ build_init_MS_Global (io_ms_global.cpp:308-1090)
void set_noise_params(Block &N, const ModelFile &m)          // io_ms_global.cpp:1093-1190
{
    static const char *nm[3] = {"Harvey-Noise_H", "Harvey-Noise_tc", "Harvey-Noise_p"};
    for (int k = 0; k < 9; k++) { N.names[(size_t)k] = nm[k % 3]; N.pnames[(size_t)k] = k < 6 ? "Fix" : "Gaussian"; N.relax[(size_t)k] = k < 6 ? 0 : 1; }
    N.names[9] = "White_Noise_N0"; N.pnames[9] = "Gaussian"; N.relax[9] = 1;
    N.inputs = m.noise_params;
    for (int g = 0; g < 3; g++)
        if (N.inputs[(size_t)(3 * g)] <= 0 || N.inputs[(size_t)(3 * g + 1)] <= 0 || N.inputs[(size_t)(3 * g + 2)] <= 0)
            for (int k = 0; k < 3; k++) { N.pnames[(size_t)(3 * g + k)] = "Fix"; N.relax[(size_t)(3 * g + k)] = 0; N.inputs[(size_t)(3 * g + k)] = k == 2 ? 1 : 0; }
    const auto &s2 = m.noise_s2;
    for (int k = 6; k < 10; k++) N.priors[(size_t)k][0] = s2[(size_t)k][0];
    N.priors[6][1] = (s2[6][1] + s2[6][2]) * 3. / 2;
    N.priors[7][1] = (s2[7][1] + s2[7][2]) * 3. / 2;
    N.priors[8][1] = (s2[8][1] != 0) ? (s2[8][1] + s2[8][2]) * 3. / 2 : N.priors[8][0] * 0.1;
    N.priors[9][1] = (N.pnames[9] == "Uniform") ? (s2[9][1] + s2[9][2]) : N.priors[9][0] * 0.1;
    const double floor_[4] = {0.05, 0.005, 0.05, 0.0005};    // smallest relative uncertainty accepted on H3, tc3, p3, N0
    for (int k = 6; k < 10; k++)
        if ((N.priors[(size_t)k][1] / N.priors[(size_t)k][0]) <= floor_[k - 6] && N.pnames[(size_t)k] != "Fix") N.priors[(size_t)k][1] = N.priors[(size_t)k][0] * floor_[k - 6];
}
void set_width_appourchaux(Block &W, double numax, int version)     // io_ms_global.cpp:1231-1320
{
    std::vector<double> out;
    if (version == 2) out.push_back(numax);
    out.push_back(numax);                                           // nudip
    out.push_back(4. / 2150. * numax + (1. - 1000. * 4. / 2150.));  // alpha
    out.push_back(0.8 / 2150. * numax + (4.5 - 1000. * 0.8 / 2150.));
    out.push_back(3400. / 2150. * numax + (1000. - 1000. * 3400. / 2150.));
    out.push_back(2.8 / 2200. * numax + (1. - 2.8 / 2200. * 1.));
    static const char *n1[5] = {"nudip", "alpha", "Gamma_alpha", "Wdip", "DeltaGammadip"};
    static const char *n2[6] = {"numax", "nudip", "alpha", "Gamma_alpha", "Wdip", "DeltaGammadip"};
    const double fr1[5] = {0.1, 0.2, 0.2, 0.2, 0.4}, fr2[6] = {0.1, 0.1, 0.2, 0.2, 0.2, 0.4};
    for (size_t k = 0; k < out.size(); k++) {
        const std::string name = std::string("width:Appourchaux_v") + (version == 1 ? "1:" : "2:") + (version == 1 ? n1[k] : n2[k]);
        W.fill(name, "Gaussian", out[k], {out[k], out[k] * (version == 1 ? fr1[k] : fr2[k]), -9999, -9999}, (int)k, 0);
    }
}

InputData build_init_MS_Global(const ModelFile &m, double resol)
{
    const double Hmin = 1, Hmax = 10000;
    InputData out;
    Common c(m);
    bool a11_eq_a12 = true, avg_a1n = true, do_amp = false;
    int width_app = 0;
    double numax = m.numax, trunc_c = -1;
    int lmax = 0; for (int e : m.els) lmax = std::max(lmax, e);
    out.model_fullname = " ";
    for (size_t i = 0; i < c.n(); i++) {
        if (c.is(i, "model_fullname")) {
            const std::string &n = out.model_fullname = c.prior(i);
            if (n == "model_MS_Global_a1l_etaa3_HarveyLike") { a11_eq_a12 = false; avg_a1n = true; }
            else if (n == "model_MS_Global_a1n_etaa3_HarveyLike") { a11_eq_a12 = true; avg_a1n = false; }
            else if (n == "model_MS_Global_a1nl_etaa3_HarveyLike") { a11_eq_a12 = false; avg_a1n = false; }
            else if (n == "model_MS_Global_a1etaa3_AppWidth_HarveyLike_v1" || n == "model_MS_Global_a1etaa3_AppWidth_HarveyLike_v2") {
                a11_eq_a12 = avg_a1n = true; width_app = (n.back() == '1') ? 1 : 2;
                if (numax != -9999 && numax <= 0) die("model " + n + ": numax must be positive, or -9999 / absent to have it computed");
            } else if (n == "model_MS_Global_a1etaa3_HarveyLike_Classic" || n == "model_MS_Global_a1etaa3_HarveyLike_Classic_v2" || n == "model_MS_Global_a1etaa3_HarveyLike_Classic_v3" ||
                       n == "model_MS_Global_a1etaa3_HarveyLike" || n == "model_MS_Global_a1etaa3_Harvey1985" || n == "model_MS_Global_a1acta3_HarveyLike" || n == "model_MS_Global_a1acta3_Harvey1985") { a11_eq_a12 = avg_a1n = true; }
        }
        if (c.is(i, "fit_squareAmplitude_instead_Height")) { if (c.prior(i) != "bool") bad_keyword("fit_squareAmplitude_instead_Height", "bool"); do_amp = c.v(i, 0) != 0; }
    }
    if (out.model_fullname == " ") die("Model name empty. Check that the .model file contains the model_fullname variable.");
    Block Vis, Inc, Snlm, Noise, freq, height, width;
    Vis.init(lmax); Inc.init(1);
    // ---- frequencies / widths / heights: l = 0 carries the heights and widths, higher degrees only frequencies ----
    int Nf_el[4] = {0, 0, 0, 0};
    std::vector<double> f_in, fmin, fmax, w_in, h_in; std::vector<bool> f_rl, w_rl, h_rl;
    for (int el = 0; el <= lmax && el < 4; el++) {
        Degree d = collect_degree(m, el);
        Nf_el[el] = (int)d.f.size();
        f_in.insert(f_in.end(), d.f.begin(), d.f.end()); fmin.insert(fmin.end(), d.fmin.begin(), d.fmin.end()); fmax.insert(fmax.end(), d.fmax.begin(), d.fmax.end());
        f_rl.insert(f_rl.end(), d.rf.begin(), d.rf.end());
        if (el == 0) { w_in = d.w; h_in = d.h; w_rl = d.rw; h_rl = d.rh; }
    }
    std::string name_h = "Height_l0";
    if (do_amp) { name_h = "Amplitude_l0"; for (size_t i = 0; i < h_in.size(); i++) h_in[i] = (double)(PI_L * w_in[i] * h_in[i]); }
    height.init((int)h_rl.size());
    width.init(width_app == 0 ? (int)w_rl.size() : (width_app == 1 ? 5 : 6));
    freq.init((int)f_rl.size());
    // defaults (io_ms_global.cpp:522-553); the default widths take the HEIGHT values, as the reference does
    height.fill_vect(h_in, h_rl, name_h, "Jeffreys", {Hmin, Hmax, -9999., -9999.}, 0, 0, 0);
    if (width_app == 0 || w_in.size() <= (size_t)width.size()) {
        const std::vector<double> pw = {resol, m.Dnu / 3., -9999., -9999.};
        for (size_t i = 0; i < w_in.size(); i++) width.fill("Width_l0", w_rl[i] ? "Jeffreys" : "Fix", h_in[i], pw, (int)i, 0);
    }
    for (size_t i = 0; i < f_in.size(); i++) freq.fill("Frequency_l", f_rl[i] ? "GUG" : "Fix", f_in[i], {fmin[i], fmax[i], 0.01 * m.Dnu, 0.01 * m.Dnu}, (int)i, 0);
    if (numax <= 0) {                                        // height-weighted mean frequency (io_ms_global.cpp:556-592, getnumax :1213-1229)
        std::vector<double> wgt; const double vis[4] = {1.0, 1.5, 0.53, 0.08};
        for (int el = 0; el <= 3; el++) if (Nf_el[el] != 0) {
            // the reference weights every degree with the l=0 heights (it assumes as many modes per degree as l=0 modes)
            for (int k = 0; k < Nf_el[el]; k++) { const double hk = (size_t)k < height.inputs.size() ? height.inputs[(size_t)k] : 0.0; wgt.push_back(el == 0 ? hk : hk * vis[el]); }
        }
        double s = 0, H = 0; for (size_t k = 0; k < wgt.size(); k++) s = s + freq.inputs[k] * wgt[k];
        for (double v : wgt) H += v;
        numax = s / H;
    }
    Snlm.init(snlm_size(a11_eq_a12, avg_a1n, Nf_el));
    out.extra_priors[0] = 1; out.extra_priors[1] = 2.; out.extra_priors[2] = 0.2; out.extra_priors[3] = 0;
    const double rho = mean_density(m.Dnu);
    RotState rot;
    for (size_t i = 0; i < c.n(); i++) {
        const std::string &pr = c.prior(i);
        if (c.is(i, "freq_smoothness", "Freq_smoothness")) { if (pr != "bool") bad_keyword("freq_smoothness", "bool"); out.extra_priors[0] = c.v(i, 0); out.extra_priors[1] = c.v(i, 1); }
        if (c.is(i, "trunc_c")) { if (pr != "Fix") bad_keyword("trunc_c", "Fix"); trunc_c = c.v(i, 0); }
        if (c.is(i, "Frequency", "frequency")) {
            if (pr != "GUG" && pr != "Uniform") bad_keyword(m.common_names[i], "GUG or Uniform");
            for (size_t p = 0; p < f_in.size(); p++) {
                const std::vector<double> pv = (pr == "GUG") ? std::vector<double>{fmin[p], fmax[p], c.v(i, 3), c.v(i, 4)} : std::vector<double>{fmin[p], fmax[p], -9999, -9999};
                freq.fill("Frequency_l", f_rl[p] ? pr : "Fix", f_in[p], pv, (int)p, 0);
            }
        }
        if (c.is(i, "height", "Height") || c.is(i, "amplitude", "Amplitude")) {
            if (pr == "Fix_Auto") bad_keyword(m.common_names[i], "Fix_Auto");
            for (size_t p = 0; p < h_in.size(); p++) { if (h_rl[p]) height.fill(name_h, pr, h_in[p], c.row(i), (int)p, 0); else height.fill(name_h, "Fix", h_in[p], c.row(i), (int)p, 1); }
        }
        if (c.is(i, "width", "Width") && width_app == 0) {
            std::string pn = pr; std::vector<double> pv = c.row(i);
            if (pr == "Fix_Auto") { pn = "Jeffreys"; pv = {resol, m.Dnu / 3., -9999., -9999.}; }
            for (size_t p = 0; p < w_in.size(); p++) { if (w_rl[p]) width.fill("Width_l", pn, w_in[p], pv, (int)p, 0); else width.fill("Width_l", "Fix", w_in[p], c.row(i), (int)p, 1); }
        }
        rotation_keywords(c, i, Snlm, Inc, a11_eq_a12, avg_a1n, Nf_el, 1.0 /* eta uses rho whatever Dnu */, rho, rot);
        for (int l = 1; l <= 3; l++) {
            const std::string lo = "visibility_l" + std::to_string(l), up = "Visibility_l" + std::to_string(l);
            if (!c.is(i, lo.c_str(), up.c_str())) continue;
            if (pr == "Fix_Auto") bad_keyword(lo, "Fix_Auto");
            if (lmax >= l) Vis.fill(up, pr, c.v(i, 0), c.row(i), l - 1, 1);
        }
    }
    if (rot.a1cosi != rot.a1sini) die("Both 'sqrt(splitting_a1).sini' and 'sqrt(splitting_a1).cosi' keywords must appear");
    const std::string &mn = out.model_fullname;
    if (!rot.a1cosi) {
        if (mn == "model_MS_Global_a1etaa3_HarveyLike" || mn == "model_MS_Global_a1etaa3_Harvey1985" || width_app != 0) project_a1_inclination(Snlm, Inc);
        if (mn == "model_MS_Global_a1etaa3_HarveyLike_Classic_v2") {          // 2 + 3 + 4 free height ratios (io_ms_global.cpp:904-919)
            const double inc0 = Inc.inputs[0];
            Inc.init(9);
            int ind = 0;
            for (int el = 1; el <= lmax; el++) { auto r = amplitude_ratio(el, inc0); for (int em = 0; em <= el; em++) Inc.fill("Inc:H" + std::to_string(el) + "," + std::to_string(em), "Uniform", r[(size_t)(el + em)], {0, 1, -9999., -9999.}, ind++, 0); }
            out.extra_priors[3] = 1;
        }
        if (mn == "model_MS_Global_a1etaa3_HarveyLike_Classic_v3") {          // one height per (n, l, |m|) (io_ms_global.cpp:920-951)
            const double inc0 = Inc.inputs[0]; const std::vector<double> vis0 = Vis.inputs;
            for (int el = 1; el < lmax; el++) Vis.fill("Empty", "Fix", 0, {}, el - 1, 0);
            Inc.init(Nf_el[1] * 2 + Nf_el[2] * 3 + Nf_el[3] * 4);
            int ind = 0;
            for (int el = 1; el <= lmax; el++) { auto r = amplitude_ratio(el, inc0);
                for (int en = 0; en < Nf_el[el]; en++) for (int em = 0; em <= el; em++)
                    Inc.fill("Inc: H" + std::to_string(en) + "," + std::to_string(el) + "," + std::to_string(em), "Jeffreys", height.inputs.at((size_t)en) * vis0[(size_t)(el - 1)] * r[(size_t)(el + em)], {Hmin, Hmax, -9999., -9999.}, ind++, 0); }
            out.extra_priors[3] = 2;
        }
    } else {
        if (mn == "model_MS_Global_a1etaa3_HarveyLike_Classic") die("model_MS_Global_a1etaa3_HarveyLike_Classic cannot be used with sqrt(splitting_a1).cosi / .sini: use Inclination and Splitting_a1");
        Inc.fill("Empty", "Fix", 0, Inc.prior_col(0), 0, 1);
        Snlm.fill("Empty", "Fix", 0, Snlm.prior_col(0), 0, 1);
    }
    Noise.init(10);
    set_noise_params(Noise, m);
    if (width_app) set_width_appourchaux(width, numax, width_app);
    out.plength = {(int)h_in.size(), lmax, Nf_el[0], Nf_el[1], Nf_el[2], Nf_el[3], Snlm.size(), (int)w_in.size(), Noise.size(), Inc.size(), 2};
    const auto &pl = out.plength;
    std::vector<int> off = {0, pl[0], pl[0] + pl[1]};
    off.push_back(off[2] + pl[2] + pl[3] + pl[4] + pl[5]); off.push_back(off[3] + pl[6]); off.push_back(off[4] + pl[7]); off.push_back(off[5] + pl[8]);
    if (width.size() > pl[7]) die("the Appourchaux width models need at least " + std::to_string(width.size()) + " l=0 modes in the .model file: their width parameters live in the l=0 width slots (io_ms_global.cpp:1001, :1025-1032)");
    finish_vector(out, {&height, &Vis, &freq, &Snlm, &width, &Noise, &Inc}, off, trunc_c, do_amp);
    return out;
}

// ------------------------------------------------------------------ build_init_local (io_local.cpp:303-1158)
template <typename T> std::vector<T> keep_in_range(const std::vector<T> &v, const std::vector<double> &f, double lo, double hi)   // filter_range, string_handler.cpp:486-529
{
    std::vector<T> o; for (size_t i = 0; i < f.size(); i++) if (f[i] > lo && f[i] < hi) o.push_back(v[i]); return o;
}
InputData build_init_local(const ModelFile &m, double resol)
{
    const double Hmin = 1, Hmax = 10000;
    InputData out;
    Common c(m);
    bool do_amp = false;
    double trunc_c = -1;
    int lmax = 0; for (int e : m.els) lmax = std::max(lmax, e);
    out.model_fullname = " ";
    for (size_t i = 0; i < c.n(); i++) {
        if (c.is(i, "model_fullname")) out.model_fullname = c.prior(i);
        if (c.is(i, "fit_squareAmplitude_instead_Height")) { if (c.prior(i) != "bool") bad_keyword("fit_squareAmplitude_instead_Height", "bool"); do_amp = c.v(i, 0) != 0; }
    }
    if (out.model_fullname == " ") die("Model name empty. Check that the .model file contains the model_fullname variable.");
    const bool hnlm = out.model_fullname == "model_MS_local_Hnlm";
    Degree D[4];
    int Nf_el[4] = {0, 0, 0, 0};
    for (int el = 0; el <= lmax && el < 4; el++) {
        bool any = false; for (int e : m.els) any = any || e == el;
        if (!any) continue;
        Degree d = collect_degree(m, el);
        const std::vector<double> f0 = d.f;                  // only the modes strictly inside the analysed slice are kept
        const double lo = m.freq_range[0], hi = m.freq_range[1];
        D[el].f = keep_in_range(d.f, f0, lo, hi); D[el].h = keep_in_range(d.h, f0, lo, hi); D[el].w = keep_in_range(d.w, f0, lo, hi);
        D[el].rf = keep_in_range(d.rf, f0, lo, hi); D[el].rh = keep_in_range(d.rh, f0, lo, hi); D[el].rw = keep_in_range(d.rw, f0, lo, hi);
        D[el].fmin = keep_in_range(d.fmin, f0, lo, hi); D[el].fmax = keep_in_range(d.fmax, f0, lo, hi);
        Nf_el[el] = (int)D[el].f.size();
    }
    const int Ntot = Nf_el[0] + Nf_el[1] + Nf_el[2] + Nf_el[3];
    if (Ntot == 0) die("No mode of the .model file lies inside the analysed frequency range");
    std::string name_h = "Height_l";
    if (do_amp) { name_h = "Amplitude_l"; for (int el = 0; el < 4; el++) for (size_t i = 0; i < D[el].h.size(); i++) D[el].h[i] = (double)(PI_L * D[el].w[i] * D[el].h[i]); }
    int offs[4] = {0, Nf_el[0], Nf_el[0] + Nf_el[1], Nf_el[0] + Nf_el[1] + Nf_el[2]};
    Block Inc, Snlm, Noise, freq, height, width;
    Inc.init(1);
    const std::vector<double> ph = {Hmin, Hmax, -9999., -9999.};
    if (hnlm) { height.init(Nf_el[0] + Nf_el[1] * 2 + Nf_el[2] * 3 + Nf_el[3] * 4); height.fill_vect(D[0].h, D[0].rh, name_h, "Jeffreys", ph, 0, 0, 0); }
    else { height.init(Ntot); for (int el = 0; el < 4; el++) height.fill_vect(D[el].h, D[el].rh, name_h, "Jeffreys", ph, offs[el], 0, 0); }
    width.init(Ntot); freq.init(Ntot);
    const std::vector<double> pw = {resol, m.Dnu > 0 ? m.Dnu / 3. : 20., -9999., -9999.};
    for (int el = 0; el < 4; el++) width.fill_vect(D[el].w, D[el].rw, "Width_l", "Jeffreys", pw, offs[el], 0, 0);
    for (int el = 0; el < 4; el++) for (size_t i = 0; i < D[el].f.size(); i++) {
        const double s = 0.01 * std::abs(D[el].fmax[i] - D[el].fmin[i]);
        freq.fill("Frequency_l", D[el].rf[i] ? "GUG" : "Fix", D[el].f[i], {D[el].fmin[i], D[el].fmax[i], s, s}, (int)i + offs[el], 0);
    }
    Snlm.init(6);
    out.extra_priors[0] = 0; out.extra_priors[1] = 0; out.extra_priors[2] = 0.2; out.extra_priors[3] = 0;
    const double rho = mean_density(m.Dnu);
    RotState rot;
    int pos_prior_height = -1;
    int p0 = offs[3];                                        // value the reference's p0 holds when the keyword loop starts (io_local.cpp:656)
    for (size_t i = 0; i < c.n(); i++) {
        const std::string &pr = c.prior(i);
        if (c.is(i, "trunc_c")) { if (pr != "Fix") bad_keyword("trunc_c", "Fix"); trunc_c = c.v(i, 0); }
        if (c.is(i, "height", "Height") || c.is(i, "amplitude", "Amplitude")) {
            const bool amp = c.is(i, "amplitude", "Amplitude");
            if (pr == "Fix_Auto") {                          // Jeffreys between h / v0 and h * v1, per mode
                for (int el = 0; el < (hnlm ? 1 : 4); el++) {
                    for (size_t k = 0; k < D[el].h.size(); k++) {
                        const double hk = D[el].h[k];
                        const std::vector<double> pv = amp ? std::vector<double>{(double)(PI_L * m.Dnu / 3. * hk / c.v(i, 0)), (double)(PI_L * m.Dnu / 3. * hk * c.v(i, 1)), -9999, -9999}
                                                           : std::vector<double>{hk / c.v(i, 0), hk * c.v(i, 1), -9999, -9999};
                        height.fill(name_h, D[el].rh[k] ? "Jeffreys" : "Fix", hk, pv, (int)k + (hnlm ? p0 : offs[el]), 0);
                    }
                }
                if (hnlm) pos_prior_height = (int)i; else p0 = offs[3];
            } else if (hnlm) { pos_prior_height = (int)i; height.fill_vect(D[0].h, D[0].rh, name_h, pr, c.row(i), p0, 0, 0); }
            else { for (int el = 0; el < 4; el++) height.fill_vect(D[el].h, D[el].rh, name_h, pr, c.row(i), offs[el], 1, 1); p0 = offs[3]; }
        }
        if (c.is(i, "width", "Width")) {
            std::string pn = pr; std::vector<double> pv = c.row(i);
            if (pr == "Fix_Auto") { pn = "Jeffreys"; pv = {resol, m.Dnu > 0 ? m.Dnu / 3 : 20., -9999, -9999}; }
            for (int el = 0; el < 4; el++) width.fill_vect(D[el].w, D[el].rw, "Width_l", pn, pv, offs[el], 0, 1);
            p0 = offs[3];
        }
        rotation_keywords(c, i, Snlm, Inc, true, true, Nf_el, m.Dnu, rho, rot);
        // the reference's single p0 variable is left at the slot index of the last keyword; only the Hnlm height branch reads it
        if (c.is(i, "splitting_a1", "Splitting_a1") || c.is(i, "inclination", "Inclination")) p0 = 0;
        if (c.is(i, "asphericity_eta", "Asphericity_eta") && pr != "Fix_Auto") p0 = 1;
        if (c.is(i, "splitting_a3", "Splitting_a3")) p0 = 2;
        if (c.is(i, "asymetry", "Asymetry")) p0 = 5;
        if (c.is(i, "sqrt(splitting_a1).cosi")) p0 = 3;
        if (c.is(i, "sqrt(splitting_a1).sini")) p0 = 4;
    }
    if (rot.a1cosi != rot.a1sini) die("Both 'sqrt(splitting_a1).sini' and 'sqrt(splitting_a1).cosi' keywords must appear");
    if (!rot.a1cosi) {
        if (out.model_fullname == "model_MS_local_basic") project_a1_inclination(Snlm, Inc);
        if (hnlm) {                                          // one height per (n, l, |m|), from the visibilities at the input inclination (io_local.cpp:962-1003)
            const double inc0 = Inc.inputs[0];
            Inc.fill("Empty", "Fix", 0, Inc.prior_col(0), 0, 1);
            int ind = (int)D[0].h.size();
            std::vector<double> pv = {Hmin, Hmax, -9999., -9999.}; std::string pn = "Jeffreys";
            if (pos_prior_height > 0) { pv = c.row((size_t)pos_prior_height); pn = c.prior((size_t)pos_prior_height); }
            for (int el = 1; el <= lmax && el < 4; el++) { auto r = amplitude_ratio(el, inc0);
                for (int en = 0; en < Nf_el[el]; en++) for (int em = 0; em <= el; em++)
                    height.fill("H(" + std::to_string(en) + "," + std::to_string(el) + "," + std::to_string(em) + ")", pn, D[el].h[(size_t)en] * r[(size_t)(el + em)], pv, ind++, 0); }
            out.extra_priors[3] = 2;
        }
    } else {
        Inc.fill("Empty", "Fix", 0, Inc.prior_col(0), 0, 1);
        Snlm.fill("Empty", "Fix", 0, Snlm.prior_col(0), 0, 1);
    }
    // white noise: mean / 0.5 min / 1.5 max of the Harvey background at the two ends of the slice (set_noise_params_local, io_local.cpp:1160-1212)
    Noise.init(1);
    Noise.names[0] = "White_Noise_N0"; Noise.pnames[0] = "Uniform"; Noise.relax[0] = 1;
    {
        std::vector<double> np;
        for (double v : m.noise_params) { if (v == -1) np.push_back(0); if (v >= 0) np.push_back(v); }
        int n_minus2 = 0; for (double v : m.noise_params) if (v > -2 - 1e-6 && v < -2 + 1e-6) n_minus2++;
        np.resize(m.noise_params.size() - (size_t)n_minus2, -9999.0);
        const int Nh = ((int)np.size() - 1) / 3;
        auto nv = harvey_like(np, m.freq_range, 2, Nh);
        Noise.inputs[0] = (nv[0] + nv[1]) / 2;
        Noise.priors[0][0] = std::min(nv[0], nv[1]) * 0.5; Noise.priors[0][1] = std::max(nv[0], nv[1]) * 1.5;
    }
    const int nh = hnlm ? Nf_el[0] + 2 * Nf_el[1] + 3 * Nf_el[2] + 4 * Nf_el[3] : Ntot;
    out.plength = {nh, 0, Nf_el[0], Nf_el[1], Nf_el[2], Nf_el[3], Snlm.size(), Ntot, Noise.size(), Inc.size(), 2};
    const auto &pl = out.plength;
    std::vector<int> off = {0, pl[0] + pl[1]};
    off.push_back(off[1] + pl[2] + pl[3] + pl[4] + pl[5]); off.push_back(off[2] + pl[6]); off.push_back(off[3] + pl[7]); off.push_back(off[4] + pl[8]);
    finish_vector(out, {&height, &freq, &Snlm, &width, &Noise, &Inc}, off, trunc_c, do_amp);
    return out;
}

// ------------------------------------------------------------------ configuration files
struct ListFile { std::vector<int> ids; std::vector<std::string> names; };
ListFile read_listfile(const std::string &file)              // Config::read_listfiles, config.cpp:1780-1840
{
    std::ifstream in(file.c_str());
    if (!in.is_open()) die("Could not open " + file);
    ListFile l; std::string line;
    while (std::getline(in, line)) {
        line = trim(line);
        if (line.empty() || line[0] == '#') continue;
        auto w = split(line, " \t");
        if (w.size() < 2) continue;
        l.ids.push_back(to_int(w[0])); l.names.push_back(w[1]);
    }
    return l;
}
int lookup(const ListFile &l, const std::string &name, const char *what)
{
    int v = -9999; bool ok = false;
    for (size_t i = 0; i < l.names.size(); i++) if (l.names[i] == name) { v = l.ids[i]; ok = true; }     // the last match wins, like the reference's loops
    if (!ok) die(std::string("unknown ") + what + " '" + name + "' (not in the *_ctrl.list file)");
    return v;
}
// "key = value ; comment" lines (Config::format_line, config.cpp:622-668); group lines start with '!'
bool key_value(const std::string &raw, std::string &key, std::string &val)
{
    const std::string s = trim(raw);
    if (s.empty() || s[0] == '#' || s[0] == '!' || s[0] == '/') return false;
    const size_t sc = s.find(';');
    if (sc == std::string::npos) die("configuration line without the ';' terminator: " + s);
    const std::string body = trim(s.substr(0, sc));
    const size_t eq = body.find('=');
    if (eq == std::string::npos) return false;
    key = trim(body.substr(0, eq)); val = trim(body.substr(eq + 1));
    return true;
}

}  // namespace

struct tg_frontend {
    std::string workdir;
    // config_default.cfg: group -> key -> value (the two output_root_name / output_dir keys live in different groups)
    std::map<std::string, std::map<std::string, std::string>> cfg;
    std::vector<std::string> err_names; std::vector<double> err_frac, err_off;
    ListFile models, priors, likelihoods, primepriors;
    // config_presets.cfg
    bool force_manual = false; std::string manual_file, models_dir, out_dir;
    std::vector<std::string> processing, core_out, core_in; std::vector<int> Nsamples, restore; std::vector<double> c0;
    int first_process = 0, last_process = 0;
    std::vector<std::vector<std::string>> table_ids;
    std::vector<long> Nt_learn0; long dN_mixing0 = 1;
    // ---- state of the last setup ----
    bool ready = false;
    InputData in; std::vector<int> pswitch; std::vector<double> priors_flat;
    std::vector<double> x, y, sig; bool have_sig = false;
    std::vector<double> Tcoefs, init_err, sigma0;
    std::vector<long> Nt_learn, periods_learn;
    int model_case = 0, like_case = 0, prior_case = 0, Nchains = 0, Nslices = 1;
    long dN_mixing = 1;
    double likelihood_params = 0, c0_phase = 0, eps1 = 0, eps2 = 0, A1 = 0, target = 0, lambda = 0;
    std::string star_id, phase, model_name, like_name, prior_name, dir_out, restore_dir, diags_dir, root, rin, rout;
    std::string s_params, s_stat, s_ptemp, s_acc, s_evid;
    tg_fe_run run{};

    const std::string &get(const char *grp, const char *key) const
    {
        static const std::string empty;
        auto g = cfg.find(grp); if (g == cfg.end()) return empty;
        auto k = g->second.find(key); return k == g->second.end() ? empty : k->second;
    }
    std::string need(const char *grp, const char *key) const { auto g = cfg.find(grp); if (g == cfg.end() || !g->second.count(key)) die(std::string("config_default.cfg: keyword '") + key + "' of group " + grp + " is missing"); return g->second.at(key); }
};

namespace {

void read_default_cfg(tg_frontend &fe, const std::string &file)   // Config::read_cfg_file, config.cpp:809-1319
{
    std::ifstream in(file.c_str());
    if (!in.is_open()) die("Could not open " + file);
    std::string line, grp;
    while (std::getline(in, line)) {
        const std::string s = trim(line);
        if (s.rfind("/END", 0) == 0) break;
        if (!s.empty() && s[0] == '!') { grp = trim(s.substr(0, s.find(':'))); continue; }
        std::string k, v;
        if (key_value(line, k, v)) fe.cfg[grp][k] = v;
    }
    for (const char *g : {"!MALA", "!Modeling", "!Data", "!Outputs", "!Diagnostics"}) if (!fe.cfg.count(g)) die(std::string("config_default.cfg: group ") + g + " is missing");
}
void read_errors_cfg(tg_frontend &fe, const std::string &file)    // Config::read_defautlerrors, config.cpp:1611-1662
{
    std::ifstream in(file.c_str());
    if (!in.is_open()) die("Could not open " + file);
    std::string line;
    while (std::getline(in, line)) {
        line = trim(line);
        if (line.empty() || line[0] == '#') continue;
        auto w = split(line, " \t");
        if (w.size() < 3) continue;
        fe.err_names.push_back(trim(w[0])); fe.err_frac.push_back(to_dbl(w[1])); fe.err_off.push_back(to_dbl(w[2]));
    }
}
void read_presets(tg_frontend &fe, const std::string &file)       // Config_presets::read_cfg_file_presets, config_presets.cpp:226-383
{
    std::ifstream in(file.c_str());
    if (!in.is_open()) die("Could not open the master configuration file " + file);
    std::string line; int found = 0;
    auto list = [](const std::string &v) { std::vector<std::string> o; for (auto &w : split(v, ",")) o.push_back(trim(w)); return o; };
    while (std::getline(in, line)) {
        if (trim(line).rfind("/END", 0) == 0) break;
        std::string k, v;
        if (!key_value(line, k, v)) continue;
        found++;
        if (k == "force_manual_config") fe.force_manual = to_bool(v);
        else if (k == "manual_config_file") fe.manual_file = v;
        else if (k == "cfg_models_dir") fe.models_dir = v;
        else if (k == "cfg_out_dir") fe.out_dir = v;
        else if (k == "processing") fe.processing = list(v);
        else if (k == "Nsamples") { for (auto &w : list(v)) if (w != "") fe.Nsamples.push_back(to_int(w)); }
        else if (k == "c0") { for (auto &w : list(v)) fe.c0.push_back(to_dbl(w)); }
        else if (k == "restore") { for (auto &w : list(v)) if (w != "") fe.restore.push_back(to_int(w)); }
        else if (k == "core_out") fe.core_out = list(v);
        else if (k == "core_in") fe.core_in = list(v);
        else if (k == "start_index_processing") fe.first_process = (int)to_dbl(v);
        else if (k == "last_index_processing") fe.last_process = (int)to_dbl(v);
        else if (k == "table_ids") {
            auto sz = to_dbls(v, ",");
            if (sz.size() < 2) die("config_presets.cfg: table_ids needs 'rows, columns'");
            for (int i = 0; i < (int)sz[0]; i++) {
                if (!std::getline(in, line)) die("config_presets.cfg: table_ids announces more rows than the file holds");
                std::string row = trim(line); const size_t sc = row.find(';'); if (sc != std::string::npos) row = row.substr(0, sc);
                auto w = split(row, " ");
                if ((int)w.size() < (int)sz[1]) die("config_presets.cfg: a table_ids row has too few columns");
                fe.table_ids.push_back(std::vector<std::string>(w.begin(), w.begin() + (int)sz[1]));
            }
        } else found--;
    }
    if (found != 13) die("config_presets.cfg: 13 keywords expected (force_manual_config, manual_config_file, cfg_models_dir, cfg_out_dir, processing, Nsamples, c0, restore, core_out, core_in, start_index_processing, last_index_processing, table_ids)");
    const size_t np = fe.processing.size();
    if (fe.Nsamples.size() < np || fe.c0.size() < np || fe.restore.size() < np || fe.core_out.size() < np || fe.core_in.size() < np) die("config_presets.cfg: every phase needs Nsamples, c0, restore, core_out and core_in");
}

struct DataFile { std::vector<std::vector<double>> cols; long rows = 0; };
DataFile read_data(const std::string &file)                 // Config::read_data_ascii_Ncols, config.cpp:509-620
{
    Lines f(file);
    if (!f.ok) die("Could not open the data file " + file);
    f.next();
    std::string l = trim(f.cur);
    while (first_char(l) == "#" && !f.eof()) { f.next(); l = trim(f.cur); }                      // header
    bool labels = false, units = false;
    if (first_char(l) == "!") { labels = true; f.next(); l = trim(f.cur); }                      // labels
    if (first_char(l) == "*") { units = true; }                                                  // units
    if (labels || units) { if (units) { f.next(); } l = trim(f.cur); }
    // NB the reference reads one more line when either a label or a unit line exists (config.cpp:573-575); with
    // both present (the usual case) that is the line after the units, with labels only the first data row is skipped.
    if (labels && !units) { f.next(); }
    DataFile d;
    const long cap = 1000000;                                // data_Maxsize, config.cpp:531
    size_t ncols = 0;
    while (!f.eof()) {
        auto w = split(trim(f.cur), " \t");
        if (d.rows == 0) { ncols = w.size(); d.cols.assign(ncols, std::vector<double>()); }
        if (d.rows >= cap) die("data file longer than 1000000 rows (the reference's reader cap)");
        for (size_t i = 0; i < ncols; i++) {
            long double v; double x;
            if (i < w.size() && (std::istringstream(w[i]) >> v)) x = (double)v; else x = std::nan("");
            d.cols[i].push_back(x);
        }
        d.rows++;
        f.next();
    }
    return d;
}
void mkdirs(const std::string &root, const std::string &id)  // Config_presets::generate_default_dirtree, config_presets.cpp:411-423
{
    for (const char *s : {"", "/diags", "/diags/pdfs", "/restore", "/outputs"}) { const std::string p = root + "/" + id + s; mkdir(p.c_str(), 0777); }
}
std::vector<long> to_longs(const std::string &v) { std::vector<long> o; for (auto &w : split(v, ",")) if (trim(w) != "") o.push_back((long)to_dbl(w)); return o; }

void setup(tg_frontend &fe, int star, int slice, int ph)
{
    if (fe.force_manual) die("force_manual_config=1 is not supported: the presets drive the run");
    if (star < 0 || star >= (int)fe.table_ids.size()) die("star index outside table_ids");
    if (ph < 0 || ph >= (int)fe.processing.size()) die("phase index outside 'processing'");
    fe.ready = false;
    const std::string id = fe.table_ids[(size_t)star][0];
    fe.star_id = id; fe.phase = fe.processing[(size_t)ph];
    // ---- Config_presets::apply_presets, config_presets.cpp:35-190 ----
    fe.dir_out = fe.out_dir + "/" + id + "/outputs/"; fe.restore_dir = fe.out_dir + "/" + id + "/restore/"; fe.diags_dir = fe.out_dir + "/" + id + "/diags/";
    mkdirs(fe.out_dir, id);
    const long ns = fe.Nsamples[(size_t)ph];
    fe.Nt_learn = fe.Nt_learn0; fe.dN_mixing = fe.dN_mixing0;
    if (fe.phase == "Burn-in") fe.Nt_learn.back() = ns + 1;                       // never stop learning
    else if (fe.phase == "Learning") { fe.Nt_learn.back() = ns + 1; fe.dN_mixing = ns + 1; }          // ... and never mix the chains
    else if (fe.phase == "Acquire") { for (size_t i = 0; i < fe.Nt_learn.size(); i++) fe.Nt_learn[i] = ns + 1 + (long)i; }   // never learn
    else die("unknown phase '" + fe.phase + "' (allowed: Burn-in, Learning, Acquire)");
    const std::string modelfile = fe.models_dir + id + ".model", datafile = fe.models_dir + id + ".data";
    { int n = 0; std::ifstream in(modelfile.c_str()); if (!in.is_open()) die("Unable to open the file: " + modelfile); std::string l; while (std::getline(in, l)) { if (first_char(trim(l)) == "*") n++; } fe.Nslices = n; }
    const std::string sl = fe.Nslices == 1 ? "" : std::to_string(slice + 1) + "_";
    fe.root = id + "_" + sl + fe.core_out[(size_t)ph] + "_";
    fe.rin = id + "_" + sl + "restore_" + fe.core_in[(size_t)ph] + "_"; fe.rout = id + "_" + sl + "restore_" + fe.core_out[(size_t)ph] + "_";
    const int rs = fe.restore[(size_t)ph];
    if (rs < 0 || rs > 3) die("config_presets.cfg: restore must be 0, 1, 2 or 3");
    tg_fe_run &r = fe.run;
    r.do_restore_variables = rs >= 1; r.do_restore_proposal = rs >= 2; r.do_restore_last_index = rs == 3; r.erase_old_files = rs != 3;
    r.do_restore_proposal_mean = to_bool(fe.get("!Outputs", "do_restore_proposal_mean"));
    r.Nsamples = ns; r.Nbuffer = (long)to_dbl(fe.need("!Outputs", "Nbuffer"));
    fe.c0_phase = fe.c0[(size_t)ph];
    // ---- Config::setup, config.cpp:80-189 ----
    DataFile d = read_data(datafile);
    const int xc = to_int(fe.need("!Data", "x_col")), yc = to_int(fe.need("!Data", "y_col")), sc = to_int(fe.need("!Data", "ysig_col"));
    if (xc < 0 || xc >= (int)d.cols.size() || yc < 0 || yc >= (int)d.cols.size() || d.rows < 3) die("data file: x_col / y_col outside the columns of " + datafile);
    const double resol = d.cols[(size_t)xc][2] - d.cols[(size_t)xc][1];
    fe.prior_name = fe.need("!Modeling", "prior_fct_name");
    fe.like_name = fe.need("!Modeling", "likelihood_fct_name");
    fe.likelihood_params = to_dbl(fe.need("!Modeling", "likelihood_params"));
    if (fe.table_ids[(size_t)star].size() > 1) { /* the second column of table_ids documents likelihood_params; the reference does not apply it */ }
    ModelFile mf;
    if (fe.prior_name == "io_MS_Global") { mf = read_model_file(modelfile, -1); fe.in = build_init_MS_Global(mf, resol); }
    else if (fe.prior_name == "io_local") { mf = read_model_file(modelfile, slice); fe.in = build_init_local(mf, resol); }
    else die("prior_fct_name '" + fe.prior_name + "': only io_MS_Global and io_local .model files are read by this front end");
    fe.model_name = fe.in.model_fullname;
    fe.pswitch.clear();
    for (auto &n : fe.in.all.pnames) fe.pswitch.push_back(lookup(fe.primepriors, n, "prior"));
    fe.model_case = lookup(fe.models, fe.model_name, "model"); fe.like_case = lookup(fe.likelihoods, fe.like_name, "likelihood"); fe.prior_case = lookup(fe.priors, fe.prior_name, "prior function");
    const long N = d.rows; const auto &xs = d.cols[(size_t)xc];
    long imin = 0; while (imin < N && xs[(size_t)imin] < mf.freq_range[0]) imin++;
    if (imin >= N) die("the frequency range of the .model file lies beyond the data");
    long imax = imin; while (imax < N && xs[(size_t)imax] < mf.freq_range[1]) imax++;
    fe.x.assign(xs.begin() + imin, xs.begin() + imax); fe.y.assign(d.cols[(size_t)yc].begin() + imin, d.cols[(size_t)yc].begin() + imax);
    fe.have_sig = sc >= 0 && sc < (int)d.cols.size();
    if (fe.have_sig) fe.sig.assign(d.cols[(size_t)sc].begin() + imin, d.cols[(size_t)sc].begin() + imax); else fe.sig.assign(fe.x.size(), 1.0);
    // ---- MALA::MALA + init_proposal, MALA.cpp:53-112, :246-289 ----
    fe.Nchains = (int)to_dbl(fe.need("!MALA", "Nchains")); fe.lambda = to_dbl(fe.need("!MALA", "lambda_temp"));
    fe.eps1 = to_dbl(fe.need("!MALA", "epsilon1")); fe.eps2 = to_dbl(fe.need("!MALA", "epsilon2")); fe.A1 = to_dbl(fe.need("!MALA", "A1")); fe.target = to_dbl(fe.need("!MALA", "target_acceptance"));
    fe.periods_learn = to_longs(fe.need("!MALA", "periods_learn"));
    if (fe.Nchains < 1 || fe.periods_learn.size() + 1 != fe.Nt_learn.size()) die("config_default.cfg: Nchains >= 1 and periods_learn.size() == Nt_learn.size() - 1 are required");
    fe.Tcoefs.resize((size_t)fe.Nchains); for (int mch = 0; mch < fe.Nchains; mch++) fe.Tcoefs[(size_t)mch] = std::pow(fe.lambda, mch);
    fe.init_err.clear(); fe.priors_flat.clear();
    const Block &a = fe.in.all;
    int nv = 0;
    for (int i = 0; i < a.size(); i++) {
        for (int k = 0; k < 4; k++) fe.priors_flat.push_back(a.priors[(size_t)i][(size_t)k]);
        if (a.relax[(size_t)i] != 1) continue;
        double e = 1;                                        // names without an entry in errors_default.cfg keep an error of 1
        for (size_t j = 0; j < fe.err_names.size(); j++) if (a.names[(size_t)i] == fe.err_names[j]) e = a.inputs[(size_t)i] * fe.err_frac[j] + fe.err_off[j];
        fe.init_err.push_back(e); nv++;
    }
    if (nv == 0) die("no free parameter in the .model file");
    fe.sigma0.resize((size_t)fe.Nchains); for (int mch = 0; mch < fe.Nchains; mch++) fe.sigma0[(size_t)mch] = std::pow(2.38, 2) * std::pow(fe.Tcoefs[(size_t)mch], 0.2) / nv;
    // ---- the rest of the run description ----
    fe.s_params = fe.need("!Outputs", "params_txt_fileout"); fe.s_stat = fe.need("!Outputs", "stat_txt_fileout"); fe.s_ptemp = fe.need("!Outputs", "parallel_tempering_txt_fileout");
    fe.s_acc = fe.need("!Outputs", "acceptance_txt_fileout"); fe.s_evid = fe.need("!Diagnostics", "file_evidence_diags");
    r.star_id = fe.star_id.c_str(); r.phase = fe.phase.c_str(); r.model_fct_name = fe.model_name.c_str(); r.likelihood_fct_name = fe.like_name.c_str(); r.prior_fct_name = fe.prior_name.c_str();
    r.dir_out = fe.dir_out.c_str(); r.restore_dir = fe.restore_dir.c_str(); r.diags_dir = fe.diags_dir.c_str(); r.output_root_name = fe.root.c_str();
    r.restore_file_in = fe.rin.c_str(); r.restore_file_out = fe.rout.c_str();
    r.params_txt_fileout = fe.s_params.c_str(); r.stat_txt_fileout = fe.s_stat.c_str(); r.parallel_tempering_txt_fileout = fe.s_ptemp.c_str(); r.acceptance_txt_fileout = fe.s_acc.c_str();
    r.file_evidence_diags = fe.s_evid.c_str();
    r.get_params = to_bool(fe.get("!Outputs", "get_params")); r.get_statcriteria = to_bool(fe.get("!Outputs", "get_statcriteria")); r.get_parallel_tempering = to_bool(fe.get("!Outputs", "get_parallel_tempering"));
    r.evidence_diags = to_bool(fe.get("!Diagnostics", "evidence_diags")); r.evidence_interpolation_factor = to_int(fe.need("!Diagnostics", "evidence_interpolation_factor"));
    r.Nslices = fe.Nslices; r.Nvars = nv; r.Ncons = a.size() - nv; r.lambda_temp = fe.lambda;
    { const std::string ff = fe.get("!Outputs", "file_format"); r.file_format = (ff == "text") ? 1 : ((ff == "debug") ? 2 : 0); }    // outputs.cpp:38-56
    fe.ready = true;
}

template <typename T> void dump_vec(std::string &o, const char *key, const std::vector<T> &v, const char *fmt)
{
    o += key; o += ":"; char b[64];
    for (const T &x : v) { snprintf(b, sizeof b, fmt, x); o += " "; o += b; }
    o += "\n";
}

}  // namespace

extern "C" const char *tg_fe_last_error(void) { return g_err.c_str(); }

#define FE_TRY(body) try { body; return TG_OK; } catch (const FeError &e) { g_err = e.msg; return TG_E_BADARG; } catch (const std::exception &e) { g_err = std::string("front end: ") + e.what(); return TG_E_BADARG; }

extern "C" int tg_fe_open(const char *workdir, tg_frontend **out)
{
    if (!workdir || !out) { g_err = "tg_fe_open: null argument"; return TG_E_BADARG; }
    tg_frontend *fe = new tg_frontend;
    try {
        fe->workdir = workdir;
        const std::string dd = fe->workdir + "/Config/default/";
        read_default_cfg(*fe, dd + "config_default.cfg");
        read_errors_cfg(*fe, dd + "errors_default.cfg");
        fe->models = read_listfile(dd + "models_ctrl.list"); fe->priors = read_listfile(dd + "priors_ctrl.list");
        fe->likelihoods = read_listfile(dd + "likelihoods_ctrl.list"); fe->primepriors = read_listfile(dd + "primepriors_ctrl.list");
        read_presets(*fe, fe->workdir + "/Config/config_presets.cfg");
        fe->Nt_learn0 = to_longs(fe->need("!MALA", "Nt_learn")); fe->dN_mixing0 = (long)to_dbl(fe->need("!MALA", "dN_mixing"));
        if (fe->Nt_learn0.size() < 2) die("config_default.cfg: Nt_learn needs at least two values");
    } catch (const FeError &e) { g_err = e.msg; delete fe; return TG_E_BADARG; }
    catch (const std::exception &e) { g_err = std::string("front end: ") + e.what(); delete fe; return TG_E_BADARG; }
    *out = fe;
    return TG_OK;
}
extern "C" void tg_fe_close(tg_frontend *fe) { delete fe; }
extern "C" int tg_fe_counts(const tg_frontend *fe, int *n_stars, int *n_phases, int *first_phase, int *last_phase)
{
    if (!fe) return TG_E_BADARG;
    if (n_stars) *n_stars = (int)fe->table_ids.size();
    if (n_phases) *n_phases = (int)fe->processing.size();
    if (first_phase) *first_phase = fe->first_process;
    if (last_phase) *last_phase = fe->last_process;
    return TG_OK;
}
extern "C" int tg_fe_nslices(tg_frontend *fe, int star, int *nslices)
{
    if (!fe || !nslices) return TG_E_BADARG;
    FE_TRY({
        if (star < 0 || star >= (int)fe->table_ids.size()) die("star index outside table_ids");
        const std::string f = fe->models_dir + fe->table_ids[(size_t)star][0] + ".model";
        std::ifstream in(f.c_str()); if (!in.is_open()) die("Unable to open the file: " + f);
        int n = 0; std::string l; while (std::getline(in, l)) if (first_char(trim(l)) == "*") n++;
        *nslices = n;
    })
}
extern "C" int tg_fe_setup(tg_frontend *fe, int star, int slice, int phase)
{
    if (!fe) return TG_E_BADARG;
    FE_TRY(setup(*fe, star, slice, phase))
}
extern "C" int tg_fe_problem(const tg_frontend *fe, tg_problem *p)
{
    if (!fe || !p || !fe->ready) { g_err = "tg_fe_problem: no successful tg_fe_setup yet"; return TG_E_BADARG; }
    memset(p, 0, sizeof *p);
    p->Nx = (long)fe->x.size(); p->x = fe->x.data(); p->y = fe->y.data(); p->sigma_y = fe->sig.data();
    p->model_case = fe->model_case; p->likelihood_case = fe->like_case; p->prior_case = fe->prior_case; p->likelihood_params = fe->likelihood_params;
    for (int k = 0; k < 11; k++) p->plength[k] = fe->in.plength[(size_t)k];
    p->Np = fe->in.all.size(); p->params0 = fe->in.all.inputs.data(); p->relax = fe->in.all.relax.data();
    p->priors = fe->priors_flat.data(); p->priors_switch = fe->pswitch.data();
    for (int k = 0; k < 4; k++) p->extra_priors[k] = fe->in.extra_priors[k];
    p->Nchains = fe->Nchains; p->Tcoefs = fe->Tcoefs.data();
    p->c0 = fe->c0_phase; p->epsilon1 = fe->eps1; p->epsilon2 = fe->eps2; p->A1 = fe->A1; p->target_acceptance = fe->target;
    p->n_learn = (int)fe->periods_learn.size(); p->Nt_learn = fe->Nt_learn.data(); p->periods_learn = fe->periods_learn.data();
    p->dN_mixing = fe->dN_mixing; p->init_err = fe->init_err.data();
    return TG_OK;
}
extern "C" int tg_fe_run_info(const tg_frontend *fe, tg_fe_run *r)
{
    if (!fe || !r || !fe->ready) { g_err = "tg_fe_run_info: no successful tg_fe_setup yet"; return TG_E_BADARG; }
    *r = fe->run;
    return TG_OK;
}
extern "C" const char *tg_fe_input_name(const tg_frontend *fe, int i) { return (fe && fe->ready && i >= 0 && i < fe->in.all.size()) ? fe->in.all.names[(size_t)i].c_str() : ""; }
extern "C" const char *tg_fe_prior_name(const tg_frontend *fe, int i) { return (fe && fe->ready && i >= 0 && i < fe->in.all.size()) ? fe->in.all.pnames[(size_t)i].c_str() : ""; }

extern "C" long tg_fe_dump(const tg_frontend *fe, char *buf, long cap)
{
    if (!fe || !fe->ready) return -1;
    std::string o; char b[512];
    const Block &a = fe->in.all;
    o += "Nslices: " + std::to_string(fe->Nslices) + "\n";
    o += "model_fct_name: " + fe->model_name + "\nlikelihood_fct_name: " + fe->like_name + "\nprior_fct_name: " + fe->prior_name + "\n";
    snprintf(b, sizeof b, "switches: %d %d %d\nlikelihood_params: %.17g\n", fe->model_case, fe->like_case, fe->prior_case, fe->likelihood_params); o += b;
    dump_vec(o, "inputs", a.inputs, "%.17g"); dump_vec(o, "relax", a.relax, "%d"); dump_vec(o, "plength", fe->in.plength, "%d");
    dump_vec(o, "extra_priors", std::vector<double>(fe->in.extra_priors, fe->in.extra_priors + 4), "%.17g");
    dump_vec(o, "priors_names_switch", fe->pswitch, "%d");
    for (int k = 0; k < 4; k++) { std::vector<double> row; for (int i = 0; i < a.size(); i++) row.push_back(a.priors[(size_t)i][(size_t)k]); snprintf(b, sizeof b, "priors_row%d", k); dump_vec(o, b, row, "%.17g"); }
    o += "inputs_names:"; for (auto &n : a.names) o += " |" + n; o += "\n";
    o += "priors_names:"; for (auto &n : a.pnames) o += " |" + n; o += "\n";
    o += "Nx: " + std::to_string(fe->x.size()) + "\n";
    double sx = 0, sy = 0; for (size_t i = 0; i < fe->x.size(); i++) { sx += fe->x[i]; sy += fe->y[i]; }
    snprintf(b, sizeof b, "data: %.17g %.17g %.17g %.17g %.17g %.17g\n", fe->x.front(), fe->x.back(), fe->y.front(), fe->y.back(), sx, sy); o += b;
    snprintf(b, sizeof b, "mala: Nchains %d dN_mixing %ld lambda_temp %.17g c0 %.17g epsilon1 %.17g epsi2 %.17g A1 %.17g target %.17g\n", fe->Nchains, fe->dN_mixing, fe->lambda, fe->c0_phase, fe->eps1, fe->eps2, fe->A1, fe->target); o += b;
    dump_vec(o, "Nt_learn", fe->Nt_learn, "%ld"); dump_vec(o, "periods_learn", fe->periods_learn, "%ld");
    snprintf(b, sizeof b, "outputs: Nsamples %ld Nbuffer %ld erase_old_files %d restore %d %d %d %d\n", fe->run.Nsamples, fe->run.Nbuffer, fe->run.erase_old_files,
             fe->run.do_restore_proposal, fe->run.do_restore_variables, fe->run.do_restore_last_index, fe->run.do_restore_proposal_mean); o += b;
    o += "output_root_name: " + fe->root + "\ndir_out: " + fe->dir_out + "\nrestore_dir: " + fe->restore_dir + "\nrestore_file_in: " + fe->rin + "\nrestore_file_out: " + fe->rout + "\n";
    dump_vec(o, "Tcoefs", fe->Tcoefs, "%.17g"); dump_vec(o, "sigma0", fe->sigma0, "%.17g");
    std::vector<double> cov; for (double e : fe->init_err) cov.push_back(e * e);
    dump_vec(o, "cov0_diag_chain0", cov, "%.17g"); dump_vec(o, "cov0_diag_chainlast", cov, "%.17g");
    if (buf && cap > 0) { const long n = std::min<long>((long)o.size(), cap - 1); memcpy(buf, o.data(), (size_t)n); buf[n] = 0; }
    return (long)o.size();
}

// ------------------------------------------------------------------ numeric diagnostics
namespace {
// interp1 / parabola / interp2 / quad_interpol, interpol.cpp:58-103 (long double arithmetic like the reference)
long double interp1(long double x, const std::vector<double> &a)
{
    const int n = (int)a.size();
    if (x <= 0) return a[0];
    if (x >= n - 1) return a[(size_t)(n - 1)];
    const int j = (int)x;
    return a[(size_t)j] + (x - j) * (a[(size_t)(j + 1)] - a[(size_t)j]);
}
long double parabola(long double x, long double f_1, long double f0, long double f1)
{
    if (x <= -1) return f_1;
    if (x >= 1) return f1;
    const long double l = f0 - x * (f_1 - f0), r = f0 + x * (f1 - f0);
    return (l + r + x * (r - l)) / 2;
}
long double interp2(long double x, const std::vector<double> &a)
{
    if (x <= .5 || x >= a.size() - 1.5) return interp1(x, a);
    const int j = (int)(x + .5);
    const long double t = 2 * (x - j);
    return parabola(t, (a[(size_t)(j - 1)] + a[(size_t)j]) / 2, a[(size_t)j], (a[(size_t)j] + a[(size_t)(j + 1)]) / 2);
}
}  // namespace

extern "C" int tg_evidence(const double *Tcoefs, int C, const double *logL, long nsamples, int factor, double *L_beta, double *evidence)
{
    if (!Tcoefs || !logL || !evidence || C < 1 || nsamples < 1 || factor < 1) { g_err = "tg_evidence: bad argument"; return TG_E_BADARG; }
    std::vector<double> Lb((size_t)C);
    for (int m = 0; m < C; m++) { double s = 0; for (long i = 0; i < nsamples; i++) s += logL[(size_t)i * C + m]; Lb[(size_t)m] = s / (double)nsamples; }
    if (L_beta) for (int m = 0; m < C; m++) L_beta[m] = Lb[(size_t)m];
    const int npts = factor * C;
    const long double step = npts > 1 ? (long double)((double)(C - 1) / (npts - 1)) : 0;
    double sum = 0;
    for (int j = 0; j < npts; j++) sum += (double)interp2(j * step, Lb);
    *evidence = sum / npts;
    return TG_OK;
}
extern "C" int tg_acceptance(const unsigned char *moved, long nsteps, int C, long Nbuffer, long Ncopy, long Nsamples_init, double *rate, double *xaxis)
{
    if (!moved || !rate || nsteps < 1 || C < 1 || Nbuffer < 1) { g_err = "tg_acceptance: bad argument"; return TG_E_BADARG; }
    for (int m = 0; m < C; m++) { long n = 0; for (long i = 0; i < nsteps; i++) n += moved[(size_t)i * C + m] ? 1 : 0; rate[m] = (double)n / (double)Nbuffer; }
    if (xaxis) *xaxis = ((double)Ncopy + 0.5) * (double)Nbuffer + (double)Nsamples_init;
    return TG_OK;
}
