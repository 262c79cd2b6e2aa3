// extern "C" entry points over the UNMODIFIED reference hot-path functions
// (/root/reference/tamcmc/sources: models.cpp, build_lorentzian.cpp, noise_models.cpp,
// function_rot.cpp, interpol.cpp, likelihoods.cpp, priors_calc.cpp, stats_dictionary.cpp,
// derivatives_handler.cpp), compiled against oracle/eigen_shim.  This file only marshals plain
// arrays into the reference's by-value VectorXd signatures and mirrors the dispatch switches of
// Model_def (model_def.cpp:210-356), which itself cannot be compiled (it needs config.h ->
// Boost/gnuplot).  Test infrastructure only: used to pin the oracle, to generate
// tests/golden/*, and as the "reference" CPU baseline of bench.py.
#include <Eigen/Dense>
#include <cmath>
#include "models.h"
#include "noise_models.h"
#include "likelihoods.h"
#include "priors_calc.h"
#include "interpol.h"
#include "derivatives_handler.h"

using Eigen::VectorXd;
using Eigen::VectorXi;
using Eigen::MatrixXd;

static VectorXd mkvec(const double *p, long n) { VectorXd v(n); for (long i = 0; i < n; i++) v[i] = p[i]; return v; }
static VectorXi mkveci(const int *p, long n) { VectorXi v(n); for (long i = 0; i < n; i++) v[i] = p[i]; return v; }

// Model_def::call_model, model_def.cpp:210-289
static VectorXd ref_call_model(int mc, const VectorXd &params, const VectorXi &plength, const VectorXd &x, int *status)
{
    *status = 0;
    switch (mc) {
    case 0: return model_Test_Gaussian(params, plength, x);
    case 1: return model_Harvey_Gaussian(params, plength, x);
    case 2: return model_MS_Global_a1etaa3_HarveyLike(params, plength, x);
    case 3: return model_MS_Global_a1etaa3_HarveyLike_Classic(params, plength, x);
    case 6: return model_MS_Global_a1l_etaa3_HarveyLike(params, plength, x);
    case 7: return model_MS_Global_a1n_etaa3_HarveyLike(params, plength, x);
    case 8: return model_MS_Global_a1nl_etaa3_HarveyLike(params, plength, x);
    case 9: return model_MS_Global_a1etaa3_AppWidth_HarveyLike_v1(params, plength, x);
    case 10: return model_MS_Global_a1etaa3_AppWidth_HarveyLike_v2(params, plength, x);
    case 11: return model_MS_local_basic(params, plength, x);
    case 12: return model_MS_Global_a1etaa3_HarveyLike_Classic_v2(params, plength, x);
    case 13: return model_MS_Global_a1etaa3_HarveyLike_Classic_v3(params, plength, x);
    case 14: return model_MS_local_Hnlm(params, plength, x);
    default: *status = 3; return VectorXd(x.size());  // cases 4/5 exit() in the reference
    }
}

extern "C" {

int ref_model(int mc, const double *params, int Np, const int *plength, const double *x, long Nx, double *out)
{
    int st;
    VectorXd m = ref_call_model(mc, mkvec(params, Np), mkveci(plength, 11), mkvec(x, Nx), &st);
    if (st) return st;
    for (long i = 0; i < Nx; i++) out[i] = m[i];
    return 0;
}

// Model_def::call_likelihood, model_def.cpp:291-320 (returns logL/T as long double -> double pair hi/lo not needed)
long double ref_likelihood(int lc, const double *y, const double *sigma_y, const double *model, long Nx,
                           double likelihood_params, double T)
{
    long double logL;
    if (lc == 0) {
        double p = likelihood_params;
        logL = likelihood_chi22p(mkvec(y, Nx), mkvec(model, Nx), p);
    } else {
        logL = likelihood_chi_square(mkvec(y, Nx), mkvec(model, Nx), mkvec(sigma_y, Nx));
    }
    return logL / T;
}

// Model_def::call_prior, model_def.cpp:322-356.  pri is 4 x Np column-major.
long double ref_prior(int pc, const double *params, int Np, const int *plength, const double *pri, const int *sw,
                      const double *extra)
{
    MatrixXd P(4, Np);
    for (int i = 0; i < Np; i++) for (int r = 0; r < 4; r++) P(r, i) = pri[4 * i + r];
    VectorXd p = mkvec(params, Np);
    VectorXi pl = mkveci(plength, 11), s = mkveci(sw, Np);
    VectorXd e = mkvec(extra, 4);
    switch (pc) {
    case 0: return priors_Test_Gaussian(p, pl, P, s);
    case 1: return priors_Harvey_Gaussian(p, pl, P, s);
    case 2: return priors_MS_Global(p, pl, P, s, e);
    case 3: return priors_local(p, pl, P, s, e);
    }
    return NAN;
}

// Model_def::generate_model (model_def.cpp:358-367) over C chains under the reference's only
// parallel region, "#pragma omp parallel for" over chains (MALA.cpp:632-633).
void ref_generate_models(int mc, int lc, int pc, const double *params, int C, int Np, const int *plength,
                         const double *x, const double *y, const double *sigma_y, long Nx,
                         const double *pri, const int *sw, const double *extra, double likelihood_params,
                         const double *Tcoefs, double *logL, double *logPrior, double *logPost)
{
    VectorXd X = mkvec(x, Nx), Y = mkvec(y, Nx), S = sigma_y ? mkvec(sigma_y, Nx) : VectorXd(0);
    VectorXi pl = mkveci(plength, 11), s = mkveci(sw, Np);
    VectorXd e = mkvec(extra, 4);
    MatrixXd P(4, Np);
    for (int i = 0; i < Np; i++) for (int r = 0; r < 4; r++) P(r, i) = pri[4 * i + r];
#pragma omp parallel for default(shared)
    for (int m = 0; m < C; m++) {
        int st;
        VectorXd p = mkvec(params + (size_t)m * Np, Np);
        VectorXd model = ref_call_model(mc, p, pl, X, &st);
        long double L;
        if (lc == 0) { double pp = likelihood_params; L = likelihood_chi22p(Y, model, pp); }
        else L = likelihood_chi_square(Y, model, S);
        logL[m] = L / Tcoefs[m];
        long double pr;
        switch (pc) {
        case 0: pr = priors_Test_Gaussian(p, pl, P, s); break;
        case 1: pr = priors_Harvey_Gaussian(p, pl, P, s); break;
        case 2: pr = priors_MS_Global(p, pl, P, s, e); break;
        default: pr = priors_local(p, pl, P, s, e); break;
        }
        logPrior[m] = pr;
        logPost[m] = logL[m] + logPrior[m];
    }
}

void ref_amplitude_ratio(int l, double beta, double *out)
{
    VectorXd v = amplitude_ratio(l, beta);
    for (int i = 0; i < 2 * l + 1; i++) out[i] = v[i];
}

double ref_lin_interpol(const double *x, const double *y, long n, double xi) { return lin_interpol(mkvec(x, n), mkvec(y, n), xi); }

void ref_frstder(const double *y, int n, double *out) { Deriv_out d = Frstder_adaptive_reggrid(mkvec(y, n)); for (int i = 0; i < n; i++) out[i] = d.deriv[i]; }
void ref_scndder(const double *y, int n, double *out) { Deriv_out d = Scndder_adaptive_reggrid(mkvec(y, n)); for (int i = 0; i < n; i++) out[i] = d.deriv[i]; }

// primitive priors (stats_dictionary.cpp), by index = primepriors_ctrl.list switch value
long double ref_logP(int kind, long double a, long double b, long double c, long double d, long double x)
{
    switch (kind) {
    case 1: return logP_uniform(a, b, x);
    case 2: return logP_gaussian(a, b, x);
    case 4: return logP_jeffrey(a, b, x);
    case 5: return logP_uniform_gaussian(a, b, c, x);
    case 6: return logP_gaussian_uniform(a, b, c, x);
    case 7: return logP_gaussian_uniform_gaussian(a, b, c, d, x);
    case 8: return logP_uniform_abs(a, b, x);
    case 9: return logP_uniform_cos(a, b, x);
    case 10: return logP_jeffrey_abs(a, b, x);
    }
    return NAN;
}
}
