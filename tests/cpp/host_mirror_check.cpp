// C++ host-side check of include/tamcmc_gpu.hpp: model_Test_Gaussian (models.cpp:2021-2034) with the
// chi_square likelihood, built by hand the way Config::setup leaves it, evaluated through
// Model_def::generate_model and sampled with MALA::execute.  Prints every number with 17 digits; the
// Python test (tests/test_gpu_cpp_mirror.py) builds the same problem through gpu.py and compares.
#include <cmath>
#include <cstdio>
#include <vector>
#include "tamcmc_gpu.hpp"

int main()
{
    const int Nx = 400, Np = 4, C = 3;
    std::vector<double> x(Nx), y(Nx), sig(Nx, 0.25);
    const double p0[Np] = {8.0, 30.0, 240.0, 1.5};                       // Hmax, Gsigma, Numax, N0
    for (int i = 0; i < Nx; i++) {
        x[i] = 50.0 + 400.0 * i / (Nx - 1);
        const double d = (x[i] - 240.0) / 30.0;                          // only + - * /: the same bits as the numpy twin
        y[i] = 1.5 + 8.0 / (1.0 + d * d) + 0.25 * (double)((i * 37) % 17 - 8) / 8.0;
    }
    const int relax[Np] = {1, 1, 1, 1}, prior_switch[Np] = {1, 1, 1, 1};  // uniform priors (primepriors_ctrl.list)
    const double priors[4 * Np] = {0.0, 50.0, -9999, -9999, 1.0, 200.0, -9999, -9999, 100.0, 400.0, -9999, -9999, 0.0, 10.0, -9999, -9999};
    const double T[C] = {1.0, 1.7, 2.89};
    const double init_err[Np] = {0.8, 1.3, 4.4, 0.0245};
    const long Nt_learn[2] = {10, 100000}, periods[1] = {1};
    tg_problem P = tg_problem();
    P.Nx = Nx; P.x = x.data(); P.y = y.data(); P.sigma_y = sig.data();
    P.model_case = 0; P.likelihood_case = 1; P.prior_case = 0; P.likelihood_params = 1.0;
    P.plength[0] = 4; P.Np = Np; P.params0 = p0; P.relax = relax; P.priors = priors; P.priors_switch = prior_switch;
    P.Nchains = C; P.chains_per_ensemble = 0; P.Tcoefs = T;
    P.c0 = 5.0; P.epsilon1 = 1e-12; P.epsilon2 = 1e-12; P.A1 = 1e14; P.target_acceptance = 0.234;
    P.n_learn = 1; P.Nt_learn = Nt_learn; P.periods_learn = periods; P.dN_mixing = 2;
    P.init_err = init_err; P.seed = 77; P.device = 0; P.flags = TG_F_FIX_SWAP_BOOKKEEPING;
    try {
        tamcmc_gpu::Model_def model(P);
        std::printf("sizes %d %d %d %d %ld\n", model.Nvars, model.Nparams, model.Nmodels, model.Nensembles, model.Nx);
        double params[C * Np], logL[C], logPr[C], logPo[C];
        int st[C];
        for (int c = 0; c < C; c++) for (int i = 0; i < Np; i++) params[c * Np + i] = p0[i] * (1.0 + 0.01 * (c + 1) * (i + 1));
        model.generate_model(params, C, T, logL, logPr, logPo, st);
        for (int c = 0; c < C; c++) std::printf("eval %d %.17g %.17g %.17g %d\n", c, logL[c], logPr[c], logPo[c], st[c]);
        const std::vector<double> m = model.call_model(params, 1, T);
        std::printf("model %.17g %.17g %.17g\n", m[0], m[Nx / 2], m[Nx - 1]);
        tamcmc_gpu::MALA mala(model);
        const long N = 60;
        tamcmc_gpu::Records rec(N, model.Nmodels, model.Nensembles, model.Nvars);
        mala.execute(0, N, &rec);
        for (long i = N - 3; i < N; i++)
            for (int c = 0; c < C; c++)
                std::printf("step %ld %d %.17g %.17g %d %d\n", i, c, rec.vars[((size_t)i * C + c) * model.Nvars], rec.logL[(size_t)i * C + c],
                            (int)rec.moved[(size_t)i * C + c], rec.swap_done[i]);
        std::vector<double> vars(C * Np), mu(C * Np), sigma(C), cov(C * Np * Np);
        model.get_state(vars.data(), mu.data(), sigma.data(), cov.data());
        std::printf("state %.17g %.17g %.17g %.17g\n", vars[0], mu[1], sigma[2], cov[5]);
    } catch (const tamcmc_gpu::Error &e) {
        std::printf("error %d %s\n", e.code, e.what());
        return e.code == TG_E_NODEVICE ? 3 : 1;
    }
    return 0;
}
