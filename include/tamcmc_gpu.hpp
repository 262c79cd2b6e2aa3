/* tamcmc_gpu.hpp -- C++ host-side mirror of the reference's operator interface for the hot path, over the
 * C-ABI of tamcmc_gpu.h (header only, C++11, no dependency beyond the C library).
 *
 * The reference drives the path through two classes (all paths relative to /root/reference/tamcmc):
 *   Model_def  headers/model_def.h:82, sources/model_def.cpp   -- generate_model / call_model per chain
 *   MALA       headers/MALA.h,         sources/MALA.cpp:540-739 -- execute(): the sampling loop
 * The classes below keep their member names and argument meaning; what was one chain per call becomes a
 * batch of chains per call, and what was exit(EXIT_FAILURE) becomes an exception (tamcmc_gpu::Error) or a
 * per-chain TG_ST_* status.  INTEGRATION.md shows where a maintainer of the reference plugs them in.
 */
#ifndef TAMCMC_GPU_HPP
#define TAMCMC_GPU_HPP

#include <stdexcept>
#include <string>
#include <vector>
#include "tamcmc_gpu.h"

namespace tamcmc_gpu {

struct Error : std::runtime_error {
    int code;
    Error(int c, const std::string &what) : std::runtime_error(what), code(c) {}
};

inline void check(int rc)
{
    if (rc != TG_OK) throw Error(rc, std::string("libtamcmc_gpu error ") + std::to_string(rc) + ": " + tg_last_error());
}

/* Model_def: owns the device copy of the data, the chain tables and the sampler state of one tempered
 * ensemble set (the reference builds two Model_def objects -- current and proposed -- plus a MALA; here one
 * handle holds both, model_def.cpp:27-181 + MALA.cpp:53-112). */
class Model_def {
public:
    int  Nmodels = 0;   /* Nchains   (model_def.h: Nmodels)  */
    int  Nparams = 0;   /* Np                                   */
    int  Nvars = 0;     /* free parameters                      */
    int  Nensembles = 0;
    long Nx = 0;

    explicit Model_def(const tg_problem &problem)
    {
        check(tg_create(&problem, &h_));
        check(tg_sizes(h_, &Nvars, &Nparams, &Nmodels, &Nensembles, &Nx));
    }
    ~Model_def() { if (h_) tg_destroy(h_); }
    Model_def(const Model_def &) = delete;
    Model_def &operator=(const Model_def &) = delete;

    /* Model_def::generate_model (model_def.cpp:358-367) for B parameter vectors params[B][Nparams] with the
     * tempering coefficients Tcoefs[B]: logLikelihood (tempered), logPrior, logPosterior, status bits. */
    void generate_model(const double *params, int B, const double *Tcoefs,
                        double *logLikelihood, double *logPrior, double *logPosterior, int *status)
    {
        check(tg_eval(h_, params, B, Tcoefs, logLikelihood, logPrior, logPosterior, status, nullptr));
    }

    /* Model_def::call_model (model_def.cpp:210-289): the model spectra model[B][Nx] of B parameter vectors. */
    std::vector<double> call_model(const double *params, int B, const double *Tcoefs)
    {
        std::vector<double> model((size_t)B * (size_t)Nx), l(B), p(B), q(B);
        std::vector<int> st(B);
        check(tg_eval(h_, params, B, Tcoefs, l.data(), p.data(), q.data(), st.data(), model.data()));
        return model;
    }

    /* restore triplet (outputs.cpp:863-1025 / config.cpp:1321-1579 / MALA::restore_proposal MALA.cpp:191-244) */
    void get_state(double *vars, double *mu, double *sigma, double *covarmat,
                   double *logLikelihood = nullptr, double *logPrior = nullptr, double *logPosterior = nullptr)
    {
        check(tg_get_state(h_, vars, mu, sigma, covarmat, logLikelihood, logPrior, logPosterior));
    }
    // buf_restore.*_mean of the reference (outputs.cpp:1758-1818); the problem must carry TG_F_RESTORE_MEANS
    long get_state_means(double *vars_mean, double *mu_mean, double *sigma_mean, double *covarmat_mean, bool reset = true)
    {
        long n = 0;
        check(tg_get_state_means(h_, vars_mean, mu_mean, sigma_mean, covarmat_mean, &n, reset ? 1 : 0));
        return n;
    }
    void set_state(const double *vars, const double *mu, const double *sigma, const double *covarmat)
    {
        check(tg_set_state(h_, vars, mu, sigma, covarmat));
    }

    tg_handle *handle() { return h_; }

private:
    tg_handle *h_ = nullptr;
};

/* Per-step records of a run, owning their storage (the arguments of Outputs::update_buffer_*, MALA.cpp:696-707). */
struct Records {
    std::vector<double> vars, logL, logPrior, logPost, Pmove, swap_Pswap;
    std::vector<unsigned char> moved;
    std::vector<int> swap_attempt, swap_chain0, swap_done, status;
    tg_records view;
    Records(long nsteps, int C, int E, int Nv)
        : vars((size_t)nsteps * C * Nv), logL((size_t)nsteps * C), logPrior((size_t)nsteps * C), logPost((size_t)nsteps * C),
          Pmove((size_t)nsteps * C), swap_Pswap((size_t)nsteps * E), moved((size_t)nsteps * C), swap_attempt((size_t)nsteps * E),
          swap_chain0((size_t)nsteps * E), swap_done((size_t)nsteps * E), status((size_t)nsteps * C)
    {
        view.vars = vars.data(); view.logL = logL.data(); view.logPrior = logPrior.data(); view.logPost = logPost.data();
        view.Pmove = Pmove.data(); view.moved = moved.data(); view.swap_attempt = swap_attempt.data();
        view.swap_chain0 = swap_chain0.data(); view.swap_Pswap = swap_Pswap.data(); view.swap_done = swap_done.data();
        view.status = status.data();
    }
};

/* MALA: the sampling loop (MALA::execute, MALA.cpp:608-692: update_position_MH, update_proposal,
 * parallel_tempering), device resident. */
class MALA {
public:
    explicit MALA(Model_def &model) : m_(model) {}
    /* iterations i0 .. i0+nsteps-1; rec may be null (no records kept) */
    void execute(long i0, long nsteps, Records *rec = nullptr)
    {
        check(tg_run(m_.handle(), i0, nsteps, rec ? &rec->view : nullptr));
    }

private:
    Model_def &m_;
};

}  // namespace tamcmc_gpu
#endif
