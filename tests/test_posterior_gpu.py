"""north_star, second correctness criterion: "posterior means and variances from full runs must agree with the reference
within Monte-Carlo error" -- on BASELINE.json's configs[0].

tests/golden/posterior_cfg1.npz (oracle/make_posterior_golden.py) holds six runs of the UNMODIFIED reference program
(`cpptamcmc execute 1 1 1 1 2`: shipped star TF_3443483_local-v3, slice 1, model_MS_local_basic, 10 tempered chains,
Burn-in 3000 -> Learning 6000 -> Acquire 20000 samples) with the problem its front end built.  The same three phases
(same presets: config_presets.cpp:68-93) run here on the GPU with four seeds; the run-averaged posterior mean of every
variable of the coldest chain must agree within 4 standard errors of the two sets of runs, the posterior widths within
the run-to-run scatter.  (The slice is noise dominated: the reference's own runs scatter by 15 % on the mode heights.)"""
import os

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu


def _phases(tb, prob, z, seed):
    ns, c0 = [int(v) for v in z["nsamples"]], z["c0"]
    n0 = [int(v) for v in z["Nt_learn0"]]
    Cn = z["Tcoefs"].size
    mk = lambda k, Nt, mix: tb.MalaConfig(Nchains=Cn, Tcoefs=z["Tcoefs"], c0=float(c0[k]), Nt_learn=Nt, periods_learn=[1, 1], dN_mixing=mix, seed=seed + k)
    # Burn-in: learn from Nt_learn[0] on, never stop, mix every step
    m = tb.GpuModelDef(prob, mk(0, n0 + [ns[0] + 1], 1))
    tb.GpuMALA(m).execute(0, ns[0], record=False)
    m.refresh_from_device()
    v = m.vars.copy()
    m.close()
    # Learning (restore = 1: variables only; fresh proposal), the chains are never mixed
    m = tb.GpuModelDef(prob, mk(1, n0 + [ns[1] + 1], ns[1] + 1))
    m.set_state(vars_=v)
    tb.GpuMALA(m).execute(0, ns[1], record=False)
    m.refresh_from_device()
    v = m.vars.copy()
    mu, sg, cov = m.get_proposal()
    m.close()
    # Acquire (restore = 2: variables and proposal), no learning
    m = tb.GpuModelDef(prob, mk(2, [ns[2] + 1, ns[2] + 2, ns[2] + 3], 1))
    m.set_state(vars_=v, mu=mu, sigma=sg, covarmat=cov)
    out = tb.GpuMALA(m).execute(0, ns[2])
    m.close()
    return out["vars"][:, 0, :], out["moved"][:, 0].mean()


def test_posterior_of_the_shipped_star_agrees_with_the_reference_program(tb, W):
    z = np.load(os.path.join(ROOT, "tests", "golden", "posterior_cfg1.npz"), allow_pickle=False)
    relax = z["relax"].astype(np.int32)
    prob = W.Problem(name="cfg1", model_case=int(z["model_case"]), like_case=int(z["like_case"]), prior_case=int(z["prior_case"]), plength=z["plength"].astype(np.int32),
                     params=z["params"], relax=relax, priors=z["priors"], priors_switch=z["priors_switch"].astype(np.int32), extra_priors=z["extra_priors"],
                     x=z["x"], y=z["y"], likelihood_params=float(z["likelihood_params"]), names=[str(n) for n in z["names"]])
    err = np.ones(relax.size)
    err[relax == 1] = z["init_err"]
    prob.err_frac, prob.err_off = np.zeros(relax.size), err
    runs = [_phases(tb, prob, z, 1000 * (r + 1)) for r in range(4)]
    g_mean = np.array([v.mean(0) for v, _ in runs])
    g_std = np.array([v.std(0) for v, _ in runs])
    r_mean, r_std = z["ref_mean"], z["ref_std"]
    names = [n for n, f in zip(prob.names, relax) if f == 1]
    # posterior means: difference of the run averages within 4 standard errors (plus 0.1 % for variables without scatter)
    se = np.sqrt(r_mean.var(0, ddof=1) / r_mean.shape[0] + g_mean.var(0, ddof=1) / g_mean.shape[0])
    d = np.abs(g_mean.mean(0) - r_mean.mean(0))
    assert np.all(d <= 4 * se + 1e-3 * np.abs(r_mean.mean(0))), "\n".join("%-26s gpu %.5g ref %.5g  se %.3g" % t for t in zip(names, g_mean.mean(0), r_mean.mean(0), se))
    # posterior widths: run-averaged standard deviations within a factor 1.6 (6 and 4 short runs of a noise-dominated slice)
    ratio = g_std.mean(0) / r_std.mean(0)
    assert np.all((ratio > 1 / 1.6) & (ratio < 1.6)), "\n".join("%-26s width gpu %.4g ref %.4g" % t for t in zip(names, g_std.mean(0), r_std.mean(0)))
    # the well-determined quantities on their own: the l=2 frequency and the white noise level
    k2, kn = names.index("Frequency_l") + 1, names.index("White_Noise_N0")
    assert abs(g_mean.mean(0)[k2] - r_mean.mean(0)[k2]) < 0.1 and abs(g_mean.mean(0)[kn] / r_mean.mean(0)[kn] - 1) < 0.05
    # acceptance of the coldest chain in the range the reference's runs span (0.02 .. 0.34)
    acc = np.mean([a for _, a in runs])
    assert 0.01 < acc < 0.6


def test_posterior_of_a_global_fit_agrees_with_the_reference_program(tb, W):
    """The same criterion on a GLOBAL main-sequence fit: tests/golden/posterior_global.npz
    (oracle/make_posterior_golden_global.py) holds four runs of the unmodified reference program -- `cpptamcmc execute 1 1 1`,
    model_MS_Global_a1etaa3_HarveyLike (models.cpp:480-637) with the io_MS_Global priors (priors_calc.cpp:24-172), 6 orders x
    l = 0, 1, 2, 38 free parameters, 44 000 bins, 10 tempered chains, Burn-in 2000 -> Learning 6000 -> Acquire 12000
    (MALA::execute, MALA.cpp:540-739; 15 minutes of 5 cores per run) -- with the problem its front end built.  The same
    phases run here with four seeds.

    What 12 000 correlated samples of a 38-dimensional posterior can and cannot show (scripts/dev/posterior_global_probe.py
    runs the GPU sampler for 200 000 samples): the frequencies, visibilities and the splitting/inclination pair are converged
    -- widths equal to the long run's, compared here within a factor 1.35; heights, widths and above all the three Harvey
    parameters + N0 are not (long-run widths 0.05 / 0.36 / 0.04 for tc / p / N0 against 0.026 / 0.20 / 0.021 in the reference's
    12 000-sample runs and 0.012 / 0.09 / 0.015 in ours: both samplers have only diffused part of the way from the common
    starting point, the reference about twice as far), so for those only a factor 3 is asserted.  The means are compared through
    their standard errors, which come from 4 + 4 runs (about 5 degrees of freedom each): root-mean-square z over the 38
    parameters below 2 (1.3 expected, 1.4 measured), no |z| above 6 (4.1 measured)."""
    z = np.load(os.path.join(ROOT, "tests", "golden", "posterior_global.npz"), allow_pickle=False)
    relax = z["relax"].astype(np.int32)
    prob = W.Problem(name="global", model_case=int(z["model_case"]), like_case=int(z["like_case"]), prior_case=int(z["prior_case"]), plength=z["plength"].astype(np.int32),
                     params=z["params"], relax=relax, priors=z["priors"], priors_switch=z["priors_switch"].astype(np.int32), extra_priors=z["extra_priors"],
                     x=z["x"], y=z["y"], likelihood_params=float(z["likelihood_params"]), names=[str(n) for n in z["names"]])
    err = np.ones(relax.size)
    err[relax == 1] = z["init_err"]
    prob.err_frac, prob.err_off = np.zeros(relax.size), err
    runs = [_phases(tb, prob, z, 7000 * (r + 1)) for r in range(4)]
    g_mean = np.array([v.mean(0) for v, _ in runs])
    g_std = np.array([v.std(0) for v, _ in runs])
    r_mean, r_std = z["ref_mean"], z["ref_std"]
    names = [n for n, f in zip(prob.names, relax) if f == 1]
    se = np.sqrt(r_mean.var(0, ddof=1) / r_mean.shape[0] + g_mean.var(0, ddof=1) / g_mean.shape[0])
    zs = (g_mean.mean(0) - r_mean.mean(0)) / se
    ratio = g_std.mean(0) / r_std.mean(0)
    report = "\n".join("%-26s gpu %.6g ref %.6g  z %+.2f  width gpu %.4g ref %.4g" % t for t in zip(names, g_mean.mean(0), r_mean.mean(0), zs, g_std.mean(0), r_std.mean(0)))
    assert np.sqrt(np.mean(zs**2)) < 2.0 and np.max(np.abs(zs)) < 6.0, report
    fast = np.array([n.startswith(("Frequency", "Visibility", "sqrt(")) for n in names])
    assert np.all((ratio[fast] > 1 / 1.35) & (ratio[fast] < 1.35)), report
    assert np.all((ratio[~fast] > 1 / 3.0) & (ratio[~fast] < 3.0)), report
    # every mode frequency within a tenth of a microhertz, the mode widths and the noise level within 8 per cent
    for k, n in enumerate(names):
        if n == "Frequency_l":
            assert abs(g_mean.mean(0)[k] - r_mean.mean(0)[k]) < 0.1, report
        if n in ("Width_l", "White_Noise_N0"):
            assert abs(g_mean.mean(0)[k] / r_mean.mean(0)[k] - 1) < 0.08, report
    acc = np.mean([a for _, a in runs])
    ref_acc = float(z["ref_acceptance"][:, 0].mean())
    assert abs(acc - ref_acc) < 0.05, (acc, ref_acc)
