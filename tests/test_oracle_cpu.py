"""CPU tests (-m "not gpu"): the oracle against the reference's own known answers, analytic
identities, the committed golden vectors (generated from the unmodified reference sources) and,
when oracle/_ref is present, against the reference build itself."""
import math

import numpy as np
import pytest

from conftest import relerr

NINF = float("-inf")


# ---- known-answer values printed in the reference's self test, stats_dictionary.cpp:252-326 ----
def test_prior_known_answers(oracle):
    L = oracle.lib
    assert float(L.to_logP_gaussian(1.0, 0.5, 4.0)) == pytest.approx(-18.225791, abs=5e-7)
    assert float(L.to_logP_uniform(0.0, 1.0, -2.0)) == NINF
    assert float(L.to_logP_uniform(0.0, 1.0, 0.3)) == 0.0
    assert float(L.to_logP_jeffrey(0.1, 1.0, -2.0)) == NINF
    assert float(L.to_logP_jeffrey(0.1, 1.0, 0.5)) == pytest.approx(-0.363766, abs=5e-7)
    assert float(L.to_logP_uniform_gaussian(1.0, 2.0, 0.1, -1.0)) == NINF
    assert float(L.to_logP_uniform_gaussian(1.0, 2.0, 0.1, 1.5)) == pytest.approx(-0.11807759, abs=5e-8)
    assert float(L.to_logP_uniform_gaussian(1.0, 2.0, 0.1, 3.0)) == pytest.approx(-50.118074, abs=5e-6)
    assert float(L.to_logP_gaussian_uniform(1.0, 2.0, 0.1, -1.0)) == pytest.approx(-200.11806, abs=5e-5)
    assert float(L.to_logP_gaussian_uniform(1.0, 2.0, 0.1, 1.5)) == pytest.approx(-0.11807759, abs=5e-8)
    assert float(L.to_logP_gaussian_uniform(1.0, 2.0, 0.1, 3.0)) == NINF
    assert float(L.to_logP_gaussian_uniform_gaussian(1.0, 2.0, 0.1, 0.4, -1.0)) == pytest.approx(-200.48651, abs=5e-5)
    assert float(L.to_logP_gaussian_uniform_gaussian(1.0, 2.0, 0.1, 0.4, 1.5)) == pytest.approx(-0.48652704, abs=5e-8)
    assert float(L.to_logP_gaussian_uniform_gaussian(1.0, 2.0, 0.1, 0.4, 3.0)) == pytest.approx(-3.6115268, abs=5e-7)


# ---- amplitude_ratio: closed forms of SURVEY 8(a) row a8 and sum rule ----
@pytest.mark.parametrize("beta", [0.0, 12.5, 37.0, 55.0, 90.0])
def test_amplitude_ratio_closed_forms(oracle, beta):
    b = math.radians(beta)
    s, c = math.sin(b), math.cos(b)
    v1 = oracle.amplitude_ratio(1, beta)
    v2 = oracle.amplitude_ratio(2, beta)
    v3 = oracle.amplitude_ratio(3, beta)
    np.testing.assert_allclose(v1, [0.5 * s**2, c**2, 0.5 * s**2], atol=1e-14)
    np.testing.assert_allclose(v2, [3 / 8 * s**4, 1.5 * s**2 * c**2, 0.25 * (3 * c**2 - 1) ** 2, 1.5 * s**2 * c**2, 3 / 8 * s**4], atol=1e-14)
    np.testing.assert_allclose(v3, [5 / 16 * s**6, 15 / 8 * c**2 * s**4, 3 / 16 * s**2 * (5 * c**2 - 1) ** 2,
                                    0.25 * (5 * c**3 - 3 * c) ** 2, 3 / 16 * s**2 * (5 * c**2 - 1) ** 2,
                                    15 / 8 * c**2 * s**4, 5 / 16 * s**6], atol=1e-14)
    for v in (v1, v2, v3):
        assert abs(v.sum() - 1.0) < 1e-14


# ---- derivatives: the x^2+10 demo of derivatives_handler.cpp:431-514 (d/dx=2x, d2=2 on a unit grid) ----
def test_derivatives_parabola(oracle):
    x = np.arange(12, dtype=np.float64)
    y = x**2 + 10
    d1, d2 = oracle.frstder(y), oracle.scndder(y)
    np.testing.assert_allclose(d1[1:-1], 2 * x[1:-1], atol=1e-12)
    assert d1[0] == y[1] - y[0] and d1[-1] == y[-1] - y[-2]
    np.testing.assert_allclose(d2, 2.0, atol=1e-12)


def test_lin_interpol(oracle):
    x = np.array([1.0, 2.0, 4.0, 8.0]); y = np.array([10.0, 20.0, 0.0, 4.0])
    assert oracle.lin_interpol(x, y, 1.5) == pytest.approx(15.0)
    assert oracle.lin_interpol(x, y, 3.0) == pytest.approx(10.0)
    assert oracle.lin_interpol(x, y, 8.0) == pytest.approx(4.0)
    assert oracle.lin_interpol(x, y, 0.0) == pytest.approx(0.0)       # extrapolated from the first pair
    assert oracle.lin_interpol(x, y, 10.0) == pytest.approx(6.0)      # extrapolated from the last pair


# ---- truncation window quirks Q1/Q2 ----
def test_window_branches_and_clamps(oracle):
    x = 1000.0 + 0.01 * np.arange(10000)
    step = x[1] - x[0]
    st, lo, hi = oracle.window(x, 1050.0, 1.4, 2.0, 1, step, 20.0)        # Gamma>=1, fs>=1: c*(l*fs+Gamma)=68
    assert st == 0 and lo == 0 and hi == 10000
    st, lo, hi = oracle.window(x, 1050.0, 0.4, 0.5, 2, step, 5.0)         # Gamma<=1, fs<=1: c*(l+1)=15
    assert st == 0 and (lo, hi) == (int(math.floor((1035.0 - x[0]) / step)), int(math.ceil((1065.0 - x[0]) / step)))
    st, lo, hi = oracle.window(x, 1050.0, 1.0, 1.0, 0, step, 2.0)         # every block fires, the last wins: 2.2*c
    assert st == 0 and (lo, hi) == (int(math.floor((1050 - 4.4 - x[0]) / step)), int(math.ceil((1050 + 4.4 - x[0]) / step)))
    st, lo, hi = oracle.window(x, 900.0, 0.4, 0.5, 0, step, 2.0)          # wholly below the grid: pmax = x0 + c
    assert st == 0 and lo == 0 and hi == int(math.ceil(2.0 / step))
    st, lo, hi = oracle.window(x, 1200.0, 0.4, 0.5, 0, step, 2.0)         # wholly above: pmin = x_last - c
    assert st == 0 and hi == 10000 and lo == int(math.floor((x[-1] - 2.0 - x[0]) / step))
    st, _, _ = oracle.window(x, 1050.0, float("nan"), 0.5, 1, step, 20.0)  # NaN splitting: no block fires
    assert st == 2
    st, _, _ = oracle.window(x, 1050.0, 0.4, 0.5, 1, step, -1.0)          # negative c: empty window (reference exit()s)
    assert st == 1


# ---- golden vectors produced by the unmodified reference sources ----
def test_golden_covers_every_live_model_case(golden_problems):
    cases = {g.prob.model_case for g in golden_problems.values()}
    assert cases == {0, 1, 2, 3, 6, 7, 8, 9, 10, 11, 12, 13, 14}
    assert any(k.endswith("cfg1-TF_3443483-slice1") for k in golden_problems)


def test_oracle_matches_golden(oracle, golden_problems):
    for name, g in golden_problems.items():
        p = g.prob
        st, m, _ = oracle.model(p.model_case, p.params, p.plength, p.x)
        assert st == 0, name
        assert np.array_equal(m, g.ref_model), (name, relerr(m, g.ref_model))     # bit-exact restatement
        logL = oracle.likelihood(p.like_case, p.y, p.sigma_y, m, p.likelihood_params, g.T)
        assert abs(logL - g.ref_logL) <= 1e-12 * abs(g.ref_logL), name            # only the summation order differs
        pr, pst = oracle.prior(p.prior_case, p.params, p.plength, p.priors, p.priors_switch, p.extra_priors)
        assert pst == 0 and pr == g.ref_logPrior, name


def test_oracle_matches_reference_build_on_perturbed_points(oracle, reflib, W, golden_problems):
    if not reflib.available:
        pytest.skip("oracle/_ref not built here (needs /root/reference)")
    for name, g in golden_problems.items():
        p = g.prob
        P = W.perturbed_params(W.Problem(**{**p.__dict__, "err_frac": np.full(p.Np, 0.01), "err_off": np.full(p.Np, 1e-3)}), 4, scale=1.0)
        for row in P[1:]:
            st, m, _ = oracle.model(p.model_case, row, p.plength, p.x)
            st2, m2 = reflib.model(p.model_case, row, p.plength, p.x)
            assert st == 0 and st2 == 0
            assert np.array_equal(m, m2, equal_nan=True), name
            a = oracle.prior(p.prior_case, row, p.plength, p.priors, p.priors_switch, p.extra_priors)[0]
            b = reflib.prior(p.prior_case, row, p.plength, p.priors, p.priors_switch, p.extra_priors)
            assert a == b or (math.isnan(a) and math.isnan(b)), name


def test_batched_generate_model_matches_single(oracle, W):
    prob = W.small_global(3, Nx=1500, step=0.2)
    st, m, _ = oracle.model(prob.model_case, prob.params, prob.plength, prob.x)
    prob.y = m * (-np.log(W.philox_uniform(m.size, 5)))
    P = W.perturbed_params(prob, 6)
    T = W.tcoefs(6)
    r = oracle.generate_models(prob, P, T, want_model=True)
    for c in range(6):
        st, mc, L = oracle.model(prob.model_case, P[c], prob.plength, prob.x)
        assert np.array_equal(mc, r["model"][c]) and L == r["L"][c]
        assert r["logL"][c] == oracle.likelihood(0, prob.y, None, mc, 1.0, T[c])
        assert r["logPost"][c] == r["logL"][c] + r["logPrior"][c]


def test_dead_model_cases_are_flagged(oracle, W):
    prob = W.small_global(3, Nx=500, step=0.5)
    for mc in (4, 5):
        st, _, _ = oracle.model(mc, prob.params, prob.plength, prob.x)
        assert st == 4
    st, _, _ = oracle.model(99, prob.params, prob.plength, prob.x)
    assert st == 3
