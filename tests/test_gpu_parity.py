"""GPU parity tests (-m gpu): the CUDA path, called through the C-ABI, against the CPU oracle on
the same inputs and against the committed golden vectors of the unmodified reference sources.

Tolerance (BASELINE.json north_star): model spectrum and logL within 1e-10 relative in fp64.  The
tests assert 1e-12 (observed: ~1e-15), i.e. two orders of margin."""
import numpy as np
import pytest

from conftest import relerr

pytestmark = pytest.mark.gpu
TOL = 1e-12


def _with_y(prob, oracle, W, seed=77):
    st, m, _ = oracle.model(prob.model_case, prob.params, prob.plength, prob.x)
    assert st == 0
    prob.y = m * (-np.log(W.philox_uniform(m.size, seed)))
    return prob


def _check(md, oracle, prob, P, T, want_model=True, tol=TOL):
    ref = oracle.generate_models(prob, P, T, want_model=want_model)
    r = md.evaluate(P, T, want_model=want_model)
    assert np.array_equal(r["status"] & 3, np.where(ref["status"] == 1, 1, np.where(ref["status"] == 2, 2, 0)))
    ok = ref["status"] == 0
    if want_model:
        assert relerr(r["model"][ok], ref["model"][ok]) < tol
    assert relerr(r["logL"][ok], ref["logL"][ok]) < tol
    fin = np.isfinite(ref["logPrior"])
    assert np.array_equal(np.isfinite(r["logPrior"]), fin)
    assert np.allclose(r["logPrior"][fin], ref["logPrior"][fin], rtol=1e-12, atol=1e-11)
    assert np.array_equal(r["logPrior"][~fin & ~np.isnan(ref["logPrior"])], ref["logPrior"][~fin & ~np.isnan(ref["logPrior"])])
    L, _ = md.last_work(P.shape[0])
    assert np.array_equal(L[ok], ref["L"][ok])           # identical truncation windows, bin for bin
    return r, ref


def test_golden_vectors_of_the_reference(tb, golden_problems):
    """Every live model case + the shipped cfg1 data slice: GPU vs the reference's own outputs."""
    for name, g in golden_problems.items():
        p = g.prob
        md = tb.GpuModelDef(p, tb.MalaConfig(Nchains=2))
        r = md.evaluate(p.params[None, :], np.array([g.T]), want_model=True)
        assert r["status"][0] == 0, name
        assert relerr(r["model"][0], g.ref_model) < TOL, name
        assert abs(r["logL"][0] - g.ref_logL) <= TOL * abs(g.ref_logL), name
        assert abs(r["logPrior"][0] - g.ref_logPrior) <= 1e-12 * abs(g.ref_logPrior) + 1e-12, name
        assert r["logPost"][0] == r["logL"][0] + r["logPrior"][0]
        md.close()


@pytest.mark.parametrize("mc", [2, 3, 6, 7, 8, 9, 10, 12, 13])
def test_global_family_vs_oracle(tb, W, oracle, mc):
    prob = _with_y(W.small_global(mc), oracle, W)
    md = tb.GpuModelDef(prob, tb.MalaConfig(Nchains=4))
    P = W.perturbed_params(prob, 7, scale=1.0)
    _check(md, oracle, prob, P, W.tcoefs(7))
    md.close()


@pytest.mark.parametrize("kw", [dict(asym=20.0, do_amp=1, a3=0.02), dict(trunc_c=10000.0), dict(asym=-8.0),
                                dict(lmax=2), dict(lmax=1), dict(norders=9, Nx=9000)])
def test_global_variants_vs_oracle(tb, W, oracle, kw):
    prob = _with_y(W.small_global(3, **kw), oracle, W)
    md = tb.GpuModelDef(prob, tb.MalaConfig(Nchains=3))
    _check(md, oracle, prob, W.perturbed_params(prob, 5, scale=1.0), W.tcoefs(5))
    md.close()


@pytest.mark.parametrize("mc,kw", [(11, {}), (14, {}), (11, dict(asym=15.0, do_amp=1)), (14, dict(do_amp=1))])
def test_local_family_vs_oracle(tb, W, oracle, mc, kw):
    prob = _with_y(W.small_local(mc, **kw), oracle, W)
    md = tb.GpuModelDef(prob, tb.MalaConfig(Nchains=10))
    _check(md, oracle, prob, W.perturbed_params(prob, 10, scale=1.0), W.tcoefs(10))
    md.close()


@pytest.mark.parametrize("mc,lc", [(0, 1), (0, 0), (1, 0), (1, 1)])
def test_gaussian_models_and_both_likelihoods(tb, W, oracle, mc, lc):
    prob = W.gauss_problem(mc, like_case=lc)
    if lc == 0:
        prob.y = np.abs(prob.y) + 0.1
    md = tb.GpuModelDef(prob, tb.MalaConfig(Nchains=2))
    _check(md, oracle, prob, W.perturbed_params(prob, 4, scale=0.5), W.tcoefs(4))
    md.close()


def test_cfg2_s_local(tb, W, oracle):
    prob = _with_y(W.s_local(), oracle, W, seed=12345)
    md = tb.GpuModelDef(prob, tb.MalaConfig(Nchains=10))
    _check(md, oracle, prob, W.perturbed_params(prob, 10), W.tcoefs(10))
    md.close()


@pytest.mark.parametrize("Nx", [2, 3, 255, 256, 257, 1023, 1025, 4099])
def test_ragged_sizes_and_tail_tiles(tb, W, oracle, Nx):
    prob = _with_y(W.small_local(11, Nx=Nx, step=30.0 / max(Nx, 2)), oracle, W)
    md = tb.GpuModelDef(prob, tb.MalaConfig(Nchains=2))
    for bpt in (1, 2, 4):
        md.set_tuning(2, bpt, -1)
        _check(md, oracle, prob, W.perturbed_params(prob, 3, scale=0.5), W.tcoefs(3))
    md.close()


def test_tuning_variants_give_the_same_numbers(tb, W, oracle):
    prob = _with_y(W.small_global(3, Nx=20000, step=0.02), oracle, W)
    md = tb.GpuModelDef(prob, tb.MalaConfig(Nchains=6))
    P, T = W.perturbed_params(prob, 6), W.tcoefs(6)
    base = None
    for bpt in (1, 2, 4):
        for cpb in (1, 4, 6):
            md.set_tuning(cpb, bpt, -1)
            r = md.evaluate(P, T)
            if base is None:
                base = r
            assert relerr(r["logL"], base["logL"]) < 1e-13
    _check(md, oracle, prob, P, T)
    md.close()


def test_batch_larger_than_chain_count_and_repeatability(tb, W, oracle):
    prob = _with_y(W.small_global(2, Nx=3000, step=0.1), oracle, W)
    md = tb.GpuModelDef(prob, tb.MalaConfig(Nchains=2))
    P, T = W.perturbed_params(prob, 33, scale=1.0), W.tcoefs(33) % 7 + 1
    r1, _ = _check(md, oracle, prob, P, T, want_model=False)
    r2 = md.evaluate(P, T)
    assert np.array_equal(r1["logL"], r2["logL"])         # fixed reduction order: bit-reproducible
    md.close()


def test_failure_flags_instead_of_exit(tb, W, oracle):
    """Empty window (reference: exit) and NaN splitting (reference: undefined) become per-chain flags."""
    prob = _with_y(W.small_local(11), oracle, W)
    md = tb.GpuModelDef(prob, tb.MalaConfig(Nchains=2))
    P = np.tile(prob.params, (4, 1))
    iI = prob.plength[:9].sum()
    P[1, iI + 1] = -3.0                      # trunc_c < 0 => imax-imin <= 0
    P[2, prob.plength[0] + prob.plength[2:6].sum() + 3] = np.nan     # sqrt(a1)cos i = NaN => a1 NaN
    r = md.evaluate(P, np.ones(4))
    ref = oracle.generate_models(prob, P, np.ones(4))
    assert r["status"][0] == 0 and r["status"][3] == 0
    assert (r["status"][1] & tb.ST_EMPTY_WINDOW) and ref["status"][1] == 1
    assert (r["status"][2] & tb.ST_NAN_WINDOW) and ref["status"][2] == 2
    assert np.isnan(r["logL"][1]) and np.isnan(r["logL"][2]) and (r["status"][1] & tb.ST_NAN_LOGL)
    assert relerr(r["logL"][[0, 3]], ref["logL"][[0, 3]]) < TOL
    md.close()


def test_prior_hard_limits(tb, W, oracle):
    prob = _with_y(W.small_global(3), oracle, W)
    md = tb.GpuModelDef(prob, tb.MalaConfig(Nchains=2))
    pl = prob.plength
    P = np.tile(prob.params, (5, 1))
    P[1, pl[0]] = -0.1                                        # negative visibility => -inf (priors_calc.cpp:53-57)
    iS = pl[0] + pl[1] + pl[2:6].sum()
    P[2, iS + 2] = 0.5 * P[2, iS]                             # |a3/a1| >= 0.2 => -inf (:137-139)
    P[3, pl[0] + pl[1]] += 3.5                                # frequency outside its GUG core (+-2): finite Gaussian penalty
    P[4, 0] = 2000.0                                          # height beyond the Jeffreys range => -inf
    r, ref = _check(md, oracle, prob, P, np.ones(5), want_model=False)
    assert r["logPrior"][1] == -np.inf and r["logPrior"][2] == -np.inf and r["logPrior"][4] == -np.inf
    assert np.isfinite(r["logPrior"][3]) and r["logPrior"][3] < r["logPrior"][0] - 5
    md.close()


def test_cfg3_full_size_vs_oracle_and_properties(tb, W, oracle):
    """1M-bin S-global, 10 chains: direct oracle comparison plus size-independent properties."""
    prob = _with_y(W.s_global(), oracle, W, seed=12345)
    C = 10
    md = tb.GpuModelDef(prob, tb.MalaConfig(Nchains=C))
    P, T = W.perturbed_params(prob, C), W.tcoefs(C)
    r, ref = _check(md, oracle, prob, P, T, want_model=False)
    # tempering: logL(T) * T is temperature independent
    r1 = md.evaluate(P, np.ones(C))
    assert relerr(r["logL"] * T, r1["logL"]) < 1e-14
    # additivity over bins: the logL of the full spectrum equals the sum over two bin shards
    S = tb.sharding
    tot = np.zeros(C)
    for rank in range(2):
        desc, sl = S.shard_descriptor(prob.x, 2, rank)
        sub = W.Problem(**{**prob.__dict__, "x": prob.x[sl].copy(), "y": prob.y[sl].copy()})
        mds = tb.GpuModelDef(sub, tb.MalaConfig(Nchains=C), shard=desc)
        tot += mds.evaluate(P, np.ones(C))["logL"]
        mds.close()
    assert relerr(tot, r1["logL"]) < 1e-13
    # model spectrum of one chain against the oracle, bin for bin
    m = md.call_model_explicit(P[3:4])
    assert relerr(m, oracle.model(prob.model_case, P[3], prob.plength, prob.x)[1]) < TOL
    md.close()


@pytest.mark.parametrize("case", ["cfg3", "general_p", "sharp_knee"])
def test_background_moments_equal_the_bin_by_bin_path(tb, W, oracle, case, monkeypatch):
    """Outside every truncation window the fused kernel takes sum y/b and sum ln b of a tile from Chebyshev moments of the
    data (tg_eval.cuh: tg_moment_tiles) whenever its polynomial check passes, and bin by bin otherwise.  The two paths must
    agree to rounding, and both with the oracle -- for integer Harvey exponents, for non-integer ones (the exp/log branch)
    and for a background whose knee is so sharp that most tiles must fail the check."""
    prob = W.s_global(Nx=400000, step=0.02)
    ip = [i for i, n in enumerate(prob.names) if n == "Harvey-Noise_p"]
    itc = [i for i, n in enumerate(prob.names) if n == "Harvey-Noise_tc"]
    if case == "general_p":
        prob.params[ip[1]], prob.params[ip[2]] = 3.37, 1.83
    if case == "sharp_knee":
        prob.params[ip[2]] = 40.0                    # (x / 4000 microHz)^40: a cliff in the middle of the background-only part
        prob.params[itc[2]] = 0.25
    prob = _with_y(prob, oracle, W, seed=5)
    C = 6
    P, T = W.perturbed_params(prob, C, scale=1.0), W.tcoefs(C, 1.3)
    res = {}
    for mode in ("0", "1"):
        monkeypatch.setenv("TG_MOMENTS", mode)
        md = tb.GpuModelDef(prob, tb.MalaConfig(Nchains=C))
        md.evaluate(P, T)                            # first launch flags the tiles whose check fails ...
        md.set_tuning(light_subtiles=9); md.set_tuning(light_subtiles=8)       # ... the work items are rebuilt around them
        res[mode] = md.evaluate(P, T)
        md.close()
    ref = oracle.generate_models(prob, P, T)
    assert np.array_equal(res["0"]["status"], res["1"]["status"])
    assert relerr(res["1"]["logL"], res["0"]["logL"]) < 2e-14
    assert relerr(res["1"]["logL"], ref["logL"]) < TOL and relerr(res["0"]["logL"], ref["logL"]) < TOL


def test_background_moments_over_the_prior_ranges(tb, W, oracle, monkeypatch):
    """the acceptance of a moment tile rests on an a-posteriori check at five points, so the path is compared with
    the bin-by-bin one for backgrounds drawn over (and beyond) the prior ranges of the Harvey parameters -- heights over five
    decades, time scales from far below to far above the spectrum's range (knees inside, left and right of it), exponents
    1 ... 4.5 (integer and not), white noise from negligible to dominant -- and with the oracle."""
    prob = W.s_global(Nx=400000, step=0.02)
    prob = _with_y(prob, oracle, W, seed=17)
    iH = [i for i, n in enumerate(prob.names) if n == "Harvey-Noise_H"]
    itc = [i for i, n in enumerate(prob.names) if n == "Harvey-Noise_tc"]
    ip = [i for i, n in enumerate(prob.names) if n == "Harvey-Noise_p"]
    iN = [i for i, n in enumerate(prob.names) if n == "White_Noise_N0"]
    Cn = 24
    g = np.random.Generator(np.random.Philox(key=2025))
    P = np.tile(prob.params, (Cn, 1))
    for c in range(1, Cn):
        for k in range(len(iH)):
            if prob.params[itc[k]] == 0.0 and g.random() < 0.5:
                continue                                   # an inactive term stays inactive in half of the draws
            P[c, iH[k]] = 10.0 ** g.uniform(-2, 3)
            P[c, itc[k]] = 10.0 ** g.uniform(-2.5, 2.5)     # knee at 1e3 / tc microHz: from 3e5 (far right) down to 3 (far left)
            P[c, ip[k]] = [1.0, 2.0, 3.0, 4.0, g.uniform(1.0, 4.5), g.uniform(1.0, 4.5)][int(g.integers(0, 6))]
        P[c, iN[0]] = 10.0 ** g.uniform(-3, 1)
    T = np.ones(Cn)
    res = {}
    for mode in ("0", "1"):
        monkeypatch.setenv("TG_MOMENTS", mode)
        md = tb.GpuModelDef(prob, tb.MalaConfig(Nchains=Cn))
        res[mode] = md.evaluate(P, T)
        md.close()
    ref = oracle.generate_models(prob, P, T)
    assert np.array_equal(res["0"]["status"], res["1"]["status"]) and np.array_equal(res["1"]["status"], ref["status"])
    ok = ref["status"] == 0
    assert ok.sum() >= Cn - 2
    assert relerr(res["1"]["logL"][ok], res["0"]["logL"][ok]) < 2e-14
    assert relerr(res["1"]["logL"][ok], ref["logL"][ok]) < TOL


@pytest.mark.parametrize("case", ["cfg3", "hot_chains", "narrow_modes", "general_p"])
def test_far_field_path_equals_the_exact_path(tb, W, oracle, case, monkeypatch):
    """Under the truncation windows (build_lorentzian.cpp:365-460) the fused kernel takes, per 512-bin micro-tile, a polynomial
    interpolant of the components whose poles are far (prepared by the pre-pass tg_k_ff, tg_farfield.cuh) plus the exact
    records of the near ones, instead of every component bin by bin.  Both paths must agree to rounding and both with the
    oracle: at the usual perturbation, with hot-chain excursions three times wider, with line widths five times narrower (more
    components inside the near radius of a micro-tile edge) and with non-integer Harvey exponents under the windows."""
    import ctypes as C_
    prob = W.s_global(Nx=400000, step=0.02)
    if case == "narrow_modes":
        iw = [i for i, n in enumerate(prob.names) if n.startswith("Width")]
        assert iw
        prob.params[iw] *= 0.2
    if case == "general_p":
        ip = [i for i, n in enumerate(prob.names) if n == "Harvey-Noise_p"]
        prob.params[ip[1]], prob.params[ip[2]] = 3.37, 1.83
    prob = _with_y(prob, oracle, W, seed=9)
    Cn = 8
    P, T = W.perturbed_params(prob, Cn, scale=3.0 if case == "hot_chains" else 1.0, seed=31), W.tcoefs(Cn, 1.3)
    res, summ = {}, None
    for mode in ("0", "1"):
        monkeypatch.setenv("TG_FARFIELD", mode)
        md = tb.GpuModelDef(prob, tb.MalaConfig(Nchains=Cn))
        res[mode] = md.evaluate(P, T)
        if mode == "1":
            out = (C_.c_long * 6)()
            assert md.lib.lib.tg_dbg_ff_summary(md.h, Cn, out) == 0
            summ = list(out)
        md.close()
    ref = oracle.generate_models(prob, P, T)
    # the path was taken: micro-tiles prepared, components interpolated, none fell back to the component-by-component loop
    assert summ[0] > 0 and summ[3] > 10 * summ[0] and summ[1] == 0, summ
    assert np.array_equal(res["0"]["status"], res["1"]["status"]) and np.array_equal(res["1"]["status"], ref["status"])
    ok = ref["status"] == 0
    assert ok.sum() >= 2
    assert relerr(res["1"]["logL"][ok], res["0"]["logL"][ok]) < 2e-14
    assert relerr(res["1"]["logL"][ok], ref["logL"][ok]) < TOL and relerr(res["0"]["logL"][ok], ref["logL"][ok]) < TOL


def test_async_eval_equals_the_blocking_call(tb, W, oracle):
    """tg_eval_async / tg_eval_wait (two staging slots, MALA.cpp:632-655 with the proposal on the host): a pipeline of calls
    through alternating slots returns, call by call, the bits of tg_eval; misuse of a slot is refused."""
    prob = W.small_global(3, Nx=20000, step=0.06)
    st, m, _ = oracle.model(prob.model_case, prob.params, prob.plength, prob.x)
    prob.y = m * (-np.log(W.philox_uniform(m.size, 5)))
    Cn, K = 6, 7
    md = tb.GpuModelDef(prob, tb.MalaConfig(Nchains=Cn))
    T = W.tcoefs(Cn)
    Ps = [W.perturbed_params(prob, Cn, scale=0.2 + 0.1 * k, seed=20 + k) for k in range(K)]
    ref = [md.evaluate(P, T) for P in Ps]
    got = [None] * K
    md.evaluate_async(Ps[0], T, 0)
    for k in range(1, K):
        md.evaluate_async(Ps[k], T, k & 1)
        got[k - 1] = md.evaluate_wait(Cn, (k - 1) & 1)
    got[K - 1] = md.evaluate_wait(Cn, (K - 1) & 1)
    for k in range(K):
        for key in ("logL", "logPrior", "logPost", "status"):
            assert np.array_equal(got[k][key], ref[k][key], equal_nan=True), (k, key)
    orc = oracle.generate_models(prob, Ps[3], T)
    ok = orc["status"] == 0
    assert relerr(got[3]["logL"][ok], orc["logL"][ok]) < 1e-12
    with pytest.raises(tb.TgError):
        md.evaluate_wait(Cn, 0)                      # nothing outstanding
    md.evaluate_async(Ps[0], T, 1)
    with pytest.raises(tb.TgError):
        md.evaluate_async(Ps[1], T, 1)               # slot 1 not collected yet
    with pytest.raises(tb.TgError):
        md.evaluate_async(Ps[1][:3], T[:3], 0)       # another batch size while slot 1 is in flight
    assert np.array_equal(md.evaluate_wait(Cn, 1)["logL"], ref[0]["logL"])
    md.close()
