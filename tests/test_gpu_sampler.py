"""GPU tests of the device-resident sampler (MALA::execute replacement): bookkeeping invariants,
reproducibility / resume, and posterior agreement with (a) the Fisher-matrix prediction and (b) a
CPU sampler that follows MALA.cpp with the oracle as its model -- within Monte-Carlo error, the
only parity that is meaningful for the sampler (reference RNG is rand()+time(NULL), quirk Q10)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _gauss(W):
    return W.gauss_problem(0, like_case=1, Nx=400)


def test_records_and_invariants(tb, W):
    prob = _gauss(W)
    cfg = tb.MalaConfig(Nchains=6, Nt_learn=[50, 400], periods_learn=[1], c0=5.0, dN_mixing=2, seed=99,
                        flags=tb.gpu.F_FIX_SWAP_BOOKKEEPING)
    md = tb.GpuModelDef(prob, cfg)
    out = tb.GpuMALA(md).execute(0, 600)
    assert np.all(out["status"] == 0)
    assert np.allclose(out["logPost"], out["logL"] + out["logPrior"], rtol=0, atol=1e-9)
    assert np.all((out["Pmove"] >= 0) & (out["Pmove"] <= 1))
    # swap cadence: every dN_mixing steps, never at i == 0 (MALA.cpp:676)
    att = out["swap_attempt"][:, 0]
    assert att[0] == 0 and np.array_equal(att[1:], (np.arange(1, 600) % 2 == 0).astype(np.int32))
    assert np.all(out["swap_chain0"][att == 1] >= 0) and np.all(out["swap_chain0"][att == 1] < 5)
    assert np.all(out["swap_chain0"][att == 0] == -1) and np.all(out["swap_done"][att == 0] == 0)
    assert 0.05 < out["swap_done"][att == 1].mean() <= 1.0
    # a chain that did not move and was not swapped keeps its variables
    v = out["vars"]
    for i in range(1, 600):
        sw = set()
        if out["swap_done"][i, 0]:
            a = out["swap_chain0"][i, 0]; sw = {a, a + 1}
        for c in range(6):
            if not out["moved"][i, c] and c not in sw:
                assert np.array_equal(v[i, c], v[i - 1, c])
    # host mirror == last record
    assert np.array_equal(md.vars, v[-1]) and np.array_equal(md.logLikelihood, out["logL"][-1])
    mu, sigma, cov = md.get_proposal()
    assert np.all(sigma > 0) and np.all(np.linalg.eigvalsh(cov[0]) > -1e-12)
    md.close()


def test_reproducible_and_resumable(tb, W):
    prob = _gauss(W)
    mk = lambda seed: tb.MalaConfig(Nchains=4, Nt_learn=[20, 10000], periods_learn=[1], c0=5.0, seed=seed,
                                    flags=tb.gpu.F_FIX_SWAP_BOOKKEEPING)
    a = tb.GpuModelDef(prob, mk(5)); ra = tb.GpuMALA(a).execute(0, 150)
    b = tb.GpuModelDef(prob, mk(5)); rb1 = tb.GpuMALA(b).execute(0, 100)
    for k in ("vars", "logL", "moved", "swap_done"):
        assert np.array_equal(ra[k][:100], rb1[k])
    # restore into a fresh handle (restore files: vars, mu, sigma, covarmat) and continue at i0=100
    mu, sigma, cov = b.get_proposal()
    c = tb.GpuModelDef(prob, mk(5))
    c.set_state(vars_=b.vars, mu=mu, sigma=sigma, covarmat=cov)
    rc = tb.GpuMALA(c).execute(100, 50)
    for k in ("vars", "logL", "logPost", "moved", "swap_done", "Pmove"):
        assert np.array_equal(ra[k][100:], rc[k]), k
    d = tb.GpuModelDef(prob, mk(6)); rd = tb.GpuMALA(d).execute(0, 100)
    assert not np.array_equal(rd["vars"], rb1["vars"])
    for m in (a, b, c, d):
        m.close()


def test_graph_and_direct_launch_agree(tb, W):
    prob = _gauss(W)
    cfg = tb.MalaConfig(Nchains=4, Nt_learn=[10, 10000], periods_learn=[2], c0=5.0, seed=3)
    a = tb.GpuModelDef(prob, cfg); a.set_tuning(0, 0, 1); ra = tb.GpuMALA(a).execute(0, 100)
    b = tb.GpuModelDef(prob, cfg); b.set_tuning(0, 0, 0); rb = tb.GpuMALA(b).execute(0, 100)
    for k in ra:
        assert np.array_equal(ra[k], rb[k]), k
    a.close(); b.close()


@pytest.mark.parametrize("case", ["gauss", "local", "global"])
def test_programmatic_dependent_launch_changes_nothing(tb, W, case, monkeypatch):
    """The step kernels overlap their launch with the predecessor's tail (griddepcontrol, tg_internal.cuh); the records of
    a run must be the same bits with the overlap turned off (TG_PDL=0 is read when the handle is created)."""
    prob = {"gauss": lambda: _gauss(W), "local": lambda: W.s_local(), "global": lambda: W.s_global(Nx=100000, step=0.05, norders=4, numax=2500.0, dnu=120.0)}[case]()
    cfg = tb.MalaConfig(Nchains=10, Nt_learn=[20, 10000], periods_learn=[1], c0=2.0, dN_mixing=1, seed=77)
    res = []
    for pdl in ("0", "1"):
        monkeypatch.setenv("TG_PDL", pdl)
        md = tb.GpuModelDef(prob, cfg)
        res.append(tb.GpuMALA(md).execute(0, 160))
        md.close()
    for k in res[0]:
        assert np.array_equal(res[0][k], res[1][k], equal_nan=res[0][k].dtype.kind == "f"), k
    assert np.all(np.isfinite(res[1]["logPost"][:, 0]))


def _cpu_sampler(oracle, prob, cfg, nsteps, rng):
    """MALA.cpp:608-692 with numpy RNG and the oracle as generate_model."""
    C = cfg.Nchains
    T = cfg.tcoefs()
    idx = prob.index_to_relax
    Nv = idx.size
    params = np.tile(prob.params, (C, 1))
    x = params[:, idx].copy()
    cur = oracle.generate_models(prob, params, T)
    logL, logPr, logPo = cur["logL"].copy(), cur["logPrior"].copy(), cur["logPost"].copy()
    err = prob.init_errors()
    cov = np.tile(np.diag(err**2), (C, 1, 1))
    sigma = 2.38**2 * T**0.2 / Nv
    mu = x.copy()
    keep = np.empty((nsteps, Nv))
    for i in range(nsteps):
        gamma = cfg.c0 / (1.0 + i)
        prop = params.copy()
        xn = np.empty_like(x)
        for m in range(C):
            L = np.linalg.cholesky((cov[m] + cfg.epsilon2 * np.eye(Nv)) * sigma[m])
            xn[m] = x[m] + L @ rng.standard_normal(Nv)
        prop[:, idx] = xn
        new = oracle.generate_models(prob, prop, T)
        for m in range(C):
            if np.isnan(new["logL"][m]) or new["logPost"][m] == -np.inf:
                r = 0.0
            else:
                r = min(1.0, np.exp(new["logPost"][m] - logPo[m]))
            if rng.random() <= r:
                x[m], params[m] = xn[m], prop[m]
                logL[m], logPr[m], logPo[m] = new["logL"][m], new["logPrior"][m], new["logPost"][m]
            learn = any(cfg.Nt_learn[l] <= i < cfg.Nt_learn[l + 1] and i % cfg.periods_learn[l] == 0 for l in range(len(cfg.periods_learn)))
            if learn:
                mu[m] = mu[m] + gamma * (x[m] - mu[m])
                d = x[m] - mu[m]
                cov[m] = cov[m] + gamma * (np.outer(d, d) - cov[m])
                sigma[m] = min(max(sigma[m] + gamma * (r - cfg.target_acceptance), cfg.epsilon1), cfg.A1)
        if i % cfg.dN_mixing == 0 and i != 0:
            A = rng.integers(0, C - 1); B = A + 1
            LA_TB, LB_TA = logL[A] * T[A] / T[B], logL[B] * T[B] / T[A]
            rT = min(1.0, np.exp(LA_TB + LB_TA - logL[A] - logL[B]))
            if rng.random() <= rT:
                x[[A, B]] = x[[B, A]]; params[[A, B]] = params[[B, A]]
                prA, prB = logPr[A], logPr[B]
                logL[A], logPr[A], logPo[A] = LB_TA, prB, LB_TA + prB
                logL[B], logPr[B], logPo[B] = LA_TB, prA, LA_TB + prA
        keep[i] = x[0]
    return keep


def test_posterior_matches_fisher_and_cpu_sampler(tb, W, oracle):
    prob = _gauss(W)
    cfg = tb.MalaConfig(Nchains=4, Nt_learn=[100, 4000], periods_learn=[1], c0=5.0, dN_mixing=1, seed=2024,
                        flags=tb.gpu.F_FIX_SWAP_BOOKKEEPING)
    # Fisher prediction around the maximum: cov = (J^T J / sigma^2)^-1 (chi-square likelihood, flat priors)
    p0 = prob.params.copy()
    J = np.empty((prob.x.size, 4))
    for k in range(4):
        h = 1e-5 * max(1.0, abs(p0[k]))
        a, b = p0.copy(), p0.copy(); a[k] += h; b[k] -= h
        J[:, k] = (oracle.model(0, a, prob.plength, prob.x)[1] - oracle.model(0, b, prob.plength, prob.x)[1]) / (2 * h)
    # likelihood_chi_square is -sum((y-m)^2/sigma^2) (no 1/2): posterior covariance = (2 J^T J / sigma^2)^-1
    fisher_cov = np.linalg.inv(2.0 * (J.T / prob.sigma_y**2) @ J)
    md = tb.GpuModelDef(prob, cfg)
    mala = tb.GpuMALA(md)
    mala.execute(0, 4000, record=False)
    out = mala.execute(4000, 30000)
    s = out["vars"][:, 0, :]
    acc = out["moved"][:, 0].mean()
    assert 0.1 < acc < 0.45, acc                       # adapted towards target_acceptance = 0.234
    gpu_mean, gpu_std = s.mean(axis=0), s.std(axis=0)
    fstd = np.sqrt(np.diag(fisher_cov))
    assert np.all(np.abs(gpu_std / fstd - 1.0) < 0.2), (gpu_std, fstd)
    # CPU sampler (same algorithm, oracle model)
    cpu = _cpu_sampler(oracle, prob, cfg, 14000, np.random.default_rng(7))[4000:]
    cpu_mean, cpu_std = cpu.mean(axis=0), cpu.std(axis=0)
    assert np.all(np.abs(gpu_mean - cpu_mean) < 0.35 * fstd), (gpu_mean, cpu_mean, fstd)
    assert np.all(np.abs(gpu_std / cpu_std - 1.0) < 0.3), (gpu_std, cpu_std)
    # hotter chains are broader
    hot = out["vars"][:, 3, :].std(axis=0)
    assert np.all(hot > gpu_std)
    md.close()


def test_global_fit_runs_and_learns(tb, W, oracle):
    """A reduced global MS fit: the sampler runs, adapts and keeps the chains at high posterior."""
    prob = W.small_global(3, Nx=4000, step=0.08)
    st, m, _ = oracle.model(prob.model_case, prob.params, prob.plength, prob.x)
    prob.y = m * (-np.log(W.philox_uniform(m.size, 8)))
    cfg = tb.MalaConfig(Nchains=10, Nt_learn=[100, 100000], periods_learn=[1], c0=2.0, seed=1)
    md = tb.GpuModelDef(prob, cfg)
    l0 = md.logLikelihood[0]
    out = tb.GpuMALA(md).execute(0, 3000)
    assert np.all((out["status"] & ~tb.ST_NAN_LOGL & ~tb.ST_EMPTY_WINDOW) == 0)
    late = out["moved"][2000:, 0].mean()
    assert 0.03 < late < 0.6, late
    assert out["logL"][-500:, 0].mean() > l0 - 5 * prob.Nv     # stays near the mode of the posterior
    assert np.all(np.diff(out["logL"][-500:].mean(axis=0) * cfg.tcoefs()) < 0)   # hotter chains sit at lower untempered logL
    md.close()


@pytest.mark.gpu
@pytest.mark.parametrize("norders,path", [(14, "registers (Nv <= 96)"), (18, "shared-memory matrix"), (30, "in place in HBM")])
def test_adaptation_equals_a_sequential_replay(tb, W, norders, path):
    """(mu, Sigma, sigma) after N adaptive steps == MALA::update_proposal (MALA.cpp:292-315) replayed on the
    recorded chain: the speculative two-candidate scheme is exact, on all three factorisation paths."""
    prob = W.s_global(Nx=100000, step=0.05, norders=norders, numax=2500.0, dnu=120.0)
    prob.y = np.abs(np.random.default_rng(3).standard_normal(prob.x.size)) + 0.1
    n0, N = 5, 40
    cfg = tb.MalaConfig(Nchains=3, Nt_learn=[n0, 10**6], periods_learn=[1], c0=2.0, seed=11, dN_mixing=0)
    md = tb.GpuModelDef(prob, cfg)
    Nv = md.Nvars
    assert (Nv <= 96) == (norders == 14) and (Nv > 170) == (norders == 30), Nv
    mu, sigma, cov = [a.copy() for a in md.get_proposal()]
    out = tb.GpuMALA(md).execute(0, N)
    for i in range(n0, N):
        gamma = cfg.c0 / (1.0 + i)
        for c in range(3):
            x = out["vars"][i, c]
            mu[c] = mu[c] + gamma * (x - mu[c])
            d = x - mu[c]
            cov[c] = cov[c] + gamma * (np.outer(d, d) - cov[c])
            sigma[c] = min(max(sigma[c] + gamma * (out["Pmove"][i, c] - cfg.target_acceptance), cfg.epsilon1), cfg.A1)
    mu_g, sigma_g, cov_g = md.get_proposal()
    assert np.allclose(mu_g, mu, rtol=1e-12, atol=0)
    assert np.allclose(sigma_g, sigma, rtol=1e-12, atol=0)
    scale = np.sqrt(np.einsum("cii->ci", cov))
    # x - mu cancels ~5 digits (frequencies ~2500 with steps ~0.01) and the GPU contracts a + g*(b - a) into an
    # FMA, so last-ulp differences of mu show up at ~1e-11 of sqrt(S_ii S_kk); a wrong candidate would be O(1)
    assert np.max(np.abs(cov_g - cov) / (scale[:, :, None] * scale[:, None, :])) < 1e-8
    assert np.array_equal(cov_g, np.swapaxes(cov_g, 1, 2))                      # kept exactly symmetric
    assert np.all(out["moved"][n0:].mean(axis=0) > 0.0)                          # the factors are usable: chains move
    assert np.all((out["status"] & tb.ST_CHOL_FAIL) == 0)
    md.close()


@pytest.mark.gpu
def test_restore_mirrors_the_lower_triangle(tb, W):
    """llt() reads the lower triangle only (MALA.cpp:342): a restore file with a garbage upper triangle gives
    the same chain as the symmetric one."""
    prob = _gauss(W)
    cfg = tb.MalaConfig(Nchains=2, Nt_learn=[0, 10**6], periods_learn=[1], c0=2.0, seed=2)
    a = tb.GpuModelDef(prob, cfg); b = tb.GpuModelDef(prob, cfg)
    mu, sigma, cov = a.get_proposal()
    bad = cov.copy()
    iu = np.triu_indices(cov.shape[1], 1)
    bad[:, iu[0], iu[1]] += 0.37
    b.set_state(covarmat=bad)
    ra = tb.GpuMALA(a).execute(0, 30); rb = tb.GpuMALA(b).execute(0, 30)
    assert np.array_equal(ra["vars"], rb["vars"])
    a.close(); b.close()


def test_resume_through_the_restore_files(tb, W, tmp_path):
    """Run 80 steps, write the outputs and the restore triplet in the reference's formats, read them back into
    a fresh handle and continue: the continuation equals the uninterrupted run bit for bit."""
    prob = _gauss(W)
    mk = lambda: tb.MalaConfig(Nchains=4, Nt_learn=[20, 10000], periods_learn=[1], c0=5.0, seed=8, flags=tb.gpu.F_FIX_SWAP_BOOKKEEPING)
    a = tb.GpuModelDef(prob, mk()); ra = tb.GpuMALA(a).execute(0, 120)
    b = tb.GpuModelDef(prob, mk()); rb = tb.GpuMALA(b).execute(0, 80)
    out = tb.outputs.Outputs(prob, b.Tcoefs, str(tmp_path), root="g", Nsamples=120)
    out.append(rb); out.write_restore(b, 80); out.close()
    rs = tb.outputs.read_restore(str(tmp_path / "g_restore_"))
    assert rs["iteration"] == 80 and np.array_equal(rs["vars"], rb["vars"][-1])
    c = tb.GpuModelDef(prob, mk())
    c.set_state(vars_=rs["vars"], mu=rs["mu"], sigma=rs["sigma"], covarmat=rs["covarmat"])
    rc = tb.GpuMALA(c).execute(rs["iteration"], 40)
    for k in ("vars", "logL", "logPost", "moved", "swap_done", "Pmove"):
        assert np.array_equal(ra[k][80:], rc[k]), k
    # and the appended binary files hold the whole chain
    out2 = tb.outputs.Outputs(prob, b.Tcoefs, str(tmp_path), root="g", Nsamples=120, Nsamples_sofar=80, erase_old_files=False)
    out2.append(rc); out2.close()
    v0 = np.fromfile(str(tmp_path / "g_params_chain-0.bin"), dtype=np.float64).reshape(-1, a.Nvars)
    assert np.array_equal(v0, ra["vars"][:, 0])
    for m in (a, b, c):
        m.close()


@pytest.mark.parametrize("problem,per_ensemble", [("gauss", 0), ("global", 0), ("gauss", 2)])
def test_restore_means_equal_a_step_by_step_average(tb, W, oracle, problem, per_ensemble):
    """buf_restore.*_mean of the reference (Outputs::update_buffer_restore, outputs.cpp:1758-1818: sums of sigma, mu, covarmat
    and vars of every step, divided when the restore files are written; used by MALA::restore_proposal when
    do_restore_proposal_mean = 1, MALA.cpp:206-232).  The device keeps the sums (TG_F_RESTORE_MEANS); they must equal the average
    of the states a second handle with the same seed shows step after step -- learning on, swaps on, one and two ensembles."""
    if problem == "gauss":
        prob = _gauss(W)
    else:
        prob = W.small_global(3, Nx=3000, step=0.06)
        st, m, _ = oracle.model(prob.model_case, prob.params, prob.plength, prob.x)
        prob.y = m * (-np.log(W.philox_uniform(m.size, 4)))
    n, Cn = 37, 4
    mk = lambda flags: tb.MalaConfig(Nchains=Cn, chains_per_ensemble=per_ensemble, Nt_learn=[5, 10000], periods_learn=[1], c0=3.0, seed=21, flags=flags)
    a = tb.GpuModelDef(prob, mk(tb.gpu.F_RESTORE_MEANS))
    assert a.get_state_means(reset=False)["count"] == 0
    ra = tb.GpuMALA(a).execute(0, n)
    ma = a.get_state_means(reset=True)
    b = tb.GpuModelDef(prob, mk(0))
    with pytest.raises(tb.TgError):
        b.get_state_means()
    run_b = tb.GpuMALA(b)
    acc = dict(vars=0.0, mu=0.0, sigma=0.0, covarmat=0.0)
    for i in range(n):
        rb = run_b.execute(i, 1)
        assert np.array_equal(rb["vars"][0], ra["vars"][i])          # the flag does not touch the trajectory
        mu, sigma, cov = b.get_proposal()
        acc["vars"] = acc["vars"] + b.vars; acc["mu"] = acc["mu"] + mu; acc["sigma"] = acc["sigma"] + sigma; acc["covarmat"] = acc["covarmat"] + cov
    assert ma["count"] == n
    for k in acc:
        ref = acc[k] / n
        assert np.allclose(ma[k], ref, rtol=1e-12, atol=1e-13 * np.max(np.abs(ref))), k
    assert np.allclose(ma["vars"], ra["vars"].mean(axis=0), rtol=1e-12, atol=1e-300)
    # a new average starts after the reset: nothing summed -> the current values; then the next block alone
    cur = a.get_state_means(reset=False)
    mu, sigma, cov = a.get_proposal()
    assert cur["count"] == 0 and np.array_equal(cur["mu"], mu) and np.array_equal(cur["sigma"], sigma) and np.array_equal(cur["covarmat"], cov)
    r2 = tb.GpuMALA(a).execute(n, 20)
    m2 = a.get_state_means(reset=True)
    assert m2["count"] == 20 and np.allclose(m2["vars"], r2["vars"].mean(axis=0), rtol=1e-12, atol=1e-300)
    a.close(); b.close()


def test_chunked_run_on_a_multi_item_spectrum(tb, W, oracle):
    """on a spectrum that spans many work items (moment tiles, far-field units, bin-by-bin runs) the partition follows
    the chains' windows and is rebuilt after long calls, and the grouping of the per-(chain, item) partial sums changes logL in
    its last bits.  What holds, and is asserted: a run cut into calls of different lengths (a rebuild in between) takes the same
    accept / swap decisions as the uninterrupted run and its states agree to 1e-9 over 112 steps; the same sequence of calls
    repeated gives the same bits."""
    prob = W.s_global(Nx=300000, step=0.028)
    st, m, _ = oracle.model(prob.model_case, prob.params, prob.plength, prob.x)
    prob.y = m * (-np.log(W.philox_uniform(m.size, 8)))
    mk = lambda: tb.MalaConfig(Nchains=6, lambda_temp=1.7, c0=2.0, Nt_learn=[0, 10**9], periods_learn=[1], dN_mixing=1, seed=77)
    a = tb.GpuModelDef(prob, mk()); ra = tb.GpuMALA(a).execute(10000, 112)
    outs = []
    for rep in range(2):
        b = tb.GpuModelDef(prob, mk()); run = tb.GpuMALA(b)
        parts = [run.execute(10000, 70), run.execute(10070, 26), run.execute(10096, 16)]
        outs.append({k: np.concatenate([p[k] for p in parts]) for k in ("vars", "logL", "Pmove", "moved", "swap_done", "swap_chain0")})
        b.close()
    for k in outs[0]:
        assert np.array_equal(outs[0][k], outs[1][k]), k                     # same calls, same bits
    rb = outs[0]
    assert np.array_equal(ra["moved"], rb["moved"]) and np.array_equal(ra["swap_done"], rb["swap_done"]) and np.array_equal(ra["swap_chain0"], rb["swap_chain0"])
    assert 0.05 < ra["moved"].mean() < 0.95
    assert np.allclose(ra["vars"], rb["vars"], rtol=1e-9, atol=1e-12) and np.allclose(ra["logL"], rb["logL"], rtol=1e-9)
    a.close()
