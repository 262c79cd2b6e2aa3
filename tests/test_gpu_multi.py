"""GPU tests of the batched configurations of BASELINE.json that the reference cannot run in one process:

* configs[3]: 1024 chains in one handle = 128 replicated tempered ensembles of 8 (the reference caps Nchains at 24,
  MALA.cpp:565; swaps stay inside an ensemble, MALA.cpp:381-445);
* configs[4]: several stars per GPU as ONE launch group (tg_batch_*; the reference runs stars as independent processes,
  scripts/slurm/job_array.sh:8,24, main.cpp:111-127).

Everything is compared bit for bit: a batched run must be the same trajectory as the independent runs it replaces.  The
Metropolis / tempering decisions are also replayed on the host from a numpy Philox4x32-10 (the device's generator,
tg_internal.cuh) and the recorded likelihoods, with the reference's stale-prior bookkeeping (quirk Q8, MALA.cpp:417,:428)
written out."""
import numpy as np
import pytest

from conftest import relerr

pytestmark = pytest.mark.gpu

TAG_MH, TAG_PT = 0x4D485F55, 0x50545F55


def philox4x32_10(c0, c1, c2, c3, k0, k1):
    """Philox4x32-10 (Salmon et al. 2011) on uint64 numpy arrays holding 32-bit words."""
    M0, M1, m32 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57), np.uint64(0xFFFFFFFF)
    c0, c1, c2, c3 = [np.asarray(v, dtype=np.uint64) & m32 for v in (c0, c1, c2, c3)]
    k0, k1 = np.uint64(k0) & m32, np.uint64(k1) & m32
    for _ in range(10):
        p0, p1 = M0 * c0, M1 * c2
        hi0, lo0, hi1, lo1 = p0 >> np.uint64(32), p0 & m32, p1 >> np.uint64(32), p1 & m32
        c0, c1, c2, c3 = hi1 ^ c1 ^ k0, lo1, hi0 ^ c3 ^ k1, lo0
        k0, k1 = (k0 + np.uint64(0x9E3779B9)) & m32, (k1 + np.uint64(0xBB67AE85)) & m32
    return c0, c1, c2, c3


def u01(hi, lo):
    return ((hi >> np.uint64(5)).astype(np.float64) * 67108864.0 + (lo >> np.uint64(6)).astype(np.float64) + 0.5) / 9007199254740992.0


def draw(seed, tag, it, idx):
    """(uniform, third word) of the counter (it, idx) under the key (seed, tag) -- tg_sampler.cuh."""
    it = np.asarray(it, dtype=np.uint64)
    r = philox4x32_10(it & np.uint64(0xFFFFFFFF), it >> np.uint64(32), np.asarray(idx, dtype=np.uint64), np.zeros_like(it),
                      (seed & 0xFFFFFFFF) ^ tag, seed >> 32)
    return u01(r[0], r[1]), r[2]


def _small(W, oracle, mc=3, Nx=1000, seed=8, **kw):
    """A reduced global fit on <= 1024 bins: one work item, so the likelihood sum is the same bits however many chains share a launch."""
    prob = W.small_global(mc, Nx=Nx, step=0.25 if Nx <= 1024 else 0.12, **kw)
    st, m, _ = oracle.model(prob.model_case, prob.params, prob.plength, prob.x)
    assert st == 0
    prob.y = m * (-np.log(W.philox_uniform(m.size, seed)))
    return prob


def _unswap(out, Ce):
    """moved / Pmove as the Metropolis step left them (the swap exchanges them with the chains, MALA.cpp:419-431)."""
    moved, Pmove = out["moved"].copy(), out["Pmove"].copy()
    n, E = out["swap_done"].shape
    for i, e in zip(*np.nonzero(out["swap_done"])):
        a = e * Ce + out["swap_chain0"][i, e]
        moved[i, [a, a + 1]] = moved[i, [a + 1, a]]
        Pmove[i, [a, a + 1]] = Pmove[i, [a + 1, a]]
    return moved, Pmove


def test_b1024_parity_at_full_size(tb, W, oracle):
    """configs[3]: 1024 parameter vectors on the 1M-bin S-global spectrum in ONE call against the oracle (1e-12)."""
    prob = W.s_global()
    st, m, _ = oracle.model(prob.model_case, prob.params, prob.plength, prob.x)
    prob.y = m * (-np.log(W.philox_uniform(m.size, 12345)))
    B = 1024
    md = tb.GpuModelDef(prob, tb.MalaConfig(Nchains=16, chains_per_ensemble=8))
    P, T = W.perturbed_params(prob, B, scale=0.5), W.replicated_tcoefs(B, 8)
    r = md.evaluate(P, T)
    ref = oracle.generate_models(prob, P, T)
    assert np.array_equal(r["status"] & 3, ref["status"] & 3)
    ok = ref["status"] == 0
    assert ok.sum() > 1000
    assert relerr(r["logL"][ok], ref["logL"][ok]) < 1e-12
    fin = np.isfinite(ref["logPrior"])
    assert np.array_equal(np.isfinite(r["logPrior"]), fin) and np.allclose(r["logPrior"][fin], ref["logPrior"][fin], rtol=1e-12, atol=1e-11)
    L, _ = md.last_work(B)
    assert np.array_equal(L[ok], ref["L"][ok])               # truncation windows identical, bin for bin, for all 1024
    md.close()


@pytest.mark.parametrize("E", [2, 16, 128])
def test_multi_ensemble_sampler_records(tb, W, oracle, E):
    """chains_per_ensemble = 8, E ensembles in one handle: swap cadence and records per ensemble, chains that neither moved nor
    were swapped keep their variables, every Metropolis and tempering decision equals its host replay, and the reference's
    stale-prior bookkeeping (Q8) is what the records show."""
    Ce, C, N, mix, seed = 8, 8 * E, 120, 3, 4242
    prob = _small(W, oracle)
    cfg = tb.MalaConfig(Nchains=C, chains_per_ensemble=Ce, Nt_learn=[10, 10**6], periods_learn=[1], c0=2.0, dN_mixing=mix, seed=seed)
    md = tb.GpuModelDef(prob, cfg)
    assert md.Nensembles == E
    T = cfg.tcoefs()
    assert np.array_equal(T.reshape(E, Ce), np.tile(1.7 ** np.arange(Ce), (E, 1)))
    out = tb.GpuMALA(md).execute(0, N)
    it = np.arange(N)
    att = ((it % mix == 0) & (it != 0)).astype(np.int32)
    assert np.array_equal(out["swap_attempt"], np.tile(att[:, None], (1, E)))
    assert np.all(out["swap_chain0"][att == 0] == -1) and np.all(out["swap_done"][att == 0] == 0)
    ch0 = out["swap_chain0"][att == 1]
    assert ch0.min() >= 0 and ch0.max() <= Ce - 2
    # tempering pair of (iteration, ensemble) = third Philox word mod (Ce - 1)
    ee, ii = np.meshgrid(np.arange(E), it, indexing="xy")
    u_pt, w2 = draw(seed, TAG_PT, ii, ee)
    assert np.array_equal(out["swap_chain0"][att == 1], (w2 % np.uint64(Ce - 1)).astype(np.int32)[att == 1])
    assert 0.02 < out["swap_done"][att == 1].mean() < 1.0
    # Metropolis decisions: moved == (u <= Pmove) with the device's uniform of (iteration, chain)
    moved, Pmove = _unswap(out, Ce)
    cc, ii2 = np.meshgrid(np.arange(C), it, indexing="xy")
    u_mh, _ = draw(seed, TAG_MH, ii2, cc)
    assert np.array_equal(moved.astype(bool), u_mh <= Pmove)
    v, lL, lPr, lPo = out["vars"], out["logL"], out["logPrior"], out["logPost"]
    for i in range(1, N):
        sw = np.zeros(C, dtype=bool)
        for e in range(E):
            a = e * Ce + out["swap_chain0"][i, e]
            if out["swap_done"][i, e]:
                sw[[a, a + 1]] = True
                # rows exchanged inside the ensemble, never across its boundary
                assert a // Ce == e and (a + 1) // Ce == e
                # Q8 (default): logPosterior[B] = logL_A_TB + logPrior[A] AFTER logPrior[A] became B's prior (MALA.cpp:417,:428)
                assert lPo[i, a] == lL[i, a] + lPr[i, a]
                assert lPo[i, a + 1] == lL[i, a + 1] + lPr[i, a]
                # the decision itself: u <= rT, with the likelihoods before the exchange recovered from the exchanged ones
                LB, LA = lL[i, a] * T[a] / T[a + 1], lL[i, a + 1] * T[a + 1] / T[a]
                rT = min(1.0, np.exp(LA * T[a] / T[a + 1] + LB * T[a + 1] / T[a] - LA - LB))
                assert u_pt[i, e] <= rT * (1 + 1e-9) and abs(out["swap_Pswap"][i, e] - rT) <= 1e-9 * rT
            elif att[i]:
                LA, LB = lL[i, a], lL[i, a + 1]
                rT = min(1.0, np.exp(LA * T[a] / T[a + 1] + LB * T[a + 1] / T[a] - LA - LB))
                assert u_pt[i, e] > rT * (1 - 1e-9) and out["swap_Pswap"][i, e] == 0.0
        keep = ~sw & ~out["moved"][i].astype(bool)
        assert np.array_equal(v[i][keep], v[i - 1][keep])
        # an accepted move that was not exchanged: Pmove = min(1, exp(logPost' - logPost)) with the logPost the chain CARRIED
        # (stale after a swap) -- the next decision of a swapped chain is where Q8 acts
        mv = ~sw & out["moved"][i].astype(bool)
        pm = np.minimum(1.0, np.exp(lPo[i][mv] - lPo[i - 1][mv]))
        assert np.allclose(out["Pmove"][i][mv], pm, rtol=1e-11, atol=0)
    assert np.array_equal(md.vars, v[-1])
    md.close()


def test_ensemble_of_a_batched_handle_equals_a_single_ensemble_handle(tb, W, oracle):
    """Ensemble e of a 16-ensemble handle is, bit for bit, the run of a single-ensemble handle whose Philox counters are
    offset by (e*8 chains, e ensembles): replicated ensembles are independent samplers sharing nothing but launches."""
    Ce, E, N = 8, 16, 150
    prob = _small(W, oracle, seed=21)
    mk = lambda C, **kw: tb.MalaConfig(Nchains=C, chains_per_ensemble=Ce, Nt_learn=[5, 10**6], periods_learn=[1], c0=2.0, dN_mixing=1, seed=31337, **kw)
    big = tb.GpuModelDef(prob, mk(Ce * E)); rb = tb.GpuMALA(big).execute(0, N)
    assert rb["swap_done"].sum() > 20 and rb["moved"].mean() > 0.02
    for e in (0, 5, 15):
        one = tb.GpuModelDef(prob, mk(Ce, philox_chain0=e * Ce, philox_ensemble0=e)); ro = tb.GpuMALA(one).execute(0, N)
        sl = slice(e * Ce, (e + 1) * Ce)
        for k in ("vars", "logL", "logPrior", "logPost", "Pmove", "moved", "status"):
            assert np.array_equal(rb[k][:, sl], ro[k], equal_nan=rb[k].dtype.kind == "f"), (e, k)
        for k in ("swap_attempt", "swap_chain0", "swap_Pswap", "swap_done"):
            assert np.array_equal(rb[k][:, e], ro[k][:, 0]), (e, k)
        mb, sb, cb = big.get_proposal(); mo, so, co = one.get_proposal()
        assert np.array_equal(mb[sl], mo) and np.array_equal(sb[sl], so) and np.array_equal(cb[sl], co)
        one.close()
    big.close()


def test_multi_ensemble_resume(tb, W, oracle):
    """Resume of a multi-ensemble handle from (vars, mu, sigma, covarmat), the content of the restore triplet.  Like the
    reference (model_def.cpp:104-128) the restore re-evaluates the current points, so the likelihood a swapped chain carried
    (L*T_B/T_A, MALA.cpp:405-410) comes back as -p*sum/T_A: a last-bit difference that enters sigma through Pmove.  Every
    decision is the same; the continuous records agree to rounding.  (With quirk Q8 replicated the resumed run would also
    lose the stale logPosterior of swapped chains, as the reference does -- hence the flag.)"""
    Ce, E = 8, 4
    prob = _small(W, oracle, seed=5)
    mk = lambda: tb.MalaConfig(Nchains=Ce * E, chains_per_ensemble=Ce, Nt_learn=[5, 10**6], periods_learn=[1], c0=2.0, dN_mixing=2, seed=9,
                               flags=tb.gpu.F_FIX_SWAP_BOOKKEEPING)
    a = tb.GpuModelDef(prob, mk()); ra = tb.GpuMALA(a).execute(0, 90)
    b = tb.GpuModelDef(prob, mk()); rb = tb.GpuMALA(b).execute(0, 50)
    for k in ra:
        assert np.array_equal(ra[k][:50], rb[k], equal_nan=ra[k].dtype.kind == "f"), k      # same seed, same bits
    mu, sg, cov = b.get_proposal()
    c = tb.GpuModelDef(prob, mk()); c.set_state(vars_=b.vars, mu=mu, sigma=sg, covarmat=cov)
    rc = tb.GpuMALA(c).execute(50, 40)
    assert ra["swap_done"][50:].sum() > 3 and ra["moved"][50:].mean() > 0.02
    for k in ("moved", "swap_done", "swap_chain0", "swap_attempt", "status"):
        assert np.array_equal(ra[k][50:], rc[k]), k
    for k in ("vars", "logL", "logPrior", "logPost", "Pmove"):
        assert np.allclose(ra[k][50:], rc[k], rtol=1e-9, atol=1e-9), k
    for m in (a, b, c):
        m.close()


@pytest.mark.parametrize("norders,frac", [(4, 0.5), (14, 0.5), (18, 0.5), (4, 1e-4)])
def test_adaptation_norm_clamps_p2_p3(tb, W, norders, frac):
    """A1 below ||mu|| and ||Sigma||_F: p3 rescales mu and p2 rescales Sigma at every step (MALA.cpp:155-176), p1 caps sigma
    (:135-153).  (mu, Sigma, sigma) after N adaptive steps == update_proposal (MALA.cpp:292-315) replayed in numpy WITH the
    clamps, on the register (Nv <= 96) and the shared-memory factorisation paths."""
    prob = W.s_global(Nx=60000, step=0.08, norders=norders, numax=2500.0, dnu=120.0) if norders > 4 else W.small_global(3, Nx=3000, step=0.1)
    prob.y = np.abs(np.random.default_rng(3).standard_normal(prob.x.size)) + 0.1
    err = prob.init_errors()
    x0 = prob.params[prob.index_to_relax]
    A1 = frac * min(np.linalg.norm(x0), np.sqrt(np.sum(err**4)))
    n0, N = 3, 30
    cfg = tb.MalaConfig(Nchains=3, Nt_learn=[n0, 10**6], periods_learn=[1], c0=2.0, seed=11, dN_mixing=0, A1=float(A1))
    md = tb.GpuModelDef(prob, cfg)
    mu, sigma, cov = [a.copy() for a in md.get_proposal()]
    out = tb.GpuMALA(md).execute(0, N)
    hits = np.zeros(3, dtype=int)
    moved, Pmove = out["moved"], out["Pmove"]
    for i in range(n0, N):
        gamma = cfg.c0 / (1.0 + i)
        for c in range(3):
            x = out["vars"][i, c]
            m3 = mu[c] + gamma * (x - mu[c])
            n3 = np.linalg.norm(m3)
            if n3 > A1:
                m3 = m3 * (A1 / n3); hits[0] += 1
            mu[c] = m3
            d = x - mu[c]
            s2 = cov[c] + gamma * (np.outer(d, d) - cov[c])
            n2 = np.linalg.norm(s2)
            if n2 > A1:
                s2 = s2 * (A1 / n2); hits[1] += 1
            cov[c] = s2
            s1 = sigma[c] + gamma * (Pmove[i, c] - cfg.target_acceptance)
            if s1 > A1:
                s1 = A1; hits[2] += 1
            sigma[c] = max(s1, cfg.epsilon1)
    assert hits[0] == 3 * (N - n0) and hits[1] == 3 * (N - n0), hits          # both norm clamps fired at every step
    if frac < 1e-3:
        assert hits[2] > 0                                                     # and sigma sat at A1
    mu_g, sigma_g, cov_g = md.get_proposal()
    assert np.allclose(mu_g, mu, rtol=1e-11, atol=0)
    assert np.allclose(sigma_g, sigma, rtol=1e-12, atol=0)
    assert abs(np.linalg.norm(mu_g[0]) / A1 - 1) < 1e-12 and abs(np.linalg.norm(cov_g[0]) / A1 - 1) < 1e-12
    scale = np.sqrt(np.einsum("cii->ci", np.abs(cov))) + 1e-300
    assert np.max(np.abs(cov_g - cov) / (scale[:, :, None] * scale[:, None, :])) < 1e-8
    assert np.array_equal(cov_g, np.swapaxes(cov_g, 1, 2))
    md.close()


def _three_stars(W, oracle):
    a = _small(W, oracle, mc=3, Nx=5000, seed=1)                       # several work items, global model
    b = _small(W, oracle, mc=2, Nx=2500, seed=2, norders=5)            # another model case, other Np / Nv
    c = W.s_local()                                                    # local model (io_local priors), 10k bins
    st, m, _ = oracle.model(c.model_case, c.params, c.plength, c.x)
    c.y = m * (-np.log(W.philox_uniform(m.size, 3)))
    return [a, b, c]


@pytest.mark.parametrize("use_graph", [1, 0])
def test_launch_group_of_stars_equals_the_individual_handles(tb, W, oracle, use_graph):
    """configs[4]: three different stars (different spectra, model cases, Np, Nv, chain counts) advanced as one launch group
    give, star by star and bit for bit, the records and the proposal state of tg_run on each handle alone -- over several
    calls (so the work items are rebuilt in between) and with the steps replayed from the CUDA graph or launched one by one."""
    stars = _three_stars(W, oracle)
    cfgs = [tb.MalaConfig(Nchains=c, Nt_learn=[8, 10**6], periods_learn=[1], c0=2.0, dN_mixing=1, seed=100 + k) for k, c in enumerate((10, 6, 10))]
    chunks = (70, 37, 80)

    def fresh():
        ms = [tb.GpuModelDef(p, c) for p, c in zip(stars, cfgs)]
        for m in ms:
            m.set_tuning(use_graph=use_graph)
        return ms
    solo = fresh()
    ref = [[] for _ in solo]
    for k, m in enumerate(solo):
        i0 = 0
        for n in chunks:
            ref[k].append(tb.GpuMALA(m).execute(i0, n)); i0 += n
    grp = fresh()
    batch = tb.GpuBatch(grp)
    ns, ct, nu = batch.sizes()
    assert ns == 3 and ct == 26 and nu >= 3
    i0 = 0
    for j, n in enumerate(chunks):
        outs = batch.execute(i0, n); i0 += n
        for k in range(3):
            for key in outs[k]:
                assert np.array_equal(outs[k][key], ref[k][j][key], equal_nan=outs[k][key].dtype.kind == "f"), (k, j, key)
    for k in range(3):
        for a, b in zip(grp[k].get_proposal(), solo[k].get_proposal()):
            assert np.array_equal(a, b)
        assert ref[k][-1]["moved"].mean() > 0.01
    # a handle of the group is still usable on its own afterwards, and continues the same trajectory
    r1 = tb.GpuMALA(grp[0]).execute(i0, 20); r2 = tb.GpuMALA(solo[0]).execute(i0, 20)
    assert np.array_equal(r1["vars"], r2["vars"])
    batch.close()
    for m in grp + solo:
        m.close()


def test_launch_group_outliving_its_stars_is_destroyed_safely(tb, W, oracle):
    """The stars must outlive the group (include/tamcmc_gpu.h), but a binding may tear them down in any order at interpreter
    exit: destroying the group after its stars must not touch them (it gives the stars their own unit shapes back only while
    they exist).  And a star that leaves a destroyed group runs on alone, on its own trajectory."""
    stars = _three_stars(W, oracle)[:2]
    mk = lambda: [tb.GpuModelDef(p, tb.MalaConfig(Nchains=6, Nt_learn=[8, 10**6], periods_learn=[1], c0=2.0, seed=7 + k)) for k, p in enumerate(stars)]
    ms = mk()
    batch = tb.GpuBatch(ms)
    batch.execute(0, 20, record=False)
    for m in ms:
        m.close()                      # stars first ...
    batch.close()                      # ... then the group
    ms, solo = mk(), mk()
    batch = tb.GpuBatch(ms)
    a = batch.execute(0, 30)
    batch.close()                      # the group first: its stars get their own unit shapes back
    b = tb.GpuMALA(ms[0]).execute(30, 30)
    c0, c1 = tb.GpuMALA(solo[0]).execute(0, 30), tb.GpuMALA(solo[0]).execute(30, 30)
    assert np.array_equal(a[0]["vars"], c0["vars"]) and np.array_equal(b["vars"], c1["vars"])
    for m in ms + solo:
        m.close()


def test_launch_group_rejects_bad_input(tb, W, oracle):
    prob = _small(W, oracle)
    m = tb.GpuModelDef(prob, tb.MalaConfig(Nchains=2))
    with pytest.raises(tb.TgError):
        tb.GpuBatch([m, m])
    m.close()
