"""CPU tests of the host logic: parameter layouts, sampler building blocks (against numpy) and the
multi-GPU partitioning helpers."""
import numpy as np
import pytest


def test_global_layout_matches_reference_unpacking(W):
    p = W.s_global(Nx=2000)
    pl = p.plength
    assert list(pl) == [14, 3, 14, 14, 14, 14, 6, 14, 10, 1, 2] and p.Np == 106 and pl.sum() == 106
    Nmax, lmax = pl[0], pl[1]
    Nf = pl[2:6].sum()
    iS = Nmax + lmax + Nf
    iI = iS + pl[6] + pl[7] + pl[8]
    assert p.names[Nmax] == "Visibility_l1" and p.names[iS] == "Splitting_a1" and p.names[iS + 5] == "Lorentzian_asymetry"
    assert p.names[iS + pl[6]] == "Width_l" and p.names[iI] == "Inclination"
    assert p.params[iI + 1] == 20.0 and p.params[iI + 2] == 0.0              # trunc_c, do_amp (io_ms_global.cpp:1040-1050)
    assert p.priors.shape == (106, 4) and p.priors_switch.shape == (106,)
    assert np.all(p.priors_switch[p.relax == 0] == 0)
    f0 = p.params[Nmax + lmax: Nmax + lmax + 14]
    assert np.all(np.diff(f0) > 0)                                            # lin_interpol assumes an ascending table
    assert 90 <= p.Nv <= 96


def test_local_layout(W):
    p = W.s_local()
    assert list(p.plength) == [3, 0, 1, 1, 1, 0, 6, 3, 1, 1, 2] and p.Np == 19
    iS = 3 + 0 + 3
    assert p.params[iS + 3] == pytest.approx(np.cos(np.pi / 4)) and p.params[iS + 4] == pytest.approx(np.sin(np.pi / 4))


def test_tcoefs(W):
    np.testing.assert_allclose(W.tcoefs(4), [1, 1.7, 1.7**2, 1.7**3])
    T = W.replicated_tcoefs(16, per_ensemble=8)
    np.testing.assert_allclose(T[:8], T[8:])


def test_cholesky_and_update_proposal_against_numpy(oracle):
    g = np.random.default_rng(3)
    n = 23
    A = g.standard_normal((n, n)); A = A @ A.T + n * np.eye(n)
    st, L = oracle.cholesky(A)
    assert st == 0
    np.testing.assert_allclose(L, np.linalg.cholesky(A), rtol=1e-12, atol=1e-12)
    st, _ = oracle.cholesky(-A)
    assert st == 1
    # Robbins-Monro step, MALA.cpp:292-315
    x = g.standard_normal(n); mu = g.standard_normal(n); cov = A.copy(); sigma = 0.3
    gamma, target, acc = 0.05, 0.234, 0.6
    mu2, cov2, s2 = oracle.update_proposal(x, acc, gamma, target, 1e-12, 1e14, mu, cov, sigma)
    mu_ref = mu + gamma * (x - mu)
    d = x - mu_ref
    np.testing.assert_allclose(mu2, mu_ref, rtol=1e-14)
    np.testing.assert_allclose(cov2, cov + gamma * (np.outer(d, d) - cov), rtol=1e-13)
    assert s2 == pytest.approx(sigma + gamma * (acc - target))
    # truncations p1/p2/p3 (MALA.cpp:135-176)
    mu3, cov3, s3 = oracle.update_proposal(x, 0.0, 1.0, 0.9, 1e-3, 2.0, mu, cov, 0.1)
    assert s3 == 1e-3
    assert np.linalg.norm(mu3) <= 2.0 * (1 + 1e-12) and np.linalg.norm(cov3) <= 2.0 * (1 + 1e-12)


def test_mh_and_pt_ratios(oracle):
    L = oracle.lib
    assert L.to_mh_ratio(-10.0, -12.0, -11.0) == pytest.approx(np.exp(-1.0))
    assert L.to_mh_ratio(-10.0, -5.0, -11.0) == 1.0
    assert L.to_mh_ratio(float("nan"), -5.0, -11.0) == 0.0                   # NaN logL => reject (MALA.cpp:507-509)
    assert L.to_mh_ratio(-10.0, float("-inf"), -11.0) == 0.0                 # -inf posterior => reject (:476-478)
    TA, TB = 1.0, 1.7
    LA, LB = -100.0, -70.0            # tempered log-likelihoods
    r = L.to_pt_ratio(LA, LB, TA, TB)
    assert r == pytest.approx(min(1.0, np.exp(LA * TA / TB + LB * TB / TA - LA - LB)))


def test_partitioning(tb):
    S = tb.sharding
    assert S.partition_units(64, 8, 3) == list(range(3, 64, 8))
    assert sorted(sum([S.partition_units(10, 4, r) for r in range(4)], [])) == list(range(10))
    for Nx, G in ((1000000, 8), (10, 3), (7, 8), (10000019, 4)):
        cuts = [S.bin_shard(Nx, G, r) for r in range(G)]
        assert cuts[0][0] == 0 and cuts[-1][1] == Nx
        assert all(cuts[r][1] == cuts[r + 1][0] for r in range(G - 1))
        sizes = [b - a for a, b in cuts]
        assert max(sizes) - min(sizes) <= 1
