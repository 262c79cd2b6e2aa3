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


def test_background_moment_identity_and_its_check():
    """The identity behind the fused kernel's background-only path (tg_eval.cuh, tg_moment_tiles), restated in numpy:
    with c the Chebyshev coefficients of 1/b on a tile, sum_i y_i / b(x_i) = sum_k c_k * sum_i y_i T_k(t_i) to rounding
    when b is resolved by degree 8 -- and when it is not, the check points (extrema of T_9) see an error of the size of
    the true one, so the kernel falls back to the bin-by-bin sum."""
    n, nb = 9, 4096
    tn = np.cos(np.pi * (np.arange(n) + 0.5) / n)
    Wm = (2.0 / n) * np.cos(np.outer(np.arange(n), np.pi * (np.arange(n) + 0.5) / n)); Wm[0] *= 0.5
    tc = np.cos(np.pi * np.array([0, 2, 4, 6, 9]) / n)
    cheb = lambda c, t: c @ np.cos(np.outer(np.arange(n), np.arccos(np.clip(t, -1, 1))))
    b = lambda x: 0.134 + 11.05 / (1 + (0.04967 * x) ** 4) + 0.9357 / (1 + (0.0013516 * x) ** 2)
    g = np.random.default_rng(3)
    for x0, resolved in ((2000.0, True), (600.0, True), (10.0, False)):
        x = x0 + 0.0085 * np.arange(nb)
        y = b(x) * g.exponential(size=nb)
        xm, xh = 0.5 * (x[0] + x[-1]), 0.5 * (x[-1] - x[0])
        t = (x - xm) / xh
        T = np.cos(np.outer(np.arange(n), np.arccos(np.clip(t, -1, 1))))
        YT, ST = T @ y, T.sum(axis=1)
        bn = b(xm + xh * tn)
        cr, cl = Wm @ (1.0 / bn), Wm @ np.log(bn)
        s1, s2 = np.sum(y / b(x)), np.sum(np.log(b(x)))
        e1, e2 = abs(cr @ YT - s1) / abs(s1), abs(cl @ ST - s2) / abs(s2)
        bc = b(xm + xh * tc)
        chk = max(np.max(np.abs(cheb(cr, tc) - 1.0 / bc) * bc), np.max(np.abs(cheb(cl, tc) - np.log(bc)) / np.maximum(1.0, np.abs(np.log(bc)))))
        if resolved:
            assert e1 < 5e-15 and e2 < 5e-15 and chk < 2e-14, (x0, e1, e2, chk)
        else:
            assert chk > 1e-10 and chk > 0.05 * max(e1, e2), (x0, e1, e2, chk)      # the check fires: this tile goes bin by bin
