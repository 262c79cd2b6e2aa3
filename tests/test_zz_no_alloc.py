"""GPU test of tg_eval_noalloc (include/tamcmc_gpu.h): the batched generate_model call that allocates nothing -- the scratch of
the fused kernel lives in a file-scope __device__ array of the library.  It has no counterpart in the reference;
it is checked against the existing entry point tg_eval, through the C-ABI, at the tolerance
the parity tests use (1e-12 relative on logL), on two calls in a row (the second finds the first call's contents in the
shared scratch), and the limits of the static scratch are rejected with a clear message.

Collected last on purpose; it passes alone in a fresh process and after the other tests."""
import ctypes as C

import numpy as np
import pytest

from conftest import relerr

pytestmark = pytest.mark.gpu

TOL = 1e-12


def _problem(W, oracle, Nx=20000, seed=17):
    prob = W.small_global(3, Nx=Nx, step=0.06)
    st, m, _ = oracle.model(prob.model_case, prob.params, prob.plength, prob.x)
    assert st == 0
    prob.y = m * (-np.log(W.philox_uniform(m.size, seed)))
    return prob


def _free_bytes():
    import torch
    return torch.cuda.mem_get_info(0)[0]


def _same(r, ref):
    assert np.array_equal(r["status"], ref["status"])
    ok = ref["status"] == 0
    assert ok.any()
    assert relerr(r["logL"][ok], ref["logL"][ok]) < TOL
    fin = np.isfinite(ref["logPrior"])
    assert np.array_equal(np.isfinite(r["logPrior"]), fin)
    assert np.allclose(r["logPrior"][fin], ref["logPrior"][fin], rtol=TOL, atol=1e-11)
    assert np.allclose(r["logPost"][ok & fin], ref["logPost"][ok & fin], rtol=TOL, atol=1e-11)


def test_noalloc_equals_tg_eval_on_two_calls_in_a_row(tb, W, oracle):
    prob = _problem(W, oracle)
    Cn = 12
    md = tb.GpuModelDef(prob, tb.MalaConfig(Nchains=Cn))
    T = W.tcoefs(Cn)
    P1 = W.perturbed_params(prob, Cn, scale=0.5)
    P2 = W.perturbed_params(prob, Cn, scale=1.5)[::-1].copy()
    assert not np.array_equal(P1, P2)
    ref1, ref2 = md.evaluate(P1, T), md.evaluate(P2, T)              # the existing entry point
    assert not np.array_equal(ref1["logL"], ref2["logL"])
    # and the oracle, so that both entry points are pinned, not only equal
    orc = oracle.generate_models(prob, P1, T)
    ok = orc["status"] == 0
    assert relerr(ref1["logL"][ok], orc["logL"][ok]) < TOL
    md.evaluate_noalloc(P2, T)                                       # first use on this device: symbol address, lazy module state
    md.sync()
    _free_bytes()                                                    # (torch's own first query)
    free0 = _free_bytes()
    r1 = md.evaluate_noalloc(P1, T)                                  # two calls in a row: the second finds the first one's scratch
    r2 = md.evaluate_noalloc(P2, T)
    r1b = md.evaluate_noalloc(P1, T)
    assert _free_bytes() == free0                                    # nothing was allocated (or freed) on the device
    _same(r1, ref1)
    _same(r2, ref2)
    for k in r1:
        assert np.array_equal(r1[k], r1b[k], equal_nan=True), k     # and nothing of call 2 leaked into call 3
    # a smaller batch than the handle's tables hold, after a larger one
    r3 = md.evaluate_noalloc(P1[:5], T[:5])
    _same(r3, {k: v[:5] for k, v in ref1.items() if v is not None})
    # the sampler of the same handle is untouched by the shared scratch
    out = tb.GpuMALA(md).execute(0, 20)
    assert np.all(np.isfinite(out["logL"]))
    _same(md.evaluate_noalloc(P2, T), ref2)
    md.close()


def test_noalloc_second_handle_shares_the_scratch(tb, W, oracle):
    """Two handles (different spectra, sizes) used one after the other on the same device share the array."""
    pa, pb = _problem(W, oracle, Nx=20000, seed=3), _problem(W, oracle, Nx=7000, seed=4)
    a, b = tb.GpuModelDef(pa, tb.MalaConfig(Nchains=6)), tb.GpuModelDef(pb, tb.MalaConfig(Nchains=9))
    Pa, Pb = W.perturbed_params(pa, 6, scale=0.7), W.perturbed_params(pb, 9, scale=0.7)
    Ta, Tb = W.tcoefs(6), W.tcoefs(9)
    ra, rb = a.evaluate(Pa, Ta), b.evaluate(Pb, Tb)
    for _ in range(2):
        _same(a.evaluate_noalloc(Pa, Ta), ra)
        _same(b.evaluate_noalloc(Pb, Tb), rb)
    a.close(); b.close()


def test_noalloc_rejects_sizes_beyond_the_static_scratch(tb, W, oracle):
    prob = _problem(W, oracle, Nx=4000)
    md = tb.GpuModelDef(prob, tb.MalaConfig(Nchains=4))
    P, T = W.perturbed_params(prob, 5), np.ones(5)
    with pytest.raises(tb.TgError) as e:                             # more vectors than the handle's tables hold
        md.evaluate_noalloc(P, T)
    assert "never allocates" in str(e.value)
    lib = md.lib.lib
    assert lib.tg_eval_noalloc(md.h, None, 4, T.ctypes.data, None, None, None, None) == -1
    md.close()
    # B x work items beyond TG_NOALLOC_MAX_PAIRS: one work item per 1024-bin sub-tile (TG_F_NO_TILE_ORDER = 2), 1024 chains
    big = W.small_global(3, Nx=300000, step=0.01)
    big.y = np.ones(big.x.size)
    Cn = 1024
    mb = tb.GpuModelDef(big, tb.MalaConfig(Nchains=Cn, chains_per_ensemble=8, flags=2))
    Pb, Tb = np.tile(big.params, (Cn, 1)), np.ones(Cn)
    with pytest.raises(tb.TgError) as e:
        mb.evaluate_noalloc(Pb, Tb)
    assert "TG_NOALLOC_MAX_PAIRS" in str(e.value)
    r = mb.evaluate_noalloc(Pb[:800], Tb[:800])                      # 800 x 293 <= 262144: accepted
    assert np.all(r["status"] == 0) and np.allclose(r["logL"], r["logL"][0], rtol=1e-13, atol=0)
    assert relerr(r["logL"][:8], mb.evaluate(Pb[:8], Tb[:8])["logL"]) < TOL
    mb.close()
