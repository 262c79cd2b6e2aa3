"""The C++ host mirror (include/tamcmc_gpu.hpp: Model_def / MALA over the C-ABI) against the Python mirror."""
import os
import subprocess
import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "cpp", "host_mirror_check.cpp")
LIBDIR = os.path.join(ROOT, "tamcmc-c-_b200", "csrc")


def _build(tmp_path):
    exe = str(tmp_path / "host_mirror_check")
    subprocess.run(["g++", "-std=c++11", "-O1", "-I" + os.path.join(ROOT, "include"), SRC, "-o", exe,
                    "-L" + LIBDIR, "-ltamcmc_gpu", "-Wl,-rpath," + LIBDIR], check=True)
    return exe


def test_cpp_mirror_builds_and_fails_loudly_without_a_gpu(tb, tmp_path):
    tb.load_library()                                   # builds the shared library if needed
    exe = _build(tmp_path)
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present: covered by the gpu test")
    r = subprocess.run([exe], capture_output=True, text=True)
    assert r.returncode == 3 and "no CUDA device" in r.stdout


@pytest.mark.gpu
def test_cpp_mirror_equals_python_mirror(tb, W, tmp_path):
    exe = _build(tmp_path)
    out = subprocess.run([exe], capture_output=True, text=True, check=True).stdout.split("\n")
    rows = [ln.split() for ln in out if ln]
    # the same problem through gpu.py
    prob = W.gauss_problem(0, like_case=1, Nx=400)
    p0 = prob.params
    prob.x = 50.0 + 400.0 * np.arange(400) / 399
    d = (prob.x - 240.0) / 30.0
    prob.y = 1.5 + 8.0 / (1.0 + d * d) + 0.25 * ((np.arange(400) * 37) % 17 - 8).astype(np.float64) / 8.0
    prob.sigma_y = 0.25 * np.ones(400)
    prob.err_frac = np.zeros(4); prob.err_off = np.array([0.8, 1.3, 4.4, 0.0245])
    cfg = tb.MalaConfig(Nchains=3, Tcoefs=np.array([1.0, 1.7, 2.89]), Nt_learn=[10, 100000], periods_learn=[1], c0=5.0,
                        dN_mixing=2, seed=77, flags=tb.gpu.F_FIX_SWAP_BOOKKEEPING)
    md = tb.GpuModelDef(prob, cfg)
    assert [int(v) for v in rows[0][1:]] == [md.Nvars, md.Nparams, md.Nmodels, 1, 400]
    P = np.array([[p0[i] * (1.0 + 0.01 * (c + 1) * (i + 1)) for i in range(4)] for c in range(3)])
    ev = md.evaluate(P, cfg.tcoefs())
    for c in range(3):
        r = rows[1 + c]
        assert [float(r[2]), float(r[3]), float(r[4])] == [ev["logL"][c], ev["logPrior"][c], ev["logPost"][c]] and int(r[5]) == ev["status"][c]
    m = md.evaluate(P[:1], cfg.tcoefs()[:1], want_model=True)["model"][0]
    assert [float(v) for v in rows[4][1:]] == [m[0], m[200], m[399]]
    rec = tb.GpuMALA(md).execute(0, 60)
    k = 5
    for i in range(57, 60):
        for c in range(3):
            r = rows[k]; k += 1
            assert (float(r[3]), float(r[4]), int(r[5]), int(r[6])) == (rec["vars"][i, c, 0], rec["logL"][i, c], int(rec["moved"][i, c]), int(rec["swap_done"][i, 0]))
    mu, sigma, cov = md.get_proposal()
    assert [float(v) for v in rows[k][1:]] == [md.vars[0, 0], mu[0, 1], sigma[2], cov[0, 1, 1]]
    md.close()
