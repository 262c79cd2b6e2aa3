"""tools/getmodel_gpu.cpp -- the reference's getmodel tool (tools/getmodel.cpp) on the CUDA path: data file +
params file + model name -> model spectrum, checked against the model spectra of the unmodified reference
build stored in tests/golden/ (all live model cases)."""
import os
import subprocess
import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIBDIR = os.path.join(ROOT, "tamcmc-c-_b200", "csrc")
NAMES = ["Test_Gaussian", "model_Harvey_Gaussian", "model_MS_Global_a1etaa3_HarveyLike", "model_MS_Global_a1etaa3_HarveyLike_Classic",
         "model_MS_Global_a1etaa3_Harvey1985", "model_MS_Global_a1acta3_Harvey1985", "model_MS_Global_a1l_etaa3_HarveyLike",
         "model_MS_Global_a1n_etaa3_HarveyLike", "model_MS_Global_a1nl_etaa3_HarveyLike", "model_MS_Global_a1etaa3_AppWidth_HarveyLike_v1",
         "model_MS_Global_a1etaa3_AppWidth_HarveyLike_v2", "model_MS_local_basic", "model_MS_Global_a1etaa3_HarveyLike_Classic_v2",
         "model_MS_Global_a1etaa3_HarveyLike_Classic_v3", "model_MS_local_Hnlm"]


def _build(tb, tmp_path):
    tb.load_library()
    exe = str(tmp_path / "getmodel_gpu")
    subprocess.run(["g++", "-std=c++11", "-O1", "-I" + os.path.join(ROOT, "include"), os.path.join(ROOT, "tools", "getmodel_gpu.cpp"),
                    "-o", exe, "-L" + LIBDIR, "-ltamcmc_gpu", "-Wl,-rpath," + LIBDIR], check=True)
    return exe


def _write_inputs(tmp_path, gp, tag):
    p = gp.prob
    fdata, fpar = str(tmp_path / (tag + ".data")), str(tmp_path / (tag + ".params"))
    with open(fdata, "w") as f:                       # the layout of test/inputs/TF_3443483_local-v3.data
        f.write("# File written by tests/test_getmodel_tool.py\n# freq_range=[ %f , %f ]\n" % (p.x[0], p.x[-1]))
        f.write("!            frequency               power\n*            (microHz)     (ppm^2/microHz)\n")
        for a, b in zip(p.x, p.y):
            f.write("%.17g %.17g\n" % (a, b))
    with open(fpar, "w") as f:
        f.write("# plength, then one parameter vector per line\n")
        f.write(" ".join(str(int(v)) for v in p.plength) + "\n")
        f.write(" ".join("%.17g" % v for v in p.params) + "\n")
    return fdata, fpar


def test_tool_builds_and_fails_loudly_without_a_gpu(tb, golden_problems, tmp_path):
    exe = _build(tb, tmp_path)
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present: covered by the gpu test")
    name, gp = sorted(golden_problems.items())[0]
    fdata, fpar = _write_inputs(tmp_path, gp, "a")
    r = subprocess.run([exe, fdata, fpar, NAMES[gp.prob.model_case], str(tmp_path / "out.ascii")], capture_output=True, text=True)
    assert r.returncode == 3 and "no CUDA device" in r.stdout


@pytest.mark.gpu
def test_getmodel_matches_the_reference_spectra(tb, golden_problems, tmp_path):
    exe = _build(tb, tmp_path)
    seen = set()
    for k, (name, gp) in enumerate(sorted(golden_problems.items())):
        p = gp.prob
        fdata, fpar = _write_inputs(tmp_path, gp, "g%d" % k)
        fout = str(tmp_path / ("g%d.ascii" % k))
        r = subprocess.run([exe, fdata, fpar, NAMES[p.model_case], fout], capture_output=True, text=True)
        assert r.returncode == 0 and "successfully written" in r.stdout, (name, r.stdout)
        a = np.loadtxt(fout)
        assert a.shape == (p.x.size, 3)
        assert np.allclose(a[:, 0], p.x, rtol=1e-11) and np.allclose(a[:, 1], p.y, rtol=1e-11)
        rel = np.max(np.abs(a[:, 2] - gp.ref_model) / np.abs(gp.ref_model))
        assert rel < 1e-10, (name, rel)               # 12 printed digits; the CUDA path itself is at ~1e-15
        seen.add(p.model_case)
    assert seen >= {0, 1, 2, 3, 6, 7, 8, 9, 10, 11, 12, 13, 14}
