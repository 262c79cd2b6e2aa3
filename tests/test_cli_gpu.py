"""`cpptamcmc execute` on the CUDA path (tools/cpptamcmc_gpu.cpp over csrc/tg_frontend.cpp): a synthetic star written in the
reference's file formats is processed through Burn-in -> Learning -> Acquire with the restore chain of config_presets.cfg,
and the reference's output files are checked.  Also: the problem the front end builds, evaluated on the GPU, equals the
oracle (the parity test of the `.model/.data/.cfg` row)."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIBDIR = os.path.join(ROOT, "tamcmc-c-_b200", "csrc")
pytestmark = pytest.mark.gpu

MODES = [(0, 1030.0, 1029.0, 1031.0, 1.0, 10.0), (1, 1055.0, 1054.0, 1056.0, 1.0, 15.0), (2, 1024.0, 1023.0, 1025.0, 1.0, 5.0),
         (0, 1130.0, 1129.0, 1131.0, 1.2, 9.0), (1, 1155.0, 1154.0, 1156.0, 1.2, 12.0), (2, 1124.0, 1123.0, 1125.0, 1.2, 4.0)]
RANGES = [(1010.0, 1070.0), (1110.0, 1170.0)]


def _build_cli(tmp_path):
    exe = str(tmp_path / "cpptamcmc_gpu")
    subprocess.run(["g++", "-std=c++11", "-O1", "-I" + os.path.join(ROOT, "include"), os.path.join(ROOT, "tools", "cpptamcmc_gpu.cpp"), "-o", exe,
                    "-L" + LIBDIR, "-ltamcmc_gpu", "-Wl,-rpath," + LIBDIR], check=True)
    return exe


def _star(tb, tmp_path, nsamples=(600, 600, 800), last_phase=2):
    """Working directory + a two-slice local star whose spectrum is drawn from the model the .model file describes."""
    ct = tb.config_tree
    md, out, wd = str(tmp_path / "in"), str(tmp_path / "out"), str(tmp_path / "wd")
    os.makedirs(md, exist_ok=True)
    x = 1000.0 + 0.01 * np.arange(20000)
    ct.write_model_file(os.path.join(md, "star.model"), MODES, RANGES, "model_MS_local_basic", Dnu=100.0, N0=1.0, trunc_c=20.0, a1=1.0, inclination=45.0)
    ct.write_data_file(os.path.join(md, "star.data"), x, np.ones_like(x))
    ct.write_config_tree(wd, md, out, ["star"], prior_fct="io_local", Nsamples=nsamples, c0=(1.8, 1.7, 0.0), restore=(0, 1, 2), last_phase=last_phase,
                         Nbuffer=500, Nt_learn=(100, 300, 100000))
    y = np.ones_like(x)
    fe = tb.frontend.FrontEnd(wd)
    truth = {}
    for sl in range(2):
        fe.setup(0, sl, 0)
        prob, cfg = fe.problem()
        mdl = tb.GpuModelDef(prob, tb.MalaConfig(Nchains=1))
        m = mdl.evaluate(prob.params[None, :], np.ones(1), want_model=True)["model"][0]
        mdl.close()
        i0 = int(np.searchsorted(x, prob.x[0] - 1e-9))
        y[i0:i0 + m.size] = m * (-np.log(tb.workloads.philox_uniform(m.size, 77 + sl)))
        truth[sl] = prob
    fe.close()
    ct.write_data_file(os.path.join(md, "star.data"), x, y)
    return wd, out, truth


def _hdr(path):
    d = {}
    for ln in open(path):
        if ln.startswith("!"):
            k, v = ln[1:].split("=", 1)
            d[k.strip()] = v.strip()
    return d


def test_front_end_problem_on_the_gpu_equals_the_oracle(tb, oracle, tmp_path):
    wd, _, _ = _star(tb, tmp_path)
    fe = tb.frontend.FrontEnd(wd)
    for sl in range(2):
        run = fe.setup(0, sl, 0)
        prob, cfg = fe.problem()
        assert run["model_fct_name"] == "model_MS_local_basic" and prob.model_case == 11 and prob.prior_case == 3
        Cn = 4
        P, T = tb.workloads.perturbed_params(prob, Cn, scale=0.05), tb.workloads.tcoefs(Cn)
        P[0] = prob.params
        mdl = tb.GpuModelDef(prob, tb.MalaConfig(Nchains=Cn))
        r = mdl.evaluate(P, T, want_model=True)
        ref = oracle.generate_models(prob, P, T, want_model=True)
        mdl.close()
        rel = lambda a, b: float(np.max(np.abs(a - b) / np.maximum(np.abs(b), 1e-300)))
        assert rel(r["model"], ref["model"]) < 1e-12 and rel(r["logL"], ref["logL"]) < 1e-12        # north_star: 1e-10
        assert np.allclose(r["logPrior"], ref["logPrior"], rtol=1e-12, atol=1e-10)
        assert np.isfinite(r["logPost"][0])
    fe.close()


def test_cli_three_phases_write_the_reference_files(tb, tmp_path):
    exe = _build_cli(tmp_path)
    wd, out, truth = _star(tb, tmp_path)
    env = dict(os.environ, TAMCMC_SEED="4242")
    r = subprocess.run([exe, "execute", "1", "1", "1", "1", "2"], cwd=wd, capture_output=True, text=True, env=env, timeout=600)
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-500:]
    assert "Phase 3/3" in r.stdout and "Slice 2/2" not in r.stdout          # `1 2` = slice 1 only (main.cpp:116-126)
    o, rs, dg = (os.path.join(out, "star", d) for d in ("outputs", "restore", "diags"))
    prob = truth[0]
    Nv, Cn = prob.Nv, 10
    for core, ns in (("B", 600), ("L", 600), ("A", 800)):
        root = os.path.join(o, "star_1_%s_" % core)
        h = _hdr(root + "params.hdr")
        assert int(h["Nchains"]) == Cn and int(h["Nvars"]) == Nv and int(h["Nsamples"]) == ns
        assert [int(v) for v in h["relax"].split()] == list(prob.relax) and [int(v) for v in h["plength"].split()] == list(prob.plength)
        assert h["variable_names"].split()[0] == "Height_l"
        for m in range(Cn):
            assert os.path.getsize(root + "params_chain-%d.bin" % m) == ns * Nv * 8
        stat = np.fromfile(root + "stat_criteria.bin").reshape(ns, 3 * Cn)
        assert np.all(np.isfinite(stat))
        if core == "L":                      # no swaps in this phase: logPosterior = logLikelihood + logPrior for every chain (model_def.cpp:365)
            assert np.allclose(stat[:, 2 * Cn:], stat[:, :Cn] + stat[:, Cn:2 * Cn], rtol=1e-12)
        assert os.path.getsize(root + "parallel_tempering.bin") == ns * 14
        acc = [ln for ln in open(root + "acceptance.txt") if not ln.startswith(("#", "!"))]
        assert len(acc) == -(-ns // 500) and len(acc[0].split()) == 1 + Cn
        ev = [ln for ln in open(os.path.join(dg, "star_1_%s_evidence.txt" % core)) if not ln.startswith(("#", "!"))]
        assert len(ev) == -(-ns // 500) and np.isfinite(float(ev[-1].split()[-1]))
        rst = tb.outputs.read_restore(os.path.join(rs, "star_1_restore_%s_" % core))
        assert rst["iteration"] == ns and rst["vars"].shape == (Cn, Nv) and rst["covarmat"].shape == (Cn, Nv, Nv)
    # Learning never mixes the chains (dN_mixing = Nsamples + 1), Burn-in and Acquire do
    pt = np.fromfile(os.path.join(o, "star_1_L_parallel_tempering.bin"), dtype=np.uint8).reshape(600, 14)
    assert pt[:, 0].sum() == 0
    assert np.fromfile(os.path.join(o, "star_1_A_parallel_tempering.bin"), dtype=np.uint8).reshape(800, 14)[1:, 0].all()
    # Learning starts where Burn-in stopped (restore = 1: variables only)
    vb = tb.outputs.read_restore(os.path.join(rs, "star_1_restore_B_"))["vars"]
    first_l = np.fromfile(os.path.join(o, "star_1_L_params_chain-0.bin")).reshape(600, Nv)[0]
    moved_far = np.abs(first_l - vb[0]) / np.maximum(np.abs(vb[0]), 1e-3)
    assert np.all(moved_far < 0.5)
    # Acquire: the coldest chain sits on the truth (the l=0 mode, S/N 10, within 0.15 microHz; the weak l=2 multiplet within 0.5)
    va = np.fromfile(os.path.join(o, "star_1_A_params_chain-0.bin")).reshape(800, Nv)
    names = [prob.names[i] for i in prob.index_to_relax]
    for k, n in enumerate(names):
        if n == "Frequency_l":
            f_true = prob.params[prob.index_to_relax[k]]
            assert abs(va[:, k].mean() - f_true) < (0.15 if abs(f_true - 1030.0) < 1e-6 else 0.5)
    # the *_mean blocks of the restore files are averages over the last buffer (Outputs::update_buffer_restore, outputs.cpp:1758-1818):
    # vars_mean = mean of the last 100 samples of the params files (Nbuffer = 500: 600 = 500 + 100), and while the proposal is
    # being learnt mu_mean / sigmas_mean are not the last values
    root_l = os.path.join(rs, "star_1_restore_L_")
    last_l, mean_l = tb.outputs.read_restore(root_l), tb.outputs.read_restore_means(root_l)
    for m in range(Cn):
        vm = np.fromfile(os.path.join(o, "star_1_L_params_chain-%d.bin" % m)).reshape(600, Nv)[500:].mean(axis=0)
        assert np.allclose(mean_l["vars"][m], vm, rtol=1e-12, atol=1e-300)
    assert not np.allclose(mean_l["mu"], last_l["mu"], rtol=1e-9) and not np.allclose(mean_l["covarmat"], last_l["covarmat"], rtol=1e-9)
    assert np.all(mean_l["sigma"] > 0) and np.all(np.linalg.eigvalsh(mean_l["covarmat"][0]) > -1e-12)
    # same seed, same files
    r2 = subprocess.run([exe, "execute", "1", "1", "1", "1", "2"], cwd=wd, capture_output=True, text=True, env=env, timeout=600)
    assert r2.returncode == 0
    assert np.array_equal(va, np.fromfile(os.path.join(o, "star_1_A_params_chain-0.bin")).reshape(800, Nv))
