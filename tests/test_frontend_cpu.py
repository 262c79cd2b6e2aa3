"""Input front end (csrc/tg_frontend.cpp) and numeric diagnostics (rank 3) on the CPU.

The checker is the UNMODIFIED reference front end (Config, Config_presets, io_ms_global, io_local, MALA constructor)
compiled against the Eigen-API shim: ``oracle/_ref/ref_frontend_dump`` prints Input_Data / Data / MALA configuration
with 17 digits and the same fields are compared one by one.  Cases: the only complete star the reference ships
(test/inputs/TF_3443483_local-v3, all 8 slices, model_MS_local_basic), every global `.model` file it ships, the
global model family obtained by renaming model_fullname, and trees written by tamcmc_b200.config_tree (read by both).
Tests that need the reference tree are skipped where /root/reference does not exist."""
import os
import shutil
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
DUMP = os.path.join(ROOT, "oracle", "_ref", "ref_frontend_dump")
REFBIN = os.path.join(ROOT, "oracle", "_ref", "cpptamcmc")
have_ref = os.path.isdir(os.path.join(REF, "Config")) and os.path.exists(DUMP)
needs_ref = pytest.mark.skipif(not have_ref, reason="reference tree / oracle/_ref/ref_frontend_dump absent")
TEXT = ("inputs_names", "priors_names", "model_fct_name", "likelihood_fct_name", "prior_fct_name", "output_root_name", "dir_out", "restore_dir",
        "restore_file_in", "restore_file_out")
LOCAL_ID = "TF_3443483_local-v3"
GLOBAL_MODELS = ["model_MS_Global_a1etaa3_HarveyLike_Classic", "model_MS_Global_a1etaa3_HarveyLike_Classic_v2", "model_MS_Global_a1etaa3_HarveyLike_Classic_v3",
                 "model_MS_Global_a1etaa3_HarveyLike", "model_MS_Global_a1l_etaa3_HarveyLike", "model_MS_Global_a1n_etaa3_HarveyLike",
                 "model_MS_Global_a1nl_etaa3_HarveyLike"]


def _parse(txt):
    return {k.strip(): v.strip() for k, v in (ln.split(":", 1) for ln in txt.splitlines() if ":" in ln)}


def _same(a, b, key, tol=1e-12):
    if key in TEXT:
        return a == b
    ta, tb_ = a.split(), b.split()
    if len(ta) != len(tb_):
        return False
    for u, v in zip(ta, tb_):
        try:
            fu, fv = float(u), float(v)
        except ValueError:
            if u != v:
                return False
            continue
        if fu != fv and not abs(fu - fv) <= tol * max(abs(fu), abs(fv)):
            return False
    return True


def _compare(tb, workdir, star, sl, phase):
    ref = subprocess.run([DUMP, str(star + 1), str(sl + 1), str(phase)], cwd=workdir, capture_output=True, text=True)
    assert ref.returncode == 0, "reference front end failed: " + ref.stdout[-400:] + ref.stderr[-400:]
    fe = tb.frontend.FrontEnd(workdir)
    fe.setup(star, sl, phase)
    mine, r = _parse(fe.dump()), _parse(ref.stdout)
    fe.close()
    assert len(r) >= 30
    bad = [k for k in r if k not in mine or not _same(mine[k], r[k], k)]
    assert not bad, "fields differ from the reference front end: %s\n mine: %s\n ref : %s" % (bad, mine.get(bad[0], "")[:400], r[bad[0]][:400])


def _ref_tree(tmp_path, models_dir, ids, prior_fct, nsamples="200     ,  300  , 400"):
    """A working directory with the reference's OWN Config/ files, edited the way SURVEY.md 8 (cfg1) says."""
    wd = str(tmp_path / "wd")
    shutil.copytree(os.path.join(REF, "Config"), os.path.join(wd, "Config"))
    for r, _, fs in os.walk(wd):
        os.chmod(r, 0o755)
        for f in fs:
            os.chmod(os.path.join(r, f), 0o644)
    os.makedirs(os.path.join(wd, "tmp"))
    out = str(tmp_path / "out")
    os.makedirs(out)
    p = os.path.join(wd, "Config", "config_presets.cfg")
    s = open(p).read()
    s = s.replace("cfg_models_dir=/Users/obenomar/Documents/TAMCMC-C-work/data/inputs-tests/;", "cfg_models_dir=%s/;" % models_dir.rstrip("/"))
    s = s.replace("cfg_out_dir=/Users/obenomar/Documents/TAMCMC-C-work/data/outputs-tests;", "cfg_out_dir=%s;" % out)
    s = s.replace("Nsamples        = 5000     ,  50000  , 500000;", "Nsamples        = %s;" % nsamples)
    s = s.replace("table_ids=2, 2;", "table_ids=%d, 2;" % len(ids)).replace("TF_10528917   1;\nTF_11519450   1;\n", "".join("%s 1;\n" % i for i in ids))
    open(p, "w").write(s)
    p = os.path.join(wd, "Config", "default", "config_default.cfg")
    s = open(p).read().replace("prior_fct_name=io_local;", "prior_fct_name=%s;" % prior_fct).replace("model_fct_name=io_local;", "model_fct_name=%s;" % prior_fct)
    open(p, "w").write(s)
    return wd, out


def _synthetic_data(tb, model_path, data_path, seed):
    txt = open(model_path).read()
    rg = [ln for ln in txt.splitlines() if ln.strip().startswith("*")]
    lo, hi = float(rg[0].split()[1]), float(rg[-1].split()[2])
    x = np.arange(max(lo - 30, 1.0), hi + 30, 0.05)
    y = -np.log(np.random.default_rng(seed).random(x.size)) * (1 + 5 * np.exp(-0.5 * ((x - 0.5 * (lo + hi)) / (0.2 * (hi - lo))) ** 2))
    tb.config_tree.write_data_file(data_path, x, y)


@needs_ref
def test_shipped_star_all_slices_equal_the_reference_front_end(tb, tmp_path):
    wd, _ = _ref_tree(tmp_path, os.path.join(REF, "test", "inputs"), [LOCAL_ID], "io_local")
    fe = tb.frontend.FrontEnd(wd)
    assert fe.counts() == dict(n_stars=1, n_phases=3, first_phase=0, last_phase=0) and fe.nslices(0) == 8
    fe.close()
    for sl in range(8):
        _compare(tb, wd, 0, sl, 0)


@needs_ref
def test_global_model_family_equals_the_reference_front_end(tb, tmp_path):
    src = os.path.join(REF, "test", "inputs")
    md = str(tmp_path / "inputs")
    os.makedirs(md)
    base = open(os.path.join(src, "kplr008379927_kasoc-psd_slc_v2_1000.model")).read()
    ids = []
    for k, m in enumerate(GLOBAL_MODELS):
        s = base.replace("model_MS_Global_a1etaa3_HarveyLike_Classic_v2", m)
        if k == 3:
            s = s.replace("Height            Jeffreys", "fit_squareAmplitude_instead_Height bool 1\n                   Height            Jeffreys")
        if k == 0:
            s = s.replace("Inclination             Uniform         55.000000", "trunc_c Fix 25.0\n Inclination             Uniform         95.000000")
        open(os.path.join(md, "g%d.model" % k), "w").write(s)
        ids.append("g%d" % k)
    for extra in ("00088.0.model", "00088.1.model", "Bugs-Yuting/02199.0.model", "Bugs-Yuting/02194.0.model"):
        if os.path.exists(os.path.join(src, extra)):
            name = "e_" + os.path.basename(extra).replace(".model", "").replace(".", "_")
            shutil.copy(os.path.join(src, extra), os.path.join(md, name + ".model"))
            os.chmod(os.path.join(md, name + ".model"), 0o644)
            ids.append(name)
    for k, i in enumerate(ids):
        _synthetic_data(tb, os.path.join(md, i + ".model"), os.path.join(md, i + ".data"), k)
    wd, _ = _ref_tree(tmp_path, md, ids, "io_MS_Global")
    for st in range(len(ids)):
        _compare(tb, wd, st, 0, 0)


@needs_ref
def test_generated_tree_is_read_alike_by_both_front_ends(tb, tmp_path):
    """Files written by tamcmc_b200.config_tree (the ones the GPU tests use) through both front ends: local basic,
    local Hnlm with amplitudes, global Classic; Burn-in and Acquire phases."""
    md, out, wd = str(tmp_path / "in"), str(tmp_path / "out"), str(tmp_path / "wd")
    os.makedirs(md)
    ct = tb.config_tree
    modes = [(0, 1030.0, 1029.0, 1031.0, 1.0, 10.0), (1, 1055.0, 1054.0, 1056.0, 1.0, 15.0), (2, 1024.0, 1023.0, 1025.0, 1.0, 5.0),
             (0, 1130.0, 1129.0, 1131.0, 1.2, 9.0), (1, 1155.0, 1154.0, 1156.0, 1.2, 12.0), (2, 1124.0, 1123.0, 1125.0, 1.2, 4.0)]
    x = 1000.0 + 0.01 * np.arange(20000)
    y = -np.log(np.random.default_rng(5).random(x.size))
    for name, full, extra in (("loc", "model_MS_local_basic", ()), ("hnlm", "model_MS_local_Hnlm", ("fit_squareAmplitude_instead_Height bool 1",))):
        ct.write_model_file(os.path.join(md, name + ".model"), modes, [(1010.0, 1070.0), (1110.0, 1170.0)], full, Dnu=100.0, N0=1.0, trunc_c=20.0, extra_common=extra)
        ct.write_data_file(os.path.join(md, name + ".data"), x, y)
    ct.write_config_tree(wd, md, out, ["loc", "hnlm"], prior_fct="io_local", Nsamples=(50, 60, 70), restore=(0, 0, 0))
    for st in range(2):
        for sl in range(2):
            _compare(tb, wd, st, sl, 0)
    _compare(tb, wd, 0, 1, 2)
    wd2 = str(tmp_path / "wd2")
    ct.write_model_file(os.path.join(md, "glob.model"), modes, [(1005.0, 1180.0)], "model_MS_Global_a1etaa3_HarveyLike_Classic", Dnu=100.0, C_l=30.0,
                        noise=((0, 0, 1), (11.0, 49.6, 4.0), (0.9, 1.35, 2.0)), N0=0.13, a1=1.4, inclination=55.0)
    ct.write_data_file(os.path.join(md, "glob.data"), x, y)
    ct.write_config_tree(wd2, md, out, ["glob"], prior_fct="io_MS_Global", Nsamples=(50, 60, 70), restore=(0, 0, 0))
    _compare(tb, wd2, 0, 0, 0)
    _compare(tb, wd2, 0, 0, 1)


@needs_ref
def test_front_end_problem_reproduces_the_reference_program_start(tb, oracle, tmp_path):
    """cfg1 of BASELINE.json end to end on the CPU side: the WHOLE unmodified reference program
    (`cpptamcmc execute 1 1 1 1 2`, slice 1 of the shipped star) prints logLikelihood / logPrior / logPosterior of its
    starting point; the problem built by this repo's front end, evaluated by the oracle, gives the same numbers; the
    evidence the reference writes equals tg_evidence on the logL columns of its own stat_criteria file."""
    wd, out = _ref_tree(tmp_path, os.path.join(REF, "test", "inputs"), [LOCAL_ID], "io_local", nsamples="120     ,  300  , 400")
    r = subprocess.run([REFBIN, "execute", "1", "1", "1", "1", "2"], cwd=wd, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-500:]
    val = {}
    for ln in r.stdout.splitlines():
        for key in ("logLikelihood[0]=", "logPrior[0]=", "logPosterior[0]="):
            if key in ln and key not in val:
                val[key] = float(ln.split("=")[1])
    fe = tb.frontend.FrontEnd(wd)
    run = fe.setup(0, 0, 0)
    prob, cfg = fe.problem()
    fe.close()
    res = oracle.generate_models(prob, prob.params[None, :], np.ones(1))
    assert abs(res["logL"][0] - val["logLikelihood[0]="]) < 2e-5 * abs(res["logL"][0])         # the reference prints 6 digits
    assert abs(res["logPrior"][0] - val["logPrior[0]="]) < 2e-5 * abs(res["logPrior"][0])
    assert abs(res["logPost"][0] - val["logPosterior[0]="]) < 2e-5 * abs(res["logPost"][0])
    # evidence: the reference's file against tg_evidence on the reference's own samples
    base = os.path.join(out, LOCAL_ID)
    stat = np.fromfile(os.path.join(base, "outputs", run["output_root_name"] + "stat_criteria.bin"))
    Cn = cfg.Nchains
    stat = stat.reshape(-1, 3 * Cn)
    last = [ln for ln in open(os.path.join(base, "diags", run["output_root_name"] + "evidence.txt")) if not ln.startswith(("#", "!"))][-1].split()
    # Diagnostics::read_bin_matrix_dbl (diagnostics.cpp:677-720) drops the last row it reads: the reference's evidence is over rows[:-1]
    Lb, ev = tb.frontend.evidence(cfg.Tcoefs, stat[:-1, :Cn], 1000)
    assert np.allclose(Lb, [float(v) for v in last[1:1 + Cn]], rtol=2e-9) and abs(ev - float(last[-1])) < 2e-9 * abs(ev)


def test_evidence_and_acceptance_follow_the_reference_formulas(tb):
    rng = np.random.default_rng(0)
    Cn, n = 6, 500
    T = 1.7 ** np.arange(Cn)
    logL = -1000.0 / T[None, :] + rng.standard_normal((n, Cn))
    Lb, ev = tb.frontend.evidence(T, logL, 50)
    assert np.allclose(Lb, logL.mean(axis=0), rtol=1e-13)

    def interp1(x, a):                       # interpol.cpp:74-82
        if x <= 0:
            return a[0]
        if x >= len(a) - 1:
            return a[-1]
        j = int(x)
        return a[j] + (x - j) * (a[j + 1] - a[j])

    def interp2(x, a):                       # interpol.cpp:96-103
        if x <= .5 or x >= len(a) - 1.5:
            return interp1(x, a)
        j = int(x + .5)
        t = 2 * (x - j)
        f_1, f0, f1 = (a[j - 1] + a[j]) / 2, a[j], (a[j] + a[j + 1]) / 2
        l, r = f0 - t * (f_1 - f0), f0 + t * (f1 - f0)
        return (l + r + t * (r - l)) / 2
    m = 50 * Cn
    step = (Cn - 1) / (m - 1)
    assert abs(ev - np.mean([interp2(j * step, Lb) for j in range(m)])) < 1e-10 * abs(ev)
    moved = (rng.random((n, Cn)) < 0.3).astype(np.uint8)
    xa, rate = tb.frontend.acceptance(moved, Nbuffer=1000, Ncopy=2, Nsamples_init=100)
    assert xa == 2.5 * 1000 + 100 and np.allclose(rate, moved.sum(axis=0) / 1000.0)


def test_front_end_errors_are_statuses_not_exits(tb, tmp_path):
    with pytest.raises(tb.TgError, match="Could not open"):
        tb.frontend.FrontEnd(str(tmp_path))
    ct = tb.config_tree
    md, out, wd = str(tmp_path / "in"), str(tmp_path / "out"), str(tmp_path / "wd")
    os.makedirs(md)
    ct.write_config_tree(wd, md, out, ["nothing"], prior_fct="io_local")
    fe = tb.frontend.FrontEnd(wd)
    with pytest.raises(tb.TgError, match="Unable to open the file"):
        fe.setup(0, 0, 0)
    with pytest.raises(tb.TgError, match="star index"):
        fe.setup(3, 0, 0)
    fe.close()


def test_cli_builds_and_fails_loudly_without_a_gpu(tb, tmp_path):
    """tools/cpptamcmc_gpu.cpp: `execute 0` only loads the configuration; `execute 1 ...` needs the GPU (exit code 3)."""
    import torch
    tb.load_library()
    libdir = os.path.join(ROOT, "tamcmc-c-_b200", "csrc")
    exe = str(tmp_path / "cpptamcmc_gpu")
    subprocess.run(["g++", "-std=c++11", "-O1", "-I" + os.path.join(ROOT, "include"), os.path.join(ROOT, "tools", "cpptamcmc_gpu.cpp"), "-o", exe,
                    "-L" + libdir, "-ltamcmc_gpu", "-Wl,-rpath," + libdir], check=True)
    ct = tb.config_tree
    md, out, wd = str(tmp_path / "in"), str(tmp_path / "out"), str(tmp_path / "wd")
    os.makedirs(md)
    x = 1000.0 + 0.01 * np.arange(8000)
    ct.write_model_file(os.path.join(md, "s.model"), [(0, 1030.0, 1029.0, 1031.0, 1.0, 10.0), (2, 1024.0, 1023.0, 1025.0, 1.0, 5.0)], [(1010.0, 1050.0)],
                        "model_MS_local_basic", Dnu=100.0)
    ct.write_data_file(os.path.join(md, "s.data"), x, np.ones_like(x))
    ct.write_config_tree(wd, md, out, ["s"], prior_fct="io_local", Nsamples=(20, 20, 20))
    r = subprocess.run([exe, "execute", "0"], cwd=wd, capture_output=True, text=True)
    assert r.returncode == 0 and "execute=0" in r.stdout
    assert subprocess.run([exe, "bogus"], cwd=wd, capture_output=True).returncode == 2
    if not torch.cuda.is_available():
        r = subprocess.run([exe, "execute", "1", "1", "1"], cwd=wd, capture_output=True, text=True)
        assert r.returncode == 3 and "no CUDA device" in r.stdout
