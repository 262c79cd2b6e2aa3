"""On-disk formats of the reference: the files written by csrc/tg_outputs.cpp are
read back (a) by the reference's own, unmodified tools/read_stats.cpp (oracle/_ref/getstats), (b) by
independent numpy readers that follow the reference's readers (diagnostics.cpp:677-760, bin2txt_params.cpp)
and (c) by the restore reader, checked against the keyword grammar of config.cpp:1321-1579."""
import os
import subprocess
import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GETSTATS = os.path.join(ROOT, "oracle", "_ref", "getstats")


def _records(n, Cn, Nv, seed=0):
    g = np.random.default_rng(seed)
    return dict(vars=g.standard_normal((n, Cn, Nv)) * 10.0 ** g.integers(-3, 4, (1, 1, Nv)),
                logL=-1e3 * g.random((n, Cn)), logPrior=-10 * g.random((n, Cn)), logPost=-1e3 * g.random((n, Cn)),
                Pmove=g.random((n, Cn)), moved=(g.random((n, Cn)) < 0.3).astype(np.uint8),
                swap_attempt=(np.arange(n)[:, None] % 2 == 0).astype(np.int32), swap_chain0=g.integers(-1, Cn - 1, (n, 1)).astype(np.int32),
                swap_Pswap=g.random((n, 1)), swap_done=(g.random((n, 1)) < 0.5).astype(np.int32), status=np.zeros((n, Cn), dtype=np.int32))


def _hdr(path):
    d = {}
    for ln in open(path):
        if ln.startswith("!"):
            k, v = ln[1:].split("=", 1)
            d[k.strip()] = v.strip()
    return d


def test_binary_files_and_headers(tb, W, tmp_path):
    prob = W.small_global(3, Nx=2000)
    Cn, Nv = 4, prob.index_to_relax.size
    T = W.tcoefs(Cn)
    out = tb.outputs.Outputs(prob, T, str(tmp_path), root="star_A", Nsamples=50)
    r1, r2 = _records(30, Cn, Nv, 1), _records(20, Cn, Nv, 2)
    out.append(r1); out.append(r2)
    out.close()
    rec = {k: np.concatenate([r1[k], r2[k]]) for k in r1}
    # params: one raw f64 [sample][Nvars] file per chain (outputs.cpp:1307-1320; reader diagnostics.cpp:677-720)
    for c in range(Cn):
        a = np.fromfile(str(tmp_path / ("star_A_params_chain-%d.bin" % c)), dtype=np.float64).reshape(-1, Nv)
        assert np.array_equal(a, rec["vars"][:, c, :])
    h = _hdr(str(tmp_path / "star_A_params.hdr"))
    assert (int(h["Nsamples"]), int(h["Nchains"]), int(h["Nsamples_done"]), int(h["Nvars"])) == (50, Cn, 50, Nv)
    assert int(h["Ncons"]) == prob.params.size - Nv
    assert [int(v) for v in h["relax"].split()] == list(prob.relax)
    assert [int(v) for v in h["plength"].split()] == list(prob.plength)
    assert h["variable_names"].split() == " ".join(prob.names[i] for i in prob.index_to_relax).split()
    cons = np.flatnonzero(prob.relax == 0)
    assert h["constant_names"].split() == " ".join(prob.names[i] for i in cons).split()     # some names contain blanks, as in the reference
    assert np.allclose([float(v) for v in h["constant_values"].split()], prob.params[cons], rtol=1e-5)
    # stat criteria: per sample logL[C] | logPrior[C] | logPost[C] (outputs.cpp:1524-1538)
    s = np.fromfile(str(tmp_path / "star_A_stat_criteria.bin"), dtype=np.float64).reshape(-1, 3, Cn)
    assert np.array_equal(s[:, 0], rec["logL"]) and np.array_equal(s[:, 1], rec["logPrior"]) and np.array_equal(s[:, 2], rec["logPost"])
    # parallel tempering: packed {bool, int, double, bool} = 14 bytes per sample (outputs.cpp:1383-1392)
    dt = np.dtype([("attempt", "u1"), ("chain0", "<i4"), ("Pswitch", "<f8"), ("switched", "u1")])
    assert dt.itemsize == 14
    p = np.fromfile(str(tmp_path / "star_A_parallel_tempering.bin"), dtype=dt)
    assert np.array_equal(p["attempt"], rec["swap_attempt"][:, 0]) and np.array_equal(p["chain0"], rec["swap_chain0"][:, 0])
    assert np.array_equal(p["Pswitch"], rec["swap_Pswap"][:, 0]) and np.array_equal(p["switched"], rec["swap_done"][:, 0])
    assert np.allclose([float(v) for v in _hdr(str(tmp_path / "star_A_parallel_tempering.hdr"))["Tcoefs"].split()], T, rtol=1e-5)


@pytest.mark.skipif(not os.path.exists(GETSTATS), reason="oracle/_ref/getstats (reference tool) not built")
def test_reference_getstats_reads_our_stat_file(tb, W, tmp_path):
    prob = W.small_global(3, Nx=2000)
    Cn, Nv = 3, prob.index_to_relax.size
    out = tb.outputs.Outputs(prob, W.tcoefs(Cn), str(tmp_path), root="s", Nsamples=40)
    rec = _records(40, Cn, Nv, 5)
    out.append(rec); out.close()
    for chain in range(Cn):
        txt = str(tmp_path / ("stats_%d.txt" % chain))
        r = subprocess.run([GETSTATS, str(tmp_path / "s_stat_criteria"), str(chain), txt], capture_output=True, text=True)
        assert r.returncode == 0 and "All done successfully" in r.stdout, r.stdout
        a = np.loadtxt(txt, comments="#")
        assert a.shape == (40, 3)
        for k, key in enumerate(("logL", "logPrior", "logPost")):
            assert np.allclose(a[:, k], rec[key][:, chain], rtol=1e-9, atol=0)      # the tool prints 10 significant digits


def test_append_mode_and_restore_round_trip(tb, W, tmp_path):
    prob = W.small_global(3, Nx=2000)
    Cn, Nv = 3, prob.index_to_relax.size
    T = W.tcoefs(Cn)
    a = tb.outputs.Outputs(prob, T, str(tmp_path), root="r", Nsamples=30)
    r1 = _records(10, Cn, Nv, 1)
    a.append(r1)
    g = np.random.default_rng(9)
    vars_, mu, sigma = r1["vars"][-1], g.standard_normal((Cn, Nv)), g.random(Cn) + 0.1
    B = g.standard_normal((Cn, Nv, Nv)); cov = B @ np.swapaxes(B, 1, 2) * 1e-3
    a.write_restore_arrays(10, vars_, mu, sigma, cov)
    a.close()
    # the restore triplet comes back bit for bit (17 significant digits) and carries the iteration
    rs = tb.outputs.read_restore(str(tmp_path / "r_restore_"))
    assert rs["iteration"] == 10
    for k, ref in (("vars", vars_), ("mu", mu), ("sigma", sigma), ("covarmat", cov)):
        assert np.array_equal(rs[k], ref), k
    assert tb.outputs.read_restore(str(tmp_path / "r_restore_"), do_restore_last_index=False)["iteration"] == 0
    # keyword grammar of config.cpp:1321-1579: 6 keywords in file 1, 4 in file 2, 2 in file 3
    kw = lambda n: [ln.split("=")[0] for ln in open(str(tmp_path / ("r_restore_%d.dat" % n))) if ln.startswith("!")]
    assert [k for k in kw(1) if k in ("! Nvars", "! Nchains", "! iteration", "! variable_names", "! vars", "! vars_mean")].__len__() == 6
    assert [k for k in kw(2) if k in ("! sigmas", "! mus", "! sigmas_mean", "! mus_mean")].__len__() == 4
    assert [k for k in kw(3) if k in ("! covarmats", "! covarmats_mean")].__len__() == 2
    lines3 = [ln.strip() for ln in open(str(tmp_path / "r_restore_3.dat"))]
    assert lines3[lines3.index("! covarmats=") + 1] == "*0" and lines3.count("*%d" % (Cn - 1)) == 2
    # erase_old_files = 0: the data files grow, the headers are left as they were (outputs.cpp:1254)
    b = tb.outputs.Outputs(prob, T, str(tmp_path), root="r", Nsamples=30, Nsamples_sofar=10, erase_old_files=False)
    r2 = _records(5, Cn, Nv, 2)
    b.append(r2); b.close()
    v0 = np.fromfile(str(tmp_path / "r_params_chain-0.bin"), dtype=np.float64).reshape(-1, Nv)
    assert np.array_equal(v0, np.concatenate([r1["vars"][:, 0], r2["vars"][:, 0]]))
    assert int(_hdr(str(tmp_path / "r_params.hdr"))["Nsamples_done"]) == 10


def test_restore_mean_blocks_round_trip(tb, W, tmp_path):
    """buf_restore.*_mean (outputs.cpp:1758-1818): written from the device's running means, read back by tg_restore_read_means
    (MALA::restore_proposal with do_restore_proposal_mean = 1, MALA.cpp:206-232); the last values are untouched by them."""
    prob = W.small_global(3, Nx=2000)
    Cn, Nv = 3, prob.index_to_relax.size
    a = tb.outputs.Outputs(prob, W.tcoefs(Cn), str(tmp_path), root="m", Nsamples=30)
    g = np.random.default_rng(11)
    last = dict(vars=g.standard_normal((Cn, Nv)), mu=g.standard_normal((Cn, Nv)), sigma=g.random(Cn) + 0.1, covarmat=g.standard_normal((Cn, Nv, Nv)))
    mean = dict(vars=g.standard_normal((Cn, Nv)), mu=g.standard_normal((Cn, Nv)), sigma=g.random(Cn) + 0.1, covarmat=g.standard_normal((Cn, Nv, Nv)))
    a.write_restore_arrays(7, last["vars"], last["mu"], last["sigma"], last["covarmat"], means=mean)
    a.close()
    root = str(tmp_path / "m_restore_")
    rl, rm = tb.outputs.read_restore(root), tb.outputs.read_restore_means(root)
    assert rl["iteration"] == 7 and rm["iteration"] == 7
    for k in ("vars", "mu", "sigma", "covarmat"):
        assert np.array_equal(rl[k], last[k]) and np.array_equal(rm[k], mean[k]), k


REF = "/root/reference"
REFBIN = os.path.join(ROOT, "oracle", "_ref", "cpptamcmc")
LOCAL_ID = "TF_3443483_local-v3"


@pytest.mark.skipif(not (os.path.isdir(os.path.join(REF, "Config")) and os.path.exists(REFBIN)), reason="reference tree / oracle/_ref/cpptamcmc absent")
def test_reference_program_restores_from_our_mean_blocks(tb, tmp_path):
    """a restore triplet written by this repo, read by the UNMODIFIED reference program with
    do_restore_proposal_mean = 1 -- its Acquire phase prints the sigma / mu it starts from (MALA::restore_proposal,
    MALA.cpp:206-244): they are the *_mean blocks we wrote, not the last values."""
    import shutil, subprocess
    from test_frontend_cpu import _ref_tree
    wd, out = _ref_tree(tmp_path, os.path.join(REF, "test", "inputs"), [LOCAL_ID], "io_local", nsamples="40     ,  30  , 40")
    pcfg = os.path.join(wd, "Config", "default", "config_default.cfg")
    s = open(pcfg).read()
    assert "do_restore_proposal_mean=0;" in s
    open(pcfg, "w").write(s.replace("do_restore_proposal_mean=0;", "do_restore_proposal_mean=1;"))
    ppre = os.path.join(wd, "Config", "config_presets.cfg")
    s = open(ppre).read()
    import re
    s = re.sub(r"start_index_processing\s*=\s*\d+;", "start_index_processing=2;", s)
    s = re.sub(r"last_index_processing\s*=\s*\d+;", "last_index_processing=2;", s)
    open(ppre, "w").write(s)
    fe = tb.frontend.FrontEnd(wd)
    run = fe.setup(0, 0, 2)                                        # star 0, slice 0, Acquire: restores variables and proposal from the Learning files
    prob, cfg = fe.problem()
    fe.close()
    assert run["do_restore_proposal"] and run["do_restore_proposal_mean"]
    Cn, Nv = cfg.Nchains, prob.index_to_relax.size
    g = np.random.default_rng(3)
    x0 = prob.params[prob.index_to_relax]
    last = dict(vars=np.tile(x0, (Cn, 1)), mu=np.tile(x0, (Cn, 1)) * (1 + 1e-3 * g.standard_normal((Cn, Nv))), sigma=0.5 + g.random(Cn))
    B = 1e-2 * g.standard_normal((Cn, Nv, Nv)); last["covarmat"] = B @ np.swapaxes(B, 1, 2) + 1e-4 * np.eye(Nv)
    mean = dict(vars=last["vars"], mu=last["mu"] * 1.01, sigma=last["sigma"] + 3.0, covarmat=2.0 * last["covarmat"])
    o = tb.outputs.Outputs(prob, cfg.Tcoefs, str(tmp_path / "w"), root="w", Nsamples=40)
    o.write_restore_arrays(40, last["vars"], last["mu"], last["sigma"], last["covarmat"], means=mean)
    o.close()
    os.makedirs(run["restore_dir"], exist_ok=True)
    for k in (1, 2, 3):
        shutil.copy(str(tmp_path / "w" / ("w_restore_%d.dat" % k)), run["restore_dir"] + run["restore_file_in"] + "%d.dat" % k)
    r = subprocess.run([REFBIN, "execute", "1", "1", "1", "1", "2"], cwd=wd, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-800:]
    assert "Use the averaged values over Nbuffer" in r.stdout
    sig = [float(ln.split("=")[1]) for ln in r.stdout.splitlines() if ln.startswith("scaling factor sigma[")]
    assert len(sig) >= Cn and np.allclose(sig[:Cn], mean["sigma"], rtol=2e-5), (sig[:Cn], mean["sigma"])
    mus = [np.array([float(v) for v in ln.split("=")[1].split()]) for ln in r.stdout.splitlines() if ln.startswith("Initial Vectors mu[")]
    assert len(mus) >= Cn and all(np.allclose(mus[c], mean["mu"][c], rtol=2e-5, atol=1e-12) for c in range(Cn))


@pytest.mark.skipif(not (os.path.isdir(os.path.join(REF, "Config")) and os.path.exists(REFBIN)), reason="reference tree / oracle/_ref/cpptamcmc absent")
def test_ascii_outputs_equal_the_reference_programs(tb, tmp_path):
    """file_format = text / debug (outputs.cpp:510-613, :615-676, :791-860).  In debug mode the UNMODIFIED reference program
    writes every sample twice, binary (.dbg) and ASCII (.dbg.txt): its binary records, fed to this repo's writer in the same mode,
    must give its ASCII files back byte for byte (headers included) -- and the binary ones."""
    import re, subprocess
    from test_frontend_cpu import _ref_tree
    wd, out = _ref_tree(tmp_path, os.path.join(REF, "test", "inputs"), [LOCAL_ID], "io_local", nsamples="25     ,  30  , 40")
    pcfg = os.path.join(wd, "Config", "default", "config_default.cfg")
    s = open(pcfg).read()
    s2 = re.sub(r"file_format\s*=\s*\w+;", "file_format=debug;", s)
    assert s2 != s
    open(pcfg, "w").write(s2)
    r = subprocess.run([REFBIN, "execute", "1", "1", "1", "1", "2"], cwd=wd, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-800:]
    fe = tb.frontend.FrontEnd(wd)
    run = fe.setup(0, 0, 0)
    prob, cfg = fe.problem()
    fe.close()
    assert run["file_format"] == 2
    Cn, Nv, n = cfg.Nchains, prob.index_to_relax.size, 25
    base = os.path.join(out, LOCAL_ID, "outputs", run["output_root_name"])
    rec = dict(vars=np.stack([np.fromfile(base + "params_chain-%d.dbg" % c).reshape(n, Nv) for c in range(Cn)], axis=1))
    st = np.fromfile(base + "stat_criteria.dbg").reshape(n, 3 * Cn)
    rec.update(logL=st[:, :Cn].copy(), logPrior=st[:, Cn:2 * Cn].copy(), logPost=st[:, 2 * Cn:].copy())
    pt = np.fromfile(base + "parallel_tempering.dbg", dtype=np.dtype([("a", "u1"), ("c", "<i4"), ("p", "<f8"), ("s", "u1")]))
    assert pt.size == n
    rec.update(swap_attempt=pt["a"].astype(np.int32)[:, None], swap_chain0=pt["c"].astype(np.int32)[:, None], swap_Pswap=pt["p"].astype(np.float64)[:, None],
               swap_done=pt["s"].astype(np.int32)[:, None], Pmove=np.zeros((n, Cn)), moved=np.zeros((n, Cn), dtype=np.uint8), status=np.zeros((n, Cn), dtype=np.int32))
    root = run["output_root_name"].rstrip("_")
    o = tb.outputs.Outputs(prob, cfg.Tcoefs, str(tmp_path / "mine"), root=root, Nsamples=25, file_format="debug")
    o.append({k: rec[k][:10] for k in rec})                   # two blocks: the header goes with the first only
    o.append({k: rec[k][10:] for k in rec})
    o.close()
    mine = str(tmp_path / "mine" / (root + "_"))
    names = ["params_chain-%d" % c for c in range(Cn)] + ["stat_criteria", "parallel_tempering"]
    for nm in names:
        assert open(mine + nm + ".dbg", "rb").read() == open(base + nm + ".dbg", "rb").read(), nm
        a, b = open(mine + nm + ".dbg.txt").read(), open(base + nm + ".dbg.txt").read()
        assert a == b, (nm, [(x, y) for x, y in zip(a.splitlines(), b.splitlines()) if x != y][:3])
    # text mode: the same lines under the extension txt, and no binary files
    o = tb.outputs.Outputs(prob, cfg.Tcoefs, str(tmp_path / "txt"), root=root, Nsamples=25, file_format="text")
    o.append(rec); o.close()
    for nm in names:
        assert open(str(tmp_path / "txt" / (root + "_" + nm + ".txt"))).read() == open(base + nm + ".dbg.txt").read()
    assert not [f for f in os.listdir(str(tmp_path / "txt")) if f.endswith((".bin", ".hdr", ".dbg"))]


def test_header_declares_every_outputs_symbol(tb):
    import re
    decl = set(re.findall(r"\b(tg_\w+)\s*\(", open(os.path.join(ROOT, "include", "tamcmc_outputs.h")).read()))
    lib = tb.load_library()
    for name in decl:
        assert hasattr(lib.lib, name), name
    assert decl == set(tb.gpu.EXPORTS_OUT)
