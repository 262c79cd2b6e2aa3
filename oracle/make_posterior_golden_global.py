#!/usr/bin/env python
"""Generates tests/golden/posterior_global.npz: posterior summaries of the UNMODIFIED reference program on a reduced GLOBAL
main-sequence fit -- model_MS_Global_a1etaa3_HarveyLike (models.cpp:480-637) with the io_MS_Global priors
(priors_calc.cpp:24-172), 6 radial orders x l = 0, 1, 2, two Harvey-like noise terms + white noise, about 40 free
parameters, a synthetic spectrum of 44 000 bins in the reference's file formats (pattern of the shipped
test/inputs/kplr008379927_kasoc-psd_slc_v2_1000.model, written by tamcmc_b200.config_tree) -- run through
Burn-in -> Learning -> Acquire by `cpptamcmc execute 1 1 1` (MALA::execute, MALA.cpp:540-739), several times (the reference
seeds rand() with time(NULL)).  The file also holds the problem the reference front end built (oracle/_ref/ref_frontend_dump)
and the spectrum, so that the GPU test can run the same three phases where /root/reference does not exist.

Test infrastructure.  Needs /root/reference and `make -C oracle full`.
Usage: python oracle/make_posterior_golden_global.py [runs] [omp_threads]"""
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
NSAMPLES = tuple(int(v) for v in os.environ.get("TG_GOLDEN_NS", "2000,6000,12000").split(","))
OUT = os.environ.get("TG_GOLDEN_OUT", os.path.join(ROOT, "tests", "golden", "posterior_global.npz"))
STAR = "synthglobal"
DNU, STEP, LO, HI = 70.0, 0.01, 1090.0, 1530.0
NORD = 6


def modes():
    rows = []
    n = np.arange(NORD)
    f0 = 1120.0 + DNU * n + 0.15 * (n - 2.5) ** 2
    G = np.linspace(0.9, 1.8, NORD)
    H = 14.0 * np.exp(-0.5 * ((f0 - f0.mean()) / 150.0) ** 2)
    for l, (df, hv) in enumerate([(0.0, 1.0), (DNU / 2 - 3.0, 1.5), (-6.0, 0.53)]):
        for k in range(NORD):
            f = f0[k] + df
            rows.append((l, f, f - 3.0, f + 3.0, G[k], H[k] * hv))
    return rows


def build_tree(tb, t):
    ct = tb.config_tree
    md, out, wd = os.path.join(t, "in"), os.path.join(t, "out"), os.path.join(t, "wd")
    os.makedirs(md)
    x = LO - 20.0 + STEP * np.arange(int(round((HI - LO + 40.0) / STEP)))
    noise = ((0.0, 0.0, 1.0), (3.0, 2.5, 4.0), (0.5, 0.3, 2.0))
    ct.write_model_file(os.path.join(md, STAR + ".model"), modes(), [(LO, HI)], "model_MS_Global_a1etaa3_HarveyLike", Dnu=DNU, C_l=0.0, noise=noise, N0=0.4,
                        a1=1.2, inclination=50.0, height_prior=("Jeffreys", 1.0, 1000.0), trunc_c=20.0)
    ct.write_data_file(os.path.join(md, STAR + ".data"), x, np.ones_like(x))
    ct.write_config_tree(wd, md, out, [STAR], prior_fct="io_MS_Global", Nsamples=NSAMPLES, c0=(1.8, 1.7, 0.0), restore=(0, 1, 2), last_phase=2,
                         Nbuffer=4000, Nt_learn=(500, 1500, 100000))
    return md, out, wd, x


def main():
    runs = int(sys.argv[1]) if len(sys.argv) > 1 else 4
    threads = sys.argv[2] if len(sys.argv) > 2 else "5"
    import tamcmc_b200_loader
    tb = tamcmc_b200_loader.load()
    from oracle.oracle import Oracle
    O = Oracle()
    means, stds, acc = [], [], []
    with tempfile.TemporaryDirectory() as t:
        md, out, wd, x = build_tree(tb, t)
        # the truth: the model the .model file describes, through this repo's front end and the oracle; y = m * Exp(1)
        fe = tb.frontend.FrontEnd(wd)
        fe.setup(0, 0, 0)
        prob, _ = fe.problem()
        fe.close()
        st, m, _ = O.model(prob.model_case, prob.params, prob.plength, prob.x)
        assert st == 0
        y = np.ones_like(x)
        i0 = int(np.searchsorted(x, prob.x[0] - 1e-9))
        y[i0:i0 + m.size] = m * (-np.log(tb.workloads.philox_uniform(m.size, 2024)))
        tb.config_tree.write_data_file(os.path.join(md, STAR + ".data"), x, y)
        env = dict(os.environ, OMP_NUM_THREADS=threads)
        dump = subprocess.run([os.path.join(ROOT, "oracle", "_ref", "ref_frontend_dump"), "1", "1", "0"], cwd=wd, check=True, capture_output=True, text=True).stdout
        for r in range(runs):
            t0 = time.time()
            subprocess.run([os.path.join(ROOT, "oracle", "_ref", "cpptamcmc"), "execute", "1", "1", "1"], cwd=wd, check=True, env=env,
                           stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
            o = os.path.join(out, STAR, "outputs")
            base = [os.path.join(o, f[:-len("params.hdr")]) for f in os.listdir(o) if f.endswith("_A_params.hdr")][0]
            hdr = {k.strip(): v.strip() for k, v in (ln[1:].split("=", 1) for ln in open(base + "params.hdr") if ln.startswith("!"))}
            nv = int(hdr["Nvars"])
            v = np.fromfile(base + "params_chain-0.bin").reshape(-1, nv)
            means.append(v.mean(0)); stds.append(v.std(0))
            rows = [ln.split() for ln in open(base + "acceptance.txt") if not ln.startswith(("#", "!"))]
            acc.append(np.mean([[float(a) for a in row[1:]] for row in rows], axis=0))
            print("run %d: %.0f s, acceptance(chain 0) %.3f, mean[0:3] %s" % (r, time.time() - t0, acc[-1][0], means[-1][:3]), flush=True)
            time.sleep(1.1)                      # the reference seeds with time(NULL)
        d = np.loadtxt(os.path.join(md, STAR + ".data"), comments=("#", "!", "*"))
    f = {k.strip(): v.strip() for k, v in (ln.split(":", 1) for ln in dump.splitlines() if ":" in ln)}
    num = lambda k: np.array([float(a) for a in f[k].split()])
    sel = (d[:, 0] >= LO) & (d[:, 0] <= HI)
    xs, ys = d[sel, 0], d[sel, 1]
    nx = int(f["Nx"])
    if xs.size != nx:                            # the reference's crop (config.cpp:96-125) decides the end points
        j0 = int(np.argmin(np.abs(d[:, 0] - num("data")[0])))
        xs, ys = d[j0:j0 + nx, 0], d[j0:j0 + nx, 1]
    assert xs.size == nx and xs[0] == num("data")[0] and xs[-1] == num("data")[1], (xs.size, nx, xs[0], xs[-1], num("data"))
    sw = [int(a) for a in f["switches"].split()]
    np.savez_compressed(OUT,
                        x=xs, y=ys, params=num("inputs"), relax=num("relax").astype(np.int32), plength=num("plength").astype(np.int32), extra_priors=num("extra_priors"),
                        priors=np.stack([num("priors_row%d" % k) for k in range(4)], axis=1), priors_switch=num("priors_names_switch").astype(np.int32),
                        model_case=sw[0], like_case=sw[1], prior_case=sw[2], likelihood_params=float(f["likelihood_params"]),
                        init_err=np.sqrt(num("cov0_diag_chain0")), Tcoefs=num("Tcoefs"), nsamples=np.array(NSAMPLES), c0=np.array([1.8, 1.7, 0.0]),
                        Nt_learn0=np.array([500, 1500]), names=np.array([n.strip() for n in f["inputs_names"].split("|") if n.strip()]),
                        ref_mean=np.array(means), ref_std=np.array(stds), ref_acceptance=np.array(acc))
    print("wrote %s: %d reference runs, Nv=%d, Nx=%d" % (OUT, runs, nv, nx))


if __name__ == "__main__":
    main()
