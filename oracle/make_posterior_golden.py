#!/usr/bin/env python
"""Generates tests/golden/posterior_cfg1.npz: posterior summaries of the UNMODIFIED reference program on BASELINE.json's
configs[0] -- `cpptamcmc execute 1 1 1 1 2` on the star the reference ships (test/inputs/TF_3443483_local-v3, slice 1,
model_MS_local_basic, 10 tempered chains) through Burn-in -> Learning -> Acquire -- together with the problem the
reference front end built for it (oracle/_ref/ref_frontend_dump) and the cropped spectrum, so that the GPU test can run
the same three phases where /root/reference does not exist.

Test infrastructure.  Needs /root/reference and `make -C oracle full`.  Usage: python oracle/make_posterior_golden.py [runs]"""
import os
import shutil
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
STAR = "TF_3443483_local-v3"
NSAMPLES = (3000, 6000, 20000)


def tree(wd, out):
    shutil.copytree(os.path.join(REF, "Config"), os.path.join(wd, "Config"))
    for r, _, fs in os.walk(wd):
        os.chmod(r, 0o755)
        for f in fs:
            os.chmod(os.path.join(r, f), 0o644)
    os.makedirs(os.path.join(wd, "tmp"))
    os.makedirs(out)
    p = os.path.join(wd, "Config", "config_presets.cfg")
    s = open(p).read()
    s = s.replace("cfg_models_dir=/Users/obenomar/Documents/TAMCMC-C-work/data/inputs-tests/;", "cfg_models_dir=%s/test/inputs/;" % REF)
    s = s.replace("cfg_out_dir=/Users/obenomar/Documents/TAMCMC-C-work/data/outputs-tests;", "cfg_out_dir=%s;" % out)
    s = s.replace("Nsamples        = 5000     ,  50000  , 500000;", "Nsamples        = %d     ,  %d  , %d;" % NSAMPLES)
    s = s.replace("last_index_processing=0;", "last_index_processing=2;")
    s = s.replace("table_ids=2, 2;", "table_ids=1, 2;").replace("TF_10528917   1;\nTF_11519450   1;\n", STAR + " 1;\n")
    open(p, "w").write(s)


def main():
    runs = int(sys.argv[1]) if len(sys.argv) > 1 else 6
    means, stds, acc = [], [], []
    dump = None
    for r in range(runs):
        with tempfile.TemporaryDirectory() as t:
            wd, out = os.path.join(t, "wd"), os.path.join(t, "out")
            tree(wd, out)
            subprocess.run([os.path.join(ROOT, "oracle", "_ref", "cpptamcmc"), "execute", "1", "1", "1", "1", "2"], cwd=wd, check=True,
                           stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
            base = os.path.join(out, STAR, "outputs", STAR + "_1_A_")
            hdr = {k.strip(): v.strip() for k, v in (ln[1:].split("=", 1) for ln in open(base + "params.hdr") if ln.startswith("!"))}
            nv = int(hdr["Nvars"])
            v = np.fromfile(base + "params_chain-0.bin").reshape(-1, nv)
            means.append(v.mean(0)); stds.append(v.std(0))
            rows = [ln.split() for ln in open(base + "acceptance.txt") if not ln.startswith(("#", "!"))]
            acc.append(np.mean([[float(a) for a in row[1:]] for row in rows], axis=0))
            if dump is None:
                dump = subprocess.run([os.path.join(ROOT, "oracle", "_ref", "ref_frontend_dump"), "1", "1", "0"], cwd=wd, check=True, capture_output=True, text=True).stdout
                # the cropped spectrum, as Config::setup crops it (config.cpp:96-125)
                d = np.loadtxt(os.path.join(REF, "test", "inputs", STAR + ".data"), comments=("#", "!", "*"))
        print("run %d: mean(f0) %.4f  acceptance(chain 0) %.3f" % (r, means[-1][2], acc[-1][0]))
    f = {k.strip(): v.strip() for k, v in (ln.split(":", 1) for ln in dump.splitlines() if ":" in ln)}
    num = lambda k: np.array([float(a) for a in f[k].split()])
    lo, hi = 94.30, 102.20                                       # slice 1 of the .model file
    sel = (d[:, 0] >= lo) & (d[:, 0] < hi)
    x, y = d[sel, 0], d[sel, 1]
    assert x.size == int(f["Nx"]) and x[0] == num("data")[0] and x[-1] == num("data")[1]
    sw = [int(a) for a in f["switches"].split()]
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "posterior_cfg1.npz"),
                        x=x, y=y, params=num("inputs"), relax=num("relax").astype(np.int32), plength=num("plength").astype(np.int32), extra_priors=num("extra_priors"),
                        priors=np.stack([num("priors_row%d" % k) for k in range(4)], axis=1), priors_switch=num("priors_names_switch").astype(np.int32),
                        model_case=sw[0], like_case=sw[1], prior_case=sw[2], likelihood_params=float(f["likelihood_params"]),
                        init_err=np.sqrt(num("cov0_diag_chain0")), Tcoefs=num("Tcoefs"), nsamples=np.array(NSAMPLES), c0=np.array([1.8, 1.7, 0.0]),
                        Nt_learn0=np.array([500, 1500]), names=np.array([n.strip() for n in f["inputs_names"].split("|") if n.strip()]),
                        ref_mean=np.array(means), ref_std=np.array(stds), ref_acceptance=np.array(acc))
    print("wrote tests/golden/posterior_cfg1.npz: %d reference runs, Nv=%d" % (runs, nv))


if __name__ == "__main__":
    main()
