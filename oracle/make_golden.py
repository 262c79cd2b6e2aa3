"""Generates tests/golden/golden_ref.npz from the UNMODIFIED reference sources (oracle/_ref).

Run in the container that has /root/reference (the GPU box does not):
    python oracle/make_golden.py
Every entry holds the full inputs (params, plength, priors, x, y, ...) and the outputs the
reference functions produced for them: the model spectrum (models.cpp via call_model), the
tempered log-likelihood (likelihoods.cpp via call_likelihood, T as stored) and the log-prior
(priors_calc.cpp via call_prior).  The cfg1 entry uses slice 1 (94.30-102.20 microHz) of the only
data file the reference ships, test/inputs/TF_3443483_local-v3.data, with the mode parameters of
test/inputs/TF_3443483_local-v3.model.  Test infrastructure only.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path = [p for p in sys.path if os.path.abspath(p or '.') != os.path.join(ROOT, 'oracle')]
sys.path.insert(0, ROOT)
import tamcmc_b200_loader  # noqa: E402
from oracle.oracle import RefLib  # noqa: E402

REF_DATA = "/root/reference/test/inputs/TF_3443483_local-v3.data"


def cfg1_problem(W):
    d = np.loadtxt(REF_DATA, comments=["#", "!", "*"])
    sel = (d[:, 0] >= 94.30) & (d[:, 0] <= 102.20)
    x, y = d[sel, 0].copy(), d[sel, 1].copy()
    # modes of slice 1 in TF_3443483_local-v3.model: l=0 @99.10560 (G=0.30864,H=1199.06221), l=2 @97.77160 (0.30646, 635.50294)
    modes = [(0, 99.10560, 1199.06221, 0.30864), (2, 97.77160, 635.50294, 0.30646)]
    return W._local_problem("cfg1-TF_3443483-slice1", x, modes, a1=0.4, inc=45.0, N0=126.0, trunc_c=20.0,
                            seed=1, make_y=None, y=y)


def problems(W):
    P = [W.small_global(mc, Nx=2000, step=0.15) for mc in (2, 3, 6, 7, 8, 9, 10, 12, 13)]
    P += [W.small_global(3, Nx=2000, step=0.15, asym=20.0, do_amp=1, a3=0.02),
          W.small_global(2, Nx=2000, step=0.15, trunc_c=10000.0),
          W.small_global(6, Nx=2000, step=0.15, asym=-12.0),
          W.small_local(11), W.small_local(14), W.small_local(11, asym=15.0, do_amp=1), W.small_local(14, do_amp=1),
          W.gauss_problem(0, like_case=1), W.gauss_problem(1, like_case=0), W.s_local(Nx=3000)]
    if os.path.exists(REF_DATA):
        P.append(cfg1_problem(W))
    return P


def main():
    tb = tamcmc_b200_loader.load()
    W = tb.workloads
    R = RefLib()
    assert R.available, "oracle/_ref is not built (needs /root/reference)"
    out = {}
    names = []
    for k, p in enumerate(problems(W)):
        st, m = R.model(p.model_case, p.params, p.plength, p.x)
        assert st == 0
        if p.name.startswith("cfg1") or p.name.startswith("gauss"):
            y = p.y
        else:
            y = m * (-np.log(W.philox_uniform(m.size, 1000 + k)))
        T = 1.7 ** (k % 4)
        logL = R.likelihood(p.like_case, y, p.sigma_y, m, p.likelihood_params, T)
        pr = R.prior(p.prior_case, p.params, p.plength, p.priors, p.priors_switch, p.extra_priors)
        key = "%02d_%s" % (k, p.name)
        names.append(key)
        for f, v in dict(model_case=p.model_case, like_case=p.like_case, prior_case=p.prior_case, plength=p.plength,
                         params=p.params, relax=p.relax, priors=p.priors, priors_switch=p.priors_switch,
                         extra_priors=p.extra_priors, x=p.x, y=y, sigma_y=(p.sigma_y if p.sigma_y is not None else np.zeros(0)),
                         likelihood_params=p.likelihood_params, T=T, ref_model=m, ref_logL=logL, ref_logPrior=pr).items():
            out[key + "/" + f] = np.asarray(v)
        print(key, "Nx", p.x.size, "logL", logL, "prior", pr)
    out["names"] = np.array(names)
    dst = os.path.join(ROOT, "tests", "golden", "golden_ref.npz")
    np.savez_compressed(dst, **out)
    print("wrote", dst, os.path.getsize(dst), "bytes")


if __name__ == "__main__":
    main()
