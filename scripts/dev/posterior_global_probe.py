"""Developer probe: widths of the global-fit posterior on the GPU for run lengths the reference cannot afford."""
import sys, os, numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import tamcmc_b200_loader
tb = tamcmc_b200_loader.load(); W = tb.workloads
import test_posterior_gpu as T
z = dict(np.load(os.path.join(ROOT, "tests", "golden", "posterior_global.npz"), allow_pickle=False))
relax = z["relax"].astype(np.int32)
prob = W.Problem(name="global", model_case=int(z["model_case"]), like_case=int(z["like_case"]), prior_case=int(z["prior_case"]), plength=z["plength"].astype(np.int32),
                 params=z["params"], relax=relax, priors=z["priors"], priors_switch=z["priors_switch"].astype(np.int32), extra_priors=z["extra_priors"],
                 x=z["x"], y=z["y"], likelihood_params=float(z["likelihood_params"]), names=[str(n) for n in z["names"]])
err = np.ones(relax.size); err[relax == 1] = z["init_err"]
prob.err_frac, prob.err_off = np.zeros(relax.size), err
names = [n for n, f in zip(prob.names, relax) if f == 1]
for ns in ((2000, 6000, 12000), (2000, 6000, 200000), (2000, 30000, 200000)):
    z2 = dict(z); z2["nsamples"] = np.array(ns)
    res = [T._phases(tb, prob, z2, 911 * (r + 1)) for r in range(3)]
    sd = np.array([v.std(0) for v, _ in res]).mean(0); mn = np.array([v.mean(0) for v, _ in res])
    print("nsamples", ns, "acceptance", [round(float(a), 3) for _, a in res])
    for k in (0, 5, 8, 20, 28, 33, 34, 35, 36, 37):
        print("   %-22s width %.4g (ref 12000-sample runs %.4g)  run scatter of the mean %.3g (ref %.3g)" % (names[k], sd[k], z["ref_std"].mean(0)[k], mn[:, k].std(ddof=1), z["ref_mean"][:, k].std(ddof=1)))
