"""Developer check: steps per second of the sampler on the shipped star's slice (configs[0]: 973 bins, 9 free parameters, 10 chains),
the problem the reference's own front end built (tests/golden/posterior_cfg1.npz)."""
import sys, os, numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import tamcmc_b200_loader
tb = tamcmc_b200_loader.load(); W = tb.workloads
z = np.load(os.path.join(ROOT, "tests", "golden", "posterior_cfg1.npz"), allow_pickle=False)
relax = z["relax"].astype(np.int32)
prob = W.Problem(name="cfg1", model_case=int(z["model_case"]), like_case=int(z["like_case"]), prior_case=int(z["prior_case"]), plength=z["plength"].astype(np.int32),
                 params=z["params"], relax=relax, priors=z["priors"], priors_switch=z["priors_switch"].astype(np.int32), extra_priors=z["extra_priors"],
                 x=z["x"], y=z["y"], likelihood_params=float(z["likelihood_params"]))
md = tb.GpuModelDef(prob, tb.MalaConfig(Nchains=10, lambda_temp=1.7, c0=2.0, Nt_learn=[0, 10**9], periods_learn=[1], dN_mixing=1, seed=1))
run = tb.GpuMALA(md)
run.execute(10000, 64, record=False)
for n in (1000, 4000):
    run.execute(20000, n, record=False)
    ms = md.last_run_ms()
    print("cfg1 (Nx=%d, Nv=%d, 10 chains): %d steps, %.2f us per step, %.0f chain-steps/s" % (prob.x.size, md.Nvars, n, ms / n * 1e3, 10 * n / (ms * 1e-3)))
md.close()
