"""Small driver for ncu captures of the fused kernel: cfg3, C chains (argv[1], default 10), 6 resident evaluations."""
import sys, os, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import tamcmc_b200_loader
tb = tamcmc_b200_loader.load()
W = tb.workloads
C = int(sys.argv[1]) if len(sys.argv) > 1 else 10
prob = W.s_global(Nx=1000000)
prob.y = np.abs(np.random.default_rng(1).standard_normal(prob.x.size)) + 0.1
P, T = W.perturbed_params(prob, C), W.tcoefs(C)
md = tb.GpuModelDef(prob, tb.MalaConfig(Nchains=C, chains_per_ensemble=10 if C % 10 == 0 else 8))
md.eval_upload(P, T)
for _ in range(6): md.eval_resident(C)
md.sync()
print("ok", md.work_profile())
md.close()
