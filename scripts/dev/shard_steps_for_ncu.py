"""One shard of S-10M (argv: world rank) as a plain handle, 24 sampler steps without graphs (for an ncu launch list)."""
import sys, os, copy, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import tamcmc_b200_loader
tb = tamcmc_b200_loader.load(); W, S = tb.workloads, tb.sharding
world, rank = int(sys.argv[1]), int(sys.argv[2])
prob = W.s_global(Nx=10_000_000, step=8500.0 / 10_000_000)
desc, sl = S.balanced_shard_descriptor(prob, world, rank, 17.0)
sub = copy.copy(prob); sub.x, sub.y = np.ascontiguousarray(prob.x[sl]), np.ascontiguousarray(prob.y[sl])
md = tb.GpuModelDef(sub, tb.MalaConfig(Nchains=10, Nt_learn=[10, 1000000], periods_learn=[1], c0=1.7, seed=2024), shard=desc if world > 1 else None)
md.set_tuning(use_graph=0)
print("bins", sub.x.size, "profile", md.work_profile())
tb.GpuMALA(md).execute(0, 24, record=False)
print("ok")
