"""Step time of one shard of S-10M as a plain handle (no exchange): argv world rank"""
import sys, os, copy, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import tamcmc_b200_loader
tb = tamcmc_b200_loader.load(); W, S = tb.workloads, tb.sharding
world, rank = int(sys.argv[1]), int(sys.argv[2])
prob = W.s_global(Nx=10_000_000, step=8500.0 / 10_000_000)
desc, sl = S.balanced_shard_descriptor(prob, world, rank, 17.0)
sub = copy.copy(prob); sub.x, sub.y = np.ascontiguousarray(prob.x[sl]), np.ascontiguousarray(prob.y[sl])
md = tb.GpuModelDef(sub, tb.MalaConfig(Nchains=10, Nt_learn=[10, 1000000], periods_learn=[1], c0=1.7, seed=2024), shard=desc if world > 1 else None)
m = tb.GpuMALA(md)
m.execute(0, 72, record=False)
md.timer_start(); m.execute(72, 200, record=False); ms = md.timer_stop_ms()
print("world %d rank %d: %.1f us per step (plain handle, graphs)" % (world, rank, ms / 200 * 1e3))
md.kernel_time(reset=1); m.execute(272, 50, record=False); n, mse, _ = md.kernel_time(reset=2); print("   eval kernel %.1f us, pre-pass %.1f us" % (mse / n * 1e3, md.prepass_time() / n * 1e3))
