"""Developer check: step time of the global model on a 4096-bin spectrum = cost of everything but the fused kernel's bins."""
import sys, os, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import tamcmc_b200_loader
tb = tamcmc_b200_loader.load()
W = tb.workloads
p = W.s_global(Nx=4096, step=2.0)
p.y = np.abs(np.random.default_rng(1).standard_normal(p.x.size)) + 0.1
for learn in (True, False):
    md = tb.GpuModelDef(p, tb.MalaConfig(Nchains=10, lambda_temp=1.7, c0=2.0, Nt_learn=[100, 10**9] if learn else [10**8, 10**9], periods_learn=[1], dN_mixing=1, seed=7))
    ml = tb.GpuMALA(md)
    ml.execute(0, 300, record=False)
    md.sync(); md.timer_start(); ml.execute(300, 3000, record=False); ms = md.timer_stop_ms()
    md.kernel_time(reset=1); ml.execute(3300, 64, record=False); n, mk, _ = md.kernel_time(reset=2)
    print("learning=%s: %.2f us/step, fused kernel %.2f us" % (learn, ms / 3000 * 1e3, mk / n * 1e3))
    md.close()
