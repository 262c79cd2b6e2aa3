"""Developer check: step time of the S-local configuration (10k bins, 10 chains)."""
import sys, os, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import tamcmc_b200_loader
tb = tamcmc_b200_loader.load()
W = tb.workloads
def make_y(p):
    t = tb.GpuModelDef(p, tb.MalaConfig(Nchains=1)); m = t.call_model(0); t.close(); return m
p1 = W.s_local(make_y=make_y)
md = tb.GpuModelDef(p1, tb.MalaConfig(Nchains=10, lambda_temp=1.7, c0=2.0, Nt_learn=[100, 10**9], periods_learn=[1], dN_mixing=1, seed=7))
ml = tb.GpuMALA(md)
ml.execute(0, 300, record=False)
md.sync(); md.timer_start(); ml.execute(300, 3000, record=False); ms = md.timer_stop_ms()
md.kernel_time(reset=1); ml.execute(3300, 64, record=False); n, mk, _ = md.kernel_time(reset=2)
print("S-local: %.2f us/step, fused kernel %.2f us" % (ms / 3000 * 1e3, mk / n * 1e3))
