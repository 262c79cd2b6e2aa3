"""work profile of cfg3 at the stages bench.py goes through (developer scratch)"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
import tamcmc_b200_loader
tb = tamcmc_b200_loader.load()
W = tb.workloads
def make_y(p):
    tmp = tb.GpuModelDef(p, tb.MalaConfig(Nchains=1)); m = tmp.call_model(0); tmp.close(); return m
p = W.s_global(Nx=1_000_000, model_case=3, seed=12345, make_y=make_y)
md = tb.GpuModelDef(p, tb.MalaConfig(Nchains=10, lambda_temp=1.7, c0=2.0, Nt_learn=[0, 10**9], periods_learn=[1], dN_mixing=1, seed=4242))
print("after create", md.work_profile())
r = tb.GpuMALA(md)
r.execute(0, 32, record=False); print("after 32", md.work_profile())
md.kernel_time(reset=2); r.execute(32, 20, record=False); print("after 20", md.work_profile())
md.kernel_time(reset=1); r.execute(52, 20, record=False); print(md.kernel_time(reset=2)); print("after timing", md.work_profile())
md.evaluate(md.params, md.Tcoefs); print("after eval", md.work_profile())
r.execute(72, 200, record=False); print("after 200", md.work_profile())
md.kernel_time(reset=1); r.execute(272, 100, record=False); n, ms, _ = md.kernel_time(reset=2); print("kernel us", ms / n * 1e3)
