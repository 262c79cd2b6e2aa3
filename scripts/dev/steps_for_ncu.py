"""cfg3 with C chains (argv[1]) and S stars (argv[2]): 48 sampler steps without graphs (for an ncu launch list)."""
import sys, os, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import tamcmc_b200_loader
tb = tamcmc_b200_loader.load(); W = tb.workloads
C = int(sys.argv[1]) if len(sys.argv) > 1 else 10
S = int(sys.argv[2]) if len(sys.argv) > 2 else 1
def make_y(p):
    tmp = tb.GpuModelDef(p, tb.MalaConfig(Nchains=1)); m = tmp.call_model(0); tmp.close(); return m
mds = []
for s in range(S):
    p = W.s_global(Nx=1000000, seed=12345 + s, make_y=make_y)
    md = tb.GpuModelDef(p, tb.MalaConfig(Nchains=C, lambda_temp=1.7, c0=2.0, Nt_learn=[0, 10**9], periods_learn=[1], dN_mixing=1, seed=4242 + s))
    md.set_tuning(use_graph=0)
    mds.append(md)
run = tb.GpuBatch(mds) if S > 1 else tb.GpuMALA(mds[0])
run.execute(10000, 48, record=False)
print("ok")
