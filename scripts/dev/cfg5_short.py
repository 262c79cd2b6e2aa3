"""Developer check: S stars x 10 chains as one launch group under the driver's short protocol -- 32 warm-up steps, then calls of 64
steps timed one by one; prints the counters of partitions / graph captures / probes / argument uploads after every call."""
import sys, os, ctypes as C, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import tamcmc_b200_loader
tb = tamcmc_b200_loader.load(); W = tb.workloads
S = int(sys.argv[1]) if len(sys.argv) > 1 else 8
Cn = int(sys.argv[2]) if len(sys.argv) > 2 else 10
def make_y(p):
    tmp = tb.GpuModelDef(p, tb.MalaConfig(Nchains=1)); m = tmp.call_model(0); tmp.close(); return m
mds = []
for s in range(S):
    p = W.s_global(Nx=1000000, seed=12345 + s, make_y=make_y)
    mds.append(tb.GpuModelDef(p, tb.MalaConfig(Nchains=Cn, lambda_temp=1.7, c0=2.0, Nt_learn=([10**8, 10**9] if os.environ.get("TG_NOLEARN") else [0, 10**9]), periods_learn=[1], dN_mixing=1, seed=4242 + s)))
run = tb.GpuBatch(mds) if S > 1 else tb.GpuMALA(mds[0])
tm = run if S > 1 else mds[0]
lib = mds[0].lib.lib
cnt = (C.c_long * 4)()
def counters():
    lib.tg_dbg_counters(cnt); return list(cnt)
print("after create", counters())
i = 10000
run.execute(i, 32, record=False); i += 32
print("after warm-up", counters())
for n in (64, 64, 64, 20, 20, 1000, 64, 20, 20, 1000, 64, 20, 20):
    mds[0].sync(); tm.timer_start(); run.execute(i, n, record=False); ms = tm.timer_stop_ms(); i += n
    print("%5d steps: %.1f us/step  counters %s" % (n, ms / n * 1e3, counters()))
