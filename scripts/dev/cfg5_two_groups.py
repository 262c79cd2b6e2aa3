"""Developer check: 8 stars x 10 chains per GPU as ONE launch group vs TWO groups of 4 driven by two host threads (their streams
overlap on the GPU: one group's proposal / decision kernels run beside the other's fused kernel)."""
import sys, os, time, threading, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import tamcmc_b200_loader
tb = tamcmc_b200_loader.load(); W = tb.workloads
S, Cn, K = 8, 10, int(sys.argv[1]) if len(sys.argv) > 1 else 512
def make_y(p):
    tmp = tb.GpuModelDef(p, tb.MalaConfig(Nchains=1)); m = tmp.call_model(0); tmp.close(); return m
def stars():
    return [tb.GpuModelDef(W.s_global(Nx=1000000, seed=12345 + s, make_y=make_y),
                           tb.MalaConfig(Nchains=Cn, lambda_temp=1.7, c0=2.0, Nt_learn=[0, 10**9], periods_learn=[1], dN_mixing=1, seed=4242 + s)) for s in range(S)]
def timed(groups, i0, n):
    for m in sum([g.models for g in groups], []): m.sync()
    t0 = time.perf_counter()
    th = [threading.Thread(target=g.execute, args=(i0, n), kwargs=dict(record=False)) for g in groups]
    for t in th: t.start()
    for t in th: t.join()
    for m in sum([g.models for g in groups], []): m.sync()
    return (time.perf_counter() - t0) / n * 1e6
for split in (1, 2, 4):
    mds = stars()
    groups = [tb.GpuBatch(mds[k * (S // split):(k + 1) * (S // split)]) for k in range(split)]
    timed(groups, 10000, 64)
    ts = [timed(groups, 10064 + r * K, K) for r in range(3)]
    print("%d launch group(s) of %d stars: %.1f us per step of all %d chains (runs: %s)" % (split, S // split, min(ts), S * Cn, ", ".join("%.1f" % t for t in ts)), flush=True)
    for g in groups: g.close()
    for m in mds: m.close()
