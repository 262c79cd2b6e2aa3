"""Developer check: fused-kernel time and roofline fraction vs number of chains (cfg3 -> cfg4)."""
import sys, os, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import tamcmc_b200_loader
tb = tamcmc_b200_loader.load()
W = tb.workloads
prob = W.s_global(Nx=1000000)
lib = tb.load_library()
peak, _ = lib.measure_fp64_peak(0)
print("FP64 peak %.2f TFLOP/s" % peak)
for C in [int(a) for a in (sys.argv[1:] or ["10", "20", "40", "80", "160", "320", "1024"])]:
    P, T = W.perturbed_params(prob, C), W.tcoefs(C)
    md = tb.GpuModelDef(prob, tb.MalaConfig(Nchains=C, chains_per_ensemble=10 if C % 10 == 0 else 8))
    md.eval_upload(P, T)
    for _ in range(3): md.eval_resident(C)
    md.kernel_time(reset=1)
    for _ in range(10): md.eval_resident(C)
    n, ms, tot = md.kernel_time(reset=2)
    L, K = md.last_work(C)
    F = float(np.sum(8.0 * L + prob.x.size * (5.0 * K + 5.0)))
    t = ms / n * 1e-3
    print("C=%4d: eval kernel %.1f us  (%.2f us/chain)  achieved %.2f TFLOP/s  frac %.3f" % (C, t * 1e6, t * 1e6 / C, F / t / 1e12, F / t / 1e12 / peak))
    md.close()
