"""Developer check of the far-field path: logL with TG_FARFIELD on / off, fused-kernel time per chain count (developer scratch)."""
import sys, os, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import tamcmc_b200_loader
tb = tamcmc_b200_loader.load()
W = tb.workloads
def make_y(p):
    tmp = tb.GpuModelDef(p, tb.MalaConfig(Nchains=1)); m = tmp.call_model(0); tmp.close(); return m
prob = W.s_global(Nx=1000000, make_y=make_y)
res = {}
for C in [int(a) for a in (sys.argv[1:] or ["10", "80"])]:
    P, T = W.perturbed_params(prob, C), W.tcoefs(C)
    for mode in ("0", "1"):
        os.environ["TG_FARFIELD"] = mode
        md = tb.GpuModelDef(prob, tb.MalaConfig(Nchains=C, chains_per_ensemble=10 if C % 10 == 0 else 8))
        r = md.evaluate(P, T)
        md.eval_upload(P, T)
        for _ in range(3): md.eval_resident(C)
        md.kernel_time(reset=1)
        for _ in range(20): md.eval_resident(C)
        n, ms, tot = md.kernel_time(reset=2)
        res[(C, mode)] = r
        print("C=%4d farfield=%s: eval kernel %.1f us (%.2f us/chain) units %d" % (C, mode, ms / n * 1e3, ms / n * 1e3 / C, md.work_profile()[2]), flush=True)
        md.close()
    a, b = res[(C, "0")], res[(C, "1")]
    print("   max rel diff logL (ff on vs off): %.3e   status equal: %s" % (float(np.max(np.abs(a["logL"] - b["logL"]) / np.abs(a["logL"]))), np.array_equal(a["status"], b["status"])))
