"""Developer sweep of the scheduling knobs of the fused kernel with the far-field path (developer scratch)."""
import sys, os, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import tamcmc_b200_loader
tb = tamcmc_b200_loader.load()
W = tb.workloads
def make_y(p):
    tmp = tb.GpuModelDef(p, tb.MalaConfig(Nchains=1)); m = tmp.call_model(0); tmp.close(); return m
prob = W.s_global(Nx=1000000, make_y=make_y)
ref = {}
KEYS = ("TG_FARFIELD", "TG_HEAVY_SUB", "TG_HYBRID_PCT", "TG_HYBRID_GROUP", "TG_HEAVY_GROUP", "TG_FF_R2")
def run(C, env):
    for k in KEYS:
        os.environ.pop(k, None)
    os.environ.update(env)
    P, T = W.perturbed_params(prob, C), W.tcoefs(C)
    md = tb.GpuModelDef(prob, tb.MalaConfig(Nchains=C, chains_per_ensemble=10 if C % 10 == 0 else 8))
    r = md.evaluate(P, T)
    md.eval_upload(P, T)
    for _ in range(3): md.eval_resident(C)
    md.kernel_time(reset=1)
    for _ in range(20): md.eval_resident(C)
    n, ms, tot = md.kernel_time(reset=2)
    if C not in ref: ref[C] = r["logL"]
    d = float(np.max(np.abs(r["logL"] - ref[C]) / np.abs(ref[C])))
    print("C=%4d %-70s kernel %.1f us (%.2f us/chain) units %d  dlogL %.1e" % (C, str(env), ms / n * 1e3, ms / n * 1e3 / C, md.work_profile()[2], d), flush=True)
    md.close()
if __name__ == "__main__":
    run(10, {"TG_FARFIELD": "0"})
    for hs in ("8", "4", "2"):
        for pct, hg in (("0", "1"), ("100", "2"), ("70", "3"), ("100", "3"), ("100", "5")):
            run(10, {"TG_HEAVY_SUB": hs, "TG_HYBRID_PCT": pct, "TG_HYBRID_GROUP": hg})
    run(10, {"TG_HEAVY_SUB": "8", "TG_HYBRID_PCT": "0", "TG_FF_R2": "25"})
    run(10, {"TG_HEAVY_SUB": "8", "TG_HYBRID_PCT": "0", "TG_FF_R2": "49"})
    run(80, {"TG_FARFIELD": "0"})
    for hs in ("8", "4"):
        for hg in ("2", "3", "4", "5"):
            run(80, {"TG_HEAVY_SUB": hs, "TG_HYBRID_PCT": "100", "TG_HYBRID_GROUP": hg})
    run(80, {"TG_HEAVY_SUB": "8", "TG_HYBRID_PCT": "0"})
