"""Developer check: far radius of the far-field path (TG_FF_R2) against accuracy and the counts of exact records."""
import sys, os, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import tamcmc_b200_loader
tb = tamcmc_b200_loader.load(); W = tb.workloads
def make_y(p):
    tmp = tb.GpuModelDef(p, tb.MalaConfig(Nchains=1)); m = tmp.call_model(0); tmp.close(); return m
prob = W.s_global(Nx=1000000, make_y=make_y)
C = 10
P, T = W.perturbed_params(prob, C), W.tcoefs(C)
os.environ["TG_FARFIELD"] = "0"
md = tb.GpuModelDef(prob, tb.MalaConfig(Nchains=C)); ref = md.evaluate(P, T)["logL"]; md.close()
os.environ["TG_FARFIELD"] = "1"
for r2 in sys.argv[1:]:
    os.environ["TG_FF_R2"] = r2
    md = tb.GpuModelDef(prob, tb.MalaConfig(Nchains=C))
    r = md.evaluate(P, T)["logL"]
    prof = md.farfield_profile(C)
    print("R2=%-6s max rel diff vs exact path %.3e   micro-tiles %d marked %d records %.2f per micro-tile, far %.1f" % (r2, float(np.max(np.abs(r - ref) / np.abs(ref))), prof[0], prof[1], prof[2] / prof[0], prof[3] / prof[0]))
    md.close()
