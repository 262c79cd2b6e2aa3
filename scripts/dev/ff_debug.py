import sys, os, ctypes as C, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import tamcmc_b200_loader
tb = tamcmc_b200_loader.load(); W = tb.workloads
def make_y(p):
    tmp = tb.GpuModelDef(p, tb.MalaConfig(Nchains=1)); m = tmp.call_model(0); tmp.close(); return m
prob = W.s_global(Nx=1000000, make_y=make_y)
Cn = int(sys.argv[1]) if len(sys.argv) > 1 else 10
P, T = W.perturbed_params(prob, Cn), W.tcoefs(Cn)
md = tb.GpuModelDef(prob, tb.MalaConfig(Nchains=Cn))
r = md.evaluate(P, T)
out = (C.c_long * 6)()
lib = md.lib.lib
lib.tg_dbg_ff_summary.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_long)]
rc = lib.tg_dbg_ff_summary(md.h, Cn, out)
print("rc", rc, "micro-tiles %d fallbacks %d records %d (%.2f per micro-tile, max %d) far comps %.1f per micro-tile, noise folded %d" % (out[0], out[1], out[2], out[2] / max(out[0], 1), out[5], out[3] / max(out[0], 1), out[4]))
print(md.work_profile())
