import sys, os, ctypes as C, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import tamcmc_b200_loader
tb = tamcmc_b200_loader.load()
W = tb.workloads
lib = tb.gpu.GpuLibrary(os.path.join(os.path.dirname(os.path.abspath(__file__)), "libtamcmc_gpu_prof.so"))
prob = W.s_global(Nx=1000000, model_case=3, seed=12345, make_y=None) if False else W.s_global(Nx=1000000)
prob.y = np.ones_like(prob.x)
cfg = tb.MalaConfig(Nchains=10, Nt_learn=[5, 10**9], periods_learn=[1], c0=2.0)
md = tb.GpuModelDef(prob, cfg, lib=lib)
md.set_tuning(0, 0, 0)
tb.GpuMALA(md).execute(0, 20, record=False)
out = (C.c_longlong * 32)()
lib.lib.tg_dbg_clocks.argtypes = [C.POINTER(C.c_longlong)]
lib.lib.tg_dbg_clocks(out)
v = np.array(out[:10], dtype=np.int64)
names = ["start", "rng done", "matvec done", "scatter done", "stage params", "A done", "B done", "C done", "D done", "E done"]
for k in range(1, 10):
    print("%-14s +%7d cycles (%.2f us)" % (names[k], v[k] - v[k - 1], (v[k] - v[k - 1]) / 1965.0))
print("total %.2f us" % ((v[9] - v[0]) / 1965.0))

out2 = (C.c_longlong * 16)()
lib.lib.tg_dbg_specs.argtypes = [C.POINTER(C.c_longlong)]
lib.lib.tg_dbg_specs(out2)
w = np.array(out2[:6], dtype=np.int64)
for k, nm in enumerate(["mu update", "cov update", "stage matrix", "cholesky", "write factor"]):
    print("spec %-14s +%7d cycles (%.2f us)" % (nm, w[k + 1] - w[k], (w[k + 1] - w[k]) / 1965.0))
