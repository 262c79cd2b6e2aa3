"""Developer check (TG_PROFILE build): timeline of one step of cfg3 from %globaltimer stamps in the three kernels."""
import sys, os, ctypes as C, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import tamcmc_b200_loader
tb = tamcmc_b200_loader.load()
W = tb.workloads
here = os.path.dirname(os.path.abspath(__file__))
lib = tb.gpu.GpuLibrary(os.path.join(here, "libtamcmc_gpu_prof.so"))
prob = W.s_global(Nx=1000000)
prob.y = np.abs(np.random.default_rng(1).standard_normal(prob.x.size)) + 0.1
md = tb.GpuModelDef(prob, tb.MalaConfig(Nchains=10, lambda_temp=1.7, c0=2.0, Nt_learn=[100, 10**9], periods_learn=[1], dN_mixing=1, seed=4242), lib=lib)
ml = tb.GpuMALA(md)
ml.execute(0, 400, record=False)
for rep in range(3):
    ml.execute(400 + 160 * rep, 160, record=False); md.sync()
    tl = (C.c_longlong * 8)(); lib.lib.tg_dbg_timeline.argtypes = [C.POINTER(C.c_longlong)]; lib.lib.tg_dbg_timeline(tl)
    buf = (C.c_longlong * (8192 * 6))(); lib.lib.tg_dbg_ctas.argtypes = [C.POINTER(C.c_longlong)]; lib.lib.tg_dbg_ctas(buf)
    a = np.array(buf, dtype=np.int64).reshape(8192, 6); a = a[a[:, 5] > 0]; a = a[a[:, 4] >= a[:, 4].max() - 150000]
    e0, e1 = a[:, 4].min(), a[:, 5].max()
    t = np.array(tl[:], dtype=np.int64)
    us = lambda x: x / 1e3
    print("prev finish end -> propose start %.2f | propose %.2f | propose end -> first eval CTA %.2f | eval span %.2f | last eval CTA -> finish past wait %.2f (finish resident %.2f earlier) | finish %.2f | side kernel start +%.2f after propose start"
          % (us(t[0] - t[7]), us(t[1] - t[0]), us(e0 - t[1]), us(e1 - e0), us(t[3] - e1), us(t[3] - t[2]), us(t[4] - t[3]), us(t[5] - t[0])))
    print("   step total (propose start -> finish end) %.2f us" % us(t[4] - t[0]))
