"""Developer check: where the time of one tg_eval call goes (host enqueue, GPU work, sync)."""
import sys, os, time, ctypes as C, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import tamcmc_b200_loader
tb = tamcmc_b200_loader.load()
W = tb.workloads
prob = W.s_global(Nx=1000000)
prob.y = np.abs(np.random.default_rng(1).standard_normal(prob.x.size)) + 0.1
cfg = tb.MalaConfig(Nchains=10, lambda_temp=1.7, seed=1)
md = tb.GpuModelDef(prob, cfg)
lib = md.lib.lib
B = 10
P = W.perturbed_params(prob, B); T = W.tcoefs(B)
o = [np.empty(B) for _ in range(3)]; st = np.zeros(B, np.int32)
args = (md.h, P.ctypes.data, B, T.ctypes.data, o[0].ctypes.data, o[1].ctypes.data, o[2].ctypes.data, st.ctypes.data, None)
for _ in range(20): lib.tg_eval(*args)
N = 500
t0 = time.perf_counter()
for _ in range(N): lib.tg_eval(*args)
t1 = time.perf_counter()
print("tg_eval full call: %.1f us" % ((t1 - t0) / N * 1e6))
md.sync()
t0 = time.perf_counter()
for _ in range(N): lib.tg_eval_upload(md.h, P.ctypes.data, B, T.ctypes.data); lib.tg_eval_resident(md.h, B)
te = time.perf_counter()
md.sync()
t1 = time.perf_counter()
print("upload+resident, no fetch: enqueue %.1f us/call, total %.1f us/call (GPU back-to-back)" % ((te - t0) / N * 1e6, (t1 - t0) / N * 1e6))
md.sync()
t0 = time.perf_counter()
for _ in range(N): lib.tg_eval_resident(md.h, B)
md.sync()
t1 = time.perf_counter()
print("resident only: %.1f us/call (GPU back-to-back)" % ((t1 - t0) / N * 1e6))
md.kernel_time(reset=1)
for _ in range(100): lib.tg_eval_resident(md.h, B)
n, ms, _ = md.kernel_time(reset=2)
print("tg_k_eval alone: %.1f us" % (ms / n * 1e3))
md.close()
