"""Developer check (run under gpurun): parity vs the oracle + a few sampler steps + timings."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import tamcmc_b200_loader
tb = tamcmc_b200_loader.load()
from oracle.oracle import Oracle
W = tb.workloads
O = Oracle()
lib = tb.load_library()
print("fp64 peak TFLOP/s:", lib.measure_fp64_peak(0))

def rel(a, b):
    return np.max(np.abs(a - b) / np.maximum(np.abs(b), 1e-300))

for mk in (lambda: W.s_local(), lambda: W.s_global(Nx=200000), lambda: W.s_global(Nx=1000000)):
    prob = mk()
    st, m0, L0 = O.model(prob.model_case, prob.params, prob.plength, prob.x)
    prob.y = m0 * (-np.log(W.philox_uniform(m0.size, 12345)))
    C = 10
    T = W.tcoefs(C)
    md = tb.GpuModelDef(prob, tb.MalaConfig(Nchains=C))
    P = W.perturbed_params(prob, C)
    ref = O.generate_models(prob, P, T, want_model=True)
    t = time.time(); r = md.evaluate(P, T, want_model=True); dt = time.time() - t
    print(prob.name, prob.x.size, "status", r["status"], ref["status"])
    print("  model max rel", rel(r["model"], ref["model"]))
    print("  logL rel", np.abs(r["logL"] - ref["logL"]) / np.abs(ref["logL"]))
    print("  prior abs", np.abs(r["logPrior"] - ref["logPrior"]))
    L, K = md.last_work(C); print("  work L", L[:3], ref["L"][:3], "K", K, " eval wall %.1f ms" % (dt * 1e3))
    # resident timing
    md.eval_upload(P, T); md.kernel_time(reset=1)
    for _ in range(3): md.eval_resident(C)
    md.sync(); md.kernel_time(reset=1)
    t = time.time()
    for _ in range(20): md.eval_resident(C)
    md.sync(); dt = (time.time() - t) / 20
    n, ms, tot = md.kernel_time(reset=2)
    print("  resident eval: %.1f us/batch wall; eval kernel %.1f us avg over %d launches" % (dt * 1e6, ms / n * 1e3, n))
    for ls in (2, 4, 8):
        for hg in (1, 2, 3, 5):
            md.set_tuning(5, ls, -1, hg); md.kernel_time(reset=1)
            for _ in range(10): md.eval_resident(C)
            n, ms, tot = md.kernel_time(reset=2)
            print("    light_sub=%d heavy_group=%d: eval kernel %.1f us" % (ls, hg, ms / n * 1e3))
    md.set_tuning(5, 4, -1, 3)
    # sampler
    mala = tb.GpuMALA(md)
    t = time.time(); out = mala.execute(0, 64); dt = time.time() - t
    print("  run 64 steps: %.2f ms/step; acceptance" % (dt / 64 * 1e3), out["moved"].mean(axis=0))
    t = time.time(); out = mala.execute(64, 512); dt = time.time() - t
    print("  run 512 steps: %.3f ms/step  chain-steps/s %.0f; acc" % (dt / 512 * 1e3, 512 * C / dt), out["moved"].mean(axis=0), "swaps", out["swap_done"].mean())
    print("  logL cur", md.logLikelihood[:3], "status any", np.bitwise_or.reduce(out["status"].ravel()))
    md.close()
