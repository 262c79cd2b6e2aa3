"""Developer check: fused-kernel time vs resident CTAs per SM (register-capped builds) and chain-group size."""
import sys, os, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import tamcmc_b200_loader
tb = tamcmc_b200_loader.load()
W = tb.workloads
here = os.path.dirname(os.path.abspath(__file__))
prob = W.s_global(Nx=1000000)
prob.y = np.abs(np.random.default_rng(1).standard_normal(prob.x.size)) + 0.1
C = 10
P, T = W.perturbed_params(prob, C), W.tcoefs(C)
for name in sys.argv[1:]:
    lib = tb.gpu.GpuLibrary(os.path.join(here, name)) if name != "default" else tb.load_library()
    md = tb.GpuModelDef(prob, tb.MalaConfig(Nchains=C), lib=lib)
    md.eval_upload(P, T)
    for cpb in (2, 3, 4, 5):
        md.set_tuning(cpb, 8, -1, 2); md.kernel_time(reset=1)
        for _ in range(20): md.eval_resident(C)
        n, ms, tot = md.kernel_time(reset=2)
        print("%-12s cpb=%d: eval kernel %.1f us" % (name, cpb, ms / n * 1e3))
    md.close()
