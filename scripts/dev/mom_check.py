"""Developer check: background-only tiles through the moments of the data (TG_MOMENTS) against the bin-by-bin path."""
import sys, os, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import tamcmc_b200_loader
tb = tamcmc_b200_loader.load()
W = tb.workloads
prob = W.s_global(Nx=1000000)
prob.y = -np.log(W.philox_uniform(prob.x.size, 5)) * (0.2 + 10.0 / (1 + (prob.x / 300.0) ** 2))
for Cn in (10, 64):
    P, T = W.perturbed_params(prob, Cn, scale=1.0), W.tcoefs(Cn, 1.2)
    res = {}
    for mode in ("0", "1"):
        os.environ["TG_MOMENTS"] = mode
        md = tb.GpuModelDef(prob, tb.MalaConfig(Nchains=Cn))
        tune = [int(v) for v in os.environ.get('TG_TUNE', '0,0,0').split(',')]   # chains_per_block, heavy_group, light_subtiles
        if any(tune): md.set_tuning(chains_per_block=tune[0], heavy_group=tune[1], light_subtiles=tune[2])
        r = md.evaluate(P, T)
        md.eval_upload(P, T)
        for _ in range(3): md.eval_resident(Cn)
        md.sync()
        L = tune[2] if tune[2] else 8
        md.set_tuning(light_subtiles=L + 1); md.set_tuning(light_subtiles=L)      # rebuild the work items with the failure flags of the warm-up
        for _ in range(3): md.eval_resident(Cn)
        md.sync(); md.kernel_time(reset=1)
        for _ in range(20): md.eval_resident(Cn)
        n, ms, _ = md.kernel_time(reset=2)
        res[mode] = (r["logL"].copy(), ms / max(n, 1) * 1e3)
        md.close()
    d = np.abs(res["1"][0] - res["0"][0]) / np.abs(res["0"][0])
    print("C=%d: bin-by-bin %.1f us, with moments %.1f us; logL rel diff max %.2e" % (Cn, res["0"][1], res["1"][1], d.max()))
