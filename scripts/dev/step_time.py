"""Developer check: ms/step of tg_run with learning on/off and graph on/off."""
import sys, os, time, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import tamcmc_b200_loader
tb = tamcmc_b200_loader.load()
W = tb.workloads
prob = W.s_global(Nx=1000000)
prob.y = np.abs(np.random.default_rng(1).standard_normal(prob.x.size)) + 0.1
for learn in (True, False):
    for graph in (1,):
        cfg = tb.MalaConfig(Nchains=10, lambda_temp=1.7, c0=2.0, Nt_learn=[100, 10**9] if learn else [10**8, 10**9], periods_learn=[1], dN_mixing=1, seed=4242)
        md = tb.GpuModelDef(prob, cfg)
        md.set_tuning(0, 0, graph)
        mala = tb.GpuMALA(md)
        mala.execute(0, 300, record=False)
        md.sync(); md.timer_start()
        mala.execute(300, 1000, record=False)
        ms = md.timer_stop_ms()
        md.sync(); md.timer_start()
        mala.execute(1300, 1000, record=False)
        ms2 = md.timer_stop_ms()
        print("learning=%s graph=%d: %.4f %.4f ms/step" % (learn, graph, ms / 1000, ms2 / 1000))
        md.close()
