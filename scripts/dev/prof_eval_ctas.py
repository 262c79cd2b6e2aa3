"""Developer check: per-CTA timeline of the fused kernel (TG_PROFILE build)."""
import sys, os, ctypes as C, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import tamcmc_b200_loader
tb = tamcmc_b200_loader.load()
W = tb.workloads
here = os.path.dirname(os.path.abspath(__file__))
lib = tb.gpu.GpuLibrary(os.path.join(here, sys.argv[1] if len(sys.argv) > 1 else "libtamcmc_gpu_prof.so"))
prob = W.s_global(Nx=1000000)
prob.y = np.abs(np.random.default_rng(1).standard_normal(prob.x.size)) + 0.1
Cn = int(os.environ.get('TG_C', '10'))
P, T = W.perturbed_params(prob, Cn), W.tcoefs(Cn)
md = tb.GpuModelDef(prob, tb.MalaConfig(Nchains=Cn), lib=lib)
tune = [int(v) for v in os.environ.get('TG_TUNE', '0,0,0').split(',')]   # chains_per_block, heavy_group, light_subtiles
md.set_tuning(chains_per_block=tune[0], heavy_group=tune[1], light_subtiles=tune[2])
md.eval_upload(P, T)
for _ in range(5): md.eval_resident(Cn)
md.sync()
lib.lib.tg_sync(md.h)
buf = (C.c_longlong * (8192 * 6))()
lib.lib.tg_dbg_ctas.argtypes = [C.POINTER(C.c_longlong)]
lib.lib.tg_dbg_ctas(buf)
a = np.array(buf, dtype=np.int64).reshape(8192, 6); a = a[a[:, 5] > 0]; a = a[a[:, 4] >= a[:, 4].max() - 2000000]   # last launch only
t0, t1, t2 = a[:, 0], a[:, 1], a[:, 2]
sm = a[:, 3] >> 32; tile = a[:, 3] & 0xffffffff
g0, g1 = a[:, 4], a[:, 5]
dur = (g1 - g0) / 1e3
print("kernel span %.1f us (globaltimer)" % ((g1.max() - g0.min()) / 1e3))
print("CTA duration us: min %.2f median %.2f mean %.2f p90 %.2f max %.2f" % (dur.min(), np.median(dur), dur.mean(), np.percentile(dur, 90), dur.max()))
ph1 = (t1 - t0) / 1965.0
print("phase-1 us: median %.2f mean %.2f max %.2f" % (np.median(ph1), ph1.mean(), ph1.max()))
order = np.argsort(-dur)
ph2 = (t2 - t1) / 1965.0
nsubs = np.round(ph2 / np.median(ph2[dur < np.median(dur) * 1.5]) ).astype(int)
for lo_, hi_ in ((0, 350000), (350000, 10**9)):
    sel = (tile >= lo_) & (tile < hi_)
    print("items with first bin in [%d,%d): n=%d, dur median %.2f mean %.2f" % (lo_, hi_, sel.sum(), np.median(dur[sel]), dur[sel].mean()))
print("top CTAs (dur, tile, sm, start):", [(round(float(dur[k]), 1), int(tile[k]), int(sm[k]), round(float((g0[k] - g0.min()) / 1e3), 1)) for k in order[:8]])
busy = np.zeros(148)
for k in range(len(a)): busy[sm[k]] += dur[k]
print("per-SM sum of CTA durations us: min %.1f mean %.1f max %.1f (2 resident -> /2)" % (busy.min(), busy.mean(), busy.max()))
last = np.array([(g1[sm == s].max() - g0.min()) / 1e3 for s in range(148) if np.any(sm == s)])
print("per-SM finish time us: min %.1f mean %.1f max %.1f" % (last.min(), last.mean(), last.max()))
heavy = dur > 3 * np.median(dur)
print("heavy CTAs: %d, mean dur %.1f us; light: %d, mean dur %.2f us" % (heavy.sum(), dur[heavy].mean(), (~heavy).sum(), dur[~heavy].mean()))
# histogram
h, e = np.histogram(dur, bins=[0, 1, 2, 3, 4, 6, 8, 12, 16, 24, 32, 48, 64, 1000])
print("hist", list(zip(e[:-1].tolist(), h.tolist())))

f = 1965.0
setup = (t1 - t0) / f
print("setup (phase 0) us: mean %.2f  share of CTA time %.3f" % (setup.mean(), setup.sum() / dur.sum()))
start = (g0 - g0.min()) / 1e3; end = (g1 - g0.min()) / 1e3
print("CTA start times us: first wave (<1us) %d CTAs, last start %.1f; last end %.1f" % ((start < 1).sum(), start.max(), end.max()))
slots_busy = dur.sum() / (148 * 2)
print("sum of CTA durations / 296 slots = %.1f us  (efficiency vs span %.3f)" % (slots_busy, slots_busy / end.max()))
# idle tail per SM: span - finish time of the SM
print("idle tail per SM us: mean %.2f max %.2f" % ((end.max() - last).mean(), (end.max() - last).max()))
# far-field path: where a unit's time goes (thread 0's clock: stage + barrier + fetch issue | interpolant + exact records | background, likelihood, running sums)
fb = (C.c_longlong * (8192 * 4))()
lib.lib.tg_dbg_ffstats.argtypes = [C.POINTER(C.c_longlong)]
lib.lib.tg_dbg_ffstats(fb)
ffs = np.array(fb, dtype=np.int64).reshape(8192, 4)
full = np.array(buf, dtype=np.int64).reshape(8192, 6)
ok = (full[:, 5] > 0) & (ffs[:, 3] > 0)
hv = ffs[ok]; dd = (full[ok, 5] - full[ok, 4]) / 1e3; tot = (full[ok, 2] - full[ok, 0]) / f
print("far-field units: %d, %.1f (pair, chain) steps each, duration mean %.2f us (clock %.2f): in the steps, mean over warps %.2f us, slowest warp %.2f us (staging waits %.2f us per warp)"
      % (ok.sum(), hv[:, 3].mean(), dd.mean(), tot.mean(), hv[:, 0].mean() / 8 / f, hv[:, 1].mean() / f, hv[:, 2].mean() / 8 / f))
# units outside the far-field path (background-only: moment tiles, bin-by-bin runs): how long, where
oth = (full[:, 5] > 0) & (ffs[:, 3] == 0)
do = (full[oth, 5] - full[oth, 4]) / 1e3; to = full[oth, 3] & 0xffffffff; so = (full[oth, 4] - full[full[:, 5] > 0, 4].min()) / 1e3
order = np.argsort(-do)
print("other units: %d, duration mean %.2f max %.2f; sum %.0f us (far-field units: sum %.0f us)" % (oth.sum(), do.mean(), do.max(), do.sum(), dd.sum()))
print("longest other units (dur us, first bin, start us):", [(round(float(do[k]), 1), int(to[k]), round(float(so[k]), 1)) for k in order[:16]])
ffd = (full[ok, 5] - full[ok, 4]) / 1e3; fft = full[ok, 3] & 0xffffffff; ffst = (full[ok, 4] - full[full[:, 5] > 0, 4].min()) / 1e3
order = np.argsort(-ffd)
print("longest far-field units (dur us, first bin, start us, mean warp, slowest warp):", [(round(float(ffd[k]), 1), int(fft[k]), round(float(ffst[k]), 1), round(float(hv[k, 0] / 8 / f), 1), round(float(hv[k, 1] / f), 1)) for k in order[:8]])
