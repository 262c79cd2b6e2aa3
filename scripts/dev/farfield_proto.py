"""Prototype (numpy, no GPU): inside the truncation windows, the components whose centre lies more than kappa half-widths of
a 256/512-bin micro-tile away (and whose window covers the micro-tile) plus the background form a smooth function of x;
one Chebyshev interpolant per (micro-tile, chain) replaces their per-bin evaluation and only the near components stay
exact.  Prints the number of near components and the worst interpolation error relative to the model value.
Result on cfg3 ("what comes next"): 13 nodes, kappa = 6: 1.8 near of 29 components, error 6e-15."""
import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import tamcmc_b200_loader
tb = tamcmc_b200_loader.load(); W = tb.workloads
prob = W.s_global()
p = prob.params; no = 14
H = p[:no]; V = np.r_[1.0, p[no:no + 3]]
f = p[no + 3:no + 3 + 4 * no].reshape(4, no)
Wd = p[no + 3 + 4 * no + 6:no + 3 + 4 * no + 6 + no]; a1 = 1.4
comps = np.array([(f[l, k] + m * a1, Wd[k], H[k] * V[l] / (2 * l + 1)) for l in range(4) for k in range(no) for m in range(-l, l + 1)])
x = prob.x
bg = lambda xx: 0.13392108 + 11.049588 / (1 + (1e-3 * 49.669854 * xx) ** 4) + 0.93569041 / (1 + (1e-3 * 1.3516447 * xx) ** 2)
lor = lambda xx, c: c[2] / (1 + 4 * ((xx - c[0]) / c[1]) ** 2)


def cheb(n):
    k = np.arange(n); Wm = (2.0 / n) * np.cos(np.outer(k, np.pi * (np.arange(n) + 0.5) / n)); Wm[0] *= 0.5
    return Wm, np.cos(np.pi * (np.arange(n) + 0.5) / n)


for MT in (256, 512):
    for n in (9, 13):
        Wm, tn = cheb(n)
        for kappa in (6, 10, 16):
            worst, nnear, nall = 0.0, [], []
            for t0 in range(300 * 1024, 300 * 1024 + 40 * MT, MT):
                xs = x[t0:t0 + MT].astype(np.longdouble); xm = (xs[0] + xs[-1]) / 2; xh = (xs[-1] - xs[0]) / 2
                inwin = np.abs(comps[:, 0] - float(xm)) < 110
                near = inwin & (np.abs(comps[:, 0] - float(xm)) < kappa * float(xh)); far = inwin & ~near
                nnear.append(near.sum()); nall.append(inwin.sum())
                sfar = lambda xx: bg(xx) + sum(lor(xx, c) for c in comps[far])
                co = Wm @ np.asarray(sfar(xm + xh * tn), dtype=np.float64)
                t = np.asarray((xs - xm) / xh, dtype=np.float64)
                ex = sfar(xs); tot = ex + sum(lor(xs, c) for c in comps[near])
                val = co @ np.cos(np.outer(np.arange(n), np.arccos(np.clip(t, -1, 1))))
                worst = max(worst, float(np.max(np.abs(val - np.asarray(ex, dtype=np.float64)) / np.asarray(tot, dtype=np.float64))))
            print("micro-tile %d bins, %d nodes, kappa %d: near %.1f of %.1f components, worst error %.1e" % (MT, n, kappa, np.mean(nnear), np.mean(nall), worst))
