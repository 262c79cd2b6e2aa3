"""Multi-GPU partitioning of the hot path; host logic, no collectives here.

(i)  Independent tempered ensembles (stars, slices, replicated ensembles): ensemble e -> rank e % G,
     every chain of an ensemble on one GPU (tempering swaps stay intra-GPU), no data-path collective.
     The reference does the same with a SLURM array of processes (scripts/slurm/job_array.sh).
(ii) Bin-sharded variant for very long spectra: contiguous bin ranges per rank; every rank holds the
     same chain state and Philox streams, computes per-chain partial (sum y/m, sum ln m) over its
     bins and the 2*C doubles are all-reduced (sum) once per step.
"""
import numpy as np


def partition_units(n_units, world, rank):
    """Round-robin assignment of independent ensembles/stars to ranks: unit e -> rank e % world."""
    return [e for e in range(n_units) if e % world == rank]


def bin_shard(Nx, world, rank):
    """Contiguous [lo, hi) bin range of a rank; sizes differ by at most one bin."""
    base, rem = divmod(Nx, world)
    lo = rank * base + min(rank, rem)
    hi = lo + base + (1 if rank < rem else 0)
    return lo, hi


def shard_descriptor(x, world, rank):
    """Arguments of GpuModelDef(shard=...) + the local slice for the bin-sharded variant."""
    lo, hi = bin_shard(x.size, world, rank)
    return dict(Nx_global=int(x.size), shard_lo=int(lo), x_first=float(x[0]), x_second=float(x[1]),
                x_last=float(x[-1])), slice(lo, hi)


def window_region(prob, margin=1.0):
    """[lo, hi] in x units that holds every truncation window of the initial parameter vector of a global / local MS
    problem (build_lorentzian.cpp:377-412: half-width c*(l*f_s+Gamma)), from the parameter layout of
    io_ms_global.cpp:996-1050.  A scheduling estimate only."""
    pl, p = [int(v) for v in prob.plength], np.asarray(prob.params, dtype=np.float64)
    o_f = pl[0] + pl[1]
    nf = sum(pl[2:6])
    o_s = o_f + nf
    o_w = o_s + pl[6]
    f = p[o_f:o_f + nf]
    if f.size == 0:
        return float(prob.x[0]), float(prob.x[-1])
    a1 = max(abs(p[o_s]), p[o_s + 3] ** 2 + p[o_s + 4] ** 2, 1.0) if pl[6] >= 5 else 1.0
    gam = max(float(np.max(np.abs(p[o_w:o_w + pl[7]]))) if pl[7] else 1.0, 1.0)
    c = p[-2] if p[-2] > 0 else 10000.0
    half = margin * c * (3.0 * a1 + gam)
    return float(f.min() - half), float(f.max() + half)


def balanced_bin_shard(x, world, rank, heavy_lo, heavy_hi, heavy_weight=17.0):
    """Contiguous [lo, hi) bin range of a rank with EQUAL ESTIMATED COST instead of equal length: bins between heavy_lo and
    heavy_hi (where truncation windows lie) count heavy_weight times a noise-only bin (about 44 Lorentzian components
    per bin at 6.25 FP64 instructions against 17 for background + likelihood on the S-global star)."""
    w = np.ones(x.size)
    w[(x >= heavy_lo) & (x <= heavy_hi)] = heavy_weight
    cw = np.concatenate([[0.0], np.cumsum(w)])
    cuts = [int(np.searchsorted(cw, cw[-1] * k / world)) for k in range(world + 1)]
    cuts[0], cuts[-1] = 0, x.size
    return cuts[rank], max(cuts[rank + 1], cuts[rank] + 2)


def balanced_shard_descriptor(prob, world, rank, heavy_weight=17.0):
    """shard_descriptor with cost-balanced cuts (see balanced_bin_shard)."""
    x = prob.x
    lo_f, hi_f = window_region(prob)
    lo, hi = balanced_bin_shard(x, world, rank, lo_f, hi_f, heavy_weight)
    return dict(Nx_global=int(x.size), shard_lo=int(lo), x_first=float(x[0]), x_second=float(x[1]), x_last=float(x[-1])), slice(lo, hi)


def logl_from_sums(s1, s2, p, T, like_case=0):
    """likelihoods.cpp:25 + model_def.cpp:302 applied to (all-reduced) partial sums."""
    s1, s2 = np.asarray(s1), np.asarray(s2)
    f = -float(int(p)) * (s1 + s2) if like_case == 0 else -s1
    return f / np.asarray(T)
