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


def logl_from_sums(s1, s2, p, T, like_case=0):
    """likelihoods.cpp:25 + model_def.cpp:302 applied to (all-reduced) partial sums."""
    s1, s2 = np.asarray(s1), np.asarray(s2)
    f = -float(int(p)) * (s1 + s2) if like_case == 0 else -s1
    return f / np.asarray(T)
