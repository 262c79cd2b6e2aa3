"""CPU test of the N>1 host logic with torch.distributed (gloo, world_size 2): ensemble partitioning
needs no data-path collective; the bin-sharded variant all-reduces 2*C partial sums per step."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import tamcmc_b200_loader
    tb = tamcmc_b200_loader.load()
    from oracle.oracle import Oracle
    W, S, O = tb.workloads, tb.sharding, Oracle()
    prob = W.small_global(3, Nx=3001, step=0.1)
    st, m, _ = O.model(prob.model_case, prob.params, prob.plength, prob.x)
    prob.y = m * (-np.log(W.philox_uniform(m.size, 4)))
    C = 5
    P, T = W.perturbed_params(prob, C), W.tcoefs(C)
    full = O.generate_models(prob, P, T, want_model=True)
    # (ii) bin-sharded: partial sums over this rank's bins, all-reduced
    desc, sl = S.shard_descriptor(prob.x, world, rank)
    part = np.zeros(2 * C)
    for c in range(C):
        mm = full["model"][c][sl]
        part[2 * c] = np.sum(prob.y[sl] * (1.0 / mm)); part[2 * c + 1] = np.sum(np.log(mm))
    t = torch.from_numpy(part)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    logL = S.logl_from_sums(t.numpy()[0::2], t.numpy()[1::2], prob.likelihood_params, T)
    ok_shard = bool(np.allclose(logL, full["logL"], rtol=1e-12))
    # (i) ensemble partition: units are disjoint and complete, only a final gather
    mine = S.partition_units(7, world, rank)
    got = [None] * world
    dist.all_gather_object(got, mine)
    ok_units = sorted(sum(got, [])) == list(range(7)) and all(len(set(a) & set(b)) == 0 for i, a in enumerate(got) for b in got[i + 1:])
    if rank == 0:
        q.put((ok_shard, ok_units, desc["shard_lo"]))
    dist.barrier()
    dist.destroy_process_group()


def test_gloo_world2_sharding():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = q.get(timeout=240)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert res[0] and res[1] and res[2] == 0
