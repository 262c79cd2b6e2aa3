#!/usr/bin/env python
"""Bin-sharded variant of the hot path over NCCL (BASELINE.json configs[4] "bin-sharded NCCL variant
for a 10M-bin spectrum"): every rank holds a contiguous range of bins of ONE long spectrum and the same chain state and
Philox streams; per step the 2*C partial sums (sum y/m, sum ln m) are all-reduced (ncclAllReduce, fp64 sum) between
tg_step_begin and tg_step_end, after which every rank takes the same Metropolis / adaptation / tempering decisions.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 \
        tools/bin_sharded_nccl.py [--nx 10000000] [--steps 200] [--check-steps 40]

Checks (a run fails loudly otherwise): (1) after every checked step all ranks hold bit-identical logL / vars (the
replicas never diverge); (2) the sharded trajectory equals the one of the unsharded single-GPU run of the same seed on
rank 0 (every accept / swap decision identical, logL to 1e-8: only the summation order over bins differs, and the
step-size adaptation feeds those last-bit differences back into the positions).  Prints one JSON line with
the device-timed throughput (max over ranks)."""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import tamcmc_b200_loader  # noqa: E402


class _DevBuf:
    """A device buffer owned by libtamcmc_gpu.so seen through the CUDA array interface (no copy)."""

    def __init__(self, ptr, n):
        self.__cuda_array_interface__ = dict(shape=(n,), typestr="<f8", data=(ptr, False), version=3, strides=None)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--nx", type=int, default=10_000_000)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--check-steps", type=int, default=40)
    ap.add_argument("--chains", type=int, default=10)
    ap.add_argument("--equal-length", action="store_true", help="equal-length shards instead of cost-balanced ones")
    ap.add_argument("--host-driven", action="store_true", help="the round-1 form: tg_step_begin / all_reduce by torch / tg_step_end with two host synchronisations per step "
                    "(default: the all-reduce inside the library's captured step graph, tg_shard_attach)")
    ap.add_argument("--nccl", action="store_true", help="ncclAllReduce inside the library's step graph (tg_shard_attach) instead of the peer-memory exchange kernel (tg_shard_ipc_*)")
    ap.add_argument("--heavy-weight", type=float, default=17.0, help="cost of a bin under truncation windows in units of a background-only bin (measured: 17 .. 2000 give the same step time on 2 GPUs)")
    a = ap.parse_args()
    import torch
    import torch.distributed as dist
    tb = tamcmc_b200_loader.load()
    W, S = tb.workloads, tb.sharding
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    # S-10M of SURVEY.md 8(d): the S-global star on a ten times finer grid (step 0.00085 microHz for 10M bins)
    prob = W.s_global(Nx=a.nx, step=8500.0 / a.nx)
    full_x, full_y = prob.x, prob.y
    # contiguous bin ranges of equal estimated cost (the truncation windows of the modes make some bins ~17x dearer)
    desc, sl = (S.balanced_shard_descriptor(prob, world, rank, a.heavy_weight) if not a.equal_length else S.shard_descriptor(full_x, world, rank))
    cfg = tb.MalaConfig(Nchains=a.chains, Nt_learn=[10, 1000000], periods_learn=[1], c0=1.7, seed=2024)
    import copy
    sub = copy.copy(prob)
    sub.x, sub.y = np.ascontiguousarray(full_x[sl]), np.ascontiguousarray(full_y[sl])
    md = tb.GpuModelDef(sub, cfg, device=local, shard=desc if world > 1 else None)
    mala = tb.GpuMALA(md)
    ptr, n = mala.partials_ptr()
    part = torch.as_tensor(_DevBuf(ptr, n), device="cuda:%d" % local)

    in_library = world > 1 and not a.host_driven
    if in_library and not a.nccl:
        # peer memory: every rank exports its exchange buffer (CUDA IPC handle, 64 bytes), torch.distributed gathers the handles
        mine = torch.frombuffer(bytearray(md.shard_ipc_export(world)), dtype=torch.uint8).cuda()
        allh = [torch.zeros(64, dtype=torch.uint8, device="cuda") for _ in range(world)]
        dist.all_gather(allh, mine)
        md.shard_ipc_attach(b"".join(bytes(t.cpu().numpy().tobytes()) for t in allh), world, rank)
    elif in_library:
        # the library's own communicator: rank 0 draws the id, torch.distributed carries the 128 bytes, every rank attaches
        uid = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            uid.copy_(torch.frombuffer(bytearray(tb.gpu.shard_unique_id()), dtype=torch.uint8))
        dist.broadcast(uid, src=0)
        md.shard_attach(bytes(uid.cpu().numpy().tobytes()), world, rank)
    elif world > 1:
        # the Model_def constructor evaluated every chain on this rank's bins only: install the all-reduced start
        md.sync()
        dist.all_reduce(part, op=dist.ReduceOp.SUM)
        torch.cuda.synchronize()
        mala.install_current()

    def step(i):
        if in_library or world == 1:
            mala.execute(i, 1, record=False)                        # one step of tg_run: the all-reduce is inside
            return
        mala.step_begin(i)
        md.sync()
        dist.all_reduce(part, op=dist.ReduceOp.SUM)
        torch.cuda.synchronize()
        mala.step_end(i)

    traj = []
    for i in range(a.check_steps):
        step(i)
        md.sync(); md.refresh_from_device()
        traj.append((md.logLikelihood.copy(), md.vars.copy()))
        if world > 1:                                              # (1) replicas identical, bit for bit
            mine = torch.from_numpy(np.concatenate([md.logLikelihood, md.vars.ravel()])).cuda()
            ref = mine.clone()
            dist.broadcast(ref, src=0)
            assert torch.equal(mine, ref), "rank %d diverged from rank 0 at step %d" % (rank, i)
    # timed steps
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    if in_library or world == 1:
        mala.execute(a.check_steps, 32, record=False)                # (the step graphs are replayed from here on)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        md.timer_start()
        mala.execute(a.check_steps + 32, a.steps, record=False)      # ONE call: no host synchronisation per step
        ms_lib = md.timer_stop_ms()
    e0.record()
    if not (in_library or world == 1):
        for i in range(a.check_steps, a.check_steps + a.steps):
            step(i)
    md.sync()
    e1.record(); torch.cuda.synchronize()
    ms = torch.tensor([ms_lib if (in_library or world == 1) else e0.elapsed_time(e1)], device="cuda")
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ok = True
    if rank == 0:                                                  # (2) against the unsharded run of the same seed
        if world > 1 and a.nx <= 20_000_000:
            ref_md = tb.GpuModelDef(prob, cfg, device=local)
            rm = tb.GpuMALA(ref_md)
            out = rm.execute(0, a.check_steps)
            for i, (l, v) in enumerate(traj):                    # equal positions after every step <=> identical accept / swap decisions
                # (the step-size adaptation feeds the last-bit differences of logL back into the positions: drift ~1e-10 per step)
                same = np.allclose(l, out["logL"][i], rtol=1e-8, atol=0) and np.allclose(v, out["vars"][i], rtol=1e-7, atol=1e-12)
                if not same:
                    ok = False
                    print("mismatch with the unsharded run at step %d: logL %s vs %s" % (i, l[:3], out["logL"][i][:3]), file=sys.stderr)
                    break
            ref_md.close()
        print(json.dumps({"metric": "tempered chain-steps/s, bin-sharded variant (one spectrum over all GPUs, all-reduce of 2*C doubles per step)",
                          "value": a.chains * a.steps / (float(ms.item()) * 1e-3), "unit": "chain-steps/s", "n_gpus": world, "steps": a.steps,
                          "ms_per_step": float(ms.item()) / a.steps, "Nx": a.nx, "chains": a.chains, "bins_per_gpu": int(sub.x.size),
                          "replicas_identical_steps": a.check_steps, "equals_unsharded_run": bool(ok), "collective": (("ncclAllReduce fp64 sum, %d doubles, inside the library's step graph" if a.nccl else "peer-memory exchange kernel (tg_k_xchg over CUDA IPC / NVLink), %d doubles per rank, inside the library's step graph") if in_library else "nccl all_reduce by torch, fp64 sum, %d doubles, host-driven") % n}))
    if in_library and not a.nccl:
        md.shard_status()
    if world > 1:
        dist.barrier()                                             # nobody unmaps a buffer a peer may still write
    md.close()
    if world > 1:
        dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
