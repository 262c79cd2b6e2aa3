#!/usr/bin/env python
"""bench.py -- tempered chain-steps/s of the TAMCMC hot path on B200 (BASELINE.json metric).

Workload (N=1): cfg3 of BASELINE.json -- model_MS_Global_a1etaa3_HarveyLike_Classic, 106 parameters,
synthetic 1 000 000-bin spectrum (SURVEY.md 8(d) "S-global"), one tempered ensemble of 10 chains.
A "step" is one MCMC iteration of every chain: proposal, fused model+likelihood over all bins,
prior, Metropolis accept, Robbins-Monro adaptation, tempering swap.  N>1: one independent
ensemble (a different synthetic star) per GPU, no data-path collective (weak scaling).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

Prints ONE JSON line (rank 0).  ``--impl reference`` times the reference's own CPU implementation
of the path (oracle/_ref: unmodified reference sources, OpenMP over chains) on the host cores.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

C_CHAINS = 10
NX = 1_000_000
METRIC = "tempered chain-steps/s (model_MS_Global 1M-bin spectrum, 10 chains per ensemble)"


def synth_problem(W, make_y, rank=0):
    """S-global; rank r gets another star (seed, numax/Dnu jitter) -- cfg5's recipe."""
    g = np.random.Generator(np.random.Philox(key=777 + rank))
    jit = (1.0 + 0.1 * (2 * g.random(2) - 1)) if rank > 0 else np.ones(2)
    return W.s_global(Nx=NX, model_case=3, seed=12345 + rank, make_y=make_y, numax=2700.0 * jit[0], dnu=120.186 * jit[1])


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (recipe of B200_PROFILING.md)."""

    Q = "index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown," \
        "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, gpu_index):
        self.rows = []
        self.proc = None
        self.idx = gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.idx), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
                for k, nm in enumerate(names):
                    if r[5 + k].lower().startswith("active"):
                        reasons.add(nm)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": float(max(mx)) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def load_traffic():
    """dram bytes per launch of the fused kernel from the committed ncu summary (profiles/), if any."""
    p = os.path.join(ROOT, "profiles", "eval_kernel_ncu_summary.json")
    if os.path.exists(p):
        try:
            return json.load(open(p)).get("dram_bytes_per_launch")
        except Exception:
            return None
    return None


def run_reference(args, rank, world):
    """CPU arm: Model_def::generate_model for the C chains of one step, OpenMP over chains
    (MALA.cpp:632), unmodified reference sources (oracle/_ref) -- or the oracle port if absent."""
    if rank != 0:
        return
    # all the host threads the box has (torchrun exports OMP_NUM_THREADS=1 to its workers)
    cores = os.cpu_count() or 1
    os.environ["OMP_NUM_THREADS"] = str(cores)
    try:
        import ctypes
        ctypes.CDLL("libgomp.so.1").omp_set_num_threads(cores)
    except OSError:
        pass
    import tamcmc_b200_loader
    tb = tamcmc_b200_loader.load()
    W = tb.workloads
    from oracle.oracle import Oracle, RefLib
    O, R = Oracle(), RefLib()
    prob = synth_problem(W, make_y=lambda p: O.model(p.model_case, p.params, p.plength, p.x)[1])
    P, T = W.perturbed_params(prob, C_CHAINS), W.tcoefs(C_CHAINS)
    kind = "reference" if R.available else "port"
    fn = (lambda: R.generate_models(prob, P, T)) if R.available else (lambda: O.generate_models(prob, P, T))
    cores = os.cpu_count() or 1
    for _ in range(args.warmup):
        fn()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        fn()
    dt = time.perf_counter() - t0
    val = C_CHAINS * args.steps / dt
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": "chain-steps/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "cfg3 S-global: model_MS_Global_a1etaa3_HarveyLike_Classic, 106 params, Nx=1000000, C=10 chains",
                       "step": "generate_model (model+logL+prior) for the 10 chains of one MCMC iteration"},
            "cpu_baseline": {"value": val, "unit": "chain-steps/s", "cores": min(cores, C_CHAINS), "kind": kind,
                             "sample": "%d full steps (10 chains x 1M bins each), OpenMP over chains, %d host cores" % (args.steps, cores)},
            "e2e": {"value": val, "unit": "chain-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "bin_evals_per_s": val * NX}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--warmup", type=int, default=200)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--cpu-steps", type=int, default=4)
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        if args.steps > 60:
            args.steps = 60          # bounded sample: each step is ~1 s of all-core CPU work
        args.warmup = min(args.warmup, 3)
        run_reference(args, rank, world)
        return

    import torch
    import tamcmc_b200_loader
    tb = tamcmc_b200_loader.load()
    W = tb.workloads
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback in the product path)")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        if os.environ.get("NCCL_DEBUG", "").upper() in ("", "VERSION"):
            os.environ["NCCL_DEBUG"] = "WARN"          # keep stdout to the one JSON line
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    lib = tb.load_library()
    peak_tflops, _ = lib.measure_fp64_peak(local_rank)

    # ---- synthetic star of this rank: y = m_true * Exp(1), m_true from the GPU model path itself ----
    def make_y(p):
        tmp = tb.GpuModelDef(p, tb.MalaConfig(Nchains=1), device=local_rank)
        m = tmp.call_model(0)
        tmp.close()
        return m
    prob = synth_problem(W, make_y, rank)
    cfg = tb.MalaConfig(Nchains=C_CHAINS, lambda_temp=1.7, c0=2.0, Nt_learn=[100, 10**9], periods_learn=[1],
                        dN_mixing=1, seed=4242 + rank)
    md = tb.GpuModelDef(prob, cfg, device=local_rank)
    mala = tb.GpuMALA(md)
    W_steps, K = max(args.warmup, 3), args.steps

    def barrier():
        md.sync()
        if dist is not None:
            dist.barrier()
        md.sync()

    # ---- warm-up (learning on from i=100, as in the Learning phase of the reference) ----
    mala.execute(0, W_steps, record=False)
    i0 = W_steps
    # ---- timed region 1: `value` -- device-resident loop, nothing crosses PCIe ----
    clocks = ClockSampler(local_rank)
    clocks.start()
    md.kernel_time(reset=2)
    barrier()
    md.timer_start()
    mala.execute(i0, K, record=False)
    ms = md.timer_stop_ms()
    barrier()
    n_eval, _, launches = md.kernel_time(reset=0)
    i0 += K
    clk = clocks.stop()
    if dist is not None:
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    value = world * C_CHAINS * K / (ms * 1e-3)

    # ---- timed region 2: same steps with CUDA events around every launch of the fused kernel ----
    md.kernel_time(reset=1)
    Kr = min(K, 500)
    mala.execute(i0, Kr, record=False)
    n_ev, ms_ev, _ = md.kernel_time(reset=2)
    i0 += Kr
    t_kernel = ms_ev / max(n_ev, 1) * 1e-3
    fin = md.evaluate(md.params, md.Tcoefs)
    Lw, Kh = md.last_work(C_CHAINS)
    F_chain = 8.0 * Lw + NX * (5.0 * Kh + 5.0)                 # F = 8L + Nx(5K+1+4), asym == 0
    F_launch = float(F_chain.sum())
    achieved = F_launch / t_kernel / 1e12
    # x and y once per launch (ln x is only read for non-integer Harvey exponents; cfg3 has p = 4, 2) + per-chain tables
    hbm_bytes = 16.0 * NX + C_CHAINS * (md.Nparams * 8 + 224 * 48 + 44 * 8)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = peaks.get("hbm_gbs", 6650.0)
    roofline = {"bound": "fp64", "achieved": achieved, "peak": peak_tflops, "unit": "TFLOP/s", "frac": achieved / peak_tflops,
                "traffic": load_traffic(), "kernel": "tg_k_eval", "avg_launch_us": t_kernel * 1e6, "launches_timed": n_ev,
                "algorithmic_flop_per_launch": F_launch, "L_component_bins_per_chain": float(Lw.mean()), "K_harvey": Kh,
                "kernel_share_of_step": t_kernel / (ms * 1e-3 / K),
                "peak_source": "DFMA microkernel measured in this run (MEASURED_PEAKS.json has no FP64 figure)",
                "hbm": {"achieved": hbm_bytes / t_kernel / 1e9, "peak": hbm_peak, "unit": "GB/s",
                        "frac": hbm_bytes / t_kernel / 1e9 / hbm_peak, "algorithmic_bytes_per_launch": hbm_bytes,
                        "peak_source": "MEASURED_PEAKS.json" if "hbm_gbs" in peaks else "fallback"}}

    # ---- e2e: the reference-facing call (batched Model_def::generate_model) with HOST buffers ----
    g = np.random.Generator(np.random.Philox(key=99 + rank))
    idx = md.index_to_relax
    err = prob.init_errors()
    Pcur = md.params.copy()
    Ke = min(K, 400)
    # the host sampler's proposals for every step, in pinned host memory (what an unmodified host-side
    # MALA::execute would hand to generate_model); every call copies its own C x Np block to the device
    pin_params = torch.empty((Ke + 3, C_CHAINS, md.Nparams), dtype=torch.float64).pin_memory()
    pin_T = torch.from_numpy(md.Tcoefs.copy()).pin_memory()
    Ph, Th = pin_params.numpy(), pin_T.numpy()
    Ph[:] = Pcur[None, :, :]
    Ph[:, :, idx] += 0.05 * err[None, None, :] * g.standard_normal((Ke + 3, C_CHAINS, idx.size))
    outs = [np.empty(C_CHAINS) for _ in range(3)]
    st = np.zeros(C_CHAINS, dtype=np.int32)
    step_bytes = C_CHAINS * md.Nparams * 8
    p0, pT = Ph.ctypes.data, Th.ctypes.data
    po = [o.ctypes.data for o in outs] + [st.ctypes.data]
    tg_eval = lib.lib.tg_eval

    def host_step(k):
        rc = tg_eval(md.h, p0 + k * step_bytes, C_CHAINS, pT, po[0], po[1], po[2], po[3], None)
        if rc != 0:
            lib.check(rc)
    for k in range(3):
        host_step(k)
    barrier()
    t0 = time.perf_counter()
    for k in range(Ke):
        host_step(3 + k)
    dt_e2e = time.perf_counter() - t0
    if dist is not None:
        t = torch.tensor([dt_e2e], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt_e2e = float(t.item())
    e2e = {"value": world * C_CHAINS * Ke / dt_e2e, "unit": "chain-steps/s",
           "h2d_bytes_per_step": C_CHAINS * md.Nparams * 8 + C_CHAINS * 8, "d2h_bytes_per_step": C_CHAINS * (3 * 8 + 4),
           "call": "tg_eval == batched Model_def::generate_model, pinned host params in, logL/logPrior/logPost/status out",
           "steps": Ke}

    # ---- the other single-GPU configurations of BASELINE.json, same device-resident sampler, short runs (N=1 only;
    #      reported beside the headline, not part of it) ----
    other = None
    if rank == 0 and world == 1:
        other = {}

        def short_run(p, c, warm, steps):
            m2 = tb.GpuModelDef(p, c, device=local_rank)
            ml = tb.GpuMALA(m2)
            ml.execute(0, warm, record=False)
            m2.sync(); m2.timer_start()
            ml.execute(warm, steps, record=False)
            t_ms = m2.timer_stop_ms()
            m2.kernel_time(reset=1)
            ml.execute(warm + steps, min(steps, 64), record=False)
            n2, ms2, _ = m2.kernel_time(reset=2)
            m2.close()
            return t_ms / steps, ms2 / max(n2, 1)
        p1 = W.s_local(make_y=make_y)
        ms1, k1 = short_run(p1, tb.MalaConfig(Nchains=10, lambda_temp=1.7, c0=2.0, Nt_learn=[100, 10**9], periods_learn=[1], dN_mixing=1, seed=7), 200, 2000)
        other["configs[1] S-local: model_MS_local_basic, 10k bins, 10 chains"] = {"chain_steps_per_s": 10 / (ms1 * 1e-3), "ms_per_step": ms1, "fused_kernel_us": k1 * 1e3}
        C4 = 1024
        ms4, k4 = short_run(prob, tb.MalaConfig(Nchains=C4, chains_per_ensemble=8, lambda_temp=1.7, c0=2.0, Nt_learn=[20, 10**9], periods_learn=[1], dN_mixing=1, seed=11), 32, 64)
        other["configs[3] S-global with 1024 batched chains (128 tempered ensembles of 8)"] = {
            "chain_steps_per_s": C4 / (ms4 * 1e-3), "ms_per_step": ms4, "fused_kernel_us": k4 * 1e3,
            "roofline_frac": F_launch / C_CHAINS * C4 / (k4 * 1e-3) / 1e12 / peak_tflops}

    # ---- CPU baseline on this box's host cores (rank 0, N=1 only; bounded sample) ----
    cpu = None
    if rank == 0 and world == 1:
        from oracle.oracle import Oracle, RefLib
        O, R = Oracle(), RefLib()
        Pc, Tc = W.perturbed_params(prob, C_CHAINS), W.tcoefs(C_CHAINS)
        fn = (lambda: R.generate_models(prob, Pc, Tc)) if R.available else (lambda: O.generate_models(prob, Pc, Tc))
        fn()
        t0 = time.perf_counter()
        for _ in range(args.cpu_steps):
            fn()
        dtc = time.perf_counter() - t0
        t0 = time.perf_counter()
        for _ in range(args.cpu_steps):
            O.generate_models(prob, Pc, Tc)
        dtp = time.perf_counter() - t0
        ncpu = os.cpu_count() or 1
        cpu = {"value": C_CHAINS * args.cpu_steps / dtc, "unit": "chain-steps/s", "cores": min(ncpu, C_CHAINS),
               "kind": "reference" if R.available else "port",
               "sample": "%d full steps (10 chains x 1M bins), generate_model under OpenMP over chains, box has %d cores" % (args.cpu_steps, ncpu),
               "lean_port_value": C_CHAINS * args.cpu_steps / dtp}

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": "chain-steps/s", "n_gpus": world, "steps": K, "warmup": W_steps,
                "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
                "data": "synthetic",
                "config": {"workload": "cfg3 S-global: model_MS_Global_a1etaa3_HarveyLike_Classic, 106 params (%d free), Nx=1000000, C=10 chains per GPU, lambda=1.7, learning on" % md.Nvars,
                           "l2": "x, y = 16 MB per launch (< L2) but the timed kernel is FP64-bound; no L2 flush (steady-state MCMC reuses the spectrum every step by design)",
                           "per_gpu": "one independent tempered ensemble (star) per GPU, no collective"},
                "clocks": clk, "e2e": e2e, "gpu_launches": int(launches), "roofline": roofline, "cpu_baseline": cpu,
                "bin_evals_per_s": value * NX, "other_configs": other}
        print(json.dumps(line), flush=True)
    md.close()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
