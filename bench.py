#!/usr/bin/env python
"""bench.py -- tempered chain-steps/s of the TAMCMC hot path on B200 (BASELINE.json metric).

Workload (N=1): cfg3 of BASELINE.json -- model_MS_Global_a1etaa3_HarveyLike_Classic, 106 parameters,
synthetic 1 000 000-bin spectrum (SURVEY.md 8(d) "S-global"), one tempered ensemble of 10 chains.
A "step" is one MCMC iteration of every chain: proposal, fused model+likelihood over all bins,
prior, Metropolis accept, Robbins-Monro adaptation, tempering swap.  N>1: one independent
ensemble (a different synthetic star) per GPU, no data-path collective (weak scaling).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
                    [--config cfg3|cfg2|cfg4|cfg5] [--stars-per-gpu S] [--no-cfg5] [--extra-out PATH]

Prints ONE JSON line (rank 0).  ``--impl reference`` times the reference's own CPU implementation
of the path (oracle/_ref: unmodified reference sources, OpenMP over chains) on the host cores.

The default line is cfg3 (the headline) and carries, under "cfg5_8_stars_per_gpu", the same metric with EIGHT stars of
10 chains per GPU advanced as one launch group (configs[4] of BASELINE.json: 64 stars over 8 GPUs).  ``--config`` selects
another configuration of BASELINE.json as the line's workload (builder runs, committed under profiles/).
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
STEPS_PER_GRAPH = 16
I0 = 10000          # first iteration of the bench: a run restored into its Learning phase (gamma = c0/(1+i) = 2e-4, MALA.cpp:641-652); at i = 0
                    # gamma would be 2 and the first Robbins-Monro updates would leave an indefinite covariance
METRIC = "tempered chain-steps/s (model_MS_Global 1M-bin spectrum, 10 chains per ensemble)"


def synth_problem(W, make_y, star=0, jitter=True):
    """S-global; star s > 0 is another star (noise seed, and numax/Dnu jitter -- cfg5's recipe -- unless jitter is off: the headline's
    one star per GPU is the SAME star observed with independent noise on every GPU, i.e. equal work per GPU, so that the N-GPU value
    measures the machine and not the spread of window widths between stars)."""
    g = np.random.Generator(np.random.Philox(key=777 + star))
    jit = (1.0 + 0.1 * (2 * g.random(2) - 1)) if (star > 0 and jitter) else np.ones(2)
    return W.s_global(Nx=NX, model_case=3, seed=12345 + star, make_y=make_y, numax=2700.0 * jit[0], dnu=120.186 * jit[1])


def workload_config(nv, stars_per_gpu=1):
    """config of the JSON line -- the same dict in both arms."""
    return {"workload": "cfg3 S-global: model_MS_Global_a1etaa3_HarveyLike_Classic, 106 params (%d free), Nx=1000000, C=10 chains per star, "
                        "%d star(s) per GPU, lambda=1.7, learning on (iterations from %d)" % (nv, stars_per_gpu, I0),
            "step": "one MCMC iteration of every chain; the reference arm and e2e time its dominant part, generate_model (model+logL+prior) for the chains of one iteration",
            "l2": "x, y = 16 MB per launch and star (< L2) but the timed kernel is FP64-bound; no L2 flush (steady-state MCMC reuses the spectrum every step by design)",
            "per_gpu": "independent tempered ensembles (stars) per GPU, no collective; with one star per GPU every GPU holds the same star with its own noise realisation "
                       "and Philox key (equal work per GPU), with several stars per GPU the stars differ (numax, Dnu within 10 %)"}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (recipe of B200_PROFILING.md); one sampler per node."""

    Q = "index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown," \
        "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, gpu_index, enabled=True):
        self.rows = []
        self.proc = None
        self.idx = gpu_index
        self.enabled = enabled

    def start(self):
        if not self.enabled:
            return
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
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["not sampled on this rank" if not self.enabled else "nvidia-smi unavailable"]}
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
    # bounded sample: a step is ~0.5 s of all-core CPU work; the driver's --steps 20 --warmup 5 run as given
    steps, warm = min(args.steps, 60), min(args.warmup, 10)
    for _ in range(warm):
        fn()
    t0 = time.perf_counter()
    for _ in range(steps):
        fn()
    dt = time.perf_counter() - t0
    val = C_CHAINS * steps / dt
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": "chain-steps/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / steps * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(prob.Nv),
            "cpu_baseline": {"value": val, "unit": "chain-steps/s", "cores": min(cores, C_CHAINS), "kind": kind,
                             "sample": "%d full steps after %d warm-up steps (10 chains x 1M bins each: generate_model of one star), OpenMP over chains as in MALA.cpp:632, %d host cores; "
                                       "unmodified reference sources compiled against the repo's Eigen-API shim" % (steps, warm, cores)},
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
    ap.add_argument("--config", default="cfg3", choices=["cfg2", "cfg3", "cfg4", "cfg5"])
    ap.add_argument("--stars-per-gpu", type=int, default=0, help="cfg5: stars (tempered ensembles of 10 chains, each its own spectrum) per GPU as one launch group")
    ap.add_argument("--no-cfg5", action="store_true", help="skip the 8-stars-per-GPU measurement carried by the default line")
    ap.add_argument("--no-bin-sharded", action="store_true", help="N > 1: skip the bin-sharded 10 M-bin measurement carried by the default line")
    ap.add_argument("--extra-out", default=None, help="also write the JSON line to this file")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
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
            os.environ["NCCL_DEBUG"] = "WARN"
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")     # stdout carries the one JSON line and nothing else
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    lib = tb.load_library()
    peak_tflops, _ = lib.measure_fp64_peak(local_rank)
    props = torch.cuda.get_device_properties(local_rank)

    # ---- synthetic stars of this rank: y = m_true * Exp(1), m_true from the GPU model path itself ----
    def make_y(p):
        tmp = tb.GpuModelDef(p, tb.MalaConfig(Nchains=1), device=local_rank)
        m = tmp.call_model(0)
        tmp.close()
        return m

    def mala_cfg(C, seed, per_ensemble=0):
        # learning on from the first step (Learning phase of the reference, MALA.cpp:641-652): 4 launches per step,
        # the adaptation CTAs run beside the proposal and the fused kernel
        return tb.MalaConfig(Nchains=C, chains_per_ensemble=per_ensemble, lambda_temp=1.7, c0=2.0, Nt_learn=[0, 10**9], periods_learn=[1],
                             dN_mixing=1, seed=seed)

    cfg_name = args.config
    S = args.stars_per_gpu if args.stars_per_gpu > 0 else (8 if cfg_name == "cfg5" else 1)
    if S > 1:
        cfg_name = "cfg5"
    C_star = {"cfg2": 10, "cfg3": C_CHAINS, "cfg4": 1024, "cfg5": C_CHAINS}[cfg_name]
    nx_star = 10000 if cfg_name == "cfg2" else NX
    W_run, K = max(args.warmup, 2 * STEPS_PER_GRAPH), args.steps       # warm-up covers two whole step graphs

    def build_group(n_stars, first_star):
        mds = []
        for s in range(n_stars):
            star = first_star + s
            if cfg_name == "cfg2":
                p = W.s_local(make_y=make_y)
            else:
                p = synth_problem(W, make_y, star, jitter=(n_stars > 1))
            mds.append(tb.GpuModelDef(p, mala_cfg(C_star, 4242 + star, 8 if cfg_name == "cfg4" else 0), device=local_rank))
        return mds, (tb.GpuBatch(mds) if n_stars > 1 else None)

    def barrier(md):
        md.sync()
        if dist is not None:
            dist.barrier()
        md.sync()

    def timed_run(mds, grp, warm, steps, clock_rank=False):
        """warm-up, then `steps` steps timed on the device (CUDA events on the group's stream), max over ranks."""
        md0 = mds[0]
        runner = grp if grp is not None else tb.GpuMALA(md0)
        runner.execute(I0, warm, record=False)
        clocks = ClockSampler(local_rank, enabled=clock_rank)
        clocks.start()
        (grp or md0).kernel_time(reset=2)
        barrier(md0)
        runner.execute(I0 + warm, steps, record=False)
        # CUDA events recorded by the call itself on its stream, right before the first launch and right after the last step: exactly
        # the K steps on the device's clock, none of this script's host code in between
        ms = (grp or md0).last_run_ms()
        barrier(md0)
        _, _, launches = (grp or md0).kernel_time(reset=0)
        clk = clocks.stop()
        if dist is not None:
            t = torch.tensor([ms], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        # the same steps with CUDA events around every launch of the fused kernel
        (grp or md0).kernel_time(reset=1)
        Kr = min(steps, 200)
        runner.execute(I0 + warm + steps, Kr, record=False)
        n_ev, ms_ev, _ = (grp or md0).kernel_time(reset=2)
        ms_ff = (grp or md0).prepass_time()        # the far-field pre-pass launches of the same steps
        return ms, launches, clk, ms_ev / max(n_ev, 1) * 1e-3, n_ev, ms_ff / max(n_ev, 1) * 1e-3

    def algorithmic_work(mds):
        """F = 8 L + Nx (5 K + 5) per chain (asym == 0).  Executed by the fused kernel: under the truncation windows
        (micro-tiles the far-field pre-pass prepared) a Horner recurrence of 15 FMA and 8 flop per exact record per bin (+ the
        likelihood's 5, + the 5 K of the background where the interpolant does not hold it), the component-by-component 8 L only
        where the path is off; the background term bin by bin only where bins are taken one by one, 14 exact evaluations per
        (moment tile, chain) elsewhere.  Executed by the pre-pass: 8 flop per far component and node (16 nodes), and per
        micro-tile the node-to-coefficient products (2 x 16 x 16), the change of basis (2 x 72) and the pair's Horner (2 x 15 x 8)."""
        F = Fx = Fpre = 0.0
        Lm, Kh = [], 0
        ff_stats = [0, 0, 0, 0, 0]
        ff_bins_per_mt = 0
        for md in mds:
            md.evaluate(md.params, md.Tcoefs)
            Lw, Kh = md.last_work(md.Nmodels)
            bins_direct, mom_tiles, _ = md.work_profile()
            mt, marked, recs, nfar, folded, mtb = md.farfield_profile(md.Nmodels)
            ff_stats = [a + b for a, b in zip(ff_stats, (mt, marked, recs, nfar, folded))]
            ff_bins_per_mt = mtb
            F += float((8.0 * Lw + md.Nx * (5.0 * Kh + 5.0)).sum())
            mom = mom_tiles * (14 * (5.0 * Kh + 5.0) + 2 * 2 * 14 * 9 + 4 * 9)
            if mt > 0 and marked == 0:
                ff_bins = float(mtb) * mt / md.Nmodels                    # per chain
                per_chain = (bins_direct - ff_bins) * (5.0 * Kh + 5.0) + mom
                Fx += md.Nmodels * per_chain + float(mtb) * (mt * (30.0 + 5.0) + 8.0 * recs + (mt - folded) * 5.0 * Kh)
                Fpre += 16 * 8.0 * nfar + mt * (2 * 16 * 16 + 2 * 72 + 2 * 15 * 8)
            else:
                Fx += float((8.0 * Lw + bins_direct * (5.0 * Kh + 5.0) + mom).sum())
            Lm.append(float(Lw.mean()))
        return F, Fx, float(np.mean(Lm)), Kh, Fpre, ff_stats + [ff_bins_per_mt]

    mds, grp = build_group(S, rank * S)
    md = mds[0]
    ms, launches, clk, t_kernel, n_ev, t_pre = timed_run(mds, grp, W_run, K, clock_rank=(local_rank == 0))
    chains_gpu = S * C_star
    value = world * chains_gpu * K / (ms * 1e-3)
    F_launch, F_exec, L_mean, Kh, F_pre, ff_stats = algorithmic_work(mds)
    achieved = F_launch / t_kernel / 1e12
    peak_nominal = props.multi_processor_count * 64 * 2 * (clk.get("sm_max_mhz") or 1965.0) * 1e6 / 1e12     # SMs x 64 DFMA lanes x 2 flop x clock
    # x and y once per launch and star (ln x is only read for non-integer Harvey exponents; cfg3 has p = 4, 2) + per-chain tables
    hbm_bytes = S * 16.0 * nx_star + chains_gpu * (md.Nparams * 8 + 224 * 48 + 44 * 8)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = peaks.get("hbm_gbs", 6650.0)
    roofline = {"bound": "fp64", "achieved": achieved, "peak": peak_tflops, "unit": "TFLOP/s", "frac": achieved / peak_tflops,
                "frac_executed": F_exec / t_kernel / 1e12 / peak_tflops, "peak_nominal": peak_nominal, "frac_of_nominal": achieved / peak_nominal,
                "traffic": load_traffic(), "kernel": "tg_k_eval", "avg_launch_us": t_kernel * 1e6, "launches_timed": n_ev,
                "algorithmic_flop_per_launch": F_launch, "executed_flop_per_launch": F_exec,
                "L_component_bins_per_chain": L_mean, "K_harvey": Kh,
                "kernel_share_of_step": t_kernel / (ms * 1e-3 / K),
                "prepass_kernel": "tg_k_ff", "prepass_avg_launch_us": t_pre * 1e6,
                "frac_with_prepass": F_launch / (t_kernel + t_pre) / 1e12 / peak_tflops,
                "frac_executed_with_prepass": (F_exec + F_pre) / (t_kernel + t_pre) / 1e12 / peak_tflops,
                "prepass_executed_flop_per_launch": F_pre,
                "far_field": {"micro_tiles": ff_stats[0], "marked": ff_stats[1], "exact_records": ff_stats[2], "far_components": ff_stats[3], "background_folded": ff_stats[4], "bins_per_micro_tile": ff_stats[5],
                              "note": "per launch, all chains: micro-tiles under the truncation windows prepared by tg_k_ff; exact records = components still evaluated "
                                      "exactly per bin there; far components = (component, micro-tile) pairs replaced by the interpolant"},
                "model_kernels_share_of_step": (t_kernel + t_pre) / (ms * 1e-3 / K),
                "prepass_note": "tg_k_ff prepares, per chain and 512-bin micro-tile under the truncation windows, the Chebyshev interpolant of the far components and the "
                                "records of the near ones; frac counts the fused kernel alone, frac_with_prepass the two launches that evaluate the model",
                "peak_source": "DFMA microkernel measured in this run (MEASURED_PEAKS.json has no FP64 figure); peak_nominal = SMs x 64 lanes x 2 x max SM clock",
                "frac_executed_note": "flops the fused kernel actually runs / its time / peak: under the truncation windows a 15-FMA Horner recurrence + 8 flop per exact record "
                                      "per bin instead of 8 flop per overlapping component (frac counts those 8 L as if they were run: it measures the algorithmic "
                                      "gain, not the pipe), the Nx(5K+5) background term only over bins taken one by one, 14 exact background evaluations per "
                                      "(moment tile, chain) elsewhere",
                "hbm": {"achieved": hbm_bytes / t_kernel / 1e9, "peak": hbm_peak, "unit": "GB/s",
                        "frac": hbm_bytes / t_kernel / 1e9 / hbm_peak, "algorithmic_bytes_per_launch": hbm_bytes,
                        "peak_source": "MEASURED_PEAKS.json" if "hbm_gbs" in peaks else "fallback"}}

    # ---- e2e: the reference-facing call (batched Model_def::generate_model) with HOST buffers, star after star ----
    g = np.random.Generator(np.random.Philox(key=99 + rank))
    Ke = min(max(K, 100), 400) if S == 1 else min(K, 100)      # at least 100 calls: a pipeline of 20 would mostly measure its own filling
    tg_eval = lib.lib.tg_eval
    calls = []
    for m_ in mds:
        idx = m_.index_to_relax
        err = m_.prob.init_errors()
        # the host sampler's proposals for every step, in pinned host memory (what an unmodified host-side
        # MALA::execute would hand to generate_model); every call copies its own C x Np block to the device
        pin_params = torch.empty((Ke + 3, m_.Nmodels, m_.Nparams), dtype=torch.float64).pin_memory()
        pin_T = torch.from_numpy(m_.Tcoefs.copy()).pin_memory()
        Ph, Th = pin_params.numpy(), pin_T.numpy()
        Ph[:] = m_.params[None, :, :]
        Ph[:, :, idx] += 0.05 * err[None, None, :] * g.standard_normal((Ke + 3, m_.Nmodels, idx.size))
        outs = [np.empty(m_.Nmodels) for _ in range(3)]
        st = np.zeros(m_.Nmodels, dtype=np.int32)
        calls.append((m_.h, Ph.ctypes.data, m_.Nmodels * m_.Nparams * 8, m_.Nmodels, Th.ctypes.data, [o.ctypes.data for o in outs] + [st.ctypes.data], (pin_params, pin_T, outs, st)))

    def host_step(k):
        for (h, p0, nb, cn, pT, po, _) in calls:
            rc = tg_eval(h, p0 + k * nb, cn, pT, po[0], po[1], po[2], po[3], None)
            if rc != 0:
                lib.check(rc)
    for k in range(3):
        host_step(k)
    barrier(md)
    t0 = time.perf_counter()
    for k in range(Ke):
        host_step(3 + k)
    dt_sync = time.perf_counter() - t0
    # the same calls in two halves (tg_eval_async / tg_eval_wait, two staging slots per star): the host hands over the vectors of
    # step k+1 while the GPU evaluates step k -- what a host-side sampler that prepares its next proposal during the evaluation sees
    ev_async, ev_wait = lib.lib.tg_eval_async, lib.lib.tg_eval_wait

    def submit(k):
        for (h, p0, nb, cn, pT, po, _) in calls:
            rc = ev_async(h, p0 + k * nb, cn, pT, k & 1)
            if rc != 0:
                lib.check(rc)

    def collect(k):
        for (h, p0, nb, cn, pT, po, _) in calls:
            rc = ev_wait(h, k & 1, po[0], po[1], po[2], po[3])
            if rc != 0:
                lib.check(rc)
    # both slots once before the clock starts: the second slot's tables, stream and graph are made at its first use
    submit(0); submit(1); collect(0); collect(1)
    barrier(md)
    t0 = time.perf_counter()
    submit(3)
    for k in range(1, Ke):
        submit(3 + k)
        collect(3 + k - 1)
    collect(3 + Ke - 1)
    dt_e2e = time.perf_counter() - t0
    if dist is not None:
        t = torch.tensor([dt_e2e], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt_e2e = float(t.item())
    e2e = {"value": world * chains_gpu * Ke / dt_e2e, "unit": "chain-steps/s",
           "h2d_bytes_per_step": chains_gpu * md.Nparams * 8 + chains_gpu * 8, "d2h_bytes_per_step": chains_gpu * (3 * 8 + 4),
           "call": "tg_eval_async + tg_eval_wait == batched Model_def::generate_model in two halves through two slots (own tables, staging buffers and stream each): "
                   "pinned host params in, logL/logPrior/logPost/status out, one call per star and step; the host submits step k+1 while step k is evaluated, "
                   "so the copy-in, prologue and pre-pass of one call run beside the fused kernel of the other",
           "blocking_call_value": world * chains_gpu * Ke / dt_sync, "blocking_call": "tg_eval (one blocking call per star and step)",
           "steps": Ke}
    nv = md.Nvars
    for m_ in mds:
        m_.sync()

    # ---- cfg5 beside the headline: EIGHT stars of 10 chains per GPU as one launch group (same run, same GPU) ----
    cfg5 = None
    if cfg_name == "cfg3" and not args.no_cfg5:
        if grp is not None:
            grp.close()
        for m_ in mds:
            m_.close()
        S5 = 8
        mds5, grp5 = build_group(S5, 1000 + rank * S5)
        ms5, launches5, _, tk5, n5, tp5 = timed_run(mds5, grp5, 2 * STEPS_PER_GRAPH, max(K, 4 * STEPS_PER_GRAPH))
        K5 = max(K, 4 * STEPS_PER_GRAPH)
        F5, Fx5, _, _, Fp5, _ = algorithmic_work(mds5)
        cfg5 = {"value": world * S5 * C_CHAINS * K5 / (ms5 * 1e-3), "unit": "chain-steps/s", "stars_per_gpu": S5, "chains_per_gpu": S5 * C_CHAINS,
                "steps": K5, "ms_per_step": ms5 / K5, "fused_kernel_us": tk5 * 1e6, "units_per_launch": grp5.sizes()[2],
                "roofline_frac": F5 / tk5 / 1e12 / peak_tflops, "roofline_frac_executed": Fx5 / tk5 / 1e12 / peak_tflops,
                "prepass_us": tp5 * 1e6, "roofline_frac_with_prepass": F5 / (tk5 + tp5) / 1e12 / peak_tflops,
                "roofline_frac_executed_with_prepass": (Fx5 + Fp5) / (tk5 + tp5) / 1e12 / peak_tflops,
                "kernel_share_of_step": tk5 / (ms5 * 1e-3 / K5), "per_gpu_vs_one_star": (S5 * C_CHAINS * K5 / (ms5 * 1e-3)) / (chains_gpu * K / (ms * 1e-3)),
                "gpu_launches": int(launches5),
                "what": "configs[4] of BASELINE.json: 8 different synthetic stars (own spectrum, parameters, Philox key) x 10 tempered chains per GPU, "
                        "ONE proposal / fused / decision launch per step for all of them (tg_batch_run)"}
        grp5.close()
        for m_ in mds5:
            m_.close()
        mds, grp = [], None

    # ---- N > 1: the bin-sharded variant beside the headline -- ONE spectrum of 10 M bins over all ranks, 10 chains, the 2*C partial
    #      sums exchanged every step by the library's peer-memory kernel inside its step graph (tools/bin_sharded_nccl.py checks that the
    #      replicas stay bit-identical and equal the unsharded run); rank 0 also times the unsharded spectrum on its own GPU ----
    bin_sharded = None
    if dist is not None and cfg_name == "cfg3" and not args.no_bin_sharded and world <= 8:
        import copy
        Sh = tb.sharding
        nx10 = 10_000_000
        p10 = W.s_global(Nx=nx10, step=8500.0 / nx10)
        desc, sl = Sh.balanced_shard_descriptor(p10, world, rank, 17.0)
        sub = copy.copy(p10)
        sub.x, sub.y = np.ascontiguousarray(p10.x[sl]), np.ascontiguousarray(p10.y[sl])
        cfg10 = tb.MalaConfig(Nchains=C_CHAINS, Nt_learn=[0, 10**9], periods_learn=[1], c0=2.0, lambda_temp=1.7, seed=2024)
        ms10 = tb.GpuModelDef(sub, cfg10, device=local_rank, shard=desc)
        mine = torch.frombuffer(bytearray(ms10.shard_ipc_export(world)), dtype=torch.uint8).cuda()
        allh = [torch.zeros(64, dtype=torch.uint8, device="cuda") for _ in range(world)]
        dist.all_gather(allh, mine)
        ms10.shard_ipc_attach(b"".join(bytes(t.cpu().numpy().tobytes()) for t in allh), world, rank)
        r10 = tb.GpuMALA(ms10)
        r10.execute(I0, 2 * STEPS_PER_GRAPH, record=False)
        barrier(ms10)
        r10.execute(I0 + 2 * STEPS_PER_GRAPH, 96, record=False)
        t10 = torch.tensor([ms10.last_run_ms()], dtype=torch.float64, device="cuda")
        dist.all_reduce(t10, op=dist.ReduceOp.MAX)
        ms10.shard_status()
        dist.barrier()
        ms10.close()
        one = None
        if rank == 0:
            m1 = tb.GpuModelDef(p10, cfg10, device=local_rank)
            r1 = tb.GpuMALA(m1)
            r1.execute(I0, 2 * STEPS_PER_GRAPH, record=False)
            r1.execute(I0 + 2 * STEPS_PER_GRAPH, 96, record=False)
            one = m1.last_run_ms() / 96
            m1.close()
        dist.barrier()
        bin_sharded = {"workload": "S-10M: one 10 000 000-bin spectrum over %d GPUs (contiguous bin ranges of equal estimated cost), 10 chains replicated" % world,
                       "ms_per_step": float(t10.item()) / 96, "chain_steps_per_s": C_CHAINS * 96 / (float(t10.item()) * 1e-3), "steps": 96,
                       "one_gpu_ms_per_step": one, "speedup_vs_one_gpu": (one / (float(t10.item()) / 96)) if one else None,
                       "exchange": "tg_k_xchg: 2*C doubles per rank and step over NVLink peer memory (CUDA IPC), inside the captured step graph; "
                                   "ncclAllReduce variant: tools/bin_sharded_nccl.py --nccl"}

    # ---- CPU baseline on this box's host cores (rank 0, N=1 only; bounded sample) ----
    cpu = None
    if rank == 0 and world == 1 and cfg_name in ("cfg3", "cfg5"):
        from oracle.oracle import Oracle, RefLib
        O, R = Oracle(), RefLib()
        prob = synth_problem(W, make_y=lambda p: O.model(p.model_case, p.params, p.plength, p.x)[1])
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
               "sample": "%d full steps (10 chains x 1M bins), generate_model under OpenMP over chains, box has %d cores; unmodified reference sources "
                         "compiled against the repo's Eigen-API shim (lean_port_value: the repo's plain-C restatement of the same arithmetic, same threads)" % (args.cpu_steps, ncpu),
               "lean_port_value": C_CHAINS * args.cpu_steps / dtp}

    if rank == 0:
        config = workload_config(nv, S)
        if cfg_name == "cfg2":
            config["workload"] = "cfg2 S-local: model_MS_local_basic, 10k bins, 10 chains, learning on"
        if cfg_name == "cfg4":
            config["workload"] = "cfg4 S-global with 1024 batched chains (128 replicated tempered ensembles of 8), Nx=1000000, learning on"
        config["warmup_steps_run"] = W_run
        line = {"metric": METRIC, "value": value, "unit": "chain-steps/s", "n_gpus": world, "steps": K, "warmup": args.warmup,
                "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
                "data": "synthetic", "config": config,
                "clocks": clk, "e2e": e2e, "gpu_launches": int(launches), "roofline": roofline, "cpu_baseline": cpu,
                "bin_evals_per_s": value * nx_star, "cfg5_8_stars_per_gpu": cfg5, "bin_sharded_10M": bin_sharded}
        txt = json.dumps(line)
        print(txt, flush=True)
        if args.extra_out:
            with open(args.extra_out, "w") as f:
                f.write(txt + "\n")
    if grp is not None:
        grp.close()
    for m_ in mds:
        m_.close()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
