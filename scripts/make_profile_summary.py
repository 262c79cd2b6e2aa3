#!/usr/bin/env python
"""Copies one capture of scripts/gpu_round.sh from gpurun_out/ into profiles/ and writes its summaries (no GPU needed).
usage: make_profile_summary.py TAG"""
import collections, csv, json, os, shutil, subprocess, sys
T = sys.argv[1]
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G, P = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")
for f in ("bench.json", "bench_reference.json", "bench_2000steps.json", "gputests.log", "launches.csv"):
    shutil.copy(os.path.join(G, "%s_%s" % (T, f)), os.path.join(P, "%s_%s" % (T, f)))
S = os.path.join(ROOT, "scripts", "ncu_summary.py")
def summary(rep, kern, out, which=0):
    subprocess.run([sys.executable, S, os.path.join(G, rep), kern, os.path.join(P, out), str(which)], check=True, stdout=subprocess.DEVNULL)
step = "%s_step_full.ncu-rep" % T
txt = subprocess.run(["ncu", "-i", os.path.join(G, step), "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(txt.splitlines())); head = rows[0]; ik, ig = head.index("Kernel Name"), head.index("launch__grid_size")
prop = [(i, int(float(r[ig]))) for i, r in enumerate(r for r in rows[2:] if "tg_k_propose" in r[ik])]
summary(step, "tg_k_eval", "%s_eval_kernel_ncu_summary" % T)
summary(step, "tg_k_ff", "%s_prepass_kernel_ncu_summary" % T)
summary(step, "tg_k_finish", "%s_finish_kernel_ncu_summary" % T)
main = min(prop, key=lambda t: t[1])[0]; side = max(prop, key=lambda t: t[1])[0]
summary(step, "tg_k_propose", "%s_propose_kernel_ncu_summary" % T, main)
summary(step, "tg_k_propose", "%s_propose_side_kernel_ncu_summary" % T, side)
c5 = "%s_cfg5_full.ncu-rep" % T
if os.path.exists(os.path.join(G, c5)):
    summary(c5, "tg_k_eval", "%s_eval_cfg5_80chains_ncu_summary" % T)
    summary(c5, "tg_k_ff(", "%s_prepass_cfg5_80chains_ncu_summary" % T)
rows = [r for r in csv.reader(open(os.path.join(P, "%s_launches.csv" % T))) if len(r) > 10 and r[0].isdigit()]
t = collections.OrderedDict()
for r in rows:
    t.setdefault(r[4].split("(")[0], []).append(float(r[-1]) / 1e3)
tot = sum(sum(v) for v in t.values())
b, b2, ref = (json.load(open(os.path.join(P, "%s_%s.json" % (T, k)))) for k in ("bench", "bench_2000steps", "bench_reference"))
r, c5j = b["roofline"], b.get("cfg5_8_stars_per_gpu") or {}
with open(os.path.join(P, "%s_launches.md" % T), "w") as f:
    f.write("# Capture %s: launch list of `python bench.py --steps 40 --warmup 40 --no-cfg5`\n\n" % T)
    f.write("`ncu --metrics gpu__time_duration.sum --clock-control none -s 300 -c 400` (per-launch times are cold-cache and serialised: the SHARES are what must agree with the live run). "
            "A sampler step is five launches: `tg_k_propose` twice (proposal + prologue of every chain on the main stream; the adaptation candidates and log-prior CTAs as a second launch on a side stream, "
            "concurrent with everything up to the decision in the live run), the far-field pre-pass `tg_k_ff`, the fused model + likelihood kernel `tg_k_eval`, the decision kernel `tg_k_finish`. "
            "The other kernels in the window belong to the `tg_eval` (e2e) leg of the bench: `tg_k_prologue`, `tg_k_ff`, `tg_k_eval`, `tg_k_reduce_finalize` per call.\n\n")
    f.write("| kernel | launches | avg us | share of captured time |\n|---|---|---|---|\n")
    for k, v in t.items():
        f.write("| `%s` | %d | %.1f | %.3f |\n" % (k, len(v), sum(v) / len(v), sum(v) / tot))
    f.write("\nLive run of the same build, driver protocol `--steps 20 --warmup 5` (`profiles/%s_bench.json`, CUDA events, no profiler): %.1f k chain-steps/s, %.4f ms/step; `tg_k_eval` %.1f us per launch "
            "(share %.2f of the step; algorithmic flops / time / peak = %.2f, executed flops / time / peak = %.2f), `tg_k_ff` %.1f us timed alone with events (less inside the live step, where it overlaps its neighbours: `scripts/dev/step_timeline.py`), "
            "the two kernels that evaluate the model together %.2f of the step. 2000-step run (`%s_bench_2000steps.json`): %.1f k chain-steps/s. e2e through `tg_eval_async`/`tg_eval_wait` with pinned host buffers: %.1f k chain-steps/s "
            "(blocking `tg_eval`: %.1f k). Reference arm on the same box (`%s_bench_reference.json`): %.1f chain-steps/s on %d of the host cores; lean C port %.0f. " %
            (T, b["value"] / 1e3, b["ms_per_step"], r["avg_launch_us"], r["kernel_share_of_step"], r["frac"], r["frac_executed"], r["prepass_avg_launch_us"], r["model_kernels_share_of_step"],
             T, b2["value"] / 1e3, b["e2e"]["value"] / 1e3, b["e2e"]["blocking_call_value"] / 1e3, T, ref["value"], ref["cpu_baseline"]["cores"], b["cpu_baseline"]["lean_port_value"]))
    if c5j:
        f.write("8 stars x 10 chains per GPU as one launch group (cfg5): %.1f k chain-steps/s, %.4f ms per step of 80 chains, fused kernel %.1f us (algorithmic fraction %.2f, executed %.2f), pre-pass %.1f us; %.2f x the one-star figure. " %
                (c5j["value"] / 1e3, c5j["ms_per_step"], c5j["fused_kernel_us"], c5j["roofline_frac"], c5j["roofline_frac_executed"], c5j["prepass_us"], c5j["per_gpu_vs_one_star"]))
    f.write("GPU tests on the same box: `%s_gputests.log` (%s).\n" % (T, open(os.path.join(P, "%s_gputests.log" % T)).read().strip().splitlines()[-2].strip()))
print(open(os.path.join(P, "%s_launches.md" % T)).read())
