"""Print the measured tables of BASELINE.md from the JSON evidence committed under profiles/."""
import glob
import json
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
P = os.path.join(ROOT, "profiles")


def last_json(path):
    lines = [l for l in open(path).read().splitlines() if l.startswith("{")]
    return json.loads(lines[-1]) if lines else None


def bench_table():
    rows = []
    for f in sorted(glob.glob(os.path.join(P, "r1_bench_*.json"))):
        d = last_json(f)
        if not d or "value" not in d:
            continue
        rows.append((d.get("impl", "ours"), d["n_gpus"], d["config"].get("transport", "-"), d["value"], d["ms_per_step"],
                     d["e2e"]["value"], d["clocks"]["sm_mhz"], ",".join(d["clocks"]["reasons"]) or "none", os.path.basename(f)))
    rows.sort(key=lambda r: (r[0], r[1], r[2]))
    print("| arm | GPUs | aggregation transport | impressions/s (device-timed) | ms/step | impressions/s (end to end) | SM MHz | throttle | evidence |")
    print("|---|---|---|---|---|---|---|---|---|")
    for r in rows:
        print(f"| {r[0]} | {r[1]} | {r[2]} | {r[3]:,.0f} | {r[4]:.2f} | {r[5]:,.0f} | {r[6]:.0f} | {r[7]} | profiles/{r[8]} |")


def sweep_table():
    print("| GPUs | size | fused grad avg (clip+noise+mean) µs | NCCL all-reduce + clip/noise kernels µs | fused FedAvg + redistribute µs (bus GB/s) | NCCL all-reduce + scale µs (bus GB/s) | max abs err vs NCCL |")
    print("|---|---|---|---|---|---|---|")
    for f in sorted(glob.glob(os.path.join(P, "agg_sweep_w*.json"))):
        tag = os.path.basename(f)[len("agg_sweep_"):-5]
        for r in json.load(open(f)):
            g = r.get("fused_grad_clip_noise_us_b128", r["fused_grad_clip_noise_us"])
            pa = r.get("fused_param_avg_us_b128", r["fused_param_avg_us"])
            pb = r.get("fused_param_busbw_gbs_b128", r["fused_param_busbw_gbs"])
            print(f"| {tag} | {r['mb']:g} MB | {g:.1f} | {r['nccl_plus_clip_noise_kernels_us']:.1f} | {pa:.1f} ({pb:.0f}) | "
                  f"{r['nccl_allreduce_mean_us']:.1f} ({r['nccl_busbw_gbs']:.0f}) | {r['max_abs_err_vs_nccl']:.1e} |")


if __name__ == "__main__":
    bench_table()
    print()
    sweep_table()
