"""Turn `ncu --set full` reports (gpurun_out/ncu/*.ncu-rep) into committed, readable summaries under profiles/.

Runs on the CPU box: `ncu -i` only reads the report. For every report: duration, DRAM traffic, tensor-pipe and DRAM
utilisation, occupancy limits, the warp-stall mix and the hottest instructions from the source page, and the
achieved fraction of the measured roofline (MEASURED_PEAKS.json) from the algorithmic FLOPs / bytes of the launch.
"""
import csv
import io
import json
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PEAKS = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) \
    else {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}
M = int(os.environ.get("FEDREC_PROFILE_M", "27000"))

# algorithmic work of each profiled launch: (FLOPs, bytes that must cross HBM, note)
WORK = {
    "gemm_ffn1": (2.0 * M * 3072 * 768, M * 768 * 2 + 3072 * 768 * 2 + M * 3072 * 2, f"M={M} N=3072 K=768 bias+GELU"),
    "gemm_ffn2": (2.0 * M * 768 * 3072, M * 3072 * 2 + 768 * 3072 * 2 + 2 * M * 768 * 2, f"M={M} N=768 K=3072 bias+residual"),
    "gemm_qkv": (2.0 * M * 2304 * 768, M * 768 * 2 + 2304 * 768 * 2 + M * 2304 * 2, f"M={M} N=2304 K=768 bias"),
    "gemm_out": (2.0 * M * 768 * 768, M * 768 * 2 * 3 + 768 * 768 * 2, f"M={M} N=768 K=768 bias+residual"),
    "gemm_fold_ffn1": (2.0 * M * 3072 * 768, M * 768 * 2 + 3072 * 768 * 2 + M * 3072 * 2 + M * 96, f"M={M} N=3072 K=768 LN(A) folded + bias + GELU"),
    "gemm_fold_ffn2": (2.0 * M * 768 * 3072, M * 3072 * 2 + 768 * 3072 * 2 + 2 * M * 768 * 2 + 2 * M * 96, f"M={M} N=768 K=3072 bias + LN(residual) + row stats"),
    "gemm_fold_qkv": (2.0 * M * 2304 * 768, M * 768 * 2 + 2304 * 768 * 2 + M * 2304 * 2 + M * 96, f"M={M} N=2304 K=768 LN(A) folded + bias"),
    "gemm_fold_out": (2.0 * M * 768 * 768, M * 768 * 2 * 3 + 768 * 768 * 2 + 2 * M * 96, f"M={M} N=768 K=768 bias + LN(residual) + row stats"),
    "attention": (4.0 * (M // 30) * 12 * 30 * 30 * 64, M * 2304 * 2 + M * 768 * 2, f"{M // 30} titles x 12 heads, T=30"),
    "gemm_grad": (2.0 * M * 384 * 768, M * (384 + 768) * 2 + 384 * 768 * 4, f"dW[384,768] += dY^T X, K={M}"),
    "layernorm": (0.0, M * 768 * 2 * 2, f"{M} rows x 768 bf16"),
    "ue_attention": (2.0 * 50 * 20 * 50 * 50 * 20 * 2, 2500 * 1200 * 2 + 2500 * 400 * 2 + 50 * 20 * 50 * 8,
                     "50 users x 20 heads, L=50, dk=20 (useful FLOPs; the padded 128x128x32 tiles execute 13x more)"),
    "ue_attention_bwd": (2.0 * 50 * 20 * 50 * 50 * 20 * 5, 2 * 2500 * 1200 * 2 + 2500 * 400 * 2 + 50 * 20 * 50 * 8,
                         "50 users x 20 heads, L=50, dk=20, five products per (user, head)"),
    "aggregate": (0.0, 8 * 1_164_896 * 4 * 3, "8 logical ranks on one GPU, 4.66 MB each, clip + noise + mean"),
}


def ncu_csv(rep, page):
    out = subprocess.run(["ncu", "-i", rep, "--page", page, "--csv"], capture_output=True, text=True).stdout
    lines = [l for l in out.splitlines() if l.startswith('"')]
    return list(csv.reader(io.StringIO("\n".join(lines))))


def raw_metrics(rep):
    rows = ncu_csv(rep, "raw")
    if len(rows) < 3:
        return []
    hdr, units = rows[0], rows[1]
    return [dict(zip(hdr, r)) for r in rows[2:]], dict(zip(hdr, units))


def source_stalls(rep):
    rows = ncu_csv(rep, "source")
    # the source page repeats a two-line header per kernel; take the first kernel block
    hdr_idx = next((i for i, r in enumerate(rows) if r and r[0] == "Address"), None)
    if hdr_idx is None:
        return {}, []
    hdr = rows[hdr_idx]
    data = []
    for r in rows[hdr_idx + 1:]:
        if r and r[0] == "Kernel Name":
            break
        if len(r) == len(hdr):
            data.append(r)
    ix = {h: i for i, h in enumerate(hdr)}
    stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
    tot = {s: 0 for s in stalls}
    for r in data:
        for s in stalls:
            try:
                tot[s] += int(r[ix[s]])
            except ValueError:
                pass

    def samples(r):
        try:
            return int(r[ix["# Samples"]])
        except ValueError:
            return 0
    top = sorted(data, key=lambda r: -samples(r))[:8]
    hot = []
    for r in top:
        st = {s: int(r[ix[s]] or 0) for s in stalls}
        main = max(st, key=st.get)
        hot.append((samples(r), r[ix["Source"]].strip()[:100], main))
    return tot, hot


KEYS = [
    ("gpu__time_duration.sum", "duration"),
    ("launch__grid_size", "grid"), ("launch__block_size", "block"), ("launch__cluster_size", "cluster"),
    ("launch__registers_per_thread", "registers/thread"),
    ("launch__shared_mem_per_block_dynamic", "dynamic smem/block"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved occupancy %"),
    ("dram__bytes_read.sum", "DRAM read"), ("dram__bytes_write.sum", "DRAM write"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM throughput % of peak"),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor pipe active %"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM throughput %"),
    ("lts__t_sector_hit_rate.pct", "L2 hit rate %"),
    ("l1tex__m_xbar2l1tex_read_bytes.sum", "L2→SM bytes"),
    ("sm__inst_executed.sum", "warp instructions"),
]


def main():
    src = os.path.join(ROOT, "gpurun_out", "ncu")
    dst = os.path.join(ROOT, "profiles")
    os.makedirs(dst, exist_ok=True)
    index = []
    for f in sorted(os.listdir(src)):
        if not f.endswith(".ncu-rep"):
            continue
        name = f[:-8]
        rep = os.path.join(src, f)
        res = raw_metrics(rep)
        if not res:
            continue
        kernels, units = res
        lines = [f"# ncu capture: {name}", "",
                 "`ncu --set full --clock-control none --import-source on`, one B200, after the same command exited 0 without ncu.",
                 "Source: `scripts/profile_target.py " + name + "`; summarised by `scripts/make_profiles.py`.", ""]
        for k in kernels:
            kn = k.get("Kernel Name", "?")
            lines += [f"## `{kn[:160]}`", ""]
            dur_ns = None
            for key, label in KEYS:
                if key in k and k[key] != "":
                    lines.append(f"- {label}: {k[key]} {units.get(key, '')}")
                    if key == "gpu__time_duration.sum":
                        v = float(k[key].replace(",", ""))
                        u = units.get(key, "ns")
                        dur_ns = v * {"ns": 1, "us": 1e3, "ms": 1e6, "s": 1e9}.get(u, 1)
            if name in WORK and dur_ns and len(kernels) == 1:
                fl, by, note = WORK[name]
                lines.append(f"- work: {note}")
                if fl:
                    tf = fl / dur_ns / 1e3
                    lines.append(f"- achieved: {tf:.0f} TFLOP/s = {tf / PEAKS['bf16_tflops']:.2f} of measured cuBLAS bf16 peak "
                                 f"({PEAKS['bf16_tflops']:.0f} TFLOP/s)")
                gbs = by / dur_ns
                lines.append(f"- algorithmic HBM traffic {by / 1e6:.1f} MB → {gbs:.0f} GB/s = {gbs / PEAKS['hbm_gbs']:.2f} of measured "
                             f"copy bandwidth ({PEAKS['hbm_gbs']:.0f} GB/s)")
                lines.append("- (captured with a cold L2 under ncu serialisation; the in-step numbers are in BASELINE.md)")
            lines.append("")
        if len(kernels) == 1:
            tot, hot = source_stalls(rep)
            s = sum(tot.values())
            if s:
                lines += ["### warp-stall mix (PC samples)", ""]
                for st, v in sorted(tot.items(), key=lambda kv: -kv[1])[:7]:
                    lines.append(f"- {st}: {100.0 * v / s:.1f}%")
                lines += ["", "### hottest instructions", "", "| samples | SASS | dominant stall |", "|---|---|---|"]
                for n, srcline, st in hot:
                    lines.append(f"| {n} | `{srcline}` | {st} |")
                lines.append("")
        open(os.path.join(dst, f"ncu_{name}.md"), "w").write("\n".join(lines))
        index.append(name)
    print("wrote", index)


if __name__ == "__main__":
    main()
