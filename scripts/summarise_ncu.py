"""Summarise gpurun_out ncu captures into profiles/ (tracked).
    python scripts/summarise_ncu.py <tag> [rep=gpurun_out/band_tc.ncu-rep] [launches=gpurun_out/launches.csv]
"""
import collections
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KEEP = ["gpu__time_duration.sum", "sm__cycles_elapsed.max", "launch__registers_per_thread", "launch__grid_size",
        "launch__block_size", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_subpipe_imma_cycles_active_realtime.avg",
        "sm__pipe_fp64_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct", "lts__t_bytes.sum",
        "l1tex__data_pipe_tc_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
        "sm__pipe_shared_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.per_cycle_active",
        "smsp__inst_executed.sum", "sm__inst_executed_pipe_uniform.sum"]


def main():
    tag = sys.argv[1]
    rep = sys.argv[2] if len(sys.argv) > 2 else os.path.join(ROOT, "gpurun_out", "band_tc.ncu-rep")
    launches = sys.argv[3] if len(sys.argv) > 3 else os.path.join(ROOT, "gpurun_out", "launches.csv")
    out = []
    if os.path.exists(rep):
        raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
        rows = list(csv.reader(io.StringIO(raw)))
        hdr, units, data = rows[0], rows[1], rows[2:]
        name_col = hdr.index("Kernel Name")
        out.append(f"# ncu --set full --clock-control none ({os.path.basename(rep)}), one column per captured launch")
        out.append("kernel: " + ", ".join(r[name_col].split("(")[0] for r in data))
        traffic = []
        for i, h in enumerate(hdr):
            if h in KEEP or any(h.endswith("." + k) or h.endswith(k) for k in KEEP if "tensor" in k or "fp64" in k):
                out.append(f"{h} [{units[i]}]: " + ", ".join(r[i] for r in data))
        rd, wr = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
        scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}
        for r in data:
            traffic.append(float(r[rd]) * scale[units[rd]] + float(r[wr]) * scale[units[wr]])
        out.append("dram bytes per launch (read+write): " + ", ".join(f"{t:.4g}" for t in traffic))
        json.dump({"dram_bytes_per_launch": sum(traffic) / len(traffic), "source": f"profiles/{tag}_band_tc_ncu.txt"},
                  open(os.path.join(ROOT, "profiles", "band_traffic.json"), "w"))
        open(os.path.join(ROOT, "profiles", f"{tag}_band_tc_ncu.txt"), "w").write("\n".join(out) + "\n")
    if os.path.exists(launches):
        lines = [l for l in open(launches) if not l.startswith("==")]
        rows = list(csv.DictReader(io.StringIO("".join(lines))))
        tot = collections.Counter()
        cnt = collections.Counter()
        for r in rows:
            if r.get("Metric Name") != "gpu__time_duration.sum":
                continue
            v = float(r["Metric Value"].replace(",", ""))
            u = r["Metric Unit"]
            v *= {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(u, 1e-6)
            k = r["Kernel Name"].split("(")[0]
            tot[k] += v
            cnt[k] += 1
        s = sum(tot.values())
        o = ["# ncu --metrics gpu__time_duration.sum --clock-control none: every launch of `bench.py --steps 1 --warmup 3 "
             "--no-cpu-baseline` (cold-cache, serialised: compare SHARES)", f"{'kernel':60s} {'launches':>8s} {'ms':>10s} {'share':>7s}"]
        for k, v in tot.most_common():
            o.append(f"{k[:60]:60s} {cnt[k]:8d} {v:10.3f} {100 * v / s:6.1f}%")
        # the step's kernels (the int8 peak probe and the plan-time kernels are the engine's too, but not part of a step)
        ours = [k for k in tot if any(t in k for t in ("band_", "unpack_", "count_", "stats_", "finalize_", "zero_rows", "chunk_mfree", "publish_rows"))]
        so = sum(tot[k] for k in ours)
        o.append("")
        o.append("# shares among the engine's own kernels (the synthetic-data generator's torch kernels excluded)")
        for k in sorted(ours, key=lambda k: -tot[k]):
            o.append(f"{k[:60]:60s} {cnt[k]:8d} {tot[k]:10.3f} {100 * tot[k] / so:6.1f}%")
        open(os.path.join(ROOT, "profiles", f"{tag}_launches.txt"), "w").write("\n".join(o) + "\n")
        print("\n".join(o[-8:]))
    print("\n".join(out))


if __name__ == "__main__":
    main()
