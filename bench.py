#!/usr/bin/env python
"""Benchmark of the `ld` LD-score hot path (BASELINE.json metric: ordered window SNP-pairs/s,
additive + dominance).

    python bench.py [--gpus N --steps K --warmup W] [--impl reference]

Workload at N=1: BASELINE.json configs[1] -- synthetic chr22-sized bfile, 10,000 organisms x
100,000 SNPs, 2% missing, --ld-wind-cm 1 -maf 0.0001 --std-thr 1e-5.  With N ranks (torchrun)
every rank scores its own bfile of that shape (one chromosome per GPU, seed 2 + rank; no
data-path collective), so scaling is weak.

A step = one pass of the hot path over the rank's bfile:
  value : packed .bed bytes already resident in HBM; count -> stats -> unpack -> banded int8
          tcgen05 GEMM + FP64 epilogue -> finalize (nldsc_ld_session_run)
  e2e   : the reference-facing C-ABI call nldsc_ld_calculate_packed on a pinned HOST buffer --
          planning, device allocation, H2D, all kernels, D2H of the seven result arrays.
The oracle / compiled reference is only ever the checker or the timed CPU baseline here.
"""
from __future__ import annotations

import argparse
import ctypes
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

N_ORG, N_SNP, MISSING, WIND_CM, MAF, STD_THR = 10_000, 100_000, 0.02, 1.0, 1e-4, 1e-5
WORKLOAD = ("configs[1]: synthetic chr22-sized bfile N=10,000 x 100,000 SNPs, 2% missing, "
            "--ld-wind-cm 1 -maf 0.0001 --std-thr 1e-5")
METRIC = "window SNP-pairs/s (add+dom)"
UNIT = "pairs/s"


def _peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        return None


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self.proc = index, [], None

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.index)], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def __exit__(self, *a):
        if self.proc:
            time.sleep(0.15)
            self.proc.terminate()
            self.t.join(timeout=2)

    def summary(self):
        sm, mx, reasons = [], 0.0, set()
        for r in self.rows:
            try:
                sm.append(float(r[1])); mx = max(mx, float(r[2]))
            except Exception:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        # "under load" = the upper half of the samples (idle gaps between steps pull the median down)
        sm.sort()
        return {"sm_mhz": float(np.median(sm[len(sm) // 2:])), "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


def make_workload(rank: int, device: str):
    """-> (pinned host tensor [n_snp, bps] uint8, cM positions)."""
    import torch
    from nldsc_b200 import synth
    spec = synth.SynthSpec(n_org=N_ORG, n_snp=N_SNP, missing=MISSING, seed=2 + rank)
    dev_packed = synth.generate_packed_cuda(spec, device)
    host = torch.empty(dev_packed.shape, dtype=torch.uint8, pin_memory=True)
    host.copy_(dev_packed)
    del dev_packed
    torch.cuda.empty_cache()
    cm, _ = synth.genetic_map(spec)
    return host, cm, spec


def host_cores() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def time_reference(packed_np, cm, n_org, n_rows, threads=None, flavour=32):
    """Reference's own C++ (oracle/_ref, stock float build) on the first n_rows SNPs.
    -> (pairs, seconds, kind)."""
    from oracle import ref
    if not ref.available(flavour):
        raise RuntimeError("oracle/_ref/libref_ld32.so is not built")
    # torchrun exports OMP_NUM_THREADS=1 to every rank; the reference arm is meant to use all host cores
    ctypes.CDLL("libgomp.so.1", mode=ctypes.RTLD_GLOBAL).omp_set_num_threads(threads or host_cores())
    from nldsc_b200 import synth
    with tempfile.TemporaryDirectory() as d:
        with open(os.path.join(d, "s.bed"), "wb") as f:
            f.write(synth.BED_MAGIC)
            f.write(np.ascontiguousarray(packed_np[:n_rows]).tobytes())
        t0 = time.perf_counter()
        r = ref.calculate(os.path.join(d, "s.bed"), n_rows, n_org, cm[:n_rows], WIND_CM, MAF, STD_THR, 1.0 / N_SNP,
                          flavour=flavour)
        dt = time.perf_counter() - t0
    ws = r["l2_ws"]
    return int(ws[ws >= 0].sum()), dt


def run_reference_arm(args):
    """--impl reference: the reference's CPU implementation of the path on the host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from nldsc_b200 import synth
    cores = host_cores()
    # bounded sample: calibrate on 768 SNPs, then size each step for ~150 s over all steps
    spec = synth.SynthSpec(n_org=N_ORG, n_snp=N_SNP, missing=MISSING, seed=2)
    cm, _ = synth.genetic_map(spec)
    budget = 150.0 / max(1, args.steps + args.warmup)
    cal_rows = 768
    packed = synth.generate_packed(spec, (0, 4096))
    pairs, dt = time_reference(packed, cm, N_ORG, cal_rows)
    rate = pairs / dt
    rows = int(min(4096, max(cal_rows, budget * rate / 700.0 + 700)))
    for _ in range(args.warmup):
        time_reference(packed, cm, N_ORG, rows)
    tot_pairs, tot_t = 0, 0.0
    for _ in range(args.steps):
        pairs, dt = time_reference(packed, cm, N_ORG, rows)
        tot_pairs += pairs
        tot_t += dt
    v = tot_pairs / tot_t
    sample = f"first {rows} SNPs of the workload bfile (all N=10,000 organisms), {tot_pairs // args.steps} pairs per step"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * tot_t / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "sample": sample},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "reference", "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference_arm(args)

    import torch
    import torch.distributed as dist
    from nldsc_b200 import build, capi
    build.build_native()                      # no-op when the in-tree .so is current
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a B200: there is no CPU fallback for the LD-score path")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
    os.environ["NLDSC_DEVICE"] = str(local)
    os.environ["NLDSC_GPUS"] = "1"
    warmup = max(args.warmup, 3)

    host, cm, spec = make_workload(rank, f"cuda:{local}")
    rsq_thr = 1.0 / N_SNP
    bps = host.shape[1]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=f"cuda:{local}")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=f"cuda:{local}")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    # ---- value: inputs resident in HBM ---------------------------------------------------------
    sess = capi.Session(local, N_SNP, N_ORG, cm, WIND_CM, MAF, STD_THR, rsq_thr)
    sess.upload(host)
    sess.sync()
    for _ in range(warmup):
        sess.run()
    sess.sync()
    res = sess.download()
    ws = res["l2_ws"]
    pairs_rank = int(ws[ws >= 0].astype(np.int64).sum())
    dev_ms = []
    with ClockSampler(local) as clk:
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            sess.run()
        sess.sync()
        barrier()
        wall = time.perf_counter() - t0
    st = sess.stats()
    # per-step device time of the last step and per-kernel-family split (CUDA events on the launching stream)
    dev_ms = st["ms_total"]
    launches = st["kernel_launches"] * args.steps
    wall = max_over_ranks(wall)
    pairs_total = sum_over_ranks(float(pairs_rank))
    value = pairs_total * args.steps / wall
    clocks = clk.summary()

    # ---- roofline of the dominant kernel (banded int8 GEMM + epilogue) -------------------------
    peaks = _peaks()
    ops = pairs_rank * 12.0 * N_ORG               # 12*N int8 ops per ordered pair with missing data
    band_s = st["ms_band"] * 1e-3
    if peaks:
        peak = 2.0 * peaks["bf16_tflops"]
        peak_note = ("2 x measured burst bf16 (MEASURED_PEAKS.json); int8 dense = 2 x bf16 nominal. Understates the int8 pipe: "
                     "the band kernel sustains 4,060 TOP/s executed on a wide-window probe (profiles/r01g_band_probe.json)")
    else:
        peak = 2.0 * 1590.0
        peak_note = "2 x fallback bf16 1.59 PFLOP/s (B200_PROFILING.md)"
    achieved = ops / band_s / 1e12 if band_s > 0 else 0.0
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "band_traffic.json")
    if os.path.exists(tpath):
        try:
            traffic = json.load(open(tpath)).get("dram_bytes_per_launch")
        except Exception:
            traffic = None
    roofline = {"bound": "tensor", "kernel": {0: "band_simt_kernel", 1: "band_tc_kernel", 2: "band_tc_wide_kernel"}[st["band_impl"]], "achieved": achieved, "peak": peak, "unit": "TOP/s",
                "frac": achieved / peak, "traffic": traffic, "peak_source": peak_note,
                "ms_band": st["ms_band"], "ms_unpack": st["ms_unpack"], "ms_count": st["ms_count"],
                "ms_stats": st["ms_stats"], "ms_finalize": st["ms_finalize"], "ms_step_device": dev_ms,
                "tiles": st["tiles"], "chunks": st["chunks"],
                # one band launch per chunk: `achieved` = ops per launch / mean launch time = ops per step / ms_band
                "launches_per_step": st["chunks"], "ops_per_launch": ops / max(1, st["chunks"]),
                "ms_per_launch": st["ms_band"] / max(1, st["chunks"]),
                "executed_frac": (st["tiles"] * 2.0 * 256 * 192 * st["npad"] / band_s / 1e12 / peak) if band_s > 0 else 0.0}
    # secondary kernel, HBM-bound: time unpack ALONE (one step with the two streams serialised; beside the band
    # kernel its span is stretched by the contention).  Algorithmic bytes per SNP: ceil(N/4) read + 4 planes * npad written.
    os.environ["NLDSC_OVERLAP"] = "0"
    sess.run()
    sess.sync()
    os.environ.pop("NLDSC_OVERLAP")
    st1 = sess.stats()
    ub = float(st1["load_end"] - st1["load_begin"]) * (bps + 4.0 * st1["npad"])
    hbm_peak = peaks["hbm_gbs"] if peaks else 6500.0
    if st1["ms_unpack"] > 0:
        roofline["unpack"] = {"bound": "hbm", "kernel": "unpack_kernel", "achieved": ub / (st1["ms_unpack"] * 1e-3) / 1e9,
                              "peak": hbm_peak, "unit": "GB/s", "frac": ub / (st1["ms_unpack"] * 1e-3) / 1e9 / hbm_peak,
                              "ms_alone": st1["ms_unpack"], "ms_beside_band": st["ms_unpack"]}
    sess.close()

    # ---- e2e: host buffer -> C ABI -> host results ----------------------------------------------
    out = capi.alloc_result(N_SNP)
    for _ in range(2):
        capi.calculate_packed(host, N_SNP, N_ORG, cm, WIND_CM, MAF, STD_THR, rsq_thr, out=out)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        capi.calculate_packed(host, N_SNP, N_ORG, cm, WIND_CM, MAF, STD_THR, rsq_thr, out=out)
    barrier()
    e2e_wall = max_over_ranks(time.perf_counter() - t0)
    e2e = {"value": pairs_total * args.steps / e2e_wall, "unit": UNIT, "ms_per_step": 1e3 * e2e_wall / args.steps,
           "h2d_bytes_per_step": int(N_SNP * bps + 8 * N_SNP + 32 * st["tiles"]),
           "d2h_bytes_per_step": int(N_SNP * (4 * 8 + 3 * 4))}
    for k in res:
        assert np.array_equal(res[k], out[0][k], equal_nan=True), "e2e result differs from resident run: " + k

    # ---- CPU baseline: the reference's own C++ on this box's host cores (rank 0, N=1 only) ------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            rows = 3072
            pk = host[:rows].numpy()
            cpairs, cdt = time_reference(pk, cm, N_ORG, rows)
            c1pairs, c1dt = time_reference(pk, cm, N_ORG, 512, threads=1)
            cpu = {"value": cpairs / cdt, "unit": UNIT, "cores": host_cores(), "kind": "reference",
                   "single_thread_value": c1pairs / c1dt,
                   "sample": f"first {rows} SNPs of the workload bfile, {cpairs} pairs, {cdt:.1f} s; stock float32 "
                             "reference source (oracle/_ref/libref_ld32.so), OpenMP over all host cores"}
            # parity spot check of the timed GPU run against that same reference output is in tests/
        except Exception as e:                     # the checker is optional for the number, never for the product
            cpu = {"value": None, "unit": UNIT, "cores": host_cores(), "kind": "reference",
                   "sample": f"unavailable: {e}"}

    if rank == 0:
        print(json.dumps({
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": warmup,
            "ms_per_step": 1e3 * wall / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "int8 cross-counts (int32 accumulate) + f64 epilogue", "data": "synthetic",
            "config": {"workload": WORKLOAD, "pairs_per_step_per_gpu": pairs_rank, "parallelism": f"shard{world}",
                       "l2": "inputs larger than L2 (250 MB packed + 4 GB unpacked planes per step vs 126 MB)"},
            "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": launches, "clocks": clocks,
        }))
    if world > 1:
        dist.destroy_process_group()


def _only_json_on_stdout():
    """Libraries (NCCL prints its version banner on stdout) must not add lines to the one-JSON-line contract:
    everything written to fd 1 goes to stderr; the returned file is the real stdout, for the JSON line alone."""
    sys.stdout.flush()
    real = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    return real


if __name__ == "__main__":
    _real_stdout = _only_json_on_stdout()
    _print = print

    def print(*a, **k):          # noqa: A001 -- the JSON line (the only print in this file) goes to the real stdout
        k.setdefault("file", _real_stdout)
        _print(*a, **k)
        _real_stdout.flush()

    main()
