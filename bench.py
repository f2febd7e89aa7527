#!/usr/bin/env python
"""Benchmark of the `ld` LD-score hot path (BASELINE.json metric: ordered window SNP-pairs/s,
additive + dominance).

    python bench.py [--gpus N --steps K --warmup W] [--impl reference]

Workload at N=1: BASELINE.json configs[1] -- synthetic chr22-sized bfile, 10,000 organisms x
100,000 SNPs, 2% missing, --ld-wind-cm 1 -maf 0.0001 --std-thr 1e-5.  With N ranks (torchrun)
every rank scores its own bfile of that shape (one chromosome per GPU, seed 2 + rank; no
data-path collective), so `value` / `e2e` scale weakly.

A step = one pass of the hot path over the rank's bfile:
  value : packed .bed bytes already resident in HBM; count -> stats -> unpack -> banded int8
          tcgen05 GEMM + FP64 epilogue -> finalize (nldsc_ld_session_run)
  e2e   : the reference-facing C-ABI call nldsc_ld_calculate_packed on a pinned HOST buffer --
          planning, device allocation, H2D, all kernels, D2H of the seven result arrays;
          e2e.bed_* = the same through nldsc_ld_calculate on a .bed file (what the CLI reaches)
Extra records:
  roofline        band kernel against the int8 tensor peak MEASURED in this run (back-to-back tcgen05 int8 MMAs,
                  nldsc_ld_debug_i8_peak); roofline.e2e = the H2D floor of the e2e call at this N
  parity_check    N=1: the timed GPU output against the reference's own C++ on the same bfile (the cpu_baseline run)
                  and against the FP64 oracle on random rows
  strong          N>1: ONE bfile (a cut of configs[4]: N=100,000, 10 cM window) scored by rank 0 through
                  nldsc_ld_calculate_packed with NLDSC_GPUS=N -- the north_star partition, contiguous SNP shards +
                  window halo -- against the same call on one GPU, with a bitwise comparison of the outputs
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
ARMA_NOTE = ("Armadillo stand-in (oracle/arma_shim): the reference's own sources compiled against a 100-line header whose "
             "dot / mean are 2-accumulator omp-simd float loops, not a real Armadillo + OpenBLAS sdot_")
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
                                          "-lms", "50", "-i", str(self.index)], stdout=subprocess.PIPE, text=True)
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


def cpu_model() -> str:
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name"):
                return line.split(":", 1)[1].strip()
    except OSError:
        pass
    return "unknown"


def use_real_blas() -> None:
    """Point the reference build at a real OpenBLAS sdot when the image has one (what Armadillo calls for n > 32 when
    linked to BLAS, reference CMakeLists.txt:24,43); read once per process, at the first dot."""
    if "NLDSC_REF_BLAS" not in os.environ:
        from oracle import ref
        lib = ref.find_openblas()
        os.environ["NLDSC_REF_BLAS"] = lib or ""


def blas_note() -> str:
    from oracle import ref
    name = ref.blas_name(32)
    lib = os.environ.get("NLDSC_REF_BLAS", "")
    if name.endswith("sdot"):
        return (f"Armadillo stand-in header (oracle/arma_shim) over the reference's own sources; arma::dot -> {name} of a real "
                f"OpenBLAS ({os.path.basename(lib)}), single-threaded per call inside the reference's OpenMP loop; arma::mean = omp-simd loop")
    return ARMA_NOTE


def time_reference(packed_np, cm, n_org, n_rows, threads=None, flavour=32, want_result=False):
    """Reference's own C++ (oracle/_ref, stock float build) on the first n_rows SNPs.
    -> (pairs, seconds[, the reference's seven output arrays])."""
    from oracle import ref
    if not ref.available(flavour):
        raise RuntimeError("oracle/_ref/libref_ld32.so is not built")
    use_real_blas()
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
    pairs = int(ws[ws >= 0].astype(np.int64).sum())
    return (pairs, dt, r) if want_result else (pairs, dt)


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
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "cpu_model": cpu_model(), "kind": "reference",
                         "sample": sample, "blas": blas_note()},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


def h2d_probe(host, local, barrier, reps=4):
    """Plain pinned -> device copies of the rank's packed buffer (torch copy engine path): the PCIe ceiling the e2e
    call cannot beat.  -> GB/s"""
    import torch
    dev = torch.empty(host.shape, dtype=torch.uint8, device=f"cuda:{local}")
    dev.copy_(host, non_blocking=True)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(reps):
        dev.copy_(host, non_blocking=True)
    e1.record()
    torch.cuda.synchronize()
    gbs = host.numel() * reps / (e0.elapsed_time(e1) * 1e-3) / 1e9
    del dev
    return gbs


def bed_path_leg(host, cm, rsq_thr, res, steps):
    """The entry point the reference CLI reaches: nldsc_ld_calculate on a .bed file -- page-cache warm, then once
    after the kernel was asked to drop the file's pages (posix_fadvise DONTNEED after fsync)."""
    from nldsc_b200 import capi, synth
    out = {}
    d = tempfile.mkdtemp(prefix="nldsc_bench_", dir=os.environ.get("NLDSC_BENCH_TMP"))
    path = os.path.join(d, "w.bed")
    try:
        with open(path, "wb") as f:
            f.write(synth.BED_MAGIC)
            f.write(host.numpy().tobytes())
            f.flush()
            os.fsync(f.fileno())
        a = (N_SNP, N_ORG, cm, WIND_CM, MAF, STD_THR, rsq_thr)
        for _ in range(2):
            got = capi.calculate(path, *a)
        n = max(3, min(steps, 10))
        t0 = time.perf_counter()
        for _ in range(n):
            got = capi.calculate(path, *a)
        out["bed_ms_warm"] = 1e3 * (time.perf_counter() - t0) / n
        for k in res:
            assert np.array_equal(res[k], got[k], equal_nan=True), ".bed path result differs from resident run: " + k
        fd = os.open(path, os.O_RDONLY)
        try:
            os.posix_fadvise(fd, 0, 0, os.POSIX_FADV_DONTNEED)
        finally:
            os.close(fd)
        t0 = time.perf_counter()
        capi.calculate(path, *a)
        out["bed_ms_cold"] = 1e3 * (time.perf_counter() - t0)
        out["bed_note"] = ("nldsc_ld_calculate on a 250 MB .bed under " + d.rsplit("/", 1)[0] + ": warm = page cache, cold = one call "
                           "after posix_fadvise(DONTNEED) (a tmpfs / overlay may ignore it; then cold == warm)")
    finally:
        try:
            os.remove(path)
            os.rmdir(d)
        except OSError:
            pass
    return out


def cpu_and_parity_leg(host, cm, rsq_thr, res):
    """The reference's own C++ on this box's host cores over (a bounded prefix of) the timed bfile, and the timed GPU
    output checked against what it returned; then the FP64 oracle on random rows of the whole file."""
    from nldsc_b200 import capi
    from oracle import ld_oracle
    cores = host_cores()
    pk = host.numpy()
    cal_pairs, cal_dt = time_reference(pk, cm, N_ORG, 768)
    budget_s = float(os.environ.get("NLDSC_BENCH_CPU_SECONDS", "30"))
    rows = int(min(N_SNP, max(3072, budget_s * (cal_pairs / cal_dt) / 640.0)))
    cpairs, cdt, ref = time_reference(pk, cm, N_ORG, rows, want_result=True)
    c1pairs, c1dt = time_reference(pk, cm, N_ORG, 512, threads=1)
    whole = rows == N_SNP
    cpu = {"value": cpairs / cdt, "unit": UNIT, "cores": cores, "cpu_model": cpu_model(), "kind": "reference",
           "single_thread_value": c1pairs / c1dt, "blas": blas_note(),
           "sample": (("the WHOLE workload bfile" if whole else f"first {rows} SNPs of the workload bfile") +
                      f", {cpairs} pairs, {cdt:.1f} s; stock float32 reference source (oracle/_ref/libref_ld32.so), "
                      "OpenMP over all host cores")}
    # ---- the timed GPU output against that run (float32 reference: filters and counts exact, scores within float noise)
    lo, hi = capi.plan_windows(cm, WIND_CM)
    inside = np.nonzero(hi < rows)[0]                       # rows whose whole window lies inside the prefix
    keep_g = ~np.isnan(res["l2"][inside])
    keep_r = ~np.isnan(ref["l2"][inside])
    k = inside[keep_g & keep_r]
    chk = {"reference_rows": int(len(inside)),
           "retained_list_equal": bool(np.array_equal(keep_g, keep_r)),
           "maf_bit_equal": bool(np.array_equal(res["maf"][inside], ref["maf"][inside], equal_nan=True)),
           "wsa_equal": bool(np.array_equal(res["l2_ws"][inside], ref["l2_ws"][inside])),
           "wsd_equal": bool(np.array_equal(res["l2d_ws"][inside], ref["l2d_ws"][inside])),
           "wsde_mismatches": int(np.sum(res["l2d_wse"][inside] != ref["l2d_wse"][inside])),
           "l2_max_rel_vs_float32_reference": float(np.max(np.abs(res["l2"][k] - ref["l2"][k]) / np.abs(ref["l2"][k]))) if len(k) else None,
           "l2d_max_abs_vs_float32_reference": float(np.max(np.abs(res["l2d"][k] - ref["l2d"][k]))) if len(k) else None,
           "note": "WSDE may differ where r2adj(dom) sits within float32 noise of rsq_thr (SURVEY 8c); L2/L2D differ from the "
                   "stock build by its float32 rounding, the 1e-9 bar is against the FP64 oracle below"}
    # ---- FP64 oracle (restatement pinned to the FP64 build of the reference's source) on random rows of the whole file
    rng = np.random.default_rng(20260)
    n_rows = int(os.environ.get("NLDSC_BENCH_ORACLE_ROWS", "64"))
    centres = rng.choice(N_SNP, size=max(1, n_rows // 8), replace=False)
    pick = np.unique(np.clip(np.concatenate([c + rng.choice(200, 8, replace=False) for c in centres]), 0, N_SNP - 1))
    t0 = time.perf_counter()
    want = ld_oracle.ld_scores_rows(pk, N_ORG, cm, WIND_CM, MAF, STD_THR, rsq_thr, pick, max_span=2600)
    r = want["rows"]
    kw = want["keep"]
    chk["oracle64_rows"] = int(len(r))
    chk["oracle64_seconds"] = time.perf_counter() - t0
    chk["oracle64_retained_equal"] = bool(np.array_equal(~np.isnan(res["l2"][r]), kw))
    chk["oracle64_maf_bit_equal"] = bool(np.array_equal(res["maf"][r], want["maf"], equal_nan=True))
    chk["oracle64_counts_equal"] = bool(all(np.array_equal(res[q][r][kw], want[q][kw]) for q in ("l2_ws", "l2d_ws", "l2d_wse")))
    if kw.any():
        chk["oracle64_l2_max_rel"] = float(np.max(np.abs(res["l2"][r][kw] - want["l2"][kw]) / np.maximum(np.abs(want["l2"][kw]), 1.0)))
        chk["oracle64_l2d_max_err"] = float(np.max(np.abs(res["l2d"][r][kw] - want["l2d"][kw]) / np.maximum(np.abs(want["l2d"][kw]), 1.0)))
    chk["ok"] = bool(chk["retained_list_equal"] and chk["maf_bit_equal"] and chk["wsa_equal"] and chk["wsd_equal"] and
                     chk["oracle64_retained_equal"] and chk["oracle64_maf_bit_equal"] and chk["oracle64_counts_equal"] and
                     chk.get("oracle64_l2_max_rel", 0.0) <= 1e-9 and chk.get("oracle64_l2d_max_err", 0.0) <= 1e-9)
    assert chk["ok"], "parity check of the timed run failed: " + json.dumps(chk)
    return cpu, chk


def strong_leg(n_gpus: int):
    """ONE bfile over n_gpus GPUs: contiguous SNP shards with a window halo (run_sharded), no collective.  Rank 0 runs
    this alone, after the process group is gone, through the same C-ABI call a user makes."""
    import torch
    from nldsc_b200 import capi, synth
    n_org = 100_000
    n_snp = int(os.environ.get("NLDSC_BENCH_STRONG_SNPS", "300000"))
    wind = 10.0
    spec = synth.SynthSpec(n_org=n_org, n_snp=n_snp, missing=0.02, seed=5, density=1000.0)
    t0 = time.perf_counter()
    dev_packed = synth.generate_packed_cuda(spec, "cuda:0")
    host = torch.empty(dev_packed.shape, dtype=torch.uint8, pin_memory=True)
    host.copy_(dev_packed)
    del dev_packed
    torch.cuda.empty_cache()
    gen_s = time.perf_counter() - t0
    cm, _ = synth.genetic_map(spec)
    a = (n_snp, n_org, cm, wind, MAF, STD_THR, 1.0 / n_snp)
    os.environ["NLDSC_DEVICE"] = "0"

    capi.report_device_times(True)            # per-shard device times in the record (0.2 ms of calls that take 0.2-2 s)

    def timed(gpus):
        os.environ["NLDSC_GPUS"] = str(gpus)
        out = capi.alloc_result(n_snp)
        capi.calculate_packed(host, *a, out=out)                   # first call: arenas, pinned staging, pools
        t = time.perf_counter()
        capi.calculate_packed(host, *a, out=out)
        dt = time.perf_counter() - t
        return out[0], dt, capi.last_call_report()

    one, t1, _ = timed(1)
    capi.arena_trim(0)
    many, tn, rep = timed(n_gpus)
    ws = one["l2_ws"]
    pairs = int(ws[ws >= 0].astype(np.int64).sum())
    equal = all(np.array_equal(one[k], many[k], equal_nan=True) for k in one)
    return {"workload": f"configs[4]: N={n_org:,} x {n_snp:,} SNPs, dense map (1,000 SNPs/cM), --ld-wind-cm 10, 2% missing; one bfile, "
                        f"host buffer -> nldsc_ld_calculate_packed -> host arrays, NLDSC_GPUS={n_gpus} vs NLDSC_GPUS=1",
            "n_gpus": n_gpus, "pairs": pairs, "ms_1gpu": 1e3 * t1, "ms_ngpu": 1e3 * tn, "speedup": t1 / tn,
            "efficiency": t1 / tn / n_gpus, "pairs_per_s_1gpu": pairs / t1, "pairs_per_s_ngpu": pairs / tn,
            "bitwise_equal_to_1gpu": bool(equal),
            "halo_rows_per_shard": [int(r["load_end"] - r["load_begin"] - (r["row_end"] - r["row_begin"])) for r in rep],
            "owned_rows_per_shard": [int(r["row_end"] - r["row_begin"]) for r in rep],
            "shard_ms_wall": [round(r["ms_wall"], 2) for r in rep],
            "shard_ms_device_run": [round(r["ms_dev_run"], 2) for r in rep],
            "shard_ms_device_band": [round(r["ms_dev_band"], 2) for r in rep],
            "shard_ms_h2d": [round(r["ms_dev_upload"], 2) for r in rep],
            "timing": "host wall clock around the C-ABI call (its shards are timed on their devices: shard_ms_device_*)",
            "generate_s": gen_s}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-strong", action="store_true", help="N>1: skip the one-bfile strong-scaling record")
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

    def reduce(x: float, op) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=f"cuda:{local}")
        dist.all_reduce(t, op=op)
        return float(t.item())

    RO = dist.ReduceOp if world > 1 else None
    max_over_ranks = lambda x: reduce(x, RO.MAX) if world > 1 else x      # noqa: E731
    min_over_ranks = lambda x: reduce(x, RO.MIN) if world > 1 else x      # noqa: E731
    sum_over_ranks = lambda x: reduce(x, RO.SUM) if world > 1 else x      # noqa: E731

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

    # ---- the int8 tensor-pipe ceiling of THIS box, measured now ----------------------
    # back-to-back tcgen05.mma.kind::i8 of the band kernel's shape (M=256, N=192, K=32; 8 per commit like the wide
    # kernel) on operands resident in shared memory; ~50 ms per launch so the clocks settle as they do under the kernel
    with ClockSampler(local) as pclk:
        pk192 = capi.debug_i8_peak(192, 16, 50_000, reps=4)       # the band kernel's instruction shape
        pk192_8 = capi.debug_i8_peak(192, 8, 100_000, reps=4)     # same, a commit every 8 MMAs (as the wide kernel issues them)
        pk256 = capi.debug_i8_peak(256, 8, 75_000, reps=4)
        # the pipe the band kernel actually runs on when the planes are fp4: tcgen05.mma.kind::mxf4, unit scale factors
        fp4_192 = capi.debug_fp4_probe(400_000, 192, 16, all_pairs=True)
        fp4_256 = capi.debug_fp4_probe(400_000, 256, 8, all_pairs=True)
    peak_clk = pclk.summary()

    # ---- roofline of the dominant kernel (banded int8 GEMM + epilogue) -------------------------
    peaks = _peaks()
    ops = pairs_rank * 12.0 * N_ORG               # 12*N int8 ops per ordered pair with missing data
    band_s = st["ms_band"] * 1e-3
    peak_int8 = max(pk192["tops"], pk256["tops"])      # the int8 pipe's ceiling, whichever shape shows it
    peak_fp4 = max(fp4_192["tops_equiv"], fp4_256["tops_equiv"])
    fp4 = bool(st["fp4"])
    # the roofline of the kernel is the pipe it runs on: fp4 (mxf4) when the planes are e2m1 nibbles, int8 otherwise; the
    # fraction of the INT8 peak -- north_star's yardstick -- is printed beside it in either case
    peak = peak_fp4 if fp4 else peak_int8
    peak_2x_bf16 = 2.0 * (peaks["bf16_tflops"] if peaks else 1590.0)
    achieved = ops / band_s / 1e12 if band_s > 0 else 0.0
    k_org = st["npad"] * (2 if fp4 else 1)             # padded organisms of a plane row (fp4 planes: two per byte)
    # executed MACs x 2: 256 x 192 per 64-SNP j-block of an i-pair, or 256 x 480 per 160-column supertile (band_x160_body)
    x160 = int(st.get("xtiles_run", 0)) > 0
    executed_units = st["xtiles_run"] * 480.0 if x160 else st["tiles"] * 192.0
    executed = (executed_units * 2.0 * 256 * k_org / band_s / 1e12) if band_s > 0 else 0.0
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "band_traffic.json")
    if os.path.exists(tpath):
        try:
            traffic = json.load(open(tpath)).get("dram_bytes_per_launch")
        except Exception:
            traffic = None
    roofline = {"bound": "tensor", "kernel": "band_tc_x160_kernel" if x160 else {0: "band_simt_kernel", 1: "band_tc_kernel", 2: "band_tc_wide_kernel"}[st["band_impl"]],
                "achieved": achieved, "peak": peak, "unit": "TOP/s", "frac": achieved / peak, "traffic": traffic,
                "peak_source": ("measured in this run: back-to-back tcgen05.mma.cta_group::2." +
                                ("kind::mxf4 (e2m1 operands holding 0/1/2, scale factors 1.0, exact integer sums; M=256, K=64)" if fp4
                                 else "kind::i8 (M=256, K=32)") +
                                " from shared memory on all 74 CTA pairs, no loads, no epilogue; the larger of N=192 with 16 MMAs per "
                                "commit and N=256 with 8; ops counted as int8-equivalent multiply-adds x 2 (12*N per ordered pair); "
                                "MEASURED_PEAKS.json has no int8 / fp4 entry"),
                "pipe": "fp4 (tcgen05.mma.kind::mxf4.block_scale, unit scales)" if fp4 else "int8 (tcgen05.mma.kind::i8)",
                "peak_int8": peak_int8, "frac_of_int8_peak": achieved / peak_int8, "peak_fp4": peak_fp4,
                "peak_sm_mhz": peak_clk["sm_mhz"], "peak_n192_16_per_commit": pk192["tops"], "peak_n192_8_per_commit": pk192_8["tops"],
                "peak_n256": pk256["tops"],
                "peak_2x_bf16": peak_2x_bf16, "frac_of_2x_bf16": achieved / peak_2x_bf16,
                "peak_nominal": 4500.0, "frac_of_nominal": achieved / 4500.0,
                "body": "band_x160_body (2 i-blocks x 160 j-SNPs, N = 256 + 224)" if x160 else "band_wide_body (2 i-blocks x 2 j-blocks, N = 192 + 192)",
                "executed": executed, "executed_frac": executed / peak, "executed_over_algorithmic": executed / achieved if achieved else None,
                "ms_band": st["ms_band"], "ms_unpack": st["ms_unpack"], "ms_count": st["ms_count"],
                "ms_stats": st["ms_stats"], "ms_finalize": st["ms_finalize"], "ms_step_device": dev_ms,
                "tiles": st["tiles"], "chunks": st["chunks"],
                # one band launch per chunk: `achieved` = ops per launch / mean launch time = ops per step / ms_band
                "launches_per_step": st["chunks"], "ops_per_launch": ops / max(1, st["chunks"]),
                "ms_per_launch": st["ms_band"] / max(1, st["chunks"])}
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
        roofline["unpack"] = {"bound": "hbm", "kernel": "unpack_fp4_kernel" if st1.get("fp4") else "unpack_kernel", "achieved": ub / (st1["ms_unpack"] * 1e-3) / 1e9,
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
    capi.report_device_times(True)            # one more call, untimed, for the device-side breakdown of the call
    capi.calculate_packed(host, N_SNP, N_ORG, cm, WIND_CM, MAF, STD_THR, rsq_thr, out=out)
    capi.report_device_times(False)
    rep = capi.last_call_report()[0]
    e2e = {"value": pairs_total * args.steps / e2e_wall, "unit": UNIT, "ms_per_step": 1e3 * e2e_wall / args.steps,
           "h2d_bytes_per_step": int(N_SNP * bps + 8 * N_SNP + 32 * st["tiles"]),
           "d2h_bytes_per_step": int(N_SNP * (4 * 8 + 3 * 4)),
           "last_call_ms": {k: round(rep[k], 3) for k in ("ms_wall", "ms_open", "ms_upload_start", "ms_plan", "ms_run_issue",
                                                           "ms_download", "ms_free", "ms_dev_upload", "ms_dev_run")}}
    for k in res:
        assert np.array_equal(res[k], out[0][k], equal_nan=True), "e2e result differs from resident run: " + k

    # ---- the floor under e2e: the packed bytes over PCIe, alone and with all N ranks copying at once ----
    if world > 1:
        alone = 0.0
        for r in range(world):                      # one rank at a time
            if r == rank:
                alone = h2d_probe(host, local, lambda: torch.cuda.synchronize())
            dist.barrier()
        alone = min_over_ranks(alone)
    else:
        alone = h2d_probe(host, local, barrier)
    n_way = min_over_ranks(h2d_probe(host, local, barrier))
    h2d_bytes = float(N_SNP * bps)
    floor_ms = h2d_bytes / (n_way * 1e9) * 1e3
    roofline["e2e"] = {"bound": "pcie", "h2d_bytes": int(h2d_bytes), "h2d_gbs_alone": alone, "h2d_gbs_n_way": n_way,
                       "h2d_floor_ms": floor_ms, "frac": floor_ms / e2e["ms_per_step"],
                       "note": "slowest rank's pinned-buffer copy rate, one rank at a time / all ranks at once; frac = the share of "
                               "the e2e call that is this floor (1.0 = the call costs no more than moving its input)"}

    # ---- the .bed entry point (rank 0 at N=1) ------------------------------------------------------
    if rank == 0 and world == 1:
        try:
            e2e.update(bed_path_leg(host, cm, rsq_thr, res, args.steps))
        except OSError as e:
            e2e["bed_note"] = f"unavailable: {e}"

    # ---- CPU baseline + parity of the timed run (rank 0, N=1 only) --------------------------------
    cpu, parity = None, None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        from oracle import ref
        if ref.available(32):
            cpu, parity = cpu_and_parity_leg(host, cm, rsq_thr, res)
        else:                                      # the checker is optional for the number, never for the product
            cpu = {"value": None, "unit": UNIT, "cores": host_cores(), "kind": "reference",
                   "sample": "unavailable: oracle/_ref/libref_ld32.so is not built"}

    # ---- strong scaling of ONE bfile over the N GPUs (rank 0 alone; the other ranks are done) ----------
    strong = None
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
        del host
        capi.arena_trim(-1)
        torch.cuda.empty_cache()
        if rank != 0:
            return
        if not args.no_strong:
            time.sleep(2.0)                        # the other ranks' processes exit and give their GPUs' memory back
            strong = strong_leg(world)

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": warmup,
            "ms_per_step": 1e3 * wall / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": ("exact integer cross-counts on the fp4 tensor pipe (e2m1 operands 0/1/2, unit scales, f32 accumulate) + f64 epilogue"
                      if st["fp4"] else "int8 cross-counts (int32 accumulate) + f64 epilogue"), "data": "synthetic",
            "config": {"workload": WORKLOAD, "pairs_per_step_per_gpu": pairs_rank, "parallelism": f"shard{world}",
                       "l2": "inputs larger than L2 (250 MB packed + 4 GB unpacked planes per step vs 126 MB)"},
            "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": launches, "clocks": clocks,
        }
        if parity is not None:
            line["parity_check"] = parity
        if strong is not None:
            line["strong"] = strong
        print(json.dumps(line))


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
