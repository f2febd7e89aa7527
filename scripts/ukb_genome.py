"""BASELINE.json configs[3] / north_star Target: UK-Biobank-scale synthetic genome -- N=315,599 organisms x 22 chromosomes
(~1.2M SNPs, SNP counts proportional to the chromosomes' genetic lengths), 2% missing, 1 cM window, additive + dominance --
scored END TO END on the GPUs of this box, no file involved: every chromosome is generated on a GPU straight into pinned
host memory (94.7 GB in total) and scored from there through the C ABI (nldsc_ld_calculate_packed_on: H2D, all kernels,
D2H of the seven arrays), whole chromosomes dealt largest-first to one host thread per GPU (the genome driver's schedule).

    python scripts/ukb_genome.py [gpus] [total_snps]

Two modes, chosen from the host memory the box has:
  resident  all chromosomes generated first (timed separately), then ONE timed scoring phase over all GPUs
  stream    each GPU thread generates, scores and drops one chromosome after the other (bounded pinned memory);
            scoring time is summed per GPU and the end-to-end figure is the busiest GPU's sum
The FP64 oracle is run on >= 20 random (chromosome, row) slices (row + its whole window), and chromosome 22 is also scored
through the real .bed entry point: page-cache warm, cold (posix_fadvise DONTNEED) and cold with O_DIRECT.
Writes gpurun_out/r02_ukb_genome_<gpus>gpu.json."""
import dataclasses
import json
import os
import queue
import shutil
import sys
import tempfile
import threading
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from nldsc_b200 import capi, synth
from oracle import ld_oracle

gpus = int(sys.argv[1]) if len(sys.argv) > 1 else torch.cuda.device_count()
total_snps = int(sys.argv[2]) if len(sys.argv) > 2 else 1_200_000
N = 315_599
WIND, MAF, STD = 1.0, 1e-4, 1e-5
CM_LEN = [286, 269, 223, 215, 204, 192, 187, 168, 166, 181, 158, 175, 126, 120, 141, 134, 128, 118, 108, 108, 62, 73]
CHROMS = list(range(1, 23))
SIZES = {c: max(1024, int(total_snps * CM_LEN[c - 1] / sum(CM_LEN)) // 64 * 64) for c in CHROMS}
BPS = (N + 3) // 4
ORACLE_ROWS = {c: (2 if c <= 2 else 1) for c in CHROMS}          # 24 slices
if os.environ.get("NLDSC_UKB_ORACLE_CHROMS"):                    # a re-run that checks fewer slices: only these chromosomes
    _only = {int(x) for x in os.environ["NLDSC_UKB_ORACLE_CHROMS"].split(",")}
    ORACLE_ROWS = {c: (1 if c in _only else 0) for c in CHROMS}


def mem_available() -> int:
    avail = 0
    for line in open("/proc/meminfo"):
        if line.startswith("MemAvailable"):
            avail = int(line.split()[1]) * 1024
    try:
        lim = open("/sys/fs/cgroup/memory.max").read().strip()
        if lim != "max":
            used = int(open("/sys/fs/cgroup/memory.current").read())
            avail = min(avail, int(lim) - used)
    except OSError:
        pass
    return avail


def spec_of(c):
    return dataclasses.replace(synth.SynthSpec(n_org=N, n_snp=SIZES[c], missing=0.02, seed=40 + c), chrom=c)


total_bytes = sum(SIZES.values()) * BPS
largest = max(SIZES.values()) * BPS
avail = mem_available()
mode = os.environ.get("NLDSC_UKB_MODE") or ("resident" if avail > 1.25 * total_bytes + (24 << 30) else "stream")
n_threads = gpus
if mode == "stream":
    n_threads = max(1, min(gpus, int(0.6 * avail // (largest + (4 << 30)))))
print(f"[ukb] {gpus} GPUs, {sum(SIZES.values())} SNPs, {total_bytes / 1e9:.1f} GB packed, host memory available "
      f"{avail / 1e9:.0f} GB -> mode {mode}, {n_threads} GPU threads", file=sys.stderr)

capi.warmup(0, n_threads)
capi.report_device_times(True)
order = sorted(CHROMS, key=lambda c: -SIZES[c])
lock = threading.Lock()
oracle_gate = threading.Semaphore(6)
records, checks, errors = {}, [], []
hosts = {}


def generate(c, device, buf=None):
    spec = spec_of(c)
    t0 = time.perf_counter()
    dev = synth.generate_packed_cuda(spec, f"cuda:{device}")
    if buf is None:
        buf = torch.empty(dev.shape, dtype=torch.uint8, pin_memory=True)
    host = buf.view(-1)[: dev.numel()].view(dev.shape)
    host.copy_(dev)
    torch.cuda.synchronize(device)
    del dev
    torch.cuda.empty_cache()
    cm, _ = synth.genetic_map(spec)
    return host, cm, time.perf_counter() - t0


def score(c, device, host, cm):
    n = SIZES[c]
    out = capi.alloc_result(n)
    t0 = time.perf_counter()
    got = capi.calculate_packed_on(host, n, N, cm, WIND, MAF, STD, 1.0 / n, device, 1, out=out)
    dt = time.perf_counter() - t0
    rep = capi.last_call_report()[0]
    keep = got["l2_ws"] >= 0
    rec = {"chr": c, "device": device, "snps": n, "bytes": n * BPS, "seconds": dt, "kept": int(keep.sum()),
           "pairs": int(got["l2_ws"][keep].astype(np.int64).sum()), "h2d_gbs": n * BPS / max(rep["ms_dev_upload"], 1e-3) / 1e6,
           "ms_dev_upload": rep["ms_dev_upload"], "ms_dev_run": rep["ms_dev_run"], "ms_dev_band": rep["ms_dev_band"]}
    return got, rec


def oracle_check(c, host, cm, got):
    n = SIZES[c]
    rng = np.random.default_rng(900 + c)
    rows = rng.choice(np.nonzero(got["l2_ws"] >= 0)[0], ORACLE_ROWS[c], replace=False)
    hn = host.numpy()
    for r in rows:
        with oracle_gate:
            t0 = time.perf_counter()
            w = ld_oracle.ld_scores_rows(hn, N, cm, WIND, MAF, STD, 1.0 / n, [int(r)], max_span=1)
            secs = time.perf_counter() - t0
        chk = {"chr": c, "row": int(r), "window": int(w["l2_ws"][0]),
               "counts_equal": bool(all(got[k][r] == w[k][0] for k in ("l2_ws", "l2d_ws", "l2d_wse"))),
               "maf_bit_equal": bool(got["maf"][r] == w["maf"][0]),
               "l2_rel_err": float(abs(got["l2"][r] - w["l2"][0]) / max(abs(w["l2"][0]), 1.0)),
               "l2d_err": float(abs(got["l2d"][r] - w["l2d"][0]) / max(abs(w["l2d"][0]), 1.0)),
               "rstd_err": float(abs(got["residuals_std"][r] - w["residuals_std"][0])), "oracle_seconds": secs}
        with lock:
            checks.append(chk)


t_all = time.perf_counter()
gen_s_total = 0.0
if mode == "resident":
    # ---- phase 1: generation (not part of the scored time), round-robin over the GPUs
    t0 = time.perf_counter()
    q = queue.Queue()
    for c in order:
        q.put(c)

    def gen_worker(device):
        while True:
            try:
                c = q.get_nowait()
            except queue.Empty:
                return
            try:
                host, cm, s = generate(c, device)
                with lock:
                    hosts[c] = (host, cm, s)
            except Exception as e:
                with lock:
                    errors.append(f"generate chr{c}: {e}")

    th = [threading.Thread(target=gen_worker, args=(d,)) for d in range(gpus)]
    [t.start() for t in th]
    [t.join() for t in th]
    gen_s_total = time.perf_counter() - t0
    if errors:
        raise SystemExit("\n".join(errors))
    # ---- phase 2: THE timed region -- all 22 chromosomes, host memory -> GPUs -> host arrays
    results = {}
    q = queue.Queue()
    busy = {d: 0.0 for d in range(gpus)}

    def score_worker(device):
        while True:
            try:
                c = q.get_nowait()
            except queue.Empty:
                return
            try:
                host, cm, _ = hosts[c]
                got, rec = score(c, device, host, cm)
                with lock:
                    results[c] = got
                    records[c] = rec
                    busy[device] += rec["seconds"]
            except Exception as e:
                with lock:
                    errors.append(f"score chr{c}: {e}")

    walls = []
    for rep in range(2):                                   # first pass pays arenas / first-touch, second is the figure
        for c in order:
            q.put(c)
        for d in busy:
            busy[d] = 0.0
        t0 = time.perf_counter()
        th = [threading.Thread(target=score_worker, args=(d,)) for d in range(gpus)]
        [t.start() for t in th]
        [t.join() for t in th]
        walls.append(time.perf_counter() - t0)
        if errors:
            raise SystemExit("\n".join(errors))
    score_wall = walls[-1]
    t0 = time.perf_counter()
    th = [threading.Thread(target=lambda cs=cs: [oracle_check(c, hosts[c][0], hosts[c][1], results[c]) for c in cs])
          for cs in (order[k::6] for k in range(6))]
    [t.start() for t in th]
    [t.join() for t in th]
    oracle_s = time.perf_counter() - t0
else:
    q = queue.Queue()
    for c in order:
        q.put(c)
    busy = {d: 0.0 for d in range(n_threads)}
    walls = []

    def stream_worker(device):
        buf = torch.empty(largest, dtype=torch.uint8, pin_memory=True)
        first = True
        while True:
            try:
                c = q.get_nowait()
            except queue.Empty:
                return
            try:
                host, cm, s = generate(c, device, buf)
                if first:                                   # the device's arenas, pinned plan staging etc.
                    score(c, device, host, cm)
                    first = False
                got, rec = score(c, device, host, cm)
                rec["generate_s"] = s
                with lock:
                    records[c] = rec
                    busy[device] += rec["seconds"]
                oracle_check(c, host, cm, got)
            except Exception as e:
                with lock:
                    errors.append(f"chr{c}: {e}")

    th = [threading.Thread(target=stream_worker, args=(d,)) for d in range(n_threads)]
    [t.start() for t in th]
    [t.join() for t in th]
    if errors:
        raise SystemExit("\n".join(errors))
    score_wall = max(busy.values())
    gen_s_total = sum(r.get("generate_s", 0.0) for r in records.values())
    oracle_s = sum(c["oracle_seconds"] for c in checks)
total_wall = time.perf_counter() - t_all

# ---- one chromosome through the real .bed entry point: warm, cold, cold + O_DIRECT --------------------------
bed = {}
try:
    c = 22
    tmp = tempfile.mkdtemp(prefix="ukb_", dir=os.environ.get("NLDSC_BENCH_TMP"))
    free = shutil.disk_usage(tmp).free
    if free > SIZES[c] * BPS * 1.2:
        host, cm, _ = hosts[c] if c in hosts else generate(c, 0)
        path = os.path.join(tmp, "chr22.bed")
        with open(path, "wb") as f:
            f.write(synth.BED_MAGIC)
            hn = host.numpy()
            for r0 in range(0, SIZES[c], 2048):
                f.write(hn[r0:r0 + 2048].tobytes())
            f.flush()
            os.fsync(f.fileno())
        a = (SIZES[c], N, cm, WIND, MAF, STD, 1.0 / SIZES[c])
        os.environ["NLDSC_DEVICE"], os.environ["NLDSC_GPUS"] = "0", "1"
        gb = SIZES[c] * BPS / 1e9

        def drop():
            fd = os.open(path, os.O_RDONLY)
            os.posix_fadvise(fd, 0, 0, os.POSIX_FADV_DONTNEED)
            os.close(fd)

        def timed():
            t0 = time.perf_counter()
            r = capi.calculate(path, *a)
            return time.perf_counter() - t0, r
        timed()
        t_warm, r_warm = timed()
        drop()
        t_cold, _ = timed()
        os.environ["NLDSC_ODIRECT"] = "1"
        drop()
        t_direct, r_direct = timed()
        t_direct2, _ = timed()
        os.environ.pop("NLDSC_ODIRECT")
        mem = capi.calculate_packed_on(host, *a, 0, 1)
        bed = {"chr": c, "GB": gb, "warm_s": t_warm, "warm_gbs": gb / t_warm, "cold_s": t_cold, "cold_gbs": gb / t_cold,
               "odirect_cold_s": t_direct, "odirect_cold_gbs": gb / t_direct, "odirect_again_s": t_direct2,
               "odirect_again_gbs": gb / t_direct2, "fs": tmp,
               "bitwise_equal_to_host_buffer": bool(all(np.array_equal(mem[k], r_warm[k], equal_nan=True) and
                                                        np.array_equal(mem[k], r_direct[k], equal_nan=True) for k in mem))}
    else:
        bed = {"skipped": f"{free / 1e9:.1f} GB free under {tmp}"}
    shutil.rmtree(tmp, ignore_errors=True)
except Exception as e:                                       # the storage figure is a side record
    bed = {"error": str(e)}

pairs = sum(r["pairs"] for r in records.values())
rec = {"config": f"configs[3]: N={N:,} x 22 chromosomes, {sum(SIZES.values()):,} SNPs, 2% missing, 1 cM, additive + dominance",
       "gpus": gpus, "gpu_threads": n_threads, "mode": mode, "packed_GB": total_bytes / 1e9, "host_mem_available_GB": avail / 1e9,
       "score_wall_s": score_wall, "score_wall_all_passes_s": walls if mode == "resident" else None,
       "pairs": pairs, "pairs_per_s": pairs / score_wall, "aggregate_h2d_gbs": total_bytes / score_wall / 1e9,
       "algorithmic_int8_tops": pairs * 12.0 * N / score_wall / 1e12,
       "per_gpu_busy_s": {str(k): v for k, v in busy.items()}, "generate_s": gen_s_total, "oracle_s": oracle_s,
       "total_wall_s": total_wall,
       "oracle": {"slices": len(checks), "all_counts_equal": all(c["counts_equal"] for c in checks),
                  "all_maf_bit_equal": all(c["maf_bit_equal"] for c in checks),
                  "max_l2_rel_err": max(c["l2_rel_err"] for c in checks), "max_l2d_err": max(c["l2d_err"] for c in checks),
                  "max_rstd_err": max(c["rstd_err"] for c in checks), "rows": checks},
       "bed_path_chr22": bed,
       "per_chromosome": [records[c] for c in CHROMS],
       "what_is_timed": ("resident: wall clock over ALL chromosomes, pinned host memory -> nldsc_ld_calculate_packed_on -> host arrays, "
                         "second pass (the first pays the device arenas); stream: the busiest GPU's sum of the same calls")}
os.makedirs("gpurun_out", exist_ok=True)
json.dump(rec, open(f"gpurun_out/r02_ukb_genome_{gpus}gpu.json", "w"), indent=1)
rec.pop("per_chromosome")
rec["oracle"].pop("rows")
print(json.dumps(rec))
