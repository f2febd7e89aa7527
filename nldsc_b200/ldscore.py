"""Host-side mirror of the reference's `ld` Python driver, on top of the B200 `_ldscore` module.

The reference's own Python layer (nldsc/ldscore/routine.py, common.py, nldsc/__main__.py) is the
caller of the hot path and stays unchanged when the `_ldscore` extension is swapped (see
INTEGRATION.md).  This module restates just enough of it, under the same names and with the same
argument meaning and error behaviour, that the path can be driven and tested end to end without
the reference tree: `python -m nldsc_b200 ld --bfile ... --ld-wind-cm ... -maf ... --std-thr ...`.

  estimate_lds                   ldscore/routine.py:51-102
  make_output (column layout)    ldscore/routine.py:32-48
  window / threshold validation  ldscore/common.py:10-36,146-182
  .bim / .fam parsing            ldscore/common.py:86-89,114-117,130-133

Beside the mirror (things the reference does not have):
  result_arrays                  zero-copy NumPy views instead of per-read list copies (ldscore.cpp:43-51)
  m_counts / write_m_files       the `.M` / `.M_5_50` side files h2's LDScoreReader looks for
                                 (h2/common.py:115-144), with its own fallback formula for MD
  estimate_lds_genome            all chromosome bfiles of a genome in one process, spread over the GPUs
"""
from __future__ import annotations

import os
import queue
import re
import threading

import numpy as np
import pandas as pd

BIM_COLUMNS = ("CHR", "SNP", "CM", "BP", "A1", "A2")


class NLDSCParameterError(ValueError):
    """Same role as the reference's core.common.NLDSCParameterError."""


def _window(ld_wind, metric: str) -> tuple[float, str]:
    w = float(ld_wind)
    if metric == "kbp":
        w, metric = w * 1000.0, "bp"
    if metric not in ("bp", "cm"):
        raise NLDSCParameterError("Invalid metric")
    if w <= 0:
        raise NLDSCParameterError("The ld-window must be greater than 0")
    if metric == "bp" and w > 5e6:
        raise NLDSCParameterError("The ld-window cannot be larger than 5 Mbp")
    if metric == "cm" and w > 100:
        raise NLDSCParameterError("The ld-window cannot be larger than 100 cm")
    return w, metric


def _unit_interval(value, hi: float, what: str) -> float:
    v = float(value)
    if not (0 <= v < hi):
        raise NLDSCParameterError(f"{what} must be between 0 and {hi:g}!")
    return v


def _bfile_paths(bfile: str) -> tuple[str, str, str]:
    path = os.path.abspath(bfile)
    root, ext = os.path.splitext(path)
    if ext in (".bed", ".bim", ".fam"):
        path = root
    paths = tuple(path + e for e in (".bed", ".bim", ".fam"))
    for p in paths:
        if not os.path.exists(p):
            raise FileNotFoundError(f'No such file: "{p}"')
    return paths


def _module():
    try:
        from . import _ldscore
    except ImportError as e:          # same contract as the reference's routines.py:5-10 stub
        raise RuntimeError("nldsc_b200._ldscore is not built: run `python -m nldsc_b200.build` "
                           "(there is no CPU fallback)") from e
    return _ldscore


_FIELDS = ("l2", "l2d", "maf", "residuals_std", "l2_ws", "l2d_ws", "l2d_wse")


def result_arrays(ld) -> dict:
    """The seven result vectors as NumPy arrays.  With this repo's `_ldscore` they are views into the
    result object (`as_arrays`); with the reference's module each list attribute is converted once."""
    lds = _module()
    if hasattr(lds, "as_arrays") and isinstance(ld, lds.LDScoreResult):
        return lds.as_arrays(ld)
    return {k: np.asarray(getattr(ld, k), dtype=np.float64 if k in _FIELDS[:4] else np.int64) for k in _FIELDS}


def make_output(bim: pd.DataFrame, ld, *, extra: bool = False) -> pd.DataFrame:
    """CHR SNP BP L2 L2D [MAF WSA WSD WSDE RSTD] -- the layout h2's LDScoreReader reads by name."""
    a = ld if isinstance(ld, dict) else result_arrays(ld)
    out = pd.DataFrame({"CHR": bim["CHR"], "SNP": bim["SNP"], "BP": bim["BP"],
                        "L2": np.asarray(a["l2"], dtype=np.float64), "L2D": np.asarray(a["l2d"], dtype=np.float64)})
    if extra:
        out["MAF"] = np.asarray(a["maf"], dtype=np.float64)
        out["WSA"] = np.asarray(a["l2_ws"], dtype=np.int64)
        out["WSD"] = np.asarray(a["l2d_ws"], dtype=np.int64)
        out["WSDE"] = np.asarray(a["l2d_wse"], dtype=np.int64)
        out["RSTD"] = np.asarray(a["residuals_std"], dtype=np.float64)
    return out


def m_counts(arrays: dict, snp_ids=None, maf_min: float | None = None, first=None) -> tuple[int, int]:
    """(M, MD) as h2 computes them when no side file exists (h2/common.py:126-130): M = rows of the .L2
    file that survive `dropna().drop_duplicates('SNP')`, MD = int(M * mean(WSDE / WSA)) over those rows.
    `maf_min` = 0.05 gives the `.M_5_50` flavour (common SNPs only).  Rows with WSA = 0 (a retained SNP
    with an empty window) are left out of the mean, where the reference would divide by zero."""
    l2 = np.asarray(arrays["l2"], dtype=np.float64)
    keep = ~np.isnan(l2) & ~np.isnan(np.asarray(arrays["l2d"], dtype=np.float64))
    if first is None and snp_ids is not None:
        first = ~pd.Series(np.asarray(snp_ids)).duplicated().to_numpy()
    if first is not None:                       # rows that survive drop_duplicates('SNP')
        keep &= np.asarray(first, dtype=bool)
    if maf_min is not None:
        keep &= np.asarray(arrays["maf"], dtype=np.float64) > maf_min
    m = int(keep.sum())
    wsa = np.asarray(arrays["l2_ws"], dtype=np.float64)[keep]
    wsde = np.asarray(arrays["l2d_wse"], dtype=np.float64)[keep]
    ok = wsa > 0
    md = int(m * float(np.mean(wsde[ok] / wsa[ok]))) if ok.any() else 0
    return m, md


def write_m_files(l2_path: str, arrays: dict, snp_ids=None, first=None) -> dict:
    """`<stem>.M` and `<stem>.M_5_50` next to an .L2 file, in the tab-separated `M MD` layout that
    LDScoreReader._read_m parses (h2/common.py:141-144)."""
    stem = os.path.splitext(l2_path)[0]
    out = {}
    for suffix, maf_min in ((".M", None), (".M_5_50", 0.05)):
        m, md = m_counts(arrays, snp_ids, maf_min, first)
        with open(stem + suffix, "w") as f:
            f.write(f"M\tMD\n{m}\t{md}\n")
        out[suffix] = (m, md)
    return out


def estimate_lds(bfile: str, ld_wind: float, wind_metric: str, maf_thr: float = 1e-5, std_thr: float = 1e-5,
                 rsq_thr: float | None = None, *, out: str | None = None, extra: bool = False,
                 summary: bool = False, verbose: int = 0, device: int | None = None, gpus: int = 1,
                 m_files: bool = False):
    lds = _module()
    bed, bim_path, fam_path = _bfile_paths(bfile)
    bim = pd.read_csv(bim_path, sep="\t", names=BIM_COLUMNS)
    if len(bim["CHR"].unique()) != 1:
        raise NLDSCParameterError("The current version of the program can only work with one chromosome in one file.")
    with open(fam_path, "rb") as f:          # number of lines, without iterating 300,000 of them under the GIL
        fam = f.read()
    n_org = fam.count(b"\n") + (1 if fam and not fam.endswith(b"\n") else 0)
    wind, metric = _window(ld_wind, wind_metric)
    maf = _unit_interval(maf_thr, 1.0, "Minor allele frequency")
    std = _unit_interval(std_thr, 1.0, "standard deviation threshold")
    rsq = _unit_interval(1.0 / len(bim) if rsq_thr is None else rsq_thr, 0.1, "r-squared threshold")
    positions = bim["CM" if metric == "cm" else "BP"].to_numpy(dtype=np.float64)
    params = lds.LDScoreParams(bfile=bed, n_snp=len(bim), n_org=n_org, ld_wind=wind, maf=maf, std_thr=std,
                               rsq_thr=rsq, positions=positions)
    ld = lds.calculate(params) if device is None else lds.calculate_on(params, device, gpus)
    arrays = result_arrays(ld)
    if summary:
        d = pd.DataFrame({"L2": arrays["l2"], "L2D": arrays["l2d"], "MAF": arrays["maf"]})
        print("L2/L2D/MAF Correlation matrix\n" + str(d.corr()))
        print(f"- Number of additive non-null LD: {d['L2'].count()}\n"
              f"- Number of non-additive non-null LD: {d['L2D'].count()}")
    if out and hasattr(lds, "write_l2") and bim["BP"].dtype.kind in "iu":
        # same bytes as the pandas writer below (tests/test_host.py), without its seconds at genome scale
        lds.write_l2(out, bim["CHR"].astype(str).tolist(), bim["SNP"].astype(str).tolist(), bim["BP"].to_numpy(np.int64),
                     arrays["l2"], arrays["l2d"], arrays["maf"], arrays["residuals_std"], arrays["l2_ws"],
                     arrays["l2d_ws"], arrays["l2d_wse"], extra)
        if m_files:
            write_m_files(out, arrays, bim["SNP"])
        return None
    out_df = make_output(bim, arrays, extra=extra)
    if out:
        out_df.to_csv(out, sep="\t", index=False, float_format="%.5f")
        if m_files:
            write_m_files(out, arrays, bim["SNP"])
        return None
    return out_df


def _genome_bfiles(bfile: str) -> list[tuple[str, str]]:
    """`prefix@suffix` (ldsc convention: @ = chromosome number) -> [(chromosome label, bfile)] for the
    chromosomes whose .bed exists; a plain bfile gives one entry labelled by its file name."""
    if "@" not in bfile:
        return [(os.path.basename(bfile), bfile)]
    found = []
    for c in [str(i) for i in range(1, 23)] + ["X"]:
        cand = bfile.replace("@", c)
        if os.path.exists(cand + ".bed"):
            found.append((c, cand))
    if not found:
        raise FileNotFoundError(f'No chromosome bfile matches "{bfile}"')
    return found


def _parse_bfile_meta(path: str) -> dict:
    """Everything the scoring call needs from the .bim / .fam of one chromosome (runs in a pool thread beside the
    others).  With this repo's `_ldscore` the .bim is parsed natively (`read_bim`: GIL released, the CHR / SNP strings
    stay on the C++ side for the .L2 writer); otherwise by pandas, as the reference does (ldscore/common.py:86-89)."""
    import time
    t0 = time.perf_counter()
    lds = _module()
    bed, bim_path, fam_path = _bfile_paths(path)
    with open(fam_path, "rb") as f:
        fam = f.read()
    n_org = fam.count(b"\n") + (1 if fam and not fam.endswith(b"\n") else 0)
    m = {"bed": bed, "n_org": n_org, "bed_bytes": os.path.getsize(bed)}
    if hasattr(lds, "read_bim"):
        tbl = lds.read_bim(bim_path)
        if len(tbl.chromosomes()) != 1:
            raise NLDSCParameterError("The current version of the program can only work with one chromosome in one file.")
        m.update(table=tbl if tbl.bp_integral else None, n_snp=len(tbl), cm=tbl.cm(), bp_f=tbl.bp_float(),
                 first=tbl.first_occurrence())
        if not tbl.bp_integral:
            m.update(chr=[tbl.chromosomes()[0]] * len(tbl), snp=tbl.snps(), bp=None)
    else:
        bim = pd.read_csv(bim_path, sep="\t", names=BIM_COLUMNS)
        if len(bim["CHR"].unique()) != 1:
            raise NLDSCParameterError("The current version of the program can only work with one chromosome in one file.")
        m.update(table=None, n_snp=len(bim), chr=bim["CHR"].astype(str).tolist(), snp=bim["SNP"].astype(str).tolist(),
                 bp=bim["BP"].to_numpy(np.int64) if bim["BP"].dtype.kind in "iu" else None,
                 cm=bim["CM"].to_numpy(dtype=np.float64), bp_f=bim["BP"].to_numpy(dtype=np.float64),
                 first=~bim["SNP"].duplicated().to_numpy())
    m["seconds"] = time.perf_counter() - t0
    return m


def estimate_lds_genome(bfile: str, ld_wind: float, wind_metric: str, maf_thr: float = 1e-5, std_thr: float = 1e-5,
                        rsq_thr: float | None = None, *, out_dir: str, extra: bool = True, gpus: int | None = None,
                        first_device: int = 0, parse_workers: int | None = None) -> pd.DataFrame:
    """Score every chromosome bfile of a genome in ONE process.

    The reference scores one chromosome per invocation (ldscore/common.py:114-117) and never writes the
    `.M` files `h2` looks for.  Here
      1. every .bim / .fam is parsed by a thread pool (pandas' C parser releases the GIL), largest chromosome first,
         while the GPUs are warmed up in parallel (`_ldscore.warmup`: contexts, pinned staging, kernel images);
      2. the chromosomes are dealt, largest .bed first, to one host thread per GPU (`_ldscore.calculate_on`, GIL
         released, no collective: chromosomes are independent); NLDSC_READERS is set to cores / GPUs so that the
         .bed reader threads of all GPUs together do not oversubscribe the host;
      3. every result goes to a writer pool: `<out_dir>/<chr>.L2` (native writer, GIL released) + `.M`, `.M_5_50`,
         the folder layout LDScoreReader._read_folder globs (h2/common.py:98-112).
    Returns one row per chromosome: CHR, file, SNPs, M, MD, seconds (scoring call), device; `.attrs["phases"]` holds
    the wall time of every phase (parse, warmup, score, write) and `.attrs["per_gpu_busy_s"]` the scoring seconds per GPU."""
    import time
    from concurrent.futures import ThreadPoolExecutor
    lds = _module()
    t_start = time.perf_counter()
    jobs = _genome_bfiles(bfile)
    os.makedirs(out_dir, exist_ok=True)
    if gpus is None:
        gpus = int(os.environ.get("NLDSC_GPUS", "1")) if os.environ.get("NLDSC_GPUS", "1") != "all" else 64
    n_dev = max(1, min(int(gpus), len(jobs), max(1, lds.device_count() - first_device)))
    wind, metric = _window(ld_wind, wind_metric)
    maf = _unit_interval(maf_thr, 1.0, "Minor allele frequency")
    std = _unit_interval(std_thr, 1.0, "standard deviation threshold")
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    own_readers = "NLDSC_READERS" not in os.environ and n_dev > 1
    if own_readers:
        os.environ["NLDSC_READERS"] = str(max(1, min(8, cores // n_dev)))

    # ---- 1. parse (processes) beside the device warm-up; the GPU threads below start on a chromosome as soon as ITS
    # parse is done, so only the first parse is on the critical path
    phases = {}
    t0 = time.perf_counter()
    warm = threading.Thread(target=lambda: lds.warmup(first_device, n_dev) if hasattr(lds, "warmup") else None)
    warm.start()
    order = sorted(jobs, key=lambda j: -os.path.getsize(j[1] + ".bed"))          # longest first
    # threads, not processes: pandas' C parser releases the GIL, and a process pool costs more to start (fork of a
    # process that holds CUDA contexts) and to return 100,000-string lists from (pickle) than the parsing itself --
    # measured on configs[2]: the last parse landed at 0.15-0.29 s of a 0.55 s job (profiles/r02h_cfg3_genome_*.json)
    n_parse = max(1, min(len(jobs), parse_workers if parse_workers is not None else max(1, min(8, cores - n_dev))))
    pool = ThreadPoolExecutor(max_workers=n_parse)
    meta_f = {label: pool.submit(_parse_bfile_meta, path) for label, path in order}
    parse_done = [0.0]

    def _note_parse(_f):
        parse_done[0] = max(parse_done[0], time.perf_counter() - t0)
    for f in meta_f.values():
        f.add_done_callback(_note_parse)
    warm.join()
    phases["warmup_s"] = time.perf_counter() - t0
    todo: queue.Queue = queue.Queue()
    for j in order:
        todo.put(j)
    rows, errors, lock = [], [], threading.Lock()
    busy = {first_device + d: 0.0 for d in range(n_dev)}
    writers = ThreadPoolExecutor(max_workers=max(2, min(8, cores // 2)))
    pending = []

    def write(label, m, arrays, seconds, device):
        t = time.perf_counter()
        l2_path = os.path.join(out_dir, f"{label}.L2")
        if m.get("table") is not None:
            lds.write_l2_bim(l2_path, m["table"], arrays["l2"], arrays["l2d"], arrays["maf"], arrays["residuals_std"],
                             arrays["l2_ws"], arrays["l2d_ws"], arrays["l2d_wse"], extra)
        elif hasattr(lds, "write_l2") and m["bp"] is not None:
            lds.write_l2(l2_path, m["chr"], m["snp"], m["bp"], arrays["l2"], arrays["l2d"], arrays["maf"],
                         arrays["residuals_std"], arrays["l2_ws"], arrays["l2d_ws"], arrays["l2d_wse"], extra)
        else:
            bim = pd.DataFrame({"CHR": m["chr"], "SNP": m["snp"], "BP": m["bp_f"]})
            make_output(bim, arrays, extra=extra).to_csv(l2_path, sep="\t", index=False, float_format="%.5f")
        mm = write_m_files(l2_path, arrays, first=m["first"])
        with lock:
            rows.append({"CHR": label, "file": l2_path, "SNPs": m["n_snp"], "M": mm[".M"][0], "MD": mm[".M"][1],
                         "seconds": seconds, "device": device, "write_seconds": time.perf_counter() - t})

    def worker(device: int) -> None:
        while True:
            try:
                label, path = todo.get_nowait()
            except queue.Empty:
                return
            try:
                m = meta_f[label].result()
                rsq = _unit_interval(1.0 / m["n_snp"] if rsq_thr is None else rsq_thr, 0.1, "r-squared threshold")
                params = lds.LDScoreParams(bfile=m["bed"], n_snp=m["n_snp"], n_org=m["n_org"], ld_wind=wind, maf=maf,
                                           std_thr=std, rsq_thr=rsq, positions=m["cm"] if metric == "cm" else m["bp_f"])
                t0 = time.perf_counter()
                ld = lds.calculate_on(params, device, 1)
                dt = time.perf_counter() - t0
                with lock:
                    busy[device] += dt
                pending.append(writers.submit(write, label, m, result_arrays(ld), dt, device))
            except Exception as e:          # one bad chromosome must not hide behind the others
                with lock:
                    errors.append((label, e))

    t0 = time.perf_counter()
    threads = [threading.Thread(target=worker, args=(first_device + d,)) for d in range(n_dev)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    phases["score_s"] = time.perf_counter() - t0
    if own_readers:
        os.environ.pop("NLDSC_READERS", None)
    phases["parse_s"] = parse_done[0]                # when the last parse finished, from the start (overlaps the scoring)
    pool.shutdown()
    t0 = time.perf_counter()
    for f in pending:
        try:
            f.result()
        except Exception as e:
            errors.append(("write", e))
    writers.shutdown()
    phases["write_tail_s"] = time.perf_counter() - t0
    phases["total_s"] = time.perf_counter() - t_start
    if errors:
        label, e = errors[0]
        raise RuntimeError(f"chromosome {label}: {e}") from e
    pos = {label: i for i, (label, _) in enumerate(jobs)}
    df = pd.DataFrame(sorted(rows, key=lambda r: pos[r["CHR"]]))
    df.attrs["phases"] = phases
    df.attrs["per_gpu_busy_s"] = busy
    if os.environ.get("NLDSC_TRACE"):
        import sys
        print(f"[nldsc_ld] genome: {phases}  busy per GPU: {busy}", file=sys.stderr)
    return df
