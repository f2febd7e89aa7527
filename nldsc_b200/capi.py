"""ctypes binding of the C ABI declared in include/nldsc_ld.h (libnldsc_ld.so).

There is no CPU fallback: importing works anywhere, but every compute call needs the
in-tree CUDA library and a B200; a missing library raises at load time, a missing GPU
raises RuntimeError from the call (the library reports NLDSC_ERR_RUNTIME).
Error mapping follows the reference's pybind11 module: NLDSC_ERR_INVALID -> ValueError
(std::invalid_argument, e.g. the PLINK magic check of stream.h:88-102), anything else ->
RuntimeError.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("NLDSC_LIB") or os.path.join(HERE, "libnldsc_ld.so")      # NLDSC_LIB: an A/B build of the same sources

EXPORTS = [
    "nldsc_ld_calculate", "nldsc_ld_calculate_packed", "nldsc_ld_session_create",
    "nldsc_ld_session_upload", "nldsc_ld_session_run", "nldsc_ld_session_download",
    "nldsc_ld_session_sync", "nldsc_ld_session_stats", "nldsc_ld_session_destroy",
    "nldsc_ld_debug_unpack", "nldsc_ld_debug_cross_counts", "nldsc_ld_device_count",
    "nldsc_ld_plan_windows", "nldsc_ld_plan_shards", "nldsc_ld_debug_band_probe", "nldsc_ld_debug_plan_windows_device",
    "nldsc_ld_calculate_on", "nldsc_ld_debug_i8_peak", "nldsc_ld_last_call_report", "nldsc_ld_arena_trim",
    "nldsc_ld_calculate_packed_on", "nldsc_ld_warmup", "nldsc_ld_debug_h2d_probe", "nldsc_ld_debug_fp4_probe",
    "nldsc_ld_debug_unpack_fp4", "nldsc_ld_report_device_times", "nldsc_ld_snp_stats_static",
    "nldsc_ld_debug_plan_xtiles",
]
STATIC_MAX_SNPS = 262144      # NLDSC_STATIC_MAX_SNPS of include/nldsc_ld.h


class Params(C.Structure):
    _fields_ = [("bedfile", C.c_char_p), ("n_snp", C.c_int32), ("n_org", C.c_int32),
                ("ld_wind", C.c_double), ("positions", C.POINTER(C.c_double)),
                ("maf", C.c_double), ("std_thr", C.c_double), ("rsq_thr", C.c_double)]


class Result(C.Structure):
    _fields_ = [("l2", C.POINTER(C.c_double)), ("l2d", C.POINTER(C.c_double)),
                ("maf", C.POINTER(C.c_double)), ("residuals_std", C.POINTER(C.c_double)),
                ("l2_ws", C.POINTER(C.c_int32)), ("l2d_ws", C.POINTER(C.c_int32)),
                ("l2d_wse", C.POINTER(C.c_int32))]


class Stats(C.Structure):
    _fields_ = [("pairs", C.c_int64), ("tiles", C.c_int64), ("chunks", C.c_int32),
                ("kernel_launches", C.c_int32), ("load_begin", C.c_int32), ("load_end", C.c_int32),
                ("npad", C.c_int32), ("band_impl", C.c_int32), ("ms_count", C.c_float),
                ("ms_stats", C.c_float), ("ms_unpack", C.c_float), ("ms_band", C.c_float),
                ("ms_finalize", C.c_float), ("ms_total", C.c_float), ("mfree_chunks", C.c_int32), ("fp4", C.c_int32),
                ("qtiles", C.c_int64), ("mfree_qtiles", C.c_int64), ("mfree_tiles_skipped", C.c_int64),
                ("xtiles", C.c_int64), ("xtiles_run", C.c_int64)]


class ShardReport(C.Structure):
    _fields_ = [("device", C.c_int32), ("row_begin", C.c_int32), ("row_end", C.c_int32), ("load_begin", C.c_int32),
                ("load_end", C.c_int32), ("chunks", C.c_int32), ("pairs", C.c_int64), ("bytes_h2d", C.c_int64),
                ("ms_open", C.c_float), ("ms_upload_start", C.c_float), ("ms_plan", C.c_float),
                ("ms_run_issue", C.c_float), ("ms_download", C.c_float), ("ms_free", C.c_float),
                ("ms_wall", C.c_float), ("ms_dev_upload", C.c_float), ("ms_dev_run", C.c_float),
                ("ms_dev_band", C.c_float), ("ms_dev_unpack", C.c_float)]


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -m nldsc_b200.build` "
                "(there is no CPU fallback for the LD-score path)")
        L = C.CDLL(LIB_PATH)
        for name in EXPORTS:
            getattr(L, name)
        L.nldsc_ld_session_destroy.restype = None
        _lib = L
    return _lib


def _raise(rc: int, err) -> None:
    if rc == 0:
        return
    msg = err.value.decode(errors="replace")
    if rc == 1:
        raise ValueError(msg)
    raise RuntimeError(msg)


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def _ip(a):
    return a.ctypes.data_as(C.POINTER(C.c_int32))


def make_params(n_snp, n_org, positions, ld_wind, maf, std_thr, rsq_thr, bedfile=None):
    pos = np.ascontiguousarray(positions, dtype=np.float64)
    if pos.shape != (n_snp,):
        raise ValueError("positions must hold n_snp coordinates")
    p = Params(os.fsencode(bedfile) if bedfile is not None else None, int(n_snp), int(n_org),
               float(ld_wind), _dp(pos), float(maf), float(std_thr), float(rsq_thr))
    p._keep = pos
    return p


def alloc_result(n):
    out = {"l2": np.empty(n, np.float64), "l2d": np.empty(n, np.float64),
           "maf": np.empty(n, np.float64), "residuals_std": np.empty(n, np.float64),
           "l2_ws": np.empty(n, np.int32), "l2d_ws": np.empty(n, np.int32),
           "l2d_wse": np.empty(n, np.int32)}
    r = Result(_dp(out["l2"]), _dp(out["l2d"]), _dp(out["maf"]), _dp(out["residuals_std"]),
               _ip(out["l2_ws"]), _ip(out["l2d_ws"]), _ip(out["l2d_wse"]))
    return out, r


def _host_ptr(packed) -> int:
    """Address of a C-contiguous uint8 host buffer (numpy array or torch CPU tensor)."""
    if hasattr(packed, "data_ptr"):
        return packed.data_ptr()
    a = np.ascontiguousarray(packed, dtype=np.uint8)
    _host_ptr.keep = a
    return a.ctypes.data


def device_count() -> int:
    n = C.c_int32(0)
    err = C.create_string_buffer(512)
    _raise(lib().nldsc_ld_device_count(C.byref(n), err, C.c_size_t(512)), err)
    return n.value


def plan_windows(positions, ld_wind):
    """Inclusive [lo, hi] window of every SNP (host-only)."""
    n = len(positions)
    p = make_params(n, 1, positions, ld_wind, 0.0, 0.0, 0.0)
    lo, hi = np.empty(n, np.int32), np.empty(n, np.int32)
    err = C.create_string_buffer(512)
    _raise(lib().nldsc_ld_plan_windows(C.byref(p), _ip(lo), _ip(hi), err, C.c_size_t(512)), err)
    return lo, hi


def plan_windows_device(positions, ld_wind):
    """plan_windows computed by the device kernel of the calculate path (parity hook)."""
    n = len(positions)
    p = make_params(n, 1, positions, ld_wind, 0.0, 0.0, 0.0)
    lo, hi = np.empty(n, np.int32), np.empty(n, np.int32)
    err = C.create_string_buffer(512)
    _raise(lib().nldsc_ld_debug_plan_windows_device(C.byref(p), _ip(lo), _ip(hi), err, C.c_size_t(512)), err)
    return lo, hi


def plan_shards(positions, ld_wind, parts):
    """Row cuts [parts + 1] of the contiguous shards used for `parts` GPUs (host-only)."""
    n = len(positions)
    p = make_params(n, 1, positions, ld_wind, 0.0, 0.0, 0.0)
    cuts = np.empty(parts + 1, np.int32)
    err = C.create_string_buffer(512)
    _raise(lib().nldsc_ld_plan_shards(C.byref(p), C.c_int32(parts), _ip(cuts), err, C.c_size_t(512)), err)
    return cuts


def calculate(bedfile, n_snp, n_org, positions, ld_wind, maf, std_thr, rsq_thr) -> dict:
    p = make_params(n_snp, n_org, positions, ld_wind, maf, std_thr, rsq_thr, bedfile)
    out, r = alloc_result(n_snp)
    err = C.create_string_buffer(512)
    _raise(lib().nldsc_ld_calculate(C.byref(p), C.byref(r), err, C.c_size_t(512)), err)
    return out


def calculate_on(bedfile, n_snp, n_org, positions, ld_wind, maf, std_thr, rsq_thr, first_device, n_devices=1) -> dict:
    """nldsc_ld_calculate on GPUs first_device .. first_device + n_devices - 1 (thread-safe placement)."""
    p = make_params(n_snp, n_org, positions, ld_wind, maf, std_thr, rsq_thr, bedfile)
    out, r = alloc_result(n_snp)
    err = C.create_string_buffer(512)
    _raise(lib().nldsc_ld_calculate_on(C.byref(p), C.byref(r), C.c_int32(first_device), C.c_int32(n_devices), err,
                                       C.c_size_t(512)), err)
    return out


def calculate_packed(packed, n_snp, n_org, positions, ld_wind, maf, std_thr, rsq_thr, out=None) -> dict:
    p = make_params(n_snp, n_org, positions, ld_wind, maf, std_thr, rsq_thr)
    if out is None:
        out, r = alloc_result(n_snp)
    else:
        out, r = out
    err = C.create_string_buffer(512)
    ptr = C.c_void_p(_host_ptr(packed))
    _raise(lib().nldsc_ld_calculate_packed(ptr, C.byref(p), C.byref(r), err, C.c_size_t(512)), err)
    return out


def calculate_packed_on(packed, n_snp, n_org, positions, ld_wind, maf, std_thr, rsq_thr, first_device, n_devices=1,
                        out=None) -> dict:
    """nldsc_ld_calculate_packed on GPUs first_device .. first_device + n_devices - 1 (thread-safe placement)."""
    p = make_params(n_snp, n_org, positions, ld_wind, maf, std_thr, rsq_thr)
    out, r = alloc_result(n_snp) if out is None else out
    err = C.create_string_buffer(512)
    ptr = C.c_void_p(_host_ptr(packed))
    _raise(lib().nldsc_ld_calculate_packed_on(ptr, C.byref(p), C.byref(r), C.c_int32(first_device), C.c_int32(n_devices),
                                              err, C.c_size_t(512)), err)
    return out


def warmup(first_device: int = 0, n_devices: int = 1) -> None:
    """Contexts, pinned staging, stream pools and kernel images of the devices, prepared in parallel."""
    err = C.create_string_buffer(512)
    _raise(lib().nldsc_ld_warmup(C.c_int32(first_device), C.c_int32(n_devices), err, C.c_size_t(512)), err)


class Session:
    """One shard on one GPU, device-resident between calls (nldsc_ld_session_*)."""

    def __init__(self, device, n_snp, n_org, positions, ld_wind, maf, std_thr, rsq_thr,
                 row_begin=0, row_end=None):
        self._h = C.c_void_p(None)
        self._p = make_params(n_snp, n_org, positions, ld_wind, maf, std_thr, rsq_thr)
        self.n_snp, self.n_org = int(n_snp), int(n_org)
        self.row_begin = int(row_begin)
        self.row_end = int(n_snp if row_end is None else row_end)
        err = C.create_string_buffer(512)
        _raise(lib().nldsc_ld_session_create(C.byref(self._h), C.c_int(device), C.byref(self._p),
                                             C.c_int32(self.row_begin), C.c_int32(self.row_end),
                                             err, C.c_size_t(512)), err)
        self.bps = (self.n_org + 3) // 4

    def stats(self) -> dict:
        st = Stats()
        if lib().nldsc_ld_session_stats(self._h, C.byref(st)) != 0:
            raise RuntimeError("nldsc_ld_session_stats failed")
        return {k: getattr(st, k) for k, _ in Stats._fields_}

    def upload(self, packed_all) -> None:
        """packed_all: the whole bfile's rows [n_snp, bps] on the host; only this shard's
        rows [load_begin, load_end) are copied (asynchronously, chunk by chunk)."""
        st = self.stats()
        base = _host_ptr(packed_all) + st["load_begin"] * self.bps
        err = C.create_string_buffer(512)
        _raise(lib().nldsc_ld_session_upload(self._h, C.c_void_p(base), None, err, C.c_size_t(512)), err)

    def run(self) -> None:
        err = C.create_string_buffer(512)
        _raise(lib().nldsc_ld_session_run(self._h, None, err, C.c_size_t(512)), err)

    def sync(self) -> None:
        err = C.create_string_buffer(512)
        _raise(lib().nldsc_ld_session_sync(self._h, err, C.c_size_t(512)), err)

    def download(self, out=None) -> dict:
        if out is None:
            out = alloc_result(self.row_end - self.row_begin)
        arrays, r = out
        err = C.create_string_buffer(512)
        _raise(lib().nldsc_ld_session_download(self._h, C.byref(r), None, err, C.c_size_t(512)), err)
        return arrays

    def close(self) -> None:
        if self._h:
            lib().nldsc_ld_session_destroy(self._h)
            self._h = C.c_void_p(None)

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def debug_unpack(packed, n_org):
    """-> planes int8 [4, n_snp, npad] (g, m, h, a), counts int32 [n_snp, 3] (het, hom2, missing)."""
    a = np.ascontiguousarray(packed, dtype=np.uint8)
    n_snp = a.shape[0]
    npad = C.c_int32(0)
    err = C.create_string_buffer(512)
    _raise(lib().nldsc_ld_debug_unpack(C.c_void_p(a.ctypes.data), C.c_int32(n_snp), C.c_int32(n_org),
                                       None, None, C.byref(npad), err, C.c_size_t(512)), err)
    planes = np.empty((4, n_snp, npad.value), np.int8)
    counts = np.empty((n_snp, 3), np.int32)
    _raise(lib().nldsc_ld_debug_unpack(C.c_void_p(a.ctypes.data), C.c_int32(n_snp), C.c_int32(n_org),
                                       C.c_void_p(planes.ctypes.data), C.c_void_p(counts.ctypes.data),
                                       C.byref(npad), err, C.c_size_t(512)), err)
    return planes, counts


def debug_cross_counts(packed, n_org, i_block, j_block) -> np.ndarray:
    """Raw band-kernel accumulators for one 64x64 SNP block pair -> int64 [2, 3, 64, 64]
    ordered (g, m) x (m, h, a), the layout of oracle.ld_oracle.cross_counts."""
    a = np.ascontiguousarray(packed, dtype=np.uint8)
    acc = np.empty((128, 192), np.int32)
    err = C.create_string_buffer(512)
    _raise(lib().nldsc_ld_debug_cross_counts(C.c_void_p(a.ctypes.data), C.c_int32(a.shape[0]),
                                             C.c_int32(n_org), C.c_int32(i_block), C.c_int32(j_block),
                                             C.c_void_p(acc.ctypes.data), err, C.c_size_t(512)), err)
    # row 32*(s/16) + 16*p + s%16, column 64*q + t
    r = acc.reshape(4, 2, 16, 3, 64)            # [s/16, p, s%16, q, t]
    return r.transpose(1, 3, 0, 2, 4).reshape(2, 3, 64, 64).astype(np.int64)


def debug_band_probe(mode: int, n_org: int, n_tiles: int, reps: int = 5) -> dict:
    """Band kernel on synthetic supertiles: mode 0 real, 1 MMA only, 2 TMA only.
    -> {ms, tops} where tops counts the int8 ops the MMAs of one launch perform (2 * 256 * 192 * npad per supertile)."""
    ms = C.c_float(0)
    err = C.create_string_buffer(512)
    _raise(lib().nldsc_ld_debug_band_probe(C.c_int32(mode), C.c_int32(n_org), C.c_int32(n_tiles), C.c_int32(reps),
                                           C.byref(ms), err, C.c_size_t(512)), err)
    npad = (n_org + 127) // 128 * 128
    ops = 2.0 * 256 * 192 * npad * n_tiles * (2 if mode >= 5 else 1)      # modes 5, 6: wide supertiles, two j-blocks each
    return {"ms": ms.value, "tops": ops / (ms.value * 1e-3) / 1e12, "gbs_operands": n_tiles * 2 * (40960 if mode >= 5 else 28672) * (npad // 128) / (ms.value * 1e-3) / 1e9}


def debug_i8_peak(n_cols: int = 192, per_commit: int = 8, iters: int = 20000, reps: int = 3, depth: int = 4) -> dict:
    """Back-to-back tcgen05 int8 MMAs (M=256, N=n_cols, K=32) from shared memory on every CTA pair: the tensor-pipe
    ceiling of the box for the band kernel's instruction shape.  -> {ms, tops}"""
    ms, tops = C.c_float(0), C.c_double(0)
    err = C.create_string_buffer(512)
    _raise(lib().nldsc_ld_debug_i8_peak(C.c_int32(n_cols), C.c_int32(per_commit), C.c_int32(depth), C.c_int32(iters), C.c_int32(reps),
                                        C.byref(ms), C.byref(tops), err, C.c_size_t(512)), err)
    return {"ms": ms.value, "tops": tops.value, "n_cols": n_cols, "per_commit": per_commit, "depth": depth, "iters": iters}


def last_call_report() -> list[dict]:
    """Per-shard record of the last calculate* call made on this thread (nldsc_ld_last_call_report)."""
    n = C.c_int32(0)
    lib().nldsc_ld_last_call_report(None, C.c_int32(0), C.byref(n))
    arr = (ShardReport * max(1, n.value))()
    lib().nldsc_ld_last_call_report(arr, C.c_int32(n.value), C.byref(n))
    return [{k: getattr(arr[i], k) for k, _ in ShardReport._fields_} for i in range(n.value)]


def arena_trim(device: int = -1) -> None:
    """Give the device memory cached between calls back to the driver (all devices when device < 0)."""
    lib().nldsc_ld_arena_trim(C.c_int32(device))


def debug_h2d_probe(nbytes: int, n_devices: int, write_combined: bool = False, first_device: int = 0, reps: int = 4) -> list[float]:
    """GB/s per device of n_devices simultaneous pinned-host -> HBM copies (default or write-combined pinned memory)."""
    out = (C.c_double * n_devices)()
    err = C.create_string_buffer(512)
    _raise(lib().nldsc_ld_debug_h2d_probe(C.c_int64(nbytes), C.c_int32(1 if write_combined else 0), C.c_int32(first_device),
                                          C.c_int32(n_devices), C.c_int32(reps), out, err, C.c_size_t(512)), err)
    return [float(x) for x in out]


def debug_fp4_probe(n_mma: int, n_cols: int = 256, per_commit: int = 8, all_pairs: bool = False) -> dict:
    """Accumulating fp4 (mxf4, scale 1.0) MMAs on known integer patterns -> accumulator columns 0..15 of the first CTA pair."""
    out = np.zeros((2, 128, 16), np.float32)
    ms = C.c_float(0)
    err = C.create_string_buffer(512)
    _raise(lib().nldsc_ld_debug_fp4_probe(C.c_int32(n_mma), C.c_int32(n_cols), C.c_int32(per_commit), C.c_int32(1 if all_pairs else 0),
                                          out.ctypes.data_as(C.POINTER(C.c_float)), C.byref(ms), err, C.c_size_t(512)), err)
    pairs = 74 if all_pairs else 1
    return {"acc": out, "ms": ms.value, "tops_equiv": pairs * n_mma * 2.0 * 256 * n_cols * 64 / (ms.value * 1e-3) / 1e12}


def debug_unpack_fp4(packed, n_org):
    """-> planes int8-valued [4, n_snp, 2 * npad_bytes] (g, m, h, a) decoded from the fp4 planes' nibbles, counts [n_snp, 3]."""
    a = np.ascontiguousarray(packed, dtype=np.uint8)
    n_snp = a.shape[0]
    npad = C.c_int32(0)
    err = C.create_string_buffer(512)
    _raise(lib().nldsc_ld_debug_unpack_fp4(C.c_void_p(a.ctypes.data), C.c_int32(n_snp), C.c_int32(n_org),
                                           None, None, C.byref(npad), err, C.c_size_t(512)), err)
    raw = np.empty((4, n_snp, npad.value), np.uint8)
    counts = np.empty((n_snp, 3), np.int32)
    _raise(lib().nldsc_ld_debug_unpack_fp4(C.c_void_p(a.ctypes.data), C.c_int32(n_snp), C.c_int32(n_org),
                                           C.c_void_p(raw.ctypes.data), C.c_void_p(counts.ctypes.data),
                                           C.byref(npad), err, C.c_size_t(512)), err)
    nib = np.stack([raw & 0xF, raw >> 4], -1).reshape(4, n_snp, 2 * npad.value)      # organism 2i = low nibble of byte i
    value = np.full(16, -1, np.int8)
    value[[0x0, 0x2, 0x4]] = [0, 1, 2]                                               # e2m1: 0.0, 1.0, 2.0
    return value[nib], counts


def report_device_times(on: bool = True) -> None:
    """Fill the ms_dev_* fields of last_call_report() from now on (costs 0.15-0.2 ms per call and shard)."""
    lib().nldsc_ld_report_device_times(C.c_int32(1 if on else 0))


def snp_stats_static(d_packed: int, n_snp: int, n_org: int, maf: float, std_thr: float, d_counts_out: int, d_maf_out: int,
                     d_rstd_out: int, device: int = 0, stream: int = 0) -> None:
    """nldsc_ld_snp_stats_static: per-SNP statistics of packed rows that are already on the device, written to device arrays
    the caller owns; allocates nothing (static scratch, at most STATIC_MAX_SNPS rows per call).  Every pointer argument is a
    DEVICE address (an int, e.g. tensor.data_ptr()); d_counts_out may be 0.  The work is queued on `stream`, not synchronised."""
    err = C.create_string_buffer(512)
    _raise(lib().nldsc_ld_snp_stats_static(C.c_void_p(d_packed), C.c_int32(n_snp), C.c_int32(n_org), C.c_double(maf),
                                           C.c_double(std_thr), C.c_void_p(d_counts_out or None), C.c_void_p(d_maf_out),
                                           C.c_void_p(d_rstd_out), C.c_int32(device), C.c_void_p(stream or None), err,
                                           C.c_size_t(512)), err)


def debug_plan_xtiles(jb_lo, jb_hi, raster: int = 8) -> np.ndarray:
    """Host-only: the 160-column supertiles of one chunk -> int32 [n_tiles, 8] = a_row[2], i0[2], j0, slot, n_grp, odd."""
    lo = np.ascontiguousarray(jb_lo, dtype=np.int32)
    hi = np.ascontiguousarray(jb_hi, dtype=np.int32)
    n = C.c_int32(0)
    err = C.create_string_buffer(512)
    _raise(lib().nldsc_ld_debug_plan_xtiles(_ip(lo), _ip(hi), C.c_int32(lo.size), C.c_int32(raster), None, C.c_int32(0),
                                            C.byref(n), err, C.c_size_t(512)), err)
    out = np.empty((n.value, 8), np.int32)
    _raise(lib().nldsc_ld_debug_plan_xtiles(_ip(lo), _ip(hi), C.c_int32(lo.size), C.c_int32(raster), _ip(out), C.c_int32(n.value),
                                            C.byref(n), err, C.c_size_t(512)), err)
    return out
