"""CPU: the C-ABI library loads and exports what include/nldsc_ld.h declares, the host-side
planning matches the reference's window semantics, and the error behaviour of the boundary
(no compute happens here: there is no GPU and no CPU fallback)."""
import importlib
import os
import re
import struct

import numpy as np
import pytest

from nldsc_b200 import build, capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module", autouse=True)
def built():
    build.build_module()


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "nldsc_ld.h")).read()
    declared = set(re.findall(r"\b(nldsc_ld_[a-z0-9_]+)\s*\(", hdr))
    declared -= {"nldsc_ld_params", "nldsc_ld_result", "nldsc_ld_stats", "nldsc_ld_session"}
    assert len(declared) >= 12
    L = capi.lib()
    for name in sorted(declared):
        assert hasattr(L, name), name
    assert declared == set(capi.EXPORTS)


def test_windows_match_reference_semantics():
    rng = np.random.default_rng(0)
    pos = np.sort(rng.uniform(0, 5, 3000))
    pos[100:110] = pos[100]                       # ties
    for wind in (0.0, 0.01, 0.3, 10.0):
        lo, hi = capi.plan_windows(pos, wind)
        for i in rng.integers(0, 3000, 200):
            inw = np.nonzero(np.abs(pos - pos[i]) <= wind)[0]     # tools.h:43-44
            assert lo[i] == inw[0] and hi[i] == inw[-1]


def test_shard_cuts_cover_rows_once():
    pos = np.cumsum(np.random.default_rng(1).exponential(1 / 340, 20000))
    for parts in (1, 2, 3, 8):
        cuts = capi.plan_shards(pos, 1.0, parts)
        assert cuts[0] == 0 and cuts[-1] == 20000 and np.all(np.diff(cuts) >= 0)
        assert np.all(cuts[1:-1] % 64 == 0)
        lo, hi = capi.plan_windows(pos, 1.0)
        area = np.add.reduceat((hi - lo + 1).astype(np.int64), cuts[:-1][np.diff(cuts) > 0])
        assert area.max() / area.mean() < 1.15


@pytest.mark.parametrize("seed", range(6))
def test_x160_supertiles_score_every_column_group_once(seed):
    """The 160-column supertiles of a chunk (plan_xtiles_chunk, host-only): every 32-SNP column group of every i-pair's
    j-range is scored by exactly one supertile, into the (slot, half) the 64-block bodies and finalize use for it, and the
    groups a supertile carries beyond the range (n_grp < 5) are never scored."""
    rng = np.random.default_rng(300 + seed)
    n_ib = int(rng.integers(1, 40))
    width = rng.integers(1, 30, size=n_ib)
    lo = np.sort(rng.integers(0, 50, size=n_ib)).astype(np.int32)
    hi = np.maximum.accumulate(lo + width).astype(np.int32)            # non-decreasing, >= lo
    raster = int(rng.choice([1, 2, 8]))
    t = capi.debug_plan_xtiles(lo, hi, raster)
    slot_next = 0
    for p, k in enumerate(range(0, n_ib, 2)):
        lo_k = int(lo[k])
        hi_k = int(hi[k + 1] if k + 1 < n_ib else hi[k])
        mine = t[t[:, 2] == 64 * k]                                     # i0[0] identifies the i-pair
        assert np.all(mine[:, 3] == (64 * (k + 1) if k + 1 < n_ib else 0x3FFFFFC0))
        seen = {}
        for a_row0, a_row1, i0, i1, j0, slot, n_grp, odd in mine.tolist():
            assert j0 % 32 == 0 and 1 <= n_grp <= 5 and odd == (j0 // 32) & 1 and a_row0 == i0
            for g in range(n_grp):
                g32 = j0 // 32 + g                                     # absolute 32-SNP group
                assert g32 not in seen, "a column group scored twice"
                seen[g32] = (slot + ((odd + g) >> 1), (odd + g) & 1)   # where band_x160_body writes its partial sum
        want = {g32: (slot_next + (g32 // 2 - lo_k), g32 & 1) for g32 in range(2 * lo_k, 2 * (hi_k + 1))}
        assert seen == want
        slot_next += hi_k - lo_k + 1
    assert len(t) == sum((2 * (int(hi[k + 1] if k + 1 < n_ib else hi[k]) - int(lo[k]) + 1) + 4) // 5 for k in range(0, n_ib, 2))
    with pytest.raises(ValueError):
        capi.debug_plan_xtiles([3, 2], [4, 5])


def test_invalid_positions_rejected():
    with pytest.raises(ValueError):
        capi.plan_windows(np.array([0.0, 2.0, 1.0]), 1.0)
    with pytest.raises(ValueError):
        capi.plan_windows(np.array([-1.0, 2.0]), 1.0)


def test_bad_magic_is_value_error(tmp_path):
    """reference stream.h:88-102 throws std::invalid_argument -> ValueError."""
    p = tmp_path / "bad.bed"
    p.write_bytes(bytes([0x6C, 0x1B, 0x00]) + bytes(10))
    with pytest.raises(ValueError, match="magic"):
        capi.calculate(str(p), 5, 8, np.arange(5.0), 1.0, 0.01, 1e-5, 0.1)


def test_truncated_bed_is_value_error(tmp_path):
    p = tmp_path / "short.bed"
    p.write_bytes(bytes([0x6C, 0x1B, 0x01]) + bytes(9))
    with pytest.raises(ValueError, match="size"):
        capi.calculate(str(p), 5, 8, np.arange(5.0), 1.0, 0.01, 1e-5, 0.1)


def test_missing_file_is_runtime_error():
    with pytest.raises(RuntimeError):
        capi.calculate("/nonexistent/x.bed", 5, 8, np.arange(5.0), 1.0, 0.01, 1e-5, 0.1)


def test_ldscore_module_surface():
    """Same names / kwargs / attributes as reference ldscore.cpp:17-54."""
    m = importlib.import_module("nldsc_b200._ldscore")
    p = m.LDScoreParams("/x.bed", n_snp=2, n_org=4, ld_wind=1.0, maf=0.1, std_thr=1e-5, rsq_thr=0.1,
                        positions=[0.0, 1.0])
    assert (p.bedfile, p.n_snp, p.n_org, p.ld_wind, p.maf, p.std_thr, p.rsq_thr, p.positions) == \
        ("/x.bed", 2, 4, 1.0, 0.1, 1e-5, 0.1, [0.0, 1.0])
    with pytest.raises(TypeError):
        m.LDScoreParams("/x.bed", 2, 4, 1.0, 0.1, 1e-5, 0.1, [0.0, 1.0])     # kw_only, ldscore.cpp:23
    r = m.LDScoreResult()
    for f in ("l2", "l2d", "maf", "residuals_std", "l2_ws", "l2d_ws", "l2d_wse"):
        assert getattr(r, f) == []
    q = m.LDScoreParams()
    q.positions = [1.0]
    assert q.positions == [1.0]
    with pytest.raises(RuntimeError):
        m.calculate(p)                                                       # no such file


def test_result_arrays_are_views_and_m_files_follow_h2(tmp_path):
    """NumPy views of the result vectors; `.M` files in LDScoreReader's layout with its
    own fallback formula (reference h2/common.py:126-130,141-144)."""
    import pandas as pd
    from nldsc_b200 import ldscore
    m = importlib.import_module("nldsc_b200._ldscore")
    r = m.LDScoreResult()
    r.l2 = [1.5, float("nan"), 2.5, 3.0]
    r.l2d = [0.1, float("nan"), 0.2, 0.3]
    r.maf = [0.2, 0.0, 0.01, 0.4]
    r.residuals_std = [0.5, float("nan"), 0.5, 0.5]
    r.l2_ws, r.l2d_ws, r.l2d_wse = [4, -1, 2, 0], [4, -1, 2, 0], [2, -1, 2, 0]
    a = m.as_arrays(r)
    assert a["l2"].dtype == np.float64 and a["l2_ws"].dtype == np.int32
    assert not a["l2"].flags.owndata and a["l2"].base is not None
    np.testing.assert_array_equal(a["l2_ws"], [4, -1, 2, 0])
    # h2's fallback on the same rows: dropna, M = len, MD = int(M * mean(WSDE / WSA)); the WSA = 0 row is left out of the mean
    assert ldscore.m_counts(a) == (3, int(3 * np.mean([2 / 4, 2 / 2])))
    assert ldscore.m_counts(a, maf_min=0.05) == (2, int(2 * (2 / 4)))
    assert ldscore.m_counts(a, snp_ids=["a", "b", "a", "c"]) == (2, int(2 * (2 / 4)))
    out = ldscore.write_m_files(str(tmp_path / "22.L2"), a)
    got = pd.read_csv(tmp_path / "22.M", sep="\t")          # LDScoreReader._read_m
    assert (int(got["M"].iloc[0]), int(got["MD"].iloc[0])) == out[".M"] == (3, 2)
    assert (tmp_path / "22.M_5_50").read_text() == "M\tMD\n2\t1\n"


@pytest.mark.parametrize("extra", [False, True])
def test_l2_writer_matches_pandas_bytes(tmp_path, extra):
    """`_ldscore.write_l2` = the reference's to_csv(sep='\\t', index=False, float_format='%.5f') (routine.py:99), byte for byte."""
    import pandas as pd
    from nldsc_b200 import ldscore
    m = importlib.import_module("nldsc_b200._ldscore")
    rng = np.random.default_rng(5)
    n = 2000
    a = {"l2": rng.normal(50, 30, n), "l2d": rng.normal(0, 1e-3, n), "maf": rng.uniform(0, 0.5, n).astype(np.float32).astype(np.float64),
         "residuals_std": rng.uniform(0, 1, n), "l2_ws": rng.integers(0, 2000, n).astype(np.int32),
         "l2d_ws": rng.integers(0, 2000, n).astype(np.int32), "l2d_wse": rng.integers(0, 2000, n).astype(np.int32)}
    drop = rng.random(n) < 0.1
    for k in ("l2", "l2d", "residuals_std"):
        a[k][drop] = np.nan
    for k in ("l2_ws", "l2d_ws", "l2d_wse"):
        a[k][drop] = -1
    a["l2d"][:4] = [0.0, -0.0, -4e-6, 0.000005]          # rounding and signed-zero cases
    a["l2"][4:7] = [1e6 + 0.123456, 0.999995, 123.4]
    # the writer formats %.5f itself and falls back to printf near rounding boundaries and for huge values: all of those
    a["l2"][10:20] = [0.5e-5, 1.5e-5, 2.5e-5, -2.5e-5, 1e12, -123456789.123455, 1e16, np.inf, -np.inf, 99999.999995]
    a["l2d"][10:400] = (rng.integers(-10 ** 7, 10 ** 7, 390) + 0.5) * 1e-5      # as close to a tie as doubles get
    bim = pd.DataFrame({"CHR": 22, "SNP": [f"rs{i}" for i in range(n)], "BP": np.arange(n, dtype=np.int64) * 997 + 1})
    ldscore.make_output(bim, a, extra=extra).to_csv(tmp_path / "pandas.L2", sep="\t", index=False, float_format="%.5f")
    m.write_l2(str(tmp_path / "native.L2"), bim["CHR"].astype(str).tolist(), bim["SNP"].tolist(), bim["BP"].to_numpy(),
               a["l2"], a["l2d"], a["maf"], a["residuals_std"], a["l2_ws"], a["l2d_ws"], a["l2d_wse"], extra)
    assert (tmp_path / "native.L2").read_bytes() == (tmp_path / "pandas.L2").read_bytes()


def test_native_bim_table(tmp_path):
    """`_ldscore.read_bim` (the .bim parsed and kept on the C++ side for the genome driver) against pandas, the way the
    reference reads it (ldscore/common.py:86-89), and `write_l2_bim` against the list-based writer."""
    import pandas as pd
    m = importlib.import_module("nldsc_b200._ldscore")
    rng = np.random.default_rng(8)
    n = 5000
    cm = np.round(np.cumsum(rng.exponential(1 / 340, n)), 8)
    ids = [f"rs{i}" for i in range(n)]
    ids[77] = ids[3]                                          # a duplicated SNP id
    with open(tmp_path / "t.bim", "w") as f:
        for i in range(n):
            sep = "\t" if i % 3 else " "                       # PLINK allows either
            f.write(sep.join(["7", ids[i], f"{cm[i]:.8f}", str(1000 + 37 * i), "A", "C"]) + ("\r\n" if i == 5 else "\n"))
    t = m.read_bim(str(tmp_path / "t.bim"))
    assert len(t) == n and t.chromosomes() == ["7"] and t.bp_integral and t.snp(77) == ids[3]
    assert np.array_equal(t.cm(), np.array([float(f"{v:.8f}") for v in cm]))
    assert np.array_equal(t.bp(), 1000 + 37 * np.arange(n)) and np.array_equal(t.bp_float(), t.bp().astype(float))
    assert np.array_equal(t.first_occurrence(), ~pd.Series(ids).duplicated().to_numpy())
    a = {k: rng.normal(size=n) for k in ("l2", "l2d", "maf", "residuals_std")}
    w = {k: rng.integers(-1, 900, n).astype(np.int32) for k in ("l2_ws", "l2d_ws", "l2d_wse")}
    for extra in (False, True):
        m.write_l2_bim(str(tmp_path / "a.L2"), t, a["l2"], a["l2d"], a["maf"], a["residuals_std"], w["l2_ws"], w["l2d_ws"], w["l2d_wse"], extra)
        m.write_l2(str(tmp_path / "b.L2"), ["7"] * n, ids, t.bp(), a["l2"], a["l2d"], a["maf"], a["residuals_std"], w["l2_ws"],
                   w["l2d_ws"], w["l2d_wse"], extra)
        assert (tmp_path / "a.L2").read_bytes() == (tmp_path / "b.L2").read_bytes()
    (tmp_path / "two.bim").write_text("1\trs1\t0.1\t5\tA\tC\n2\trs2\t0.2\t9\tA\tC\n")
    assert m.read_bim(str(tmp_path / "two.bim")).chromosomes() == ["1", "2"]
    (tmp_path / "bad.bim").write_text("1\trs1\n")
    with pytest.raises(RuntimeError):
        m.read_bim(str(tmp_path / "bad.bim"))


def test_genome_bfile_pattern(tmp_path):
    from nldsc_b200 import ldscore
    for c in (2, 11, 21):
        (tmp_path / f"g.{c}.bed").write_bytes(b"x" * c)
    got = ldscore._genome_bfiles(str(tmp_path / "g.@"))
    assert [c for c, _ in got] == ["2", "11", "21"]
    with pytest.raises(FileNotFoundError):
        ldscore._genome_bfiles(str(tmp_path / "h.@"))
    with pytest.raises(ValueError):
        capi.calculate_on("/x.bed", 2, 4, np.arange(2.0), 1.0, 0.1, 1e-5, 0.1, first_device=-1)


@pytest.mark.skipif(os.path.exists("/dev/nvidiactl"), reason="GPU box")
def test_no_cpu_fallback(tmp_path):
    """Without a GPU the product path fails loudly instead of computing on the host."""
    p = tmp_path / "ok.bed"
    p.write_bytes(bytes([0x6C, 0x1B, 0x01]) + bytes(2 * 5))
    with pytest.raises(RuntimeError, match="CUDA"):
        capi.calculate(str(p), 5, 8, np.arange(5.0), 1.0, 0.01, 1e-5, 0.1)
    with pytest.raises(RuntimeError, match="CUDA"):
        capi.calculate_packed(np.zeros((5, 2), np.uint8), 5, 8, np.arange(5.0), 1.0, 0.01, 1e-5, 0.1)
