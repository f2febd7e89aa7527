"""GPU (-m gpu): the CUDA path, called through the C ABI, against the CPU oracle
(oracle/ld_oracle.py) and the committed golden vectors of the compiled reference.

Bar (BASELINE.json): filter decisions, retained-SNP lists, MAF and window counts bit-exact;
L2 <= 1e-9 relative, L2D <= 1e-9 absolute-or-relative-to-max(|L2D|, 1), against the FP64
build of the reference's own source (oracle-64)."""
import importlib
import os

import numpy as np
import pytest

from conftest import RESULT_KEYS, assert_ld_close
from nldsc_b200 import build, capi, synth
from oracle import ld_oracle

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module", autouse=True)
def built():
    build.build_module()
    assert capi.device_count() >= 1


def _ref64(g):
    return {k: g["ref64_" + k] for k in RESULT_KEYS}


def _run(g, **kw):
    return capi.calculate_packed(g["packed"], g["packed"].shape[0], g["n_org"], g["positions"], g["ld_wind"],
                                 g["maf"], g["std_thr"], g["rsq_thr"], **kw)


def _case(n_org, n_snp, missing, seed, wind=1.0, density=340.0, maf=1e-4, std_thr=1e-5):
    spec = synth.SynthSpec(n_org=n_org, n_snp=n_snp, missing=missing, seed=seed, density=density)
    packed = synth.generate_packed(spec)
    cm, _ = synth.genetic_map(spec)
    return {"packed": packed, "n_org": n_org, "positions": cm, "ld_wind": wind, "maf": maf, "std_thr": std_thr,
            "rsq_thr": 1.0 / n_snp, "name": f"n{n_org}_m{n_snp}"}


def _oracle(g):
    return ld_oracle.ld_scores(g["packed"], g["n_org"], g["positions"], g["ld_wind"], g["maf"], g["std_thr"],
                               g["rsq_thr"])


# ---- unpack + count ------------------------------------------------------------------------

@pytest.mark.parametrize("n_org", [1, 3, 4, 127, 128, 129, 1001, 1002, 1003, 4096])
def test_unpack_and_count_match_decoder(n_org):
    """BedStreamReader::read (stream.h:43-69) incl. the N%4 tail of SURVEY F4."""
    rng = np.random.default_rng(n_org)
    n_snp = 37
    bps = (n_org + 3) // 4
    packed = rng.integers(0, 256, (n_snp, bps), dtype=np.uint8)      # every code, every tail pattern
    codes = ld_oracle.decode_bed_msb_first(packed, n_org)
    planes, counts = capi.debug_unpack(packed, n_org)
    want = np.stack([(codes == 2).sum(1), (codes == 3).sum(1), (codes == 1).sum(1)], 1)
    assert np.array_equal(counts, want)
    # planes hold the same multiset of samples per SNP (order inside a byte differs)
    g, m, h, a = (planes[p].astype(np.int64) for p in range(4))
    assert np.array_equal(h.sum(1), want[:, 0]) and np.array_equal(a.sum(1), want[:, 1])
    assert np.array_equal(m.sum(1), want[:, 2]) and np.array_equal(g, h + 2 * a)
    assert np.all((h + a + m) <= 1)
    assert not planes[:, :, bps * 4:].any()                          # K padding is zero
    # exact placement: slot (byte j, LSB pair k) -> organism 4j+k
    lsb = np.stack([(packed >> (2 * k)) & 3 for k in range(4)], 2).reshape(n_snp, bps * 4)
    valid = np.ones(bps * 4, bool)
    if n_org % 4:
        valid[(bps - 1) * 4:(bps - 1) * 4 + (4 - n_org % 4)] = False  # low pairs of the last byte are dropped
    assert np.array_equal(h[:, :bps * 4], (lsb == 2) & valid)
    assert np.array_equal(a[:, :bps * 4], (lsb == 3) & valid)
    assert np.array_equal(m[:, :bps * 4], (lsb == 1) & valid)


@pytest.mark.parametrize("n_org", [1, 3, 4, 31, 32, 33, 255, 256, 257, 1003, 4099])
def test_fp4_unpack_equals_int8_unpack(n_org):
    """The fp4 planes (e2m1 nibbles for tcgen05.mma.kind::mxf4) hold exactly the values of the int8 planes, organism for organism,
    zero padding included, and the fused genotype counts agree."""
    rng = np.random.default_rng(100 + n_org)
    packed = rng.integers(0, 256, (41, (n_org + 3) // 4), dtype=np.uint8)
    p8, c8 = capi.debug_unpack(packed, n_org)
    p4, c4 = capi.debug_unpack_fp4(packed, n_org)
    assert np.array_equal(c8, c4)
    assert p4.min() >= 0                                   # only the nibbles 0x0, 0x2, 0x4 occur
    k = min(p8.shape[2], p4.shape[2])
    assert np.array_equal(p8[:, :, :k], p4[:, :, :k])
    assert not p8[:, :, k:].any() and not p4[:, :, k:].any()


def test_device_windows_equal_host_windows():
    """plan.cu windows_kernel vs the host sweep (tools.h:43-44 predicate), incl. ties, zero and huge windows."""
    rng = np.random.default_rng(3)
    pos = np.sort(rng.uniform(0, 40, 70001))
    pos[500:540] = pos[500]
    pos[-3:] = pos[-3]
    for wind in (0.0, 1e-9, 0.37, 1.0, 1e6):
        lo_h, hi_h = capi.plan_windows(pos, wind)
        lo_d, hi_d = capi.plan_windows_device(pos, wind)
        assert np.array_equal(lo_h, lo_d) and np.array_equal(hi_h, hi_d), wind
    bp = np.sort(rng.integers(1, 10**8, 5000)).astype(np.float64)
    for wind in (0.0, 1000.0, 5e6):
        assert all(np.array_equal(a, b) for a, b in zip(capi.plan_windows(bp, wind), capi.plan_windows_device(bp, wind)))


# ---- band GEMM: exact integer cross-counts -----------------------------------------------

@pytest.mark.parametrize("impl", ["simt", "tc"])
@pytest.mark.parametrize("n_org,n_snp", [(1003, 200), (130, 64), (5000, 130)])
def test_cross_counts_exact(impl, n_org, n_snp, monkeypatch):
    monkeypatch.setenv("NLDSC_BAND", impl)
    rng = np.random.default_rng(n_org + n_snp)
    packed = rng.integers(0, 256, (n_snp, (n_org + 3) // 4), dtype=np.uint8)
    codes = ld_oracle.decode_bed_msb_first(packed, n_org)
    nb = (n_snp + 63) // 64
    pad = np.zeros((nb * 64, n_org), np.uint8)
    pad[:n_snp] = codes
    for ib, jb in {(0, 0), (0, nb - 1), (nb - 1, 0), (nb - 1, nb - 1)}:
        got = capi.debug_cross_counts(packed, n_org, ib, jb)
        want = ld_oracle.cross_counts(pad[ib * 64:ib * 64 + 64], pad[jb * 64:jb * 64 + 64])
        assert np.array_equal(got, want), (impl, ib, jb)


# ---- end to end ------------------------------------------------------------------------------

@pytest.mark.parametrize("impl", ["simt", "tc", "tc_narrow"])
def test_golden_reference_parity(golden, impl, monkeypatch):
    monkeypatch.setenv("NLDSC_BAND", impl.split("_")[0])
    if impl == "tc_narrow":
        monkeypatch.setenv("NLDSC_WIDE", "0")
    got = _run(golden)
    assert_ld_close(got, _ref64(golden), label=f"{golden['name']}[{impl}]")


def test_cfg1_full_parity():
    """BASELINE.json configs[0]: N=1,000 x 10,000 SNPs, 1 cM, no missing -- every SNP vs the oracle."""
    g = _case(1000, 10000, 0.0, seed=1)
    assert_ld_close(_run(g), _oracle(g), label="cfg1")


def test_missing_and_ragged_parity():
    g = _case(2051, 3000, 0.02, seed=2)           # n % 4 == 3, 2% missing (configs[1] shape, scaled)
    assert_ld_close(_run(g), _oracle(g), label="ragged")


def test_tc_agrees_with_simt_yardstick(monkeypatch):
    """Same integer cross-counts, same epilogue; only the FP64 summation order inside a tile differs."""
    g = _case(777, 1500, 0.03, seed=5)
    monkeypatch.setenv("NLDSC_BAND", "tc")
    a = _run(g)
    monkeypatch.setenv("NLDSC_BAND", "simt")
    b = _run(g)
    assert_ld_close(a, b, rtol=1e-12, label="tc vs simt")


def test_chunked_ring_equals_single_pass_bitwise(monkeypatch):
    """Small chunks and a ring barely wider than one window (rows wrap many times)."""
    g = _case(600, 4000, 0.02, seed=6, wind=0.5)
    base = _run(g)
    monkeypatch.setenv("NLDSC_CHUNK_ROWS", "128")
    monkeypatch.setenv("NLDSC_PLANE_BYTES", str(640 * 4 * 640))      # 640 SNP rows of 4 x 640 bytes
    small = _run(g)
    for k in RESULT_KEYS:
        assert np.array_equal(base[k], small[k], equal_nan=True), k


def test_sharded_equals_single_bitwise():
    """contiguous SNP ranges + halo, no exchange; gathered == single-GPU bitwise."""
    g = _case(900, 3000, 0.02, seed=7)
    n_snp = 3000
    base = _run(g)
    cuts = capi.plan_shards(g["positions"], g["ld_wind"], 3)
    got = {k: np.empty_like(base[k]) for k in RESULT_KEYS}
    for k in range(3):
        with capi.Session(0, n_snp, g["n_org"], g["positions"], g["ld_wind"], g["maf"], g["std_thr"], g["rsq_thr"],
                          row_begin=int(cuts[k]), row_end=int(cuts[k + 1])) as s:
            s.upload(g["packed"])
            s.run()
            part = s.download()
            st = s.stats()
            assert st["load_begin"] <= cuts[k] and st["load_end"] >= cuts[k + 1]
            if k == 1:
                assert st["load_begin"] > 0 and st["load_end"] < n_snp      # really a halo, not the whole file
        for key in RESULT_KEYS:
            got[key][cuts[k]:cuts[k + 1]] = part[key]
    for key in RESULT_KEYS:
        assert np.array_equal(base[key], got[key], equal_nan=True), key
    # unaligned cut (not a multiple of 64)
    with capi.Session(0, n_snp, g["n_org"], g["positions"], g["ld_wind"], g["maf"], g["std_thr"], g["rsq_thr"],
                      row_begin=1000, row_end=1777) as s:
        s.upload(g["packed"])
        s.run()
        part = s.download()
    for key in RESULT_KEYS:
        assert np.array_equal(base[key][1000:1777], part[key], equal_nan=True), key


@pytest.mark.parametrize("shards", [2, 3, 5])
def test_run_sharded_matches_single(shards, monkeypatch, tmp_path):
    """run_sharded (session.cu): one host thread per shard, halo reads, host-side gather -- bitwise equal to the
    one-shard call, from a host buffer and from a .bed.  One shard per GPU when the box has `shards` GPUs
    (NLDSC_GPUS); otherwise the NLDSC_TEST_SHARDS hook deals the same shards round-robin to the GPUs there are, so
    this never skips."""
    g = _case(900, 3000, 0.02, seed=8)
    base = _run(g)
    with open(tmp_path / "s.bed", "wb") as f:
        f.write(synth.BED_MAGIC)
        f.write(g["packed"].tobytes())
    if capi.device_count() >= shards:
        monkeypatch.setenv("NLDSC_GPUS", str(shards))
    else:
        monkeypatch.setenv("NLDSC_TEST_SHARDS", str(shards))
    sharded = _run(g)
    from_file = capi.calculate(str(tmp_path / "s.bed"), 3000, g["n_org"], g["positions"], g["ld_wind"], g["maf"],
                               g["std_thr"], g["rsq_thr"])
    for key in RESULT_KEYS:
        assert np.array_equal(base[key], sharded[key], equal_nan=True), key
        assert np.array_equal(base[key], from_file[key], equal_nan=True), key
    cuts = capi.plan_shards(g["positions"], g["ld_wind"], shards)
    assert len(set(cuts.tolist())) == shards + 1                 # every shard owns rows: the cut really is `shards` wide


def test_degenerate_snps():
    """Monomorphic SNPs are dropped (NaN / -1), collinear-dominance SNPs (no hom-second, residual
    std exactly 0) are kept but not dominance-usable, equal positions share a window."""
    g = _case(400, 700, 0.0, seed=9)
    codes = ld_oracle.decode_bed_msb_first(g["packed"], 400).copy()
    codes[5] = 0                                   # monomorphic
    codes[6][codes[6] == 3] = 2                    # no hom-second -> dominance coding collinear
    g["packed"] = synth.pack_codes(codes)
    # pack_codes is LSB-first, the decoder MSB-first: only the multiset per SNP matters
    g["positions"] = g["positions"].copy()
    g["positions"][300:320] = g["positions"][300]
    got, want = _run(g), _oracle(g)
    assert_ld_close(got, want, label="degenerate")
    assert np.isnan(got["l2"][5]) and got["l2_ws"][5] == -1 and got["maf"][5] == 0.0
    assert got["residuals_std"][6] == 0.0 and got["l2_ws"][6] >= 0


def test_unused_snps_negative_positions(monkeypatch):
    """position < 0 = "SNP not used" (tools.h:16-24): dropped row with NaN / -1 outputs and NaN MAF, in nobody's
    window; the others are scored as if it were not in the file.  Leading, trailing and interior unused SNPs, also
    through the sharded path."""
    g = _case(500, 1300, 0.02, seed=13)
    pos = g["positions"].copy()
    unused = np.r_[0:3, 400, 401, 777, 1290:1300]
    pos[unused] = -1.0
    pos[5] = -0.25
    g["positions"] = pos
    got, want = _run(g), _oracle(g)
    assert_ld_close(got, want, label="unused")
    assert np.all(np.isnan(got["maf"][unused])) and np.all(got["l2_ws"][unused] == -1) and np.isnan(got["l2"][5])
    # the used SNPs score exactly as in the file without the unused rows (same windows, same neighbours)
    used = pos >= 0
    sub = dict(g, packed=g["packed"][used], positions=pos[used])
    alone = _run(sub)
    for k in ("l2_ws", "l2d_ws", "l2d_wse", "maf"):
        assert np.array_equal(got[k][used], alone[k], equal_nan=True), k
    assert np.allclose(got["l2"][used], alone["l2"], rtol=1e-12, atol=0, equal_nan=True)
    monkeypatch.setenv("NLDSC_TEST_SHARDS", "3")
    sharded = _run(g)
    for k in RESULT_KEYS:
        assert np.array_equal(got[k], sharded[k], equal_nan=True), k
    with pytest.raises(ValueError, match="non-decreasing"):
        bad = dict(g, positions=np.r_[pos[:600], pos[10], pos[601:]])
        _run(bad)


def test_bfile_path_and_module(tmp_path):
    """nldsc_ld_calculate on a real .bed, and the pybind11 `_ldscore` drop-in on top of it."""
    spec = synth.SynthSpec(n_org=501, n_snp=900, missing=0.02, seed=12)
    packed = synth.generate_packed(spec)
    info = synth.write_bfile(str(tmp_path / "t"), spec, packed)
    want = ld_oracle.ld_scores(packed, 501, info["cm"], 1.0, 1e-4, 1e-5, 1 / 900)
    got = capi.calculate(str(tmp_path / "t.bed"), 900, 501, info["cm"], 1.0, 1e-4, 1e-5, 1 / 900)
    assert_ld_close(got, want, label="bfile")
    m = importlib.import_module("nldsc_b200._ldscore")
    r = m.calculate(m.LDScoreParams(str(tmp_path / "t.bed"), n_snp=900, n_org=501, ld_wind=1.0, maf=1e-4,
                                    std_thr=1e-5, rsq_thr=1 / 900, positions=list(info["cm"])))
    for key in RESULT_KEYS:
        assert np.array_equal(np.asarray(getattr(r, key)), got[key], equal_nan=True), key
    assert isinstance(r.l2, list) and isinstance(r.l2_ws[0], int)


def test_genome_driver_writes_l2_and_m_files(tmp_path):
    """three chromosome bfiles scored in one process by two worker threads
    (`_ldscore.calculate_on`), each `.L2` equal to the single-chromosome path and each `.M` equal to
    h2's own fallback formula on that file (reference h2/common.py:126-130)."""
    import dataclasses
    import pandas as pd
    from nldsc_b200 import ldscore
    specs = {}
    for c, (n_snp, seed) in {"1": (700, 31), "2": (500, 32), "21": (300, 33)}.items():
        spec = dataclasses.replace(synth.SynthSpec(n_org=403, n_snp=n_snp, missing=0.02, seed=seed), chrom=int(c))
        packed = synth.generate_packed(spec, workers=1)
        info = synth.write_bfile(str(tmp_path / f"g.{c}"), spec, packed)
        specs[c] = (packed, info)
    summary = ldscore.estimate_lds_genome(str(tmp_path / "g.@"), 1.0, "cm", 1e-4, 1e-5, out_dir=str(tmp_path / "out"),
                                          gpus=2 if capi.device_count() >= 2 else 1)
    assert list(summary["CHR"]) == ["1", "2", "21"]
    for c, (packed, info) in specs.items():
        n_snp = info["n_snp"]
        want = ld_oracle.ld_scores(packed, 403, info["cm"], 1.0, 1e-4, 1e-5, 1.0 / n_snp)
        df = pd.read_csv(tmp_path / "out" / f"{c}.L2", sep="\t")
        assert list(df.columns) == ["CHR", "SNP", "BP", "L2", "L2D", "MAF", "WSA", "WSD", "WSDE", "RSTD"]
        keep = want["keep"]
        assert np.array_equal(df["L2"].notna().to_numpy(), keep)
        assert np.array_equal(df["WSA"].to_numpy(), want["l2_ws"])
        assert np.allclose(df["L2"].to_numpy()[keep], want["l2"][keep], rtol=0, atol=6e-6)     # %.5f in the file
        # h2's fallback on the file it will read
        score = df.dropna().drop_duplicates(subset="SNP")
        m = len(score["L2"])
        md = int(m * (score["WSDE"] / score["WSA"]).mean())
        got = pd.read_csv(tmp_path / "out" / f"{c}.M", sep="\t")
        assert (int(got["M"].iloc[0]), int(got["MD"].iloc[0])) == (m, md)
        row = summary[summary["CHR"] == c].iloc[0]
        assert (row["M"], row["MD"], row["SNPs"]) == (m, md, n_snp)


def test_multi_piece_bed_read_equals_host_buffer(tmp_path, monkeypatch):
    """The .bed path (parallel reader threads, several 8 MB pieces, pinned slot ring) must hand the kernels the
    same bytes as a host buffer: results bitwise equal to nldsc_ld_calculate_packed on a 44 MB file."""
    rng = np.random.default_rng(17)
    n_org, n_snp = 20001, 8800
    bps = (n_org + 3) // 4
    # mostly hom-first / het / hom-second, a few missing codes (01); any byte pattern is a valid .bed row
    codes = rng.choice(np.array([0, 2, 3, 1], dtype=np.uint8), size=(n_snp, bps * 4), p=[0.55, 0.3, 0.13, 0.02])
    packed = (codes[:, 0::4] | (codes[:, 1::4] << 2) | (codes[:, 2::4] << 4) | (codes[:, 3::4] << 6)).astype(np.uint8)
    pos = np.cumsum(rng.exponential(1 / 340.0, n_snp))
    with open(tmp_path / "big.bed", "wb") as f:
        f.write(synth.BED_MAGIC)
        f.write(packed.tobytes())
    args = (n_snp, n_org, pos, 0.3, 1e-4, 1e-5, 1.0 / n_snp)
    mem = capi.calculate_packed(packed, *args)
    for lag in ("0", "3000", "20000"):      # reader 0 held back: its pieces reach the copy stream after later ones
        monkeypatch.setenv("NLDSC_TEST_READER0_LAG_US", lag)
        got = capi.calculate(str(tmp_path / "big.bed"), *args)
        for key in RESULT_KEYS:
            assert np.array_equal(got[key], mem[key], equal_nan=True), key
    assert (mem["l2_ws"] >= 0).sum() > 8000


def _session_stats(g, **env):
    with capi.Session(0, g["packed"].shape[0], g["n_org"], g["positions"], g["ld_wind"], g["maf"], g["std_thr"], g["rsq_thr"]) as s:
        s.upload(g["packed"])
        s.run()
        out = s.download()
        return out, s.stats()


def test_x160_body_equals_wide_body_bitwise(monkeypatch):
    """band_x160_body (one i-pair x five 32-SNP column groups, two mxf4 MMAs of N = 256 + 224 per 64 organisms; the default
    for the 6-plane work from N = 4,096) writes the same partial sums, group by group, as the 2 x 2-block supertiles:
    bit-equal scores, oracle bar met -- on ragged ends, a wrapping ring, shards, and beside missing-free chunks."""
    g = _case(1003, 3000, 0.03, seed=29, wind=0.7)
    monkeypatch.setenv("NLDSC_X160", "0")
    wide, stw = _session_stats(g)
    assert stw["xtiles"] == 0 and stw["xtiles_run"] == 0
    monkeypatch.setenv("NLDSC_X160", "1")
    x, stx = _session_stats(g)
    assert stx["xtiles_run"] == stx["xtiles"] > 0
    assert_ld_close(x, _oracle(g), label="x160")
    for k in RESULT_KEYS:
        assert np.array_equal(x[k], wide[k], equal_nan=True), k
    # the default picks it by N
    monkeypatch.delenv("NLDSC_X160")
    big = _case(4100, 700, 0.02, seed=30)
    d, std_ = _session_stats(big)
    assert std_["xtiles_run"] > 0
    assert_ld_close(d, _oracle(big), label="x160 default")
    monkeypatch.setenv("NLDSC_X160", "0")
    w = _run(big)
    for k in RESULT_KEYS:
        assert np.array_equal(d[k], w[k], equal_nan=True), k
    # ragged ends, tiny shards, a wrapping ring, several shards
    for n_snp, env in ((65, {"NLDSC_CHUNK_ROWS": "64"}), (129, {"NLDSC_CHUNK_ROWS": "64"}), (333, {"NLDSC_CHUNK_ROWS": "128"}),
                       (961, {"NLDSC_CHUNK_ROWS": "192"}), (2100, {"NLDSC_TEST_SHARDS": "3"}), (1500, {"NLDSC_OVERLAP": "0"})):
        h = _case(517, n_snp, 0.02, seed=60 + n_snp, wind=0.3)
        for k, v in env.items():
            monkeypatch.setenv(k, v)
        monkeypatch.setenv("NLDSC_X160", "1")
        a = _run(h)
        assert_ld_close(a, _oracle(h), label=f"x160 ragged {n_snp}")
        monkeypatch.setenv("NLDSC_X160", "0")
        b = _run(h)
        for k in RESULT_KEYS:
            assert np.array_equal(a[k], b[k], equal_nan=True), (n_snp, k)
        for k in env:
            monkeypatch.delenv(k)
    # missing genotypes in SNPs 1,400-1,500 only: missing-free chunks take the 2-plane body, the others the 160-column one
    m = _case(1003, 3000, 0.0, seed=23, wind=0.7)
    codes = ld_oracle.decode_bed_msb_first(m["packed"], 1003).copy()
    rng = np.random.default_rng(5)
    codes[1400:1500][rng.random((100, 1003)) < 0.03] = 1
    mixed = dict(m, packed=synth.pack_codes(codes))
    monkeypatch.setenv("NLDSC_CHUNK_ROWS", "256")
    monkeypatch.setenv("NLDSC_X160", "1")
    a, sta = _session_stats(mixed)
    assert 0 < sta["mfree_chunks"] < sta["chunks"] and 0 < sta["xtiles_run"] < sta["xtiles"], sta
    assert_ld_close(a, _oracle(mixed), label="x160 mixed")
    monkeypatch.setenv("NLDSC_X160", "0")
    b = _run(mixed)
    for k in RESULT_KEYS:
        assert np.array_equal(a[k], b[k], equal_nan=True), k


def test_missing_free_kernel_equals_six_plane_bitwise(monkeypatch):
    """band_tc_mfree_kernel (g x {h, a}: the reference's own two dot products, tools.h:87-92) takes the chunks whose
    columns hold no missing genotype; it must give the bits the 6-plane kernels give (NLDSC_MFREE=0), meet the oracle
    bar, and hand chunks WITH missing genotypes back to the 6-plane kernel -- also in one and the same run."""
    g = _case(1003, 3000, 0.0, seed=23, wind=0.7)
    out, st = _session_stats(g)
    assert st["mfree_chunks"] == st["chunks"] >= 1 and st["mfree_qtiles"] > 0
    assert_ld_close(out, _oracle(g), label="mfree")
    monkeypatch.setenv("NLDSC_MFREE", "0")
    six, st6 = _session_stats(g)
    assert st6["mfree_chunks"] == 0
    for k in RESULT_KEYS:
        assert np.array_equal(out[k], six[k], equal_nan=True), k
    monkeypatch.delenv("NLDSC_MFREE")
    # missing genotypes in SNPs 1,400-1,500 only: with 256-row chunks some chunks are missing-free, others are not
    codes = ld_oracle.decode_bed_msb_first(g["packed"], 1003).copy()
    rng = np.random.default_rng(5)
    codes[1400:1500][rng.random((100, 1003)) < 0.03] = 1
    mixed = dict(g, packed=synth.pack_codes(codes))
    want = _oracle(mixed)
    for chunk_rows in ("256", "512"):
        monkeypatch.setenv("NLDSC_CHUNK_ROWS", chunk_rows)
        monkeypatch.delenv("NLDSC_MFREE", raising=False)
        a, sta = _session_stats(mixed)
        assert 0 < sta["mfree_chunks"] < sta["chunks"], sta
        assert_ld_close(a, want, label="mixed " + chunk_rows)
        monkeypatch.setenv("NLDSC_MFREE", "0")
        b, _ = _session_stats(mixed)
        for k in RESULT_KEYS:
            assert np.array_equal(a[k], b[k], equal_nan=True), (chunk_rows, k)
    # ragged ends: odd numbers of i-blocks and j-blocks, tiny shards, a wrapping ring
    monkeypatch.delenv("NLDSC_MFREE", raising=False)
    for n_snp, chunk_rows in ((65, "64"), (129, "64"), (333, "128"), (961, "192")):
        h = _case(517, n_snp, 0.0, seed=40 + n_snp, wind=0.3)
        monkeypatch.setenv("NLDSC_CHUNK_ROWS", chunk_rows)
        monkeypatch.delenv("NLDSC_MFREE", raising=False)
        a = _run(h)
        assert_ld_close(a, _oracle(h), label=f"mfree ragged {n_snp}")
        monkeypatch.setenv("NLDSC_MFREE", "0")
        b = _run(h)
        for k in RESULT_KEYS:
            assert np.array_equal(a[k], b[k], equal_nan=True), (n_snp, k)


def test_wide_supertiles_equal_narrow_bitwise(monkeypatch):
    """band_tc_wide_kernel (two j-blocks per A tile, eight epilogue warps; the default) and band_tc_kernel (one j-block,
    double-buffered accumulators; NLDSC_WIDE=0) put their partial sums into the same slots: they must agree bit for
    bit, also with chunks and a wrapping ring, and pass the oracle bar."""
    g = _case(1003, 3000, 0.02, seed=21, wind=0.7)
    wide = _run(g)
    assert_ld_close(wide, _oracle(g), label="wide")
    monkeypatch.setenv("NLDSC_WIDE", "0")
    narrow = _run(g)
    for k in RESULT_KEYS:
        assert np.array_equal(narrow[k], wide[k], equal_nan=True), k
    for w in ("0", "1"):
        monkeypatch.setenv("NLDSC_WIDE", w)
        monkeypatch.setenv("NLDSC_CHUNK_ROWS", "256")
        monkeypatch.setenv("NLDSC_PLANE_BYTES", str(1024 * 4 * 1024))
        small = _run(g)
        for k in RESULT_KEYS:
            assert np.array_equal(wide[k], small[k], equal_nan=True), (w, k)


# ---- BASELINE.json configs at full / near-full size: oracle on many rows ---------------------------------------
# The workloads are drawn on the GPU (synth.generate_packed_cuda: torch RNG + elementwise ops, plumbing only) because the
# NumPy generator needs minutes at these sizes; the oracle then scores random rows, each against its complete window.

def _gpu_case(n_org, n_snp, missing, seed, density, wind):
    import torch
    spec = synth.SynthSpec(n_org=n_org, n_snp=n_snp, missing=missing, seed=seed, density=density)
    packed = synth.generate_packed_cuda(spec, "cuda:0").cpu().numpy()
    torch.cuda.empty_cache()
    cm, _ = synth.genetic_map(spec)
    return {"packed": packed, "n_org": n_org, "positions": cm, "ld_wind": wind, "maf": 1e-4, "std_thr": 1e-5,
            "rsq_thr": 1.0 / n_snp}


def _check_rows(g, out, n_clusters, per_cluster, spread, max_span, seed=0):
    """Oracle on n_clusters x per_cluster random rows (rows of a cluster lie within `spread` SNPs of each other, so
    they share one pass over the union of their windows)."""
    n_snp = g["packed"].shape[0]
    rng = np.random.default_rng(seed)
    centres = rng.choice(max(1, n_snp - spread), size=n_clusters, replace=False)
    rows = np.unique(np.concatenate([c + rng.choice(spread, per_cluster, replace=False) for c in centres]))
    want = ld_oracle.ld_scores_rows(g["packed"], g["n_org"], g["positions"], g["ld_wind"], g["maf"], g["std_thr"],
                                    g["rsq_thr"], rows, max_span=max_span)
    r, kw = want["rows"], want["keep"]
    assert np.array_equal(~np.isnan(out["l2"][r]), kw), "retained-SNP list"
    assert np.array_equal(out["maf"][r], want["maf"], equal_nan=True), "maf bit-exact"
    for k in ("l2_ws", "l2d_ws", "l2d_wse"):
        assert np.array_equal(out[k][r][kw], want[k][kw]), k
    assert np.all(out["l2_ws"][r][~kw] == -1)
    l2e = np.max(np.abs(out["l2"][r][kw] - want["l2"][kw]) / np.maximum(np.abs(want["l2"][kw]), 1.0))
    l2de = np.max(np.abs(out["l2d"][r][kw] - want["l2d"][kw]) / np.maximum(np.abs(want["l2d"][kw]), 1.0))
    rse = np.max(np.abs(out["residuals_std"][r][kw] - want["residuals_std"][kw]))
    assert l2e <= 1e-9 and l2de <= 1e-9 and rse <= 1e-9, (l2e, l2de, rse)
    return int(kw.sum())


def _window_count_property(g, out):
    keep = out["l2_ws"] >= 0
    lo, hi = capi.plan_windows(g["positions"], g["ld_wind"])
    ck = np.concatenate([[0], np.cumsum(keep)])
    assert np.array_equal(out["l2_ws"][keep], (ck[hi + 1] - ck[lo])[keep] - 1)          # ldscalc.h:52, stream.h:142-155
    assert np.all(out["l2d_ws"][keep] <= out["l2_ws"][keep]) and np.all(out["l2d_wse"][keep] <= out["l2d_ws"][keep])


def test_cfg2_full_size_oracle_rows():
    """BASELINE.json configs[1] IN FULL -- the benchmark's workload: N=10,000 x 100,000 SNPs, 2% missing, 1 cM --
    with the oracle on >= 200 random rows spread over the whole chromosome."""
    g = _gpu_case(10_000, 100_000, 0.02, seed=2, density=340.0, wind=1.0)
    out = _run(g)
    _window_count_property(g, out)
    assert _check_rows(g, out, n_clusters=30, per_cluster=8, spread=240, max_span=2400) >= 200


def test_cfg4_ukb_slice_oracle_rows():
    """BASELINE.json configs[3] shape (N=315,599, N % 4 == 3, 1 cM): a slice of NLDSC_TEST_UKB_SNPS SNPs (default 4,096;
    the full genome is scripts/ukb_genome.py), oracle on >= 20 rows."""
    n_snp = int(os.environ.get("NLDSC_TEST_UKB_SNPS", "4096"))
    g = _gpu_case(315_599, n_snp, 0.02, seed=7, density=340.0, wind=1.0)
    out = _run(g)
    _window_count_property(g, out)
    assert _check_rows(g, out, n_clusters=2, per_cluster=12, spread=64, max_span=1600, seed=1) >= 20


def test_cfg5_wide_window_slice_oracle_rows():
    """BASELINE.json configs[4] shape (N=100,000, dense map, 10 cM window ~ 20,000 SNPs): a slice of
    NLDSC_TEST_CFG5_SNPS SNPs (default 6,144), oracle on >= 20 rows with windows thousands of SNPs wide."""
    n_snp = int(os.environ.get("NLDSC_TEST_CFG5_SNPS", "6144"))
    g = _gpu_case(100_000, n_snp, 0.02, seed=5, density=1000.0, wind=10.0)
    out = _run(g)
    _window_count_property(g, out)
    assert _check_rows(g, out, n_clusters=2, per_cluster=12, spread=48, max_span=30000, seed=2) >= 20
