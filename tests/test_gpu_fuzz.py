"""GPU (-m gpu): seeded random shapes x execution knobs against the CPU oracle.

Every case draws organisms, SNPs, missingness, map density and window at random, then runs the C-ABI path
under one randomly chosen execution variant (tiny chunks, a plane ring barely wider than a window, one
stream, one band stream, SIMT band kernel, the narrow tcgen05 kernel, 2 or 3 shards -- one per GPU when the box has
as many, otherwise through the NLDSC_TEST_SHARDS hook, which runs the same thread-per-shard / halo / gather code with the
shards dealt round-robin to the GPUs there are).  All variants must meet the
parity bar against the oracle, and be bitwise equal to the default variant of the same case."""
import os

import numpy as np
import pytest

from conftest import RESULT_KEYS, assert_ld_close
from nldsc_b200 import build, capi, synth
from oracle import ld_oracle

pytestmark = pytest.mark.gpu

VARIANTS = [
    {},
    {"NLDSC_CHUNK_ROWS": "64"},
    {"NLDSC_CHUNK_ROWS": "192", "NLDSC_PLANE_BYTES": "ring"},
    {"NLDSC_OVERLAP": "0"},
    {"NLDSC_BAND_STREAMS": "1", "NLDSC_CHUNK_ROWS": "128"},
    {"NLDSC_BAND": "simt"},
    {"NLDSC_GPUS": "2"},
    {"NLDSC_WIDE": "0"},
    {"NLDSC_WIDE": "0", "NLDSC_CHUNK_ROWS": "128", "NLDSC_PLANE_BYTES": "ring"},
    {"NLDSC_CHUNK_ROWS": "128", "NLDSC_PLANE_BYTES": "ring", "NLDSC_GPUS": "2"},
    {"NLDSC_TEST_SHARDS": "3"},
    {"NLDSC_FUSED": "1"},
    {"NLDSC_FUSED": "1", "NLDSC_CHUNK_ROWS": "64"},
    {"NLDSC_MFREE": "0"},
    {"NLDSC_MFREE": "0", "NLDSC_CHUNK_ROWS": "128"},
    {"NLDSC_FP4": "0"},
    {"NLDSC_FP4": "0", "NLDSC_CHUNK_ROWS": "128", "NLDSC_PLANE_BYTES": "ring"},
    {"NLDSC_FP4": "0", "NLDSC_MFREE": "0"},
    {"NLDSC_X160": "1"},                               # the 160-column supertiles (default only from N = 4,096)
    {"NLDSC_X160": "1", "NLDSC_CHUNK_ROWS": "128", "NLDSC_PLANE_BYTES": "ring"},
    {"NLDSC_X160": "1", "NLDSC_TEST_SHARDS": "3", "NLDSC_MFREE": "0"},
]


def _plane_row_bytes(n_org, variant):
    """Bytes of one plane row: int8 planes (narrow kernel, SIMT yardstick, NLDSC_FP4=0) pad N to 128 organisms, one byte each;
    fp4 planes pad to 256 organisms, two per byte."""
    int8 = variant.get("NLDSC_FP4") == "0" or variant.get("NLDSC_WIDE") == "0" or variant.get("NLDSC_BAND") == "simt"
    return (n_org + 127) // 128 * 128 if int8 else (n_org + 255) // 256 * 128


def _shards_for_this_box(variant):
    """NLDSC_GPUS=k needs k devices; on a smaller box the same k-shard cut runs through NLDSC_TEST_SHARDS
    (threads, halos and the gather are the same code; only the device every shard lands on differs)."""
    k = variant.get("NLDSC_GPUS")
    if k is not None and capi.device_count() < int(k):
        variant.pop("NLDSC_GPUS")
        variant["NLDSC_TEST_SHARDS"] = k
    return variant


@pytest.fixture(scope="module", autouse=True)
def built():
    build.build_module()


def _draw(seed):
    rng = np.random.default_rng(1000 + seed)
    n_org = int(rng.choice([3, 5, 64, 127, 128, 129, 255, 300, 513, 777, 1001, 1026, 1500]))
    n_snp = int(rng.choice([1, 2, 63, 64, 65, 127, 129, 200, 333, 640, 900]))
    missing = float(rng.choice([0.0, 0.0, 0.01, 0.05, 0.3]))
    density = float(rng.choice([20.0, 100.0, 340.0, 2000.0]))
    wind = float(rng.choice([0.01, 0.1, 0.5, 1.0, 50.0]))
    return n_org, n_snp, missing, density, wind


# NLDSC_FUZZ_SEEDS=500 turns the same test into a soak run (scripts/soak.sh)
@pytest.mark.parametrize("seed", range(int(os.environ.get("NLDSC_FUZZ_SEEDS", "50"))))
def test_random_case_matches_oracle(seed, monkeypatch):
    n_org, n_snp, missing, density, wind = _draw(seed)
    spec = synth.SynthSpec(n_org=n_org, n_snp=n_snp, missing=missing, seed=500 + seed, density=density)
    packed = synth.generate_packed(spec, workers=1)
    cm, _ = synth.genetic_map(spec)
    args = (n_snp, n_org, cm, wind, 1e-4, 1e-5, 1.0 / max(n_snp, 11))
    want = ld_oracle.ld_scores(packed, n_org, cm, wind, 1e-4, 1e-5, 1.0 / max(n_snp, 11))
    base = capi.calculate_packed(packed, *args)
    label = f"seed {seed}: N={n_org} M={n_snp} miss={missing} dens={density} wind={wind}"
    # r2adj(dom) sits within rounding of rsq_thr for some pairs at tiny N: WSDE may differ there by design of the bar
    assert_ld_close(base, want, exact_wsde=n_org >= 64, label=label)
    variant = dict(VARIANTS[seed % len(VARIANTS)])
    if variant.get("NLDSC_PLANE_BYTES") == "ring":
        lo, hi = capi.plan_windows(cm, wind)
        span = int(np.max(hi - lo)) + 1
        rows = (span // 64 + 6) * 64                               # a block pair's columns, and little more
        variant["NLDSC_PLANE_BYTES"] = str(rows * 4 * _plane_row_bytes(n_org, variant))
    variant = _shards_for_this_box(variant)
    for k, v in variant.items():
        monkeypatch.setenv(k, v)
    got = capi.calculate_packed(packed, *args)
    if variant.get("NLDSC_BAND") == "simt":
        assert_ld_close(got, want, exact_wsde=n_org >= 64, label=label + " " + str(variant))
    else:
        for key in RESULT_KEYS:
            assert np.array_equal(got[key], base[key], equal_nan=True), f"{label} {variant}: {key} differs from the default run"


MID_VARIANTS = [
    {"NLDSC_CHUNK_ROWS": "512"},
    {"NLDSC_CHUNK_ROWS": "1024", "NLDSC_PLANE_BYTES": "ring"},
    {"NLDSC_WIDE": "0"},
    {"NLDSC_OVERLAP": "0", "NLDSC_CHUNK_ROWS": "2048"},
    {"NLDSC_GPUS": "2"},
    {"NLDSC_BAND_STREAMS": "1", "NLDSC_WIDE_OVERLAP": "0"},
    {"NLDSC_FIRST_BYTES": "1000000", "NLDSC_CHUNK_BYTES": "3000000"},
    {"NLDSC_TEST_SHARDS": "4", "NLDSC_CHUNK_ROWS": "1024"},
    {"NLDSC_FUSED": "1", "NLDSC_CHUNK_ROWS": "1024"},
    {"NLDSC_MFREE": "0"},
    {"NLDSC_FP4": "0"},
    {"NLDSC_FP4": "0", "NLDSC_CHUNK_ROWS": "1024", "NLDSC_PLANE_BYTES": "ring"},
    {"NLDSC_X160": "1"},
    {"NLDSC_X160": "1", "NLDSC_CHUNK_ROWS": "1024", "NLDSC_PLANE_BYTES": "ring"},
]


# Multi-chunk shapes (thousands of SNPs, windows hundreds of SNPs wide): every execution variant must
# reproduce the default run bit for bit, and 12 random rows must meet the oracle's bar.
# Four cases by default; NLDSC_FUZZ_MID=30 is the soak setting (scripts/soak.sh).
@pytest.mark.parametrize("seed", range(int(os.environ.get("NLDSC_FUZZ_MID", "4"))))
def test_midsize_case_variants_bitwise(seed, monkeypatch):
    rng = np.random.default_rng(7000 + seed)
    n_org = int(rng.choice([500, 1027, 2048, 3001, 5000]))
    n_snp = int(rng.choice([3000, 7777, 12000, 20000]))
    density = float(rng.choice([340.0, 2000.0, 6000.0]))
    wind = float(rng.choice([0.05, 0.2, 1.0]))
    spec = synth.SynthSpec(n_org=n_org, n_snp=n_snp, missing=float(rng.choice([0.0, 0.02])), seed=900 + seed,
                           density=density)
    packed = synth.generate_packed(spec, workers=4)
    cm, _ = synth.genetic_map(spec)
    args = (n_snp, n_org, cm, wind, 1e-4, 1e-5, 1.0 / n_snp)
    label = f"mid seed {seed}: N={n_org} M={n_snp} dens={density} wind={wind}"
    base = capi.calculate_packed(packed, *args)
    rows = np.sort(rng.choice(n_snp, size=12, replace=False))
    want = ld_oracle.ld_scores(packed, n_org, cm, wind, 1e-4, 1e-5, 1.0 / n_snp, rows=rows)
    assert np.array_equal(np.isnan(base["l2"]), ~want["keep"]), label + " retained-SNP list"
    assert np.array_equal(base["maf"], want["maf"], equal_nan=True), label + " maf"
    for r in rows:
        if not want["keep"][r]:
            continue
        assert base["l2_ws"][r] == want["l2_ws"][r] and base["l2d_ws"][r] == want["l2d_ws"][r], f"{label} row {r} counts"
        assert base["l2d_wse"][r] == want["l2d_wse"][r], f"{label} row {r} WSDE"
        assert abs(base["l2"][r] - want["l2"][r]) <= 1e-9 * max(abs(want["l2"][r]), 1.0), f"{label} row {r} L2"
        assert abs(base["l2d"][r] - want["l2d"][r]) <= 1e-9 * max(abs(want["l2d"][r]), 1.0), f"{label} row {r} L2D"
    for variant in (MID_VARIANTS[seed % len(MID_VARIANTS)], MID_VARIANTS[(seed + 3) % len(MID_VARIANTS)]):
        variant = dict(variant)
        if variant.get("NLDSC_PLANE_BYTES") == "ring":
            lo, hi = capi.plan_windows(cm, wind)
            rows_ring = ((int(np.max(hi - lo)) + 1) // 64 + 6) * 64 + 2048
            variant["NLDSC_PLANE_BYTES"] = str(rows_ring * 4 * _plane_row_bytes(n_org, variant))
        variant = _shards_for_this_box(variant)
        with monkeypatch.context() as mp:
            for k, v in variant.items():
                mp.setenv(k, v)
            got = capi.calculate_packed(packed, *args)
        for key in RESULT_KEYS:
            assert np.array_equal(got[key], base[key], equal_nan=True), f"{label} {variant}: {key} differs from the default run"


# The oracle the cases above are graded against is the NumPy restatement; these go to the COMPILED reference itself
# (oracle/_ref/libref_ld64.so: the reference's own sources, float storage promoted to double -- built by oracle/Makefile in the
# build container and shipped to the GPU box with the tree), through a real .bed file.
@pytest.mark.parametrize("seed", range(6))
def test_random_case_matches_compiled_reference(seed, tmp_path):
    from oracle import ref
    if not ref.available(64):
        pytest.skip("oracle/_ref/libref_ld64.so is not built on this box")
    rng = np.random.default_rng(4000 + seed)
    n_org = int(rng.choice([127, 300, 513, 1001, 1500]))
    n_snp = int(rng.choice([129, 333, 640, 900]))
    missing = float(rng.choice([0.0, 0.02, 0.1]))
    wind = float(rng.choice([0.1, 0.5, 1.0]))
    spec = synth.SynthSpec(n_org=n_org, n_snp=n_snp, missing=missing, seed=4500 + seed, density=float(rng.choice([100.0, 340.0])))
    packed = synth.generate_packed(spec, workers=1)
    cm, _ = synth.genetic_map(spec)
    bed = tmp_path / "c.bed"
    with open(bed, "wb") as f:
        f.write(synth.BED_MAGIC)
        f.write(packed.tobytes())
    a = (n_snp, n_org, cm, wind, 1e-4, 1e-5, 1.0 / n_snp)
    want = ref.calculate(str(bed), *a, flavour=64)
    label = f"compiled reference, seed {seed}: N={n_org} M={n_snp} miss={missing} wind={wind}"
    assert_ld_close(capi.calculate(str(bed), *a), want, label=label + " (.bed)")
    assert_ld_close(capi.calculate_packed(packed, *a), want, label=label + " (packed)")
