"""CPU: the NumPy restatement (oracle/ld_oracle.py) against the golden vectors produced by the
reference's own compiled source (tests/golden/make_golden.py), and -- when oracle/_ref is
built -- against that compiled reference live."""
import os
import tempfile

import numpy as np
import pytest

from conftest import assert_ld_close
from nldsc_b200 import synth
from oracle import ld_oracle, ref


def _ref64(g):
    return {k: g["ref64_" + k] for k in ("l2", "l2d", "maf", "residuals_std", "l2_ws", "l2d_ws", "l2d_wse")}


def test_restatement_matches_reference_golden(golden):
    g = golden
    o = ld_oracle.ld_scores(g["packed"], g["n_org"], g["positions"], g["ld_wind"], g["maf"], g["std_thr"],
                            g["rsq_thr"])
    assert_ld_close(o, _ref64(g), rtol=1e-12, label=g["name"])


def test_float_reference_agrees_on_filters(golden):
    """oracle-32 (stock float source) and oracle-64 take the same filter decisions."""
    g = golden
    for k in ("maf", "l2_ws", "l2d_ws"):
        assert np.array_equal(g["ref32_" + k], g["ref64_" + k], equal_nan=True)
    k = ~np.isnan(g["ref64_l2"])
    assert np.max(np.abs(g["ref32_l2"][k] - g["ref64_l2"][k]) / g["ref64_l2"][k]) < 1e-3


def test_decode_is_msb_first_with_high_pair_tail():
    """stream.h:55-67 / SURVEY F4: sample 4j+i <- bit pair (7-2i, 6-2i); the last byte gives its HIGH pairs."""
    packed = np.array([[0b11100100, 0b00011011]], dtype=np.uint8)
    assert ld_oracle.decode_bed_msb_first(packed, 8).tolist() == [[3, 2, 1, 0, 0, 1, 2, 3]]
    assert ld_oracle.decode_bed_msb_first(packed, 6).tolist() == [[3, 2, 1, 0, 0, 1]]


@pytest.mark.skipif(not ref.available(64), reason="oracle/_ref not built (needs /root/reference)")
def test_restatement_matches_live_reference():
    spec = synth.SynthSpec(n_org=515, n_snp=400, missing=0.03, seed=21)
    packed = synth.generate_packed(spec, workers=1)
    with tempfile.TemporaryDirectory() as d:
        info = synth.write_bfile(os.path.join(d, "t"), spec, packed)
        r = ref.calculate(os.path.join(d, "t.bed"), 400, 515, info["cm"], 0.7, 1e-4, 1e-5, 1 / 400, flavour=64)
    o = ld_oracle.ld_scores(packed, 515, info["cm"], 0.7, 1e-4, 1e-5, 1 / 400)
    assert_ld_close(o, r, rtol=1e-12, label="live")


def test_cross_count_closed_form_matches_restatement(golden):
    """The exact-integer form the GPU path uses: six cross-counts + per-SNP
    counts reproduce the restatement's correlations."""
    g = golden
    n = g["n_org"]
    codes = ld_oracle.decode_bed_msb_first(g["packed"], n)
    st = ld_oracle.snp_stats(codes, g["maf"])
    keep = np.nonzero(st["keep"])[0][:40]
    if len(keep) < 2:
        pytest.skip("too few kept SNPs")
    G = ld_oracle.cross_counts(codes[keep], codes[keep]).astype(np.float64)
    nh, na, nm = (st[k][keep] for k in ("n_h", "n_a", "n_m"))
    mu = ((nh + 2 * na) / (n - nm)).astype(np.float32).astype(np.float64)
    xbar = (nh + 2 * na + mu * nm) / n
    sd = np.sqrt((nh + 4 * na + mu * mu * nm) / n - xbar * xbar)
    Ggg = G[0, 1] + 2 * G[0, 2]
    Gmg = G[1, 1] + 2 * G[1, 2]
    sxx = Ggg + mu[None, :] * G[0, 0] + mu[:, None] * (Gmg + mu[None, :] * G[1, 0])
    r = (sxx - n * xbar[:, None] * xbar[None, :]) / (n * sd[:, None] * sd[None, :])
    want = st["A"][keep] @ st["A"][keep].T / n
    assert np.max(np.abs(r - want)) < 1e-12


def test_row_sliced_oracle_equals_whole_file_oracle():
    """ld_scores_rows (what the full-size GPU tests and bench.py check against) must agree with ld_scores, itself
    pinned to the compiled reference above -- for any clustering of the rows and any block size."""
    from nldsc_b200 import synth
    spec = synth.SynthSpec(n_org=701, n_snp=2000, missing=0.03, seed=14)
    packed = synth.generate_packed(spec, workers=1)
    cm, _ = synth.genetic_map(spec)
    full = ld_oracle.ld_scores(packed, 701, cm, 0.8, 1e-4, 1e-5, 1 / 2000)
    rows = [0, 3, 700, 701, 702, 1500, 1999] + list(np.nonzero(~full["keep"])[0][:2])
    for max_span, block in ((1, 64), (900, 257), (4096, 1024)):
        got = ld_oracle.ld_scores_rows(packed, 701, cm, 0.8, 1e-4, 1e-5, 1 / 2000, rows, max_span=max_span, block=block)
        r = got["rows"]
        assert np.array_equal(got["keep"], full["keep"][r])
        for k in ("l2_ws", "l2d_ws", "l2d_wse"):
            assert np.array_equal(got[k], full[k][r]), k
        assert np.array_equal(got["maf"], full["maf"][r], equal_nan=True)
        for k in ("l2", "l2d", "residuals_std"):
            assert np.allclose(got[k], full[k][r], rtol=1e-12, atol=1e-12, equal_nan=True), k
