"""compute-sanitizer target: the smallest golden cases through the C ABI under the execution variants
that exercise the hand-rolled protocols -- wide kernel (default), narrow kernel, a wrapping plane ring with tiny chunks
(cross-stream ring hazards), the .bed reader threads, and 3 shards on one device -- each checked against the golden
vectors of the compiled reference.   usage: compute-sanitizer --tool <memcheck|racecheck|synccheck> python scripts/sanitize.py"""
import os
import sys
import tempfile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np

from conftest import RESULT_KEYS, assert_ld_close, load_golden
from nldsc_b200 import capi, synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
VARIANTS = [{}, {"NLDSC_WIDE": "0"}, {"NLDSC_CHUNK_ROWS": "128", "NLDSC_PLANE_BYTES": "ring"},
            {"NLDSC_TEST_SHARDS": "3"}, {"NLDSC_OVERLAP": "0"}, {"NLDSC_MFREE": "0"}, {"NLDSC_FUSED": "1"}]
for name in ("n1001_m300_miss", "n1003_m600_miss", "n1000_m768_nomiss"):
    g = load_golden(os.path.join(ROOT, "tests", "golden", name + ".npz"))
    a = (g["packed"].shape[0], g["n_org"], g["positions"], g["ld_wind"], g["maf"], g["std_thr"], g["rsq_thr"])
    want = {k: g["ref64_" + k] for k in RESULT_KEYS}
    for v in VARIANTS:
        for k in list(os.environ):
            if k.startswith("NLDSC_") and k not in ("NLDSC_DEVICE",):
                os.environ.pop(k)
        v = dict(v)
        if v.get("NLDSC_PLANE_BYTES") == "ring":             # a ring barely wider than one block pair's columns: rows wrap
            lo, hi = capi.plan_windows(g["positions"], g["ld_wind"])
            rows = ((int(np.max(hi - lo)) + 1) // 64 + 6) * 64
            v["NLDSC_PLANE_BYTES"] = str(rows * 4 * ((g["n_org"] + 255) // 256 * 128))   # fp4 planes: two organisms per byte
        os.environ.update(v)
        got = capi.calculate_packed(g["packed"], *a)
        assert_ld_close(got, want, label=f"{name} {v}")
        print("ok", name, v, flush=True)
    with tempfile.TemporaryDirectory() as d:
        with open(os.path.join(d, "s.bed"), "wb") as f:
            f.write(synth.BED_MAGIC)
            f.write(g["packed"].tobytes())
        got = capi.calculate(os.path.join(d, "s.bed"), *a)
        assert_ld_close(got, want, label=name + " .bed")
        print("ok", name, ".bed", flush=True)
print("sanitize targets done")
