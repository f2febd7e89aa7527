"""Development check of the 160-column supertiles (NLDSC_X160=1): every case must reproduce the default run bit for bit.
    python scripts/x160_check.py            # small and mid cases, prints one line per case, exits 1 on a mismatch"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

from nldsc_b200 import capi, synth

KEYS = ("l2", "l2d", "maf", "residuals_std", "l2_ws", "l2d_ws", "l2d_wse")
CASES = [  # n_org, n_snp, missing, wind, density, extra env
    (500, 1300, 0.02, 1.0, 340.0, {}),
    (130, 700, 0.05, 1.0, 340.0, {}),
    (1003, 3000, 0.03, 0.7, 340.0, {}),
    (50, 65, 0.10, 1.0, 340.0, {}),
    (517, 129, 0.02, 0.3, 340.0, {"NLDSC_CHUNK_ROWS": "64"}),
    (517, 961, 0.02, 0.3, 340.0, {"NLDSC_CHUNK_ROWS": "192"}),
    (2048, 7777, 0.02, 0.2, 2000.0, {"NLDSC_CHUNK_ROWS": "512"}),
    (1027, 12000, 0.02, 0.05, 6000.0, {"NLDSC_CHUNK_ROWS": "1024", "ring": "1"}),
    (3001, 7777, 0.02, 1.0, 340.0, {"NLDSC_TEST_SHARDS": "3"}),
    (5000, 20000, 0.02, 1.0, 340.0, {}),
    (1003, 3000, 0.0, 0.7, 340.0, {}),                      # missing-free: the x160 body must stay out of it
]


def run(packed, n_org, cm, wind, env):
    for k, v in env.items():
        os.environ[k] = v
    try:
        n = packed.shape[0]
        if "NLDSC_TEST_SHARDS" in env:
            return capi.calculate_packed(packed, n, n_org, cm, wind, 1e-4, 1e-5, 1.0 / n), None
        with capi.Session(0, n, n_org, cm, wind, 1e-4, 1e-5, 1.0 / n) as s:
            s.upload(packed)
            s.run()
            return s.download(), s.stats()
    finally:
        for k in env:
            os.environ.pop(k, None)


# the variant under test, as environment settings (X160_CHECK_ENV="NLDSC_X160=1,NLDSC_CHUNK_ROWS=256" ...)
FORCE = dict(kv.split("=", 1) for kv in os.environ.get("X160_CHECK_ENV", "NLDSC_X160=1").split(","))
bad = 0
for n_org, n_snp, missing, wind, density, extra in CASES:
    spec = synth.SynthSpec(n_org=n_org, n_snp=n_snp, missing=missing, seed=n_org + n_snp, density=density)
    packed = synth.generate_packed(spec, workers=4)
    cm, _ = synth.genetic_map(spec)
    env = dict(extra)
    if env.pop("ring", None):
        lo, hi = capi.plan_windows(cm, wind)
        rows = ((int(np.max(hi - lo)) + 1) // 64 + 6) * 64
        env["NLDSC_PLANE_BYTES"] = str(rows * 4 * ((n_org + 255) // 256 * 128))
    base, st0 = run(packed, n_org, cm, wind, dict(env, NLDSC_X160="0"))
    got, st1 = run(packed, n_org, cm, wind, dict(env, **FORCE))
    diffs = {k: int(np.sum(~((got[k] == base[k]) | (np.isnan(got[k].astype(float)) & np.isnan(base[k].astype(float)))))) for k in KEYS}
    ok = all(v == 0 for v in diffs.values())
    bad += 0 if ok else 1
    rel = float(np.nanmax(np.abs(got["l2"] - base["l2"]) / np.maximum(np.abs(base["l2"]), 1.0)))
    xt = (st1 or {}).get("xtiles"), (st1 or {}).get("xtiles_run"), (st0 or {}).get("tiles")
    print(f"N={n_org} M={n_snp} miss={missing} wind={wind} {extra}: {'OK' if ok else 'MISMATCH ' + str(diffs)}  l2 max rel {rel:.2e}  "
          f"xtiles/run/slots {xt}", flush=True)
sys.exit(1 if bad else 0)
