"""On-box ceilings of the band kernel's tile shape: MMA-only, TMA-only and the real kernel."""
import json, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nldsc_b200 import capi
out = {}
for n_org in (10000, 100000):
    tiles = 74 * (40 if n_org == 10000 else 6)
    for mode, name in ((1, "mma_only"), (2, "tma_only"), (0, "real"), (3, "epilogue_only"), (4, "real_full_epilogue"), (6, "wide_real")):
        r = capi.debug_band_probe(mode, n_org, tiles, reps=5)
        out[f"n{n_org}_{name}"] = r
        print(n_org, name, r, flush=True)
json.dump(out, open("gpurun_out/probe.json", "w"), indent=1)
# epilogue cost per supertile: one K-step (n_org = 128) so the MMA / TMA time is negligible
for mode, name in ((0, "k1_cheap_epilogue"), (3, "k1_epilogue_only"), (4, "k1_full"), (6, "k1_wide_cheap_epilogue")):
    r = capi.debug_band_probe(mode, 128, 74 * 200, reps=5)
    r["us_per_supertile"] = r["ms"] * 1e3 / 200
    out[name] = r
    print(name, r, flush=True)
json.dump(out, open("gpurun_out/probe.json", "w"), indent=1)
