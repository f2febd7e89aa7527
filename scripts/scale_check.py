"""Large-shape checks on one B200 (BASELINE.json configs[3] and configs[4] shapes, SNP count cut to fit a
few seconds): end-to-end time through the C ABI from a host buffer, size-independent properties, and the
CPU oracle on a few random rows with their full windows.  Writes gpurun_out/scale_check.json."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from nldsc_b200 import capi, synth
from oracle import ld_oracle

CASES = {
    # BASELINE.json configs[0]: the reference's own CPU-runnable case
    "cfg1":       dict(n_org=1000, n_snp=10000, missing=0.0, density=340.0, wind=1.0, oracle_rows=3),
    # reference-panel sizes (1000 Genomes: 2,504 samples) and a mid-size cohort: which band kernel wins where
    "kg_panel":   dict(n_org=2504, n_snp=100000, missing=0.0, density=340.0, wind=1.0, oracle_rows=0),
    "mid_cohort": dict(n_org=5000, n_snp=100000, missing=0.02, density=340.0, wind=1.0, oracle_rows=0),
    "ukb_slice":  dict(n_org=315599, n_snp=40000, missing=0.02, density=340.0, wind=1.0, oracle_rows=2),
    "ukb_ring":   dict(n_org=315599, n_snp=40000, missing=0.02, density=340.0, wind=1.0, oracle_rows=0,
                       env={"NLDSC_PLANE_BYTES": "12e9"}),
    # BASELINE.json configs[4] at full size: 7.5 GB .bed, interior window ~20,000 SNPs, ~5.8e9 ordered pairs
    "cfg5_full":  dict(n_org=100000, n_snp=300000, missing=0.02, density=1000.0, wind=10.0, oracle_rows=1),
    "cfg5_slice": dict(n_org=100000, n_snp=49152, missing=0.02, density=1000.0, wind=10.0, oracle_rows=0),
    "wide_band":  dict(n_org=100000, n_snp=40000, missing=0.02, density=200.0, wind=10.0, oracle_rows=1),
}
out = {}
kept_results = {}
for name in (sys.argv[1:] or [k for k in CASES if k not in ("cfg5_full", "cfg5_slice", "cfg1", "kg_panel", "mid_cohort")]):
    c = CASES[name]
    for k, v in c.get("env", {}).items():
        os.environ[k] = v
    spec = synth.SynthSpec(n_org=c["n_org"], n_snp=c["n_snp"], missing=c["missing"], seed=7, density=c["density"])
    cm, _ = synth.genetic_map(spec)
    t0 = time.perf_counter()
    dev = synth.generate_packed_cuda(spec)
    host = torch.empty(dev.shape, dtype=torch.uint8, pin_memory=True)
    host.copy_(dev); torch.cuda.synchronize(); del dev; torch.cuda.empty_cache()
    gen_s = time.perf_counter() - t0
    args = (c["n_snp"], c["n_org"], cm, c["wind"], 1e-4, 1e-5, 1.0 / c["n_snp"])
    res = capi.alloc_result(c["n_snp"])
    times = []
    for rep in range(3):
        t0 = time.perf_counter()
        got = capi.calculate_packed(host, *args, out=res)
        times.append(time.perf_counter() - t0)
    keep = got["l2_ws"] >= 0
    pairs = int(got["l2_ws"][keep].sum())
    lo, hi = capi.plan_windows(cm, c["wind"])
    ck = np.concatenate([[0], np.cumsum(keep)])
    ok_wsa = bool(np.array_equal(got["l2_ws"][keep], (ck[hi + 1] - ck[lo])[keep] - 1))
    ok_order = bool(np.all(got["l2d_ws"][keep] <= got["l2_ws"][keep]) and np.all(got["l2d_wse"][keep] <= got["l2d_ws"][keep]))
    ok_floor = bool(np.all(got["l2"][keep] >= 1.0 - got["l2_ws"][keep] / (c["n_org"] - 2) - 1e-9))
    rec = {"n_org": c["n_org"], "n_snp": c["n_snp"], "wind_cm": c["wind"], "pairs": pairs,
           "mean_window": pairs / max(1, int(keep.sum())), "kept": int(keep.sum()), "gen_s": gen_s,
           "e2e_s": min(times), "e2e_all_s": times, "pairs_per_s": pairs / min(times),
           "int8_ops_per_s_algorithmic": pairs * 12.0 * c["n_org"] / min(times),
           "bed_bytes": int(host.numel()), "wsa_property": ok_wsa, "order_property": ok_order, "floor_property": ok_floor,
           "env": c.get("env", {})}
    # kernel-only: shard resident in HBM, per-family CUDA-event times of one run (as bench.py's `value` leg)
    sess = capi.Session(0, c["n_snp"], c["n_org"], cm, c["wind"], 1e-4, 1e-5, 1.0 / c["n_snp"])
    sess.upload(host); sess.sync()
    for _ in range(3):
        sess.run()
    sess.sync()
    st = sess.stats()
    sess.close()
    ops = pairs * 12.0 * c["n_org"]
    rec["resident"] = {"ms_step": st["ms_total"], "ms_band": st["ms_band"], "ms_unpack": st["ms_unpack"], "chunks": st["chunks"],
                       "tiles": st["tiles"], "band_algorithmic_top_s": ops / (st["ms_band"] * 1e-3) / 1e12,
                       "band_executed_top_s": st["tiles"] * 2 * 128 * 192 * 2.0 * st["npad"] / (st["ms_band"] * 1e-3) / 1e12,
                       "step_algorithmic_top_s": ops / (st["ms_total"] * 1e-3) / 1e12}
    kept_results[name] = {k: v.copy() for k, v in got.items()}
    if name == "ukb_ring" and "ukb_slice" in kept_results:      # a 12 GB ring that wraps vs all rows resident
        rec["bitwise_equal_to_ukb_slice"] = bool(all(np.array_equal(got[k], kept_results["ukb_slice"][k], equal_nan=True) for k in got))
    hn = host.numpy()
    rng = np.random.default_rng(1)
    errs = []
    avail_gb = int([l for l in open("/proc/meminfo") if l.startswith("MemAvailable")][0].split()[1]) / 1e6
    rec["host_mem_available_gb"] = avail_gb
    if avail_gb < 96:
        c = dict(c, oracle_rows=0)
    for i in rng.choice(np.nonzero(keep)[0], c["oracle_rows"], replace=False) if c["oracle_rows"] else []:
        a, b = int(lo[i]), int(hi[i]) + 1
        t0 = time.perf_counter()
        sub = ld_oracle.ld_scores(hn[a:b], c["n_org"], cm[a:b], c["wind"], 1e-4, 1e-5, 1.0 / c["n_snp"], rows=[i - a])
        errs.append({"row": int(i), "window": b - a, "oracle_s": time.perf_counter() - t0,
                     "l2_rel": float(abs(sub["l2"][i - a] - got["l2"][i]) / abs(sub["l2"][i - a])),
                     "l2d_abs": float(abs(sub["l2d"][i - a] - got["l2d"][i])),
                     "counts_equal": bool(sub["l2_ws"][i - a] == got["l2_ws"][i] and sub["l2d_ws"][i - a] == got["l2d_ws"][i]
                                          and sub["l2d_wse"][i - a] == got["l2d_wse"][i])})
    rec["oracle"] = errs
    out[name] = rec
    print(name, json.dumps(rec), flush=True)
    for k in c.get("env", {}):
        os.environ.pop(k, None)
    del host
os.makedirs("gpurun_out", exist_ok=True)
json.dump(out, open("gpurun_out/scale_check%s.json" % ("_" + "_".join(sys.argv[1:]) if sys.argv[1:] else ""), "w"), indent=1)
