"""Same-box A/B of the END-TO-END call on BASELINE.json configs[4] (N=100,000 x 300,000 SNPs, dense map, 10 cM, 2 % missing):
host buffer -> nldsc_ld_calculate_packed -> host arrays on ONE GPU, once per environment variant given on the command line.
    python scripts/cfg4_e2e_ab.py wide=NLDSC_X160=0 x160=NLDSC_X160=1"""
import hashlib, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from nldsc_b200 import capi, synth

n_org, n_snp = 100_000, int(os.environ.get("NLDSC_BENCH_STRONG_SNPS", "300000"))
spec = synth.SynthSpec(n_org=n_org, n_snp=n_snp, missing=0.02, seed=5, density=1000.0)
dev = synth.generate_packed_cuda(spec, "cuda:0")
host = torch.empty(dev.shape, dtype=torch.uint8, pin_memory=True)
host.copy_(dev)
del dev
torch.cuda.empty_cache()
cm, _ = synth.genetic_map(spec)
a = (n_snp, n_org, cm, 10.0, 1e-4, 1e-5, 1.0 / n_snp)
capi.report_device_times(True)
out = capi.alloc_result(n_snp)
res = {}
for arg in sys.argv[1:] or ["default"]:
    name, _, envs = arg.partition("=")
    kv = dict(x.split("=", 1) for x in envs.split(",") if x)
    os.environ.update(kv)
    capi.calculate_packed(host, *a, out=out)
    ts = []
    for _ in range(2):
        t = time.perf_counter()
        capi.calculate_packed(host, *a, out=out)
        ts.append(time.perf_counter() - t)
    r = capi.last_call_report()[0]
    sha = hashlib.sha256(b"".join(np.ascontiguousarray(out[0][k]).tobytes() for k in sorted(out[0]))).hexdigest()[:12]
    res[name] = {"ms": [round(1e3 * x, 1) for x in ts], "ms_dev_run": r["ms_dev_run"], "ms_dev_band": r["ms_dev_band"],
                 "ms_dev_unpack": r["ms_dev_unpack"], "ms_dev_upload": r["ms_dev_upload"], "chunks": r["chunks"], "sha": sha}
    print(name, res[name], flush=True)
    for k in kv:
        os.environ.pop(k)
    capi.arena_trim(0)
os.makedirs("gpurun_out", exist_ok=True)
json.dump(res, open(os.path.join("gpurun_out", os.environ.get("AB_OUT", "cfg4_e2e_ab.json")), "w"), indent=1)
