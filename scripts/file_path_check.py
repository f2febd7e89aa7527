"""nldsc_ld_calculate (reads the .bed itself) against nldsc_ld_calculate_packed (pinned host buffer) on one
chromosome of configs[2] shape: how much of the wall time is the file read."""
import json, os, sys, tempfile, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from nldsc_b200 import capi, synth
N, M = 50_000, int(sys.argv[1]) if len(sys.argv) > 1 else 100_000
spec = synth.SynthSpec(n_org=N, n_snp=M, missing=0.02, seed=31)
cm, _ = synth.genetic_map(spec)
dev = synth.generate_packed_cuda(spec)
host = torch.empty(dev.shape, dtype=torch.uint8, pin_memory=True); host.copy_(dev); torch.cuda.synchronize(); del dev
d = tempfile.mkdtemp(prefix="fp_")
with open(f"{d}/x.bed", "wb") as f:
    f.write(synth.BED_MAGIC); f.write(host.numpy().tobytes())
args = (M, N, cm, 1.0, 1e-4, 1e-5, 1.0 / M)
out = {}
for name, fn in (("packed_pinned", lambda: capi.calculate_packed(host, *args)), ("bed_file", lambda: capi.calculate(f"{d}/x.bed", *args))):
    ts = []
    for _ in range(4):
        t0 = time.perf_counter(); r = fn(); ts.append(time.perf_counter() - t0)
    out[name] = {"s": ts, "gb_per_s": host.numel() / min(ts) / 1e9}
    if name == "packed_pinned":
        ref = r
    else:
        out[name]["bitwise_equal_to_packed"] = bool(all(np.array_equal(r[k], ref[k], equal_nan=True) for k in ref))
    print(name, out[name], flush=True)
os.environ["NLDSC_TRACE"] = "1"
capi.calculate(f"{d}/x.bed", *args)
os.makedirs("gpurun_out", exist_ok=True)
json.dump(out, open("gpurun_out/file_path.json", "w"), indent=1)
os.remove(f"{d}/x.bed")
