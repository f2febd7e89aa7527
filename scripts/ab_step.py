"""Same-box A/B of the resident step on BASELINE.json configs[1]: every variant (environment knobs, or an A/B build of the
library through NLDSC_LIB) runs in its own subprocess on the same generated bfile; prints ms per step (CUDA events, median
of 30) and the band / unpack split.   usage: ab_step.py name=ENV1=v,ENV2=v ...   (name alone = defaults)"""
import json, os, subprocess, sys, tempfile
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CHILD = r'''
import json, os, sys
sys.path.insert(0, %r)
import numpy as np
from nldsc_b200 import capi, synth
n_org, n_snp = int(os.environ.get("AB_N", 10000)), int(os.environ.get("AB_M", 100000))
packed = np.load(sys.argv[1], mmap_mode="r")
spec = synth.SynthSpec(n_org=n_org, n_snp=n_snp, missing=float(os.environ.get("AB_MISSING", 0.02)), seed=2, density=float(os.environ.get("AB_DENSITY", 340)))
cm, _ = synth.genetic_map(spec)
s = capi.Session(0, n_snp, n_org, cm, float(os.environ.get("AB_WIND", 1.0)), 1e-4, 1e-5, 1.0 / n_snp)
s.upload(np.ascontiguousarray(packed)); s.sync()
for _ in range(5): s.run()
s.sync()
tot, band, unp = [], [], []
for _ in range(int(os.environ.get("AB_REPS", 30))):
    s.run(); s.sync(); st = s.stats()
    tot.append(st["ms_total"]); band.append(st["ms_band"]); unp.append(st["ms_unpack"])
r = s.download()
import hashlib
h = hashlib.sha256(b"".join(np.ascontiguousarray(r[k]).tobytes() for k in sorted(r))).hexdigest()[:12]
print(json.dumps({"ms_step": float(np.median(tot)), "ms_step_min": float(np.min(tot)), "ms_band": float(np.median(band)),
                  "ms_unpack": float(np.median(unp)), "launches": st["kernel_launches"], "chunks": st["chunks"],
                  "mfree_chunks": st["mfree_chunks"], "pairs": int(r["l2_ws"][r["l2_ws"] >= 0].astype(np.int64).sum()), "sha": h}))
''' % ROOT
sys.path.insert(0, ROOT)
import numpy as np
import torch
from nldsc_b200 import synth
n_org, n_snp = int(os.environ.get("AB_N", 10000)), int(os.environ.get("AB_M", 100000))
spec = synth.SynthSpec(n_org=n_org, n_snp=n_snp, missing=float(os.environ.get("AB_MISSING", 0.02)), seed=2, density=float(os.environ.get("AB_DENSITY", 340)))
path = os.path.join(tempfile.gettempdir(), f"ab_{n_org}_{n_snp}.npy")
np.save(path, synth.generate_packed_cuda(spec, "cuda:0").cpu().numpy())
out = {}
for arg in sys.argv[1:] or ["default"]:
    name, _, envs = arg.partition("=")
    env = dict(os.environ)
    for kv in filter(None, envs.split(",")):
        k, _, v = kv.partition("=")
        env[k] = v
    r = subprocess.run([sys.executable, "-c", CHILD, path], env=env, capture_output=True, text=True)
    try:
        out[name] = json.loads(r.stdout.strip().splitlines()[-1])
    except Exception:
        out[name] = {"error": (r.stderr or r.stdout)[-400:]}
    print(name, out[name], flush=True)
os.remove(path)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", os.environ.get("AB_OUT", "ab_step.json")), "w"), indent=1)
