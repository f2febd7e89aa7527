"""NLDSC_TRACE=1 timeline of the end-to-end call on BASELINE.json configs[1] (host buffer -> nldsc_ld_calculate_packed -> host arrays):
where the milliseconds beside the PCIe floor go.  usage: python scripts/e2e_trace.py"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from nldsc_b200 import capi, synth
spec = synth.SynthSpec(n_org=10000, n_snp=100000, missing=0.02, seed=2)
dev = synth.generate_packed_cuda(spec, "cuda:0")
host = torch.empty(dev.shape, dtype=torch.uint8, pin_memory=True); host.copy_(dev); del dev
cm, _ = synth.genetic_map(spec)
a = (100000, 10000, cm, 1.0, 1e-4, 1e-5, 1e-5)
out = capi.alloc_result(100000)
for _ in range(3):
    capi.calculate_packed(host, *a, out=out)
ts = []
for _ in range(20):
    t = time.perf_counter(); capi.calculate_packed(host, *a, out=out); ts.append(time.perf_counter() - t)
print("e2e ms median %.3f min %.3f" % (1e3 * float(np.median(ts)), 1e3 * min(ts)))
capi.report_device_times(True); capi.calculate_packed(host, *a, out=out); capi.report_device_times(False)
print(capi.last_call_report()[0])
os.environ["NLDSC_TRACE"] = "1"
capi.calculate_packed(host, *a, out=out)
