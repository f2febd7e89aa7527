"""e2e phase split of nldsc_ld_calculate_packed on the bench workload (NLDSC_TRACE)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["NLDSC_TRACE"] = "1"
import bench
from nldsc_b200 import capi
host, cm, spec = bench.make_workload(0, "cuda:0")
out = capi.alloc_result(bench.N_SNP)
for i in range(5):
    t0 = time.perf_counter()
    capi.calculate_packed(host, bench.N_SNP, bench.N_ORG, cm, bench.WIND_CM, bench.MAF, bench.STD_THR, 1.0 / bench.N_SNP, out=out)
    print(f"call {i}: {1e3 * (time.perf_counter() - t0):.2f} ms", flush=True)
