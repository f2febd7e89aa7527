"""BASELINE.json configs[2]: N=50,000 x 5 chromosome bfiles of 100,000 SNPs, 2% missing, 1 cM window, scored by the
genome driver (one process, one host thread per GPU, .bed files read from disk) on the GPUs of this box.
Writes gpurun_out/cfg3_genome_<gpus>gpu.json.   usage: cfg3_genome.py [gpus[,gpus...]] [n_snp_per_chromosome]
(several GPU counts: the bfiles are generated once and scored with each count in turn)
With NLDSC_GENOME=ukb: configs[3] shape instead -- N=315,599, 22 chromosomes with SNP counts proportional to their genetic
lengths, `n_snp_per_chromosome` x 22 SNPs in total (the full 1.2M-SNP genome is 94.7 GB; this is a cut that fits a scratch disk)."""
import dataclasses, json, os, shutil, sys, tempfile, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import pandas as pd
import torch
from nldsc_b200 import ldscore, synth

gpu_counts = [int(x) for x in sys.argv[1].split(",")] if len(sys.argv) > 1 else [torch.cuda.device_count()]
n_snp = int(sys.argv[2]) if len(sys.argv) > 2 else 100_000
N = 50_000
CHROMS = list(range(1, 6))
SIZES = {c: n_snp for c in CHROMS}
TAG = "cfg3_genome"
if os.environ.get("NLDSC_GENOME") == "ukb":
    N = 315_599
    cm_len = [286, 269, 223, 215, 204, 192, 187, 168, 166, 181, 158, 175, 126, 120, 141, 134, 128, 118, 108, 108, 62, 73]
    CHROMS = list(range(1, 23))
    SIZES = {c: max(1024, int(n_snp * 22 * cm_len[c - 1] / sum(cm_len)) // 64 * 64) for c in CHROMS}
    TAG = "cfg4_genome"
d = tempfile.mkdtemp(prefix="cfg3_")
t0 = time.perf_counter()
for c in CHROMS:
    n_snp = SIZES[c]
    spec = dataclasses.replace(synth.SynthSpec(n_org=N, n_snp=n_snp, missing=0.02, seed=30 + c), chrom=c)
    packed = synth.generate_packed_cuda(spec).cpu().numpy()
    cm, bp = synth.genetic_map(spec)
    with open(f"{d}/g.{c}.bed", "wb") as f:
        f.write(synth.BED_MAGIC); f.write(packed.tobytes())
    pd.DataFrame({"c": c, "s": [f"rs{c}_{k}" for k in range(n_snp)], "cm": np.round(cm, 8), "bp": bp, "a": "A", "b": "C"}
                 ).to_csv(f"{d}/g.{c}.bim", sep="\t", header=False, index=False)
    with open(f"{d}/g.{c}.fam", "w") as f:
        f.writelines(f"F{i}\tI{i}\t0\t0\t0\t-9\n" for i in range(N))
    del packed
torch.cuda.empty_cache()
gen_s = time.perf_counter() - t0
import hashlib
for gpus in gpu_counts:
    runs, digests = [], []
    for rep in range(2):
        shutil.rmtree(f"{d}/out", ignore_errors=True)
        t0 = time.perf_counter()
        summ = ldscore.estimate_lds_genome(f"{d}/g.@", 1.0, "cm", 1e-4, 1e-5, out_dir=f"{d}/out", gpus=gpus)
        runs.append(time.perf_counter() - t0)
        digests.append(hashlib.sha256(b"".join(open(f"{d}/out/{c}.L2", "rb").read() for c in CHROMS)).hexdigest())
    pairs = 0
    for c in CHROMS:
        df = pd.read_csv(f"{d}/out/{c}.L2", sep="\t")
        pairs += int(df["WSA"][df["WSA"] >= 0].sum())
    total_snps = sum(SIZES.values())
    phases = dict(summ.attrs.get("phases", {}))
    busy = {str(k): v for k, v in summ.attrs.get("per_gpu_busy_s", {}).items()}
    rec = {"phases_last_run": phases, "per_gpu_busy_s_last_run": busy, "config": "%s: N=%d x %d chromosomes, %d SNPs in total, 2%% missing, 1 cM" % (TAG, N, len(CHROMS), total_snps), "gpus": gpus,
           "bed_bytes": total_snps * ((N + 3) // 4),
           "pairs": pairs, "wall_s": runs, "outputs_identical_across_runs": digests[0] == digests[1], "l2_sha256": digests[0], "pairs_per_s_wall": pairs / min(runs), "generate_s": gen_s,
           "per_chromosome": summ.drop(columns=["file"]).to_dict(orient="records"),
           "note": "wall time of estimate_lds_genome: .bim/.fam parsing, .bed read from the page cache, GPU scoring, %.5f .L2 + .M files written"}
    print(json.dumps(rec))
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(rec, open("gpurun_out/%s_%dgpu.json" % (TAG, gpus), "w"), indent=1)

shutil.rmtree(d, ignore_errors=True)
