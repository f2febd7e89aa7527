"""Deterministic synthetic PLINK bfiles for the `ld` path.

The reference ships no sample data, so parity tests and the
benchmark generate their own.  The model follows SURVEY.md section 8(d):

* per-SNP second-allele frequency p_j ~ U(0.01, 0.5); ``mono_frac`` of the SNPs
  monomorphic (p = 0) and ``rare_frac`` rare (p = 8e-4) to exercise the MAF and
  the collinear-dominance filters (encoder.h:120-122, ldscalc.h:40);
* two haplotypes per organism from an AR(1) latent Gaussian along the SNP axis
  (rho = 0.95), restarted every ``CHUNK`` SNPs so any chunk can be regenerated
  independently (and in parallel) from ``(seed, chunk)``;
* missing genotypes (code 01) i.i.d. at ``missing``;
* map: cM = cumsum Exp(mean 1/density); BP = floor(cM * 1e6) + 1.

The .bed written is PLINK-spec (LSB-first bit pairs, SNP-major, magic
6c 1b 01).  The reference decodes it MSB-first (stream.h:55-67) --
that is the reference's business; the file on disk is a normal bfile.
"""
from __future__ import annotations

import os
from concurrent.futures import ProcessPoolExecutor
from dataclasses import dataclass

import numpy as np

CHUNK = 4096
BED_MAGIC = bytes([0x6C, 0x1B, 0x01])


@dataclass
class SynthSpec:
    n_org: int
    n_snp: int
    missing: float = 0.0
    seed: int = 1
    density: float = 340.0      # SNPs per cM
    mono_frac: float = 0.03
    rare_frac: float = 0.03
    rho: float = 0.95
    chrom: int = 22


def _norm_ppf(p: np.ndarray) -> np.ndarray:
    from scipy.special import ndtri
    return ndtri(p)


def snp_frequencies(spec: SynthSpec) -> np.ndarray:
    """Second-allele frequency of every SNP (depends on seed only)."""
    rng = np.random.default_rng([spec.seed, 0xF0E])
    p = rng.uniform(0.01, 0.5, spec.n_snp)
    u = rng.uniform(0.0, 1.0, spec.n_snp)
    p[u < spec.mono_frac] = 0.0
    p[(u >= spec.mono_frac) & (u < spec.mono_frac + spec.rare_frac)] = 8e-4
    return p


def genetic_map(spec: SynthSpec) -> tuple[np.ndarray, np.ndarray]:
    rng = np.random.default_rng([spec.seed, 0x3A9])
    cm = np.cumsum(rng.exponential(1.0 / spec.density, spec.n_snp))
    bp = np.floor(cm * 1e6).astype(np.int64) + 1
    return cm, bp


def _chunk_codes(spec: SynthSpec, chunk: int, p: np.ndarray) -> np.ndarray:
    """2-bit PLINK codes (0 hom-first, 1 missing, 2 het, 3 hom-second), shape
    [snps_in_chunk, n_org] uint8."""
    from scipy.signal import lfilter
    s0 = chunk * CHUNK
    s1 = min(spec.n_snp, s0 + CHUNK)
    m = s1 - s0
    rng = np.random.default_rng([spec.seed, 0xC0DE, chunk])
    eps = rng.standard_normal((m, 2 * spec.n_org), dtype=np.float32)
    a = np.float32(np.sqrt(1.0 - spec.rho ** 2))
    eps[1:] *= a
    z = lfilter([1.0], [1.0, -spec.rho], eps, axis=0).astype(np.float32)
    thr = np.full(m, -np.inf, dtype=np.float32)
    pc = p[s0:s1]
    nz = pc > 0
    thr[nz] = _norm_ppf(pc[nz]).astype(np.float32)
    allele = z < thr[:, None]
    g = allele[:, : spec.n_org].astype(np.uint8) + allele[:, spec.n_org:].astype(np.uint8)
    codes = np.where(g == 0, 0, np.where(g == 1, 2, 3)).astype(np.uint8)
    if spec.missing > 0:
        miss = rng.random((m, spec.n_org), dtype=np.float32) < spec.missing
        # never emit an all-missing SNP
        miss[:, 0] = False
        codes[miss] = 1
    return codes


def pack_codes(codes: np.ndarray) -> np.ndarray:
    """[m, n_org] codes -> [m, ceil(n_org/4)] PLINK bytes (sample 4k+i in bits 2i..2i+1)."""
    m, n = codes.shape
    bps = (n + 3) // 4
    pad = np.zeros((m, bps * 4), dtype=np.uint8)
    pad[:, :n] = codes
    q = pad.reshape(m, bps, 4)
    return (q[:, :, 0] | (q[:, :, 1] << 2) | (q[:, :, 2] << 4) | (q[:, :, 3] << 6)).astype(np.uint8)


def _chunk_packed(args) -> np.ndarray:
    spec, chunk, p = args
    return pack_codes(_chunk_codes(spec, chunk, p))


def generate_packed(spec: SynthSpec, snp_range: tuple[int, int] | None = None,
                    workers: int | None = None) -> np.ndarray:
    """Packed genotype bytes [n_snp (or range), ceil(n_org/4)] without the 3-byte magic."""
    p = snp_frequencies(spec)
    lo, hi = (0, spec.n_snp) if snp_range is None else snp_range
    c0, c1 = lo // CHUNK, (hi + CHUNK - 1) // CHUNK
    jobs = [(spec, c, p) for c in range(c0, c1)]
    if workers is None:
        workers = min(len(jobs), max(1, (os.cpu_count() or 2) // 2), 32)
    if workers > 1 and len(jobs) > 1:
        with ProcessPoolExecutor(max_workers=workers) as ex:
            parts = list(ex.map(_chunk_packed, jobs))
    else:
        parts = [_chunk_packed(j) for j in jobs]
    out = np.concatenate(parts, axis=0) if parts else np.zeros((0, (spec.n_org + 3) // 4), np.uint8)
    return np.ascontiguousarray(out[lo - c0 * CHUNK: hi - c0 * CHUNK])


def write_bfile(prefix: str, spec: SynthSpec, packed: np.ndarray | None = None,
                snp_range: tuple[int, int] | None = None) -> dict:
    """Write prefix.bed/.bim/.fam (tab-separated, one chromosome).  Returns the
    map arrays used, for callers that bypass the .bim parser."""
    lo, hi = (0, spec.n_snp) if snp_range is None else snp_range
    if packed is None:
        packed = generate_packed(spec, snp_range)
    cm, bp = genetic_map(spec)
    cm, bp = cm[lo:hi], bp[lo:hi]
    with open(prefix + ".bed", "wb") as f:
        f.write(BED_MAGIC)
        f.write(packed.tobytes())
    with open(prefix + ".bim", "w") as f:
        for k in range(hi - lo):
            f.write(f"{spec.chrom}\trs{lo + k}\t{cm[k]:.8f}\t{bp[k]}\tA\tC\n")
    with open(prefix + ".fam", "w") as f:
        for i in range(spec.n_org):
            f.write(f"F{i}\tI{i}\t0\t0\t0\t-9\n")
    # the .bim stores cM with 8 decimals: hand back exactly what a parser will read
    cm_rt = np.array([float(f"{v:.8f}") for v in cm])
    return {"cm": cm_rt, "bp": bp, "n_snp": hi - lo, "n_org": spec.n_org}


def generate_packed_cuda(spec: SynthSpec, device="cuda:0"):
    """Same genotype model as generate_packed, drawn on the GPU with torch (different random
    stream, so NOT byte-identical to the NumPy generator): the benchmark's workload at
    BASELINE.json sizes takes seconds instead of minutes of host time.  torch is plumbing here
    (RNG + elementwise ops); nothing of the LD-score path runs through it.
    Returns a uint8 CUDA tensor [n_snp, ceil(n_org/4)] in PLINK (LSB-first) packing."""
    import torch
    dev = torch.device(device)
    n, m = spec.n_org, spec.n_snp
    n_chunks = (m + CHUNK - 1) // CHUNK
    p = snp_frequencies(spec)
    thr = np.full(n_chunks * CHUNK, -np.inf, dtype=np.float32)
    nz = p > 0
    thr[:m][nz] = _norm_ppf(p[nz]).astype(np.float32)
    thr_t = torch.from_numpy(thr).to(dev).view(n_chunks, CHUNK)
    gen = torch.Generator(device=dev)
    gen.manual_seed(int(spec.seed) * 1000003 + 17)
    bps = (n + 3) // 4
    codes = torch.zeros((n_chunks, CHUNK, bps * 4), dtype=torch.uint8, device=dev)
    a = float(np.sqrt(1.0 - spec.rho ** 2))
    z = torch.randn((n_chunks, 2 * n), generator=gen, device=dev)
    for t in range(CHUNK):
        if t:
            z.mul_(spec.rho).add_(torch.randn((n_chunks, 2 * n), generator=gen, device=dev), alpha=a)
        al = z < thr_t[:, t:t + 1]
        g = al[:, :n].to(torch.uint8) + al[:, n:].to(torch.uint8)         # 0 / 1 / 2 second alleles
        codes[:, t, :n] = g + (g > 0).to(torch.uint8)                      # 0 -> 00, 1 -> 10 (het), 2 -> 11
    codes = codes.view(n_chunks * CHUNK, bps * 4)[:m]
    if spec.missing > 0:
        for r0 in range(0, m, 8192):
            blk = codes[r0:r0 + 8192]
            miss = torch.rand(blk.shape, generator=gen, device=dev) < spec.missing
            miss[:, 0] = False                                             # never an all-missing SNP
            miss[:, n:] = False
            blk[miss] = 1
    q = codes.view(m, bps, 4)
    return (q[:, :, 0] | (q[:, :, 1] << 2) | (q[:, :, 2] << 4) | (q[:, :, 3] << 6)).contiguous()
