"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference's `ld` hot path.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import
this module; the product path (nldsc_b200/) never does.

"restatement-64" of SURVEY.md section 8(c): an independent NumPy FP64 transcription of
the reference algorithm.  Citations are relative to
/root/reference/nldsc/ldscore/_ldscore/.

Parity pin: the reference has NO tests, golden vectors or fixtures (SURVEY.md
section 4), so this restatement is pinned against the reference's own source compiled
here (oracle/Makefile -> oracle/_ref/libref_ld64.so, the FP64-promoted build,
and libref_ld32.so, the stock float build) by tests/test_oracle.py, and against
tests/golden/*.npz produced from those builds by tests/golden/make_golden.py.

Numeric contract: everything is FP64 EXCEPT the two things that
decide filters and imputation, which stay in the reference's float32:
  * the imputation means  static_cast<float>(sum / n_obs)       encoder.h:114-115
  * the MAF path           f2 = mean/2; maf = f2<0.5 ? f2 : 1-f2  encoder.h:117-118
"""
from __future__ import annotations

import numpy as np


def decode_bed_msb_first(packed: np.ndarray, n_org: int) -> np.ndarray:
    """stream.h:43-69.  Byte j supplies samples 4j+0..4j+3 from bit pairs
    (7,6),(5,4),(3,2),(1,0) -- MSB first, the reverse of the PLINK spec -- and the
    last byte supplies only the first n_org%4 (high) pairs.
    Returns codes [n_snp, n_org] in {0,1,2,3} = {HOM_FST, MISS, HET, HOM_SND}
    (encoder.h:11-16 uses the same pairs shifted left by 6)."""
    m, bps = packed.shape
    assert bps == (n_org + 3) // 4
    out = np.empty((m, bps, 4), dtype=np.uint8)
    for i in range(4):
        out[:, :, i] = (packed >> (6 - 2 * i)) & 3
    return out.reshape(m, bps * 4)[:, :n_org]


def snp_stats(codes: np.ndarray, maf_thr: float):
    """encoder.h:91-133 + tools.h:54-85 for every SNP.  Returns dict of per-SNP
    arrays; A and D are the standardised additive vector and the standardised
    dominance residual (only rows with keep=True are meaningful)."""
    m, n = codes.shape
    het = codes == 2
    hom = codes == 3
    mis = codes == 1
    n_h = het.sum(1).astype(np.float64)
    n_a = hom.sum(1).astype(np.float64)
    n_m = mis.sum(1).astype(np.float64)
    n_obs = n - n_m
    with np.errstate(all="ignore"):
        add_sum = n_h + 2.0 * n_a                       # encoder.h:104-106
        nadd_sum = 2.0 * (n_h + n_a)
        mu_a = (add_sum / n_obs).astype(np.float32)     # encoder.h:114
        mu_d = (nadd_sum / n_obs).astype(np.float32)    # encoder.h:115
        f2 = mu_a / np.float32(2)                       # encoder.h:117 (float)
        maf = np.where(f2 < 0.5, f2, np.float32(1) - f2).astype(np.float32)  # :118
        keep = ~(maf.astype(np.float64) <= maf_thr)     # encoder.h:120 (NaN passes, F9)

    mu_a64 = mu_a.astype(np.float64)
    mu_d64 = mu_d.astype(np.float64)
    A = np.full((m, n), np.nan)
    D = np.full((m, n), np.nan)
    rstd = np.full(m, np.nan)
    for s in np.nonzero(keep)[0]:
        x = het[s] * 1.0 + hom[s] * 2.0
        y = (het[s] | hom[s]) * 2.0                     # "dominant" coding encoder.h:38-40
        x[mis[s]] = mu_a64[s]                           # encoder.h:124-129
        y[mis[s]] = mu_d64[s]
        with np.errstate(all="ignore"):
            xm, ym = x.mean(), y.mean()                 # tools.h:62-63
            slope = (x @ y / n - xm * ym) / (x @ x / n - xm * xm)   # tools.h:65-66
            res = y - slope * x                         # tools.h:67
            xc = x - x.mean()                           # tools.h:70-85 (population variance)
            A[s] = xc / np.sqrt(xc @ xc / n)
            rc = res - res.mean()
            sd = np.sqrt(rc @ rc / n)
            rstd[s] = sd
            D[s] = rc / sd
    return {"maf": maf.astype(np.float64), "keep": keep, "A": A, "D": D, "rstd": rstd,
            "n_h": n_h, "n_a": n_a, "n_m": n_m}


def ld_scores(packed: np.ndarray, n_org: int, positions: np.ndarray, ld_wind: float,
              maf_thr: float, std_thr: float, rsq_thr: float, rows=None) -> dict:
    """calculate() with the reference's "unused SNP" rule: a SNP with position < 0 is skipped in the .bed
    (stream.h:189-194 `pass`), is in nobody's window (tools.h:16-24,41-49) and keeps the initial outputs NaN / -1
    (ldscalc.h:13-21,24,57), MAF included (never decoded: encoder.h:47).  The used SNPs are scored among themselves
    by _ld_scores_used.  NB: this is the rule as the reference's filter states it; the reference's own sliding cache
    stops reading ahead at an unused SNP (stream.h:182-197) and indexes past its cache when the file starts with
    one, so its results with negative positions are not well defined and are not a parity target."""
    positions = np.asarray(positions, dtype=np.float64)
    used = positions >= 0
    if used.all():
        return _ld_scores_used(packed, n_org, positions, ld_wind, maf_thr, std_thr, rsq_thr, rows)
    idx = np.nonzero(used)[0]
    sub_rows = None
    if rows is not None:
        where = {int(g): k for k, g in enumerate(idx)}
        sub_rows = [where[int(r)] for r in rows if int(r) in where]
    sub = _ld_scores_used(packed[idx], n_org, positions[idx], ld_wind, maf_thr, std_thr, rsq_thr, sub_rows)
    m = len(positions)
    out = {"l2": np.full(m, np.nan), "l2d": np.full(m, np.nan), "maf": np.full(m, np.nan),
           "residuals_std": np.full(m, np.nan), "l2_ws": np.full(m, -1, np.int32), "l2d_ws": np.full(m, -1, np.int32),
           "l2d_wse": np.full(m, -1, np.int32), "keep": np.zeros(m, bool), "counts": np.zeros((m, 3), np.int64)}
    for k in out:
        out[k][idx] = sub[k]
    return out


def _ld_scores_used(packed: np.ndarray, n_org: int, positions: np.ndarray, ld_wind: float,
                    maf_thr: float, std_thr: float, rsq_thr: float, rows=None) -> dict:
    """calculate(), ldscalc.h:8-65, for sorted non-negative positions (the window
    of SNP i is {j != i : kept_j and |pos_j - pos_i| <= ld_wind}, stream.h:142-155
    with tools.h:41-49).  `rows`: score only these SNPs (spot checks at large sizes)."""
    positions = np.asarray(positions, dtype=np.float64)
    m = packed.shape[0]
    assert np.all(positions >= 0) and np.all(np.diff(positions) >= 0)
    codes = decode_bed_msb_first(packed, n_org)
    st = snp_stats(codes, maf_thr)
    keep, A, D, rstd = st["keep"], st["A"], st["D"], st["rstd"]
    n = float(n_org)

    l2 = np.full(m, np.nan)
    l2d = np.full(m, np.nan)
    wsa = np.full(m, -1, dtype=np.int32)
    wsd = np.full(m, -1, dtype=np.int32)
    wsde = np.full(m, -1, dtype=np.int32)
    lo = np.searchsorted(positions, positions - ld_wind, side="left")
    hi = np.searchsorted(positions, positions + ld_wind, side="right")   # exclusive
    with np.errstate(all="ignore"):
        dom_ok = rstd > std_thr                          # tools.h:31-33 (NaN -> False)
    todo = np.nonzero(keep)[0] if rows is None else [i for i in rows if keep[i]]
    for i in todo:
        js = np.arange(lo[i], hi[i])
        # the searchsorted bounds use pos -/+ wind; re-check with the reference's own
        # expression |pos_j - pos_i| <= wind (tools.h:43-44) at the two edges
        js = js[np.abs(positions[js] - positions[i]) <= ld_wind]
        js = js[keep[js] & (js != i)]
        with np.errstate(all="ignore"):
            ca = A[js] @ A[i] / n                        # tools.h:89
            r2 = ca * ca
            add = 1.0 + np.sum(1.0 - (1.0 - r2) * (n - 1) / (n - 2))    # tools.h:91, ldscalc.h:28,38
            jd = js[dom_ok[js]]
            cd = D[jd] @ A[i] / n                        # ldscalc.h:41
            r2d = cd * cd
            adj = 1.0 - (1.0 - r2d) * (n - 1) / (n - 2)
            l2[i] = add
            l2d[i] = np.sum(adj)
            wsa[i] = len(js)
            wsd[i] = len(jd)
            wsde[i] = int(np.sum(adj > rsq_thr))         # tools.h:36-38 (strict)
    rstd_out = np.where(keep, rstd, np.nan)
    return {"l2": l2, "l2d": l2d, "maf": st["maf"], "residuals_std": rstd_out,
            "l2_ws": wsa, "l2d_ws": wsd, "l2d_wse": wsde, "keep": keep,
            "counts": np.stack([st["n_h"], st["n_a"], st["n_m"]], 1).astype(np.int64)}


def cross_counts(codes_i: np.ndarray, codes_j: np.ndarray) -> np.ndarray:
    """Exact integer cross-counts G_pq(i,j) = sum_s p_i(s) q_j(s) for p in {g,m},
    q in {m,h,a} (SURVEY.md section 8a 'exact integer form').  Returns int64
    [2, 3, n_i, n_j] ordered (g,m) x (m,h,a)."""
    g = ((codes_i == 2) + 2 * (codes_i == 3)).astype(np.int64)
    mi = (codes_i == 1).astype(np.int64)
    planes_j = [(codes_j == 1), (codes_j == 2), (codes_j == 3)]
    out = np.empty((2, 3, codes_i.shape[0], codes_j.shape[0]), dtype=np.int64)
    for a, left in enumerate((g, mi)):
        for b, right in enumerate(planes_j):
            out[a, b] = left @ right.astype(np.int64).T
    return out


def ld_scores_rows(packed: np.ndarray, n_org: int, positions: np.ndarray, ld_wind: float, maf_thr: float,
                   std_thr: float, rsq_thr: float, rows, max_span: int = 4096, block: int = 1024) -> dict:
    """calculate() (ldscalc.h:8-65) for a few rows of a bfile too large to score whole.  Every requested row is scored
    against its complete window and nothing else, which is all the reference ever looks at for that row
    (stream.h:142-155, ldscalc.h:34-47).  Rows whose windows overlap share one pass over the union of their windows
    (up to `max_span` SNPs); the window SNPs are decoded and standardised `block` at a time, so memory stays at
    block x n_org doubles however wide the window is (configs[4]: 20,000 SNPs x 100,000 organisms).
    Returns the seven result arrays restricted to `rows` (ascending), plus "keep" and "rows"."""
    positions = np.asarray(positions, dtype=np.float64)
    assert np.all(positions >= 0) and np.all(np.diff(positions) >= 0)
    rows = np.asarray(sorted(int(r) for r in rows), dtype=np.int64)
    lo = np.searchsorted(positions, positions[rows] - ld_wind, side="left")
    hi = np.searchsorted(positions, positions[rows] + ld_wind, side="right")      # exclusive
    n = float(n_org)
    nr = len(rows)
    l2 = np.full(nr, np.nan); l2d = np.full(nr, np.nan)
    wsa = np.full(nr, -1, dtype=np.int32); wsd = np.full(nr, -1, dtype=np.int32); wsde = np.full(nr, -1, dtype=np.int32)
    st_r = snp_stats(decode_bed_msb_first(packed[rows], n_org), maf_thr)
    keep_r, A_r = st_r["keep"], st_r["A"]
    k = 0
    while k < nr:
        a, b, e = int(lo[k]), int(hi[k]), k + 1
        while e < nr and max(b, int(hi[e])) - a <= max_span:
            b = max(b, int(hi[e]))
            e += 1
        idx = [q for q in range(k, e) if keep_r[q]]
        add = {q: 1.0 for q in idx}                     # ldscalc.h:27-28
        dom = {q: 0.0 for q in idx}
        cnt = {q: [0, 0, 0] for q in idx}
        for s0 in range(a, b, block):
            s1 = min(b, s0 + block)
            st = snp_stats(decode_bed_msb_first(packed[s0:s1], n_org), maf_thr)
            with np.errstate(all="ignore"):
                dom_ok = st["rstd"] > std_thr                                   # tools.h:31-33 (NaN -> False)
            js_all = np.arange(s0, s1)
            for q in idx:
                i = rows[q]
                # the reference's own predicate at the edges (tools.h:43-44), kept SNPs only, not the SNP itself
                m = st["keep"] & (np.abs(positions[js_all] - positions[i]) <= ld_wind) & (js_all != i)
                if not m.any():
                    continue
                with np.errstate(all="ignore"):
                    ca = st["A"][m] @ A_r[q] / n                                # tools.h:89
                    add[q] += float(np.sum(1.0 - (1.0 - ca * ca) * (n - 1) / (n - 2)))   # tools.h:91, ldscalc.h:38
                    md = m & dom_ok
                    cd = st["D"][md] @ A_r[q] / n                               # ldscalc.h:41
                    adj = 1.0 - (1.0 - cd * cd) * (n - 1) / (n - 2)
                    dom[q] += float(np.sum(adj))
                cnt[q][0] += int(m.sum())
                cnt[q][1] += int(md.sum())
                cnt[q][2] += int(np.sum(adj > rsq_thr))                         # tools.h:36-38 (strict)
        for q in idx:
            l2[q], l2d[q] = add[q], dom[q]
            wsa[q], wsd[q], wsde[q] = cnt[q]
        k = e
    return {"l2": l2, "l2d": l2d, "maf": st_r["maf"], "residuals_std": np.where(keep_r, st_r["rstd"], np.nan),
            "l2_ws": wsa, "l2d_ws": wsd, "l2d_wse": wsde, "keep": keep_r, "rows": rows}
