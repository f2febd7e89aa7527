// FP64 epilogue shared by the tcgen05 band kernel and the SIMT yardstick:
// exact integer cross-counts -> correlations -> bias-corrected r^2 -> row sums.
//
// Replaces Math::r2_adjusted (reference tools.h:87-92) and the body of the OpenMP loop
// in calculate() (ldscalc.h:34-47).  With x = g + mu*m the imputed additive vector and
// res = ch*h + ca*a + cm*m the dominance residual:
//   sum x_i x_j   = G_gg + mu_j G_gm + mu_i (G_mg + mu_j G_mm)
//   sum x_i res_j = ch G_gh + ca G_ga + cm G_gm + mu_i (ch G_mh + ca G_ma + cm G_mm)
//   corr = (sum - n xbar_i ybar_j) / (n sd_i sd_j);  r2adj = 1 - (1 - corr^2)(n-1)/(n-2)
#pragma once
#include "common.cuh"

namespace nldsc {

struct RowCtx {
    double mu;      // mu_i
    double nxbar;   // n * xbar_i
    double w;       // inv_sd_i / n
    int32_t lo, hi; // inclusive window (shard-local SNP indices)
    int32_t gi;     // this row's SNP index
    bool keep;
};

struct RowAcc {
    double add = 0.0, dom = 0.0;
    int32_t wsa = 0, wsd = 0, wsde = 0;
};

__device__ __forceinline__ RowCtx load_row_ctx(const BandArgs& a, int32_t gi) {
    RowCtx r;
    r.gi = gi;
    if (gi < a.n_load) {
        const SnpCoef c = a.coef[gi];
        r.mu = c.mu;
        r.nxbar = a.n * c.xbar;
        r.w = c.inv_sd * a.inv_n;
        r.lo = a.win_lo[gi];
        r.hi = a.win_hi[gi];
        r.keep = (a.flags[gi] & kFlagKeep) != 0;
    } else {
        r.mu = r.nxbar = r.w = 0.0;
        r.lo = 1; r.hi = 0;
        r.keep = false;
    }
    if (!r.keep) { r.lo = 1; r.hi = 0; }     // dropped rows score nothing (ldscalc.h:25)
    return r;
}

// One ordered pair (i, j): the six cross-counts are exact integers from the int8 GEMM.
__device__ __forceinline__ void pair_eval(const BandArgs& a, const RowCtx& r, int32_t gj,
                                          int32_t Ggm, int32_t Ggh, int32_t Gga,
                                          int32_t Gmm, int32_t Gmh, int32_t Gma, RowAcc& acc) {
    if (gj < r.lo || gj > r.hi || gj == r.gi) return;        // stream.h:145-147, tools.h:41-49
    const uint8_t fj = __ldg(a.flags + gj);
    if (!(fj & kFlagKeep)) return;                            // cache_[i] is false for dropped SNPs
    const double2* cp = reinterpret_cast<const double2*>(a.coef + gj);
    const double2 c0 = __ldg(cp), c1 = __ldg(cp + 1), c2 = __ldg(cp + 2), c3 = __ldg(cp + 3);
    const double mu_j = c0.x, xbar_j = c0.y, inv_sd_j = c1.x, ch = c1.y;
    const double ca = c2.x, cm = c2.y, rbar_j = c3.x, inv_rsd_j = c3.y;

    const double dGgm = Ggm, dGgh = Ggh, dGga = Gga, dGmm = Gmm, dGmh = Gmh, dGma = Gma;
    const double dGgg = (double)(Ggh + 2 * Gga);
    const double dGmg = (double)(Gmh + 2 * Gma);

    // additive: ldscalc.h:38
    const double sa = (dGgg + mu_j * dGgm) + r.mu * (dGmg + mu_j * dGmm) - r.nxbar * xbar_j;
    const double ra = sa * r.w * inv_sd_j;
    const double r2a = ra * ra;
    acc.add += 1.0 - (1.0 - r2a) * a.adj_c;                   // tools.h:91
    acc.wsa += 1;

    // dominance: ldscalc.h:40-46
    if (fj & kFlagDom) {
        const double u0 = ch * dGgh + ca * dGga + cm * dGgm;
        const double u1 = ch * dGmh + ca * dGma + cm * dGmm;
        const double sd = (u0 + r.mu * u1) - r.nxbar * rbar_j;
        const double rd = sd * r.w * inv_rsd_j;
        const double r2d = rd * rd;
        const double adj = 1.0 - (1.0 - r2d) * a.adj_c;
        acc.dom += adj;
        acc.wsd += 1;
        acc.wsde += (adj > a.rsq_thr) ? 1 : 0;                // tools.h:36-38 (strict)
    }
}

}  // namespace nldsc
