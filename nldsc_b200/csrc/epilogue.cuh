// FP64 epilogue shared by the tcgen05 band kernels and the SIMT yardstick:
// exact integer cross-counts -> correlations -> bias-corrected r^2 -> row sums.
//
// Replaces Math::r2_adjusted (reference tools.h:87-92) and the body of the OpenMP loop
// in calculate() (ldscalc.h:34-47).  With x = g + mu*m the imputed additive vector and
// res = ch*h + ca*a + cm*m the dominance residual:
//   sum x_i x_j   = G_gg + mu_j G_gm + mu_i (G_mg + mu_j G_mm)
//   sum x_i res_j = ch G_gh + ca G_ga + cm G_gm + mu_i (ch G_mh + ca G_ma + cm G_mm)
//   corr = (sum - n xbar_i ybar_j) / (n sd_i sd_j);  r2adj = 1 - (1 - corr^2)(n-1)/(n-2)
// Measured and rejected (profiles/r02g_ab_step*.json, same box): accumulating only (sum - ...)^2 / sd_j^2 per pair and
// applying the affine bias correction once per row in finalize_kernel -- six FP64 operations per pair fewer -- was 1.4 %
// SLOWER on configs[1] (3.586 vs 3.537 ms per step): the exposed epilogue waits on memory, not on the FP64 pipe.
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

// Does any j in [j_first, j_last] fall into this row's window?  (warp-uniform skipping of the out-of-window corners
// of edge tiles: a skipped pair would have added +0.0 / 0 to every sum, so skipping changes no bit of the result)
__device__ __forceinline__ bool row_needs(const RowCtx& r, int32_t j_first, int32_t j_last) {
    return j_last >= r.lo && j_first <= r.hi;
}

// 1 - (1 - corr^2) (n-1)/(n-2)   (tools.h:90-91), one fixed sequence of operations for every kernel
__device__ __forceinline__ double r2adj(double corr, double adj_c) {
    return __fma_rn(-(1.0 - corr * corr), adj_c, 1.0);
}

// Exact int32 -> double without the conversion unit (I2F.F64 runs on the quarter-rate XU pipe): build the
// double 2^52 + 2^31 + x from its bit pattern and subtract the bias.  Exact for every int32.
__device__ __forceinline__ double i2d(int32_t x) {
    return __hiloint2double(0x43300000, x ^ static_cast<int32_t>(0x80000000u)) - 4503601774854144.0;
}

// A cross-count as a double.  kF32 = false: an int32 (int8 GEMM).  kF32 = true: the bit pattern of an FP32 accumulator of the fp4
// GEMM holding an exact integer -- ONE conversion instruction on the otherwise idle XU pipe instead of the five ALU / FP64
// instructions of float -> int -> double; the epilogue is bound by instruction issue (profiles/r02u_band_tc_ncu.txt).
template <bool kF32>
__device__ __forceinline__ double cnt2d(int32_t x) {
    if constexpr (kF32) return static_cast<double>(__int_as_float(x));
    else return i2d(x);
}

// One ordered pair (i, j): the six cross-counts are exact integers from the int8 GEMM.
// Branch-free: every pair is evaluated and its terms are SELECTED in or out, so the compiler can interleave the
// FP64 chains of the unrolled j loop (with early returns each pair cost ~1,000 cycles of exposed latency).
// A rejected pair may compute inf / NaN (a dropped SNP has 1/sd = 0 or inf); selects never propagate them.
// `fj` / `c0..c3` are SNP j's flags and its 64-byte SnpCoef as four double2 (from global memory, or from the copy
// the wide kernel stages in shared memory while its epilogue warps wait for the MMAs).
template <bool kF32 = false>
__device__ __forceinline__ void pair_eval_core(const BandArgs& a, const RowCtx& r, int32_t gj, uint8_t fj, const double2 c0,
                                               const double2 c1, const double2 c2, const double2 c3,
                                               int32_t Ggm, int32_t Ggh, int32_t Gga,
                                               int32_t Gmm, int32_t Gmh, int32_t Gma, RowAcc& acc) {
    // stream.h:145-147, tools.h:41-49: in the window, not the SNP itself; cache_[j] is false for dropped SNPs
    const bool in_win = gj >= r.lo && gj <= r.hi && gj != r.gi;
    const bool use = in_win && (fj & kFlagKeep);
    const bool use_d = use && (fj & kFlagDom);
    const double mu_j = c0.x, xbar_j = c0.y, inv_sd_j = c1.x, ch = c1.y;
    const double ca = c2.x, cm = c2.y, rbar_j = c3.x, inv_rsd_j = c3.y;

    const double dGgm = cnt2d<kF32>(Ggm), dGgh = cnt2d<kF32>(Ggh), dGga = cnt2d<kF32>(Gga);
    const double dGmm = cnt2d<kF32>(Gmm), dGmh = cnt2d<kF32>(Gmh), dGma = cnt2d<kF32>(Gma);
    // G_gg = G_gh + 2 G_ga, G_mg = G_mh + 2 G_ma: exact either way (integers far below 2^53)
    const double dGgg = kF32 ? __fma_rn(2.0, dGga, dGgh) : i2d(Ggh + 2 * Gga);
    const double dGmg = kF32 ? __fma_rn(2.0, dGma, dGmh) : i2d(Gmh + 2 * Gma);

    // Every multiply-add below is contracted EXPLICITLY: pair_eval_mfree evaluates the same expressions with the four
    // missing-genotype counts and G_gm identically zero, and must give the same bits (fma(x, 0, y) == y).
    // additive: ldscalc.h:38
    const double sa = __fma_rn(-r.nxbar, xbar_j, __fma_rn(r.mu, __fma_rn(mu_j, dGmm, dGmg), __fma_rn(mu_j, dGgm, dGgg)));
    const double ra = sa * r.w * inv_sd_j;
    const double adj_a = r2adj(ra, a.adj_c);                    // tools.h:91
    acc.add += use ? adj_a : 0.0;
    acc.wsa += use ? 1 : 0;

    // dominance: ldscalc.h:40-46
    const double u0 = __fma_rn(cm, dGgm, __fma_rn(ca, dGga, ch * dGgh));
    const double u1 = __fma_rn(cm, dGmm, __fma_rn(ca, dGma, ch * dGmh));
    const double sd = __fma_rn(-r.nxbar, rbar_j, __fma_rn(r.mu, u1, u0));
    const double rd = sd * r.w * inv_rsd_j;
    const double adj = r2adj(rd, a.adj_c);
    acc.dom += use_d ? adj : 0.0;
    acc.wsd += use_d ? 1 : 0;
    acc.wsde += (use_d && adj > a.rsq_thr) ? 1 : 0;             // tools.h:36-38 (strict)
}

// The same pair when neither SNP block holds a missing genotype: G_gm = G_mm = G_mh = G_ma = 0 exactly, so only
// G_gh and G_ga arrive (band_mfree_body).  Bitwise equal to pair_eval_core on those counts.
template <bool kF32 = false>
__device__ __forceinline__ void pair_eval_mfree(const BandArgs& a, const RowCtx& r, int32_t gj, uint8_t fj, const double2 c0,
                                                const double2 c1, const double2 c2, const double2 c3,
                                                int32_t Ggh, int32_t Gga, RowAcc& acc) {
    const bool in_win = gj >= r.lo && gj <= r.hi && gj != r.gi;
    const bool use = in_win && (fj & kFlagKeep);
    const bool use_d = use && (fj & kFlagDom);
    const double xbar_j = c0.y, inv_sd_j = c1.x, ch = c1.y, ca = c2.x, rbar_j = c3.x, inv_rsd_j = c3.y;
    const double dGgh = cnt2d<kF32>(Ggh), dGga = cnt2d<kF32>(Gga);
    const double dGgg = kF32 ? __fma_rn(2.0, dGga, dGgh) : i2d(Ggh + 2 * Gga);
    const double sa = __fma_rn(-r.nxbar, xbar_j, dGgg);
    const double ra = sa * r.w * inv_sd_j;
    const double adj_a = r2adj(ra, a.adj_c);
    acc.add += use ? adj_a : 0.0;
    acc.wsa += use ? 1 : 0;
    const double u0 = __fma_rn(ca, dGga, ch * dGgh);
    const double sd = __fma_rn(-r.nxbar, rbar_j, u0);
    const double rd = sd * r.w * inv_rsd_j;
    const double adj = r2adj(rd, a.adj_c);
    acc.dom += use_d ? adj : 0.0;
    acc.wsd += use_d ? 1 : 0;
    acc.wsde += (use_d && adj > a.rsq_thr) ? 1 : 0;
}

// SNP j's constants straight from global memory (narrow kernel, SIMT yardstick)
__device__ __forceinline__ void pair_eval(const BandArgs& a, const RowCtx& r, int32_t gj,
                                          int32_t Ggm, int32_t Ggh, int32_t Gga,
                                          int32_t Gmm, int32_t Gmh, int32_t Gma, RowAcc& acc) {
    const bool in_win = gj >= r.lo && gj <= r.hi && gj != r.gi;
    const int32_t gjc = in_win ? gj : 0;                        // a valid index: the loads below are unconditional
    const uint8_t fj = __ldg(a.flags + gjc);
    const double2* cp = reinterpret_cast<const double2*>(a.coef + gjc);
    pair_eval_core<false>(a, r, gj, fj, __ldg(cp), __ldg(cp + 1), __ldg(cp + 2), __ldg(cp + 3), Ggm, Ggh, Gga, Gmm, Gmh, Gma, acc);
}

}  // namespace nldsc
