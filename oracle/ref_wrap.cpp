// TEST INFRASTRUCTURE ONLY -- not part of the shipped product path.
//
// extern "C" wrapper around the UNMODIFIED reference driver
// (`calculate`, /root/reference/nldsc/ldscore/_ldscore/ldscalc.h:8-65), so the
// tests and bench.py's CPU-baseline leg can call it through ctypes without
// pybind11 and so the float32 and FP64-promoted builds can coexist in one
// process (different sonames, RTLD_LOCAL).  Built by oracle/Makefile into
// oracle/_ref/ only.
#include "ldscalc.h"

#include <cstring>
#include <fcntl.h>
#include <unistd.h>

extern "C" int ref_ld_calculate(const char* bedfile, int n_snp, int n_org, double ld_wind,
                                double maf, double std_thr, double rsq_thr,
                                const double* positions,
                                double* l2, double* l2d, double* maf_out, double* rstd,
                                int* l2_ws, int* l2d_ws, int* l2d_wse,
                                char* err, size_t errlen) {
    // The reference ticks a progress bar on stdout once per SNP (ldscalc.h:59).
    fflush(stdout);
    int saved = dup(1);
    int devnull = open("/dev/null", O_WRONLY);
    dup2(devnull, 1);
    int rc = 0;
    try {
        LDScoreParams p(bedfile, n_snp, n_org, ld_wind, maf, std_thr, rsq_thr,
                        std::vector<double>(positions, positions + n_snp));
        LDScoreResult r = calculate(p);
        for (int i = 0; i < n_snp; ++i) {
            l2[i] = r.l2[i]; l2d[i] = r.l2d[i]; maf_out[i] = r.maf[i];
            rstd[i] = r.residuals_std[i];
            l2_ws[i] = r.l2_ws[i]; l2d_ws[i] = r.l2d_ws[i]; l2d_wse[i] = r.l2d_wse[i];
        }
    } catch (const std::invalid_argument& e) {
        rc = 1; if (err && errlen) { strncpy(err, e.what(), errlen - 1); err[errlen - 1] = 0; }
    } catch (const std::exception& e) {
        rc = 2; if (err && errlen) { strncpy(err, e.what(), errlen - 1); err[errlen - 1] = 0; }
    }
    fflush(stdout);
    dup2(saved, 1);
    close(saved); close(devnull);
    return rc;
}

extern "C" int ref_ld_elem_bytes() { return (int)sizeof(ARMA_SHIM_ELEM); }

// which dot the float build is using right now: "shim loop (omp simd)" or the BLAS symbol it resolved (NLDSC_REF_BLAS)
extern "C" const char* ref_shim_blas_name() { return arma::detail::blas().name; }
