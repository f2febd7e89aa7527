/*
 * nldsc_ld.h -- C ABI of the B200-native LD-score engine (libnldsc_ld.so).
 *
 * Drop-in boundary for the reference's `_ldscore.calculate`
 * (reference: nldsc/ldscore/_ldscore/ldscore.cpp:53 -> ldscalc.h:8-65, called from
 * nldsc/ldscore/routine.py:77-89).  Plain pointers and sizes only; no torch, no
 * C++ types.  Every entry point returns 0 on success, NLDSC_ERR_INVALID for what
 * the reference raises as std::invalid_argument (-> Python ValueError; e.g. the
 * PLINK magic check, stream.h:88-102) and NLDSC_ERR_RUNTIME for everything the
 * reference would surface as std::runtime_error (-> RuntimeError): CUDA failures,
 * I/O errors, a missing GPU.  There is NO CPU fallback: without a usable sm_100
 * device the calls fail with NLDSC_ERR_RUNTIME and a message.
 *
 * Result conventions are the reference's (data.h:21-31, ldscalc.h:13-21,49-54,
 * stream.h:165-179): seven arrays of length n_snp aligned with the .bim; SNPs
 * dropped by the MAF filter keep l2 = l2d = NaN, l2_ws = l2d_ws = l2d_wse = -1,
 * residuals_std = NaN and carry their computed maf.
 */
#ifndef NLDSC_LD_H
#define NLDSC_LD_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NLDSC_OK 0
#define NLDSC_ERR_INVALID 1 /* std::invalid_argument in the reference -> ValueError */
#define NLDSC_ERR_RUNTIME 2 /* std::runtime_error -> RuntimeError */

/* Mirrors LDScoreParams (reference data.h:33-66); same field meaning and order. */
typedef struct nldsc_ld_params {
    const char* bedfile;     /* absolute path of the .bed (UTF-8); ignored by *_packed / session calls */
    int32_t n_snp;           /* number of SNPs (rows of the .bim) */
    int32_t n_org;           /* number of organisms (rows of the .fam) */
    double ld_wind;          /* window half-width, in the unit of `positions` (cM or bp) */
    const double* positions; /* n_snp coordinates; must be >= 0 and non-decreasing */
    double maf;              /* drop SNP iff (double)maf_float <= maf        (encoder.h:120) */
    double std_thr;          /* dominance usable iff residual std > std_thr  (tools.h:31-33) */
    double rsq_thr;          /* WSDE counts r2adj_dom > rsq_thr, strict      (tools.h:36-38) */
} nldsc_ld_params;

/* Mirrors LDScoreResult (reference data.h:21-31).  Caller-allocated arrays. */
typedef struct nldsc_ld_result {
    double* l2;            /* additive LD score                 (ldscalc.h:49) */
    double* l2d;           /* dominance LD score                (ldscalc.h:50) */
    double* maf;           /* float MAF widened to double       (stream.h:165-171) */
    double* residuals_std; /* std of the dominance residual     (stream.h:173-179) */
    int32_t* l2_ws;        /* WSA  (ldscalc.h:52) */
    int32_t* l2d_ws;       /* WSD  (ldscalc.h:53) */
    int32_t* l2d_wse;      /* WSDE (ldscalc.h:54) */
} nldsc_ld_result;

/* Replaces calculate(LDScoreParams&) (ldscalc.h:8): opens and validates the .bed
 * (magic 6c 1b 01 and, unlike the reference, the exact size 3 + n_snp*ceil(n_org/4)),
 * streams it to the GPU(s) and fills `out` (arrays of length n_snp).
 * GPUs: environment NLDSC_GPUS (default 1) contiguous SNP shards with a window
 * halo, one host thread per GPU, no collective; results gathered host-side.
 * Reading: NLDSC_READERS pread threads per shard (default 4) through the page cache, or past it with
 * NLDSC_ODIRECT=1 (O_DIRECT, 4 KB-aligned supersets of every piece) for genomes larger than host memory. */
int nldsc_ld_calculate(const nldsc_ld_params* params, nldsc_ld_result* out, char* err, size_t errlen);

/* nldsc_ld_calculate with the GPUs named by the caller instead of the environment: shards go to devices
 * first_device .. first_device + n_devices - 1.  Lets ONE process score several chromosome bfiles at the
 * same time, one host thread and one GPU (or GPU group) each -- the multi-chromosome driver of SURVEY.md
 * section 8(f)-3; the reference handles one chromosome per invocation (ldscore/common.py:114-117). */
int nldsc_ld_calculate_on(const nldsc_ld_params* params, nldsc_ld_result* out, int32_t first_device,
                          int32_t n_devices, char* err, size_t errlen);

/* Same computation from a host buffer holding the genotype bytes of the .bed
 * WITHOUT its 3-byte magic: n_snp rows of ceil(n_org/4) bytes (what
 * BedStreamReader::read consumes, stream.h:43-69). */
int nldsc_ld_calculate_packed(const uint8_t* packed, const nldsc_ld_params* params,
                              nldsc_ld_result* out, char* err, size_t errlen);

/* nldsc_ld_calculate_packed with the GPUs named by the caller (thread-safe placement, like nldsc_ld_calculate_on):
 * several host threads can each score a chromosome that is already in host memory, one GPU (group) each. */
int nldsc_ld_calculate_packed_on(const uint8_t* packed, const nldsc_ld_params* params, nldsc_ld_result* out,
                                 int32_t first_device, int32_t n_devices, char* err, size_t errlen);

/* Prepares devices first_device .. first_device + n_devices - 1 for the calls above, every device on its own host
 * thread: CUDA context, pinned staging buffers, stream pool, kernel images.  Optional -- the first call on a device
 * does the same -- but a driver that is about to keep 8 GPUs busy pays these 0.3-1 s per device once and in parallel
 * instead of inside the first chromosome of every GPU. */
int nldsc_ld_warmup(int32_t first_device, int32_t n_devices, char* err, size_t errlen);

/* ---- host-side planning (no GPU needed) --------------------------------------------------- */
/* Inclusive window [lo[i], hi[i]] of every SNP: {j : |pos_j - pos_i| <= ld_wind}, the
 * reference's SNPFilter::filter (tools.h:41-49) with the sliding-cache bounds of
 * ChunkwiseReader (stream.h:142-155,182-197) for sorted non-negative positions. */
int nldsc_ld_plan_windows(const nldsc_ld_params* params, int32_t* lo, int32_t* hi, char* err, size_t errlen);
/* Contiguous SNP ranges of (roughly) equal band area for `parts` GPUs: shard k owns rows
 * [cuts[k], cuts[k+1]); cuts has parts + 1 entries, interior cuts on multiples of 64. */
int nldsc_ld_plan_shards(const nldsc_ld_params* params, int32_t parts, int32_t* cuts, char* err, size_t errlen);

/* ---- session API: one shard on one GPU, device-resident between calls ------------- */
typedef struct nldsc_ld_session nldsc_ld_session;

typedef struct nldsc_ld_stats {
    int64_t pairs;            /* ordered in-window SNP pairs of the owned rows BEFORE the MAF filter
                                 (the scored pairs are the sum of the returned l2_ws >= 0) */
    int64_t tiles;            /* supertiles (2 i-blocks x 1 j-block of 64 SNPs) issued to the band kernel */
    int32_t chunks;           /* plane-ring refills */
    int32_t kernel_launches;  /* kernels launched by the last run() */
    int32_t load_begin;       /* first SNP row the shard needs (owned range minus halo, 64-aligned) */
    int32_t load_end;         /* one past the last SNP row needed */
    int32_t npad;             /* bytes per plane row: padded organisms with int8 planes, half of them with fp4 planes */
    int32_t band_impl;        /* 1 = tcgen05 kernel, 2 = its wide variant (two j-blocks per supertile; the default at every N), 0 = SIMT yardstick (NLDSC_BAND=simt) */
    float ms_count;           /* CUDA-event times of the last run(), per kernel family; counting is fused
                                 into the unpack kernel, so this one stays 0 */
    float ms_stats;
    float ms_unpack;
    float ms_band;
    float ms_finalize;
    float ms_total;
    /* missing-free path (band_tc_mfree_kernel: the two cross-counts g x {h, a} instead of six): */
    int32_t mfree_chunks;     /* chunks of the last run() whose columns held no missing genotype and took it */
    int32_t fp4;              /* 1 = the planes hold e2m1 (fp4) nibbles and the band kernel multiplies them with kind::mxf4 */
    int64_t qtiles;           /* missing-free supertiles planned (4 i-blocks x 2 j-blocks each), all chunks */
    int64_t mfree_qtiles;     /* ... of the chunks that took the path in the last run() */
    int64_t mfree_tiles_skipped;   /* 6-plane supertiles (wide: 2 i-blocks x <= 2 j-blocks) those chunks did NOT run */
    int64_t xtiles;           /* 160-column supertiles (2 i-blocks x five 32-SNP column groups, fp4 planes) planned, all chunks; 0 = that body is off */
    int64_t xtiles_run;       /* ... of the chunks that did not take the missing-free path in the last run() */
} nldsc_ld_stats;

/* Shard = SNP rows [row_begin,row_end) of the bfile described by `params`
 * (bedfile ignored).  Plans windows, tiles and the halo from `positions`. */
int nldsc_ld_session_create(nldsc_ld_session** out, int device, const nldsc_ld_params* params,
                            int32_t row_begin, int32_t row_end, char* err, size_t errlen);
/* Host -> device copy of packed rows [load_begin,load_end) (see stats), asynchronous
 * on `stream` (a cudaStream_t, or NULL for the session's own stream). */
int nldsc_ld_session_upload(nldsc_ld_session* s, const uint8_t* packed_rows, void* stream,
                            char* err, size_t errlen);
/* Unpack (+ genotype counts) -> per-SNP stats -> banded cross-count GEMM (tcgen05; fp4 or int8 operands) + FP64 epilogue -> finalize,
 * all asynchronous on `stream`.  Inputs must be resident (upload done). */
int nldsc_ld_session_run(nldsc_ld_session* s, void* stream, char* err, size_t errlen);
/* Device -> host copy of the owned rows' seven arrays (length row_end-row_begin),
 * then synchronises `stream`. */
int nldsc_ld_session_download(nldsc_ld_session* s, nldsc_ld_result* out, void* stream,
                              char* err, size_t errlen);
/* Waits for the session's own copy and compute streams (upload / run issued with stream = NULL). */
int nldsc_ld_session_sync(nldsc_ld_session* s, char* err, size_t errlen);
int nldsc_ld_session_stats(nldsc_ld_session* s, nldsc_ld_stats* out);
void nldsc_ld_session_destroy(nldsc_ld_session* s);

/* What the last nldsc_ld_calculate* call made on the CALLING THREAD did, one record per shard (= per GPU, or per
 * NLDSC_TEST_SHARDS shard): host wall time of every phase and, on request, the CUDA-event times of the device work;
 * the records are what the genome driver and bench.py report as per-GPU busy time. */
typedef struct nldsc_ld_shard_report {
    int32_t device;
    int32_t row_begin, row_end;      /* owned rows */
    int32_t load_begin, load_end;    /* rows read and unpacked: owned rows + window halo */
    int32_t chunks;
    int64_t pairs;                   /* ordered in-window pairs of the owned rows before the MAF filter */
    int64_t bytes_h2d;               /* packed genotype bytes this shard uploaded */
    float ms_open, ms_upload_start, ms_plan, ms_run_issue, ms_download, ms_free;   /* host phases, in call order */
    float ms_wall;                   /* the shard's thread, start to end */
    float ms_dev_upload;             /* first to last byte of the H2D copy (host-buffer uploads; 0 for the .bed readers) */
    float ms_dev_run, ms_dev_band, ms_dev_unpack;   /* run() on the device; its band / unpack kernel families */
} nldsc_ld_shard_report;
/* The ms_dev_* fields cost a hundred CUDA-event queries per shard (0.15-0.2 ms), so they are filled only after
 * nldsc_ld_report_device_times(1) (process-wide; default off) or under NLDSC_TRACE; the host phase times always are. */
int nldsc_ld_report_device_times(int32_t on);
/* n = number of shards of that call; the first min(n, max) records are written to `out` (may be NULL). */
int nldsc_ld_last_call_report(nldsc_ld_shard_report* out, int32_t max, int32_t* n);
/* Releases the device memory cached between calls (the per-device workspace arena: packed rows, plane ring, ...)
 * on `device`, or on every device when device < 0.  Blocks in use by a running call are left alone. */
int nldsc_ld_arena_trim(int32_t device);

/* ---- allocation-free entry point --------------------------------------------------- */
/* No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Per-SNP statistics of `n_snp` packed .bed rows ((n_org + 3) / 4 bytes each, no slack needed) that are already in DEVICE
 * memory on `device`: d_maf_out[n_snp] and d_rstd_out[n_snp] receive exactly what nldsc_ld_calculate* return as `maf` and
 * `residuals_std` for the same rows and thresholds (every position used), d_counts_out (int32 [n_snp][3] = n_het, n_hom2,
 * n_missing, as nldsc_ld_debug_unpack; may be NULL) the genotype counts.  All three are DEVICE pointers the caller owns;
 * d_packed must be 4-byte aligned.  The work is queued on `stream` (a cudaStream_t, NULL = default stream) and NOT
 * synchronised: the outputs are valid once the stream has reached this point.
 * The call allocates nothing -- no cudaMalloc / cudaFree / allocating library routine: its intermediate arrays (counts,
 * per-SNP coefficients, flags) live in one static __device__ array of 19.25 MiB that every GPU which loads the library
 * holds until the process exits, sized for NLDSC_STATIC_MAX_SNPS rows.  n_snp beyond that is refused with
 * NLDSC_ERR_INVALID before anything touches the GPU; split the rows over several calls.  Every call shares that array
 * (it clears what it accumulates into), so calls that OVERLAP on the GPU -- other streams, other host threads -- race:
 * serialise them.  The first call on each device looks the array's address up (cudaGetSymbolAddress) and loads the
 * kernels; make it outside a capture. */
#define NLDSC_STATIC_MAX_SNPS 262144
int nldsc_ld_snp_stats_static(const uint8_t* d_packed, int32_t n_snp, int32_t n_org, double maf, double std_thr,
                              int32_t* d_counts_out, double* d_maf_out, double* d_rstd_out, int32_t device,
                              void* stream, char* err, size_t errlen);

/* ---- parity hooks (used by tests/ only; they run the same kernels as the session) -- */
/* Unpack + count `n_snp` packed rows: planes_out = int8 [4][n_snp][npad] in plane order
 * g,m,h,a (npad returned), counts_out = int32 [n_snp][3] = n_het, n_hom2, n_missing. */
int nldsc_ld_debug_unpack(const uint8_t* packed, int32_t n_snp, int32_t n_org, int8_t* planes_out,
                          int32_t* counts_out, int32_t* npad_out, char* err, size_t errlen);
/* The same through the fp4 unpack kernel: planes_out = bytes [4][n_snp][npad] with npad = plane row bytes (returned), two
 * organisms per byte (organism 2i in the low nibble), every value an e2m1 nibble: 0 -> 0x0, 1 -> 0x2, 2 -> 0x4. */
int nldsc_ld_debug_unpack_fp4(const uint8_t* packed, int32_t n_snp, int32_t n_org, int8_t* planes_out,
                              int32_t* counts_out, int32_t* npad_out, char* err, size_t errlen);
/* Raw band-kernel accumulators for SNP block pair (i_block, j_block) (64 SNPs each):
 * acc_out = int32 [128][192]; row 32*(s/16) + 16*p + s%16 is SNP s of the i block,
 * left plane p (0 = g, 1 = m); column 64*q + t is SNP t of the j block, right plane q
 * (0 = m, 1 = h, 2 = a). */
int nldsc_ld_debug_cross_counts(const uint8_t* packed, int32_t n_snp, int32_t n_org,
                                int32_t i_block, int32_t j_block, int32_t* acc_out,
                                char* err, size_t errlen);
/* nldsc_ld_plan_windows evaluated by the device kernel the calculate entry points use (plan.cu);
 * must equal the host routine element for element. */
int nldsc_ld_debug_plan_windows_device(const nldsc_ld_params* params, int32_t* lo, int32_t* hi, char* err, size_t errlen);
/* Host-only (no GPU needed): the 160-column supertiles session planning cuts for ONE chunk of `n_iblocks` consecutive 64-SNP
 * i-blocks whose j-ranges are the 64-SNP blocks [jb_lo[k], jb_hi[k]] (non-decreasing), in the kernel's execution order with
 * raster groups of `raster` i-pairs.  tiles_out = int32 [max_tiles][8] = a_row[2], i0[2], j0, slot, n_grp, odd per tile (the XTile of
 * csrc/common.cuh; slots numbered i-pair-major, one per j-block of a pair's range); n_tiles = how many there are (also when it
 * exceeds max_tiles).  Lets the CPU suite check that every 32-SNP column group of every i-pair is scored exactly once. */
int nldsc_ld_debug_plan_xtiles(const int32_t* jb_lo, const int32_t* jb_hi, int32_t n_iblocks, int32_t raster, int32_t* tiles_out,
                               int32_t max_tiles, int32_t* n_tiles, char* err, size_t errlen);
/* Measurement hook: runs the band kernel on `n_tiles` synthetic supertiles (2 x 64 x 64 SNP pairs each,
 * K = n_org, band-like operand reuse) `reps` times and returns the mean CUDA-event time of one launch.
 * mode 0 = the real kernel (TMA + MMA + epilogue), 1 = tcgen05 MMAs only (no operand loads: the int8
 * tensor-pipe ceiling of this tile shape), 2 = TMA loads only (no MMAs: the operand-feed ceiling). */
int nldsc_ld_debug_band_probe(int32_t mode, int32_t n_org, int32_t n_tiles, int32_t reps, float* ms_out,
                              char* err, size_t errlen);
/* Measurement hook: the int8 tensor-pipe ceiling on this box.  Every CTA pair issues `iters` groups of `per_commit` (4, 8 or 16; at
 * most `depth` <= 8 groups in flight) back-to-back tcgen05.mma.cta_group::2.kind::i8 (M = 256, N = n_cols, K = 32) on operands resident in shared
 * memory -- no loads, no epilogue.  ms_out = mean time of one launch, tops_out = int8 TOP/s it sustained
 * ("the int8 peak must be measured on the box"). */
int nldsc_ld_debug_i8_peak(int32_t n_cols, int32_t per_commit, int32_t depth, int32_t iters, int32_t reps,
                           float* ms_out, double* tops_out, char* err, size_t errlen);
/* Feasibility probe (measurement only): n_mma accumulating tcgen05.mma.kind::mxf4 (fp4 e2m1 operands holding 0 / 1 / 2, all
 * scale factors 1.0; M = 256, N = n_cols, K = 64) on known operand patterns; out = float [2][128][16]: the first 16
 * accumulator columns of every row of the first CTA pair; ms_out = kernel time (all_pairs != 0: on every CTA pair, for the rate).
 * Asks whether the fp4 pipe sums small integers exactly -- the precondition for running the cross-count GEMM on it. */
int nldsc_ld_debug_fp4_probe(int32_t n_mma, int32_t n_cols, int32_t per_commit, int32_t all_pairs, float* out, float* ms_out,
                             char* err, size_t errlen);
/* Measurement hook: the host's H2D ceiling with n_devices GPUs uploading at once.  Every device (own host thread) copies a
 * pinned host buffer of `bytes` (cudaHostAllocDefault, or cudaHostAllocWriteCombined when write_combined != 0) to its HBM
 * `reps` times, all starting together; gbs_out[k] = GB/s of device first_device + k.  (what bounds the
 * end-to-end call when 4 or 8 GPUs upload at the same time.) */
int nldsc_ld_debug_h2d_probe(int64_t bytes, int32_t write_combined, int32_t first_device, int32_t n_devices, int32_t reps,
                             double* gbs_out, char* err, size_t errlen);
/* Number of usable CUDA devices (0 and an error message when none). */
int nldsc_ld_device_count(int32_t* count, char* err, size_t errlen);

#ifdef __cplusplus
}
#endif
#endif /* NLDSC_LD_H */
