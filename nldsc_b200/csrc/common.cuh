// Shared definitions for the LD-score CUDA path (sm_100a only).
#pragma once

#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <stdexcept>
#include <string>

namespace nldsc {

// ---- geometry -------------------------------------------------------------------
// SNP blocks are 64 wide and aligned on ORIGINAL .bim indices (multiples of 64), so the
// FP64 summation order of every row is independent of where shard boundaries fall.
constexpr int kBlk = 64;           // SNPs per i-block and per j-block
constexpr int kPlanes = 4;         // plane order in HBM: g, m, h, a
constexpr int kPlaneG = 0, kPlaneM = 1, kPlaneH = 2, kPlaneA = 3;
constexpr int kKAlign = 128;       // plane rows are padded to a multiple of 128 organisms

// Per-SNP constants the epilogue needs (derived from the three genotype counts only;
// reference: encoder.h:91-133, tools.h:54-85).  64 bytes, one cache-line half.
struct __align__(16) SnpCoef {
    double mu;       // float-rounded additive imputation mean, widened   (encoder.h:114)
    double xbar;     // mean of the imputed additive vector               (tools.h:81)
    double inv_sd;   // 1 / population std of the imputed additive vector (tools.h:82)
    double ch;       // dominance residual value for a het genotype: 2 - beta
    double ca;       // ... hom-second: 2 - 2 beta
    double cm;       // ... missing: mu_d - beta * mu_a
    double rbar;     // mean of the residual vector
    double inv_rsd;  // 1 / population std of the residual vector
};

constexpr uint8_t kFlagKeep = 1;   // passed the MAF filter (encoder.h:120-122)
constexpr uint8_t kFlagDom = 2;    // residual std > std_thr (tools.h:31-33, ldscalc.h:40)

// One supertile of the band: two vertically adjacent 64-SNP i-blocks x one 64-SNP j-block, scored
// by a CTA pair (cluster rank r owns i-block r; both share the j-block's operand).
constexpr int32_t kDeadBlock = 0x3FFFFFC0;   // i0 of a missing second i-block: scores nothing
struct Tile {
    int32_t a_row[2];  // rows of the two i-blocks inside the plane ring
    int32_t i0[2];     // first SNP (shard-local index) of each i-block, or kDeadBlock
    int32_t b_row;     // row of the j-block inside the plane ring
    int32_t j0;        // first SNP (shard-local index) of the j-block
    int32_t slot;      // where its partial row sums go: position in i-pair-major order (what finalize walks), which is
                       // NOT the execution order -- tiles of 8 i-pairs are interleaved by j-block so that concurrent
                       // CTA pairs share B operands in L2 as well as A (wide windows would otherwise stream B from HBM)
    int32_t b_row2;    // wide supertiles (band_tc_wide_kernel, the default): ring row of a SECOND j-block j0 + 64 scored from
                       // the same A tiles (its partials go to slot + 1), or -1.  Unused (0) by the narrow kernel.
};

// Missing-free supertile (band_mfree_body of band_tc_wide_kernel): FOUR consecutive 64-SNP i-blocks (cluster rank r owns blocks 2r, 2r+1 =
// one i-pair of the slot scheme) x the even-aligned PAIR of j-blocks starting at j0.  Without missing genotypes only the
// two cross-counts g x {h, a} are needed (SURVEY.md section 8d "4*N int8 ops"), so the 128 MMA rows of a CTA hold g of
// 128 i-SNPs and one N = 256 accumulator holds [h | a] of 128 j-SNPs: a third of the MMA work per SNP pair.
struct QTile {
    int32_t a_row[4];    // ring rows of the four i-blocks (a missing block repeats block 0's row)
    int32_t i0[4];       // first SNP (shard-local) of each i-block, or kDeadBlock
    int32_t b_row[2];    // ring rows of j-blocks j0 / 64 and j0 / 64 + 1 (a missing second block repeats the first)
    int32_t j0;          // first SNP of the j-block pair (a multiple of 128)
    int32_t slot[2][2];  // [i-pair][j-block]: partial-sum slot, or -1 where the j-block lies outside the i-pair's range
    int32_t pad;
};

// 160-column supertile (band_x160_body of band_tc_wide_kernel, fp4 planes with missing genotypes): one i-pair x FIVE
// consecutive 32-SNP column groups starting at j0 (a multiple of 32, anywhere in the i-pair's j-range).  An mxf4 MMA costs
// the same ~150 clk whatever its N, so the 480 accumulator columns of 160 j-SNPs x {m, h, a} are filled by TWO MMAs per 64
// organisms (N = 256 + 224) where two 64-SNP j-blocks take two for 128 j-SNPs: a fifth fewer instructions per SNP pair.
// Group g of the supertile is half (odd + g) & 1 of the 64-SNP j-block at slot + ((odd + g) >> 1): the same partial-sum
// groups, summed in the same order, as the 64-block supertiles -- bitwise identical scores.
struct XTile {
    int32_t a_row[2];  // ring rows of the two i-blocks
    int32_t i0[2];     // first SNP (shard-local) of each i-block, or kDeadBlock
    int32_t j0;        // first SNP of the supertile (shard-local, a multiple of 32)
    int32_t slot;      // partial-sum slot of the 64-SNP j-block that holds j0
    int32_t n_grp;     // 1..5 groups lie inside the i-pair's j-range; the others are multiplied but never scored
    int32_t odd;       // (j0 / 32) & 1
};

// Per-row partial sums a (64 x 64 SNP) tile contributes, in two halves: j-SNPs 0..31 and 32..63 of the j-block
// (the wide kernel scores the halves in different warps).  SoA, index = part_index(slot, rank, half, row).
__host__ __device__ inline int64_t part_index(int64_t slot, int rank, int half, int row) {
    return ((slot * 2 + rank) * 2 + half) * kBlk + row;
}
struct Partials {
    double* add;     // sum of r2adj(A_i, A_j)
    double* dom;     // sum of r2adj(A_i, D_j) over dominance-usable j
    int32_t* wsa;    // in-window kept j
    int32_t* wsd;    // ... that are dominance-usable
    int32_t* wsde;   // ... with r2adj_dom > rsq_thr
};

struct BandArgs {
    const Tile* tiles;
    int32_t n_tiles;
    int32_t k_steps;            // ceil(npad / 128)
    const SnpCoef* coef;        // [n_load]
    const uint8_t* flags;       // [n_load]
    const int32_t* win_lo;      // [n_load] inclusive shard-local window bounds
    const int32_t* win_hi;
    int32_t n_load;
    double n;                   // n_org
    double inv_n;
    double adj_c;               // (n - 1) / (n - 2)
    double rsq_thr;
    Partials part;
    int32_t* dump;              // optional raw accumulators [tile][rank][128][192] (parity hook)
    const QTile* qtiles;        // missing-free supertiles of the same chunk (band_mfree_body), same slots
    int32_t n_qtiles;
    const uint32_t* mfree_flag; // per chunk, written on the device after the chunk's stats: 1 = no kept SNP of the chunk's
                                // column range has a missing genotype: the launch then scores `qtiles` (2 planes) instead of
                                // `tiles` (6 planes); null = the missing-free path is off
    const uint32_t* ready_rows; // fused launch (one band launch per shard): ring rows [0, *ready_rows) are unpacked and their
                                // per-SNP statistics final; published after every chunk's unpack + stats.  null = everything is ready
    const XTile* xtiles;        // 160-column supertiles of the same chunk (band_x160_body; fp4 planes only), same slots; null = off
    int32_t n_xtiles;
    int32_t ring_rows;          // rows of the plane ring (a multiple of 64): ring row of SNP j = j % ring_rows
    int32_t fp4;                // the planes hold e2m1 nibbles (launch_unpack_fp4): the wide / missing-free bodies multiply them with
                                // tcgen05.mma.kind::mxf4 (scale factors 1.0, K = 64 per MMA, k_steps = plane row bytes / 128)
    int32_t runahead;           // wide body: start the next supertile's j-block-0 MMAs while the epilogue still reads accumulator 1
    int32_t no_skip;            // A/B: score the out-of-window column groups too (NLDSC_NOSKIP=1)
    int32_t commit2;            // experiment: release the operand ring two stages at a time (16 MMAs between commits)
    int32_t probe;              // measurement hook: 0 = normal, 1 = MMAs without TMA loads, 2 = TMA loads without MMAs, 3 = epilogue only
};

// ---- error plumbing ----------------------------------------------------------------
struct CudaError : std::runtime_error {
    using std::runtime_error::runtime_error;
};

std::string band_hang_note();   // band_tc.cu: which barrier a band kernel gave up on, if one did ("" otherwise)

inline void cuda_check(cudaError_t e, const char* what, const char* file, int line) {
    if (e != cudaSuccess) {
        char buf[512];
        snprintf(buf, sizeof(buf), "CUDA error %s at %s:%d: %s", what, file, line, cudaGetErrorString(e));
        throw CudaError(std::string(buf) + band_hang_note());
    }
}
#define CUDA_CHECK(x) ::nldsc::cuda_check((x), #x, __FILE__, __LINE__)

// ---- kernel launchers (defined in the .cu files) -------------------------------------
// counts / coef / flags / maf / rstd are indexed by shard-local row; rows [row0, row0 + n_rows) are processed
void launch_stats(const int32_t* counts, int32_t row0, int32_t n_rows, int32_t n_org, double maf_thr,
                  double std_thr, SnpCoef* coef, uint8_t* flags, double* maf_out,
                  double* rstd_out, cudaStream_t st, const uint8_t* unused = nullptr,   // unused[row] != 0: position < 0
                  uint8_t* blk_miss = nullptr);   // [row / 64] = 1 where a kept SNP has a missing genotype
void launch_unpack(const uint8_t* packed, int64_t bps, int32_t row0, int32_t n_rows,
                   int32_t n_org, int32_t npad, int8_t* planes, int64_t plane_stride,
                   int32_t dst_row0, int32_t* counts, cudaStream_t st);   // counts != null: += genotype counts (zeroed by the caller)
// the same planes as e2m1 (fp4) nibbles, two organisms per byte; npad = plane row BYTES (a multiple of 128 = 256 organisms)
void launch_unpack_fp4(const uint8_t* packed, int64_t bps, int32_t row0, int32_t n_rows,
                       int32_t n_org, int32_t npad, int8_t* planes, int64_t plane_stride,
                       int32_t dst_row0, int32_t* counts, cudaStream_t st);
void launch_windows(const double* pos, int32_t n, double wind, int32_t* lo, int32_t* hi, cudaStream_t st,
                    int32_t n_pad = 0, unsigned long long* total = nullptr);   // rows n..n_pad-1 get the empty window (1, 0)
void launch_copy16(const void* src, void* dst, size_t bytes, cudaStream_t st);   // SM copy, either side may be pinned host memory
void launch_publish_rows(uint32_t* ready_rows, uint32_t value, cudaStream_t st);   // release-store for BandArgs::ready_rows
void launch_zero_rows(int8_t* planes, int64_t plane_stride, int32_t npad, int32_t row0,
                      int32_t n_rows, cudaStream_t st);

// tcgen05 band kernel: needs the two TMA descriptors over the plane ring.
struct BandMaps;   // opaque: 2 x CUtensorMap
BandMaps* band_maps_create(const int8_t* planes, int64_t plane_rows, int32_t npad);
void band_maps_destroy(BandMaps*);
void launch_band_tc(const BandMaps* maps, const BandArgs& args, int sm_count, cudaStream_t st);
void launch_band_tc_wide(const BandMaps* maps, const BandArgs& args, int sm_count, cudaStream_t st);   // tiles carry b_row2
// blk_miss[b] != 0: 64-SNP block b holds a kept SNP with a missing genotype (written by stats_kernel).  One block:
// *flag = 1 if no block in [block0, block1) is marked, else 0.
void launch_chunk_mfree_flag(const uint8_t* blk_miss, int32_t block0, int32_t block1, uint32_t* flag, cudaStream_t st);
void install_band_kernels();   // per device: function attributes + module load of the band kernels (warm-up)
void preload_unpack_kernels(); // per device: force the (lazily loaded) images of the unpack.cu / plan.cu / band_simt.cu kernels in
void preload_plan_kernels();
void preload_finalize_kernels();
void launch_i8_peak(int iters, int n_cols, int per_commit, int depth, int sm_count, cudaStream_t st);   // measurement only
float launch_fp4_probe(int n_mma, int n_cols, int per_commit, int n_pairs, float* d_out, cudaStream_t st);   // feasibility probe, ms
void launch_band_simt(const int8_t* planes, int64_t plane_stride, int32_t npad,
                      const BandArgs& args, cudaStream_t st);

struct FinalizeArgs {
    const int32_t* tile_begin;  // [ceil(n_iblocks / 2) + 1] supertile range of every owned i-block pair
    int32_t iblock0;            // shard-local block index of the first owned block in this chunk
    int32_t n_iblocks;
    int32_t own_begin;          // shard-local first owned row
    int32_t own_end;
    Partials part;
    const uint8_t* flags;
    double* l2;                 // outputs, indexed by (local row - own_begin)
    double* l2d;
    int32_t* wsa;
    int32_t* wsd;
    int32_t* wsde;
};
void launch_finalize(const FinalizeArgs& a, cudaStream_t st);

}  // namespace nldsc
