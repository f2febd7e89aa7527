// Host-side shard state: plan (windows, halo, plane-ring chunks, band tiles) + device buffers.
#pragma once
#include "common.cuh"
#include "../../include/nldsc_ld.h"

#include <condition_variable>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

namespace nldsc {

struct InvalidArgument : std::invalid_argument {
    using std::invalid_argument::invalid_argument;
};

// Inclusive window [lo, hi] of every SNP in ORIGINAL .bim indices:
// {j : |pos_j - pos_i| <= wind}, the reference's SNPFilter::filter (tools.h:41-49)
// evaluated with its own expression.  Requires 0 <= pos, non-decreasing.
void compute_windows(const double* pos, int32_t n_snp, double wind, std::vector<int32_t>& lo,
                     std::vector<int32_t>& hi);

// A chunk = a run of owned i-blocks scored by one band launch.  SNP rows are unpacked into
// the plane ring exactly once, in order: chunk c adds rows [new_row0, new_row1).
struct Chunk {
    int32_t iblock0;      // first owned i-block (shard-local block index)
    int32_t n_iblocks;
    int32_t new_row0;     // shard-local rows this chunk adds to the ring (count, stats, unpack)
    int32_t new_row1;
    int64_t tile0;        // offset into the tile list
    int32_t n_tiles;
    int32_t tb0;          // offset into tile_begin (n_iblocks + 1 entries, relative to tile0)
    int64_t qtile0;       // missing-free supertiles of the same chunk (session::d_qtiles)
    int32_t n_qtiles;
    int64_t xtile0;       // 160-column supertiles of the same chunk (session::d_xtiles)
    int32_t n_xtiles;
    int32_t blk0, blk1;   // 64-SNP blocks [blk0, blk1) the chunk's tiles read: the range its missing-free flag covers
    int32_t slot_base;    // fused plans: first partial slot of the chunk (its tiles' slots already include it)
};

// Device workspace of one session: a bump allocator over a few large cudaMalloc blocks that are
// cached per device between calls, so a repeated nldsc_ld_calculate* pays no cudaMalloc / cudaFree
// (each a device-wide synchronisation, and tens of ms for the multi-GB plane ring).
struct Arena {
    struct Block { char* base; size_t size, used; };
    int device = -1;
    bool cached = false;          // blocks go back to the per-device cache on release
    std::vector<Block> blocks;
    void* alloc(size_t bytes);    // 512-byte aligned; grows by cudaMalloc when no block has room
    size_t capacity() const;
    size_t largest_free() const;  // the most one further allocation can take without a cudaMalloc
};
Arena* arena_acquire(int device);     // the device's cached arena, or a private one if that is in use
void arena_release(Arena* a);         // resets the bump pointers; private arenas free their blocks
size_t arena_cached_bytes(int device);
void arena_trim(int device);          // frees the cached blocks (device < 0: every device)

// Pinned staging ring for reading a .bed: kSlots slots the reader thread fills by pread and drains by
// cudaMemcpyAsync; cached per device like the Arena (pinning memory costs tens of ms).
struct HostStage {
    static constexpr int kReaders = 8;            // most reader threads a shard may use: one pread stream copies ~10 GB/s out of
                                                  // the page cache, PCIe takes 55; the default is min(8, max(2, cores / 2))
    static constexpr int kSlots = 2 * kReaders;   // two slots per reader
    int device = -1;
    bool cached = false;
    void* base = nullptr;
    size_t bytes = 0;             // kSlots equal slots
    cudaEvent_t ev[kSlots] = {};                  // slot's last H2D has landed
};
HostStage* host_stage_acquire(int device);
void host_stage_release(HostStage* h);

// Pinned staging for the small planning arrays (positions up, windows back, band plan up), moved by
// launch_copy16 instead of the copy engines; cached per device and grown on demand.
struct PlanStage {
    int device = -1;
    bool cached = false;
    char* base = nullptr;
    size_t bytes = 0;
    void reserve(size_t need);    // contents are lost when the buffer grows
};
PlanStage* plan_stage_acquire(int device);
void plan_stage_release(PlanStage* h);

enum Family { kFamCount = 0, kFamStats, kFamUnpack, kFamBand, kFamFinalize, kFamN };

}  // namespace nldsc

struct nldsc_ld_session {
    int device = 0;
    int sm_count = 0;
    int32_t n_org = 0, n_snp = 0;
    int32_t row_begin = 0, row_end = 0;       // owned rows, original indices
    int32_t load_begin = 0, load_end = 0;     // rows resident on the device, original indices
    int32_t n_load = 0, own_begin = 0, own_end = 0;   // shard-local
    int64_t bps = 0;
    int32_t npad = 0;                         // plane row bytes (int8 planes: = padded organisms; fp4 planes: half of them)
    double maf_thr = 0, std_thr = 0, rsq_thr = 0;
    bool use_simt = false;
    bool wide = false;                        // band_tc_wide_kernel: two j-blocks per supertile (the default; NLDSC_WIDE=0 = narrow kernel)
    bool mfree = false;                       // plan carries missing-free supertiles; chunks without a missing genotype take the 2-plane body of the band kernel
    nldsc::QTile* d_qtiles = nullptr;
    uint8_t* d_blk_miss = nullptr;            // [n_load_pad / 64] written by stats_kernel
    uint32_t* d_mfree_flags = nullptr;        // [chunks] 1 = the chunk is missing-free (launch_chunk_mfree_flag)
    std::vector<nldsc::QTile> h_qtiles;
    int64_t total_qtiles = 0;
    bool fp4 = false;                         // planes hold e2m1 nibbles; npad = plane row BYTES = organisms / 2, padded to 128
    bool x160 = false;                        // fp4 planes: chunks with missing genotypes take the 160-column supertiles (band_x160_body)
    nldsc::XTile* d_xtiles = nullptr;
    std::vector<nldsc::XTile> h_xtiles;
    bool fused_ok = false;                    // the plan allows ONE band launch per shard (session_plan); run() decides
    uint32_t* d_ready = nullptr;              // fused launch: ring rows [0, *d_ready) are unpacked, their statistics final
    bool upload_started = false;
    bool planned = false;

    std::vector<nldsc::Chunk> chunks;
    int64_t total_tiles = 0;
    int64_t total_pairs = 0;                  // sum over owned rows of (window size - 1) before the MAF filter
    int32_t max_chunk_tiles = 0;
    int32_t max_chunk_slots = 0;              // partial-sum slots (= supertiles of the narrow kernel) of the largest chunk
    int64_t plane_rows = 0;                   // ring capacity in SNP rows (multiple of 64)

    // device (all carved from `arena`)
    nldsc::Arena* arena = nullptr;
    uint8_t* d_packed = nullptr;
    int32_t* d_counts = nullptr;
    nldsc::SnpCoef* d_coef = nullptr;
    uint8_t* d_flags = nullptr;
    uint8_t* d_unused = nullptr;               // [n_load] 1 = position < 0 ("unused", tools.h:16-24); null when no SNP is
    std::vector<uint8_t> h_unused;
    double* d_maf = nullptr;
    double* d_rstd = nullptr;
    int32_t* d_win_lo = nullptr;
    int32_t* d_win_hi = nullptr;
    unsigned long long* d_total_pairs = nullptr;   // whole-file plan only: sum of window widths, summed by windows_kernel
    int8_t* d_planes = nullptr;
    nldsc::Tile* d_tiles = nullptr;
    int32_t* d_tile_begin = nullptr;
    nldsc::Partials part{};
    nldsc::Partials part2{};                  // odd chunks (second band stream)
    double* d_l2 = nullptr;
    double* d_l2d = nullptr;
    int32_t* d_wsa = nullptr;
    int32_t* d_wsd = nullptr;
    int32_t* d_wsde = nullptr;
    nldsc::BandMaps* maps = nullptr;
    cudaStream_t own_stream = nullptr;        // compute
    cudaStream_t copy_stream = nullptr;       // H2D of packed rows, piece by piece
    cudaStream_t unpack_stream = nullptr;     // unpack + stats of chunk c+1 run beside the band kernel of chunk c
    std::vector<cudaEvent_t> ev_unpacked;     // per chunk: planes, coefficients and flags are ready
    std::vector<cudaEvent_t> ev_band;         // per chunk: the band kernel no longer reads its ring rows
    cudaStream_t band2_stream = nullptr;      // band + finalize of odd chunks: the next launch fills the SMs the last wave of this one leaves idle
    cudaEvent_t ev_join = nullptr;
    cudaEvent_t ev_fork = nullptr;            // run() start on the compute stream, seen by the unpack stream
    int32_t piece_rows = 0;                   // rows per upload piece (multiple of 64)
    std::vector<cudaEvent_t> ev_piece;        // piece has landed in d_packed
    cudaEvent_t ev_planned = nullptr;         // windows / tiles have landed
    cudaEvent_t ev_up0 = nullptr, ev_up1 = nullptr;   // first / last byte of the upload (timing)
    // upload progress as seen by the host: run() may only wait on piece events that were recorded
    std::mutex up_mu;
    std::condition_variable up_cv;
    int32_t pieces_issued = 0;
    std::string up_error;
    std::vector<std::thread> readers;         // .bed readers (nldsc_ld_calculate only): piece k belongs to reader k % kReaders
    std::vector<char> piece_done;             // pieces_issued = length of the all-done prefix
    const char* bed_path = nullptr;           // .bed uploads: the path, for the optional O_DIRECT descriptor
    int fd_direct = -1;
    nldsc::HostStage* stage = nullptr;
    nldsc::PlanStage* pstage = nullptr;
    // host copies of the plan (sources of the asynchronous H2D copies)
    std::vector<int32_t> h_win_lo, h_win_hi, h_tile_begin;
    std::vector<nldsc::Tile> h_tiles;

    // timing of the last run(): event pairs per kernel family
    struct Span { cudaEvent_t a, b; int family; };
    std::vector<Span> spans;
    size_t spans_used = 0;
    cudaEvent_t ev_run0 = nullptr, ev_run1 = nullptr;
    int32_t launches = 0;
    bool ran = false;
};
