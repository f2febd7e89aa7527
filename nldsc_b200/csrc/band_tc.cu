// Banded int8 cross-count GEMM on the 5th-gen tensor cores + fused FP64 epilogue.
//
// Three bodies, one pipeline (TMA producer warp, one MMA-issuing thread per CTA pair, epilogue warps, mbarrier hand-offs):
//   band_tc_kernel        narrow: 2 i-blocks x 1 j-block, accumulator double-buffered        (NLDSC_WIDE=0; second implementation)
//   band_wide_body        wide:   2 i-blocks x 2 j-blocks, 6 cross-counts per pair            (the default)
//   band_mfree_body       4 i-blocks x 2 j-blocks, 2 cross-counts per pair, for chunks without a missing genotype
// The last two share the launch band_tc_wide_kernel; a per-chunk flag written on the device picks the body.  Both run on either
// operand format: int8 planes with tcgen05.mma.kind::i8, or (default, BandArgs::fp4) e2m1 planes with kind::mxf4 and unit scale
// factors -- twice the rate, the same exact integer counts.
//
// Replaces the pair loop of calculate() (reference ldscalc.h:34-47) and the two
// arma::dot calls per ordered pair in Math::r2_adjusted (tools.h:87-92).
//
// Persistent CTA PAIRS (cluster of 2 = the two SMs of a TPC, tcgen05 cta_group::2), one pair per
// TPC, walking a host-built list of supertiles that covers only the band |pos_j - pos_i| <= window.
// A supertile is two vertically adjacent 64-SNP i-blocks (one per CTA) x one 64-SNP j-block; per
// supertile one full-K (= organisms) int8 GEMM
//     D[256 x 192] (int32, TMEM: 128 lanes x 192 columns in each CTA) = A[256 x K] * B[192 x K]^T
//   A rows (per CTA, 128): per group of 16 i-SNPs, 16 rows of g (dosage 0/1/2) then 16 rows of m
//   B rows (192, half staged by each CTA): for j-SNPs 32r..32r+31 of the block, 32 rows of m_j,
//                                          32 of h_j (het), 32 of a_j (hom-second)      (r = CTA rank)
// gives the six exact cross-counts of every pair.  Sharing B between the two SMs cuts the operand
// bytes each SM pulls from L2 and reads from shared memory from 40 KB to 28 KB per 128-organism
// K-step, which is what bounded the one-CTA version (ncu: tensor pipe 46 %, L2 65 %).  Warp roles:
//   warp 0   TMA producer  (cp.async.bulk.tensor.cta_group::2, 128B-swizzled K-major tiles, 8 stages;
//                           both CTAs' bytes complete on the leader CTA's mbarrier)
//   warp 1   MMA issuer    (leader CTA only: tcgen05.mma.cta_group::2.kind::i8, M=256 N=192 K=32;
//                           commits multicast to both CTAs; 2 TMEM accumulators)
//   warps 2-5 epilogue     (tcgen05.ld -> FP64 correlations -> r2adj -> row sums), each CTA its own rows
// The accumulators are double-buffered in TMEM so the FP64 epilogue of tile t overlaps
// the MMAs of tile t+1.
#include "common.cuh"
#include "epilogue.cuh"

#include <cuda.h>
#include <mutex>
#include <type_traits>
#include <string>

#ifndef NLDSC_FP4_F2F
#define NLDSC_FP4_F2F 1      // fp4 accumulators -> double by one F2F each (0: float -> int -> double, the A/B build)
#endif

namespace nldsc {

namespace {

constexpr int kStages = 8;
constexpr int kKStep = 128;                 // organisms (bytes) per pipeline stage
constexpr int kARows = 128;                 // per CTA: 64 i-SNPs x {g, m}
constexpr int kBRows = 192;                 // per pair: 64 j-SNPs x {m, h, a}; each CTA stages half
constexpr int kBHalf = kBRows / 2;          // 96 rows = 32 j-SNPs x {m, h, a}
constexpr int kABytes = kARows * kKStep;    // 16 KB
constexpr int kBBytes = kBHalf * kKStep;    // 12 KB
constexpr int kStageBytes = kABytes + kBBytes;   // 28 KB per CTA per stage
constexpr int kTmemCols = 512;              // 2 accumulators of 192 columns at 0 and 256
constexpr int kThreads = 192;
constexpr int kSmemBytes = kStages * kStageBytes + 1024 /*align*/ + 256 /*barriers*/;
constexpr uint32_t kPeerMask = 0xFEFFFFFFu; // shared::cluster address of the same offset in the pair's even CTA

// ---- PTX wrappers -------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
// A protocol bug must end the kernel with a diagnostic, not hang the GPU box -- but a legitimately long wait (under
// cuda-gdb, compute-sanitizer, an ncu replay or heavy co-tenancy) must not be mistaken for one: the bound is
// ELAPSED TIME (%globaltimer, kHangNs), not a spin count.  Before trapping, the waiter leaves
// (site << 24 | block) in a word of mapped host memory, which the host appends to its error message.
constexpr unsigned long long kHangNs = 20ull * 1000000000ull;
__device__ uint32_t* g_hang_word = nullptr;      // set per device by band_set_hang_word()
__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
__device__ __noinline__ void hang_trap(uint32_t site) {
    if (g_hang_word != nullptr) {
        *reinterpret_cast<volatile uint32_t*>(g_hang_word) = (site << 24) | (blockIdx.x & 0xFFFFFFu) | 0x800000u;
        __threadfence_system();
    }
    __trap();
}
__device__ __forceinline__ uint32_t mbar_try(uint32_t addr, uint32_t parity) {
    uint32_t done;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(addr), "r"(parity)
        : "memory");
    return done;
}
// out of line, so that its clock and counters take no registers from the callers' hot loops
__device__ __noinline__ void mbar_wait_slow(uint32_t addr, uint32_t parity, uint32_t site, uint32_t sleep_ns) {
    const unsigned long long t0 = global_ns();
    for (uint32_t spins = 0;; ++spins) {
        if (mbar_try(addr, parity)) return;
        if (sleep_ns) __nanosleep(sleep_ns);
        if ((spins & 0x3FFu) == 0x3FFu && global_ns() - t0 > kHangNs) hang_trap(site);
    }
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity, uint32_t site = 0) {
    const uint32_t addr = smem_u32(bar);
    for (uint32_t spins = 0; spins < (1u << 16); ++spins)
        if (mbar_try(addr, parity)) return;
    mbar_wait_slow(addr, parity, site, 0);
}
// The same wait for warps that have nothing else to do for a long time (the epilogue warps during a supertile's
// K loop): sleep between polls, so that eight polling warps do not take issue slots from the one thread that
// issues the MMAs and the one that issues the TMA loads.
__device__ __forceinline__ void mbar_wait_idle(uint64_t* bar, uint32_t parity, uint32_t site = 0) {
    const uint32_t addr = smem_u32(bar);
    for (uint32_t spins = 0; spins < (1u << 14); ++spins) {
        if (mbar_try(addr, parity)) return;
        __nanosleep(256);
    }
    mbar_wait_slow(addr, parity, site, 256);
}
// Fused launch: block until ring rows [0, need) are unpacked and their statistics final (BandArgs::ready_rows, published
// by launch_publish_rows after every chunk).  `seen` caches the last value read, so a tile whose rows are already known
// to be there costs nothing.  A wait can legitimately last as long as the upload of the shard (seconds from a cold disk),
// so its bound is 10 x the barrier bound; a failed upload publishes 0xFFFFFFFF, which releases every waiter.
__device__ __noinline__ uint32_t wait_rows_slow(const uint32_t* p, uint32_t need) {
    const unsigned long long t0 = global_ns();
    for (uint32_t spins = 0;; ++spins) {
        uint32_t v;
        asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
        if (v >= need) return v;
        __nanosleep(400);
        if ((spins & 0xFFu) == 0xFFu && global_ns() - t0 > 10 * kHangNs) hang_trap(10);
    }
}
__device__ __forceinline__ void wait_rows(const uint32_t* p, uint32_t need, uint32_t& seen) {
    if (p == nullptr || seen >= need) return;
    seen = wait_rows_slow(p, need);
}
// rows a supertile reads: its i-blocks and its j-block(s)
__device__ __forceinline__ uint32_t tile_rows_needed(const Tile& tl, bool two) {
    int32_t need = tl.i0[0] + kBlk;
    if (tl.i0[1] != kDeadBlock) need = max(need, tl.i0[1] + kBlk);
    return static_cast<uint32_t>(max(need, tl.j0 + (two ? 2 : 1) * kBlk));
}

// remote arrive on the leader (even) CTA's copy of `bar`
__device__ __forceinline__ void mbar_arrive_leader(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(smem_u32(bar) & kPeerMask) : "memory");
}
// Executed by both CTAs of the pair; the bytes land in the issuing CTA's shared memory and
// complete_tx on the LEADER CTA's barrier (peer bit cleared), which the MMA issuer waits on.
__device__ __forceinline__ void tma_load_3d(void* dst, const CUtensorMap* map, uint64_t* bar,
                                            int32_t c0, int32_t c1, int32_t c2) {
    asm volatile(
        "cp.async.bulk.tensor.3d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4, %5}], [%2];"
        ::"r"(smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar) & kPeerMask),
          "r"(c0), "r"(c1), "r"(c2)
        : "memory");
}
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
// arrive on the same barrier offset in BOTH CTAs once all prior MMAs of the pair have retired
__device__ __forceinline__ void tc_commit_pair(uint64_t* bar) {
    asm volatile(
        "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
        ::"r"(smem_u32(bar)), "h"(static_cast<uint16_t>(3))
        : "memory");
}
// the same commit arriving on the executing (leader) CTA's barrier only
__device__ __forceinline__ void tc_commit_leader(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// D[tmem, both CTAs] (+)= A[smem, 128 rows per CTA] * B[smem, 96 rows per CTA]^T, int8 x int8 -> int32
__device__ __forceinline__ void tc_mma_i8(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                          uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::i8 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// K-major, 128-byte swizzle: rows at a 128 B pitch, 8-row atoms 1024 B apart.
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t smem_addr) {
    uint64_t d = 0;
    d |= static_cast<uint64_t>((smem_addr & 0x3FFFFu) >> 4);   // start address
    d |= static_cast<uint64_t>(1024u >> 4) << 32;              // stride byte offset (8 rows)
    d |= 1ull << 46;                                           // descriptor version (sm_100)
    d |= 2ull << 61;                                           // SWIZZLE_128B
    return d;
}
// instruction descriptor, kind::i8: D = s32, A/B = u8, both K-major, M = 256 (pair), N = 192
constexpr uint32_t kIdesc = (2u << 4) | (0u << 7) | (0u << 10) | ((kBRows >> 3) << 17) | (((2 * kARows) >> 4) << 24);

__device__ __forceinline__ void tmem_ld16(uint32_t taddr, int32_t (&v)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
          "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
        : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// ---- fp4 (mxf4) operands: the same cross-counts at twice the int8 rate -------------------------------------------
// Genotype planes hold 0 / 1 / 2, exact in e2m1; with every scale factor 2^0 tcgen05.mma.kind::mxf4 multiplies 64
// organisms per instruction (32 operand bytes per row, as kind::i8 does for 32) and accumulates the small integer block
// sums exactly in FP32 as long as a count stays below 2^24 (measured: nldsc_ld_debug_fp4_probe, 150,000 accumulating MMAs
// to 14,400,000 exact).  The scale factors live in TMEM; the kernels store 1.0 everywhere in columns 448..479 once.
__device__ __forceinline__ void tc_mma_mxf4(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t sfa,
                                            uint32_t sfb, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::mxf4.block_scale.block32 [%0], %1, %2, %3, [%5], [%6], p;\n\t}"
        ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate), "r"(sfa), "r"(sfb)
        : "memory");
}
constexpr uint32_t kSfCols = 448;   // TMEM columns 448..479 hold scale factors 1.0 for every lane: any SF layout reads ones

__device__ __forceinline__ void tmem_fill_sf_ones(uint32_t tmem_base, int warp_in_quarter_order, uint32_t sf_cols = kSfCols) {
    // 32 lanes of this warp's quarter x 32 columns of 0x7F7F7F7F (four E8M0 exponents of 2^0 per 32-bit column)
    const uint32_t taddr = tmem_base + sf_cols + (static_cast<uint32_t>(32 * (warp_in_quarter_order & 3)) << 16);
    const uint32_t v = 0x7F7F7F7Fu;
#pragma unroll
    for (int c = 0; c < 32; c += 16)
        asm volatile(
            "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1};"
            ::"r"(taddr + c), "r"(v)
            : "memory");
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
}

// FP32 accumulator (an exact integer below 2^22) -> int32 without the conversion pipe: add 1.5 * 2^23, read the mantissa
__device__ __forceinline__ int32_t acc_f2i(int32_t bits) {
    return __float_as_int(__int_as_float(bits) + 12582912.0f) - 0x4B400000;
}
__device__ __forceinline__ void acc_to_int(int32_t (&v)[16]) {
#pragma unroll
    for (int e = 0; e < 16; ++e) v[e] = acc_f2i(v[e]);
}
// `g_lane ? a : b` in ONE instruction: PRMT with a per-lane selector (0x3210 picks a, 0x7654 picks b).  Written as a ternary
// the compiler emits a move and a predicated move, and the 6-plane epilogues -- bound by instruction issue -- do nine of these
// per SNP pair (which three counts go to the partner lane, and which of own / received is the g row and which the m row).
// Used by the 160-column body only (-1.2 % per step at N=10,000); the 2 x 2-block body at N=2,504 was 1 % SLOWER with it
// (profiles/r02zk_ab_prmt_*.json), so the bodies that serve small N keep the ternaries.
__device__ __forceinline__ int32_t pick(uint32_t lane_sel, int32_t a, int32_t b) {
    return static_cast<int32_t>(__byte_perm(static_cast<uint32_t>(a), static_cast<uint32_t>(b), lane_sel));
}
// instruction descriptor, block-scaled (cute/arch/mma_sm100_desc.hpp InstrDescriptorBlockScaled): A / B = e2m1 (1 at bits
// 7 and 10), both K-major, N >> 3 at bit 17, scale format E8M0 (bit 23), M >> 4 at bit 24, K = 64 (bit 31 = 0)
__device__ __forceinline__ constexpr uint32_t idesc_mxf4(uint32_t n_cols) {
    return (1u << 7) | (1u << 10) | ((n_cols >> 3) << 17) | (1u << 23) | ((256u >> 4) << 24);
}

}  // namespace

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
band_tc_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
               const BandArgs args) {
    extern __shared__ uint8_t smem_raw[];
    // identical offsets in both CTAs of the pair: the MMA addresses the peer's operands and the
    // multicast commits address the peer's barriers by the leader's own offsets
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + kStages * kStageBytes);   // used in the leader CTA only
    uint64_t* empty = full + kStages;          // one per CTA
    uint64_t* tfull = empty + kStages;         // one per CTA
    uint64_t* tempty = tfull + 2;              // used in the leader CTA only
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();               // 0 = leader (issues the MMAs)
    const int pair = blockIdx.x >> 1, n_pairs = gridDim.x >> 1;

    if (warp == 0 && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&map_a)) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&map_b)) : "memory");
        for (int s = 0; s < kStages; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
        for (int b = 0; b < 2; ++b) { mbar_init(&tfull[b], 1); mbar_init(&tempty[b], 8); }   // 4 epilogue warps x 2 CTAs
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                     "r"((uint32_t)kTmemCols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    cluster_sync_all();                        // peer barriers initialised before any remote arrive / complete_tx
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===== TMA producer (both CTAs: own A block, own half of B) =====
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int t = pair; t < args.n_tiles; t += n_pairs) {
                const Tile tl = args.tiles[t];
                const int32_t a_row = tl.a_row[rank];
                const int32_t b_row = tl.b_row + 32 * static_cast<int32_t>(rank);
                for (int k = 0; k < args.k_steps; ++k) {
                    mbar_wait(&empty[stage], phase ^ 1, 1);
                    if (args.probe == 1 || args.probe == 3) {   // no loads: operands are whatever the stage holds
                        if (rank == 0) mbar_expect_tx(&full[stage], 0);
                        if (++stage == kStages) { stage = 0; phase ^= 1; }
                        continue;
                    }
                    if (rank == 0) mbar_expect_tx(&full[stage], 2 * kStageBytes);    // both CTAs' bytes
                    uint8_t* sa = smem + stage * kStageBytes;
                    uint8_t* sb = sa + kABytes;
#pragma unroll
                    for (int u = 0; u < 4; ++u)   // 16 SNPs x {g, m} = 32 rows per box
                        tma_load_3d(sa + u * 4096, &map_a, &full[stage], k * kKStep, a_row + 16 * u, kPlaneG);
                    tma_load_3d(sb, &map_b, &full[stage], k * kKStep, b_row, kPlaneM);   // 32 SNPs x {m, h, a}
                    if (++stage == kStages) { stage = 0; phase ^= 1; }
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ===== MMA issuer (leader CTA only) =====
        if (rank == 0 && lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            int n = 0;
            for (int t = pair; t < args.n_tiles; t += n_pairs, ++n) {
                const int buf = n & 1;
                mbar_wait(&tempty[buf], ((n >> 1) & 1) ^ 1, 2);       // both CTAs' epilogues drained this accumulator
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + buf * 256;
                for (int k = 0; k < args.k_steps; ++k) {
                    mbar_wait(&full[stage], phase, 3);
                    tc_fence_after();
                    const uint32_t sa = smem_u32(smem + stage * kStageBytes);
                    const uint64_t a_desc = make_smem_desc(sa);
                    const uint64_t b_desc = make_smem_desc(sa + kABytes);
                    if (args.probe < 2) {
#pragma unroll
                        for (int kk = 0; kk < kKStep / 32; ++kk)   // +32 B along K inside the swizzle atom
                            tc_mma_i8(d_tmem, a_desc + 2 * kk, b_desc + 2 * kk, kIdesc, (k | kk) != 0);
                    }
                    tc_commit_pair(&empty[stage]);                  // both CTAs' smem slots free once these MMAs retire
                    if (++stage == kStages) { stage = 0; phase ^= 1; }
                }
                tc_commit_pair(&tfull[buf]);                        // accumulator complete, in both CTAs
            }
        }
        __syncwarp();
    } else {
        // ===== epilogue: 4 warps per CTA, warp q owns TMEM lanes 32q..32q+31 of its CTA's i-block =====
        const int q = warp & 3;
        const bool is_g = lane < 16;                 // lanes 0-15 hold g rows, 16-31 the m rows of the same SNPs
        const int s = 16 * q + (lane & 15);          // SNP within the i-block
        int n = 0;
        for (int t = pair; t < args.n_tiles; t += n_pairs, ++n) {
            const int buf = n & 1;
            const Tile tl = args.tiles[t];
            const int32_t i0 = tl.i0[rank];
            const RowCtx row = load_row_ctx(args, i0 + s);
            mbar_wait(&tfull[buf], (n >> 1) & 1, 4);
            tc_fence_after();
            const uint32_t taddr = tmem_base + buf * 256 + (static_cast<uint32_t>(32 * q) << 16);
#pragma unroll 1
            for (int half = 0; half < 2; ++half) {   // j-SNPs 0-31 (staged by CTA 0), then 32-63 (CTA 1): one partial each
                RowAcc acc;
#pragma unroll 1
                for (int c2 = 0; c2 < 2; ++c2) {      // 16 j-SNPs per pass
                    const int c = 2 * half + c2;
                    if (args.dump == nullptr && !__any_sync(0xFFFFFFFFu, row_needs(row, tl.j0 + 16 * c, tl.j0 + 16 * c + 15))) continue;
                    // columns: [m h a] of j-SNPs 0-31, then [m h a] of j-SNPs 32-63
                    const uint32_t col = 96 * half + 16 * c2;
                    int32_t vm[16], vh[16], va[16];
                    tmem_ld16(taddr + col, vm);
                    tmem_ld16(taddr + col + 32, vh);
                    tmem_ld16(taddr + col + 64, va);
                    tmem_ld_wait();
                    if (args.dump != nullptr) {
                        int32_t* drow = args.dump + ((static_cast<int64_t>(tl.slot) * 2 + rank) * kARows + 32 * q + lane) * kBRows;
#pragma unroll
                        for (int e = 0; e < 16; ++e) {
                            drow[16 * c + e] = vm[e];
                            drow[64 + 16 * c + e] = vh[e];
                            drow[128 + 16 * c + e] = va[e];
                        }
                    }
#pragma unroll
                    for (int u = 0; u < 8; ++u) {
                        // g-lane scores even j, its partner m-lane scores odd j; swap what the other needs
                        const int32_t sm_ = is_g ? vm[2 * u + 1] : vm[2 * u];
                        const int32_t sh_ = is_g ? vh[2 * u + 1] : vh[2 * u];
                        const int32_t sa_ = is_g ? va[2 * u + 1] : va[2 * u];
                        const int32_t rm = __shfl_xor_sync(0xFFFFFFFFu, sm_, 16);
                        const int32_t rh = __shfl_xor_sync(0xFFFFFFFFu, sh_, 16);
                        const int32_t ra = __shfl_xor_sync(0xFFFFFFFFu, sa_, 16);
                        const int jl = 16 * c + 2 * u + (is_g ? 0 : 1);
                        const int32_t Ggm = is_g ? vm[2 * u] : rm, Gmm = is_g ? rm : vm[2 * u + 1];
                        const int32_t Ggh = is_g ? vh[2 * u] : rh, Gmh = is_g ? rh : vh[2 * u + 1];
                        const int32_t Gga = is_g ? va[2 * u] : ra, Gma = is_g ? ra : va[2 * u + 1];
                        pair_eval(args, row, tl.j0 + jl, Ggm, Ggh, Gga, Gmm, Gmh, Gma, acc);
                    }
                }
                if (half == 1) {
                    // accumulator fully read: hand it back to the leader's MMA warp
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive_leader(&tempty[buf]);
                }
                // even-j and odd-j halves of the row
                acc.add += __shfl_xor_sync(0xFFFFFFFFu, acc.add, 16);
                acc.dom += __shfl_xor_sync(0xFFFFFFFFu, acc.dom, 16);
                acc.wsa += __shfl_xor_sync(0xFFFFFFFFu, acc.wsa, 16);
                acc.wsd += __shfl_xor_sync(0xFFFFFFFFu, acc.wsd, 16);
                acc.wsde += __shfl_xor_sync(0xFFFFFFFFu, acc.wsde, 16);
                if (is_g) {
                    const int64_t o = part_index(tl.slot, static_cast<int>(rank), half, s);
                    args.part.add[o] = acc.add;
                    args.part.dom[o] = acc.dom;
                    args.part.wsa[o] = acc.wsa;
                    args.part.wsd[o] = acc.wsd;
                    args.part.wsde[o] = acc.wsde;
                }
            }
        }
    }

    // nobody leaves while the peer may still read this CTA's operands or signal its barriers
    tc_fence_before();
    cluster_sync_all();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)kTmemCols)
                     : "memory");
    }
}

// ---- wide supertiles: two j-blocks per A tile ---------------------------------------------------
// Same roles and pipeline as band_tc_kernel, but a supertile may carry a SECOND j-block (Tile::b_row2): the A
// tiles of a K-step are then multiplied with both B tiles, D0[256 x 192] in TMEM columns 0..191 and D1 in 192..383.
// Operand bytes from L2 drop from 56 KB to 40 KB per 64 columns of a CTA pair, which lifts the L2 -> SM bound the
// narrow kernel runs into.  Both accumulators belong to one supertile, so the epilogue is NOT hidden behind the
// next supertile's MMAs; eight epilogue warps (two per 32 TMEM lanes, one per half of a j-block) keep it to
// 2 x 3 us per supertile.  Measured faster than the narrow kernel at every N from 2,504 up: the default.
// Partials of the two j-blocks go to slots `slot` and `slot + 1`, exactly where the narrow kernel puts them, so both
// kernels give bitwise identical scores.
namespace {
constexpr int kWStages = 5;
constexpr int kWStageBytes = kABytes + 2 * kBBytes;          // 40 KB per CTA per stage
constexpr int kWCols = 2 * kBlk;                             // j-SNPs of a wide supertile
constexpr int kWColBytes = kWCols * static_cast<int>(sizeof(SnpCoef)) + kWCols;   // their SnpCoef + flags, staged for the epilogue
constexpr int kWSmemBytes = kWStages * kWStageBytes + 1024 + 256 + kWColBytes;
constexpr int kWThreads = 320;                               // TMA warp, MMA warp, 8 epilogue warps
}  // namespace

// 128 registers x 320 threads leave a third of the register file to the unpack blocks that run beside this kernel
// (at the 162 ptxas would take, only one unpack block fits per SM and the unpack becomes the critical path)
__device__ __forceinline__ void band_wide_body(const CUtensorMap& map_a, const CUtensorMap& map_b, const BandArgs& args) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + kWStages * kWStageBytes);   // leader CTA only
    uint64_t* empty = full + kWStages;         // one per CTA
    uint64_t* tfull = empty + kWStages;        // one per CTA
    uint64_t* tempty = tfull + 1;              // leader CTA only: one per accumulator
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);
    // the j-SNPs' constants of the supertile being scored: staged by the epilogue warps while they wait for its MMAs
    double2* col_coef = reinterpret_cast<double2*>(smem + kWStages * kWStageBytes + 256);
    uint8_t* col_flags = reinterpret_cast<uint8_t*>(col_coef) + kWCols * sizeof(SnpCoef);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    const int pair = blockIdx.x >> 1, n_pairs = gridDim.x >> 1;

    if (warp == 0 && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&map_a)) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&map_b)) : "memory");
        for (int s = 0; s < kWStages; ++s) { mbar_init(&full[s], args.probe == 1 ? 2 : 1); mbar_init(&empty[s], 1); }
        mbar_init(tfull, 1);
        mbar_init(&tempty[0], 16);             // 8 epilogue warps x 2 CTAs, per accumulator
        mbar_init(&tempty[1], 16);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                     "r"((uint32_t)kTmemCols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    if (args.fp4) {                            // scale factors 1.0 for every lane of both CTAs, before the first MMA
        if (warp >= 2 && warp < 6) tmem_fill_sf_ones(tmem_base, warp);
        tc_fence_before();
        cluster_sync_all();
        tc_fence_after();
    }

    if (warp == 0) {
        // ===== TMA producer: own A block, own half of one or two B blocks =====
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0, seen = 0;
            for (int t = pair; t < args.n_tiles; t += n_pairs) {
                const Tile tl = args.tiles[t];
                const int32_t a_row = tl.a_row[rank];
                const int32_t b_row = tl.b_row + 32 * static_cast<int32_t>(rank);
                const bool two = tl.b_row2 >= 0;
                const int32_t b_row2 = tl.b_row2 + 32 * static_cast<int32_t>(rank);
                if (args.ready_rows != nullptr && seen < tile_rows_needed(tl, two)) {
                    wait_rows(args.ready_rows, tile_rows_needed(tl, two), seen);
                    asm volatile("fence.proxy.async;" ::: "memory");      // the rows were written through the generic proxy
                }
                const uint32_t bytes = 2u * (kABytes + (two ? 2 : 1) * kBBytes);          // both CTAs' bytes
                for (int k = 0; k < args.k_steps; ++k) {
                    mbar_wait(&empty[stage], phase ^ 1, 5);
                    if (args.probe == 1) {           // tensor-pipe probe: no loads, operands are whatever the stage holds
                        // the full barrier must still wait for BOTH producers (as the bytes of both make it do normally):
                        // a peer that is not part of the handshake falls phases behind its empty barriers and loses them
                        if (rank == 0) mbar_expect_tx(&full[stage], 0); else mbar_arrive_leader(&full[stage]);
                        if (++stage == kWStages) { stage = 0; phase ^= 1; }
                        continue;
                    }
                    if (rank == 0) mbar_expect_tx(&full[stage], bytes);
                    uint8_t* sa = smem + stage * kWStageBytes;
                    uint8_t* sb = sa + kABytes;
#pragma unroll
                    for (int u = 0; u < 4; ++u)
                        tma_load_3d(sa + u * 4096, &map_a, &full[stage], k * kKStep, a_row + 16 * u, kPlaneG);
                    tma_load_3d(sb, &map_b, &full[stage], k * kKStep, b_row, kPlaneM);
                    if (two) tma_load_3d(sb + kBBytes, &map_b, &full[stage], k * kKStep, b_row2, kPlaneM);
                    if (++stage == kWStages) { stage = 0; phase ^= 1; }
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ===== MMA issuer (leader CTA only) =====
        if (rank == 0 && lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            int n = 0;
            // the MMAs of one stage into accumulator 0 (j-block 0) and / or accumulator 1 (j-block 1)
            auto issue = [&](int st, int k, bool acc0, bool acc1) {
                const uint32_t sa = smem_u32(smem + st * kWStageBytes);
                const uint64_t a_desc = make_smem_desc(sa);
                const uint64_t b_desc = make_smem_desc(sa + kABytes);
                const uint64_t b2_desc = make_smem_desc(sa + kABytes + kBBytes);
                if (args.fp4) {            // 128 operand bytes per row = 256 organisms per stage, 64 per MMA
                    const uint32_t sf = tmem_base + kSfCols;
                    if (acc0) {
#pragma unroll
                        for (int kk = 0; kk < kKStep / 32; ++kk)
                            tc_mma_mxf4(tmem_base, a_desc + 2 * kk, b_desc + 2 * kk, idesc_mxf4(kBRows), sf, sf + 16, (k | kk) != 0);
                    }
                    if (acc1) {
#pragma unroll
                        for (int kk = 0; kk < kKStep / 32; ++kk)
                            tc_mma_mxf4(tmem_base + kBRows, a_desc + 2 * kk, b2_desc + 2 * kk, idesc_mxf4(kBRows), sf, sf + 16, (k | kk) != 0);
                    }
                } else {
                    if (acc0) {
#pragma unroll
                        for (int kk = 0; kk < kKStep / 32; ++kk)
                            tc_mma_i8(tmem_base, a_desc + 2 * kk, b_desc + 2 * kk, kIdesc, (k | kk) != 0);
                    }
                    if (acc1) {
#pragma unroll
                        for (int kk = 0; kk < kKStep / 32; ++kk)
                            tc_mma_i8(tmem_base + kBRows, a_desc + 2 * kk, b2_desc + 2 * kk, kIdesc, (k | kk) != 0);
                    }
                }
            };
            for (int t = pair; t < args.n_tiles; t += n_pairs, ++n) {
                const bool two = args.tiles[t].b_row2 >= 0;
                mbar_wait(&tempty[0], (n & 1) ^ 1, 6);               // both CTAs' epilogues drained accumulator 0
                tc_fence_after();
                int k = 0;
                if (two && args.runahead) {
                    // The epilogue of the previous supertile is still reading accumulator 1 (it is exposed: both accumulators
                    // belong to one supertile).  Run ahead into accumulator 0: the stages the producer has already filled are
                    // multiplied with j-block 0 now, and with j-block 1 -- and released -- once accumulator 1 is free.
                    const int ahead = args.k_steps < kWStages - 1 ? args.k_steps : kWStages - 1;
                    int st = stage;
                    uint32_t ph = phase;
                    for (int kk = 0; kk < ahead; ++kk) {
                        mbar_wait(&full[st], ph, 7);
                        tc_fence_after();
                        issue(st, kk, true, false);
                        if (++st == kWStages) { st = 0; ph ^= 1; }
                    }
                    mbar_wait(&tempty[1], (n & 1) ^ 1, 6);
                    tc_fence_after();
                    for (; k < ahead; ++k) {
                        issue(stage, k, false, true);
                        tc_commit_pair(&empty[stage]);
                        if (++stage == kWStages) { stage = 0; phase ^= 1; }
                    }
                } else {
                    mbar_wait(&tempty[1], (n & 1) ^ 1, 6);
                    tc_fence_after();
                }
                for (; k < args.k_steps; ++k) {
                    mbar_wait(&full[stage], phase, 7);
                    tc_fence_after();
                    issue(stage, k, true, two);
                    if (!args.commit2) {
                        tc_commit_pair(&empty[stage]);
                    } else if ((k & 1) || k == args.k_steps - 1) {
                        // two stages released together: 16 MMAs between commits instead of 8
                        if (k & 1) tc_commit_pair(&empty[stage == 0 ? kWStages - 1 : stage - 1]);
                        tc_commit_pair(&empty[stage]);
                    }
                    if (++stage == kWStages) { stage = 0; phase ^= 1; }
                }
                tc_commit_pair(tfull);
            }
        }
        __syncwarp();
    } else {
        // ===== epilogue: EIGHT warps per CTA.  Warps 2-5 score j-SNPs 0..31 of every j-block (TMEM columns of the
        // half CTA 0 staged), warps 6-9 j-SNPs 32..63; warp w reads TMEM lanes 32 (w % 4) .. +31 as the hardware
        // demands.  Each writes the partial of its half, so the exposed epilogue is half as long. =====
        const int q = warp & 3;
        const int half = (warp - 2) >> 2;
        const bool is_g = lane < 16;
        const int s = 16 * q + (lane & 15);
        int n = 0;
        uint32_t seen = 0;
        for (int t = pair; t < args.n_tiles; t += n_pairs, ++n) {
            const Tile tl = args.tiles[t];
            const int32_t i0 = tl.i0[rank];
            const int n_j = tl.b_row2 >= 0 ? 2 : 1;
            // fused launch: the per-SNP coefficients / flags of the tile's rows are final only once its chunk's stats ran
            wait_rows(args.ready_rows, tile_rows_needed(tl, n_j == 2), seen);
            const RowCtx row = load_row_ctx(args, i0 + s);
            // Stage the constants of the supertile's 64 or 128 j-SNPs (64 bytes + 1 flag byte each) in shared memory
            // NOW, while the MMAs run: in the epilogue every pair needs them, and fetched pair by pair from L2 their
            // latency was what the (exposed) epilogue spent most of its time on.  Named barrier 1 = the 8 epilogue
            // warps: nobody overwrites the staging area while a slower warp still scores the previous supertile.
            asm volatile("bar.sync 1, 256;" ::: "memory");
            {
                const int te = threadIdx.x - 64;                       // 0..255 among the epilogue threads
                const int n_cols = n_j * kBlk;
                for (int e = te; e < n_cols * 4; e += 256) {           // 4 double2 per SNP
                    const int32_t gj = tl.j0 + (e >> 2);
                    double2 v = make_double2(0.0, 0.0);
                    if (gj < args.n_load) v = __ldg(reinterpret_cast<const double2*>(args.coef + gj) + (e & 3));
                    col_coef[e] = v;
                }
                if (te < n_cols) {
                    const int32_t gj = tl.j0 + te;
                    col_flags[te] = gj < args.n_load ? __ldg(args.flags + gj) : uint8_t(0);
                }
            }
            asm volatile("bar.sync 1, 256;" ::: "memory");
            mbar_wait_idle(tfull, n & 1, 8);
            tc_fence_after();
#pragma unroll 1
            for (int jj = 0; jj < n_j; ++jj) {
                RowAcc acc;
                const uint32_t taddr = tmem_base + jj * kBRows + (static_cast<uint32_t>(32 * q) << 16);
                const int32_t j0 = tl.j0 + kBlk * jj;
#pragma unroll 1
                for (int c2 = 0; c2 < 2; ++c2) {
                    const int c = 2 * half + c2;
                    // the 16 j-SNPs of this pass lie outside the window of every row of the warp (the corners of the
                    // first / last supertiles of an i-pair, a fifth of all pair slots at a 1 cM window): nothing to add
                    if (!args.no_skip && !__any_sync(0xFFFFFFFFu, row_needs(row, j0 + 16 * c, j0 + 16 * c + 15))) continue;
                    const uint32_t col = 96 * half + 16 * c2;
                    int32_t vm[16], vh[16], va[16];
                    tmem_ld16(taddr + col, vm);
                    tmem_ld16(taddr + col + 32, vh);
                    tmem_ld16(taddr + col + 64, va);
                    tmem_ld_wait();
#if !NLDSC_FP4_F2F
                    if (args.fp4) { acc_to_int(vm); acc_to_int(vh); acc_to_int(va); }    // FP32 accumulators holding exact counts
#endif
                    auto score8 = [&](auto f32) {
                    constexpr bool kF32 = decltype(f32)::value;
#pragma unroll
                    for (int u = 0; u < 8; ++u) {
                        const int32_t sm_ = is_g ? vm[2 * u + 1] : vm[2 * u];
                        const int32_t sh_ = is_g ? vh[2 * u + 1] : vh[2 * u];
                        const int32_t sa_ = is_g ? va[2 * u + 1] : va[2 * u];
                        const int32_t rm = __shfl_xor_sync(0xFFFFFFFFu, sm_, 16);
                        const int32_t rh = __shfl_xor_sync(0xFFFFFFFFu, sh_, 16);
                        const int32_t ra = __shfl_xor_sync(0xFFFFFFFFu, sa_, 16);
                        const int jl = 16 * c + 2 * u + (is_g ? 0 : 1);
                        const int32_t Ggm = is_g ? vm[2 * u] : rm, Gmm = is_g ? rm : vm[2 * u + 1];
                        const int32_t Ggh = is_g ? vh[2 * u] : rh, Gmh = is_g ? rh : vh[2 * u + 1];
                        const int32_t Gga = is_g ? va[2 * u] : ra, Gma = is_g ? ra : va[2 * u + 1];
                        const double2* cj = col_coef + 4 * (kBlk * jj + jl);
                        pair_eval_core<kF32>(args, row, j0 + jl, col_flags[kBlk * jj + jl], cj[0], cj[1], cj[2], cj[3],
                                             Ggm, Ggh, Gga, Gmm, Gmh, Gma, acc);
                    }
                    };
#if NLDSC_FP4_F2F
                    // fp4: the shuffled values are FP32 bit patterns, converted to double pair by pair (cnt2d)
                    if (args.fp4) score8(std::true_type{}); else score8(std::false_type{});
#else
                    score8(std::false_type{});
#endif
                }
                {
                    // this warp is done with accumulator jj (a one-j-block supertile never wrote the other): tell the leader's MMA warp
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) {
                        mbar_arrive_leader(&tempty[jj]);
                        if (n_j == 1) mbar_arrive_leader(&tempty[1]);
                    }
                }
                acc.add += __shfl_xor_sync(0xFFFFFFFFu, acc.add, 16);
                acc.dom += __shfl_xor_sync(0xFFFFFFFFu, acc.dom, 16);
                acc.wsa += __shfl_xor_sync(0xFFFFFFFFu, acc.wsa, 16);
                acc.wsd += __shfl_xor_sync(0xFFFFFFFFu, acc.wsd, 16);
                acc.wsde += __shfl_xor_sync(0xFFFFFFFFu, acc.wsde, 16);
                if (is_g) {
                    const int64_t o = part_index(tl.slot + jj, static_cast<int>(rank), half, s);
                    args.part.add[o] = acc.add;
                    args.part.dom[o] = acc.dom;
                    args.part.wsa[o] = acc.wsa;
                    args.part.wsd[o] = acc.wsd;
                    args.part.wsde[o] = acc.wsde;
                }
            }
        }
    }

    tc_fence_before();
    cluster_sync_all();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)kTmemCols)
                     : "memory");
    }
}
// ---- missing-free supertiles: g x {h, a} only -----------------------------------------------------
// When no kept SNP of a chunk's column range has a missing genotype, the four cross-counts with a missing-indicator
// factor are identically zero and each ordered pair needs only G_gh and G_ga -- the reference's own two dot products
// (tools.h:87-92; SURVEY.md section 8d: 4*N int8 ops per pair instead of 12*N).  Same roles and pipeline as the kernels
// above, other shapes: a CTA's 128 MMA rows hold the dosage plane g of 128 i-SNPs (two i-blocks = one i-pair of the slot
// scheme), the pair's B operand is [h ; a] of 128 j-SNPs (each CTA stages the j-block of its rank: 64 SNPs x 2 planes),
// one tcgen05.mma M = 256, N = 256, K = 32 per 32 organisms -- the instruction shape that reaches the pipe's measured
// ceiling -- and the 256-column accumulator is double-buffered in TMEM, so the epilogue of supertile t runs under the
// MMAs of t+1.  A third of the MMA work and 2/5 of the operand bytes per SNP pair.  Partials go to the slots the
// 6-plane kernels use, summed in the same order from the same expressions (pair_eval_mfree): bitwise identical scores.
namespace {
constexpr int kMStages = 6;
constexpr int kMABytes = 2 * kBlk * kKStep;                  // 16 KB: g of 128 i-SNPs
constexpr int kMBBytes = 2 * kBlk * kKStep;                  // 16 KB: [h ; a] of 64 j-SNPs
constexpr int kMStageBytes = kMABytes + kMBBytes;            // 32 KB per CTA per stage
constexpr int kMCols = 2 * kBlk;                             // j-SNPs of a supertile
constexpr int kMColBytes = kMCols * static_cast<int>(sizeof(SnpCoef)) + kMCols;
constexpr int kMSmemBytes = kMStages * kMStageBytes + 1024 + 256 + kMColBytes;
constexpr int kMThreads = 320;                               // TMA warp, MMA warp, 8 epilogue warps
constexpr uint32_t kIdescM = (2u << 4) | ((256u >> 3) << 17) | ((256u >> 4) << 24);   // s32 += u8 x u8, M = 256, N = 256
}  // namespace

__device__ __forceinline__ void band_mfree_body(const CUtensorMap& map_g, const CUtensorMap& map_ha, const BandArgs& args) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + kMStages * kMStageBytes);   // leader CTA only
    uint64_t* empty = full + kMStages;         // one per CTA
    uint64_t* tfull = empty + kMStages;        // one per CTA, two accumulator buffers
    uint64_t* tempty = tfull + 2;              // leader CTA only
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);
    double2* col_coef = reinterpret_cast<double2*>(smem + kMStages * kMStageBytes + 256);
    uint8_t* col_flags = reinterpret_cast<uint8_t*>(col_coef) + kMCols * sizeof(SnpCoef);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    const int pair = blockIdx.x >> 1, n_pairs = gridDim.x >> 1;

    if (warp == 0 && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&map_g)) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&map_ha)) : "memory");
        for (int s = 0; s < kMStages; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
        for (int b = 0; b < 2; ++b) { mbar_init(&tfull[b], 1); mbar_init(&tempty[b], 16); }   // 8 epilogue warps x 2 CTAs
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                     "r"((uint32_t)kTmemCols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    // fp4: the scale factors take TMEM columns 448..479, so there is room for ONE 256-column accumulator: single-buffered
    const bool one_buf = args.fp4 != 0;
    if (args.fp4) {
        if (warp >= 2 && warp < 6) tmem_fill_sf_ones(tmem_base, warp);
        tc_fence_before();
        cluster_sync_all();
        tc_fence_after();
    }

    if (warp == 0) {
        // ===== TMA producer: g of this CTA's two i-blocks, [h ; a] of the j-block of its rank =====
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int t = pair; t < args.n_qtiles; t += n_pairs) {
                const QTile tl = args.qtiles[t];
                const int32_t a_row0 = tl.a_row[2 * rank], a_row1 = tl.a_row[2 * rank + 1];
                const int32_t b_row = tl.b_row[rank];
                for (int k = 0; k < args.k_steps; ++k) {
                    mbar_wait(&empty[stage], phase ^ 1, 11);
                    if (rank == 0) mbar_expect_tx(&full[stage], 2 * kMStageBytes);    // both CTAs' bytes
                    uint8_t* sa = smem + stage * kMStageBytes;
                    tma_load_3d(sa, &map_g, &full[stage], k * kKStep, a_row0, kPlaneG);
                    tma_load_3d(sa + kBlk * kKStep, &map_g, &full[stage], k * kKStep, a_row1, kPlaneG);
                    tma_load_3d(sa + kMABytes, &map_ha, &full[stage], k * kKStep, b_row, kPlaneH);
                    if (++stage == kMStages) { stage = 0; phase ^= 1; }
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ===== MMA issuer (leader CTA only) =====
        if (rank == 0 && lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            int n = 0;
            for (int t = pair; t < args.n_qtiles; t += n_pairs, ++n) {
                const int buf = one_buf ? 0 : (n & 1);
                mbar_wait(&tempty[buf], ((one_buf ? n : (n >> 1)) & 1) ^ 1, 12);   // both CTAs' epilogues drained this accumulator
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + buf * 256;
                for (int k = 0; k < args.k_steps; ++k) {
                    mbar_wait(&full[stage], phase, 13);
                    tc_fence_after();
                    const uint32_t sa = smem_u32(smem + stage * kMStageBytes);
                    const uint64_t a_desc = make_smem_desc(sa);
                    const uint64_t b_desc = make_smem_desc(sa + kMABytes);
                    if (args.fp4) {
                        const uint32_t sf = tmem_base + kSfCols;
#pragma unroll
                        for (int kk = 0; kk < kKStep / 32; ++kk)
                            tc_mma_mxf4(d_tmem, a_desc + 2 * kk, b_desc + 2 * kk, idesc_mxf4(256), sf, sf + 16, (k | kk) != 0);
                    } else {
#pragma unroll
                        for (int kk = 0; kk < kKStep / 32; ++kk)
                            tc_mma_i8(d_tmem, a_desc + 2 * kk, b_desc + 2 * kk, kIdescM, (k | kk) != 0);
                    }
                    // A commit costs the tensor pipe ~220 clk (i8 peak probe: N = 256 reaches 3,400 TOP/s at 4 MMAs per commit,
                    // 4,580 at 8).  commit2: release the operand ring two stages at a time -- 8 MMAs per commit; the 6 x 32 KB
                    // ring is then three 64 KB granules deep.
                    if (!args.commit2) {
                        tc_commit_pair(&empty[stage]);
                    } else if ((k & 1) || k == args.k_steps - 1) {
                        if (k & 1) tc_commit_pair(&empty[stage == 0 ? kMStages - 1 : stage - 1]);
                        tc_commit_pair(&empty[stage]);
                    }
                    if (++stage == kMStages) { stage = 0; phase ^= 1; }
                }
                tc_commit_pair(&tfull[buf]);
            }
        }
        __syncwarp();
    } else {
        // ===== epilogue: 8 warps.  TMEM lane = i-SNP (128 per CTA); warps 2-5 score j-block 0 (columns 0..127 =
        // [h | a] of its 64 SNPs), warps 6-9 j-block 1; warp w reads lanes 32 (w % 4) .. +31 =====
        const int q = warp & 3;
        const int jj = (warp - 2) >> 2;
        const int rho = 32 * q + lane;               // row of this CTA's A tile
        const int blk = rho >> 6;                    // which i-block of the i-pair ("rank" of the slot scheme)
        int n = 0;
        for (int t = pair; t < args.n_qtiles; t += n_pairs, ++n) {
            const int buf = one_buf ? 0 : (n & 1);
            const QTile tl = args.qtiles[t];
            const int32_t i0 = tl.i0[2 * rank + blk];
            const RowCtx row = load_row_ctx(args, i0 == kDeadBlock ? kDeadBlock : i0 + (rho & 63));
            asm volatile("bar.sync 1, 256;" ::: "memory");          // nobody still reads the previous supertile's constants
            {
                const int te = threadIdx.x - 64;
                for (int e = te; e < kMCols * 4; e += 256) {
                    const int32_t gj = tl.j0 + (e >> 2);
                    double2 v = make_double2(0.0, 0.0);
                    if (gj < args.n_load) v = __ldg(reinterpret_cast<const double2*>(args.coef + gj) + (e & 3));
                    col_coef[e] = v;
                }
                if (te < kMCols) {
                    const int32_t gj = tl.j0 + te;
                    col_flags[te] = gj < args.n_load ? __ldg(args.flags + gj) : uint8_t(0);
                }
            }
            asm volatile("bar.sync 1, 256;" ::: "memory");
            mbar_wait_idle(&tfull[buf], (one_buf ? n : (n >> 1)) & 1, 14);
            tc_fence_after();
            const uint32_t taddr = tmem_base + buf * 256 + 128 * jj + (static_cast<uint32_t>(32 * q) << 16);
            const int32_t jbase = tl.j0 + kBlk * jj;
            const int32_t slot = tl.slot[rank][jj];
#pragma unroll 1
            for (int half = 0; half < 2; ++half) {
                RowAcc ev, od;                        // even / odd j: the two chains the 6-plane kernels sum a half in
#pragma unroll 1
                for (int c2 = 0; c2 < 2; ++c2) {
                    const int c = 2 * half + c2;
                    if (!__any_sync(0xFFFFFFFFu, row_needs(row, jbase + 16 * c, jbase + 16 * c + 15))) continue;
                    int32_t vh[16], va[16];
                    tmem_ld16(taddr + 16 * c, vh);
                    tmem_ld16(taddr + kBlk + 16 * c, va);
                    tmem_ld_wait();
#if !NLDSC_FP4_F2F
                    if (args.fp4) { acc_to_int(vh); acc_to_int(va); }
#endif
                    auto score16 = [&](auto f32) {
                        constexpr bool kF32 = decltype(f32)::value;
#pragma unroll
                        for (int u = 0; u < 16; ++u) {
                            const int jl = 16 * c + u;
                            const double2* cj = col_coef + 4 * (kBlk * jj + jl);
                            pair_eval_mfree<kF32>(args, row, jbase + jl, col_flags[kBlk * jj + jl], cj[0], cj[1], cj[2], cj[3], vh[u], va[u],
                                                  (u & 1) ? od : ev);
                        }
                    };
#if NLDSC_FP4_F2F
                    if (args.fp4) score16(std::true_type{}); else score16(std::false_type{});
#else
                    score16(std::false_type{});
#endif
                }
                if (half == 1) {
                    // accumulator fully read: hand it back to the leader's MMA warp
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive_leader(&tempty[buf]);
                }
                if (slot >= 0 && i0 != kDeadBlock) {
                    const int64_t o = part_index(slot, blk, half, rho & 63);
                    args.part.add[o] = ev.add + od.add;
                    args.part.dom[o] = ev.dom + od.dom;
                    args.part.wsa[o] = ev.wsa + od.wsa;
                    args.part.wsd[o] = ev.wsd + od.wsd;
                    args.part.wsde[o] = ev.wsde + od.wsde;
                }
            }
        }
    }

    tc_fence_before();
    cluster_sync_all();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)kTmemCols)
                     : "memory");
    }
}

// ---- 160-column supertiles: five 32-SNP column groups per i-pair, two MMAs per 64 organisms (fp4 planes only) --------
// See XTile (common.cuh).  Per K-step (256 organisms) a CTA stages its A block (64 i-SNPs x {g, m}: 16 KB, as the wide body
// does) and the 80 j-SNPs of its rank as five 16-SNP boxes [m16 h16 a16] (240 rows, 30 KB; four stages of 46 KB); the
// leader issues 4 x (N = 256 from each CTA's rows 0..127, into TMEM columns 0..255) + 4 x (N = 224 from rows 128..239,
// into columns 256..479); the scale factors (all 2^0) sit in columns 480..511.  B row r of rank c lands in column
// 128 c + r (r < 128) or 256 + 112 c + (r - 128), so the m / h / a counts of sixteen k = 5 c + u (j-SNPs 16 k .. 16 k + 15
// of the supertile) are 16 consecutive columns each, at rows r = 48 u + 16 plane.
// Epilogue: 8 warps as in the wide body (warp w reads TMEM lanes 32 (w % 4) ..+31 = 16 i-SNPs x {g, m}).  A partial sum is
// the sum over ONE 32-SNP group in a fixed order (that is what makes every body bitwise equal), and five groups do not
// split over two warp groups: warp group 0 scores the first sixteen of group 2, then groups 0 and 1; warp group 1 scores
// groups 3 and 4, then takes group 2's running sums over through shared memory and adds its second sixteen -- five
// sixteens each.
namespace {
constexpr int kXStages = 4;
constexpr int kXSnps = 160;                                  // j-SNPs of a supertile; each CTA stages 80
constexpr int kXBRows = 240;                                 // per CTA: 5 sixteens x {m, h, a}
constexpr int kXBoxBytes = 48 * kKStep;                      // one sixteen: 6 KB
constexpr int kXBBytes = kXBRows * kKStep;                   // 30 KB
constexpr int kXStageBytes = kABytes + kXBBytes;             // 46 KB per CTA per stage
constexpr int kXColBytes = kXSnps * static_cast<int>(sizeof(SnpCoef)) + kXSnps;   // the j-SNPs' SnpCoef + flags
constexpr int kXHandBytes = 128 * (8 + 8 + 12);              // group 2's running sums, per epilogue lane of one warp group
constexpr int kXSmemBytes = kXStages * kXStageBytes + 1024 + 256 + kXColBytes + kXHandBytes;
constexpr uint32_t kXSfCols = 480;
static_assert(kXStageBytes % 1024 == 0 && kXBoxBytes % 1024 == 0, "swizzle atoms stay aligned");
static_assert(kXColBytes % 16 == 0, "the handover area stays 16-byte aligned");
}  // namespace

// named barriers 2..5 (literal ids: a register operand would make ptxas reserve all 16 of the SM's barriers for this CTA):
// the two epilogue warps that share TMEM lane quarter q hand group 2's running sums over
__device__ __forceinline__ void hand_bar_arrive(int q) {
    switch (q) {
        case 0: asm volatile("bar.arrive 2, 64;" ::: "memory"); break;
        case 1: asm volatile("bar.arrive 3, 64;" ::: "memory"); break;
        case 2: asm volatile("bar.arrive 4, 64;" ::: "memory"); break;
        default: asm volatile("bar.arrive 5, 64;" ::: "memory"); break;
    }
}
__device__ __forceinline__ void hand_bar_sync(int q) {
    switch (q) {
        case 0: asm volatile("bar.sync 2, 64;" ::: "memory"); break;
        case 1: asm volatile("bar.sync 3, 64;" ::: "memory"); break;
        case 2: asm volatile("bar.sync 4, 64;" ::: "memory"); break;
        default: asm volatile("bar.sync 5, 64;" ::: "memory"); break;
    }
}

__device__ __forceinline__ void band_x160_body(const CUtensorMap& map_a, const CUtensorMap& map_b16, const BandArgs& args) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + kXStages * kXStageBytes);   // leader CTA only
    uint64_t* empty = full + kXStages;         // one per CTA
    uint64_t* tfull = empty + kXStages;        // one per CTA
    uint64_t* tempty = tfull + 1;              // leader CTA only
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 1);
    double2* col_coef = reinterpret_cast<double2*>(smem + kXStages * kXStageBytes + 256);
    uint8_t* col_flags = reinterpret_cast<uint8_t*>(col_coef) + kXSnps * sizeof(SnpCoef);
    double* hand_add = reinterpret_cast<double*>(reinterpret_cast<uint8_t*>(col_coef) + kXColBytes);
    double* hand_dom = hand_add + 128;
    int32_t* hand_w = reinterpret_cast<int32_t*>(hand_dom + 128);                   // [3][128]

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    const int pair = blockIdx.x >> 1, n_pairs = gridDim.x >> 1;

    if (warp == 0 && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&map_a)) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&map_b16)) : "memory");
        for (int s = 0; s < kXStages; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
        mbar_init(tfull, 1);
        mbar_init(tempty, 16);                 // 8 epilogue warps x 2 CTAs
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                     "r"((uint32_t)kTmemCols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    if (warp >= 2 && warp < 6) tmem_fill_sf_ones(tmem_base, warp, kXSfCols);
    tc_fence_before();
    cluster_sync_all();
    tc_fence_after();

    if (warp == 0) {
        // ===== TMA producer: own A block, the five sixteens of the own 80 j-SNPs =====
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int t = pair; t < args.n_xtiles; t += n_pairs) {
                const XTile tl = args.xtiles[t];
                const int32_t a_row = tl.a_row[rank];
                int32_t b_row[5];
#pragma unroll
                for (int u = 0; u < 5; ++u) b_row[u] = (tl.j0 + 80 * static_cast<int32_t>(rank) + 16 * u) % args.ring_rows;
                const uint32_t bytes = 2u * kXStageBytes;                                    // both CTAs' bytes
                for (int k = 0; k < args.k_steps; ++k) {
                    mbar_wait(&empty[stage], phase ^ 1, 15);
                    if (rank == 0) mbar_expect_tx(&full[stage], bytes);
                    uint8_t* sa = smem + stage * kXStageBytes;
                    uint8_t* sb = sa + kABytes;
#pragma unroll
                    for (int u = 0; u < 4; ++u)
                        tma_load_3d(sa + u * 4096, &map_a, &full[stage], k * kKStep, a_row + 16 * u, kPlaneG);
#pragma unroll
                    for (int u = 0; u < 5; ++u)
                        tma_load_3d(sb + u * kXBoxBytes, &map_b16, &full[stage], k * kKStep, b_row[u], kPlaneM);
                    if (++stage == kXStages) { stage = 0; phase ^= 1; }
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ===== MMA issuer (leader CTA only) =====
        if (rank == 0 && lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            int n = 0;
            const uint32_t sf = tmem_base + kXSfCols;
            for (int t = pair; t < args.n_xtiles; t += n_pairs, ++n) {
                mbar_wait(tempty, (n & 1) ^ 1, 16);                  // both CTAs' epilogues drained the accumulator
                tc_fence_after();
                for (int k = 0; k < args.k_steps; ++k) {
                    mbar_wait(&full[stage], phase, 17);
                    tc_fence_after();
                    const uint32_t sa = smem_u32(smem + stage * kXStageBytes);
                    const uint64_t a_desc = make_smem_desc(sa);
                    const uint64_t b1_desc = make_smem_desc(sa + kABytes);
                    const uint64_t b2_desc = make_smem_desc(sa + kABytes + 128 * kKStep);
#pragma unroll
                    for (int kk = 0; kk < kKStep / 32; ++kk)
                        tc_mma_mxf4(tmem_base, a_desc + 2 * kk, b1_desc + 2 * kk, idesc_mxf4(256), sf, sf + 16, (k | kk) != 0);
#pragma unroll
                    for (int kk = 0; kk < kKStep / 32; ++kk)
                        tc_mma_mxf4(tmem_base + 256, a_desc + 2 * kk, b2_desc + 2 * kk, idesc_mxf4(224), sf, sf + 16, (k | kk) != 0);
                    tc_commit_pair(&empty[stage]);
                    if (++stage == kXStages) { stage = 0; phase ^= 1; }
                }
                tc_commit_pair(tfull);
            }
        }
        __syncwarp();
    } else {
        // ===== epilogue: eight warps per CTA; q = TMEM lane quarter, wg = warp group =====
        const int q = warp & 3;
        const int wg = (warp - 2) >> 2;
        const bool is_g = lane < 16;
        const uint32_t psel = is_g ? 0x3210u : 0x7654u;      // pick(): g lanes take the first argument, m lanes the second
        const int s = 16 * q + (lane & 15);
        const int hl = 32 * q + lane;                                 // handover slot of this lane
        int n = 0;
        for (int t = pair; t < args.n_xtiles; t += n_pairs, ++n) {
            const XTile tl = args.xtiles[t];
            const int32_t i0 = tl.i0[rank];
            const RowCtx row = load_row_ctx(args, i0 + s);
            // the constants of the supertile's 160 j-SNPs, staged while the MMAs run (as in the wide body)
            asm volatile("bar.sync 1, 256;" ::: "memory");
            {
                const int te = threadIdx.x - 64;                       // 0..255 among the epilogue threads
                for (int e = te; e < kXSnps * 4; e += 256) {           // 4 double2 per SNP
                    const int32_t gj = tl.j0 + (e >> 2);
                    double2 v = make_double2(0.0, 0.0);
                    if (gj < args.n_load) v = __ldg(reinterpret_cast<const double2*>(args.coef + gj) + (e & 3));
                    col_coef[e] = v;
                }
                if (te < kXSnps) {
                    const int32_t gj = tl.j0 + te;
                    col_flags[te] = gj < args.n_load ? __ldg(args.flags + gj) : uint8_t(0);
                }
            }
            asm volatile("bar.sync 1, 256;" ::: "memory");
            mbar_wait_idle(tfull, n & 1, 18);
            tc_fence_after();
            const uint32_t taddr = tmem_base + (static_cast<uint32_t>(32 * q) << 16);
            RowAcc acc;
#pragma unroll 1
            for (int step = 0; step < 5; ++step) {
                // warp group 0: sixteen 4 (first half of group 2), then groups 0 and 1; warp group 1: groups 3 and 4, then sixteen 5
                const int k = wg == 0 ? (step == 0 ? 4 : step - 1) : (step < 4 ? 6 + step : 5);
                const int grp = k >> 1;
                if (wg == 1 && step == 4) {
                    // group 2's sums so far, from the warp of the other group that shares this lane quarter
                    hand_bar_sync(q);
                    acc.add = hand_add[hl];
                    acc.dom = hand_dom[hl];
                    acc.wsa = hand_w[hl];
                    acc.wsd = hand_w[128 + hl];
                    acc.wsde = hand_w[256 + hl];
                }
                const int32_t jk = tl.j0 + 16 * k;
                if (grp < tl.n_grp && (args.no_skip || __any_sync(0xFFFFFFFFu, row_needs(row, jk, jk + 15)))) {
                    const int c = k >= 5 ? 1 : 0, u5 = k - 5 * c;
                    int32_t vm[16], vh[16], va[16];
                    {
                        const int r_m = 48 * u5, r_h = r_m + 16, r_a = r_m + 32;
                        const uint32_t col_m = r_m < 128 ? 128 * c + r_m : 256 + 112 * c + (r_m - 128);
                        const uint32_t col_h = r_h < 128 ? 128 * c + r_h : 256 + 112 * c + (r_h - 128);
                        const uint32_t col_a = r_a < 128 ? 128 * c + r_a : 256 + 112 * c + (r_a - 128);
                        tmem_ld16(taddr + col_m, vm);
                        tmem_ld16(taddr + col_h, vh);
                        tmem_ld16(taddr + col_a, va);
                        tmem_ld_wait();
                    }
#pragma unroll
                    for (int u = 0; u < 8; ++u) {
                        // FP32 bit patterns of exact counts, converted to double pair by pair (cnt2d<true>)
                        const int32_t sm_ = pick(psel, vm[2 * u + 1], vm[2 * u]);
                        const int32_t sh_ = pick(psel, vh[2 * u + 1], vh[2 * u]);
                        const int32_t sa_ = pick(psel, va[2 * u + 1], va[2 * u]);
                        const int32_t rm = __shfl_xor_sync(0xFFFFFFFFu, sm_, 16);
                        const int32_t rh = __shfl_xor_sync(0xFFFFFFFFu, sh_, 16);
                        const int32_t ra = __shfl_xor_sync(0xFFFFFFFFu, sa_, 16);
                        const int jl = 16 * k + 2 * u + (is_g ? 0 : 1);
                        const int32_t Ggm = pick(psel, vm[2 * u], rm), Gmm = pick(psel, rm, vm[2 * u + 1]);
                        const int32_t Ggh = pick(psel, vh[2 * u], rh), Gmh = pick(psel, rh, vh[2 * u + 1]);
                        const int32_t Gga = pick(psel, va[2 * u], ra), Gma = pick(psel, ra, va[2 * u + 1]);
                        const double2* cj = col_coef + 4 * jl;
                        pair_eval_core<true>(args, row, tl.j0 + jl, col_flags[jl], cj[0], cj[1], cj[2], cj[3],
                                             Ggm, Ggh, Gga, Gmm, Gmh, Gma, acc);
                    }
                }
                if (wg == 0 && step == 0) {
                    hand_add[hl] = acc.add;
                    hand_dom[hl] = acc.dom;
                    hand_w[hl] = acc.wsa;
                    hand_w[128 + hl] = acc.wsd;
                    hand_w[256 + hl] = acc.wsde;
                    hand_bar_arrive(q);
                    acc = RowAcc();
                    continue;
                }
                const bool group_done = wg == 0 ? (step == 2 || step == 4) : (step == 1 || step >= 3);
                if (!group_done) continue;
                if (step == 4) {
                    // the last TMEM read of this warp is behind it: tell the leader's MMA warp
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive_leader(tempty);
                }
                if (grp < tl.n_grp) {
                    acc.add += __shfl_xor_sync(0xFFFFFFFFu, acc.add, 16);
                    acc.dom += __shfl_xor_sync(0xFFFFFFFFu, acc.dom, 16);
                    acc.wsa += __shfl_xor_sync(0xFFFFFFFFu, acc.wsa, 16);
                    acc.wsd += __shfl_xor_sync(0xFFFFFFFFu, acc.wsd, 16);
                    acc.wsde += __shfl_xor_sync(0xFFFFFFFFu, acc.wsde, 16);
                    if (is_g) {
                        const int gg = tl.odd + grp;
                        const int64_t o = part_index(tl.slot + (gg >> 1), static_cast<int>(rank), gg & 1, s);
                        args.part.add[o] = acc.add;
                        args.part.dom[o] = acc.dom;
                        args.part.wsa[o] = acc.wsa;
                        args.part.wsd[o] = acc.wsd;
                        args.part.wsde[o] = acc.wsde;
                    }
                }
                acc = RowAcc();
            }
        }
    }

    tc_fence_before();
    cluster_sync_all();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)kTmemCols)
                     : "memory");
    }
}

// ONE launch per chunk for both paths: the chunk's flag (written on the device after its stats) picks the body.
// (Two kernels, one of them returning at once, cost configs[1] 12 %: the idle launch still has to wait for a whole
// grid's worth of free SMs behind the other stream's persistent band kernel -- profiles/r02i_ab_missing_n10000.json.)
#ifndef NLDSC_WIDE_MAXNREG
#define NLDSC_WIDE_MAXNREG 128   // 144 / 160 measured slower (r02y): the unpack blocks no longer fit on the SM beside this kernel
#endif
__global__ void __cluster_dims__(2, 1, 1) __maxnreg__(NLDSC_WIDE_MAXNREG)
band_tc_wide_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                    const __grid_constant__ CUtensorMap map_g, const __grid_constant__ CUtensorMap map_ha, const BandArgs args) {
    // every thread of the grid reads the same word, written before this launch could start
    const bool mfree = args.mfree_flag != nullptr && *reinterpret_cast<const volatile uint32_t*>(args.mfree_flag) != 0;
    if (mfree) band_mfree_body(map_g, map_ha, args);
    else band_wide_body(map_a, map_b, args);
}

// The same launch with the 160-column body for the 6-plane work (fp4 planes).  A kernel of its own so that the register
// allocation of the kernel above stays what it was.
__global__ void __cluster_dims__(2, 1, 1) __maxnreg__(NLDSC_WIDE_MAXNREG)
band_tc_x160_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b16,
                    const __grid_constant__ CUtensorMap map_g, const __grid_constant__ CUtensorMap map_ha, const BandArgs args) {
    const bool mfree = args.mfree_flag != nullptr && *reinterpret_cast<const volatile uint32_t*>(args.mfree_flag) != 0;
    if (mfree) band_mfree_body(map_g, map_ha, args);
    else band_x160_body(map_a, map_b16, args);
}

// ---- host side: TMA descriptors + launch ------------------------------------------------
struct BandMaps {
    CUtensorMap a;   // box {128 B, 16 rows, 2 planes}
    CUtensorMap b;   // box {128 B, 32 rows, 3 planes}
    CUtensorMap g;   // box {128 B, 64 rows, 1 plane}    missing-free kernel: g of one i-block
    CUtensorMap ha;  // box {128 B, 64 rows, 2 planes}   missing-free kernel: [h ; a] of one j-block
    CUtensorMap b16; // box {128 B, 16 rows, 3 planes}   160-column supertiles: [m ; h ; a] of one sixteen
};

namespace {
using EncodeFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                              const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                              CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeFn get_encode() {
    static EncodeFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        CUDA_CHECK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres));
        if (qres != cudaDriverEntryPointSuccess || !p) throw CudaError("cuTensorMapEncodeTiled not available in this driver");
        fn = reinterpret_cast<EncodeFn>(p);
    }
    return fn;
}

void encode(CUtensorMap* m, const int8_t* planes, int64_t plane_rows, int32_t npad, uint32_t box_rows,
            uint32_t box_planes) {
    const cuuint64_t dims[3] = {(cuuint64_t)npad, (cuuint64_t)plane_rows, (cuuint64_t)kPlanes};
    const cuuint64_t strides[2] = {(cuuint64_t)npad, (cuuint64_t)npad * (cuuint64_t)plane_rows};
    const cuuint32_t box[3] = {(cuuint32_t)kKStep, box_rows, box_planes};
    const cuuint32_t estr[3] = {1, 1, 1};
    const CUresult r = get_encode()(m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, const_cast<int8_t*>(planes), dims, strides,
                                    box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                    CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        char buf[160];
        snprintf(buf, sizeof(buf), "cuTensorMapEncodeTiled failed (CUresult %d; rows %lld npad %d)", (int)r,
                 (long long)plane_rows, npad);
        throw CudaError(buf);
    }
}
}  // namespace

BandMaps* band_maps_create(const int8_t* planes, int64_t plane_rows, int32_t npad) {
    BandMaps* m = new BandMaps;
    try {
        encode(&m->a, planes, plane_rows, npad, 16, 2);
        encode(&m->b, planes, plane_rows, npad, 32, 3);
        encode(&m->g, planes, plane_rows, npad, 64, 1);
        encode(&m->ha, planes, plane_rows, npad, 64, 2);
        encode(&m->b16, planes, plane_rows, npad, 16, 3);
    } catch (...) {
        delete m;
        throw;
    }
    return m;
}
void band_maps_destroy(BandMaps* m) { delete m; }

// One word of mapped host memory per device for hang_trap()'s diagnostic; installed on first launch.
namespace {
uint32_t* g_hang_host[64] = {nullptr};
std::mutex g_hang_mu;
const char* const kSiteNames[] = {"?", "narrow/producer:empty", "narrow/mma:tmem-empty", "narrow/mma:full", "narrow/epilogue:tmem-full",
                                  "wide/producer:empty", "wide/mma:tmem-empty", "wide/mma:full", "wide/epilogue:tmem-full", "i8-peak",
                                  "rows-ready (the chunk's unpack never published)", "mfree/producer:empty", "mfree/mma:tmem-empty", "mfree/mma:full",
                                  "mfree/epilogue:tmem-full", "x160/producer:empty", "x160/mma:tmem-empty", "x160/mma:full", "x160/epilogue:tmem-full"};
void install_hang_word() {
    int dev = 0;
    CUDA_CHECK(cudaGetDevice(&dev));
    std::lock_guard<std::mutex> g(g_hang_mu);
    if (g_hang_host[dev & 63]) return;
    uint32_t* h = nullptr;
    CUDA_CHECK(cudaHostAlloc(reinterpret_cast<void**>(&h), sizeof(uint32_t), cudaHostAllocMapped | cudaHostAllocPortable));
    *h = 0;
    uint32_t* d = nullptr;
    CUDA_CHECK(cudaHostGetDevicePointer(reinterpret_cast<void**>(&d), h, 0));
    CUDA_CHECK(cudaMemcpyToSymbol(g_hang_word, &d, sizeof(d)));
    g_hang_host[dev & 63] = h;
}
}  // namespace

std::string band_hang_note() {
    std::lock_guard<std::mutex> g(g_hang_mu);
    std::string note;
    for (int d = 0; d < 64; ++d) {
        if (!g_hang_host[d]) continue;
        const uint32_t w = *reinterpret_cast<volatile uint32_t*>(g_hang_host[d]);
        if (!w) continue;
        char buf[160];
        const uint32_t site = w >> 24;
        snprintf(buf, sizeof(buf), " [band kernel on device %d gave up after 20 s at barrier %s, block %u]", d,
                 site < sizeof(kSiteNames) / sizeof(kSiteNames[0]) ? kSiteNames[site] : "?", w & 0x7FFFFFu);
        note += buf;
    }
    return note;
}

constexpr int kWideKernelSmem = kWSmemBytes > kMSmemBytes ? kWSmemBytes : kMSmemBytes;
constexpr int kX160KernelSmem = kXSmemBytes > kMSmemBytes ? kXSmemBytes : kMSmemBytes;
static_assert(kWideKernelSmem <= 232448 && kX160KernelSmem <= 232448, "dynamic shared memory of one CTA");

void install_band_kernels() {
    install_hang_word();
    CUDA_CHECK(cudaFuncSetAttribute(band_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
    CUDA_CHECK(cudaFuncSetAttribute(band_tc_wide_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kWideKernelSmem));
    CUDA_CHECK(cudaFuncSetAttribute(band_tc_x160_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kX160KernelSmem));
}

void launch_band_tc(const BandMaps* maps, const BandArgs& args, int sm_count, cudaStream_t st) {
    if (args.n_tiles <= 0) return;
    install_hang_word();
    // per device (several GPUs may be driven from one process), so not cached in a static
    CUDA_CHECK(cudaFuncSetAttribute(band_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
    const int pairs = args.n_tiles < sm_count / 2 ? args.n_tiles : sm_count / 2;     // one CTA pair per TPC
    band_tc_kernel<<<2 * pairs, kThreads, kSmemBytes, st>>>(maps->a, maps->b, args);
    CUDA_CHECK(cudaGetLastError());
}

void launch_band_tc_wide(const BandMaps* maps, const BandArgs& args, int sm_count, cudaStream_t st) {
    // the supertile list the launch walks is chosen on the device (the chunk's missing-free flag): size the grid for the longest
    const bool x160 = args.xtiles != nullptr;
    const int32_t n_6 = x160 ? args.n_xtiles : args.n_tiles;
    const int32_t n_max = args.mfree_flag != nullptr ? (n_6 > args.n_qtiles ? n_6 : args.n_qtiles) : n_6;
    if (n_max <= 0) return;
    install_hang_word();
    const int pairs = n_max < sm_count / 2 ? n_max : sm_count / 2;
    if (x160) {
        CUDA_CHECK(cudaFuncSetAttribute(band_tc_x160_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kX160KernelSmem));
        band_tc_x160_kernel<<<2 * pairs, kWThreads, kX160KernelSmem, st>>>(maps->a, maps->b16, maps->g, maps->ha, args);
        CUDA_CHECK(cudaGetLastError());
        return;
    }
    constexpr int smem = kWideKernelSmem;
    CUDA_CHECK(cudaFuncSetAttribute(band_tc_wide_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    band_tc_wide_kernel<<<2 * pairs, kWThreads, smem, st>>>(maps->a, maps->b, maps->g, maps->ha, args);
    CUDA_CHECK(cudaGetLastError());
}

// ---- int8 tensor-pipe ceiling (measurement only) ------------------------------------------------
// Back-to-back tcgen05.mma.cta_group::2.kind::i8 (M = 256, N = n_cols, K = 32) from shared memory on every CTA
// pair: no TMA, no epilogue, operands resident.  What the tensor pipe sustains on this instruction shape at the
// clocks the box runs it at -- the denominator SURVEY section 8(d) asks to be MEASURED, not derived from bf16.
// Operands are filled with genotype-like bytes (0/1/2 left, 0/1 right) so that the datapath toggles as it does on
// real planes.  `per_commit` MMAs are issued per tcgen05.commit (the wide band kernel issues 8, the narrow one 4),
// alternating between two accumulators; at most 4 commits are in flight.
template <int kPerCommit>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128, 1)
i8_peak_kernel(int iters, int n_cols, int depth_flags) {
    const int depth = depth_flags & 0xFF;          // bit 8: commit to the leader's barrier only (no multicast)
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    constexpr int kMaxDepth = 8;
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 2 * kABytes);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar + kMaxDepth);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    // A: 128 rows x 128 B, B: up to 128 rows x 128 B (n_cols / 2 used)
    for (int i = threadIdx.x; i < 2 * kABytes / 4; i += blockDim.x) {
        uint32_t h = (i * 2654435761u) ^ (blockIdx.x * 40503u);
        h ^= h >> 15; h *= 2246822519u; h ^= h >> 13;
        const uint32_t m = i < kABytes / 4 ? 0x03030303u : 0x01010101u;
        uint32_t v = h & m;
        if (i < kABytes / 4) v -= (v >> 1) & v & 0x01010101u;          // 3 -> 2: dosages 0/1/2
        reinterpret_cast<uint32_t*>(smem)[i] = v;
    }
    if (warp == 0 && lane == 0) {
        for (int d = 0; d < kMaxDepth; ++d) mbar_init(&bar[d], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                     "r"((uint32_t)kTmemCols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy fills visible to the MMA's async proxy
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    if (warp == 1 && lane == 0 && rank == 0) {
        const uint32_t idesc = (2u << 4) | ((static_cast<uint32_t>(n_cols) >> 3) << 17) | (((2 * kARows) >> 4) << 24);
        const uint64_t a_desc = make_smem_desc(smem_u32(smem));
        const uint64_t b_desc = make_smem_desc(smem_u32(smem) + kABytes);
        for (int it = 0; it < iters; ++it) {
            const int d = it % depth;
            if (it >= depth) mbar_wait(&bar[d], ((it / depth) - 1) & 1, 9);
            tc_fence_after();
#pragma unroll
            for (int m = 0; m < kPerCommit; ++m)       // 4 K-slices into accumulator 0, then 4 into accumulator 1, ...
                tc_mma_i8(tmem_base + ((m >> 2) & 1) * 256, a_desc + 2 * (m & 3), b_desc + 2 * (m & 3), idesc, 1u);
            if (depth_flags & 0x100) tc_commit_leader(&bar[d]); else tc_commit_pair(&bar[d]);
        }
        for (int it = iters > depth ? iters - depth : 0; it < iters; ++it) mbar_wait(&bar[it % depth], (it / depth) & 1, 9);
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)kTmemCols)
                     : "memory");
    }
}

void launch_i8_peak(int iters, int n_cols, int per_commit, int depth, int sm_count, cudaStream_t st) {
    const int smem = 2 * kABytes + 1024 + 256;
    auto go = [&](auto kernel) {
        CUDA_CHECK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        kernel<<<(sm_count / 2) * 2, 128, smem, st>>>(iters, n_cols, depth);
        CUDA_CHECK(cudaGetLastError());
    };
    if (per_commit == 4) go(i8_peak_kernel<4>);
    else if (per_commit == 8) go(i8_peak_kernel<8>);
    else if (per_commit == 16) go(i8_peak_kernel<16>);
    else throw CudaError("i8 peak probe: per_commit must be 4, 8 or 16");
}

// ---- fp4 feasibility probe (measurement only) -----------------------------------------------------
// Genotype planes hold 0 / 1 / 2, all exactly representable in e2m1 (fp4), and every cross-count is below 2^24: if
// tcgen05.mma.kind::mxf4 (K = 64 per instruction, 2 x the int8 rate, half the operand bytes) accumulates small integers
// EXACTLY in its FP32 accumulator, the band kernels can run on the fp4 pipe with all scale factors 1.0 and stay
// bit-exact.  This kernel answers that on the box: one CTA pair, operands resident in shared memory with known nibble
// patterns, `n_mma` accumulating MMAs (M = 256, N = n_cols), scale factors 2^0 (E8M0 127) stored once to TMEM with
// tcgen05.st, then the first 16 accumulator columns of every row are written out as floats.
// Expected with A = [1, 2, 1, 2, ...] along K in every row, B row 0 = e_0, B row 8 = e_0 + e_1 + e_2, other B rows all ones:
//   column 0 = n_mma, column 8 = 4 n_mma, every other column = 96 n_mma  (if element 2i sits in the LOW nibble of byte i).
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128, 1)
fp4_probe_kernel(int n_mma, int n_cols, int per_commit, float* out) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 2 * kABytes);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar + 8);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    // A: 128 rows x 128 bytes, every byte 0x42 (element 2i = 0x2 = 1.0, element 2i+1 = 0x4 = 2.0); B: 128 rows, bytes 0x22 (ones)
    for (int i = threadIdx.x; i < 2 * kABytes / 4; i += blockDim.x)
        reinterpret_cast<uint32_t*>(smem)[i] = i < kABytes / 4 ? 0x42424242u : 0x22222222u;
    __syncthreads();
    if (rank == 0 && threadIdx.x == 0) {
        // rows 0 and 8 of the leader's half of B: row r of a 128B-swizzled tile starts at r * 128 and rows with r % 8 == 0 are
        // stored unpermuted.  Only K elements 0..63 (bytes 0..31) are ever multiplied.
        uint8_t* b0 = smem + kABytes;
        for (int k = 0; k < 32; ++k) { b0[k] = 0; b0[8 * 128 + k] = 0; }
        b0[0] = 0x02;                    // e_0
        b0[8 * 128 + 0] = 0x22;          // e_0 + e_1
        b0[8 * 128 + 1] = 0x02;          // + e_2
    }
    if (warp == 0 && lane == 0) {
        for (int d = 0; d < 8; ++d) mbar_init(&bar[d], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                     "r"((uint32_t)kTmemCols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    tmem_fill_sf_ones(tmem_base, warp);
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    tc_fence_after();
    if (warp == 1 && lane == 0 && rank == 0) {
        // block-scaled descriptor (cute/arch/mma_sm100_desc.hpp InstrDescriptorBlockScaled): a/b format E2M1 = 1 (bits 7, 10),
        // N >> 3 at bit 17, scale format E8M0 = 1 at bit 23, M >> 4 at bit 24, K = 64 (bit 31 = 0), both K-major
        const uint32_t idesc = (1u << 7) | (1u << 10) | ((static_cast<uint32_t>(n_cols) >> 3) << 17) | (1u << 23) | ((256u >> 4) << 24);
        const uint64_t a_desc = make_smem_desc(smem_u32(smem));
        const uint64_t b_desc = make_smem_desc(smem_u32(smem) + kABytes);
        const uint32_t sfa = tmem_base + kSfCols, sfb = tmem_base + kSfCols + 16;
        int issued = 0, it = 0;
        while (issued < n_mma) {
            const int d = it & 3;
            if (it >= 4) mbar_wait(&bar[d], ((it >> 2) - 1) & 1, 9);
            tc_fence_after();
            for (int m = 0; m < per_commit && issued < n_mma; ++m, ++issued)
                tc_mma_mxf4(tmem_base, a_desc, b_desc, idesc, sfa, sfb, issued != 0);
            tc_commit_pair(&bar[d]);
            ++it;
        }
        for (int k = it > 4 ? it - 4 : 0; k < it; ++k) mbar_wait(&bar[k & 3], (k >> 2) & 1, 9);
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    tc_fence_after();
    {
        int32_t v[16];
        tmem_ld16(tmem_base + (static_cast<uint32_t>(32 * warp) << 16), v);
        tmem_ld_wait();
        float* o = out + (static_cast<size_t>(rank) * 128 + 32 * warp + lane) * 16;
#pragma unroll
        for (int e = 0; e < 16; ++e) o[e] = __int_as_float(v[e]);
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)kTmemCols)
                     : "memory");
    }
}

float launch_fp4_probe(int n_mma, int n_cols, int per_commit, int n_pairs, float* d_out, cudaStream_t st) {
    const int smem = 2 * kABytes + 1024 + 256;
    CUDA_CHECK(cudaFuncSetAttribute(fp4_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    cudaEvent_t e0, e1;
    CUDA_CHECK(cudaEventCreate(&e0));
    CUDA_CHECK(cudaEventCreate(&e1));
    CUDA_CHECK(cudaEventRecord(e0, st));
    fp4_probe_kernel<<<2 * n_pairs, 128, smem, st>>>(n_mma, n_cols, per_commit, d_out);
    CUDA_CHECK(cudaGetLastError());
    CUDA_CHECK(cudaEventRecord(e1, st));
    CUDA_CHECK(cudaEventSynchronize(e1));
    float ms = 0;
    CUDA_CHECK(cudaEventElapsedTime(&ms, e0, e1));
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    return ms;
}

}  // namespace nldsc
