// PLINK .bed 2-bit decode, per-SNP genotype counts and per-SNP statistics.
//
// Replaces (reference, nldsc/ldscore/_ldscore/):
//   BedStreamReader::read            stream.h:43-69    -> unpack_kernel (planes + genotype counts)
//   SNPInMemory::decode + ctor       encoder.h:54-60,91-133
//   Math::regression_residuals       tools.h:54-68     -> stats_kernel (closed forms over
//   Math::var_ / standardise         tools.h:70-85        the three genotype counts)
//
// Sample order: the reference assigns sample 4j+i to bit pair (7-2i,6-2i) of byte j and,
// when n_org % 4 != 0, takes only the first n_org%4 (HIGH) pairs of the last byte
// (stream.h:60-66).  Every quantity downstream is a sum over samples, so only
// the SET of (byte, pair) slots matters, not their order: the kernels below keep all four
// pairs of every full byte and the high n_org%4 pairs of the last byte, and place slot
// (byte j, LSB-pair k) at organism index 4j+k of the plane row.
//
// This translation unit is compiled with -fmad=false: the MAF path must round exactly
// like the reference's float code, and the closed forms rely on exact power-of-two
// scaling so that collinear dominance codings give a residual std of exactly zero.
#include "common.cuh"

namespace nldsc {

__device__ __forceinline__ uint32_t tail_pair_mask(int n_org) {
    // bits of the LAST byte of a row that hold real samples (high n_org%4 pairs)
    const int r = n_org & 3;
    return r == 0 ? 0xFFu : ((0xFFu << (8 - 2 * r)) & 0xFFu);
}

// One thread per SNP: everything the filters and the epilogue need, from three counts.
__global__ void __launch_bounds__(256)
stats_kernel(const int32_t* __restrict__ counts, int32_t row0, int32_t n_rows, int32_t n_org, double maf_thr,
             double std_thr, SnpCoef* __restrict__ coef, uint8_t* __restrict__ flags,
             double* __restrict__ maf_out, double* __restrict__ rstd_out, const uint8_t* __restrict__ unused,
             uint8_t* __restrict__ blk_miss) {
    int row = blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= n_rows) return;
    row += row0;
    const double nh = counts[row * 3 + 0], na = counts[row * 3 + 1], nm = counts[row * 3 + 2];
    const double n = n_org;
    const double n_obs = n - nm;
    // encoder.h:104-118 -- double sums, float means, float MAF
    const double add_sum = nh + 2.0 * na;
    const double nadd_sum = 2.0 * (nh + na);
    const float mu_a_f = (float)(add_sum / n_obs);
    const float mu_d_f = (float)(nadd_sum / n_obs);
    const float f2 = mu_a_f / 2.0f;
    const float maf_f = ((double)f2 < 0.5) ? f2 : 1.0f - f2;
    // position < 0: the reference never reads the SNP (stream.h:189-194), its MAF stays NaN (encoder.h:47)
    const bool is_unused = unused != nullptr && unused[row] != 0;
    const bool keep = !is_unused && !((double)maf_f <= maf_thr);         // NaN passes
    maf_out[row] = is_unused ? __longlong_as_double(0x7FF8000000000000LL) : (double)maf_f;

    SnpCoef c;
    uint8_t fl = 0;
    double rsd_out = __longlong_as_double(0x7FF8000000000000LL);
    if (keep) {
        const double mu = (double)mu_a_f, mud = (double)mu_d_f;   // imputed values (encoder.h:124-129)
        const double n0 = n - nh - na - nm;
        const double xbar = (nh + 2.0 * na + mu * nm) / n;
        const double ybar = (2.0 * nh + 2.0 * na + mud * nm) / n;
        const double sxx = nh + 4.0 * na + (mu * mu) * nm;
        const double sxy = 2.0 * nh + 4.0 * na + (mu * mud) * nm;
        const double beta = (sxy / n - xbar * ybar) / (sxx / n - xbar * xbar);   // tools.h:65-66
        const double ch = 2.0 - beta, ca = 2.0 - 2.0 * beta, cm = mud - beta * mu;
        // population variance, summed per genotype class (tools.h:70-73)
        const double d0 = xbar, d1 = 1.0 - xbar, d2 = 2.0 - xbar, dm = mu - xbar;
        const double var_x = (n0 * (d0 * d0) + nh * (d1 * d1) + na * (d2 * d2) + nm * (dm * dm)) / n;
        const double sd = sqrt(var_x);
        const double rbar = (nh * ch + na * ca + nm * cm) / n;
        const double e1 = ch - rbar, e2 = ca - rbar, em = cm - rbar;
        const double var_r = (n0 * (rbar * rbar) + nh * (e1 * e1) + na * (e2 * e2) + nm * (em * em)) / n;
        const double rsd = sqrt(var_r);
        c.mu = mu; c.xbar = xbar; c.inv_sd = 1.0 / sd;
        c.ch = ch; c.ca = ca; c.cm = cm; c.rbar = rbar; c.inv_rsd = 1.0 / rsd;
        fl = kFlagKeep | ((rsd > std_thr) ? kFlagDom : 0);
        rsd_out = rsd;
        // the block can take the missing-free band kernel only if none of its KEPT SNPs has a missing genotype
        // (a dropped SNP's counts never reach a score); benign race: every writer stores the same 1
        if (blk_miss != nullptr && nm > 0.0) blk_miss[row / kBlk] = 1;
    } else {
        c.mu = c.xbar = c.inv_sd = c.ch = c.ca = c.cm = c.rbar = c.inv_rsd = 0.0;
    }
    coef[row] = c;
    flags[row] = fl;
    rstd_out[row] = rsd_out;
}

#ifndef NLDSC_STREAM_STORES
#define NLDSC_STREAM_STORES 1
#endif
constexpr bool kStreamStores = NLDSC_STREAM_STORES != 0;

// byte with four 2-bit genotype codes -> one 32-bit word per plane, byte k = sample k
__device__ __forceinline__ void expand_byte(uint32_t b, uint32_t& g, uint32_t& m, uint32_t& h, uint32_t& a) {
    const uint32_t p = (b | (b << 6) | (b << 12) | (b << 18)) & 0x03030303u;
    const uint32_t lo = p & 0x01010101u, hi = (p >> 1) & 0x01010101u;
    a = hi & lo;          // 11 hom-second
    h = hi ^ a;           // 10 het
    m = lo ^ a;           // 01 missing
    g = hi + a;           // additive dosage 0/1/2
}

// Plane rows are written once and next read by the band kernel's TMA loads, usually after far more than an
// L2's worth of other rows: store them evict-first so they do not displace the operand tiles the band
// kernel of the previous chunk (running beside this kernel) is re-reading from L2.
__device__ __forceinline__ void st_plane(int8_t* p, const uint4& v) {
    if (kStreamStores) __stcs(reinterpret_cast<uint4*>(p), v);
    else *reinterpret_cast<uint4*>(p) = v;
}

// grid.x = rows, grid.y * blockDim.x covers npad/16 sixteen-organism groups.
// Each thread turns 4 packed bytes into 16 bytes of each of the 4 planes (coalesced 16-B stores) and,
// when `counts` is given, popcounts the same word: the genotype counts of the row come out of the
// pass that reads the bytes anyway (integer atomics: exact and order-independent).
__global__ void __launch_bounds__(256)
unpack_kernel(const uint8_t* __restrict__ packed, int64_t bps, int32_t row0, int32_t n_org,
              int32_t npad, int8_t* __restrict__ planes, int64_t plane_stride, int32_t dst_row0,
              int32_t* __restrict__ counts) {
    const int t = blockIdx.y * blockDim.x + threadIdx.x;      // 16-organism group
    const int64_t row = (int64_t)row0 + blockIdx.x;
    const int64_t rb = (int64_t)t * 4;                         // row-relative first byte
    uint32_t x = 0;
    if (rb < bps) {
        const int64_t addr = row * bps + rb;
        const uint32_t* __restrict__ words = reinterpret_cast<const uint32_t*>(packed);
        const int64_t w = addr >> 2;
        const int sh = (int)(addr & 3) * 8;
        const uint32_t w0 = __ldg(words + w);
        const uint32_t w1 = sh ? __ldg(words + w + 1) : 0u;   // buffer carries 8 bytes of slack
        x = __funnelshift_r(w0, w1, sh);
        const int64_t left = bps - rb;                          // bytes of this row still ahead
        if (left < 4) x &= 0xFFFFFFFFu >> (8 * (int)(4 - left));
        if (left <= 4) {                                        // holds the row's last byte
            const int s2 = 8 * (int)(left - 1);
            x &= ~(0xFFu << s2) | (tail_pair_mask(n_org) << s2);
        }
    }
    if (t * 16 < npad) {
        uint4 g, m, h, a;
        expand_byte(x & 0xFFu, g.x, m.x, h.x, a.x);
        expand_byte((x >> 8) & 0xFFu, g.y, m.y, h.y, a.y);
        expand_byte((x >> 16) & 0xFFu, g.z, m.z, h.z, a.z);
        expand_byte(x >> 24, g.w, m.w, h.w, a.w);
        const int64_t off = ((int64_t)dst_row0 + blockIdx.x) * npad + (int64_t)t * 16;
        st_plane(planes + kPlaneG * plane_stride + off, g);
        st_plane(planes + kPlaneM * plane_stride + off, m);
        st_plane(planes + kPlaneH * plane_stride + off, h);
        st_plane(planes + kPlaneA * plane_stride + off, a);
    }
    if (counts != nullptr) {                                    // block-uniform
        const uint32_t lo = x & 0x55555555u, hi = (x >> 1) & 0x55555555u;
        // three counts <= 16 each in one register: bits 0-9 het, 10-19 hom-second, 20-29 missing
        uint32_t c = __popc(hi & ~lo) | (__popc(hi & lo) << 10) | (__popc(lo & ~hi) << 20);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xFFFFFFFFu, c, o);    // <= 512 per field
        __shared__ uint32_t part[8];
        if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = c;
        __syncthreads();
        if (threadIdx.x < 3) {
            uint32_t tot = 0;
#pragma unroll
            for (int w = 0; w < 8; ++w) tot += (part[w] >> (10 * threadIdx.x)) & 0x3FFu;
            if (tot) atomicAdd(counts + row * 3 + threadIdx.x, (int32_t)tot);
        }
    }
}

// ---- fp4 planes ---------------------------------------------------------------------------------------------
// The same four planes with every genotype value stored as an e2m1 (fp4) nibble -- 0 -> 0x0, 1 -> 0x2, 2 -> 0x4 -- organism
// 2i in the LOW nibble of byte i: the operand format of tcgen05.mma.kind::mxf4, which multiplies 64 organisms per instruction
// at twice the int8 rate from half the bytes and (scale factors 1.0) sums these small integers exactly in FP32
// (nldsc_ld_debug_fp4_probe).  A plane row is npad BYTES = 2 npad organisms.
// four 2-bit codes per byte, 8 organisms (16 bits) -> eight 4-bit fields holding the code's low / high bit at bits 0 / 1
__device__ __forceinline__ uint32_t spread_codes8(uint32_t v16) {
    uint32_t t = v16 & 0xFFFFu;
    t = (t | (t << 8)) & 0x00FF00FFu;
    t = (t | (t << 4)) & 0x0F0F0F0Fu;
    t = (t | (t << 2)) & 0x33333333u;
    return t;
}
__device__ __forceinline__ void expand8_fp4(uint32_t v16, uint32_t& g, uint32_t& m, uint32_t& h, uint32_t& a) {
    const uint32_t t = spread_codes8(v16);
    const uint32_t lo = t & 0x11111111u, hi = (t >> 1) & 0x11111111u;
    const uint32_t aa = hi & lo, hh = hi ^ aa, mm = lo ^ aa;     // 11 hom-second, 10 het, 01 missing: one bit per nibble (bit 0)
    a = aa << 1;                   // 1.0 = 0b0010
    h = hh << 1;
    m = mm << 1;
    g = (hh << 1) | (aa << 2);     // dosage: 1 -> 0b0010, 2 -> 0b0100
}

// grid.x = rows, grid.y * blockDim.x covers npad / 16 groups of 32 organisms (8 packed bytes in, 16 bytes per plane out)
__global__ void __launch_bounds__(256)
unpack_fp4_kernel(const uint8_t* __restrict__ packed, int64_t bps, int32_t row0, int32_t n_org,
                  int32_t npad, int8_t* __restrict__ planes, int64_t plane_stride, int32_t dst_row0,
                  int32_t* __restrict__ counts) {
    const int t = blockIdx.y * blockDim.x + threadIdx.x;      // 32-organism group
    const int64_t row = (int64_t)row0 + blockIdx.x;
    const int64_t rb = (int64_t)t * 8;                         // row-relative first byte
    uint32_t x0 = 0, x1 = 0;
    if (rb < bps) {
        const int64_t addr = row * bps + rb;
        const uint32_t* __restrict__ words = reinterpret_cast<const uint32_t*>(packed);
        const int64_t w = addr >> 2;
        const int sh = (int)(addr & 3) * 8;
        const int64_t left = bps - rb;                          // bytes of this row still ahead (>= 1)
        const uint32_t w0 = __ldg(words + w);
        const uint32_t w1 = __ldg(words + w + 1);               // the buffer carries 16 bytes of slack behind its last row
        const uint32_t w2 = sh ? __ldg(words + w + 2) : 0u;
        x0 = __funnelshift_r(w0, w1, sh);
        x1 = __funnelshift_r(w1, w2, sh);
        if (left < 8) {                                         // bytes past the row's end are the next row's: drop them
            const unsigned long long keep = left >= 8 ? ~0ull : ((1ull << (8 * (int)left)) - 1ull);
            x0 &= (uint32_t)keep;
            x1 &= (uint32_t)(keep >> 32);
        }
        if (left <= 8) {                                        // holds the row's last byte: only its high n_org % 4 pairs are samples
            const int s2 = 8 * (int)(left - 1);
            const unsigned long long mask = ~(0xFFull << s2) | ((unsigned long long)tail_pair_mask(n_org) << s2);
            x0 &= (uint32_t)mask;
            x1 &= (uint32_t)(mask >> 32);
        }
    }
    if (t * 16 < npad) {
        uint4 g, m, h, a;
        expand8_fp4(x0, g.x, m.x, h.x, a.x);
        expand8_fp4(x0 >> 16, g.y, m.y, h.y, a.y);
        expand8_fp4(x1, g.z, m.z, h.z, a.z);
        expand8_fp4(x1 >> 16, g.w, m.w, h.w, a.w);
        const int64_t off = ((int64_t)dst_row0 + blockIdx.x) * npad + (int64_t)t * 16;
        st_plane(planes + kPlaneG * plane_stride + off, g);
        st_plane(planes + kPlaneM * plane_stride + off, m);
        st_plane(planes + kPlaneH * plane_stride + off, h);
        st_plane(planes + kPlaneA * plane_stride + off, a);
    }
    if (counts != nullptr) {                                    // block-uniform
        const uint32_t lo0 = x0 & 0x55555555u, hi0 = (x0 >> 1) & 0x55555555u;
        const uint32_t lo1 = x1 & 0x55555555u, hi1 = (x1 >> 1) & 0x55555555u;
        // three counts <= 32 each in one 64-bit register, 16 bits apart: het, hom-second, missing (a warp sums to <= 1,024 each)
        unsigned long long c = (unsigned long long)(__popc(hi0 & ~lo0) + __popc(hi1 & ~lo1)) |
                               ((unsigned long long)(__popc(hi0 & lo0) + __popc(hi1 & lo1)) << 16) |
                               ((unsigned long long)(__popc(lo0 & ~hi0) + __popc(lo1 & ~hi1)) << 32);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xFFFFFFFFu, c, o);
        __shared__ unsigned long long part[8];
        if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = c;
        __syncthreads();
        if (threadIdx.x < 3) {
            uint32_t tot = 0;
#pragma unroll
            for (int w = 0; w < 8; ++w) tot += (uint32_t)(part[w] >> (16 * threadIdx.x)) & 0xFFFFu;
            if (tot) atomicAdd(counts + row * 3 + threadIdx.x, (int32_t)tot);
        }
    }
}

__global__ void __launch_bounds__(256)
zero_rows_kernel(int8_t* __restrict__ planes, int64_t plane_stride, int32_t npad, int32_t row0) {
    const int t = blockIdx.y * blockDim.x + threadIdx.x;
    if (t * 16 >= npad) return;
    const int64_t off = ((int64_t)row0 + blockIdx.x) * npad + (int64_t)t * 16;
    const uint4 z = make_uint4(0, 0, 0, 0);
#pragma unroll
    for (int p = 0; p < kPlanes; ++p) *reinterpret_cast<uint4*>(planes + p * plane_stride + off) = z;
}

void launch_stats(const int32_t* counts, int32_t row0, int32_t n_rows, int32_t n_org, double maf_thr,
                  double std_thr, SnpCoef* coef, uint8_t* flags, double* maf_out,
                  double* rstd_out, cudaStream_t st, const uint8_t* unused, uint8_t* blk_miss) {
    if (n_rows <= 0) return;
    stats_kernel<<<(n_rows + 255) / 256, 256, 0, st>>>(counts, row0, n_rows, n_org, maf_thr, std_thr,
                                                       coef, flags, maf_out, rstd_out, unused, blk_miss);
    CUDA_CHECK(cudaGetLastError());
}

void launch_unpack(const uint8_t* packed, int64_t bps, int32_t row0, int32_t n_rows,
                   int32_t n_org, int32_t npad, int8_t* planes, int64_t plane_stride,
                   int32_t dst_row0, int32_t* counts, cudaStream_t st) {
    if (n_rows <= 0) return;
    const int groups = npad / 16;
    dim3 grid(n_rows, (groups + 255) / 256);
    unpack_kernel<<<grid, 256, 0, st>>>(packed, bps, row0, n_org, npad, planes, plane_stride, dst_row0, counts);
    CUDA_CHECK(cudaGetLastError());
}

void launch_unpack_fp4(const uint8_t* packed, int64_t bps, int32_t row0, int32_t n_rows,
                       int32_t n_org, int32_t npad, int8_t* planes, int64_t plane_stride,
                       int32_t dst_row0, int32_t* counts, cudaStream_t st) {
    if (n_rows <= 0) return;
    const int groups = npad / 16;
    dim3 grid(n_rows, (groups + 255) / 256);
    unpack_fp4_kernel<<<grid, 256, 0, st>>>(packed, bps, row0, n_org, npad, planes, plane_stride, dst_row0, counts);
    CUDA_CHECK(cudaGetLastError());
}

void launch_zero_rows(int8_t* planes, int64_t plane_stride, int32_t npad, int32_t row0,
                      int32_t n_rows, cudaStream_t st) {
    if (n_rows <= 0) return;
    const int groups = npad / 16;
    dim3 grid(n_rows, (groups + 255) / 256);
    zero_rows_kernel<<<grid, 256, 0, st>>>(planes, plane_stride, npad, row0);
    CUDA_CHECK(cudaGetLastError());
}

// *flag = 1 when no 64-SNP block in [block0, block1) holds a kept SNP with a missing genotype (one block of 256 threads)
__global__ void __launch_bounds__(256)
chunk_mfree_flag_kernel(const uint8_t* __restrict__ blk_miss, int32_t block0, int32_t block1, uint32_t* __restrict__ flag) {
    int any = 0;
    for (int32_t b = block0 + threadIdx.x; b < block1; b += blockDim.x) any |= blk_miss[b];
    any = __syncthreads_or(any);
    if (threadIdx.x == 0) *flag = any ? 0u : 1u;
}

void launch_chunk_mfree_flag(const uint8_t* blk_miss, int32_t block0, int32_t block1, uint32_t* flag, cudaStream_t st) {
    chunk_mfree_flag_kernel<<<1, 256, 0, st>>>(blk_miss, block0, block1, flag);
    CUDA_CHECK(cudaGetLastError());
}

void preload_unpack_kernels() {
    cudaFuncAttributes fa;
    CUDA_CHECK(cudaFuncGetAttributes(&fa, stats_kernel));
    CUDA_CHECK(cudaFuncGetAttributes(&fa, unpack_kernel));
    CUDA_CHECK(cudaFuncGetAttributes(&fa, unpack_fp4_kernel));
    CUDA_CHECK(cudaFuncGetAttributes(&fa, zero_rows_kernel));
    CUDA_CHECK(cudaFuncGetAttributes(&fa, chunk_mfree_flag_kernel));
}

}  // namespace nldsc
