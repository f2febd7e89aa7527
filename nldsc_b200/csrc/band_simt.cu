// (1) SIMT yardstick for the band kernel: the same tiles, the same epilogue, the same
//     partial layout, but the cross-counts come from __dp4a on CUDA cores.  It exists to
//     tell a tensor-core descriptor bug from an epilogue bug (tests compare the two and the
//     NumPy cross-counts); it is selected only by NLDSC_BAND=simt and is never a fallback.
// (2) finalize kernel: deterministic reduction of the per-tile row partials
//     (replaces the OpenMP reduction and the result stores of ldscalc.h:34-54).
#include "common.cuh"
#include "epilogue.cuh"

namespace nldsc {

namespace {
constexpr int kChunk = 64;            // organisms per shared-memory pass
constexpr int kPitch = kChunk / 4 + 1;   // words per row (+1: fewer bank conflicts)
}

// block = one (supertile, rank) = one 64x64 SNP-pair tile; 256 threads; thread (s = tid/4, cg = tid%4)
// scores SNP s of the i-block against j-SNPs cg*16 .. cg*16+15.
__global__ void __launch_bounds__(256)
band_simt_kernel(const int8_t* __restrict__ planes, int64_t plane_stride, int32_t npad, const BandArgs args) {
    __shared__ uint32_t sh[5][kBlk][kPitch];   // g_i, m_i, m_j, h_j, a_j
    const int t = blockIdx.x >> 1, rank = blockIdx.x & 1;
    const Tile tl = args.tiles[t];
    const int32_t a_row = tl.a_row[rank], i0 = tl.i0[rank];
    const int s = threadIdx.x >> 2, cg = threadIdx.x & 3;

    int32_t acc[6][16];
#pragma unroll
    for (int p = 0; p < 6; ++p)
#pragma unroll
        for (int c = 0; c < 16; ++c) acc[p][c] = 0;

    const int8_t* src[5] = {planes + kPlaneG * plane_stride + (int64_t)a_row * npad,
                            planes + kPlaneM * plane_stride + (int64_t)a_row * npad,
                            planes + kPlaneM * plane_stride + (int64_t)tl.b_row * npad,
                            planes + kPlaneH * plane_stride + (int64_t)tl.b_row * npad,
                            planes + kPlaneA * plane_stride + (int64_t)tl.b_row * npad};

    for (int k0 = 0; k0 < npad; k0 += kChunk) {
        __syncthreads();
        for (int e = threadIdx.x; e < 5 * kBlk * (kChunk / 4); e += blockDim.x) {
            const int w = e % (kChunk / 4), r = (e / (kChunk / 4)) % kBlk, p = e / (kBlk * (kChunk / 4));
            uint32_t v = 0;
            if (k0 + 4 * w < npad)
                v = *reinterpret_cast<const uint32_t*>(src[p] + (int64_t)r * npad + k0 + 4 * w);
            sh[p][r][w] = v;
        }
        __syncthreads();
#pragma unroll 4
        for (int w = 0; w < kChunk / 4; ++w) {
            const int gi = (int)sh[0][s][w], mi = (int)sh[1][s][w];
#pragma unroll
            for (int c = 0; c < 16; ++c) {
                const int j = cg * 16 + c;
                const int mj = (int)sh[2][j][w], hj = (int)sh[3][j][w], aj = (int)sh[4][j][w];
                acc[0][c] = __dp4a(gi, mj, acc[0][c]);
                acc[1][c] = __dp4a(gi, hj, acc[1][c]);
                acc[2][c] = __dp4a(gi, aj, acc[2][c]);
                acc[3][c] = __dp4a(mi, mj, acc[3][c]);
                acc[4][c] = __dp4a(mi, hj, acc[4][c]);
                acc[5][c] = __dp4a(mi, aj, acc[5][c]);
            }
        }
    }

    if (args.dump != nullptr) {
        // same layout as the tensor-core kernel's accumulator dump
        int32_t* d = args.dump + (int64_t)blockIdx.x * 128 * 192;
        const int rg = 32 * (s / 16) + (s % 16), rm = rg + 16;
#pragma unroll
        for (int c = 0; c < 16; ++c) {
            const int j = cg * 16 + c;
            d[rg * 192 + j] = acc[0][c];
            d[rg * 192 + 64 + j] = acc[1][c];
            d[rg * 192 + 128 + j] = acc[2][c];
            d[rm * 192 + j] = acc[3][c];
            d[rm * 192 + 64 + j] = acc[4][c];
            d[rm * 192 + 128 + j] = acc[5][c];
        }
    }

    const RowCtx row = load_row_ctx(args, i0 + s);
    RowAcc ra;
#pragma unroll
    for (int c = 0; c < 16; ++c)
        pair_eval(args, row, tl.j0 + cg * 16 + c, acc[0][c], acc[1][c], acc[2][c], acc[3][c], acc[4][c], acc[5][c], ra);
    // the 4 threads of a row are adjacent lanes: cg 0,1 score j 0..31 (half 0), cg 2,3 score j 32..63 (half 1)
    ra.add += __shfl_xor_sync(0xFFFFFFFFu, ra.add, 1);
    ra.dom += __shfl_xor_sync(0xFFFFFFFFu, ra.dom, 1);
    ra.wsa += __shfl_xor_sync(0xFFFFFFFFu, ra.wsa, 1);
    ra.wsd += __shfl_xor_sync(0xFFFFFFFFu, ra.wsd, 1);
    ra.wsde += __shfl_xor_sync(0xFFFFFFFFu, ra.wsde, 1);
    if ((cg & 1) == 0) {
        const int64_t o = part_index(tl.slot, rank, cg >> 1, s);
        args.part.add[o] = ra.add;
        args.part.dom[o] = ra.dom;
        args.part.wsa[o] = ra.wsa;
        args.part.wsd[o] = ra.wsd;
        args.part.wsde[o] = ra.wsde;
    }
}

void launch_band_simt(const int8_t* planes, int64_t plane_stride, int32_t npad, const BandArgs& args,
                      cudaStream_t st) {
    if (args.n_tiles <= 0) return;
    band_simt_kernel<<<2 * args.n_tiles, 256, 0, st>>>(planes, plane_stride, npad, args);
    CUDA_CHECK(cudaGetLastError());
}

// One thread per owned row: add the row's tile partials in ascending j-block order
// (fixed order => bitwise reproducible, independent of sharding), then apply the
// conventions of ldscalc.h:13-21,28,49-54.
__global__ void __launch_bounds__(256)
finalize_kernel(const FinalizeArgs a) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;          // row within this chunk's blocks
    if (idx >= a.n_iblocks * kBlk) return;
    const int b = idx / kBlk, r = idx % kBlk;
    const int row = (a.iblock0 + b) * kBlk + r;                     // shard-local SNP index
    if (row < a.own_begin || row >= a.own_end) return;
    const int o = row - a.own_begin;
    if (!(a.flags[row] & kFlagKeep)) {
        const double nan = __longlong_as_double(0x7FF8000000000000LL);
        a.l2[o] = nan; a.l2d[o] = nan;
        a.wsa[o] = -1; a.wsd[o] = -1; a.wsde[o] = -1;
        return;
    }
    double add = 1.0, dom = 0.0;                                     // ldscalc.h:27-28
    int wsa = 0, wsd = 0, wsde = 0;
    // i-block b of the chunk is rank b % 2 of supertile row b / 2
    for (int t = a.tile_begin[b >> 1]; t < a.tile_begin[(b >> 1) + 1]; ++t) {
#pragma unroll
        for (int half = 0; half < 2; ++half) {
            const int64_t p = part_index(t, b & 1, half, r);
            add += a.part.add[p];
            dom += a.part.dom[p];
            wsa += a.part.wsa[p];
            wsd += a.part.wsd[p];
            wsde += a.part.wsde[p];
        }
    }
    a.l2[o] = add; a.l2d[o] = dom;
    a.wsa[o] = wsa; a.wsd[o] = wsd; a.wsde[o] = wsde;
}

void launch_finalize(const FinalizeArgs& a, cudaStream_t st) {
    if (a.n_iblocks <= 0) return;
    const int n = a.n_iblocks * kBlk;
    finalize_kernel<<<(n + 255) / 256, 256, 0, st>>>(a);
    CUDA_CHECK(cudaGetLastError());
}

void preload_finalize_kernels() {
    cudaFuncAttributes fa;
    CUDA_CHECK(cudaFuncGetAttributes(&fa, finalize_kernel));
}

}  // namespace nldsc
