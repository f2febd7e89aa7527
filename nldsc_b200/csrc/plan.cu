// Window bounds of every SNP on the device: {j : |pos_j - pos_i| <= ld_wind}, the reference's
// SNPFilter::filter (tools.h:41-49) with the sliding-cache bounds of ChunkwiseReader
// (stream.h:142-155,182-197) for sorted non-negative positions.
//
// Same predicate, same IEEE double subtraction as the host routine compute_windows (session.cu), which
// stays the implementation behind the GPU-free planning entry points; this kernel exists because the
// host sweep is branch-misprediction bound (~16 ns per SNP) and sits on the critical path of every
// calculate call -- nothing can be scored before the band plan exists.
#include "common.cuh"

#include <algorithm>

namespace nldsc {

// Threads n <= i < n_pad write the empty window (1, 0) of a padding row; `total` (optional) receives
// sum_i (hi[i] - lo[i]) = the ordered in-window pairs before the MAF filter.
__global__ void __launch_bounds__(256)
windows_kernel(const double* __restrict__ pos, int32_t n, int32_t n_pad, double wind, int32_t* __restrict__ lo,
               int32_t* __restrict__ hi, unsigned long long* __restrict__ total) {
    const int32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    int32_t width = 0;
    if (i < n) {
        const double pi = pos[i];
        // the predicate is monotone on either side of i (sorted positions, monotone IEEE subtraction)
        int32_t x = 0, y = i;                       // smallest l in [0, i] inside the window
        while (x < y) {
            const int32_t m = x + ((y - x) >> 1);
            if (fabs(pos[m] - pi) <= wind) y = m; else x = m + 1;
        }
        lo[i] = x;
        width = -x;
        x = i; y = n - 1;                           // largest r in [i, n-1] inside the window
        while (x < y) {
            const int32_t m = x + ((y - x + 1) >> 1);
            if (fabs(pos[m] - pi) <= wind) x = m; else y = m - 1;
        }
        hi[i] = x;
        width += x;
    } else if (i < n_pad) {
        lo[i] = 1;
        hi[i] = 0;
    }
    if (total != nullptr) {                         // grid-uniform
        unsigned long long w = static_cast<unsigned long long>(width);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) w += __shfl_xor_sync(0xFFFFFFFFu, w, o);
        if ((threadIdx.x & 31) == 0 && w) atomicAdd(total, w);
    }
}

void launch_windows(const double* pos, int32_t n, double wind, int32_t* lo, int32_t* hi, cudaStream_t st,
                    int32_t n_pad, unsigned long long* total) {
    if (n <= 0) return;
    if (n_pad < n) n_pad = n;
    windows_kernel<<<(n_pad + 255) / 256, 256, 0, st>>>(pos, n, n_pad, wind, lo, hi, total);
    CUDA_CHECK(cudaGetLastError());
}

// Small-array transport that stays off the copy engines: while a shard's packed rows are queued on the
// H2D engine (hundreds of MB, FIFO), a cudaMemcpy of the band plan would wait behind all of them.  The SMs
// read / write pinned host memory directly instead (UVA), sharing PCIe with the bulk upload.
__global__ void __launch_bounds__(256)
copy16_kernel(const uint4* __restrict__ src, uint4* __restrict__ dst, size_t n16) {
    for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n16;
         i += static_cast<size_t>(gridDim.x) * blockDim.x)
        dst[i] = src[i];
}

void launch_copy16(const void* src, void* dst, size_t bytes, cudaStream_t st) {
    const size_t n16 = (bytes + 15) / 16;       // both buffers are padded to 16 bytes by their allocators
    if (n16 == 0) return;
    const int blocks = static_cast<int>(std::min<size_t>((n16 + 255) / 256, 1184));
    copy16_kernel<<<blocks, 256, 0, st>>>(static_cast<const uint4*>(src), static_cast<uint4*>(dst), n16);
    CUDA_CHECK(cudaGetLastError());
}

// One thread: *ready_rows = value with release semantics at GPU scope -- everything the stream ran before this kernel
// (unpack, stats) is visible to a band-kernel thread that reads the value with an acquire load (band_tc.cu wait_rows).
__global__ void publish_rows_kernel(uint32_t* ready_rows, uint32_t value) {
    __threadfence();
    asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(ready_rows), "r"(value) : "memory");
}

void launch_publish_rows(uint32_t* ready_rows, uint32_t value, cudaStream_t st) {
    publish_rows_kernel<<<1, 1, 0, st>>>(ready_rows, value);
    CUDA_CHECK(cudaGetLastError());
}

void preload_plan_kernels() {
    cudaFuncAttributes fa;
    CUDA_CHECK(cudaFuncGetAttributes(&fa, publish_rows_kernel));
    CUDA_CHECK(cudaFuncGetAttributes(&fa, copy16_kernel));
}

}  // namespace nldsc
