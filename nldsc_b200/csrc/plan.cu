// Window bounds of every SNP on the device: {j : |pos_j - pos_i| <= ld_wind}, the reference's
// SNPFilter::filter (tools.h:41-49) with the sliding-cache bounds of ChunkwiseReader
// (stream.h:142-155,182-197) for sorted non-negative positions.
//
// Same predicate, same IEEE double subtraction as the host routine compute_windows (session.cu), which
// stays the implementation behind the GPU-free planning entry points; this kernel exists because the
// host sweep is branch-misprediction bound (~16 ns per SNP) and sits on the critical path of every
// calculate call -- nothing can be scored before the band plan exists.
#include "common.cuh"

namespace nldsc {

__global__ void __launch_bounds__(256)
windows_kernel(const double* __restrict__ pos, int32_t n, double wind, int32_t* __restrict__ lo,
               int32_t* __restrict__ hi) {
    const int32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double pi = pos[i];
    // the predicate is monotone on either side of i (sorted positions, monotone IEEE subtraction)
    int32_t x = 0, y = i;                       // smallest l in [0, i] inside the window
    while (x < y) {
        const int32_t m = x + ((y - x) >> 1);
        if (fabs(pos[m] - pi) <= wind) y = m; else x = m + 1;
    }
    lo[i] = x;
    x = i; y = n - 1;                           // largest r in [i, n-1] inside the window
    while (x < y) {
        const int32_t m = x + ((y - x + 1) >> 1);
        if (fabs(pos[m] - pi) <= wind) x = m; else y = m - 1;
    }
    hi[i] = x;
}

void launch_windows(const double* pos, int32_t n, double wind, int32_t* lo, int32_t* hi, cudaStream_t st) {
    if (n <= 0) return;
    windows_kernel<<<(n + 255) / 256, 256, 0, st>>>(pos, n, wind, lo, hi);
    CUDA_CHECK(cudaGetLastError());
}

}  // namespace nldsc
