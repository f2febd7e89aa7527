// nldsc_ld_snp_stats_static: the per-SNP statistics (genotype counts -> MAF, residual std) of packed .bed rows that are ALREADY
// in device memory, written to device arrays the caller owns, without a single allocation -- for callers that cannot call
// cudaMalloc (code under CUDA graph capture, serving loops, harnesses that forbid it).  No counterpart in the reference.
// Nothing else in the library calls this file, and it changes no existing path.
//
// The session keeps its intermediate arrays (genotype counts, SnpCoef, flags) in arena blocks it allocates; here the same three
// arrays live in ONE file-scope __device__ array of fixed size: every GPU that loads the library holds kScratchBytes (19.25 MiB)
// until the process exits, and NLDSC_STATIC_MAX_SNPS (262,144) SNPs is the most one call can take.  All calls share the array:
//   * it keeps the last call's contents -- the counts are accumulated with atomics, so every call clears its part first;
//   * two calls that overlap on the GPU (different streams, different host threads) RACE on it.  Serialise them.
//
// The counting kernel is the counting half of unpack_kernel / unpack_fp4_kernel (unpack.cu): same slots of the packed bytes (all
// four pairs of a full byte, the high n_org % 4 pairs of a row's last byte, stream.h:60-66), same exact integer atomics, no
// planes written.  The statistics come from the session's own stats_kernel (launch_stats), reading and writing the static array.
#include "common.cuh"
#include "../../include/nldsc_ld.h"

#include <algorithm>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <stdexcept>
#include <string>

namespace nldsc {
namespace {

constexpr int32_t kMaxSnps = NLDSC_STATIC_MAX_SNPS;
constexpr size_t kCountsBytes = sizeof(int32_t) * 3 * kMaxSnps;            //  3.00 MiB   int32 [n][3] = het, hom2, missing
constexpr size_t kCoefBytes = sizeof(SnpCoef) * static_cast<size_t>(kMaxSnps);   // 16.00 MiB   what the band epilogue would read
constexpr size_t kFlagBytes = static_cast<size_t>(kMaxSnps);               //  0.25 MiB   kFlagKeep | kFlagDom
constexpr size_t kCoefOff = kCountsBytes;
constexpr size_t kFlagOff = kCoefOff + kCoefBytes;
constexpr size_t kScratchBytes = kFlagOff + kFlagBytes;
static_assert(kCoefOff % 256 == 0 && kFlagOff % 256 == 0, "sub-arrays stay 256-byte aligned");
static_assert(sizeof(SnpCoef) == 64, "scratch layout assumes 8 doubles per SNP");

// In HOST code this name is only a handle for cudaGetSymbolAddress -- never a device address (scratch_of below).
static __device__ __align__(256) unsigned char g_static_scratch[kScratchBytes];

__device__ __forceinline__ uint32_t last_byte_mask(int n_org) {
    const int r = n_org & 3;                       // bits of a row's LAST byte that hold real samples (high n_org % 4 pairs)
    return r == 0 ? 0xFFu : ((0xFFu << (8 - 2 * r)) & 0xFFu);
}

// grid.x = SNP rows, grid.y * blockDim.x threads stride over the aligned 32-bit words that overlap the row's bytes
// [row * bps, (row + 1) * bps).  Bytes of a word that belong to a neighbouring row are masked off; words that would reach
// past the end of the buffer are assembled from single bytes (the caller's buffer carries no slack).
__global__ void __launch_bounds__(256)
count_static_kernel(const uint8_t* __restrict__ packed, int64_t bps, int64_t total_bytes, int32_t n_org,
                    int32_t* __restrict__ counts) {
    const int64_t row = blockIdx.x;
    const int64_t b0 = row * bps, b1 = b0 + bps;
    const int64_t w0 = b0 >> 2, w1 = (b1 + 3) >> 2;
    const uint32_t* __restrict__ words = reinterpret_cast<const uint32_t*>(packed);
    const uint32_t tail = last_byte_mask(n_org);
    uint32_t nh = 0, na = 0, nm = 0;
    for (int64_t w = w0 + blockIdx.y * blockDim.x + threadIdx.x; w < w1; w += (int64_t)gridDim.y * blockDim.x) {
        const int64_t wb = w << 2;
        uint32_t x;
        if (wb + 4 <= total_bytes) {
            x = __ldg(words + w);
        } else {
            x = 0;
            for (int k = 0; k < 4 && wb + k < total_bytes; ++k) x |= (uint32_t)packed[wb + k] << (8 * k);
        }
        uint32_t keep = 0xFFFFFFFFu;
        if (wb < b0) keep &= 0xFFFFFFFFu << (8 * (int)(b0 - wb));             // bytes of the previous row
        if (wb + 4 > b1) keep &= 0xFFFFFFFFu >> (8 * (int)(wb + 4 - b1));     // bytes of the next row / past the end
        if (wb + 4 >= b1) {                                                     // holds the row's last byte
            const int s = 8 * (int)(b1 - 1 - wb);
            keep &= ~(0xFFu << s) | (tail << s);
        }
        x &= keep;
        const uint32_t lo = x & 0x55555555u, hi = (x >> 1) & 0x55555555u;
        na += __popc(hi & lo);      // 11 hom-second
        nh += __popc(hi & ~lo);     // 10 het
        nm += __popc(lo & ~hi);     // 01 missing
    }
    nh = __reduce_add_sync(0xFFFFFFFFu, nh);
    na = __reduce_add_sync(0xFFFFFFFFu, na);
    nm = __reduce_add_sync(0xFFFFFFFFu, nm);
    if ((threadIdx.x & 31) == 0 && (nh | na | nm)) {
        if (nh) atomicAdd(counts + row * 3 + 0, (int32_t)nh);
        if (na) atomicAdd(counts + row * 3 + 1, (int32_t)na);
        if (nm) atomicAdd(counts + row * 3 + 2, (int32_t)nm);
    }
}

// The device address of g_static_scratch on `device`, looked up once per device with that device current.
unsigned char* scratch_of(int device) {
    static std::mutex mu;
    static unsigned char* ptr[64] = {};
    if (device < 0 || device >= 64) throw std::invalid_argument("device index out of range");
    std::lock_guard<std::mutex> g(mu);
    if (ptr[device] == nullptr) {
        int cur = -1;
        CUDA_CHECK(cudaGetDevice(&cur));
        if (cur != device) throw CudaError("scratch_of: the device is not current");
        unsigned char* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_static_scratch);
        if (e != cudaSuccess || p == nullptr) {
            cudaGetLastError();
            throw CudaError(std::string("cudaGetSymbolAddress(g_static_scratch) failed: ") + cudaGetErrorString(e));
        }
        ptr[device] = p;
    }
    return ptr[device];
}

void set_msg(char* err, size_t errlen, const char* msg) {
    if (err && errlen) {
        strncpy(err, msg, errlen - 1);
        err[errlen - 1] = 0;
    }
}

}  // namespace
}  // namespace nldsc

extern "C" int nldsc_ld_snp_stats_static(const uint8_t* d_packed, int32_t n_snp, int32_t n_org, double maf, double std_thr,
                                         int32_t* d_counts_out, double* d_maf_out, double* d_rstd_out, int32_t device,
                                         void* stream, char* err, size_t errlen) {
    using namespace nldsc;
    int prev = -1;
    bool switched = false;
    int rc = NLDSC_OK;
    try {
        // arguments first: nothing below touches the GPU for a call that is going to be refused
        if (n_snp > kMaxSnps) {
            char buf[200];
            snprintf(buf, sizeof(buf), "nldsc_ld_snp_stats_static: n_snp = %d exceeds the static scratch limit of %d SNPs per call "
                                       "(NLDSC_STATIC_MAX_SNPS); split the rows over several calls", n_snp, kMaxSnps);
            throw std::invalid_argument(buf);
        }
        if (!d_packed || !d_maf_out || !d_rstd_out || n_snp <= 0 || n_org <= 0) throw std::invalid_argument("bad arguments");
        if (reinterpret_cast<uintptr_t>(d_packed) & 3) throw std::invalid_argument("d_packed must be 4-byte aligned");
        CUDA_CHECK(cudaGetDevice(&prev));
        if (prev != device) {
            CUDA_CHECK(cudaSetDevice(device));
            switched = true;
        }
        unsigned char* scratch = scratch_of(device);
        cudaStream_t st = static_cast<cudaStream_t>(stream);
        int32_t* counts = reinterpret_cast<int32_t*>(scratch);
        SnpCoef* coef = reinterpret_cast<SnpCoef*>(scratch + kCoefOff);
        uint8_t* flags = scratch + kFlagOff;
        const int64_t bps = (static_cast<int64_t>(n_org) + 3) / 4;
        // the counts are sums: clear this call's part of the shared array (coef and flags are overwritten row by row)
        CUDA_CHECK(cudaMemsetAsync(counts, 0, sizeof(int32_t) * 3 * static_cast<size_t>(n_snp), st));
        const int64_t row_words = (bps + 3) / 4 + 1;
        const unsigned gy = static_cast<unsigned>(std::min<int64_t>(64, std::max<int64_t>(1, (row_words + 1023) / 1024)));
        count_static_kernel<<<dim3(static_cast<unsigned>(n_snp), gy), 256, 0, st>>>(d_packed, bps, bps * n_snp, n_org, counts);
        CUDA_CHECK(cudaGetLastError());
        launch_stats(counts, 0, n_snp, n_org, maf, std_thr, coef, flags, d_maf_out, d_rstd_out, st);
        if (d_counts_out)
            CUDA_CHECK(cudaMemcpyAsync(d_counts_out, counts, sizeof(int32_t) * 3 * static_cast<size_t>(n_snp),
                                       cudaMemcpyDeviceToDevice, st));
    } catch (const std::invalid_argument& e) {
        set_msg(err, errlen, e.what());
        rc = NLDSC_ERR_INVALID;
    } catch (const std::exception& e) {
        set_msg(err, errlen, e.what());
        rc = NLDSC_ERR_RUNTIME;
    } catch (...) {
        set_msg(err, errlen, "unknown error");
        rc = NLDSC_ERR_RUNTIME;
    }
    if (switched) cudaSetDevice(prev);
    return rc;
}
