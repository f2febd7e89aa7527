// C ABI of the LD-score engine (include/nldsc_ld.h): shard planning, device buffers, the
// chunked  upload -> count -> stats -> unpack -> band GEMM + epilogue -> finalize  pipeline,
// and the multi-GPU driver (contiguous SNP shards with a window halo, no collective).
//
// Replaces calculate() (reference ldscalc.h:8-65) and ChunkwiseReader (stream.h:106-198):
// the reference slides a cache over the .bed one SNP at a time; here the host computes
// every SNP's window up front (compute_windows), cuts the band into 64x64 SNP-pair tiles
// and streams SNP rows through a ring of unpacked genotype planes in HBM.
#include "session.cuh"

#include <algorithm>
#include <cerrno>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <fcntl.h>
#include <limits>
#include <memory>
#include <mutex>
#include <string>
#include <sys/stat.h>
#include <thread>
#include <unistd.h>

using namespace nldsc;

namespace nldsc {

void compute_windows(const double* pos, int32_t n_snp, double wind, std::vector<int32_t>& lo,
                     std::vector<int32_t>& hi) {
    lo.resize(n_snp);
    hi.resize(n_snp);
    int32_t l = 0, r = 0;
    for (int32_t i = 0; i < n_snp; ++i) {
        // SNPFilter::filter: std::abs(positions[snd] - positions[fst]) <= ld_wind  (tools.h:43-44)
        while (l < i && !(std::abs(pos[l] - pos[i]) <= wind)) ++l;
        if (r < i) r = i;
        while (r + 1 < n_snp && std::abs(pos[r + 1] - pos[i]) <= wind) ++r;
        lo[i] = l;
        hi[i] = r;
    }
}

namespace {

void set_err(char* err, size_t errlen, const char* msg) {
    if (err && errlen) {
        strncpy(err, msg, errlen - 1);
        err[errlen - 1] = 0;
    }
}

template <class F>
int guarded(char* err, size_t errlen, F&& f) {
    try {
        f();
        return NLDSC_OK;
    } catch (const std::invalid_argument& e) {
        set_err(err, errlen, e.what());
        return NLDSC_ERR_INVALID;
    } catch (const std::exception& e) {
        set_err(err, errlen, e.what());
        return NLDSC_ERR_RUNTIME;
    } catch (...) {
        set_err(err, errlen, "unknown error");
        return NLDSC_ERR_RUNTIME;
    }
}

int usable_devices() {
    int n = 0;
    const cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) {
        cudaGetLastError();
        throw CudaError(std::string("no usable CUDA device (there is no CPU fallback): ") + cudaGetErrorString(e));
    }
    if (n <= 0) throw CudaError("no usable CUDA device (there is no CPU fallback)");
    return n;
}

void require_sm100(int device) {
    cudaDeviceProp p;
    CUDA_CHECK(cudaGetDeviceProperties(&p, device));
    if (p.major != 10) {
        char buf[160];
        snprintf(buf, sizeof(buf), "device %d is sm_%d%d; this library is built for sm_100a only", device, p.major,
                 p.minor);
        throw CudaError(buf);
    }
}

void validate_params(const nldsc_ld_params* p) {
    if (!p) throw InvalidArgument("params is null");
    if (p->n_snp <= 0) throw InvalidArgument("n_snp must be positive");
    if (p->n_org <= 0) throw InvalidArgument("n_org must be positive");
    if (!p->positions) throw InvalidArgument("positions is null");
    if (!(p->ld_wind >= 0.0)) throw InvalidArgument("ld_wind must be non-negative");
    for (int32_t i = 0; i < p->n_snp; ++i) {
        // negative ("unused") or unsorted positions are undefined behaviour in the reference
        if (!(p->positions[i] >= 0.0)) throw InvalidArgument("positions must be non-negative");
        if (i && p->positions[i] < p->positions[i - 1]) throw InvalidArgument("positions must be non-decreasing");
    }
}

template <class T>
T* dmalloc(size_t n) {
    void* p = nullptr;
    CUDA_CHECK(cudaMalloc(&p, std::max<size_t>(n, 1) * sizeof(T)));
    return static_cast<T*>(p);
}

int32_t floor64(int32_t x) { return x & ~63; }

struct SpanScope {
    nldsc_ld_session* s;
    cudaStream_t st;
    size_t idx;
    SpanScope(nldsc_ld_session* s_, cudaStream_t st_, int family) : s(s_), st(st_) {
        if (s->spans_used == s->spans.size()) {
            nldsc_ld_session::Span sp;
            CUDA_CHECK(cudaEventCreate(&sp.a));
            CUDA_CHECK(cudaEventCreate(&sp.b));
            sp.family = family;
            s->spans.push_back(sp);
        }
        idx = s->spans_used++;
        s->spans[idx].family = family;
        CUDA_CHECK(cudaEventRecord(s->spans[idx].a, st));
    }
    ~SpanScope() { cudaEventRecord(s->spans[idx].b, st); }
};

size_t env_size(const char* name, size_t dflt) {
    const char* v = getenv(name);
    if (!v || !*v) return dflt;
    char* end = nullptr;
    const double d = strtod(v, &end);
    if (end == v || !(d > 0)) return dflt;
    return static_cast<size_t>(d);
}

}  // namespace

// ---- session ------------------------------------------------------------------------------

static void session_free(nldsc_ld_session* s) {
    if (!s) return;
    cudaSetDevice(s->device);
    cudaFree(s->d_packed); cudaFree(s->d_counts); cudaFree(s->d_coef); cudaFree(s->d_flags);
    cudaFree(s->d_maf); cudaFree(s->d_rstd); cudaFree(s->d_win_lo); cudaFree(s->d_win_hi);
    cudaFree(s->d_planes); cudaFree(s->d_tiles); cudaFree(s->d_tile_begin);
    cudaFree(s->part.add); cudaFree(s->part.dom); cudaFree(s->part.wsa); cudaFree(s->part.wsd);
    cudaFree(s->part.wsde);
    cudaFree(s->d_l2); cudaFree(s->d_l2d); cudaFree(s->d_wsa); cudaFree(s->d_wsd); cudaFree(s->d_wsde);
    if (s->maps) band_maps_destroy(s->maps);
    for (auto& sp : s->spans) { cudaEventDestroy(sp.a); cudaEventDestroy(sp.b); }
    for (auto e : s->ev_uploaded) cudaEventDestroy(e);
    if (s->ev_run0) cudaEventDestroy(s->ev_run0);
    if (s->ev_run1) cudaEventDestroy(s->ev_run1);
    if (s->own_stream) cudaStreamDestroy(s->own_stream);
    if (s->copy_stream) cudaStreamDestroy(s->copy_stream);
    delete s;
}

static nldsc_ld_session* session_build(int device, const nldsc_ld_params* p, int32_t row_begin, int32_t row_end,
                                       const std::vector<int32_t>& lo, const std::vector<int32_t>& hi) {
    if (row_begin < 0 || row_end > p->n_snp || row_begin >= row_end)
        throw InvalidArgument("shard row range is empty or outside [0, n_snp)");
    const int ndev = usable_devices();
    if (device < 0 || device >= ndev) throw CudaError("CUDA device index out of range");
    CUDA_CHECK(cudaSetDevice(device));
    require_sm100(device);

    std::unique_ptr<nldsc_ld_session, void (*)(nldsc_ld_session*)> sp(new nldsc_ld_session, session_free);
    nldsc_ld_session* s = sp.get();
    s->device = device;
    CUDA_CHECK(cudaDeviceGetAttribute(&s->sm_count, cudaDevAttrMultiProcessorCount, device));
    s->n_org = p->n_org;
    s->n_snp = p->n_snp;
    s->row_begin = row_begin;
    s->row_end = row_end;
    s->maf_thr = p->maf;
    s->std_thr = p->std_thr;
    s->rsq_thr = p->rsq_thr;
    s->bps = (static_cast<int64_t>(p->n_org) + 3) / 4;
    s->npad = static_cast<int32_t>(((static_cast<int64_t>(p->n_org) + kKAlign - 1) / kKAlign) * kKAlign);
    const char* band = getenv("NLDSC_BAND");
    s->use_simt = band && strcmp(band, "simt") == 0;

    // ---- rows this shard needs: owned range widened by the window halo, 64-aligned on the left
    s->load_begin = floor64(lo[row_begin]);
    s->load_end = hi[row_end - 1] + 1;
    s->n_load = s->load_end - s->load_begin;
    s->own_begin = row_begin - s->load_begin;
    s->own_end = row_end - s->load_begin;
    const int32_t n_load = s->n_load;
    const int32_t n_load_pad = (n_load + kBlk - 1) / kBlk * kBlk;

    std::vector<int32_t> wlo(n_load), whi(n_load);
    for (int32_t i = 0; i < n_load; ++i) {
        wlo[i] = std::max(lo[i + s->load_begin] - s->load_begin, 0);
        whi[i] = std::min(hi[i + s->load_begin] - s->load_begin, n_load - 1);
    }
    for (int32_t i = s->own_begin; i < s->own_end; ++i) s->total_pairs += whi[i] - wlo[i];

    // ---- owned i-blocks and the j-block range each one needs
    const int32_t ib0 = s->own_begin / kBlk, ib1 = (s->own_end - 1) / kBlk + 1;
    const int32_t n_ib = ib1 - ib0;
    std::vector<int32_t> jb_lo(n_ib), jb_hi(n_ib);
    int32_t max_span = 0;
    for (int32_t b = ib0; b < ib1; ++b) {
        const int32_t r0 = std::max(b * kBlk, s->own_begin), r1 = std::min(b * kBlk + kBlk, s->own_end) - 1;
        jb_lo[b - ib0] = wlo[r0] / kBlk;
        jb_hi[b - ib0] = whi[r1] / kBlk;
        max_span = std::max(max_span, (jb_hi[b - ib0] - jb_lo[b - ib0] + 1) * kBlk);
    }

    // ---- fixed-size device buffers
    s->d_packed = dmalloc<uint8_t>(static_cast<size_t>(n_load) * s->bps + 16);
    s->d_counts = dmalloc<int32_t>(static_cast<size_t>(n_load_pad) * 3);
    s->d_coef = dmalloc<SnpCoef>(n_load_pad);
    s->d_flags = dmalloc<uint8_t>(n_load_pad);
    s->d_maf = dmalloc<double>(n_load_pad);
    s->d_rstd = dmalloc<double>(n_load_pad);
    s->d_win_lo = dmalloc<int32_t>(n_load_pad);
    s->d_win_hi = dmalloc<int32_t>(n_load_pad);
    CUDA_CHECK(cudaMemset(s->d_flags, 0, n_load_pad));
    CUDA_CHECK(cudaMemset(s->d_coef, 0, sizeof(SnpCoef) * n_load_pad));
    CUDA_CHECK(cudaMemset(s->d_packed + static_cast<size_t>(n_load) * s->bps, 0, 16));
    const int32_t n_own = s->own_end - s->own_begin;
    s->d_l2 = dmalloc<double>(n_own);
    s->d_l2d = dmalloc<double>(n_own);
    s->d_wsa = dmalloc<int32_t>(n_own);
    s->d_wsd = dmalloc<int32_t>(n_own);
    s->d_wsde = dmalloc<int32_t>(n_own);
    {
        std::vector<int32_t> pad_lo(n_load_pad, 1), pad_hi(n_load_pad, 0);
        std::copy(wlo.begin(), wlo.end(), pad_lo.begin());
        std::copy(whi.begin(), whi.end(), pad_hi.begin());
        CUDA_CHECK(cudaMemcpy(s->d_win_lo, pad_lo.data(), sizeof(int32_t) * n_load_pad, cudaMemcpyHostToDevice));
        CUDA_CHECK(cudaMemcpy(s->d_win_hi, pad_hi.data(), sizeof(int32_t) * n_load_pad, cudaMemcpyHostToDevice));
    }

    // ---- plane ring: as many SNP rows as the memory budget allows (4 int8 planes each)
    size_t free_b = 0, total_b = 0;
    CUDA_CHECK(cudaMemGetInfo(&free_b, &total_b));
    const size_t row_bytes = static_cast<size_t>(kPlanes) * s->npad;
    size_t budget = static_cast<size_t>(0.85 * static_cast<double>(free_b));
    budget = std::min(budget, env_size("NLDSC_PLANE_BYTES", budget));
    int64_t cap = static_cast<int64_t>(budget / row_bytes) / kBlk * kBlk;
    cap = std::min<int64_t>(cap, n_load_pad);
    if (cap < max_span) {
        char buf[200];
        snprintf(buf, sizeof(buf),
                 "window too wide for device memory: one SNP block needs %d resident SNP rows of %zu bytes, budget %zu bytes",
                 max_span, row_bytes, budget);
        throw CudaError(buf);
    }
    s->plane_rows = cap;
    s->d_planes = dmalloc<int8_t>(static_cast<size_t>(cap) * row_bytes);

    // ---- chunks: greedy runs of i-blocks whose column span fits the ring
    const int32_t chunk_rows = static_cast<int32_t>(
        std::min<size_t>(env_size("NLDSC_CHUNK_ROWS", 32768), static_cast<size_t>(std::numeric_limits<int32_t>::max() / 2)));
    std::vector<Tile> tiles;
    std::vector<int32_t> tile_begin;
    int32_t unpacked_to = jb_lo[0] * kBlk;   // rows below the first needed column never enter the ring
    for (int32_t b = 0; b < n_ib;) {
        Chunk c{};
        c.iblock0 = ib0 + b;
        c.tile0 = static_cast<int64_t>(tiles.size());
        c.tb0 = static_cast<int32_t>(tile_begin.size());
        const int32_t col0 = jb_lo[b] * kBlk;
        int32_t e = b;
        while (e < n_ib) {
            const int32_t col1 = (jb_hi[e] + 1) * kBlk;
            if (col1 - col0 > cap) break;
            if (e > b && (e - b) * kBlk >= chunk_rows) break;
            ++e;
        }
        // e > b is guaranteed by cap >= max_span
        const int32_t col1 = (jb_hi[e - 1] + 1) * kBlk;
        c.n_iblocks = e - b;
        c.new_row0 = std::max(unpacked_to, col0);
        c.new_row1 = std::max(col1, c.new_row0);
        unpacked_to = c.new_row1;
        for (int32_t k = b; k < e; ++k) {
            tile_begin.push_back(static_cast<int32_t>(tiles.size() - c.tile0));
            for (int32_t jb = jb_lo[k]; jb <= jb_hi[k]; ++jb) {
                Tile t;
                t.i0 = (ib0 + k) * kBlk;
                t.j0 = jb * kBlk;
                t.a_row = static_cast<int32_t>(t.i0 % cap);
                t.b_row = static_cast<int32_t>(t.j0 % cap);
                tiles.push_back(t);
            }
        }
        tile_begin.push_back(static_cast<int32_t>(tiles.size() - c.tile0));
        c.n_tiles = static_cast<int32_t>(tiles.size() - c.tile0);
        s->max_chunk_tiles = std::max(s->max_chunk_tiles, c.n_tiles);
        s->chunks.push_back(c);
        b = e;
    }
    s->total_tiles = static_cast<int64_t>(tiles.size());
    s->d_tiles = dmalloc<Tile>(tiles.size());
    s->d_tile_begin = dmalloc<int32_t>(tile_begin.size());
    CUDA_CHECK(cudaMemcpy(s->d_tiles, tiles.data(), sizeof(Tile) * tiles.size(), cudaMemcpyHostToDevice));
    CUDA_CHECK(cudaMemcpy(s->d_tile_begin, tile_begin.data(), sizeof(int32_t) * tile_begin.size(),
                          cudaMemcpyHostToDevice));
    const size_t np = static_cast<size_t>(s->max_chunk_tiles) * kBlk;
    s->part.add = dmalloc<double>(np);
    s->part.dom = dmalloc<double>(np);
    s->part.wsa = dmalloc<int32_t>(np);
    s->part.wsd = dmalloc<int32_t>(np);
    s->part.wsde = dmalloc<int32_t>(np);

    if (!s->use_simt) s->maps = band_maps_create(s->d_planes, s->plane_rows, s->npad);
    CUDA_CHECK(cudaStreamCreateWithFlags(&s->own_stream, cudaStreamNonBlocking));
    CUDA_CHECK(cudaStreamCreateWithFlags(&s->copy_stream, cudaStreamNonBlocking));
    s->ev_uploaded.resize(s->chunks.size());
    for (auto& e : s->ev_uploaded) CUDA_CHECK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    CUDA_CHECK(cudaEventCreate(&s->ev_run0));
    CUDA_CHECK(cudaEventCreate(&s->ev_run1));
    return sp.release();
}

// H2D of the packed rows, one copy per chunk so run() can start on chunk 0 while the rest
// of the shard is still crossing PCIe.  host_rows points at row load_begin.
static void session_upload(nldsc_ld_session* s, const uint8_t* host_rows, cudaStream_t st) {
    CUDA_CHECK(cudaSetDevice(s->device));
    cudaStream_t cs = st ? st : s->copy_stream;
    for (size_t c = 0; c < s->chunks.size(); ++c) {
        const Chunk& ch = s->chunks[c];
        const int32_t r0 = std::min(ch.new_row0, s->n_load), r1 = std::min(ch.new_row1, s->n_load);
        if (r1 > r0) {
            const size_t off = static_cast<size_t>(r0) * s->bps, len = static_cast<size_t>(r1 - r0) * s->bps;
            CUDA_CHECK(cudaMemcpyAsync(s->d_packed + off, host_rows + off, len, cudaMemcpyHostToDevice, cs));
        }
        CUDA_CHECK(cudaEventRecord(s->ev_uploaded[c], cs));
    }
    s->uploaded = true;
}

static void unpack_rows(nldsc_ld_session* s, int32_t r0, int32_t r1, cudaStream_t st) {
    // rows [r0, r1) -> ring slots (row % plane_rows); rows at or past n_load are zero padding
    const int64_t plane_stride = s->plane_rows * static_cast<int64_t>(s->npad);
    while (r0 < r1) {
        const int32_t slot = static_cast<int32_t>(r0 % s->plane_rows);
        int32_t n = static_cast<int32_t>(std::min<int64_t>(r1 - r0, s->plane_rows - slot));
        if (r0 < s->n_load) {
            n = std::min(n, s->n_load - r0);
            launch_unpack(s->d_packed, s->bps, r0, n, s->n_org, s->npad, s->d_planes, plane_stride, slot, st);
        } else {
            launch_zero_rows(s->d_planes, plane_stride, s->npad, slot, n, st);
        }
        ++s->launches;
        r0 += n;
    }
}

static void session_run(nldsc_ld_session* s, cudaStream_t st_in) {
    if (!s->uploaded) throw CudaError("session_run called before session_upload");
    CUDA_CHECK(cudaSetDevice(s->device));
    cudaStream_t st = st_in ? st_in : s->own_stream;
    s->spans_used = 0;
    s->launches = 0;
    CUDA_CHECK(cudaEventRecord(s->ev_run0, st));
    const double n = static_cast<double>(s->n_org);
    for (size_t ci = 0; ci < s->chunks.size(); ++ci) {
        const Chunk& c = s->chunks[ci];
        CUDA_CHECK(cudaStreamWaitEvent(st, s->ev_uploaded[ci], 0));
        const int32_t r0 = std::min(c.new_row0, s->n_load), r1 = std::min(c.new_row1, s->n_load);
        if (r1 > r0) {
            {
                SpanScope sc(s, st, kFamCount);
                launch_count(s->d_packed, s->bps, r0, r1 - r0, s->n_org, s->d_counts, st);
            }
            {
                SpanScope sc(s, st, kFamStats);
                launch_stats(s->d_counts, r0, r1 - r0, s->n_org, s->maf_thr, s->std_thr, s->d_coef, s->d_flags,
                             s->d_maf, s->d_rstd, st);
            }
            s->launches += 2;
        }
        {
            SpanScope sc(s, st, kFamUnpack);
            unpack_rows(s, c.new_row0, c.new_row1, st);
        }
        BandArgs a{};
        a.tiles = s->d_tiles + c.tile0;
        a.n_tiles = c.n_tiles;
        a.k_steps = s->npad / kKAlign;
        a.coef = s->d_coef;
        a.flags = s->d_flags;
        a.win_lo = s->d_win_lo;
        a.win_hi = s->d_win_hi;
        a.n_load = s->n_load;
        a.n = n;
        a.inv_n = 1.0 / n;
        a.adj_c = (n - 1.0) / (n - 2.0);
        a.rsq_thr = s->rsq_thr;
        a.part = s->part;
        a.dump = nullptr;
        {
            SpanScope sc(s, st, kFamBand);
            if (s->use_simt)
                launch_band_simt(s->d_planes, s->plane_rows * static_cast<int64_t>(s->npad), s->npad, a, st);
            else
                launch_band_tc(s->maps, a, s->sm_count, st);
            ++s->launches;
        }
        FinalizeArgs f{};
        f.tile_begin = s->d_tile_begin + c.tb0;
        f.iblock0 = c.iblock0;
        f.n_iblocks = c.n_iblocks;
        f.own_begin = s->own_begin;
        f.own_end = s->own_end;
        f.part = s->part;
        f.flags = s->d_flags;
        f.l2 = s->d_l2;
        f.l2d = s->d_l2d;
        f.wsa = s->d_wsa;
        f.wsd = s->d_wsd;
        f.wsde = s->d_wsde;
        {
            SpanScope sc(s, st, kFamFinalize);
            launch_finalize(f, st);
            ++s->launches;
        }
    }
    CUDA_CHECK(cudaEventRecord(s->ev_run1, st));
    s->ran = true;
}

static void session_download(nldsc_ld_session* s, nldsc_ld_result* out, cudaStream_t st_in) {
    if (!s->ran) throw CudaError("session_download called before session_run");
    if (!out || !out->l2 || !out->l2d || !out->maf || !out->residuals_std || !out->l2_ws || !out->l2d_ws ||
        !out->l2d_wse)
        throw InvalidArgument("result arrays must all be non-null");
    CUDA_CHECK(cudaSetDevice(s->device));
    cudaStream_t st = st_in ? st_in : s->own_stream;
    const size_t n = static_cast<size_t>(s->own_end - s->own_begin);
    CUDA_CHECK(cudaMemcpyAsync(out->l2, s->d_l2, n * sizeof(double), cudaMemcpyDeviceToHost, st));
    CUDA_CHECK(cudaMemcpyAsync(out->l2d, s->d_l2d, n * sizeof(double), cudaMemcpyDeviceToHost, st));
    CUDA_CHECK(cudaMemcpyAsync(out->maf, s->d_maf + s->own_begin, n * sizeof(double), cudaMemcpyDeviceToHost, st));
    CUDA_CHECK(cudaMemcpyAsync(out->residuals_std, s->d_rstd + s->own_begin, n * sizeof(double),
                               cudaMemcpyDeviceToHost, st));
    CUDA_CHECK(cudaMemcpyAsync(out->l2_ws, s->d_wsa, n * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    CUDA_CHECK(cudaMemcpyAsync(out->l2d_ws, s->d_wsd, n * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    CUDA_CHECK(cudaMemcpyAsync(out->l2d_wse, s->d_wsde, n * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    CUDA_CHECK(cudaStreamSynchronize(st));
}

}  // namespace nldsc

// ---- multi-GPU driver ----------------------------------------------------------------------

namespace {

// Contiguous row ranges with (roughly) equal band area = sum of window sizes.
std::vector<int32_t> split_rows(const std::vector<int32_t>& lo, const std::vector<int32_t>& hi, int parts) {
    const int32_t n = static_cast<int32_t>(lo.size());
    std::vector<int32_t> cut(parts + 1, n);
    cut[0] = 0;
    double total = 0;
    for (int32_t i = 0; i < n; ++i) total += hi[i] - lo[i] + 1;
    double acc = 0;
    int k = 1;
    for (int32_t i = 0; i < n && k < parts; ++i) {
        acc += hi[i] - lo[i] + 1;
        while (k < parts && acc >= total * k / parts) {
            // cut on a 64-row boundary so block-aligned shards waste no partial i-blocks
            int32_t c = std::min(n, (i + 1 + kBlk / 2) / kBlk * kBlk);
            cut[k] = std::max(c, cut[k - 1]);
            ++k;
        }
    }
    return cut;
}

struct ShardJob {
    int device;
    int32_t row_begin, row_end;
    int rc = NLDSC_OK;
    std::string msg;
};

int gpus_from_env(int available) {
    const char* v = getenv("NLDSC_GPUS");
    int want = 1;
    if (v && *v) {
        if (strcmp(v, "all") == 0) want = available;
        else want = atoi(v);
    }
    if (want < 1) want = 1;
    return std::min(want, available);
}

int first_device_from_env() {
    const char* v = getenv("NLDSC_DEVICE");
    return (v && *v) ? atoi(v) : 0;
}

// `fetch(row0, row1)` returns a host pointer to packed rows [row0, row1) valid until the shard is done.
template <class Fetch>
void run_sharded(const nldsc_ld_params* p, nldsc_ld_result* out, Fetch&& fetch) {
    validate_params(p);
    if (!out || !out->l2 || !out->l2d || !out->maf || !out->residuals_std || !out->l2_ws || !out->l2d_ws ||
        !out->l2d_wse)
        throw InvalidArgument("result arrays must all be non-null");
    const int ndev = usable_devices();
    std::vector<int32_t> lo, hi;
    compute_windows(p->positions, p->n_snp, p->ld_wind, lo, hi);
    const int dev0 = first_device_from_env();
    int parts = gpus_from_env(std::max(1, ndev - dev0));
    parts = std::max(1, std::min<int>(parts, (p->n_snp + kBlk - 1) / kBlk));
    const std::vector<int32_t> cut = split_rows(lo, hi, parts);

    std::vector<ShardJob> jobs;
    for (int k = 0; k < parts; ++k)
        if (cut[k + 1] > cut[k]) jobs.push_back({dev0 + k, cut[k], cut[k + 1]});

    auto work = [&](ShardJob& j) {
        nldsc_ld_session* s = nullptr;
        try {
            s = session_build(j.device, p, j.row_begin, j.row_end, lo, hi);
            const uint8_t* host = fetch(s->load_begin, s->load_end);
            session_upload(s, host, nullptr);
            session_run(s, nullptr);
            nldsc_ld_result sub = *out;
            sub.l2 += j.row_begin; sub.l2d += j.row_begin; sub.maf += j.row_begin;
            sub.residuals_std += j.row_begin; sub.l2_ws += j.row_begin; sub.l2d_ws += j.row_begin;
            sub.l2d_wse += j.row_begin;
            session_download(s, &sub, nullptr);
        } catch (const std::invalid_argument& e) {
            j.rc = NLDSC_ERR_INVALID; j.msg = e.what();
        } catch (const std::exception& e) {
            j.rc = NLDSC_ERR_RUNTIME; j.msg = e.what();
        }
        session_free(s);
    };
    if (jobs.size() == 1) {
        work(jobs[0]);
    } else {
        std::vector<std::thread> th;
        for (auto& j : jobs) th.emplace_back(work, std::ref(j));
        for (auto& t : th) t.join();
    }
    for (auto& j : jobs) {
        if (j.rc == NLDSC_ERR_INVALID) throw InvalidArgument(j.msg);
        if (j.rc != NLDSC_OK) throw CudaError(j.msg);
    }
}

}  // namespace

// ---- extern "C" ----------------------------------------------------------------------------

extern "C" {

int nldsc_ld_device_count(int32_t* count, char* err, size_t errlen) {
    if (count) *count = 0;
    return guarded(err, errlen, [&] {
        const int n = usable_devices();
        if (count) *count = n;
    });
}

int nldsc_ld_plan_windows(const nldsc_ld_params* params, int32_t* lo, int32_t* hi, char* err, size_t errlen) {
    return guarded(err, errlen, [&] {
        validate_params(params);
        if (!lo || !hi) throw InvalidArgument("lo / hi is null");
        std::vector<int32_t> l, h;
        compute_windows(params->positions, params->n_snp, params->ld_wind, l, h);
        std::copy(l.begin(), l.end(), lo);
        std::copy(h.begin(), h.end(), hi);
    });
}

int nldsc_ld_plan_shards(const nldsc_ld_params* params, int32_t parts, int32_t* cuts, char* err, size_t errlen) {
    return guarded(err, errlen, [&] {
        validate_params(params);
        if (!cuts || parts < 1) throw InvalidArgument("cuts is null or parts < 1");
        std::vector<int32_t> l, h;
        compute_windows(params->positions, params->n_snp, params->ld_wind, l, h);
        const std::vector<int32_t> c = split_rows(l, h, parts);
        std::copy(c.begin(), c.end(), cuts);
    });
}

int nldsc_ld_calculate_packed(const uint8_t* packed, const nldsc_ld_params* params, nldsc_ld_result* out, char* err,
                              size_t errlen) {
    return guarded(err, errlen, [&] {
        if (!packed) throw InvalidArgument("packed is null");
        validate_params(params);
        const int64_t bps = (static_cast<int64_t>(params->n_org) + 3) / 4;
        run_sharded(params, out, [&](int32_t r0, int32_t) { return packed + static_cast<size_t>(r0) * bps; });
    });
}

int nldsc_ld_calculate(const nldsc_ld_params* params, nldsc_ld_result* out, char* err, size_t errlen) {
    return guarded(err, errlen, [&] {
        validate_params(params);
        if (!params->bedfile) throw InvalidArgument("bedfile is null");
        const int fd = open(params->bedfile, O_RDONLY);
        if (fd < 0) throw CudaError(std::string("cannot open ") + params->bedfile + ": " + strerror(errno));
        struct Closer { int fd; ~Closer() { close(fd); } } closer{fd};
        // PLINK magic, SNP-major mode (reference stream.h:88-102 -> std::invalid_argument)
        unsigned char magic[3] = {0, 0, 0};
        if (pread(fd, magic, 3, 0) != 3 || magic[0] != 0x6C || magic[1] != 0x1B || magic[2] != 0x01)
            throw InvalidArgument("Wrong magic number in .bed file (expected 6c 1b 01, SNP-major)");
        const int64_t bps = (static_cast<int64_t>(params->n_org) + 3) / 4;
        struct stat stt;
        if (fstat(fd, &stt) != 0) throw CudaError("fstat failed on .bed file");
        const int64_t want = 3 + bps * params->n_snp;
        if (stt.st_size != want) {
            char buf[200];
            snprintf(buf, sizeof(buf), ".bed file size %lld does not match 3 + n_snp * ceil(n_org / 4) = %lld",
                     static_cast<long long>(stt.st_size), static_cast<long long>(want));
            throw InvalidArgument(buf);
        }
        // one pinned staging buffer per shard thread, filled by pread
        std::vector<void*> pinned;
        std::mutex mu;
        struct Cleanup {
            std::vector<void*>& v;
            ~Cleanup() { for (void* q : v) cudaFreeHost(q); }
        } cleanup{pinned};
        run_sharded(params, out, [&](int32_t r0, int32_t r1) -> const uint8_t* {
            const size_t len = static_cast<size_t>(r1 - r0) * bps;
            void* buf = nullptr;
            CUDA_CHECK(cudaHostAlloc(&buf, std::max<size_t>(len, 1), cudaHostAllocDefault));
            {
                std::lock_guard<std::mutex> g(mu);
                pinned.push_back(buf);
            }
            size_t done = 0;
            while (done < len) {
                const ssize_t k = pread(fd, static_cast<char*>(buf) + done, std::min<size_t>(len - done, 1u << 30),
                                        3 + static_cast<off_t>(r0) * bps + static_cast<off_t>(done));
                if (k <= 0) throw CudaError("short read on .bed file");
                done += static_cast<size_t>(k);
            }
            return static_cast<const uint8_t*>(buf);
        });
    });
}

int nldsc_ld_session_create(nldsc_ld_session** out, int device, const nldsc_ld_params* params, int32_t row_begin,
                            int32_t row_end, char* err, size_t errlen) {
    if (out) *out = nullptr;
    return guarded(err, errlen, [&] {
        if (!out) throw InvalidArgument("session out pointer is null");
        validate_params(params);
        std::vector<int32_t> lo, hi;
        compute_windows(params->positions, params->n_snp, params->ld_wind, lo, hi);
        *out = session_build(device, params, row_begin, row_end, lo, hi);
    });
}

int nldsc_ld_session_upload(nldsc_ld_session* s, const uint8_t* packed_rows, void* stream, char* err, size_t errlen) {
    return guarded(err, errlen, [&] {
        if (!s || !packed_rows) throw InvalidArgument("null session or buffer");
        session_upload(s, packed_rows, static_cast<cudaStream_t>(stream));
    });
}

int nldsc_ld_session_run(nldsc_ld_session* s, void* stream, char* err, size_t errlen) {
    return guarded(err, errlen, [&] {
        if (!s) throw InvalidArgument("null session");
        session_run(s, static_cast<cudaStream_t>(stream));
    });
}

int nldsc_ld_session_download(nldsc_ld_session* s, nldsc_ld_result* out, void* stream, char* err, size_t errlen) {
    return guarded(err, errlen, [&] {
        if (!s) throw InvalidArgument("null session");
        session_download(s, out, static_cast<cudaStream_t>(stream));
    });
}

int nldsc_ld_session_sync(nldsc_ld_session* s, char* err, size_t errlen) {
    return guarded(err, errlen, [&] {
        if (!s) throw InvalidArgument("null session");
        CUDA_CHECK(cudaSetDevice(s->device));
        CUDA_CHECK(cudaStreamSynchronize(s->copy_stream));
        CUDA_CHECK(cudaStreamSynchronize(s->own_stream));
    });
}

int nldsc_ld_session_stats(nldsc_ld_session* s, nldsc_ld_stats* o) {
    if (!s || !o) return NLDSC_ERR_INVALID;
    memset(o, 0, sizeof(*o));
    o->pairs = s->total_pairs;
    o->tiles = s->total_tiles;
    o->chunks = static_cast<int32_t>(s->chunks.size());
    o->kernel_launches = s->launches;
    o->load_begin = s->load_begin;
    o->load_end = s->load_end;
    o->npad = s->npad;
    o->band_impl = s->use_simt ? 0 : 1;
    if (s->ran) {
        cudaSetDevice(s->device);
        if (cudaEventSynchronize(s->ev_run1) != cudaSuccess) return NLDSC_ERR_RUNTIME;
        float fam[kFamN] = {0, 0, 0, 0, 0};
        for (size_t i = 0; i < s->spans_used; ++i) {
            float ms = 0;
            if (cudaEventElapsedTime(&ms, s->spans[i].a, s->spans[i].b) == cudaSuccess) fam[s->spans[i].family] += ms;
        }
        o->ms_count = fam[kFamCount];
        o->ms_stats = fam[kFamStats];
        o->ms_unpack = fam[kFamUnpack];
        o->ms_band = fam[kFamBand];
        o->ms_finalize = fam[kFamFinalize];
        cudaEventElapsedTime(&o->ms_total, s->ev_run0, s->ev_run1);
    }
    return NLDSC_OK;
}

void nldsc_ld_session_destroy(nldsc_ld_session* s) { session_free(s); }

// ---- parity hooks ----------------------------------------------------------------------------

int nldsc_ld_debug_unpack(const uint8_t* packed, int32_t n_snp, int32_t n_org, int8_t* planes_out, int32_t* counts_out,
                          int32_t* npad_out, char* err, size_t errlen) {
    return guarded(err, errlen, [&] {
        if (!packed || n_snp <= 0 || n_org <= 0) throw InvalidArgument("bad arguments");
        usable_devices();
        CUDA_CHECK(cudaSetDevice(first_device_from_env()));
        require_sm100(first_device_from_env());
        const int64_t bps = (static_cast<int64_t>(n_org) + 3) / 4;
        const int32_t npad = static_cast<int32_t>(((static_cast<int64_t>(n_org) + kKAlign - 1) / kKAlign) * kKAlign);
        if (npad_out) *npad_out = npad;
        if (!planes_out && !counts_out) return;
        uint8_t* d_packed = dmalloc<uint8_t>(static_cast<size_t>(n_snp) * bps + 16);
        int32_t* d_counts = dmalloc<int32_t>(static_cast<size_t>(n_snp) * 3);
        const size_t plane_bytes = static_cast<size_t>(n_snp) * npad;
        int8_t* d_planes = dmalloc<int8_t>(plane_bytes * kPlanes);
        struct Free { void* a; void* b; void* c; ~Free() { cudaFree(a); cudaFree(b); cudaFree(c); } } fr{d_packed, d_counts, d_planes};
        CUDA_CHECK(cudaMemset(d_packed + static_cast<size_t>(n_snp) * bps, 0, 16));
        CUDA_CHECK(cudaMemcpy(d_packed, packed, static_cast<size_t>(n_snp) * bps, cudaMemcpyHostToDevice));
        launch_count(d_packed, bps, 0, n_snp, n_org, d_counts, nullptr);
        launch_unpack(d_packed, bps, 0, n_snp, n_org, npad, d_planes, static_cast<int64_t>(plane_bytes), 0, nullptr);
        CUDA_CHECK(cudaDeviceSynchronize());
        if (counts_out)
            CUDA_CHECK(cudaMemcpy(counts_out, d_counts, sizeof(int32_t) * 3 * n_snp, cudaMemcpyDeviceToHost));
        if (planes_out) CUDA_CHECK(cudaMemcpy(planes_out, d_planes, plane_bytes * kPlanes, cudaMemcpyDeviceToHost));
    });
}

int nldsc_ld_debug_cross_counts(const uint8_t* packed, int32_t n_snp, int32_t n_org, int32_t i_block, int32_t j_block,
                                int32_t* acc_out, char* err, size_t errlen) {
    return guarded(err, errlen, [&] {
        if (!packed || !acc_out || n_snp <= 0 || n_org <= 0) throw InvalidArgument("bad arguments");
        const int32_t n_blocks = (n_snp + kBlk - 1) / kBlk;
        if (i_block < 0 || j_block < 0 || i_block >= n_blocks || j_block >= n_blocks)
            throw InvalidArgument("block index out of range");
        usable_devices();
        const int dev = first_device_from_env();
        CUDA_CHECK(cudaSetDevice(dev));
        require_sm100(dev);
        int sm_count = 0;
        CUDA_CHECK(cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev));
        const int64_t bps = (static_cast<int64_t>(n_org) + 3) / 4;
        const int32_t npad = static_cast<int32_t>(((static_cast<int64_t>(n_org) + kKAlign - 1) / kKAlign) * kKAlign);
        const int32_t rows = n_blocks * kBlk;
        uint8_t* d_packed = dmalloc<uint8_t>(static_cast<size_t>(n_snp) * bps + 16);
        const size_t plane_bytes = static_cast<size_t>(rows) * npad;
        int8_t* d_planes = dmalloc<int8_t>(plane_bytes * kPlanes);
        int32_t* d_dump = dmalloc<int32_t>(128 * 192);
        Tile* d_tile = dmalloc<Tile>(1);
        SnpCoef* d_coef = dmalloc<SnpCoef>(rows);
        uint8_t* d_flags = dmalloc<uint8_t>(rows);
        int32_t* d_w = dmalloc<int32_t>(static_cast<size_t>(rows) * 2);
        double* d_pd = dmalloc<double>(2 * kBlk);
        int32_t* d_pi = dmalloc<int32_t>(3 * kBlk);
        struct Free {
            std::vector<void*> v;
            ~Free() { for (void* q : v) cudaFree(q); }
        } fr{{d_packed, d_planes, d_dump, d_tile, d_coef, d_flags, d_w, d_pd, d_pi}};
        CUDA_CHECK(cudaMemset(d_packed + static_cast<size_t>(n_snp) * bps, 0, 16));
        CUDA_CHECK(cudaMemcpy(d_packed, packed, static_cast<size_t>(n_snp) * bps, cudaMemcpyHostToDevice));
        CUDA_CHECK(cudaMemset(d_planes, 0, plane_bytes * kPlanes));
        CUDA_CHECK(cudaMemset(d_coef, 0, sizeof(SnpCoef) * rows));
        CUDA_CHECK(cudaMemset(d_flags, 0, rows));
        CUDA_CHECK(cudaMemset(d_w, 0, sizeof(int32_t) * 2 * rows));
        launch_unpack(d_packed, bps, 0, n_snp, n_org, npad, d_planes, static_cast<int64_t>(plane_bytes), 0, nullptr);
        Tile t{i_block * kBlk, j_block * kBlk, i_block * kBlk, j_block * kBlk};
        CUDA_CHECK(cudaMemcpy(d_tile, &t, sizeof(t), cudaMemcpyHostToDevice));
        BandArgs a{};
        a.tiles = d_tile;
        a.n_tiles = 1;
        a.k_steps = npad / kKAlign;
        a.coef = d_coef;
        a.flags = d_flags;
        a.win_lo = d_w;
        a.win_hi = d_w + rows;
        a.n_load = rows;
        a.n = n_org;
        a.inv_n = 1.0 / n_org;
        a.adj_c = 1.0;
        a.rsq_thr = 0.0;
        a.part.add = d_pd;
        a.part.dom = d_pd + kBlk;
        a.part.wsa = d_pi;
        a.part.wsd = d_pi + kBlk;
        a.part.wsde = d_pi + 2 * kBlk;
        a.dump = d_dump;
        const char* band = getenv("NLDSC_BAND");
        if (band && strcmp(band, "simt") == 0) {
            launch_band_simt(d_planes, static_cast<int64_t>(plane_bytes), npad, a, nullptr);
        } else {
            BandMaps* maps = band_maps_create(d_planes, rows, npad);
            struct MapFree { BandMaps* m; ~MapFree() { band_maps_destroy(m); } } mf{maps};
            launch_band_tc(maps, a, sm_count, nullptr);
        }
        CUDA_CHECK(cudaDeviceSynchronize());
        CUDA_CHECK(cudaMemcpy(acc_out, d_dump, sizeof(int32_t) * 128 * 192, cudaMemcpyDeviceToHost));
    });
}

}  // extern "C"
