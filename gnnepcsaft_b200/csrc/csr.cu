// Batch collator: dst-sorted CSR + src-sorted view + degree histogram + packed-store collation.
// Integer-only, bit-exact with argsort(stable)/bincount.  HBM-bound integer
// work: histogram -> scan -> atomic placement -> per-row sort of edge ids (rows are tiny, and
// sorting unique ids makes the atomic placement order irrelevant => deterministic).
#include <stdarg.h>
#include <string.h>

#include <stdlib.h>
#include <stdio.h>
#include <string.h>

#include <vector>

#include "common.cuh"
#include <atomic>

namespace gnnx {

static thread_local char g_err[512] = "";
static unsigned long long g_launches = 0;
static unsigned long long g_tc_launches = 0;
void count_launches(int n) { __atomic_fetch_add(&g_launches, (unsigned long long)n, __ATOMIC_RELAXED); }
void count_tc_launches(int n) { __atomic_fetch_add(&g_tc_launches, (unsigned long long)n, __ATOMIC_RELAXED); }
int num_sms() {
    static int cached[64];                                // 0 = not queried yet; benign race (every thread stores the same value)
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) { cudaGetLastError(); return 148; }
    int v = __atomic_load_n(&cached[dev & 63], __ATOMIC_RELAXED);
    if (v > 0) return v;
    if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0) { cudaGetLastError(); return 148; }
    __atomic_store_n(&cached[dev & 63], v, __ATOMIC_RELAXED);
    return v;
}
// ---- per-kernel device timeline (diagnostic; single-threaded use)
struct TraceRec { const void* func; cudaEvent_t e0, e1; };
static bool g_trace = false;
static std::vector<TraceRec> g_trace_recs;
bool trace_on() { return g_trace; }
void trace_before(const void* func, cudaStream_t st) {
    TraceRec r; r.func = func;
    cudaEventCreate(&r.e0); cudaEventCreate(&r.e1);
    cudaEventRecord(r.e0, st);
    g_trace_recs.push_back(r);
}
void trace_after(cudaStream_t st) { cudaEventRecord(g_trace_recs.back().e1, st); }

// Gang launch of the kernels that contain a grid barrier (see launch_coop in common.cuh): -1 = read GNNX_GANG_LAUNCH once
static std::atomic<int> g_gang{-1};
bool gang_launch() {
    int v = g_gang.load(std::memory_order_relaxed);
    if (v < 0) { const char* e = getenv("GNNX_GANG_LAUNCH"); v = (e && e[0] == '1') ? 1 : 0; g_gang.store(v); }
    return v == 1;
}
void set_gang_launch(int on) { g_gang.store(on ? 1 : 0); }

bool pdl_enabled() {
    static int on = -1;
    if (on < 0) { const char* e = getenv("GNNX_PDL"); on = (e && e[0] == '0') ? 0 : 1; }
    return on == 1;
}
int set_error(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

// ------------------------------------------------------------------ histogram
template <typename IdxT>
__global__ void k_hist(const IdxT* __restrict__ src, const IdxT* __restrict__ dst, int E, int N,
                       int* __restrict__ cnt_dst, int* __restrict__ cnt_src, int* __restrict__ err,
                       int* __restrict__ pos_of_eid) {
    pdl_enter();
    for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < E; e += gridDim.x * blockDim.x) {
        long long s = (long long)src[e], d = (long long)dst[e];
        if (s < 0 || s >= N || d < 0 || d >= N) { *err = 1; if (pos_of_eid) pos_of_eid[e] = 0; continue; }
        atomicAdd(&cnt_dst[d + 1], 1);
        atomicAdd(&cnt_src[s + 1], 1);
    }
}

// ------------------------------------------------------------------ scan (3-phase, 2 arrays via blockIdx.y)
constexpr int kScanThreads = 256;
constexpr int kScanItems = 8;                     // 2048 elements per block
constexpr int kScanTile = kScanThreads * kScanItems;

__device__ __forceinline__ int block_scan_inclusive(int v, int* smem_warp) {
    // inclusive scan of one int per thread across the block (256 threads = 8 warps)
    int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        int t = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v += t;
    }
    if (lane == 31) smem_warp[w] = v;
    __syncthreads();
    if (w == 0) {
        int s = lane < (kScanThreads / 32) ? smem_warp[lane] : 0;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int t = __shfl_up_sync(0xffffffffu, s, o);
            if (lane >= o) s += t;
        }
        if (lane < (kScanThreads / 32)) smem_warp[lane] = s;
    }
    __syncthreads();
    if (w > 0) v += smem_warp[w - 1];
    __syncthreads();
    return v;
}

__global__ void k_scan_tiles(int* __restrict__ a0, int* __restrict__ a1, int n, int* __restrict__ sums, int ntiles) {
    pdl_enter();
    __shared__ int sw[32];
    int* a = blockIdx.y ? a1 : a0;
    int base = blockIdx.x * kScanTile + threadIdx.x * kScanItems;
    int v[kScanItems];
    int t = 0;
#pragma unroll
    for (int i = 0; i < kScanItems; ++i) {
        v[i] = (base + i < n) ? a[base + i] : 0;
        t += v[i];
        v[i] = t;
    }
    int incl = block_scan_inclusive(t, sw);
    int excl = incl - t;
#pragma unroll
    for (int i = 0; i < kScanItems; ++i)
        if (base + i < n) a[base + i] = v[i] + excl;
    if (threadIdx.x == kScanThreads - 1) sums[blockIdx.y * ntiles + blockIdx.x] = incl;
}

__global__ void k_scan_sums(int* __restrict__ sums, int ntiles) {
    pdl_enter();
    // one block per array; sequential over chunks of kScanThreads with a running carry
    __shared__ int sw[32];
    __shared__ int carry_s;
    int* s = sums + blockIdx.x * ntiles;
    if (threadIdx.x == 0) carry_s = 0;
    __syncthreads();
    for (int base = 0; base < ntiles; base += kScanThreads) {
        int i = base + threadIdx.x;
        int v = i < ntiles ? s[i] : 0;
        int incl = block_scan_inclusive(v, sw);
        int carry = carry_s;
        if (i < ntiles) s[i] = incl - v + carry;   // exclusive prefix of tile sums
        __syncthreads();
        if (threadIdx.x == kScanThreads - 1) carry_s = carry + incl;
        __syncthreads();
    }
}

__global__ void k_scan_add(int* __restrict__ a0, int* __restrict__ a1, int n, const int* __restrict__ sums, int ntiles) {
    pdl_enter();
    int* a = blockIdx.y ? a1 : a0;
    int add = sums[blockIdx.y * ntiles + blockIdx.x];
    int base = blockIdx.x * kScanTile + threadIdx.x * kScanItems;
#pragma unroll
    for (int i = 0; i < kScanItems; ++i)
        if (base + i < n) a[base + i] += add;
}

// ------------------------------------------------------------------ placement + per-row sort
template <typename IdxT>
__global__ void k_place(const IdxT* __restrict__ src, const IdxT* __restrict__ dst, int E, int N,
                        const int* __restrict__ rowptr, const int* __restrict__ colptr,
                        int* __restrict__ fill_dst, int* __restrict__ fill_src,
                        int* __restrict__ tmp_dst, int* __restrict__ tmp_src) {
    pdl_enter();
    for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < E; e += gridDim.x * blockDim.x) {
        long long s = (long long)src[e], d = (long long)dst[e];
        if (s < 0 || s >= N || d < 0 || d >= N) continue;      // (k_hist raised err_flag; the slot tails are zero-filled below)
        tmp_dst[rowptr[d] + atomicAdd(&fill_dst[d], 1)] = e;
        tmp_src[colptr[s] + atomicAdd(&fill_src[s], 1)] = e;
    }
}

__device__ void sort_ids(int* a, int n) {
    if (n <= 32) {                               // insertion sort: rows of molecular graphs have <= 6 entries
        for (int i = 1; i < n; ++i) {
            int v = a[i], j = i - 1;
            while (j >= 0 && a[j] > v) { a[j + 1] = a[j]; --j; }
            a[j + 1] = v;
        }
        return;
    }
    // heap sort for hub rows (any degree, in place)
    for (int start = n / 2 - 1; start >= 0; --start) {
        int root = start;
        for (;;) {
            int child = 2 * root + 1;
            if (child >= n) break;
            if (child + 1 < n && a[child] < a[child + 1]) ++child;
            if (a[root] >= a[child]) break;
            int t = a[root]; a[root] = a[child]; a[child] = t;
            root = child;
        }
    }
    for (int end = n - 1; end > 0; --end) {
        int t = a[0]; a[0] = a[end]; a[end] = t;
        int root = 0;
        for (;;) {
            int child = 2 * root + 1;
            if (child >= end) break;
            if (child + 1 < end && a[child] < a[child + 1]) ++child;
            if (a[root] >= a[child]) break;
            int t2 = a[root]; a[root] = a[child]; a[child] = t2;
            root = child;
        }
    }
}

template <typename IdxT>
__global__ void k_sort_rows_dst(const IdxT* __restrict__ src, int N, const int* __restrict__ rowptr,
                                int* __restrict__ tmp_dst, int* __restrict__ csr_src,
                                int* __restrict__ csr_eid, int* __restrict__ pos_of_eid, int E) {
    pdl_enter();
    // edges with an out-of-range endpoint were skipped (err_flag): the slots past rowptr[N] hold no edge.  They are
    // zero-filled so that no consumer ever indexes with uninitialised memory (a valid batch has rowptr[N] == E).
    for (int p = rowptr[N] + blockIdx.x * blockDim.x + threadIdx.x; p < E; p += gridDim.x * blockDim.x) {
        csr_eid[p] = 0;
        csr_src[p] = 0;
    }
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < N; i += gridDim.x * blockDim.x) {
        int b = rowptr[i], n = rowptr[i + 1] - b;
        sort_ids(tmp_dst + b, n);
        for (int k = 0; k < n; ++k) {
            int e = tmp_dst[b + k];
            csr_eid[b + k] = e;
            csr_src[b + k] = (int)src[e];
            pos_of_eid[e] = b + k;
        }
    }
}

template <typename IdxT>
__global__ void k_sort_rows_src(const IdxT* __restrict__ dst, int N, const int* __restrict__ colptr,
                                int* __restrict__ tmp_src, const int* __restrict__ pos_of_eid,
                                int* __restrict__ csc_pos, int* __restrict__ csc_dst, int E) {
    pdl_enter();
    for (int p = colptr[N] + blockIdx.x * blockDim.x + threadIdx.x; p < E; p += gridDim.x * blockDim.x) {
        csc_pos[p] = 0;
        csc_dst[p] = 0;
    }
    for (int j = blockIdx.x * blockDim.x + threadIdx.x; j < N; j += gridDim.x * blockDim.x) {
        int b = colptr[j], n = colptr[j + 1] - b;
        sort_ids(tmp_src + b, n);
        for (int k = 0; k < n; ++k) {
            int e = tmp_src[b + k];
            csc_pos[b + k] = pos_of_eid[e];
            csc_dst[b + k] = (int)dst[e];
        }
    }
}

static size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

template <typename IdxT>
static int csr_build_impl(const IdxT* src, const IdxT* dst, int64_t N64, int64_t E64, int32_t* rowptr,
                          int32_t* csr_src, int32_t* csr_eid, int32_t* pos_of_eid, int32_t* colptr,
                          int32_t* csc_pos, int32_t* csc_dst, int32_t* err_flag, void* ws, size_t ws_bytes,
                          void* stream_) {
    GNNX_REQUIRE(N64 >= 0 && E64 >= 0 && N64 < (1ll << 31) - 1 && E64 < (1ll << 31) - 1, "csr_build: N/E out of int32 range");
    GNNX_REQUIRE(rowptr && colptr && err_flag, "csr_build: null output");
    GNNX_REQUIRE(ws_bytes >= gnnx_csr_workspace_bytes(N64, E64), "csr_build: workspace too small");
    cudaStream_t st = as_stream(stream_);
    int N = (int)N64, E = (int)E64;
    int n1 = N + 1;
    int ntiles = ceil_div(n1, kScanTile);
    char* w = (char*)ws;
    int* fill_dst = (int*)w; w += align256(sizeof(int) * (size_t)n1);
    int* fill_src = (int*)w; w += align256(sizeof(int) * (size_t)n1);
    int* tmp_dst = (int*)w;  w += align256(sizeof(int) * (size_t)(E > 0 ? E : 1));
    int* tmp_src = (int*)w;  w += align256(sizeof(int) * (size_t)(E > 0 ? E : 1));
    int* sums = (int*)w;

    GNNX_CUDA(cudaMemsetAsync(rowptr, 0, sizeof(int) * (size_t)n1, st));
    GNNX_CUDA(cudaMemsetAsync(colptr, 0, sizeof(int) * (size_t)n1, st));
    GNNX_CUDA(cudaMemsetAsync(fill_dst, 0, sizeof(int) * (size_t)n1, st));
    GNNX_CUDA(cudaMemsetAsync(fill_src, 0, sizeof(int) * (size_t)n1, st));
    GNNX_CUDA(cudaMemsetAsync(err_flag, 0, sizeof(int), st));
    int eb = E > 0 ? min(ceil_div(E, 256), num_sms() * 16) : 0;
    if (E > 0) launch(k_hist<IdxT>, eb, 256, 0, st, src, dst, E, N, rowptr, colptr, err_flag, pos_of_eid);
    launch(k_scan_tiles, dim3(ntiles, 2), kScanThreads, 0, st, rowptr, colptr, n1, sums, ntiles);
    launch(k_scan_sums, 2, kScanThreads, 0, st, sums, ntiles);
    launch(k_scan_add, dim3(ntiles, 2), kScanThreads, 0, st, rowptr, colptr, n1, sums, ntiles);
    if (E > 0) {
        launch(k_place<IdxT>, eb, 256, 0, st, src, dst, E, N, rowptr, colptr, fill_dst, fill_src, tmp_dst, tmp_src);
        int nb = min(ceil_div(N, 128), num_sms() * 16);
        launch(k_sort_rows_dst<IdxT>, nb, 128, 0, st, src, N, rowptr, tmp_dst, csr_src, csr_eid, pos_of_eid, E);
        launch(k_sort_rows_src<IdxT>, nb, 128, 0, st, dst, N, colptr, tmp_src, pos_of_eid, csc_pos, csc_dst, E);
    }
    return check_launch("csr_build", E > 0 ? 7 : 3);
}

// ------------------------------------------------------------------ degree histogram / graph_ptr / misc
__global__ void k_degree_hist(const int* __restrict__ rowptr, int N, unsigned long long* __restrict__ hist,
                              int nbins, int* __restrict__ maxdeg) {
    pdl_enter();
    extern __shared__ int sh[];   // nbins counters + 1 max
    for (int i = threadIdx.x; i <= nbins; i += blockDim.x) sh[i] = 0;
    __syncthreads();
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < N; i += gridDim.x * blockDim.x) {
        int d = rowptr[i + 1] - rowptr[i];
        atomicAdd(&sh[min(d, nbins - 1)], 1);
        atomicMax(&sh[nbins], d);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < nbins; i += blockDim.x)
        if (sh[i]) atomicAdd(&hist[i], (unsigned long long)sh[i]);
    if (threadIdx.x == 0) atomicMax(maxdeg, sh[nbins]);
}

__global__ void k_graph_ptr(const int64_t* __restrict__ batch, int N, int G, int* __restrict__ gp) {
    pdl_enter();
    for (int n = blockIdx.x * blockDim.x + threadIdx.x; n <= N; n += gridDim.x * blockDim.x) {
        long long prev = n == 0 ? -1 : batch[n - 1];
        long long cur = n == N ? (long long)G : batch[n];
        if (cur > G) cur = G;
        for (long long g = prev + 1; g <= cur; ++g) gp[g] = n;   // also fills empty graphs
    }
}

// Column limits of a categorical feature matrix (vocabulary sizes); n == 0: no range check (limit 256).
struct NarrowLimits { int n; int lim[16]; };

__global__ void k_narrow(const int64_t* __restrict__ in, long long count, NarrowLimits lim, uint8_t* __restrict__ out,
                         int* __restrict__ err) {
    pdl_enter();
    bool bad = false;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < count; i += (long long)gridDim.x * blockDim.x) {
        const long long v = in[i];
        const int hi = lim.n ? lim.lim[i % lim.n] : 256;
        const bool ok = v >= 0 && v < hi;
        bad |= !ok;
        out[i] = ok ? (uint8_t)v : (uint8_t)0;            // an out-of-vocabulary id never reaches a table lookup
    }
    if (bad && err) *err = 1;
}

__global__ void k_edge_codes(const uint8_t* __restrict__ ea, const int* __restrict__ eid, int E,
                             uint8_t* __restrict__ codes, int* __restrict__ err) {
    pdl_enter();
    bool bad = false;
    for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < E; p += gridDim.x * blockDim.x) {
        const int e = eid[p];
        const bool eok = e >= 0 && e < E;
        const uint8_t* a = ea + 3ll * (eok ? e : 0);
        const bool ok = eok && a[0] < 5 && a[1] < 6 && a[2] < 2;      // OGB bond vocabularies (data/ogb_utils.py:24-33)
        bad |= !ok;
        codes[p] = ok ? (uint8_t)((a[0] * 6 + a[1]) * 2 + a[2]) : (uint8_t)0;
    }
    if (bad && err) *err = 1;
}

// one warp per selected molecule: copy its node / edge records to the batch, adding offsets
__global__ void k_collate_packed(const uint8_t* __restrict__ sx, const int16_t* __restrict__ ssrc,
                                 const int16_t* __restrict__ sdst, const uint8_t* __restrict__ sea,
                                 const int64_t* __restrict__ snp, const int64_t* __restrict__ sep,
                                 const int64_t* __restrict__ gids, int G, const int64_t* __restrict__ noff,
                                 const int64_t* __restrict__ eoff, uint8_t* __restrict__ ox,
                                 int* __restrict__ osrc, int* __restrict__ odst, uint8_t* __restrict__ oea,
                                 int* __restrict__ obatch) {
    pdl_enter();
    int lane = threadIdx.x & 31;
    int warps_per_block = blockDim.x >> 5;
    for (int g = blockIdx.x * warps_per_block + (threadIdx.x >> 5); g < G; g += gridDim.x * warps_per_block) {
        long long id = gids[g];
        long long nb = snp[id], nn = snp[id + 1] - nb;
        long long ebeg = sep[id], ne = sep[id + 1] - ebeg;
        long long no = noff[g], eo = eoff[g];
        for (long long t = lane; t < nn * 9; t += 32) ox[no * 9 + t] = sx[nb * 9 + t];
        for (long long t = lane; t < nn; t += 32) obatch[no + t] = g;
        for (long long t = lane; t < ne; t += 32) {
            osrc[eo + t] = (int)ssrc[ebeg + t] + (int)no;
            odst[eo + t] = (int)sdst[ebeg + t] + (int)no;
        }
        for (long long t = lane; t < ne * 3; t += 32) oea[eo * 3 + t] = sea[ebeg * 3 + t];
    }
}

}  // namespace gnnx

using namespace gnnx;

extern "C" {

int gnnx_abi_version(void) { return GNNX_ABI_VERSION; }
const char* gnnx_last_error(void) { return gnnx::g_err; }
int gnnx_set_gang_launch(int32_t on) { gnnx::set_gang_launch(on); return GNNX_OK; }
int64_t gnnx_launch_count(void) { return (int64_t)__atomic_load_n(&gnnx::g_launches, __ATOMIC_RELAXED); }
int64_t gnnx_tc_launch_count(void) { return (int64_t)__atomic_load_n(&gnnx::g_tc_launches, __ATOMIC_RELAXED); }
#ifndef GNNX_BUILD_DIGEST
#define GNNX_BUILD_DIGEST "unstamped"
#endif
const char* gnnx_build_digest(void) { return GNNX_BUILD_DIGEST; }

size_t gnnx_csr_workspace_bytes(int64_t N, int64_t E) {
    size_t n1 = (size_t)N + 1, e = (size_t)(E > 0 ? E : 1);
    size_t ntiles = (n1 + kScanTile - 1) / kScanTile;
    return 2 * align256(4 * n1) + 2 * align256(4 * e) + align256(4 * 2 * ntiles) + 256;
}

int gnnx_csr_build_i64(const int64_t* edge_index, int64_t N, int64_t E, int32_t* rowptr, int32_t* csr_src,
                       int32_t* csr_eid, int32_t* pos_of_eid, int32_t* colptr, int32_t* csc_pos,
                       int32_t* csc_dst, int32_t* err_flag, void* ws, size_t ws_bytes, void* stream) {
    return csr_build_impl<int64_t>(edge_index, edge_index + E, N, E, rowptr, csr_src, csr_eid, pos_of_eid,
                                   colptr, csc_pos, csc_dst, err_flag, ws, ws_bytes, stream);
}

int gnnx_csr_build_i32(const int32_t* src, const int32_t* dst, int64_t N, int64_t E, int32_t* rowptr,
                       int32_t* csr_src, int32_t* csr_eid, int32_t* pos_of_eid, int32_t* colptr,
                       int32_t* csc_pos, int32_t* csc_dst, int32_t* err_flag, void* ws, size_t ws_bytes,
                       void* stream) {
    return csr_build_impl<int32_t>(src, dst, N, E, rowptr, csr_src, csr_eid, pos_of_eid, colptr, csc_pos,
                                   csc_dst, err_flag, ws, ws_bytes, stream);
}

int gnnx_degree_hist(const int32_t* rowptr, int64_t N, int64_t* hist, int32_t nbins, int32_t* max_degree,
                     void* stream) {
    GNNX_REQUIRE(nbins >= 1 && nbins <= 4096, "degree_hist: nbins must be in [1,4096]");
    if (N == 0) return GNNX_OK;
    int nb = min(ceil_div(N, 256), num_sms() * 4);
    launch(k_degree_hist, nb, 256, sizeof(int) * (nbins + 1), as_stream(stream), 
        rowptr, (int)N, (unsigned long long*)hist, nbins, max_degree);
    return check_launch("degree_hist");
}

int gnnx_graph_ptr_from_batch(const int64_t* batch, int64_t N, int64_t G, int32_t* graph_ptr, void* stream) {
    GNNX_REQUIRE(N >= 0 && G >= 0, "graph_ptr: negative size");
    int nb = min(ceil_div(N + 1, 256), num_sms() * 8);
    launch(k_graph_ptr, nb, 256, 0, as_stream(stream), batch, (int)N, (int)G, graph_ptr);
    return check_launch("graph_ptr_from_batch");
}

int gnnx_narrow_i64_u8(const int64_t* in, int64_t rows, int32_t cols, const int32_t* limits_host, uint8_t* out,
                       int32_t* err_flag, void* stream) {
    GNNX_REQUIRE(rows >= 0 && cols >= 1, "narrow_i64_u8: bad shape");
    GNNX_REQUIRE(!limits_host || cols <= 16, "narrow_i64_u8: at most 16 checked columns");
    const long long count = (long long)rows * cols;
    if (count == 0) return GNNX_OK;
    NarrowLimits lim;
    lim.n = limits_host ? cols : 0;
    for (int k = 0; k < 16; ++k) {
        int v = (limits_host && k < cols) ? limits_host[k] : 256;
        lim.lim[k] = v < 256 ? v : 256;
    }
    int nb = min(ceil_div(count, 256), num_sms() * 16);
    launch(k_narrow, nb, 256, 0, as_stream(stream), in, count, lim, out, err_flag);
    return check_launch("narrow_i64_u8");
}

int gnnx_edge_codes(const uint8_t* ea, const int32_t* csr_eid, int64_t E, uint8_t* codes, int32_t* err_flag, void* stream) {
    if (E == 0) return GNNX_OK;
    int nb = min(ceil_div(E, 256), num_sms() * 16);
    launch(k_edge_codes, nb, 256, 0, as_stream(stream), ea, csr_eid, (int)E, codes, err_flag);
    return check_launch("edge_codes");
}

int gnnx_collate_packed(const uint8_t* sx, const int16_t* ssrc, const int16_t* sdst, const uint8_t* sea,
                        const int64_t* snp, const int64_t* sep, const int64_t* gids, int64_t G,
                        const int64_t* noff, const int64_t* eoff, uint8_t* ox, int32_t* osrc, int32_t* odst,
                        uint8_t* oea, int32_t* obatch, void* stream) {
    if (G == 0) return GNNX_OK;
    int nb = min(ceil_div(G, 8), num_sms() * 8);
    launch(k_collate_packed, nb, 256, 0, as_stream(stream), sx, ssrc, sdst, sea, snp, sep, gids, (int)G, noff, eoff,
                                                        ox, osrc, odst, oea, obatch);
    return check_launch("collate_packed");
}


int gnnx_trace_begin(void) {
    for (auto& r : g_trace_recs) { cudaEventDestroy(r.e0); cudaEventDestroy(r.e1); }
    g_trace_recs.clear();
    g_trace = true;
    return GNNX_OK;
}

// Writes "name\tkernel_us\tgap_before_us" lines (launch order; gap = end of the previous gnnx kernel -> start of this
// one on the device timeline) into buf, returns the number of bytes needed (call with cap = 0 to size the buffer).
int64_t gnnx_trace_end(char* buf, int64_t cap) {
    g_trace = false;
    cudaDeviceSynchronize();
    int64_t need = 0;
    for (size_t i = 0; i < g_trace_recs.size(); ++i) {
        const TraceRec& r = g_trace_recs[i];
        float us = 0.f, gap = 0.f;
        cudaEventElapsedTime(&us, r.e0, r.e1);
        if (i > 0) cudaEventElapsedTime(&gap, g_trace_recs[i - 1].e1, r.e0);
        const char* name = nullptr;
        if (cudaFuncGetName(&name, r.func) != cudaSuccess || !name) name = "?";
        char line[1024];
        int n = snprintf(line, sizeof line, "%s\t%.3f\t%.3f\n", name, us * 1000.f, gap * 1000.f);
        if (buf && need + n < cap) memcpy(buf + need, line, n);
        need += n;
    }
    if (buf && cap > 0) buf[need < cap ? need : cap - 1] = 0;
    cudaGetLastError();
    return need + 1;
}

}  // extern "C"
