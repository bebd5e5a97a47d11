// fp32 CUDA-core GEMM (GNNX_PREC_FP32): the parity tier for the Linear layers (rel 1e-5 against the
// fp32 oracle) and the fallback for shapes the tcgen05 path does not take (tiny / unaligned).
// 128x128x8 tiles, 256 threads, 8x8 register micro-tile, register-staged double buffering,
// deterministic split-K (partials + ordered reduce) for the weight-gradient shapes (small M*N, huge K).
#include <mutex>

#include "common.cuh"
#include <atomic>
#include <cstdlib>

namespace gnnx {

constexpr int BM = 128, BN = 128, BK = 8, TM = 8, TN = 8;

// ReLU mask of the layer below: fp32 activations, or (GNNX_GEMM_MASK_BITS) one bit per element packed in uint32 words
__device__ __forceinline__ bool mask_keeps(const float* __restrict__ mask, long long ldm, int flags, long long m, int n) {
    if (flags & GNNX_GEMM_MASK_BITS)
        return (reinterpret_cast<const uint32_t*>(mask)[m * ldm + (n >> 5)] >> (n & 31)) & 1u;
    return mask[m * ldm + n] > 0.f;
}

// grouped product: rows are bucketed (bucket_ptr[n_groups+1], boundaries multiples of 128) and row block m0 uses
// the weight block of its bucket; rows past the last boundary use the last group
__device__ __forceinline__ int group_offset(const int* __restrict__ bucket_ptr, int n_groups, int m0) {
    int g = 0;
    if (bucket_ptr)
        for (int d = 1; d < n_groups; ++d) g += (m0 >= bucket_ptr[d]) ? 1 : 0;
    return g;
}

// element (m,k) of op(A): trans ? A[k*lda + m] : A[m*lda + k]
template <bool TA, bool TB>
__global__ void __launch_bounds__(256, 2)
k_gemm_f32(const float* __restrict__ A, long long lda, const float* __restrict__ B, long long ldb,
           float* __restrict__ C, long long ldc, int M, int N, int K, int k_per_split,
           const float* __restrict__ bias, int flags, float* __restrict__ partial,
           const float* __restrict__ mask, long long ldm, const int* __restrict__ bucket_ptr, int n_groups,
           long long group_stride, const int* __restrict__ split_table) {
    pdl_enter();
    __shared__ float As[2][BK][BM + 4];
    __shared__ float Bs[2][BK][BN + 4];
    const int tid = threadIdx.x;
    const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
    // grouped dW: the split's reduction range comes from the per-batch table {begin, end, group}
    const int kbeg = split_table ? split_table[3 * blockIdx.z] : blockIdx.z * k_per_split;
    const int kend = split_table ? split_table[3 * blockIdx.z + 1] : min(K, kbeg + k_per_split);
    if (split_table && kbeg >= kend) return;            // unused split: its partial is never read
    B += group_offset(bucket_ptr, n_groups, m0) * group_stride;
    const int tx = tid % 16, ty = tid / 16;     // 16x16 threads, each 8x8 outputs (strided by 16 -> conflict-free)

    float acc[TM][TN];
#pragma unroll
    for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;

    float ra[4], rb[4];
    auto load_tiles = [&](int k0) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            int el = tid + i * 256;
            int m, k;
            if (TA) { k = el / BM; m = el % BM; } else { m = el / BK; k = el % BK; }
            int gm = m0 + m, gk = k0 + k;
            ra[i] = (gm < M && gk < kend) ? (TA ? A[(long long)gk * lda + gm] : A[(long long)gm * lda + gk]) : 0.f;
            int n, kb;
            if (TB) { n = el / BK; kb = el % BK; } else { kb = el / BN; n = el % BN; }
            int gn = n0 + n, gkb = k0 + kb;
            rb[i] = (gn < N && gkb < kend) ? (TB ? B[(long long)gn * ldb + gkb] : B[(long long)gkb * ldb + gn]) : 0.f;
        }
    };
    auto store_tiles = [&](int buf) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            int el = tid + i * 256;
            int m, k;
            if (TA) { k = el / BM; m = el % BM; } else { m = el / BK; k = el % BK; }
            As[buf][k][m] = ra[i];
            int n, kb;
            if (TB) { n = el / BK; kb = el % BK; } else { kb = el / BN; n = el % BN; }
            Bs[buf][kb][n] = rb[i];
        }
    };

    int buf = 0;
    if (kbeg < kend) {
        load_tiles(kbeg);
        store_tiles(0);
    }
    __syncthreads();
    for (int k0 = kbeg; k0 < kend; k0 += BK) {
        bool has_next = k0 + BK < kend;
        if (has_next) load_tiles(k0 + BK);
#pragma unroll
        for (int kk = 0; kk < BK; ++kk) {
            float a[TM], b[TN];
#pragma unroll
            for (int i = 0; i < TM; ++i) a[i] = As[buf][kk][ty + 16 * i];
#pragma unroll
            for (int j = 0; j < TN; ++j) b[j] = Bs[buf][kk][tx + 16 * j];
#pragma unroll
            for (int i = 0; i < TM; ++i)
#pragma unroll
                for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
        }
        if (has_next) {
            store_tiles(buf ^ 1);
            __syncthreads();
            buf ^= 1;
        }
    }

    const bool split = partial != nullptr;
#pragma unroll
    for (int i = 0; i < TM; ++i) {
        int gm = m0 + ty + 16 * i;
        if (gm >= M) continue;
#pragma unroll
        for (int j = 0; j < TN; ++j) {
            int gn = n0 + tx + 16 * j;
            if (gn >= N) continue;
            float v = acc[i][j];
            if (split) {
                partial[((long long)blockIdx.z * M + gm) * N + gn] = v;
            } else {
                if (bias) v += bias[gn];
                if (flags & GNNX_GEMM_ACCUMULATE) v += C[(long long)gm * ldc + gn];
                if (mask && !mask_keeps(mask, ldm, flags, gm, gn)) v = 0.f;
                if (flags & GNNX_GEMM_RELU) v = fmaxf(v, 0.f);
                C[(long long)gm * ldc + gn] = v;
            }
        }
    }
}

// Small products (weight-space W_lin W_last, the 60-row bond tables, the MLP head): 32x32 tiles, K walked in one pass
// by every block -- ONE launch instead of split-K partials + reduce, which is what these latency-bound shapes need.
constexpr int SB = 32, SK = 64;
template <bool TA, bool TB>
__global__ void __launch_bounds__(256)
k_gemm_f32_small(const float* __restrict__ A, long long lda, const float* __restrict__ B, long long ldb,
                 float* __restrict__ C, long long ldc, int M, int N, int K, const float* __restrict__ bias, int flags,
                 const float* __restrict__ mask, long long ldm, const int* __restrict__ bucket_ptr, int n_groups,
                 long long group_stride) {
    pdl_enter();
    __shared__ float As[SK][SB + 1];
    __shared__ float Bs[SK][SB + 1];
    const int tid = threadIdx.x;
    const int m0 = blockIdx.y * SB, n0 = blockIdx.x * SB;
    B += group_offset(bucket_ptr, n_groups, m0) * group_stride;
    const int tx = tid & 15, ty = tid >> 4;          // 16 x 16 threads, 2 x 2 outputs each (strided by 16)
    constexpr int kPer = SB * SK / 256;              // elements of each operand tile per thread
    float ra[kPer], rb[kPer];
    auto load = [&](int k0) {                        // global -> registers: all loads of a tile in flight together
#pragma unroll
        for (int i = 0; i < kPer; ++i) {
            const int el = tid + i * 256;
            int m, k;
            if (TA) { k = el / SB; m = el % SB; } else { m = el / SK; k = el % SK; }
            const int gm = m0 + m, gk = k0 + k;
            ra[i] = (gm < M && gk < K) ? (TA ? __ldg(A + (long long)gk * lda + gm) : __ldg(A + (long long)gm * lda + gk)) : 0.f;
            int n, kb;
            if (TB) { n = el / SK; kb = el % SK; } else { kb = el / SB; n = el % SB; }
            const int gn = n0 + n, gkb = k0 + kb;
            rb[i] = (gn < N && gkb < K) ? (TB ? __ldg(B + (long long)gn * ldb + gkb) : __ldg(B + (long long)gkb * ldb + gn)) : 0.f;
        }
    };
    auto stash = [&]() {                             // registers -> shared
#pragma unroll
        for (int i = 0; i < kPer; ++i) {
            const int el = tid + i * 256;
            int m, k;
            if (TA) { k = el / SB; m = el % SB; } else { m = el / SK; k = el % SK; }
            As[k][m] = ra[i];
            int n, kb;
            if (TB) { n = el / SK; kb = el % SK; } else { kb = el / SB; n = el % SB; }
            Bs[kb][n] = rb[i];
        }
    };
    float acc[2][2] = {{0.f, 0.f}, {0.f, 0.f}};
    load(0);
    for (int k0 = 0; k0 < K; k0 += SK) {
        stash();
        __syncthreads();
        if (k0 + SK < K) load(k0 + SK);              // next tile's loads overlap this tile's FMAs
#pragma unroll
        for (int kk = 0; kk < SK; ++kk) {
            const float a0 = As[kk][ty], a1 = As[kk][ty + 16], b0 = Bs[kk][tx], b1 = Bs[kk][tx + 16];
            acc[0][0] = fmaf(a0, b0, acc[0][0]); acc[0][1] = fmaf(a0, b1, acc[0][1]);
            acc[1][0] = fmaf(a1, b0, acc[1][0]); acc[1][1] = fmaf(a1, b1, acc[1][1]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const int gm = m0 + ty + 16 * i;
        if (gm >= M) continue;
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            const int gn = n0 + tx + 16 * j;
            if (gn >= N) continue;
            float v = acc[i][j];
            if (bias) v += bias[gn];
            if (flags & GNNX_GEMM_ACCUMULATE) v += C[(long long)gm * ldc + gn];
            if (mask && !mask_keeps(mask, ldm, flags, gm, gn)) v = 0.f;
            if (flags & GNNX_GEMM_RELU) v = fmaxf(v, 0.f);
            C[(long long)gm * ldc + gn] = v;
        }
    }
}
static bool simt_small(int64_t M, int64_t N, int64_t K) {
    return K <= 1024 && (long long)ceil_div(M, BM) * ceil_div(N, BN) < num_sms() && ceil_div(M, SB) <= 65535;
}

__global__ void k_splitk_reduce(const float* __restrict__ partial, int splits, int M, int N, float* __restrict__ C,
                                long long ldc, const float* __restrict__ bias, int flags, const float* __restrict__ mask,
                                long long ldm) {
    pdl_enter();
    long long total = (long long)M * N;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        int m = (int)(i / N), n = (int)(i % N);
        float v = 0.f;
        int s = 0;
        for (; s + 8 <= splits; s += 8) {          // 8 independent loads in flight, summed in split order
            float t[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) t[j] = __ldcs(partial + (long long)(s + j) * total + i);
#pragma unroll
            for (int j = 0; j < 8; ++j) v += t[j];
        }
        for (; s < splits; ++s) v += __ldcs(partial + (long long)s * total + i);
        if (bias) v += bias[n];
        if (flags & GNNX_GEMM_ACCUMULATE) v += C[(long long)m * ldc + n];
        if (mask && !mask_keeps(mask, ldm, flags, m, n)) v = 0.f;
        if (flags & GNNX_GEMM_RELU) v = fmaxf(v, 0.f);
        C[(long long)m * ldc + n] = v;
    }
}

// Ordered (deterministic) sum of `splits` partial [M,N] products + bias / accumulate / ReLU; shared with gemm_tc.cu.
int splitk_reduce_launch(const float* partial, int splits, int M, int N, float* C, long long ldc, const float* bias,
                         int flags, const float* mask, long long ldm, cudaStream_t st) {
    long long total = (long long)M * N;
    int blocks = (int)((total + 255) / 256);
    if (blocks > num_sms() * 8) blocks = num_sms() * 8;
    launch(k_splitk_reduce, blocks, 256, 0, st, partial, splits, M, N, C, ldc, bias, flags, mask, ldm);
    return check_launch("splitk_reduce");
}

int simt_splits(int64_t M, int64_t N, int64_t K) {
    if (simt_small(M, N, K)) return 1;
    long long tiles = (long long)ceil_div(M, BM) * ceil_div(N, BN);
    if (tiles >= num_sms() || K < 128) return 1;
    long long want = (2ll * num_sms() + tiles - 1) / tiles;
    long long maxs = K >= 1024 ? K / 256 : K / 32;       // small weight-space products are latency-bound: split deeper
    long long s = want < maxs ? want : maxs;
    if (s > 128) s = 128;
    return (int)(s < 1 ? 1 : s);
}

size_t simt_workspace_bytes(int64_t M, int64_t N, int64_t K) {
    int s = simt_splits(M, N, K);
    return s > 1 ? (size_t)s * M * N * sizeof(float) : 0;
}

// bits[r, w] = OR_j (C[r, 32 w + j] > 0) << j : the packed ReLU mask a forward product leaves for its backward
__global__ void k_pack_relu_bits(const float* __restrict__ C, long long ldc, int M, int N, uint32_t* __restrict__ bits) {
    pdl_enter();
    const int words = (N + 31) / 32;
    const long long total = (long long)M * words;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const long long r = i / words;
        const int c0 = (int)(i % words) * 32;
        uint32_t w = 0;
        for (int j = 0; j < 32 && c0 + j < N; ++j) w |= (C[r * ldc + c0 + j] > 0.f ? 1u : 0u) << j;
        bits[i] = w;
    }
}
int pack_relu_bits_launch(const float* C, long long ldc, int M, int N, uint32_t* bits, cudaStream_t st) {
    long long total = (long long)M * ((N + 31) / 32);
    int blocks = (int)((total + 255) / 256);
    if (blocks > num_sms() * 8) blocks = num_sms() * 8;
    launch(k_pack_relu_bits, blocks, 256, 0, st, C, ldc, M, N, bits);
    return check_launch("pack_relu_bits");
}

int gemm_simt(const float* A, int64_t lda, int ta, const float* B, int64_t ldb, int tb, float* C, int64_t ldc,
              int64_t M, int64_t N, int64_t K, const float* bias, int flags, const float* mask, int64_t ldm,
              uint32_t* bits_out, void* ws, size_t ws_bytes, cudaStream_t st, const int* bucket_ptr, int n_groups) {
    if (M == 0 || N == 0) return GNNX_OK;
    // grouped: B stacks one weight block per group -- [G*N, K] rows (tb) or [G*K, N] rows
    const long long gstride = bucket_ptr ? (tb ? (long long)N * ldb : (long long)K * ldb) : 0;
    if (simt_small(M, N, K)) {
        dim3 grid(ceil_div(N, SB), ceil_div(M, SB));
#define LAUNCH_S(TA, TB) launch(k_gemm_f32_small<TA, TB>, grid, 256, 0, st, A, lda, B, ldb, C, ldc, (int)M, (int)N, (int)K, bias, flags, mask, ldm, bucket_ptr, n_groups, gstride)
        if (ta && tb) LAUNCH_S(true, true);
        else if (ta) LAUNCH_S(true, false);
        else if (tb) LAUNCH_S(false, true);
        else LAUNCH_S(false, false);
#undef LAUNCH_S
        int rc = check_launch("gemm_simt_small");
        if (rc == GNNX_OK && bits_out) rc = pack_relu_bits_launch(C, ldc, (int)M, (int)N, bits_out, st);
        return rc;
    }
    int splits = simt_splits(M, N, K);
    float* partial = nullptr;
    if (splits > 1) {
        GNNX_REQUIRE(ws && ws_bytes >= (size_t)splits * M * N * sizeof(float), "gemm: workspace too small for split-K");
        partial = (float*)ws;
    }
    int kps = ceil_div(ceil_div(K, splits), BK) * BK;
    if (kps == 0) kps = BK;
    splits = K > 0 ? ceil_div(K, kps) : 1;
    if (splits <= 1) { partial = nullptr; splits = 1; }
    dim3 grid(ceil_div(N, BN), ceil_div(M, BM), splits);
    GNNX_REQUIRE(grid.y <= 65535 * 1 || true, "gemm: M too large");
    if (grid.y > 65535) return set_error(GNNX_ERR_UNSUPPORTED, "gemm_simt: M=%lld exceeds grid.y limit", (long long)M);
#define LAUNCH(TA, TB) launch(k_gemm_f32<TA, TB>, grid, 256, 0, st, A, lda, B, ldb, C, ldc, (int)M, (int)N, (int)K, kps, bias, flags, partial, mask, ldm, bucket_ptr, n_groups, gstride, nullptr)
    if (ta && tb) LAUNCH(true, true);
    else if (ta) LAUNCH(true, false);
    else if (tb) LAUNCH(false, true);
    else LAUNCH(false, false);
#undef LAUNCH
    if (partial)
        launch(k_splitk_reduce, min(ceil_div(M * N, 256), num_sms() * 8), 256, 0, st, partial, splits, (int)M, (int)N, C, ldc, bias, flags, mask, ldm);
    int rc = check_launch("gemm_simt", partial ? 2 : 1);
    if (rc == GNNX_OK && bits_out) rc = pack_relu_bits_launch(C, ldc, (int)M, (int)N, bits_out, st);
    return rc;
}

// Grouped dW on the fp32 tier: partial[z][Mo,No] = A[rows_z, Mo]^T B[rows_z, No] over the table's row ranges.
int gemm_simt_red_grouped(const float* A, int64_t lda, const float* B, int64_t ldb, int64_t Mo, int64_t No, int64_t R,
                          const int* split_table, int S, float* partial, cudaStream_t st) {
    dim3 grid(ceil_div(No, BN), ceil_div(Mo, BM), S);
    launch(k_gemm_f32<true, false>, grid, 256, 0, st, A, lda, B, ldb, nullptr, 0, (int)Mo, (int)No, (int)R, 0, nullptr, 0, partial,
                                                  nullptr, 0, nullptr, 1, 0, split_table);
    return check_launch("gemm_simt_red(grouped)");
}

// ------------------------------------------------------------------ column sums
constexpr int kCsRows = 8;
__global__ void k_colsum_partial(const float* __restrict__ in, long long ld, int rows, int cols, int rows_per_chunk,
                                 float* __restrict__ partial) {
    pdl_enter();
    __shared__ float sh[kCsRows][33];
    int c = blockIdx.x * 32 + threadIdx.x;
    int r0 = blockIdx.y * rows_per_chunk, r1 = min(rows, r0 + rows_per_chunk);
    float s = 0.f;
    if (c < cols)
        for (int r = r0 + threadIdx.y; r < r1; r += kCsRows) s += __ldg(in + (long long)r * ld + c);
    sh[threadIdx.y][threadIdx.x] = s;
    __syncthreads();
    if (threadIdx.y == 0 && c < cols) {
        float t = 0.f;
#pragma unroll
        for (int i = 0; i < kCsRows; ++i) t += sh[i][threadIdx.x];
        partial[(long long)blockIdx.y * cols + c] = t;
    }
}
// float4 variant (cols % 4 == 0, 16-byte aligned rows): 64 column groups x 4 row lanes, 4 rows in flight per thread
__global__ void __launch_bounds__(256)
k_colsum_partial4(const float* __restrict__ in, long long ld, int rows, int cols, int rows_per_chunk, float* __restrict__ partial) {
    pdl_enter();
    __shared__ float sh[4][256 + 4];
    const int tx = threadIdx.x & 63, ty = threadIdx.x >> 6;
    const int c = (blockIdx.x * 64 + tx) * 4;
    const int r0 = blockIdx.y * rows_per_chunk, r1 = min(rows, r0 + rows_per_chunk);
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    if (c < cols) {
        int r = r0 + ty;
        for (; r + 12 < r1; r += 16) {
            float4 v[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) v[u] = ldg4_stream(in + (long long)(r + 4 * u) * ld + c);
#pragma unroll
            for (int u = 0; u < 4; ++u) { s.x += v[u].x; s.y += v[u].y; s.z += v[u].z; s.w += v[u].w; }
        }
        for (; r < r1; r += 4) {
            float4 v = ldg4_stream(in + (long long)r * ld + c);
            s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
        }
    }
    sh[ty][tx * 4 + 0] = s.x; sh[ty][tx * 4 + 1] = s.y; sh[ty][tx * 4 + 2] = s.z; sh[ty][tx * 4 + 3] = s.w;
    __syncthreads();
    const int cc = blockIdx.x * 256 + threadIdx.x;
    if (cc < cols) partial[(long long)blockIdx.y * cols + cc] = (sh[0][threadIdx.x] + sh[1][threadIdx.x]) + (sh[2][threadIdx.x] + sh[3][threadIdx.x]);
}

// one warp per column: lanes stride over the chunks, then a fixed shuffle tree (deterministic)
__global__ void k_colsum_final(const float* __restrict__ partial, int nchunks, int cols, float* __restrict__ out) {
    pdl_enter();
    const int lane = threadIdx.x & 31;
    const int c = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (c >= cols) return;
    float s = 0.f;
    for (int ch = lane; ch < nchunks; ch += 32) s += partial[(long long)ch * cols + c];
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) s += __shfl_down_sync(0xffffffffu, s, off);
    if (lane == 0) out[c] = s;
}
int colsum_final_launch(const float* partial, int nchunks, int cols, float* out, cudaStream_t st) {
    launch(k_colsum_final, ceil_div(cols, 8), 256, 0, st, partial, nchunks, cols, out);
    return check_launch("colsum_final");
}
// partial sums, a grid barrier and the final fold in ONE launch (the host launches it only when the whole grid is
// resident): same arithmetic as k_colsum_partial4 followed by k_colsum_final, so the result is bit-identical.
__global__ void __launch_bounds__(256)
k_colsum_coop4(const float* __restrict__ in, long long ld, int rows, int cols, int rows_per_chunk, float* __restrict__ partial,
               float* __restrict__ out, int* __restrict__ counters) {
    pdl_enter();
    __shared__ float sh[4][256 + 4];
    const int tx = threadIdx.x & 63, ty = threadIdx.x >> 6;
    const int c = (blockIdx.x * 64 + tx) * 4;
    const int r0 = blockIdx.y * rows_per_chunk, r1 = min(rows, r0 + rows_per_chunk);
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    if (c < cols) {
        int r = r0 + ty;
        for (; r + 12 < r1; r += 16) {
            float4 v[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) v[u] = ldg4_stream(in + (long long)(r + 4 * u) * ld + c);
#pragma unroll
            for (int u = 0; u < 4; ++u) { s.x += v[u].x; s.y += v[u].y; s.z += v[u].z; s.w += v[u].w; }
        }
        for (; r < r1; r += 4) {
            float4 v = ldg4_stream(in + (long long)r * ld + c);
            s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
        }
    }
    sh[ty][tx * 4 + 0] = s.x; sh[ty][tx * 4 + 1] = s.y; sh[ty][tx * 4 + 2] = s.z; sh[ty][tx * 4 + 3] = s.w;
    __syncthreads();
    const int cc = blockIdx.x * 256 + threadIdx.x;
    if (cc < cols) partial[(long long)blockIdx.y * cols + cc] = (sh[0][threadIdx.x] + sh[1][threadIdx.x]) + (sh[2][threadIdx.x] + sh[3][threadIdx.x]);
    const int nctas = (int)(gridDim.x * gridDim.y), nchunks = (int)gridDim.y;
    grid_barrier(counters, nctas);
    const int lane = threadIdx.x & 31;
    const int nwarps = nctas * 8;
    for (int col = ((int)(blockIdx.y * gridDim.x + blockIdx.x)) * 8 + (threadIdx.x >> 5); col < cols; col += nwarps) {
        float a = 0.f;
        for (int ch = lane; ch < nchunks; ch += 32) a += __ldcg(partial + (long long)ch * cols + col);
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) a += __shfl_down_sync(0xffffffffu, a, off);
        if (lane == 0) out[col] = a;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        if (atomicAdd(counters + 1, 1) == nctas - 1) { counters[0] = 0; counters[1] = 0; __threadfence(); }
    }
}

int colsum_chunks(int64_t rows, int64_t cols) {
    int slabs = ceil_div(cols, 256);
    int target = max(1, (num_sms() * 4) / slabs);
    int by_rows = max(1, (int)((rows + 63) / 64));
    return min(target, by_rows);
}

}  // namespace gnnx

using namespace gnnx;

extern "C" {

size_t gnnx_colsum_workspace_bytes(int64_t rows, int64_t cols) {
    return (size_t)colsum_chunks(rows, cols) * cols * sizeof(float);
}

int gnnx_colsum(const float* in, int64_t ld, int64_t rows, int64_t cols, float* out, void* ws, size_t ws_bytes,
                void* stream) {
    if (cols == 0) return GNNX_OK;
    GNNX_REQUIRE(ws_bytes >= gnnx_colsum_workspace_bytes(rows, cols), "colsum: workspace too small");
    cudaStream_t st = as_stream(stream);
    if (rows == 0) { GNNX_CUDA(cudaMemsetAsync(out, 0, cols * sizeof(float), st)); return GNNX_OK; }
    int nch = colsum_chunks(rows, cols);
    int rpc = ceil_div(rows, nch);
    if ((cols & 3) == 0 && (ld & 3) == 0 && ((uintptr_t)in & 15) == 0) {
        // one cooperative launch when the whole grid is resident (GNNX_COLSUM_COOP=0: two launches, same bits)
        static std::atomic<int> coop{0};                      // co-resident CTAs per SM; -1: unavailable
        int k = coop.load();
        if (k == 0) {
            int n = 0;
            if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, k_colsum_coop4, 256, 0) != cudaSuccess || n < 1) { cudaGetLastError(); n = -1; }
            coop.store(n); k = n;
        }
        const char* ce = getenv("GNNX_COLSUM_COOP");
        const long long ctas = (long long)ceil_div(cols, 256) * nch;
        int* ctr = (k > 0 && !(ce && ce[0] == '0') && ctas <= (long long)num_sms() * min(k, 4)) ? stream_ticket(st) : nullptr;
        if (ctr) {
            launch_coop(k_colsum_coop4, dim3(ceil_div(cols, 256), nch), 256, 0, st, in, ld, (int)rows, (int)cols, rpc, (float*)ws, out, ctr + 24);
            return check_launch("colsum", 1);
        }
        launch(k_colsum_partial4, dim3(ceil_div(cols, 256), nch), 256, 0, st, in, ld, (int)rows, (int)cols, rpc, (float*)ws);
    }
    else
        launch(k_colsum_partial, dim3(ceil_div(cols, 32), nch), dim3(32, kCsRows), 0, st, in, ld, (int)rows, (int)cols, rpc, (float*)ws);
    launch(k_colsum_final, ceil_div(cols, 8), 256, 0, st, (const float*)ws, nch, (int)cols, out);
    return check_launch("colsum", 2);
}

}  // extern "C"

// ------------------------------------------------------------------ column sums without any allocation
// gnnx_colsum_noalloc: the chunk partials of gnnx_colsum live in a file-scope device array instead of a caller
// workspace, for callers that cannot call cudaMalloc.  Every device that loads the library holds the array until
// exit.  In host code `g_colsum_scratch` is only a stand-in for the symbol: the device address is fetched once per
// device (that device current) with cudaGetSymbolAddress and handed to the kernels.
static __device__ __align__(256) float g_colsum_scratch[GNNX_COLSUM_NOALLOC_SCRATCH_BYTES / sizeof(float)];

namespace {
constexpr int kMaxDevices = 64;
std::mutex g_scratch_mu;
float* g_scratch_ptr[kMaxDevices] = {};

// 0 and *ptr set, or an error code (message in gnnx_last_error())
int colsum_scratch(float** ptr) {
    int dev = -1;
    GNNX_CUDA(cudaGetDevice(&dev));
    GNNX_REQUIRE(dev >= 0 && dev < kMaxDevices, "colsum_noalloc: device ordinal %d out of range", dev);
    std::lock_guard<std::mutex> g(g_scratch_mu);
    if (!g_scratch_ptr[dev]) {
        void* p = nullptr;
        GNNX_CUDA(cudaGetSymbolAddress(&p, g_colsum_scratch));
        g_scratch_ptr[dev] = static_cast<float*>(p);
    }
    *ptr = g_scratch_ptr[dev];
    return GNNX_OK;
}
}  // namespace

extern "C" {

size_t gnnx_colsum_noalloc_max_bytes(void) { return GNNX_COLSUM_NOALLOC_SCRATCH_BYTES; }

int gnnx_colsum_noalloc(const float* in, int64_t ld, int64_t rows, int64_t cols, float* out, void* stream) {
    if (cols == 0) return GNNX_OK;
    GNNX_REQUIRE(rows >= 0 && cols > 0 && ld >= cols && out != nullptr && (in != nullptr || rows == 0), "colsum_noalloc: bad arguments");
    GNNX_REQUIRE(rows <= INT32_MAX && cols <= INT32_MAX, "colsum_noalloc: rows / cols beyond int32");
    const size_t need = gnnx_colsum_workspace_bytes(rows, cols);
    if (need > (size_t)GNNX_COLSUM_NOALLOC_SCRATCH_BYTES)
        return set_error(GNNX_ERR_WORKSPACE, "colsum_noalloc: [%lld x %lld] needs %zu bytes of partials, the built-in scratch holds %zu "
                         "(use gnnx_colsum with a caller workspace)", (long long)rows, (long long)cols, need,
                         (size_t)GNNX_COLSUM_NOALLOC_SCRATCH_BYTES);
    cudaStream_t st = as_stream(stream);
    if (rows == 0) { GNNX_CUDA(cudaMemsetAsync(out, 0, cols * sizeof(float), st)); return GNNX_OK; }
    float* scratch = nullptr;
    int rc = colsum_scratch(&scratch);
    if (rc != GNNX_OK) return rc;
    // every partial [nch, cols] the final pass reads is written by the partial pass of THIS call: nothing in the shared
    // array has to start at zero, and what the previous call left behind is never read
    int nch = colsum_chunks(rows, cols);
    int rpc = ceil_div(rows, nch);
    if ((cols & 3) == 0 && (ld & 3) == 0 && ((uintptr_t)in & 15) == 0)
        launch(k_colsum_partial4, dim3(ceil_div(cols, 256), nch), 256, 0, st, in, ld, (int)rows, (int)cols, rpc, scratch);
    else
        launch(k_colsum_partial, dim3(ceil_div(cols, 32), nch), dim3(32, kCsRows), 0, st, in, ld, (int)rows, (int)cols, rpc, scratch);
    launch(k_colsum_final, ceil_div(cols, 8), 256, 0, st, (const float*)scratch, nch, (int)cols, out);
    return check_launch("colsum_noalloc", 2);
}

}  // extern "C"
