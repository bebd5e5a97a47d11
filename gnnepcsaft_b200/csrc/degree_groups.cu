// Degree-grouped first post_nn Linear of PNAConv.
//
// The degree scalers of DegreeScalerAggregation are functions of the in-degree alone, and molecular graphs have a
// handful of distinct in-degrees (0..4, at most 6 with hypervalent atoms).  For every row of in-degree d
//     [x | agg | s_amp(d) agg | s_att(d) agg] W^T  =  [agg | x] W_eff(d)^T,
//     W_eff(d) = [W_id + s_amp(d) W_amp + s_att(d) W_att | W_x]
// so when the rows are bucketed by degree the K = 13F product (or its 3-wide scaler-folded form) becomes ONE grouped
// product with K = 5F per tower: a third of the flops, and no [N, 3H] intermediate.  This file holds the integer side
// (stable bucketing of the nodes by in-degree, buckets padded to `align` rows so a GEMM row block never straddles two
// groups; the row-split table of the grouped dW) and the small fp32 kernels around the grouped products (W_eff, the
// row gather that moves rows between node order and bucket order, the coefficient-weighted fold of the dW partials).
// Everything is deterministic and atomic-free on the fp32 side (the integer counters use shared-memory atomics, whose
// result does not depend on order).
#include "common.cuh"

namespace gnnx {

constexpr int kDgBlock = 256;

// ------------------------------------------------------------------ bucketing (integer)
// pass 1: cnt[d * nb + blk] = nodes of in-degree d in block blk
__global__ void __launch_bounds__(kDgBlock)
k_dg_count(const int* __restrict__ rowptr, int N, int D, int* __restrict__ cnt, int nb, int* __restrict__ err) {
    pdl_enter();
    __shared__ int sh[GNNX_MAX_GROUPS];
    if (threadIdx.x < GNNX_MAX_GROUPS) sh[threadIdx.x] = 0;
    __syncthreads();
    const int i = blockIdx.x * kDgBlock + threadIdx.x;
    if (i < N) {
        int d = rowptr[i + 1] - rowptr[i];
        if (d >= D) { d = D - 1; *err = 1; }
        atomicAdd(&sh[d], 1);
    }
    __syncthreads();
    if (threadIdx.x < D) cnt[threadIdx.x * nb + blockIdx.x] = sh[threadIdx.x];
}

// pass 2 (one block): per degree, exclusive scan of the block counts; bucket d starts at bucket_ptr[d] and is padded
// to a multiple of `align` rows.  cnt becomes the first bucket row of (degree, block).
__global__ void __launch_bounds__(1024)
k_dg_scan(int* __restrict__ cnt, int nb, int D, int align, int* __restrict__ bucket_ptr) {
    pdl_enter();
    __shared__ int sw[32];
    __shared__ int s_carry;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    int base = 0;
    if (threadIdx.x == 0) bucket_ptr[0] = 0;
    for (int d = 0; d < D; ++d) {
        if (threadIdx.x == 0) s_carry = 0;
        __syncthreads();
        for (int b0 = 0; b0 < nb; b0 += 1024) {
            const int i = b0 + threadIdx.x;
            const int v = i < nb ? cnt[d * nb + i] : 0;
            int x = v;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int y = __shfl_up_sync(0xffffffffu, x, o);
                if (lane >= o) x += y;
            }
            if (lane == 31) sw[w] = x;
            __syncthreads();
            if (w == 0) {
                int s = sw[lane];
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const int y = __shfl_up_sync(0xffffffffu, s, o);
                    if (lane >= o) s += y;
                }
                sw[lane] = s;
            }
            __syncthreads();
            const int incl = x + (w ? sw[w - 1] : 0);
            const int carry = s_carry;
            if (i < nb) cnt[d * nb + i] = base + carry + incl - v;
            __syncthreads();
            if (threadIdx.x == 1023) s_carry = carry + incl;
            __syncthreads();
        }
        base += (s_carry + align - 1) / align * align;
        if (threadIdx.x == 0) bucket_ptr[d + 1] = base;
        __syncthreads();
    }
}

// pass 3: stable placement -- nodes of one degree keep their node order inside the bucket
__global__ void __launch_bounds__(kDgBlock)
k_dg_place(const int* __restrict__ rowptr, int N, int D, const int* __restrict__ off, int nb, int* __restrict__ node_pos,
           int* __restrict__ row_node) {
    pdl_enter();
    __shared__ int wc[kDgBlock / 32][GNNX_MAX_GROUPS];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int i = blockIdx.x * kDgBlock + threadIdx.x;
    int d = -1;
    if (i < N) d = min(rowptr[i + 1] - rowptr[i], D - 1);
    const unsigned peers = __match_any_sync(0xffffffffu, d);
    const int rank = __popc(peers & ((1u << lane) - 1u));
    if (threadIdx.x < (kDgBlock / 32) * GNNX_MAX_GROUPS) (&wc[0][0])[threadIdx.x] = 0;
    __syncthreads();
    if (d >= 0 && rank == 0) wc[w][d] = __popc(peers);
    __syncthreads();
    if (d >= 0) {
        int pre = 0;
        for (int k = 0; k < w; ++k) pre += wc[k][d];
        const int p = off[d * nb + blockIdx.x] + pre + rank;
        node_pos[i] = p;
        row_node[p] = i;
    }
}

// Row-split table of the grouped dW: S splits, each {first row, end row, group}; every non-empty bucket gets at least
// one split, the rest are handed out in proportion to the bucket sizes (largest rows-per-split first); ranges are
// multiples of `unit` rows.  Unused splits are {0, 0, -1}.
__global__ void k_group_split_table(const int* __restrict__ bucket_ptr, int D, int unit, int S, int* __restrict__ table) {
    pdl_enter();
    if (threadIdx.x || blockIdx.x) return;
    int u[GNNX_MAX_GROUPS], s[GNNX_MAX_GROUPS];
    long long U = 0;
    int ne = 0;
    for (int d = 0; d < D; ++d) {
        u[d] = (bucket_ptr[d + 1] - bucket_ptr[d]) / unit;
        U += u[d];
        ne += u[d] > 0;
    }
    int used = 0;
    for (int d = 0; d < D; ++d) {
        s[d] = u[d] > 0 ? 1 + (int)((long long)u[d] * (S - ne) / U) : 0;
        if (s[d] > u[d]) s[d] = u[d];
        used += s[d];
    }
    while (used < S) {                                   // leftovers: the bucket whose splits are longest
        int best = -1;
        for (int d = 0; d < D; ++d)
            if (s[d] > 0 && s[d] < u[d] && (best < 0 || (long long)u[d] * s[best] > (long long)u[best] * s[d])) best = d;
        if (best < 0) break;
        ++s[best];
        ++used;
    }
    int z = 0;
    for (int d = 0; d < D; ++d) {
        if (!s[d]) continue;
        const int ups = (u[d] + s[d] - 1) / s[d];
        for (int j = 0; j < s[d]; ++j, ++z) {
            const int r0 = min(bucket_ptr[d + 1], bucket_ptr[d] + j * ups * unit);
            const int r1 = min(bucket_ptr[d + 1], r0 + ups * unit);
            table[3 * z] = r0; table[3 * z + 1] = r1; table[3 * z + 2] = d;
        }
    }
    for (; z < S; ++z) { table[3 * z] = 0; table[3 * z + 1] = 0; table[3 * z + 2] = -1; }
}

// ------------------------------------------------------------------ fp32 side
// weff[d] = [wz_id + s_amp(d) wz_amp + s_att(d) wz_att | wx]   ([Ho, K4 + H] per degree);  coef[d] = (1, s_amp, s_att)
__global__ void k_pna_degree_weights(const float* __restrict__ wz, const float* __restrict__ wx, int Ho, int K4, int H,
                                     float avg_deg_log, int D, float* __restrict__ weff, float* __restrict__ coef) {
    pdl_enter();
    const int Kt = K4 + H, k4 = Kt / 4;
    const long long total = (long long)D * Ho * k4;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int c = (int)(i % k4) * 4;
        const int o = (int)((i / k4) % Ho);
        const int d = (int)(i / ((long long)k4 * Ho));
        const float deg = (float)d;
        const float sa = logf(deg + 1.f) / avg_deg_log;
        const float st = avg_deg_log / logf(fmaxf(deg, 1.f) + 1.f);
        float4 v;
        if (c < K4) {
            const float4 a = ldg4(wz + (long long)o * K4 + c), b = ldg4(wz + (long long)(Ho + o) * K4 + c),
                         t = ldg4(wz + (long long)(2 * Ho + o) * K4 + c);
            v = make_float4(a.x + sa * b.x + st * t.x, a.y + sa * b.y + st * t.y, a.z + sa * b.z + st * t.z, a.w + sa * b.w + st * t.w);
        } else {
            v = ldg4(wx + (long long)o * H + (c - K4));
        }
        stg4(weff + ((long long)d * Ho + o) * Kt + c, v);
        if (o == 0 && c == 0) { coef[3 * d] = 1.f; coef[3 * d + 1] = sa; coef[3 * d + 2] = st; }
    }
}

// out[r, :] = idx[r] >= 0 ? src[idx[r], :] : 0, one warp per (row, 128-column chunk); optionally the packed
// (out > 0) bits of the rows written, in the layout GNNX_GEMM_MASK_BITS consumes
__global__ void __launch_bounds__(256)
k_gather_rows(const float* __restrict__ src, long long ld_src, const int* __restrict__ idx, int rows, int cols,
              float* __restrict__ out, long long ld_out, uint32_t* __restrict__ bits) {
    pdl_enter();
    const int lane = threadIdx.x & 31;
    const int chunks = (cols + 127) / 128, words = (cols + 31) / 32;
    const long long nwork = (long long)rows * chunks;
    const int wpb = blockDim.x >> 5;
    for (long long w = blockIdx.x * (long long)wpb + (threadIdx.x >> 5); w < nwork; w += (long long)gridDim.x * wpb) {
        const int r = (int)(w / chunks);
        const int c = (int)(w % chunks) * 128 + lane * 4;
        const bool act = c < cols;
        const int s = idx[r];
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (act && s >= 0) v = ldg4_stream(src + (long long)s * ld_src + c);
        if (act) stg4(out + (long long)r * ld_out + c, v);
        if (bits) {
            uint32_t m = 0;
            if (act) m = ((v.x > 0.f ? 1u : 0u) | (v.y > 0.f ? 2u : 0u) | (v.z > 0.f ? 4u : 0u) | (v.w > 0.f ? 8u : 0u)) << (4 * (lane & 7));
            m |= __shfl_xor_sync(0xffffffffu, m, 1);
            m |= __shfl_xor_sync(0xffffffffu, m, 2);
            m |= __shfl_xor_sync(0xffffffffu, m, 4);
            if (act && (lane & 7) == 0) bits[(long long)r * words + (c >> 5)] = m;
        }
    }
}

// out[j][i] = sum over the used splits z (in split order) of coef[group(z)][j] * partial[z][i]
__global__ void __launch_bounds__(256)
k_group_reduce(const float* __restrict__ partial, const int* __restrict__ table, int S, const float* __restrict__ coef,
               int J, long long total4, float* __restrict__ out) {
    pdl_enter();
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total4; i += (long long)gridDim.x * blockDim.x) {
        float4 acc[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[j] = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int z = 0; z < S; ++z) {
            const int g = table[3 * z + 2];
            if (g < 0 || table[3 * z] >= table[3 * z + 1]) continue;
            const float4 v = __ldcs(reinterpret_cast<const float4*>(partial) + (long long)z * total4 + i);
#pragma unroll
            for (int j = 0; j < 4; ++j)
                if (j < J) {
                    const float cf = coef[g * J + j];
                    acc[j].x = fmaf(cf, v.x, acc[j].x); acc[j].y = fmaf(cf, v.y, acc[j].y);
                    acc[j].z = fmaf(cf, v.z, acc[j].z); acc[j].w = fmaf(cf, v.w, acc[j].w);
                }
        }
#pragma unroll
        for (int j = 0; j < 4; ++j)
            if (j < J) reinterpret_cast<float4*>(out)[(long long)j * total4 + i] = acc[j];
    }
}

int group_reduce_launch(const float* partial, const int* split_table, int S, const float* coef, int J, long long total,
                        float* out, cudaStream_t st) {
    const long long total4 = total / 4;
    long long blocks = (total4 + 255) / 256;
    if (blocks > (long long)num_sms() * 8) blocks = (long long)num_sms() * 8;
    launch(k_group_reduce, (int)blocks, 256, 0, st, partial, split_table, S, coef, J, total4, out);
    return check_launch("group_reduce");
}

}  // namespace gnnx

using namespace gnnx;

extern "C" {

int64_t gnnx_degree_buckets_capacity(int64_t N, int32_t n_groups, int32_t align) {
    return (N + align - 1) / align * align + (int64_t)align * n_groups;
}

size_t gnnx_degree_buckets_workspace_bytes(int64_t N, int32_t n_groups) {
    return (size_t)n_groups * (size_t)ceil_div(N > 0 ? N : 1, kDgBlock) * sizeof(int);
}

int gnnx_degree_buckets(const int32_t* rowptr, int64_t N, int32_t n_groups, int32_t align, int32_t* node_pos,
                        int32_t* row_node, int64_t capacity, int32_t* bucket_ptr, int32_t* err_flag, void* ws,
                        size_t ws_bytes, void* stream) {
    GNNX_REQUIRE(n_groups >= 1 && n_groups <= GNNX_MAX_GROUPS, "degree_buckets: 1..%d groups", GNNX_MAX_GROUPS);
    GNNX_REQUIRE(align >= 1 && capacity >= gnnx_degree_buckets_capacity(N, n_groups, align), "degree_buckets: capacity too small");
    GNNX_REQUIRE(ws_bytes >= gnnx_degree_buckets_workspace_bytes(N, n_groups), "degree_buckets: workspace too small");
    cudaStream_t st = as_stream(stream);
    GNNX_CUDA(cudaMemsetAsync(row_node, 0xff, (size_t)capacity * sizeof(int32_t), st));
    GNNX_CUDA(cudaMemsetAsync(err_flag, 0, sizeof(int32_t), st));
    if (N == 0) { GNNX_CUDA(cudaMemsetAsync(bucket_ptr, 0, (size_t)(n_groups + 1) * sizeof(int32_t), st)); return GNNX_OK; }
    const int nb = ceil_div(N, kDgBlock);
    int* cnt = static_cast<int*>(ws);
    launch(k_dg_count, nb, kDgBlock, 0, st, rowptr, (int)N, n_groups, cnt, nb, err_flag);
    launch(k_dg_scan, 1, 1024, 0, st, cnt, nb, n_groups, align, bucket_ptr);
    launch(k_dg_place, nb, kDgBlock, 0, st, rowptr, (int)N, n_groups, cnt, nb, node_pos, row_node);
    return check_launch("degree_buckets", 3);
}

int gnnx_group_split_table(const int32_t* bucket_ptr, int32_t n_groups, int32_t unit, int32_t splits, int32_t* table,
                           void* stream) {
    GNNX_REQUIRE(n_groups >= 1 && n_groups <= GNNX_MAX_GROUPS && unit >= 1 && splits >= n_groups,
                 "group_split_table: need 1..%d groups and at least one split per group", GNNX_MAX_GROUPS);
    launch(k_group_split_table, 1, 32, 0, as_stream(stream), bucket_ptr, n_groups, unit, splits, table);
    return check_launch("group_split_table");
}

int gnnx_pna_degree_weights(const float* wz, const float* wx, int32_t h_out, int32_t k_agg, int32_t h_in, float avg_deg_log,
                            int32_t n_groups, float* weff, float* coef, void* stream) {
    GNNX_REQUIRE(k_agg % 4 == 0 && h_in % 4 == 0, "pna_degree_weights: widths must be multiples of 4");
    GNNX_REQUIRE(n_groups >= 1 && n_groups <= GNNX_MAX_GROUPS, "pna_degree_weights: 1..%d groups", GNNX_MAX_GROUPS);
    const long long total = (long long)n_groups * h_out * ((k_agg + h_in) / 4);
    long long blocks = (total + 255) / 256;
    if (blocks > (long long)num_sms() * 8) blocks = (long long)num_sms() * 8;
    launch(k_pna_degree_weights, (int)(blocks > 0 ? blocks : 1), 256, 0, as_stream(stream), wz, wx, h_out, k_agg, h_in, avg_deg_log, n_groups, weff, coef);
    return check_launch("pna_degree_weights");
}

int gnnx_gather_rows(const float* src, int64_t ld_src, const int32_t* idx, int64_t rows, int32_t cols, float* out,
                     int64_t ld_out, uint32_t* relu_bits_out, void* stream) {
    GNNX_REQUIRE(cols % 4 == 0 && ld_src % 4 == 0 && ld_out % 4 == 0, "gather_rows: widths / strides must be multiples of 4");
    GNNX_REQUIRE(((uintptr_t)src & 15) == 0 && ((uintptr_t)out & 15) == 0, "gather_rows: pointers must be 16-byte aligned");
    if (rows == 0 || cols == 0) return GNNX_OK;
    const long long nwork = rows * ((cols + 127) / 128);
    long long blocks = (nwork + 7) / 8;
    if (blocks > (long long)num_sms() * 32) blocks = (long long)num_sms() * 32;
    launch(k_gather_rows, (int)blocks, 256, 0, as_stream(stream), src, ld_src, idx, (int)rows, cols, out, ld_out, relu_bits_out);
    return check_launch("gather_rows");
}

}  // extern "C"
