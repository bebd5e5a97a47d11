// OGB AtomEncoder / BondEncoder embedding sums: out[r] = sum_k table[off_k + idx[r,k]].
// Forward: one warp per row, float4 over the hidden dim, table (<= 174 rows) stays in L1/L2.
// Backward: deterministic, atomic-free: every (column, feature) pair is owned by one thread that
// walks its row chunk in order and accumulates into a shared-memory table slab; chunk partials
// are then summed in chunk order.
#include "common.cuh"

namespace gnnx {

__global__ void k_embed_fwd(const uint8_t* __restrict__ idx, int rows, int K, const int* __restrict__ offsets,
                            const float* __restrict__ table, int H, float* __restrict__ out, NodeDropout drop) {
    pdl_enter();
    int lane = threadIdx.x & 31;
    int wpb = blockDim.x >> 5;
    for (int r = blockIdx.x * wpb + (threadIdx.x >> 5); r < rows; r += gridDim.x * wpb) {
        int my = lane < K ? offsets[lane] + (int)idx[(long long)r * K + lane] : 0;
        for (int c0 = 0; c0 < H; c0 += 128) {          // warp-uniform trip count: every lane takes part in the shuffles
            const int c = c0 + lane * 4;
            const bool active = c < H;
            float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
            for (int k = 0; k < K; ++k) {
                int tr = __shfl_sync(0xffffffffu, my, k);
                if (active) {
                    float4 v = ldg4(table + (long long)tr * H + c);
                    acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
                }
            }
            if (active) {
                const long long o = (long long)r * H + c;            // Dropout(p) on the first layer's input (models.py:209)
                acc.x = drop_apply(drop, o, acc.x); acc.y = drop_apply(drop, o + 1, acc.y);
                acc.z = drop_apply(drop, o + 2, acc.z); acc.w = drop_apply(drop, o + 3, acc.w);
                stg4(out + o, acc);
            }
        }
    }
}

constexpr int kEmbSlab = 128;  // columns per CTA in the backward: 4 per lane (ncu r2a: the 1-column version was issue-bound,
                               // 18 instructions per update at 0.5 TB/s; index / address work is now shared by 4 columns)
constexpr int kEmbFlight = 8;  // rows in flight per thread (float4 each): the row walk is a latency chain

// grid: (ceil(H/128), nchunks); block: (32, K).  smem: table_rows * 128 floats.
__global__ void k_embed_bwd_partial(const uint8_t* __restrict__ idx, int rows, int K,
                                    const int* __restrict__ offsets, int table_rows,
                                    const float* __restrict__ dout, int H, int rows_per_chunk,
                                    float* __restrict__ partial, NodeDropout drop) {
    pdl_enter();
    extern __shared__ float tab[];
    const int c = blockIdx.x * kEmbSlab + threadIdx.x * 4;
    const int k = threadIdx.y;
    const int tid = threadIdx.y * 32 + threadIdx.x, nthreads = blockDim.x * blockDim.y;
    for (int i = tid; i < table_rows * kEmbSlab; i += nthreads) tab[i] = 0.f;
    __syncthreads();
    const int r0 = blockIdx.y * rows_per_chunk;
    const int r1 = min(rows, r0 + rows_per_chunk);
    const int off = offsets[k];
    auto fetch = [&](int r) -> float4 {
        const long long o = (long long)r * H + c;
        float4 v = ldg4(dout + o);
        v.x = drop_apply(drop, o, v.x); v.y = drop_apply(drop, o + 1, v.y);
        v.z = drop_apply(drop, o + 2, v.z); v.w = drop_apply(drop, o + 3, v.w);
        return v;
    };
    auto update = [&](int row, const float4& v) {          // feature k owns its table rows: no other warp touches them
        float4* t = reinterpret_cast<float4*>(tab + row * kEmbSlab + threadIdx.x * 4);
        float4 a = *t;
        a.x += v.x; a.y += v.y; a.z += v.z; a.w += v.w;
        *t = a;
    };
    if (c < H) {                                            // H % 4 == 0: a lane's four columns are in or out together
        int r = r0;
        for (; r + kEmbFlight <= r1; r += kEmbFlight) {     // independent loads in flight, then updates in row order
            float4 v[kEmbFlight];
            int ix[kEmbFlight];
#pragma unroll
            for (int u = 0; u < kEmbFlight; ++u) {
                v[u] = fetch(r + u);
                ix[u] = off + idx[(long long)(r + u) * K + k];
            }
#pragma unroll
            for (int u = 0; u < kEmbFlight; ++u) update(ix[u], v[u]);
        }
        for (; r < r1; ++r) update(off + idx[(long long)r * K + k], fetch(r));
    }
    __syncthreads();
    float* dst = partial + (long long)blockIdx.y * table_rows * H;
    for (int i = tid; i < table_rows * kEmbSlab; i += nthreads) {
        int tr = i / kEmbSlab, cc = blockIdx.x * kEmbSlab + (i % kEmbSlab);
        if (cc < H) dst[(long long)tr * H + cc] = tab[i];
    }
}

__global__ void k_reduce_partials(const float* __restrict__ partial, int nchunks, long long count,
                                  float* __restrict__ out) {
    pdl_enter();
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < count; i += (long long)gridDim.x * blockDim.x) {
        float s = 0.f;
        for (int ch = 0; ch < nchunks; ++ch) s += partial[(long long)ch * count + i];
        out[i] = s;
    }
}

static int embed_chunks(int64_t rows, int32_t H) {
    int slabs = ceil_div(H, kEmbSlab);
    int target = max(1, (num_sms() * 2) / slabs);       // two 89 KB slabs per SM are resident; every chunk costs a [table, H] partial
    int by_rows = max(1, (int)((rows + 255) / 256));    // at least ~256 rows per chunk
    return min(target, by_rows);
}

}  // namespace gnnx

using namespace gnnx;

extern "C" {

int gnnx_embed_sum_fwd(const uint8_t* idx, int64_t rows, int32_t K, const int32_t* offsets, const float* table,
                       int32_t H, float dropout_p, int64_t seed, float* out, void* stream) {
    GNNX_REQUIRE(dropout_p >= 0.f && dropout_p <= 1.f, "embed_sum_fwd: dropout_p must be in [0, 1]");
    GNNX_REQUIRE(K >= 1 && K <= 32, "embed_sum_fwd: num_features must be in [1,32]");
    GNNX_REQUIRE(H % 4 == 0, "embed_sum_fwd: hidden must be a multiple of 4");
    if (rows == 0) return GNNX_OK;
    int nb = min(ceil_div(rows, 8), num_sms() * 8);
    launch(k_embed_fwd, nb, 256, 0, as_stream(stream), idx, (int)rows, K, offsets, table, H, out, node_dropout(dropout_p, seed));
    return check_launch("embed_sum_fwd");
}

size_t gnnx_embed_sum_bwd_workspace_bytes(int64_t rows, int32_t table_rows, int32_t H) {
    return (size_t)embed_chunks(rows, H) * table_rows * H * sizeof(float);
}

int gnnx_embed_sum_bwd(const uint8_t* idx, int64_t rows, int32_t K, const int32_t* offsets, int32_t table_rows,
                       const float* dout, int32_t H, float dropout_p, int64_t seed, float* dtable, void* ws, size_t ws_bytes,
                       void* stream) {
    GNNX_REQUIRE(K >= 1 && K <= 32, "embed_sum_bwd: num_features must be in [1,32]");
    GNNX_REQUIRE(H % 4 == 0 && ((uintptr_t)dout & 15) == 0, "embed_sum_bwd: hidden must be a multiple of 4 and dout 16-byte aligned");
    GNNX_REQUIRE(ws_bytes >= gnnx_embed_sum_bwd_workspace_bytes(rows, table_rows, H), "embed_sum_bwd: workspace too small");
    cudaStream_t st = as_stream(stream);
    long long count = (long long)table_rows * H;
    if (rows == 0) { GNNX_CUDA(cudaMemsetAsync(dtable, 0, count * sizeof(float), st)); return GNNX_OK; }
    int nchunks = embed_chunks(rows, H);
    int rpc = ceil_div(rows, nchunks);
    size_t smem = (size_t)table_rows * kEmbSlab * sizeof(float);
    GNNX_REQUIRE(smem <= 200 * 1024, "embed_sum_bwd: table too large for the shared-memory slab");
    if (smem > 48 * 1024)
        GNNX_CUDA(cudaFuncSetAttribute(k_embed_bwd_partial, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    launch(k_embed_bwd_partial, dim3(ceil_div(H, kEmbSlab), nchunks), dim3(32, K), smem, st, 
        idx, (int)rows, K, offsets, table_rows, dout, H, rpc, (float*)ws, node_dropout(dropout_p, seed));
    launch(k_reduce_partials, min(ceil_div(count, 256), num_sms() * 8), 256, 0, st, (const float*)ws, nchunks, count, dtable);
    return check_launch("embed_sum_bwd", 2);
}

}  // extern "C"
