// BatchNorm(+ReLU) over node rows, graph-level segment pooling (K9), the fused
// relative-error Huber loss (K11) and AdamW-amsgrad over the flat parameter buffer (K12).
// Reductions over rows are two-pass and deterministic: per-chunk partials (Welford/Chan for the
// variance) merged in chunk order -- no atomics.
#include "common.cuh"
#include <atomic>
#include <cstdlib>

namespace gnnx {

union F4 {
    float4 q;
    float a[4];
};


struct Welford {
    float n, mean, m2;
};
__device__ __forceinline__ void welford_add(Welford& w, float x) {
    w.n += 1.f;
    float d = x - w.mean;
    w.mean += d / w.n;
    w.m2 += d * (x - w.mean);
}
__device__ __forceinline__ void welford_merge(Welford& a, const Welford& b) {
    if (b.n == 0.f) return;
    float n = a.n + b.n;
    float d = b.mean - a.mean;
    a.mean += d * (b.n / n);
    a.m2 += b.m2 + d * d * (a.n * b.n / n);
    a.n = n;
}

static int bn_chunks(int64_t rows, int32_t C) {
    int slabs = ceil_div(C, 256);
    int target = max(1, (num_sms() * 4) / slabs);
    int by_rows = max(1, (int)((rows + 63) / 64));
    return min(target, by_rows);
}

// partial: [nchunks][3][C] = (n, mean, M2).  Block = 64 float4 column groups x 4 row lanes: a warp reads 512
// contiguous bytes of a row, 4 rows in flight per thread.  Per thread the sums are taken about a pivot (the chunk's
// first row), which keeps the single-pass variance as accurate as Welford's update without a division per element;
// lanes and chunks are then merged with Chan's formula in a fixed order.
__global__ void __launch_bounds__(256)
k_bn_stats_partial(const float* __restrict__ x, int rows, int C, int rows_per_chunk, float* __restrict__ partial,
                   const int* __restrict__ row_node) {
    pdl_enter();
    // row_node != null: x is in degree-bucket order and rows with row_node[r] < 0 are padding (not part of the batch)
    __shared__ float sh[4][3][256 + 4];
    const int tx = threadIdx.x & 63, ty = threadIdx.x >> 6;
    const int c = (blockIdx.x * 64 + tx) * 4;
    const int r0 = blockIdx.y * rows_per_chunk, r1 = min(rows, r0 + rows_per_chunk);
    F4 s, q, p;
    s.q = q.q = p.q = make_float4(0.f, 0.f, 0.f, 0.f);
    float n = 0.f;
    // the pilot must be a row of the batch: a padding row of the bucket order (the layer's response to a zero operand)
    // can lie tens of standard deviations away from the data, and q - s^2 / n then cancels (measured: 2.7e-4 on the
    // predictions of a 3-layer model).  Padding sits at the end of each bucket, so the scan is short and uniform.
    int rp = r0;
    if (row_node)
        while (rp < r1 && __ldg(row_node + rp) < 0) ++rp;
    if (c < C && rp < r1) {
        p.q = ldg4(x + (long long)rp * C + c);
        int r = r0 + ty;
        for (; r + 12 < r1; r += 16) {
            F4 v[4];
            bool ok[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                v[u].q = ldg4_stream(x + (long long)(r + 4 * u) * C + c);
                ok[u] = !row_node || row_node[r + 4 * u] >= 0;
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
#pragma unroll
                for (int j = 0; j < 4; ++j) { float d = ok[u] ? v[u].a[j] - p.a[j] : 0.f; s.a[j] += d; q.a[j] = fmaf(d, d, q.a[j]); }
                n += ok[u] ? 1.f : 0.f;
            }
        }
        for (; r < r1; r += 4) {
            F4 v;
            v.q = ldg4_stream(x + (long long)r * C + c);
            const bool ok = !row_node || row_node[r] >= 0;
#pragma unroll
            for (int j = 0; j < 4; ++j) { float d = ok ? v.a[j] - p.a[j] : 0.f; s.a[j] += d; q.a[j] = fmaf(d, d, q.a[j]); }
            n += ok ? 1.f : 0.f;
        }
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        float mean = n > 0.f ? p.a[j] + s.a[j] / n : 0.f;
        float m2 = n > 0.f ? fmaxf(q.a[j] - s.a[j] * s.a[j] / n, 0.f) : 0.f;
        sh[ty][0][tx * 4 + j] = n; sh[ty][1][tx * 4 + j] = mean; sh[ty][2][tx * 4 + j] = m2;
    }
    __syncthreads();
    const int cc = blockIdx.x * 256 + threadIdx.x;
    if (cc < C) {
        Welford t = {0.f, 0.f, 0.f};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            Welford o = {sh[i][0][threadIdx.x], sh[i][1][threadIdx.x], sh[i][2][threadIdx.x]};
            welford_merge(t, o);
        }
        float* pp = partial + (long long)blockIdx.y * 3 * C;
        pp[cc] = t.n; pp[C + cc] = t.mean; pp[2 * C + cc] = t.m2;
    }
}

// one warp per channel: lanes stride over the chunks, then a fixed shuffle tree merges the 32 lane partials
// (deterministic: the merge order depends only on nchunks)
__global__ void k_bn_stats_final(const float* __restrict__ partial, int nchunks, int C, float eps, float momentum,
                                 float* __restrict__ running_mean, float* __restrict__ running_var,
                                 float* __restrict__ save_mean, float* __restrict__ save_rstd) {
    pdl_enter();
    const int lane = threadIdx.x & 31;
    const int c = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (c >= C) return;                              // warp-uniform
    Welford t = {0.f, 0.f, 0.f};
    for (int base = 0; base < nchunks; base += 32 * 8) {           // 24 independent loads in flight, merged in chunk order
        Welford o[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            const int ch = base + u * 32 + lane;
            const float* p = partial + (long long)ch * 3 * C;
            if (ch < nchunks) { o[u].n = p[c]; o[u].mean = p[C + c]; o[u].m2 = p[2 * C + c]; }
            else { o[u].n = 0.f; o[u].mean = 0.f; o[u].m2 = 0.f; }
        }
#pragma unroll
        for (int u = 0; u < 8; ++u) welford_merge(t, o[u]);
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        Welford o = {__shfl_down_sync(0xffffffffu, t.n, off), __shfl_down_sync(0xffffffffu, t.mean, off),
                     __shfl_down_sync(0xffffffffu, t.m2, off)};
        welford_merge(t, o);
    }
    if (lane != 0) return;
    float var = t.m2 / t.n;                      // biased, used to normalise
    save_mean[c] = t.mean;
    save_rstd[c] = rsqrtf(var + eps);
    if (running_mean) running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * t.mean;
    if (running_var) {
        float unbiased = t.n > 1.f ? t.m2 / (t.n - 1.f) : var;
        running_var[c] = (1.f - momentum) * running_var[c] + momentum * unbiased;
    }
}

__global__ void k_bn_eval_stats(const float* __restrict__ rm, const float* __restrict__ rv, int C, float eps,
                                float* __restrict__ mean, float* __restrict__ rstd) {
    pdl_enter();
    int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c < C) { mean[c] = rm[c]; rstd[c] = rsqrtf(rv[c] + eps); }
}

// y = (x - mean) * rstd * w + b with a fixed operation sequence: the backward re-evaluates it to recover the ReLU mask
// (sign of y) bit for bit instead of reading y back.
__device__ __forceinline__ float bn_affine(float v, float mu, float rs, float w, float b) {
    return __fmaf_rn(__fmul_rn(__fsub_rn(v, mu), rs), w, b);
}

__global__ void k_bn_apply(const float* __restrict__ x, long long rows, int C, const float* __restrict__ weight,
                           const float* __restrict__ bias, const float* __restrict__ mean, const float* __restrict__ rstd,
                           int relu, float* __restrict__ y, const int* __restrict__ row_node, NodeDropout drop) {
    pdl_enter();
    const int c4 = C / 4;
    long long total = rows * c4;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        int c = (int)(i % c4) * 4;
        long long orow = i / c4;                      // bucket-ordered input: the result goes to the node's own row
        if (row_node) { orow = row_node[orow]; if (orow < 0) continue; }
        F4 v, mu, rs, w, b, o;
        v.q = ldg4_stream(x + i * 4);
        mu.q = ldg4(mean + c); rs.q = ldg4(rstd + c);
        w.q = weight ? ldg4(weight + c) : make_float4(1.f, 1.f, 1.f, 1.f);
        b.q = bias ? ldg4(bias + c) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            float t = bn_affine(v.a[j], mu.a[j], rs.a[j], w.a[j], b.a[j]);
            t = relu ? fmaxf(t, 0.f) : t;
            o.a[j] = drop_apply(drop, orow * C + c + j, t);     // Dropout(p) of the NEXT layer's input (models.py:209), node-order index
        }
        stg4(y + orow * C + c, o.q);
    }
}

// ------------------------------------------------------------------ training forward as ONE cooperative launch
// statistics -> (grid barrier) -> per-channel finalisation, channel c by CTA c -> (grid barrier) -> normalise + ReLU.
// Three launches become one, and the second pass re-reads rows this launch has just read: phase 1 loads them with the
// default policy (they stay in L2 as far as its 126 MB go; [101 887 x 256] fp32 is 104 MB), phase 2 streams them,
// each CTA walking its rows backwards (most recently read first).
// The grid is sized by the host so that every CTA is resident (SMs x co-resident CTAs): the barrier is an arrival
// counter in global memory that thread 0 of each CTA polls with acquire loads; a counter that never fills traps after
// ~2 s instead of hanging the GPU.  counters[0] counts arrivals (nch per barrier), counters[1] exits; the last CTA out
// zeroes both for the next launch on the stream.
__global__ void __launch_bounds__(256)
k_bn_fwd_coop(const float* __restrict__ x, int rows, int C, int rows_per_chunk, float* __restrict__ partial,
              const int* __restrict__ row_node, float eps, float momentum, float* __restrict__ running_mean,
              float* __restrict__ running_var, float* __restrict__ save_mean, float* __restrict__ save_rstd,
              const float* __restrict__ weight, const float* __restrict__ bias, int relu, float* __restrict__ y,
              NodeDropout drop, int* __restrict__ counters) {
    pdl_enter();
    __shared__ float sh[4][3][256 + 4];
    const int nch = (int)gridDim.x;
    const int tx = threadIdx.x & 63, ty = threadIdx.x >> 6;
    const int r0 = blockIdx.x * rows_per_chunk, r1 = min(rows, r0 + rows_per_chunk);
    int rp = r0;                                          // pilot: a row of the batch (see k_bn_stats_partial)
    if (row_node)
        while (rp < r1 && __ldg(row_node + rp) < 0) ++rp;
    // ---- phase 1: (n, mean, M2) of this CTA's rows, per column
    for (int c0 = 0; c0 < C; c0 += 256) {
        const int c = c0 + tx * 4;
        F4 s, q, p;
        s.q = q.q = p.q = make_float4(0.f, 0.f, 0.f, 0.f);
        float n = 0.f;
        if (c < C && rp < r1) {
            p.q = ldg4(x + (long long)rp * C + c);
            int r = r0 + ty;
            for (; r + 12 < r1; r += 16) {
                F4 v[4];
                bool ok[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    v[u].q = ldg4(x + (long long)(r + 4 * u) * C + c);
                    ok[u] = !row_node || row_node[r + 4 * u] >= 0;
                }
#pragma unroll
                for (int u = 0; u < 4; ++u) {
#pragma unroll
                    for (int j = 0; j < 4; ++j) { float d = ok[u] ? v[u].a[j] - p.a[j] : 0.f; s.a[j] += d; q.a[j] = fmaf(d, d, q.a[j]); }
                    n += ok[u] ? 1.f : 0.f;
                }
            }
            for (; r < r1; r += 4) {
                F4 v;
                v.q = ldg4(x + (long long)r * C + c);
                const bool ok = !row_node || row_node[r] >= 0;
#pragma unroll
                for (int j = 0; j < 4; ++j) { float d = ok ? v.a[j] - p.a[j] : 0.f; s.a[j] += d; q.a[j] = fmaf(d, d, q.a[j]); }
                n += ok ? 1.f : 0.f;
            }
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            float mean = n > 0.f ? p.a[j] + s.a[j] / n : 0.f;
            float m2 = n > 0.f ? fmaxf(q.a[j] - s.a[j] * s.a[j] / n, 0.f) : 0.f;
            sh[ty][0][tx * 4 + j] = n; sh[ty][1][tx * 4 + j] = mean; sh[ty][2][tx * 4 + j] = m2;
        }
        __syncthreads();
        const int cc = c0 + threadIdx.x;
        if (cc < C) {
            Welford t = {0.f, 0.f, 0.f};
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                Welford o = {sh[i][0][threadIdx.x], sh[i][1][threadIdx.x], sh[i][2][threadIdx.x]};
                welford_merge(t, o);
            }
            float* pp = partial + (long long)blockIdx.x * 3 * C;
            pp[cc] = t.n; pp[C + cc] = t.mean; pp[2 * C + cc] = t.m2;
        }
        __syncthreads();
    }
    grid_barrier(counters, nch);
    // ---- finalisation: channel c by CTA c (+ k * grid): thread t merges chunks t, t + 256, ... in order, then a fixed
    // shuffle tree per warp and the eight warps in order
    for (int c = blockIdx.x; c < C; c += nch) {
        Welford t = {0.f, 0.f, 0.f};
        for (int ch = threadIdx.x; ch < nch; ch += 256) {
            const float* p = partial + (long long)ch * 3 * C;
            Welford o = {__ldcg(p + c), __ldcg(p + C + c), __ldcg(p + 2 * C + c)};
            welford_merge(t, o);
        }
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
            Welford o = {__shfl_down_sync(0xffffffffu, t.n, off), __shfl_down_sync(0xffffffffu, t.mean, off),
                         __shfl_down_sync(0xffffffffu, t.m2, off)};
            welford_merge(t, o);
        }
        __syncthreads();
        if ((threadIdx.x & 31) == 0) { sh[0][0][threadIdx.x >> 5] = t.n; sh[0][1][threadIdx.x >> 5] = t.mean; sh[0][2][threadIdx.x >> 5] = t.m2; }
        __syncthreads();
        if (threadIdx.x == 0) {
            Welford a = {0.f, 0.f, 0.f};
            for (int w = 0; w < 8; ++w) {
                Welford o = {sh[0][0][w], sh[0][1][w], sh[0][2][w]};
                welford_merge(a, o);
            }
            const float var = a.m2 / a.n;
            save_mean[c] = a.mean;
            save_rstd[c] = rsqrtf(var + eps);
            if (running_mean) running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * a.mean;
            if (running_var) {
                const float unbiased = a.n > 1.f ? a.m2 / (a.n - 1.f) : var;
                running_var[c] = (1.f - momentum) * running_var[c] + momentum * unbiased;
            }
        }
    }
    grid_barrier(counters, 2 * nch);
    // ---- phase 2: normalise this CTA's rows (the addresses phase 1 read)
    for (int c0 = 0; c0 < C; c0 += 256) {
        const int c = c0 + tx * 4;
        if (c >= C) continue;
        F4 mu, rs, w, b;
        mu.q = __ldcg(reinterpret_cast<const float4*>(save_mean + c)); rs.q = __ldcg(reinterpret_cast<const float4*>(save_rstd + c));
        w.q = weight ? ldg4(weight + c) : make_float4(1.f, 1.f, 1.f, 1.f);
        b.q = bias ? ldg4(bias + c) : make_float4(0.f, 0.f, 0.f, 0.f);
        // last rows first: what phase 1 read most recently is what L2 still holds
        for (int r = r0 + ty + 16 * max(0, (r1 - r0 - ty + 15) / 16 - 1); r >= r0 + ty; r -= 16) {
            F4 v[4];
            long long orow[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int rr = min(r + 4 * u, r1 - 1);                  // clamped: the loads stay unconditional
                v[u].q = ldg4_stream(x + (long long)rr * C + c);
                orow[u] = r + 4 * u < r1 ? (row_node ? (long long)row_node[rr] : (long long)rr) : -1;
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                if (orow[u] < 0) continue;
                F4 o;
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    float t = bn_affine(v[u].a[j], mu.a[j], rs.a[j], w.a[j], b.a[j]);
                    t = relu ? fmaxf(t, 0.f) : t;
                    o.a[j] = drop_apply(drop, orow[u] * C + c + j, t);
                }
                stg4(y + orow[u] * C + c, o.q);
            }
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        if (atomicAdd(counters + 1, 1) == nch - 1) { counters[0] = 0; counters[1] = 0; __threadfence(); }
    }
}

// partial: [nchunks][2][C] = (sum dy', sum dy' * xhat); same thread layout as k_bn_stats_partial
__global__ void __launch_bounds__(256)
k_bn_bwd_partial(const float* __restrict__ dy, const float* __restrict__ x, const float* __restrict__ y,
                 int rows, int C, const float* __restrict__ mean, const float* __restrict__ rstd, int relu,
                 int rows_per_chunk, float* __restrict__ partial, const int* __restrict__ row_node,
                 const float* __restrict__ weight, const float* __restrict__ bias, NodeDropout drop) {
    pdl_enter();
    (void)y;                                       // the ReLU mask is re-derived from x (bn_affine): y is not read
    // row_node != null: x is in degree-bucket order (rows = bucket capacity), dy / y in node order
    __shared__ float sh[4][2][256 + 4];
    const int tx = threadIdx.x & 63, ty = threadIdx.x >> 6;
    const int c = (blockIdx.x * 64 + tx) * 4;
    const int r0 = blockIdx.y * rows_per_chunk, r1 = min(rows, r0 + rows_per_chunk);
    F4 s1, s2;
    s1.q = s2.q = make_float4(0.f, 0.f, 0.f, 0.f);
    if (c < C) {
        F4 mu, rs, wv, bv;
        mu.q = ldg4(mean + c); rs.q = ldg4(rstd + c);
        wv.q = weight ? ldg4(weight + c) : make_float4(1.f, 1.f, 1.f, 1.f);
        bv.q = bias ? ldg4(bias + c) : make_float4(0.f, 0.f, 0.f, 0.f);
        int r = r0 + ty;
        for (; r + 4 < r1; r += 8) {
            F4 g[2], xv[2];
            bool pad[2] = {false, false};
            long long oi[2] = {0, 0};
#pragma unroll
            for (int u = 0; u < 2; ++u) {
                long long o = (long long)(r + 4 * u) * C + c;
                xv[u].q = ldg4(x + o);
                if (row_node) {
                    const int node = row_node[r + 4 * u];
                    if (node < 0) { g[u].q = make_float4(0.f, 0.f, 0.f, 0.f); pad[u] = true; continue; }
                    o = (long long)node * C + c;
                }
                oi[u] = o;
                g[u].q = ldg4_stream(dy + o);
            }
#pragma unroll
            for (int u = 0; u < 2; ++u)
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const bool off = pad[u] || (relu && !(bn_affine(xv[u].a[j], mu.a[j], rs.a[j], wv.a[j], bv.a[j]) > 0.f));
                    float gg = off ? 0.f : drop_apply(drop, oi[u] + j, g[u].a[j]);
                    s1.a[j] += gg;
                    s2.a[j] = fmaf(gg, (xv[u].a[j] - mu.a[j]) * rs.a[j], s2.a[j]);
                }
        }
        for (; r < r1; r += 4) {
            long long o = (long long)r * C + c;
            F4 g, xv;
            xv.q = ldg4(x + o);
            if (row_node) {
                const int node = row_node[r];
                if (node < 0) continue;
                o = (long long)node * C + c;
            }
            g.q = ldg4_stream(dy + o);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                float gg = (relu && !(bn_affine(xv.a[j], mu.a[j], rs.a[j], wv.a[j], bv.a[j]) > 0.f)) ? 0.f : drop_apply(drop, o + j, g.a[j]);
                s1.a[j] += gg;
                s2.a[j] = fmaf(gg, (xv.a[j] - mu.a[j]) * rs.a[j], s2.a[j]);
            }
        }
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) { sh[ty][0][tx * 4 + j] = s1.a[j]; sh[ty][1][tx * 4 + j] = s2.a[j]; }
    __syncthreads();
    const int cc = blockIdx.x * 256 + threadIdx.x;
    if (cc < C) {
        float a = 0.f, b = 0.f;
#pragma unroll
        for (int i = 0; i < 4; ++i) { a += sh[i][0][threadIdx.x]; b += sh[i][1][threadIdx.x]; }
        partial[(long long)blockIdx.y * 2 * C + cc] = a;
        partial[(long long)blockIdx.y * 2 * C + C + cc] = b;
    }
}

// one warp per channel: lanes stride over the chunks (each lane sums its chunks in order), then a fixed shuffle tree
__global__ void k_bn_bwd_final(const float* __restrict__ partial, int nchunks, int C, float* __restrict__ dweight, float* __restrict__ dbias) {
    pdl_enter();
    const int lane = threadIdx.x & 31;
    const int c = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (c >= C) return;                              // warp-uniform
    float a = 0.f, b = 0.f;
    for (int ch = lane; ch < nchunks; ch += 32) {
        a += partial[(long long)ch * 2 * C + c];
        b += partial[(long long)ch * 2 * C + C + c];
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        a += __shfl_down_sync(0xffffffffu, a, off);
        b += __shfl_down_sync(0xffffffffu, b, off);
    }
    if (lane == 0) { dbias[c] = a; dweight[c] = b; }
}

__global__ void k_bn_bwd_apply(const float* __restrict__ dy, const float* __restrict__ x, const float* __restrict__ y,
                               long long rows, int C, const float* __restrict__ weight, const float* __restrict__ mean,
                               const float* __restrict__ rstd, const float* __restrict__ dweight,
                               const float* __restrict__ dbias, int relu, float* __restrict__ dx,
                               const int* __restrict__ row_node, long long n_valid, const float* __restrict__ bias,
                               NodeDropout drop, int training) {
    pdl_enter();
    (void)y;
    const int c4 = C / 4;
    long long total = rows * c4;
    // eval mode: mean / rstd are constants (running statistics), so dx = w rstd dy' with no batch-coupling terms
    const float inv_n = training ? 1.f / (float)n_valid : 0.f;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        int c = (int)(i % c4) * 4;
        long long ni = i;                             // bucket-ordered x / dx: dy and y are read at the node's own row
        if (row_node) {
            const int node = row_node[i / c4];
            if (node < 0) { stg4(dx + i * 4, make_float4(0.f, 0.f, 0.f, 0.f)); continue; }   // padding rows carry no gradient
            ni = (long long)node * c4 + (i % c4);
        }
        F4 g, xv, mu, rs, w, bb, dw, db, o;
        g.q = ldg4_stream(dy + ni * 4);
        xv.q = ldg4_stream(x + i * 4);
        mu.q = ldg4(mean + c); rs.q = ldg4(rstd + c);
        w.q = weight ? ldg4(weight + c) : make_float4(1.f, 1.f, 1.f, 1.f);
        bb.q = bias ? ldg4(bias + c) : make_float4(0.f, 0.f, 0.f, 0.f);
        dw.q = ldg4(dweight + c); db.q = ldg4(dbias + c);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            float gg = (relu && !(bn_affine(xv.a[j], mu.a[j], rs.a[j], w.a[j], bb.a[j]) > 0.f)) ? 0.f : drop_apply(drop, ni * 4 + j, g.a[j]);
            float xh = (xv.a[j] - mu.a[j]) * rs.a[j];
            o.a[j] = w.a[j] * rs.a[j] * (gg - db.a[j] * inv_n - xh * dw.a[j] * inv_n);
        }
        stg4(dx + i * 4, o.q);
    }
}

// ------------------------------------------------------------------ backward as ONE cooperative launch (see k_bn_fwd_coop)
// (sum dy', sum dy' xhat) per CTA -> barrier -> channel c folded by CTA c -> barrier -> dx of the same rows.
__global__ void __launch_bounds__(256)
k_bn_bwd_coop(const float* __restrict__ dy, const float* __restrict__ x, int rows, int C, const float* __restrict__ mean,
              const float* __restrict__ rstd, int relu, int rows_per_chunk, float* __restrict__ partial,
              const int* __restrict__ row_node, const float* __restrict__ weight, const float* __restrict__ bias,
              NodeDropout drop, float* __restrict__ dweight, float* __restrict__ dbias, float* __restrict__ dx,
              long long n_valid, int training, int* __restrict__ counters) {
    pdl_enter();
    __shared__ float sh[4][2][256 + 4];
    const int nch = (int)gridDim.x;
    const int tx = threadIdx.x & 63, ty = threadIdx.x >> 6;
    const int r0 = blockIdx.x * rows_per_chunk, r1 = min(rows, r0 + rows_per_chunk);
    for (int c0 = 0; c0 < C; c0 += 256) {
        const int c = c0 + tx * 4;
        F4 s1, s2;
        s1.q = s2.q = make_float4(0.f, 0.f, 0.f, 0.f);
        if (c < C) {
            F4 mu, rs, wv, bv;
            mu.q = ldg4(mean + c); rs.q = ldg4(rstd + c);
            wv.q = weight ? ldg4(weight + c) : make_float4(1.f, 1.f, 1.f, 1.f);
            bv.q = bias ? ldg4(bias + c) : make_float4(0.f, 0.f, 0.f, 0.f);
            for (int r = r0 + ty; r < r1; r += 8) {
                F4 g[2], xv[2];
                bool off_row[2];
                long long oi[2];
#pragma unroll
                for (int u = 0; u < 2; ++u) {
                    const int rr = min(r + 4 * u, r1 - 1);              // clamped: unconditional loads
                    xv[u].q = ldg4(x + (long long)rr * C + c);
                    const int node = row_node ? row_node[rr] : rr;
                    off_row[u] = r + 4 * u >= r1 || node < 0;
                    oi[u] = (long long)max(node, 0) * C + c;
                    g[u].q = ldg4(dy + oi[u]);
                }
#pragma unroll
                for (int u = 0; u < 2; ++u)
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const bool off = off_row[u] || (relu && !(bn_affine(xv[u].a[j], mu.a[j], rs.a[j], wv.a[j], bv.a[j]) > 0.f));
                        const float gg = off ? 0.f : drop_apply(drop, oi[u] + j, g[u].a[j]);
                        s1.a[j] += gg;
                        s2.a[j] = fmaf(gg, (xv[u].a[j] - mu.a[j]) * rs.a[j], s2.a[j]);
                    }
            }
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) { sh[ty][0][tx * 4 + j] = s1.a[j]; sh[ty][1][tx * 4 + j] = s2.a[j]; }
        __syncthreads();
        const int cc = c0 + threadIdx.x;
        if (cc < C) {
            float a = 0.f, b = 0.f;
#pragma unroll
            for (int i = 0; i < 4; ++i) { a += sh[i][0][threadIdx.x]; b += sh[i][1][threadIdx.x]; }
            partial[(long long)blockIdx.x * 2 * C + cc] = a;
            partial[(long long)blockIdx.x * 2 * C + C + cc] = b;
        }
        __syncthreads();
    }
    grid_barrier(counters, nch);
    for (int c = blockIdx.x; c < C; c += nch) {
        float a = 0.f, b = 0.f;
        for (int ch = threadIdx.x; ch < nch; ch += 256) {
            a += __ldcg(partial + (long long)ch * 2 * C + c);
            b += __ldcg(partial + (long long)ch * 2 * C + C + c);
        }
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
            a += __shfl_down_sync(0xffffffffu, a, off);
            b += __shfl_down_sync(0xffffffffu, b, off);
        }
        __syncthreads();
        if ((threadIdx.x & 31) == 0) { sh[0][0][threadIdx.x >> 5] = a; sh[0][1][threadIdx.x >> 5] = b; }
        __syncthreads();
        if (threadIdx.x == 0) {
            float sa = 0.f, sb = 0.f;
            for (int w = 0; w < 8; ++w) { sa += sh[0][0][w]; sb += sh[0][1][w]; }
            dbias[c] = sa; dweight[c] = sb;
        }
    }
    grid_barrier(counters, 2 * nch);
    const float inv_n = training ? 1.f / (float)n_valid : 0.f;
    for (int c0 = 0; c0 < C; c0 += 256) {
        const int c = c0 + tx * 4;
        if (c >= C) continue;
        F4 mu, rs, w, bb, dw, db;
        mu.q = ldg4(mean + c); rs.q = ldg4(rstd + c);
        w.q = weight ? ldg4(weight + c) : make_float4(1.f, 1.f, 1.f, 1.f);
        bb.q = bias ? ldg4(bias + c) : make_float4(0.f, 0.f, 0.f, 0.f);
        dw.q = __ldcg(reinterpret_cast<const float4*>(dweight + c)); db.q = __ldcg(reinterpret_cast<const float4*>(dbias + c));
        for (int r = r0 + ty + 8 * max(0, (r1 - r0 - ty + 7) / 8 - 1); r >= r0 + ty; r -= 8) {     // last rows first (L2)
            F4 g[2], xv[2];
            int node[2];
            long long oi[2];
#pragma unroll
            for (int u = 0; u < 2; ++u) {
                const int rr = min(r + 4 * u, r1 - 1);
                xv[u].q = ldg4_stream(x + (long long)rr * C + c);
                node[u] = row_node ? row_node[rr] : rr;
                oi[u] = (long long)max(node[u], 0) * C + c;
                g[u].q = ldg4_stream(dy + oi[u]);
            }
#pragma unroll
            for (int u = 0; u < 2; ++u) {
                if (r + 4 * u >= r1) continue;
                F4 o;
                if (node[u] < 0) {
                    o.q = make_float4(0.f, 0.f, 0.f, 0.f);               // padding rows carry no gradient
                } else {
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const float gg = (relu && !(bn_affine(xv[u].a[j], mu.a[j], rs.a[j], w.a[j], bb.a[j]) > 0.f)) ? 0.f : drop_apply(drop, oi[u] + j, g[u].a[j]);
                        const float xh = (xv[u].a[j] - mu.a[j]) * rs.a[j];
                        o.a[j] = w.a[j] * rs.a[j] * (gg - db.a[j] * inv_n - xh * dw.a[j] * inv_n);
                    }
                }
                stg4(dx + (long long)(r + 4 * u) * C + c, o.q);
            }
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        if (atomicAdd(counters + 1, 1) == nch - 1) { counters[0] = 0; counters[1] = 0; __threadfence(); }
    }
}

// ------------------------------------------------------------------ graph pooling over contiguous node ranges
// one warp per (graph, 128-col chunk); mode 0 add, 1 mean, 2 max (empty graph -> 0)
__global__ void k_pool_fwd(const float* __restrict__ x, const int* __restrict__ gp, int G, int C, int mode, float* __restrict__ out) {
    pdl_enter();
    const int lane = threadIdx.x & 31;
    const int chunks = (C + 127) / 128;
    const long long nwork = (long long)G * chunks;
    const int wpb = blockDim.x >> 5;
    for (long long w = blockIdx.x * (long long)wpb + (threadIdx.x >> 5); w < nwork; w += (long long)gridDim.x * wpb) {
        int g = (int)(w / chunks);
        int c = (int)(w % chunks) * 128 + lane * 4;
        if (c >= C) continue;
        int b = gp[g], e = gp[g + 1];
        F4 acc;
        float init = mode == 2 ? -INFINITY : 0.f;
        acc.q = make_float4(init, init, init, init);
        for (int r = b; r < e; ++r) {
            F4 v; v.q = ldg4(x + (long long)r * C + c);
#pragma unroll
            for (int j = 0; j < 4; ++j) acc.a[j] = mode == 2 ? fmaxf(acc.a[j], v.a[j]) : acc.a[j] + v.a[j];
        }
        if (mode == 1) {
            float cnt = fmaxf((float)(e - b), 1.f);
#pragma unroll
            for (int j = 0; j < 4; ++j) acc.a[j] = acc.a[j] / cnt;
        }
        if (mode == 2 && b == e) acc.q = make_float4(0.f, 0.f, 0.f, 0.f);
        stg4(out + (long long)g * C + c, acc.q);
    }
}

__global__ void k_pool_bwd(const float* __restrict__ dout, const float* __restrict__ x, const float* __restrict__ out,
                           const int* __restrict__ gp, int G, int C, int mode, float* __restrict__ dx) {
    pdl_enter();
    const int lane = threadIdx.x & 31;
    const int chunks = (C + 127) / 128;
    const long long nwork = (long long)G * chunks;
    const int wpb = blockDim.x >> 5;
    for (long long w = blockIdx.x * (long long)wpb + (threadIdx.x >> 5); w < nwork; w += (long long)gridDim.x * wpb) {
        int g = (int)(w / chunks);
        int c = (int)(w % chunks) * 128 + lane * 4;
        if (c >= C) continue;
        int b = gp[g], e = gp[g + 1];
        if (b == e) continue;
        F4 go; go.q = ldg4(dout + (long long)g * C + c);
        if (mode == 0 || mode == 1) {
            if (mode == 1) {
                float inv = 1.f / (float)(e - b);
#pragma unroll
                for (int j = 0; j < 4; ++j) go.a[j] *= inv;
            }
            for (int r = b; r < e; ++r) stg4(dx + (long long)r * C + c, go.q);
        } else {
            F4 mx; mx.q = ldg4(out + (long long)g * C + c);
            F4 nt; nt.q = make_float4(0.f, 0.f, 0.f, 0.f);
            for (int r = b; r < e; ++r) {
                F4 v; v.q = ldg4(x + (long long)r * C + c);
#pragma unroll
                for (int j = 0; j < 4; ++j) nt.a[j] += (v.a[j] == mx.a[j]) ? 1.f : 0.f;
            }
            for (int r = b; r < e; ++r) {
                F4 v, o; v.q = ldg4(x + (long long)r * C + c);
#pragma unroll
                for (int j = 0; j < 4; ++j) o.a[j] = (v.a[j] == mx.a[j]) ? go.a[j] / nt.a[j] : 0.f;
                stg4(dx + (long long)r * C + c, o.q);
            }
        }
    }
}

// ------------------------------------------------------------------ loss: Huber(delta) on the relative error + MAPE
__global__ void k_relhuber(const float* __restrict__ pred, const float* __restrict__ target, int count, float delta,
                           float* __restrict__ out2, float* __restrict__ dpred) {
    pdl_enter();
    __shared__ float sh[2][32];
    float l = 0.f, mp = 0.f;
    const float inv = 1.f / (float)count;
    for (int i = threadIdx.x; i < count; i += blockDim.x) {
        float p = pred[i], t = target[i];
        float a = (p - t) / t;
        float aa = fabsf(a);
        float g;
        if (aa <= delta) { l += 0.5f * a * a; g = a; }
        else { l += delta * (aa - 0.5f * delta); g = a > 0.f ? delta : -delta; }
        mp += fabsf(p - t) / fmaxf(fabsf(t), 1.17e-06f);
        if (dpred) dpred[i] = g / t * inv;
    }
    l = warp_sum(l); mp = warp_sum(mp);
    int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (lane == 0) { sh[0][w] = l; sh[1][w] = mp; }
    __syncthreads();
    if (w == 0) {
        int nw = blockDim.x >> 5;
        l = lane < nw ? sh[0][lane] : 0.f;
        mp = lane < nw ? sh[1][lane] : 0.f;
        l = warp_sum(l); mp = warp_sum(mp);
        if (lane == 0) { out2[0] = l * inv; out2[1] = mp * inv; }
    }
}

// ------------------------------------------------------------------ AdamW(amsgrad) on the flat buffer
__global__ void k_adamw_amsgrad(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
                                float* __restrict__ vmax, long long count, float lr, float b1, float b2, float eps, float wd,
                                float bc1, float bc2_sqrt, float gscale) {
    pdl_enter();
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < count; i += (long long)gridDim.x * blockDim.x) {
        float grad = g[i] * gscale;
        float param = p[i] * (1.f - lr * wd);
        float mi = m[i] + (grad - m[i]) * (1.f - b1);
        float vi = v[i] * b2 + grad * grad * (1.f - b2);
        float vm = fmaxf(vmax[i], vi);
        float denom = sqrtf(vm) / bc2_sqrt + eps;
        param -= (lr / bc1) * (mi / denom);
        p[i] = param; m[i] = mi; v[i] = vi; vmax[i] = vm;
    }
}

static inline int flat_grid(long long nthreads) {
    long long blocks = (nthreads + 255) / 256;
    long long cap = (long long)num_sms() * 32;
    return (int)(blocks < cap ? (blocks > 0 ? blocks : 1) : cap);
}

}  // namespace gnnx

using namespace gnnx;

extern "C" {

size_t gnnx_bn_workspace_bytes(int64_t rows, int32_t C) {
    return ((size_t)bn_chunks(rows, C) * 3 * C + 2 * (size_t)C) * sizeof(float);
}

// co-resident CTAs per SM of the cooperative kernels (0: forward, 1: backward); -1: unavailable or disabled
static int coop_ctas(int which) {
    static std::atomic<int> cached[2] = {{0}, {0}};
    int k = cached[which].load();
    if (k == 0) {
        const char* e = getenv("GNNX_BN_COOP");
        int n = 0;
        cudaError_t rc = which == 0 ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, k_bn_fwd_coop, 256, 0)
                                    : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, k_bn_bwd_coop, 256, 0);
        if ((e && e[0] == '0') || rc != cudaSuccess || n < 1) { cudaGetLastError(); n = -1; }
        cached[which].store(n); k = n;
    }
    return k;
}

static int bn_fwd_impl(const float* x, int64_t rows, int64_t n_valid, const int32_t* row_node, int32_t C, const float* weight,
                       const float* bias, float* running_mean, float* running_var, int32_t training, float momentum, float eps,
                       int32_t relu, float dropout_p, int64_t seed, float* y, float* save_mean, float* save_rstd, void* ws,
                       size_t ws_bytes, void* stream) {
    GNNX_REQUIRE(dropout_p >= 0.f && dropout_p <= 1.f, "bn_relu_fwd: dropout_p must be in [0, 1]");
    GNNX_REQUIRE(C % 4 == 0, "bn_relu_fwd: channels must be a multiple of 4");
    GNNX_REQUIRE(ws_bytes >= gnnx_bn_workspace_bytes(rows, C), "bn_relu_fwd: workspace too small");
    GNNX_REQUIRE(save_mean && save_rstd, "bn_relu_fwd: save_mean / save_rstd are required");
    if (rows == 0 || n_valid == 0) return GNNX_OK;
    cudaStream_t st = as_stream(stream);
    if (training) {
        GNNX_REQUIRE(n_valid > 1, "bn_relu_fwd: training-mode batch norm needs more than one row");
        int nch = bn_chunks(rows, C);
        int rpc = ceil_div(rows, nch);
        // one cooperative launch when every CTA of the grid can be resident (GNNX_BN_COOP=0: the three-kernel path)
        const int k = coop_ctas(0);
        if (k > 0 && rows < 0x7fffffff) {
            int cn = min(nch, num_sms() * min(k, 4));
            int* ctr = stream_ticket(st);
            if (ctr) {
                launch_coop(k_bn_fwd_coop, cn, 256, 0, st, x, (int)rows, C, ceil_div(rows, cn), (float*)ws, row_node, eps, momentum, running_mean,
                       running_var, save_mean, save_rstd, weight, bias, relu, y, node_dropout(dropout_p, seed), ctr + 8);
                return check_launch("bn_relu_fwd", 1);
            }
        }
        launch(k_bn_stats_partial, dim3(ceil_div(C, 256), nch), 256, 0, st, x, (int)rows, C, rpc, (float*)ws, row_node);
        launch(k_bn_stats_final, ceil_div(C, 8), 256, 0, st, (const float*)ws, nch, C, eps, momentum, running_mean, running_var, save_mean, save_rstd);
    } else {
        GNNX_REQUIRE(running_mean && running_var, "bn_relu_fwd: eval mode needs running statistics");
        launch(k_bn_eval_stats, ceil_div(C, 128), 128, 0, st, running_mean, running_var, C, eps, save_mean, save_rstd);
    }
    launch(k_bn_apply, flat_grid(rows * (C / 4)), 256, 0, st, x, rows, C, weight, bias, save_mean, save_rstd, relu, y, row_node,
           node_dropout(dropout_p, seed));
    return check_launch("bn_relu_fwd", training ? 3 : 2);
}

int gnnx_bn_relu_fwd(const float* x, int64_t rows, int32_t C, const float* weight, const float* bias, float* running_mean,
                     float* running_var, int32_t training, float momentum, float eps, int32_t relu, float dropout_p,
                     int64_t seed, float* y, float* save_mean, float* save_rstd, void* ws, size_t ws_bytes, void* stream) {
    return bn_fwd_impl(x, rows, rows, nullptr, C, weight, bias, running_mean, running_var, training, momentum, eps, relu,
                       dropout_p, seed, y, save_mean, save_rstd, ws, ws_bytes, stream);
}

int gnnx_bn_relu_rows_fwd(const float* x, int64_t rows, const int32_t* row_node, int64_t num_nodes, int32_t C,
                          const float* weight, const float* bias, float* running_mean, float* running_var, int32_t training,
                          float momentum, float eps, int32_t relu, float dropout_p, int64_t seed, float* y, float* save_mean,
                          float* save_rstd, void* ws, size_t ws_bytes, void* stream) {
    GNNX_REQUIRE(row_node, "bn_relu_rows_fwd: row_node is required");
    return bn_fwd_impl(x, rows, num_nodes, row_node, C, weight, bias, running_mean, running_var, training, momentum, eps, relu,
                       dropout_p, seed, y, save_mean, save_rstd, ws, ws_bytes, stream);
}

static int bn_bwd_impl(const float* dy, const float* x, const float* y, int64_t rows, int64_t n_valid, const int32_t* row_node,
                       int32_t C, const float* weight, const float* bias, const float* save_mean, const float* save_rstd, int32_t relu,
                       int32_t training, float dropout_p, int64_t seed, float* dx,
                       float* dweight, float* dbias, void* ws, size_t ws_bytes, void* stream) {
    const NodeDropout drop = node_dropout(dropout_p, seed);
    GNNX_REQUIRE(C % 4 == 0, "bn_relu_bwd: channels must be a multiple of 4");
    GNNX_REQUIRE(ws_bytes >= gnnx_bn_workspace_bytes(rows, C), "bn_relu_bwd: workspace too small");
    if (rows == 0) return GNNX_OK;
    cudaStream_t st = as_stream(stream);
    int nch = bn_chunks(rows, C);
    int rpc = ceil_div(rows, nch);
    const int k = coop_ctas(1);
    if (k > 0 && rows < 0x7fffffff) {
        int cn = min(nch, num_sms() * min(k, 4));
        int* ctr = stream_ticket(st);
        if (ctr) {
            launch_coop(k_bn_bwd_coop, cn, 256, 0, st, dy, x, (int)rows, C, save_mean, save_rstd, relu, ceil_div(rows, cn), (float*)ws, row_node,
                   weight, bias, drop, dweight, dbias, dx, (long long)n_valid, training, ctr + 8);
            return check_launch("bn_relu_bwd", 1);
        }
    }
    launch(k_bn_bwd_partial, dim3(ceil_div(C, 256), nch), 256, 0, st, dy, x, y, (int)rows, C, save_mean, save_rstd, relu, rpc, (float*)ws, row_node, weight, bias, drop);
    launch(k_bn_bwd_final, ceil_div(C, 8), 256, 0, st, (const float*)ws, nch, C, dweight, dbias);
    launch(k_bn_bwd_apply, flat_grid(rows * (C / 4)), 256, 0, st, dy, x, y, rows, C, weight, save_mean, save_rstd, dweight, dbias, relu, dx, row_node, n_valid, bias, drop, training);
    return check_launch("bn_relu_bwd", 3);
}

int gnnx_bn_relu_bwd(const float* dy, const float* x, const float* y, int64_t rows, int32_t C, const float* weight,
                     const float* bias, const float* save_mean, const float* save_rstd, int32_t relu, int32_t training,
                     float dropout_p, int64_t seed, float* dx, float* dweight, float* dbias, void* ws, size_t ws_bytes,
                     void* stream) {
    return bn_bwd_impl(dy, x, y, rows, rows, nullptr, C, weight, bias, save_mean, save_rstd, relu, training, dropout_p, seed, dx,
                       dweight, dbias, ws, ws_bytes, stream);
}

int gnnx_bn_relu_rows_bwd(const float* dy, const float* x, const float* y, int64_t rows, const int32_t* row_node,
                          int64_t num_nodes, int32_t C, const float* weight, const float* bias, const float* save_mean,
                          const float* save_rstd, int32_t relu, int32_t training, float dropout_p, int64_t seed, float* dx,
                          float* dweight, float* dbias, void* ws, size_t ws_bytes, void* stream) {
    GNNX_REQUIRE(row_node, "bn_relu_rows_bwd: row_node is required");
    return bn_bwd_impl(dy, x, y, rows, num_nodes, row_node, C, weight, bias, save_mean, save_rstd, relu, training, dropout_p, seed,
                       dx, dweight, dbias, ws, ws_bytes, stream);
}

int gnnx_segment_pool_fwd(const float* x, const int32_t* graph_ptr, int64_t G, int32_t C, int32_t mode, float* out, void* stream) {
    GNNX_REQUIRE(C % 4 == 0 && mode >= 0 && mode <= 2, "segment_pool_fwd: channels %% 4 == 0 and mode in {0,1,2}");
    if (G == 0) return GNNX_OK;
    long long nwork = G * ((C + 127) / 128);
    int nb = (int)min((nwork + 7) / 8, (long long)num_sms() * 32);
    launch(k_pool_fwd, nb, 256, 0, as_stream(stream), x, graph_ptr, (int)G, C, mode, out);
    return check_launch("segment_pool_fwd");
}

int gnnx_segment_pool_bwd(const float* dout, const float* x, const float* out, const int32_t* graph_ptr, int64_t G, int32_t C,
                          int32_t mode, float* dx, void* stream) {
    GNNX_REQUIRE(C % 4 == 0 && mode >= 0 && mode <= 2, "segment_pool_bwd: channels %% 4 == 0 and mode in {0,1,2}");
    if (G == 0) return GNNX_OK;
    long long nwork = G * ((C + 127) / 128);
    int nb = (int)min((nwork + 7) / 8, (long long)num_sms() * 32);
    launch(k_pool_bwd, nb, 256, 0, as_stream(stream), dout, x, out, graph_ptr, (int)G, C, mode, dx);
    return check_launch("segment_pool_bwd");
}

int gnnx_relhuber_fwd_bwd(const float* pred, const float* target, int64_t count, float delta, float* out2, float* dpred, void* stream) {
    GNNX_REQUIRE(count > 0 && count < (1ll << 31), "relhuber: count out of range");
    launch(k_relhuber, 1, 1024, 0, as_stream(stream), pred, target, (int)count, delta, out2, dpred);
    return check_launch("relhuber_fwd_bwd");
}

int gnnx_adamw_amsgrad(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, float* max_exp_avg_sq, int64_t count,
                       float lr, float beta1, float beta2, float eps, float weight_decay, int32_t step, float grad_scale, void* stream) {
    GNNX_REQUIRE(step >= 1, "adamw_amsgrad: step must be >= 1");
    if (count == 0) return GNNX_OK;
    double bc1 = 1.0 - pow((double)beta1, (double)step);
    double bc2 = 1.0 - pow((double)beta2, (double)step);
    launch(k_adamw_amsgrad, flat_grid(count), 256, 0, as_stream(stream), param, grad, exp_avg, exp_avg_sq, max_exp_avg_sq, count, lr, beta1,
                                                                      beta2, eps, weight_decay, (float)bc1, (float)sqrt(bc2), grad_scale);
    return check_launch("adamw_amsgrad");
}

}  // extern "C"

namespace gnnx {
// out[g] = [m, sigma, epsilon/k, kappa_AB, epsilon_AB/k, mu, n_a, n_b, MW]  (demo/utils.py:664-691)
__global__ void k_assemble_params(const float* __restrict__ msigmae, const float* __restrict__ log10assoc,
                                  const float* __restrict__ assoc, const float* __restrict__ munanb,
                                  const float* __restrict__ mw, long long G, float* __restrict__ out) {
    pdl_enter();
    for (long long g = blockIdx.x * (long long)blockDim.x + threadIdx.x; g < G; g += (long long)gridDim.x * blockDim.x) {
        float* o = out + g * 9;
        o[0] = msigmae[g * 3]; o[1] = msigmae[g * 3 + 1]; o[2] = msigmae[g * 3 + 2];
        if (log10assoc) {                            // 10 ** (pred * [-1, +1])
            o[3] = exp10f(-log10assoc[g * 2]);
            o[4] = exp10f(log10assoc[g * 2 + 1]);
        } else {
            o[3] = assoc ? assoc[g * 2] : 0.f;
            o[4] = assoc ? assoc[g * 2 + 1] : 0.f;
        }
        o[5] = munanb[g * 3]; o[6] = munanb[g * 3 + 1]; o[7] = munanb[g * 3 + 2];
        o[8] = mw[g];
    }
}
}  // namespace gnnx

namespace gnnx {
__global__ void k_dropout_keep_mask(NodeDropout d, long long count, uint8_t* __restrict__ keep) {
    pdl_enter();
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < count; i += (long long)gridDim.x * blockDim.x)
        keep[i] = (d.thr == 0u || dropout_keep(d.seed, i, d.thr)) ? 1 : 0;
}
}  // namespace gnnx

extern "C" int gnnx_dropout_keep_mask(float dropout_p, int64_t seed, int64_t count, uint8_t* keep, void* stream) {
    GNNX_REQUIRE(dropout_p >= 0.f && dropout_p <= 1.f && count >= 0, "dropout_keep_mask: bad argument");
    if (count == 0) return GNNX_OK;
    launch(gnnx::k_dropout_keep_mask, gnnx::flat_grid(count), 256, 0, gnnx::as_stream(stream), gnnx::node_dropout(dropout_p, seed), count, keep);
    return gnnx::check_launch("dropout_keep_mask");
}

extern "C" int gnnx_assemble_params(const float* msigmae, const float* log10assoc, const float* assoc, const float* munanb,
                                    const float* mw, int64_t G, float* out9, void* stream) {
    GNNX_REQUIRE(msigmae && munanb && mw && out9, "assemble_params: null operand");
    GNNX_REQUIRE(!(log10assoc && assoc), "assemble_params: give the predicted log10 association pair OR known values, not both");
    if (G == 0) return GNNX_OK;
    long long blocks = (G + 255) / 256;
    if (blocks > gnnx::num_sms() * 8) blocks = gnnx::num_sms() * 8;
    launch(gnnx::k_assemble_params, (int)blocks, 256, 0, gnnx::as_stream(stream), msigmae, log10assoc, assoc, munanb, mw, G, out9);
    return gnnx::check_launch("assemble_params");
}
