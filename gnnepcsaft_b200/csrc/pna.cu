// PNA message passing over the dst-sorted CSR: gather, fused multi-aggregator
// segmented reduction (mean|min|max|std in ONE pass over the messages), the matching backward
// scatters, and the scaler-folded combine of the first post_nn Linear.  All kernels are
// deterministic and atomic-free: one warp owns one (destination row, 128-column chunk), walks the
// row's edges in edge order with float4 loads (several rows in flight), and writes each output once.
#include <stdlib.h>
#include <algorithm>
#include <atomic>

#include "common.cuh"

namespace gnnx {

// torch: out.masked_fill(out <= math.sqrt(1e-5), 0) compares in fp32
#define GNNX_STD_MASK 0.003162277629598975f

union F4 {
    float4 q;
    float a[4];
};
__device__ __forceinline__ F4 f4_zero() { F4 r; r.q = make_float4(0.f, 0.f, 0.f, 0.f); return r; }
__device__ __forceinline__ F4 f4_fill(float v) { F4 r; r.q = make_float4(v, v, v, v); return r; }
__device__ __forceinline__ F4 f4_ld(const float* p) { F4 r; r.q = ldg4(p); return r; }
__device__ __forceinline__ F4 f4_lds(const float* p) { F4 r; r.q = ldg4_stream(p); return r; }

// ------------------------------------------------------------------ gather (K3 + first pre_nn Linear, node-side split)
// 8 lanes x 4 columns -> the uint32 word of packed (value > 0) bits for those 32 columns, on every lane of the group
__device__ __forceinline__ uint32_t pack_relu_word(const F4& o, bool act, int lane) {
    uint32_t v = 0;
    if (act) v = ((o.a[0] > 0.f ? 1u : 0u) | (o.a[1] > 0.f ? 2u : 0u) | (o.a[2] > 0.f ? 4u : 0u) | (o.a[3] > 0.f ? 8u : 0u)) << (4 * (lane & 7));
    v |= __shfl_xor_sync(0xffffffffu, v, 1);
    v |= __shfl_xor_sync(0xffffffffu, v, 2);
    v |= __shfl_xor_sync(0xffffffffu, v, 4);
    return v;
}

// V = float4 groups per lane: a warp covers 128 * V columns (V = 2: the whole row at hidden 256 -- half the index loads
// per byte and twice the bytes in flight per warp; these kernels are latency-bound).  Lane l holds columns c + 128 v ..
// +3, so every load instruction still covers 512 contiguous bytes.
template <typename RIdx, int V>
__global__ void __launch_bounds__(256, V == 2 ? 4 : 6)
k_pna_gather_fwd(const float* __restrict__ pq, long long ld_pq, const float* __restrict__ r, long long ld_r,
                 const RIdx* __restrict__ ridx, const int* __restrict__ rowptr, const int* __restrict__ csr_src,
                 int N, int H, int relu, float* __restrict__ h, uint32_t* __restrict__ bits) {
    pdl_enter();
    const int lane = threadIdx.x & 31;
    const int chunks = (H + 128 * V - 1) / (128 * V);
    const int words = (H + 31) / 32;
    const int wpb = blockDim.x >> 5;
    const int gw = blockIdx.x * wpb + (threadIdx.x >> 5);        // fixed column chunk per warp, rows strided
    const int row_stride = (gridDim.x * wpb) / chunks;           // host guarantees divisibility
    const int c = (gw % chunks) * 128 * V + lane * 4;
    bool act[V];
#pragma unroll
    for (int v = 0; v < V; ++v) act[v] = c + 128 * v < H;
    for (int i = gw / chunks; i < N; i += row_stride) {
        const int b = rowptr[i], e = rowptr[i + 1];
        if (b == e) continue;
        F4 p[V];
#pragma unroll
        for (int v = 0; v < V; ++v) p[v] = act[v] ? f4_ld(pq + (long long)i * ld_pq + c + 128 * v) : f4_zero();
        for (int k = b; k < e; k += 2) {
            const bool two = k + 1 < e;
            const int s0 = csr_src[k], s1 = two ? csr_src[k + 1] : s0;
            const long long r0 = (long long)ridx[k], r1 = two ? (long long)ridx[k + 1] : r0;
            F4 q0[V], q1[V], t0[V], t1[V];
#pragma unroll
            for (int v = 0; v < V; ++v) {
                if (act[v]) {
                    q0[v] = f4_ld(pq + (long long)s0 * ld_pq + H + c + 128 * v);
                    q1[v] = f4_ld(pq + (long long)s1 * ld_pq + H + c + 128 * v);
                    t0[v] = f4_ld(r + r0 * ld_r + c + 128 * v);
                    t1[v] = f4_ld(r + r1 * ld_r + c + 128 * v);
                } else {
                    q0[v] = q1[v] = t0[v] = t1[v] = f4_zero();
                }
            }
#pragma unroll
            for (int v = 0; v < V; ++v) {
                F4 o0 = f4_zero(), o1 = f4_zero();
                if (act[v]) {
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        o0.a[j] = p[v].a[j] + q0[v].a[j] + t0[v].a[j];
                        o1.a[j] = p[v].a[j] + q1[v].a[j] + t1[v].a[j];
                        if (relu) { o0.a[j] = fmaxf(o0.a[j], 0.f); o1.a[j] = fmaxf(o1.a[j], 0.f); }
                    }
                    stg4(h + (long long)k * H + c + 128 * v, o0.q);
                    if (two) stg4(h + (long long)(k + 1) * H + c + 128 * v, o1.q);
                }
                if (bits) {                                      // packed ReLU mask of the rows just written
                    const uint32_t w0 = pack_relu_word(o0, act[v], lane), w1 = pack_relu_word(o1, act[v], lane);
                    if (act[v] && (lane & 7) == 0) {
                        bits[(long long)k * words + ((c + 128 * v) >> 5)] = w0;
                        if (two) bits[(long long)(k + 1) * words + ((c + 128 * v) >> 5)] = w1;
                    }
                }
            }
        }
    }
}

// ------------------------------------------------------------------ the same gather with the neighbour rows STAGED through
// shared memory by the bulk-copy engine (opt-in GNNX_GATHER_BULK=1; hidden 256 only: one Q[src] row = 1 KB).
// One warp per destination; lane 0 issues one cp.async.bulk per edge (up to kBulkRows per destination; longer rows
// fall back to direct loads for the rest) into the warp's stage and arms its mbarrier with the byte count, the copies
// of the NEXT destination are issued before this one is consumed (two stages per warp).  P[dst] and the 60-row
// edge-term table are read directly (one row per destination / L1-resident).  Same arithmetic, bit-identical output.
constexpr int kBulkRows = 4;
constexpr int kBulkStage = kBulkRows * 1024;                     // bytes
constexpr int kBulkSmem = 8 * 2 * kBulkStage + 8 * 2 * 8;        // 8 warps x 2 stages + 2 mbarriers per warp

__device__ __forceinline__ uint32_t sm_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

template <typename RIdx>
__global__ void __launch_bounds__(256)
k_pna_gather_fwd_bulk(const float* __restrict__ pq, long long ld_pq, const float* __restrict__ r, long long ld_r,
                      const RIdx* __restrict__ ridx, const int* __restrict__ rowptr, const int* __restrict__ csr_src,
                      int N, int relu, float* __restrict__ h, uint32_t* __restrict__ bits) {
    constexpr int H = 256;
    extern __shared__ __align__(128) uint8_t bulk_smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint8_t* my = bulk_smem + warp * 2 * kBulkStage;
    uint64_t* bars = reinterpret_cast<uint64_t*>(bulk_smem + 8 * 2 * kBulkStage) + warp * 2;
    if (lane == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(sm_u32(bars)) : "memory");
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(sm_u32(bars + 1)) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    pdl_enter();
    const int gw = blockIdx.x * 8 + warp, nw = gridDim.x * 8;
    const int c = lane * 4;
    auto issue = [&](int node, int s) {                         // lane 0 only
        const int b = rowptr[node], e = rowptr[node + 1];
        const int cnt = min(e - b, kBulkRows);
        if (cnt <= 0) return;
        const uint32_t bar = sm_u32(bars + s);
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(cnt * 1024) : "memory");
        for (int j = 0; j < cnt; ++j) {
            const float* src = pq + (long long)csr_src[b + j] * ld_pq + H;
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                             sm_u32(my + s * kBulkStage + j * 1024)),
                         "l"(src), "r"(1024), "r"(bar)
                         : "memory");
        }
    };
    uint32_t phase[2] = {0u, 0u};
    int s = 0;
    int i = gw;
    if (i < N && lane == 0) issue(i, 0);
    for (; i < N; i += nw, s ^= 1) {
        if (i + nw < N && lane == 0) issue(i + nw, s ^ 1);
        const int b = rowptr[i], e = rowptr[i + 1];
        if (b == e) continue;
        F4 p[2];
        p[0] = f4_ld(pq + (long long)i * ld_pq + c);
        p[1] = f4_ld(pq + (long long)i * ld_pq + 128 + c);
        const uint32_t bar = sm_u32(bars + s);
        uint32_t done = 0;
        while (!done)
            asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                         : "=r"(done) : "r"(bar), "r"(phase[s]) : "memory");
        phase[s] ^= 1u;
        for (int k = b; k < e; ++k) {
            const long long rr = (long long)ridx[k];
            F4 q[2], t[2];
            if (k - b < kBulkRows) {
                const float* st = reinterpret_cast<const float*>(my + s * kBulkStage + (k - b) * 1024);
                q[0].q = *reinterpret_cast<const float4*>(st + c);
                q[1].q = *reinterpret_cast<const float4*>(st + 128 + c);
            } else {
                const float* src = pq + (long long)csr_src[k] * ld_pq + H;
                q[0] = f4_ld(src + c); q[1] = f4_ld(src + 128 + c);
            }
            t[0] = f4_ld(r + rr * ld_r + c); t[1] = f4_ld(r + rr * ld_r + 128 + c);
#pragma unroll
            for (int v = 0; v < 2; ++v) {
                F4 o;
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    o.a[j] = p[v].a[j] + q[v].a[j] + t[v].a[j];
                    if (relu) o.a[j] = fmaxf(o.a[j], 0.f);
                }
                stg4(h + (long long)k * H + c + 128 * v, o.q);
                if (bits) {
                    const uint32_t w = pack_relu_word(o, true, lane);
                    if ((lane & 7) == 0) bits[(long long)k * (H / 32) + ((c + 128 * v) >> 5)] = w;
                }
            }
        }
        __syncwarp();                                           // every lane is done with stage s before it is refilled
    }
}

// dpq[:, 0:H] by CSR rows, dpq[:, H:2H] through csc_pos
__global__ void __launch_bounds__(256)
k_pna_gather_bwd(const float* __restrict__ dh, const int* __restrict__ rowptr, const int* __restrict__ colptr,
                 const int* __restrict__ csc_pos, int N, int H, float* __restrict__ dpq, long long ld) {
    pdl_enter();
    const int lane = threadIdx.x & 31;
    const int chunks = (H + 127) / 128;
    const long long nwork = (long long)N * chunks;
    const int wpb = blockDim.x >> 5;
    for (long long w = blockIdx.x * (long long)wpb + (threadIdx.x >> 5); w < nwork; w += (long long)gridDim.x * wpb) {
        int i = (int)(w / chunks);
        int c = (int)(w % chunks) * 128 + lane * 4;
        if (c >= H) continue;
        F4 a = f4_zero();
        for (int k = rowptr[i], e = rowptr[i + 1]; k < e; ++k) {
            F4 v = f4_ld(dh + (long long)k * H + c);
#pragma unroll
            for (int j = 0; j < 4; ++j) a.a[j] += v.a[j];
        }
        stg4(dpq + (long long)i * ld + c, a.q);
        F4 s = f4_zero();
        for (int k = colptr[i], e = colptr[i + 1]; k < e; ++k) {
            F4 v = f4_ld(dh + (long long)csc_pos[k] * H + c);
#pragma unroll
            for (int j = 0; j < 4; ++j) s.a[j] += v.a[j];
        }
        stg4(dpq + (long long)i * ld + H + c, s.q);
    }
}

// Same with 8 columns per lane (a warp covers 256 columns: the whole row at hidden 256): half the index loads per byte
// and twice the bytes in flight per warp -- the 4-column version is latency-bound (ncu r2a: 18.8 long-scoreboard stalls
// per issue, 51 % of the DRAM peak).  Lane l holds columns c .. c+3 and c+128 .. c+131 so every load instruction still
// covers 512 contiguous bytes.
__global__ void __launch_bounds__(256)
k_pna_gather_bwd8(const float* __restrict__ dh, const int* __restrict__ rowptr, const int* __restrict__ colptr,
                  const int* __restrict__ csc_pos, int N, int H, float* __restrict__ dpq, long long ld) {
    pdl_enter();
    const int lane = threadIdx.x & 31;
    const int chunks = (H + 255) / 256;
    const long long nwork = (long long)N * chunks;
    const int wpb = blockDim.x >> 5;
    for (long long w = blockIdx.x * (long long)wpb + (threadIdx.x >> 5); w < nwork; w += (long long)gridDim.x * wpb) {
        const int i = (int)(w / chunks);
        const int c = (int)(w % chunks) * 256 + lane * 4;
        const bool a0 = c < H, a1 = c + 128 < H;
        if (!a0) continue;
        const int kb = rowptr[i], ke = rowptr[i + 1], sb = colptr[i], se = colptr[i + 1];
        F4 p0 = f4_zero(), p1 = f4_zero(), q0 = f4_zero(), q1 = f4_zero();
        for (int k = kb; k < ke; ++k) {
            const float* row = dh + (long long)k * H + c;
            const F4 v0 = f4_ld(row), v1 = a1 ? f4_ld(row + 128) : f4_zero();
#pragma unroll
            for (int j = 0; j < 4; ++j) { p0.a[j] += v0.a[j]; p1.a[j] += v1.a[j]; }
        }
        for (int k = sb; k < se; ++k) {
            const float* row = dh + (long long)csc_pos[k] * H + c;
            const F4 v0 = f4_ld(row), v1 = a1 ? f4_ld(row + 128) : f4_zero();
#pragma unroll
            for (int j = 0; j < 4; ++j) { q0.a[j] += v0.a[j]; q1.a[j] += v1.a[j]; }
        }
        float* o = dpq + (long long)i * ld + c;
        stg4(o, p0.q);
        stg4(o + H, q0.q);
        if (a1) { stg4(o + 128, p1.q); stg4(o + H + 128, q1.q); }
    }
}

__global__ void k_permute_rows(const float* __restrict__ in, const int* __restrict__ dst_row, long long E, int H,
                               float* __restrict__ out, long long ld_out) {
    pdl_enter();
    const int h4 = H / 4;
    long long total = E * h4;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        long long p = i / h4;
        int c = (int)(i % h4) * 4;
        stg4(out + (long long)dst_row[p] * ld_out + c, ldg4(in + p * H + c));
    }
}

// grid: (ceil(H/32), nchunks), block (32, 8): each threadIdx.y owns a private [ncodes][32] table in smem
constexpr int kCodeRows = 8;
constexpr int kCodeFlight = 12;   // 12 KB in flight per SM at 4 rows was the limiter (2.5 TB/s): Little's law wants ~3x that
__global__ void k_code_sum_partial(const float* __restrict__ dh, const uint8_t* __restrict__ codes, int rows,
                                   int ncodes, int H, int rows_per_chunk, float* __restrict__ partial) {
    pdl_enter();
    extern __shared__ float tab[];   // [kCodeRows][ncodes][32]
    int c = blockIdx.x * 32 + threadIdx.x;
    float* mine = tab + (long long)threadIdx.y * ncodes * 32;
    for (int i = threadIdx.x; i < ncodes * 32; i += 32) mine[i] = 0.f;
    __syncwarp();
    int r0 = blockIdx.y * rows_per_chunk, r1 = min(rows, r0 + rows_per_chunk);
    if (c < H) {
        int r = r0 + threadIdx.y;
        for (; r + (kCodeFlight - 1) * kCodeRows < r1; r += kCodeFlight * kCodeRows) {   // rows in flight; adds stay in row order
            int cd[kCodeFlight];
            float v[kCodeFlight];
#pragma unroll
            for (int u = 0; u < kCodeFlight; ++u) {
                cd[u] = codes[r + u * kCodeRows];
                v[u] = __ldcs(dh + (long long)(r + u * kCodeRows) * H + c);
            }
#pragma unroll
            for (int u = 0; u < kCodeFlight; ++u) mine[cd[u] * 32 + threadIdx.x] += v[u];
        }
        for (; r < r1; r += kCodeRows) mine[(int)codes[r] * 32 + threadIdx.x] += __ldg(dh + (long long)r * H + c);
    }
    __syncthreads();
    if (c < H)
        for (int code = threadIdx.y; code < ncodes; code += kCodeRows) {
            float s = 0.f;
#pragma unroll
            for (int y = 0; y < kCodeRows; ++y) s += tab[((long long)y * ncodes + code) * 32 + threadIdx.x];
            partial[((long long)blockIdx.y * ncodes + code) * H + c] = s;
        }
}
__global__ void k_sum_partials(const float* __restrict__ partial, int nchunks, long long count, float* __restrict__ out) {
    pdl_enter();
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < count; i += (long long)gridDim.x * blockDim.x) {
        float s = 0.f;
        for (int ch = 0; ch < nchunks; ++ch) s += partial[(long long)ch * count + i];
        out[i] = s;
    }
}
// ------------------------------------------------------------------ the same sum over slots ORDERED BY CODE
// `order` lists the slots sorted by code (stable), code_ptr[k] = first position of code k.  The per-batch ordering is
// built once (ops.code_order) and shared by every layer.  Chunks of kCsRowsPerChunk consecutive positions never cross a
// code boundary (chunk_ptr[k] = first chunk of code k), so a warp sums the rows of ONE code into registers -- a plain
// gathered row sum with 512-byte loads, no shared-memory tables -- and writes one partial row; a second small kernel
// adds each code's partial rows in chunk order.  Deterministic: fixed row order inside a chunk, fixed chunk order.
constexpr int kCsRowsPerChunk = 256;
__global__ void __launch_bounds__(256)
k_code_sum_sorted(const float* __restrict__ dh, const int* __restrict__ order, const int* __restrict__ code_ptr,
                  const int* __restrict__ chunk_ptr, int ncodes, int H, int max_chunks, float* __restrict__ partial) {
    pdl_enter();
    // block = (chunk, 128-column slab); warp y sums rows y * 32 .. + 31 of the chunk (8 loads in flight, row order), the
    // eight warp sums are added in warp order.  One warp walking all 256 rows took 32 dependent round trips.
    __shared__ float4 sh[8][32];
    const int lane = threadIdx.x & 31, y = threadIdx.x >> 5;
    const int slabs = (H + 127) / 128;
    const int c = blockIdx.x / slabs;
    const int col = (blockIdx.x % slabs) * 128 + lane * 4;
    if (c >= __ldg(chunk_ptr + ncodes)) return;                   // chunks in use (<= max_chunks, the launch bound); block-uniform
    (void)max_chunks;
    int lo = 0, hi = ncodes;                                      // largest k with chunk_ptr[k] <= c
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (__ldg(chunk_ptr + mid) <= c) lo = mid; else hi = mid;
    }
    const int cbeg = __ldg(code_ptr + lo) + (c - __ldg(chunk_ptr + lo)) * kCsRowsPerChunk;
    const int cend = min(cbeg + kCsRowsPerChunk, __ldg(code_ptr + lo + 1));
    const int begin = cbeg + y * (kCsRowsPerChunk / 8), end = min(begin + kCsRowsPerChunk / 8, cend);
    F4 acc = f4_zero();
    if (col < H) {
        for (int p = begin; p < end; p += 8) {
            F4 v[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const int row = __ldg(order + min(p + u, end - 1));             // clamped: loads stay unconditional
                v[u] = f4_lds(dh + (long long)row * H + col);
            }
#pragma unroll
            for (int u = 0; u < 8; ++u)
                if (p + u < end) {
#pragma unroll
                    for (int j = 0; j < 4; ++j) acc.a[j] += v[u].a[j];
                }
        }
    }
    sh[y][lane] = acc.q;
    __syncthreads();
    if (y == 0 && col < H) {
        F4 s = f4_zero();
#pragma unroll
        for (int w = 0; w < 8; ++w) {
            F4 t; t.q = sh[w][lane];
#pragma unroll
            for (int j = 0; j < 4; ++j) s.a[j] += t.a[j];
        }
        stg4(partial + (long long)c * H + col, s.q);
    }
}

// dtable[code] = sum of the code's partial rows in chunk order: block = (code, 128-column slab), warp y takes chunks
// y, y + 8, ... of the code, the eight warp sums are added in warp order.
__global__ void __launch_bounds__(256)
k_code_sum_fold(const float* __restrict__ partial, const int* __restrict__ chunk_ptr, int ncodes, int H, float* __restrict__ dtable) {
    pdl_enter();
    __shared__ float4 sh[8][32];
    const int lane = threadIdx.x & 31, y = threadIdx.x >> 5;
    const int code = blockIdx.y, col = blockIdx.x * 128 + lane * 4;
    const int c0 = __ldg(chunk_ptr + code), c1 = __ldg(chunk_ptr + code + 1);
    F4 acc = f4_zero();
    if (col < H) {
        for (int c = c0 + y; c < c1; c += 64) {
            F4 v[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) v[u] = f4_ld(partial + (long long)min(c + 8 * u, c1 - 1) * H + col);
#pragma unroll
            for (int u = 0; u < 8; ++u)
                if (c + 8 * u < c1) {
#pragma unroll
                    for (int j = 0; j < 4; ++j) acc.a[j] += v[u].a[j];
                }
        }
    }
    sh[y][lane] = acc.q;
    __syncthreads();
    if (y == 0 && col < H) {
        F4 s = f4_zero();
#pragma unroll
        for (int w = 0; w < 8; ++w) {
            F4 t; t.q = sh[w][lane];
#pragma unroll
            for (int j = 0; j < 4; ++j) s.a[j] += t.a[j];
        }
        stg4(dtable + (long long)code * H + col, s.q);
    }
}

static int code_chunks(int64_t rows, int32_t H) {
    int slabs = ceil_div(H, 32);
    int target = max(1, (num_sms() * 3) / slabs);
    int by_rows = max(1, (int)((rows + 511) / 512));
    return min(target, by_rows);
}

// ------------------------------------------------------------------ fused multi-aggregator (K5)
// m: [E, H] dst-sorted, H = T*F.  agg: [N, T, 4, F].  One warp per (row, 128-col chunk).
__device__ __forceinline__ void agg_acc(const F4& v, F4& s, F4& s2, F4& mn, F4& mx) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        s.a[j] = __fadd_rn(s.a[j], v.a[j]);                       // no FMA contraction: same op
        s2.a[j] = __fadd_rn(s2.a[j], __fmul_rn(v.a[j], v.a[j]));  // sequence as scatter_add_(x*x)
        mn.a[j] = fminf(mn.a[j], v.a[j]);
        mx.a[j] = fmaxf(mx.a[j], v.a[j]);
    }
}

__global__ void __launch_bounds__(256)
k_pna_aggregate_fwd(const float* __restrict__ m, const int* __restrict__ rowptr, int N, int T, int F,
                    float* __restrict__ agg, long long ld_agg, const int* __restrict__ row_node, int rows,
                    const float* __restrict__ x) {
    pdl_enter();
    // row_node == null: output row i = node i.  Otherwise the kernel walks the `rows` OUTPUT rows (degree-bucket order):
    // row r holds node row_node[r], or is a padding row (-1) and is zero-filled; with x the node's input row is copied
    // behind the aggregates ([agg | x], the operand of the degree-grouped post Linear).
    const int H = T * F;
    const int lane = threadIdx.x & 31;
    const int chunks = (H + 127) / 128;
    const long long nwork = (long long)(row_node ? rows : N) * chunks;
    const int wpb = blockDim.x >> 5;
    for (long long w = blockIdx.x * (long long)wpb + (threadIdx.x >> 5); w < nwork; w += (long long)gridDim.x * wpb) {
        const int r = (int)(w / chunks);
        int c = (int)(w % chunks) * 128 + lane * 4;
        if (c >= H) continue;
        const int i = row_node ? row_node[r] : r;
        if (i < 0) {
            const int t = c / F, f = c % F;
            float* o = agg + (long long)r * ld_agg + (long long)t * 4 * F + f;
            const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
            stg4_stream(o, z4); stg4_stream(o + F, z4); stg4_stream(o + 2 * F, z4); stg4_stream(o + 3 * F, z4);
            if (x) stg4_stream(agg + (long long)r * ld_agg + 4 * H + c, z4);
            continue;
        }
        if (x) stg4_stream(agg + (long long)r * ld_agg + 4 * H + c, ldg4(x + (long long)i * H + c));
        int b = rowptr[i], e = rowptr[i + 1];
        F4 s = f4_zero(), s2 = f4_zero(), mn = f4_fill(INFINITY), mx = f4_fill(-INFINITY);
        int k = b;
        for (; k + 4 <= e; k += 4) {             // 4 rows in flight per lane
            F4 v0 = f4_lds(m + (long long)(k + 0) * H + c);
            F4 v1 = f4_lds(m + (long long)(k + 1) * H + c);
            F4 v2 = f4_lds(m + (long long)(k + 2) * H + c);
            F4 v3 = f4_lds(m + (long long)(k + 3) * H + c);
            agg_acc(v0, s, s2, mn, mx); agg_acc(v1, s, s2, mn, mx);
            agg_acc(v2, s, s2, mn, mx); agg_acc(v3, s, s2, mn, mx);
        }
        if (k + 2 <= e) {
            F4 v0 = f4_lds(m + (long long)(k + 0) * H + c);
            F4 v1 = f4_lds(m + (long long)(k + 1) * H + c);
            agg_acc(v0, s, s2, mn, mx); agg_acc(v1, s, s2, mn, mx);
            k += 2;
        }
        if (k < e) {
            F4 v0 = f4_lds(m + (long long)k * H + c);
            agg_acc(v0, s, s2, mn, mx);
        }
        float cnt = fmaxf((float)(e - b), 1.f);
        F4 mean, sd;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            float mu = __fdiv_rn(s.a[j], cnt);
            float m2 = __fdiv_rn(s2.a[j], cnt);
            float var = __fsub_rn(m2, __fmul_rn(mu, mu));
            float o = __fsqrt_rn(fmaxf(var, 1e-5f));
            mean.a[j] = mu;
            sd.a[j] = (o <= GNNX_STD_MASK) ? 0.f : o;
        }
        if (b == e) { mn = f4_zero(); mx = f4_zero(); }
        int t = c / F, f = c % F;
        float* o = agg + (long long)r * ld_agg + (long long)t * 4 * F + f;
        stg4_stream(o, mean.q);
        stg4_stream(o + F, mn.q);
        stg4_stream(o + 2 * F, mx.q);
        stg4_stream(o + 3 * F, sd.q);
    }
}

constexpr int kAggRegRows = 4;   // messages of one destination row kept in registers by the backward
__global__ void __launch_bounds__(256, 3)      // 80 registers; 4 blocks per SM spill and are slower (248 vs 230 us)
k_pna_aggregate_bwd(const float* __restrict__ dagg, const float* __restrict__ m, const float* __restrict__ agg,
                    const int* __restrict__ rowptr, int N, int T, int F, float* __restrict__ dm, long long ld_agg,
                    const int* __restrict__ node_pos, float* __restrict__ dx) {
    pdl_enter();
    // node_pos != null: agg / dagg rows live at node_pos[i] (degree-bucket order, row stride ld_agg); with dx the
    // gradient of the copied input row ([agg | x] operand) is moved back to node order: dx[i] = dagg[pos, 4H:5H]
    const int H = T * F;
    const int lane = threadIdx.x & 31;
    const int chunks = (H + 127) / 128;
    const long long nwork = (long long)N * chunks;
    const int wpb = blockDim.x >> 5;
    for (long long w = blockIdx.x * (long long)wpb + (threadIdx.x >> 5); w < nwork; w += (long long)gridDim.x * wpb) {
        int i = (int)(w / chunks);
        int c = (int)(w % chunks) * 128 + lane * 4;
        if (c >= H) continue;
        int b = rowptr[i], e = rowptr[i + 1];
        const long long row = node_pos ? node_pos[i] : i;
        if (dx) stg4(dx + (long long)i * H + c, ldg4_stream(dagg + row * ld_agg + 4 * H + c));
        if (b == e) continue;
        int t = c / F, f = c % F;
        long long base = row * ld_agg + (long long)t * 4 * F + f;
        F4 gmean = f4_lds(dagg + base), gmn = f4_lds(dagg + base + F), gmx = f4_lds(dagg + base + 2 * F), gsd = f4_lds(dagg + base + 3 * F);
        // The four aggregates are RECOMPUTED from the messages with the forward kernel's operation sequence (agg_acc in
        // slot order, the same rounding calls), so they are bit-identical to what the forward stored and [N, 4H] is never
        // read back.  Rows of up to kAggRegRows messages (every molecular graph) are held in registers: one set of loads,
        // all in flight together, serves the statistics, the tie counts and the gradient.
        const int deg = e - b;
        F4 vr[kAggRegRows];
        const bool in_regs = deg <= kAggRegRows;
        if (in_regs) {
#pragma unroll
            for (int u = 0; u < kAggRegRows; ++u)
                if (u < deg) vr[u] = f4_ld(m + (long long)(b + u) * H + c);
        }
        F4 s = f4_zero(), s2 = f4_zero(), mn = f4_fill(INFINITY), mx = f4_fill(-INFINITY);
        if (in_regs) {
#pragma unroll
            for (int u = 0; u < kAggRegRows; ++u)
                if (u < deg) agg_acc(vr[u], s, s2, mn, mx);
        } else {
            for (int k = b; k < e; ++k) agg_acc(f4_ld(m + (long long)k * H + c), s, s2, mn, mx);
        }
        const float cnt = (float)deg;
        F4 mean, sd;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float mu = __fdiv_rn(s.a[j], cnt);
            const float m2 = __fdiv_rn(s2.a[j], cnt);
            const float var = __fsub_rn(m2, __fmul_rn(mu, mu));
            const float o = __fsqrt_rn(fmaxf(var, 1e-5f));
            mean.a[j] = mu;
            sd.a[j] = (o <= GNNX_STD_MASK) ? 0.f : o;
        }
        // tie counts for min / max (PyG CPU semantics: gradient split evenly among ties)
        F4 nmn = f4_zero(), nmx = f4_zero();
        if (in_regs) {
#pragma unroll
            for (int u = 0; u < kAggRegRows; ++u)
                if (u < deg) {
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        nmn.a[j] += (vr[u].a[j] == mn.a[j]) ? 1.f : 0.f;
                        nmx.a[j] += (vr[u].a[j] == mx.a[j]) ? 1.f : 0.f;
                    }
                }
        } else {
            for (int k = b; k < e; ++k) {
                F4 v = f4_ld(m + (long long)k * H + c);
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    nmn.a[j] += (v.a[j] == mn.a[j]) ? 1.f : 0.f;
                    nmx.a[j] += (v.a[j] == mx.a[j]) ? 1.f : 0.f;
                }
            }
        }
        // the kernel is issue-bound (ncu: 58 % issue-active, 3.9 long-scoreboard stalls per issue): the coefficient
        // divisions use the 2-ulp fast divide -- the tie counts are small integers and the gradients are held to
        // rel 1e-4 --, while everything that DECIDES a branch (min / max equality, the Std mask) stays exact above
        float inv_cnt = __fdividef(1.f, cnt);
        F4 cmean, cmn, cmx, csd;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            cmean.a[j] = gmean.a[j] * inv_cnt;
            cmn.a[j] = __fdividef(gmn.a[j], nmn.a[j]);
            cmx.a[j] = __fdividef(gmx.a[j], nmx.a[j]);
            csd.a[j] = sd.a[j] > 0.f ? __fdividef(gsd.a[j] * inv_cnt, sd.a[j]) : 0.f;   // d std / d m_e = (m_e - mean) / (cnt * std)
        }
        auto grad_row = [&](const F4& v, int k) {
            F4 g;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                float x = v.a[j];
                g.a[j] = cmean.a[j] + ((x == mn.a[j]) ? cmn.a[j] : 0.f) + ((x == mx.a[j]) ? cmx.a[j] : 0.f) +
                         csd.a[j] * (x - mean.a[j]);
            }
            stg4(dm + (long long)k * H + c, g.q);
        };
        if (in_regs) {
#pragma unroll
            for (int u = 0; u < kAggRegRows; ++u)
                if (u < deg) grad_row(vr[u], b + u);
        } else {
            for (int k = b; k < e; ++k) grad_row(f4_ld(m + (long long)k * H + c), k);   // L1 / L2-resident from the passes above
        }
    }
}

// ------------------------------------------------------------------ first post_nn Linear, scalers folded (K5 tail + K6 head)
__device__ __forceinline__ void pna_scalers(const int* __restrict__ rowptr, int row, float avg_deg_log, float& s_amp, float& s_att) {
    float d = (float)(rowptr[row + 1] - rowptr[row]);
    s_amp = logf(d + 1.f) / avg_deg_log;
    s_att = avg_deg_log / logf(fmaxf(d, 1.f) + 1.f);
}

__global__ void k_pna_post_combine_fwd(const float* __restrict__ z, const float* __restrict__ bias,
                                       const int* __restrict__ rowptr, float avg_deg_log, long long N, int H, int relu,
                                       float* __restrict__ out) {
    pdl_enter();
    const int h4 = H / 4;
    long long total = N * h4;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        int row = (int)(i / h4);
        int c = (int)(i % h4) * 4;
        float sa, st;
        pna_scalers(rowptr, row, avg_deg_log, sa, st);
        const float* zr = z + (long long)row * 3 * H + c;
        F4 zi = f4_lds(zr), za = f4_lds(zr + H), zt = f4_lds(zr + 2 * H);
        F4 bb = bias ? f4_ld(bias + c) : f4_zero();
        F4 o;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            float v = zi.a[j] + sa * za.a[j] + st * zt.a[j] + bb.a[j];
            o.a[j] = relu ? fmaxf(v, 0.f) : v;
        }
        stg4(out + (long long)row * H + c, o.q);
    }
}

__global__ void k_pna_post_combine_bwd(const float* __restrict__ dout, const float* __restrict__ out,
                                       const int* __restrict__ rowptr, float avg_deg_log, long long N, int H, int relu,
                                       float* __restrict__ dz) {
    pdl_enter();
    const int h4 = H / 4;
    long long total = N * h4;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        int row = (int)(i / h4);
        int c = (int)(i % h4) * 4;
        float sa, st;
        pna_scalers(rowptr, row, avg_deg_log, sa, st);
        F4 g = f4_lds(dout + (long long)row * H + c);
        if (relu) {
            F4 o = f4_lds(out + (long long)row * H + c);
#pragma unroll
            for (int j = 0; j < 4; ++j) g.a[j] = o.a[j] > 0.f ? g.a[j] : 0.f;
        }
        F4 ga, gt;
#pragma unroll
        for (int j = 0; j < 4; ++j) { ga.a[j] = sa * g.a[j]; gt.a[j] = st * g.a[j]; }
        float* zr = dz + (long long)row * 3 * H + c;
        stg4(zr, g.q);
        stg4(zr + H, ga.q);
        stg4(zr + 2 * H, gt.q);
    }
}

__global__ void k_relu_bwd(const float* __restrict__ dy, const float* __restrict__ y, long long count, float* __restrict__ dx) {
    pdl_enter();
    long long n4 = count / 4;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
        F4 g = f4_lds(dy + i * 4), o = f4_lds(y + i * 4);
#pragma unroll
        for (int j = 0; j < 4; ++j) g.a[j] = o.a[j] > 0.f ? g.a[j] : 0.f;
        stg4(dx + i * 4, g.q);
    }
    if (blockIdx.x == 0 && threadIdx.x < (count & 3)) {
        long long i = n4 * 4 + threadIdx.x;
        dx[i] = y[i] > 0.f ? dy[i] : 0.f;
    }
}

static inline int warp_grid(long long nwork) {   // 8 warps per block, grid a multiple of the SM count
    long long blocks = (nwork + 7) / 8;
    long long cap = (long long)num_sms() * 32;
    return (int)(blocks < cap ? (blocks > 0 ? blocks : 1) : cap);
}
static inline int flat_grid(long long nthreads) {
    long long blocks = (nthreads + 255) / 256;
    long long cap = (long long)num_sms() * 32;
    return (int)(blocks < cap ? (blocks > 0 ? blocks : 1) : cap);
}

}  // namespace gnnx

using namespace gnnx;

extern "C" {

int gnnx_pna_gather_fwd(const float* pq, int64_t ld_pq, const float* r, int64_t ld_r, const int32_t* ridx_i32,
                        const uint8_t* ridx_u8, const int32_t* rowptr, const int32_t* csr_src, int64_t N, int64_t E,
                        int32_t H, int32_t relu, float* h, uint32_t* relu_bits_out, void* stream) {
    GNNX_REQUIRE(H % 4 == 0 && ld_pq % 4 == 0 && ld_r % 4 == 0, "pna_gather_fwd: hidden / strides must be multiples of 4");
    GNNX_REQUIRE(E == 0 || (ridx_i32 != nullptr) != (ridx_u8 != nullptr), "pna_gather_fwd: exactly one of ridx_i32 / ridx_u8");
    if (N == 0 || E == 0) return GNNX_OK;
    static int wide = -1;
    if (wide < 0) { const char* e = getenv("GNNX_GATHER8"); wide = (e && e[0] >= '0' && e[0] <= '2') ? e[0] - '0' : 1; }
    // 8 columns per lane pay in the backward (96.5 -> 79.5 us) but not here: three operand rows per edge at V = 2 need 64
    // registers, the lost occupancy eats the gain (94.5 us against 92.7).  GNNX_GATHER8=2 selects it all the same.
    {   // opt-in: neighbour rows staged through shared memory by cp.async.bulk (measured)
        const char* be = getenv("GNNX_GATHER_BULK");
        if (be && be[0] == '1' && H == 256 && ((uintptr_t)pq & 15) == 0) {
            static std::atomic<int> attr{0};
            if (!attr.load()) {
                GNNX_CUDA(cudaFuncSetAttribute(k_pna_gather_fwd_bulk<int>, cudaFuncAttributeMaxDynamicSharedMemorySize, kBulkSmem));
                GNNX_CUDA(cudaFuncSetAttribute(k_pna_gather_fwd_bulk<uint8_t>, cudaFuncAttributeMaxDynamicSharedMemorySize, kBulkSmem));
                attr.store(1);
            }
            int g = (int)std::min<long long>((N + 7) / 8, (long long)num_sms() * 3);
            if (ridx_i32) launch(k_pna_gather_fwd_bulk<int>, g, 256, kBulkSmem, as_stream(stream), pq, ld_pq, r, ld_r, ridx_i32, rowptr, csr_src, (int)N, relu, h, relu_bits_out);
            else launch(k_pna_gather_fwd_bulk<uint8_t>, g, 256, kBulkSmem, as_stream(stream), pq, ld_pq, r, ld_r, ridx_u8, rowptr, csr_src, (int)N, relu, h, relu_bits_out);
            return check_launch("pna_gather_fwd(bulk)");
        }
    }
    const int V = (wide == 2 && H > 128) ? 2 : 1;
    const int chunks = (H + 128 * V - 1) / (128 * V);
    int grid = warp_grid((long long)N * chunks);
    grid = (grid + chunks - 1) / chunks * chunks;               // every warp keeps one column chunk
#define GATHER_FWD(T, VV, IDX) launch(k_pna_gather_fwd<T, VV>, grid, 256, 0, as_stream(stream), pq, ld_pq, r, ld_r, IDX, rowptr, csr_src, (int)N, H, relu, h, relu_bits_out)
    if (ridx_i32) { if (V == 2) GATHER_FWD(int, 2, ridx_i32); else GATHER_FWD(int, 1, ridx_i32); }
    else { if (V == 2) GATHER_FWD(uint8_t, 2, ridx_u8); else GATHER_FWD(uint8_t, 1, ridx_u8); }
#undef GATHER_FWD
    return check_launch("pna_gather_fwd");
}

int gnnx_pna_gather_bwd(const float* dh, const int32_t* rowptr, const int32_t* colptr, const int32_t* csc_pos,
                        const int32_t* csr_eid, int64_t N, int64_t E, int32_t H, float* dpq, int64_t ld_dpq,
                        float* dr_edge, int64_t ld_dr, void* stream) {
    GNNX_REQUIRE(H % 4 == 0 && ld_dpq % 4 == 0, "pna_gather_bwd: hidden / strides must be multiples of 4");
    if (N == 0) return GNNX_OK;
    static int wide = -1;
    if (wide < 0) { const char* e = getenv("GNNX_GATHER8"); wide = (e && e[0] == '0') ? 0 : 1; }
    if (wide && H > 128) {
        long long nwork = (long long)N * ((H + 255) / 256);
        launch(k_pna_gather_bwd8, warp_grid(nwork), 256, 0, as_stream(stream), dh, rowptr, colptr, csc_pos, (int)N, H, dpq, ld_dpq);
    } else {
        long long nwork = (long long)N * ((H + 127) / 128);
        launch(k_pna_gather_bwd, warp_grid(nwork), 256, 0, as_stream(stream), dh, rowptr, colptr, csc_pos, (int)N, H, dpq, ld_dpq);
    }
    if (dr_edge && E > 0) {
        GNNX_REQUIRE(ld_dr % 4 == 0 && csr_eid, "pna_gather_bwd: dr_edge needs csr_eid and an aligned stride");
        launch(k_permute_rows, flat_grid(E * (H / 4)), 256, 0, as_stream(stream), dh, csr_eid, E, H, dr_edge, ld_dr);
    }
    return check_launch("pna_gather_bwd", (dr_edge && E > 0) ? 2 : 1);
}

size_t gnnx_code_segment_sum_workspace_bytes(int64_t rows, int32_t num_codes, int32_t H) {
    return (size_t)code_chunks(rows, H) * num_codes * H * sizeof(float);
}

int gnnx_code_segment_sum(const float* dh, const uint8_t* codes, int64_t rows, int32_t ncodes, int32_t H, float* dtable,
                          void* ws, size_t ws_bytes, void* stream) {
    GNNX_REQUIRE(ncodes >= 1 && ncodes <= 256, "code_segment_sum: num_codes must be in [1,256]");
    GNNX_REQUIRE(ws_bytes >= gnnx_code_segment_sum_workspace_bytes(rows, ncodes, H), "code_segment_sum: workspace too small");
    cudaStream_t st = as_stream(stream);
    long long count = (long long)ncodes * H;
    if (rows == 0) { GNNX_CUDA(cudaMemsetAsync(dtable, 0, count * sizeof(float), st)); return GNNX_OK; }
    int nch = code_chunks(rows, H);
    int rpc = ceil_div(rows, nch);
    size_t smem = (size_t)kCodeRows * ncodes * 32 * sizeof(float);
    GNNX_REQUIRE(smem <= 200 * 1024, "code_segment_sum: too many codes for the shared-memory tables");
    if (smem > 48 * 1024)
        GNNX_CUDA(cudaFuncSetAttribute(k_code_sum_partial, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    launch(k_code_sum_partial, dim3(ceil_div(H, 32), nch), dim3(32, kCodeRows), smem, st, dh, codes, (int)rows, ncodes, H, rpc, (float*)ws);
    launch(k_sum_partials, flat_grid(count), 256, 0, st, (const float*)ws, nch, count, dtable);
    return check_launch("code_segment_sum", 2);
}

size_t gnnx_code_segment_sum_sorted_workspace_bytes(int64_t rows, int32_t num_codes, int32_t H) {
    return (size_t)(rows / kCsRowsPerChunk + num_codes + 1) * H * sizeof(float);
}

int gnnx_code_segment_sum_sorted(const float* dh, const int32_t* order, const int32_t* code_ptr, const int32_t* chunk_ptr,
                                 int64_t rows, int32_t ncodes, int32_t H, float* dtable, void* ws, size_t ws_bytes, void* stream) {
    GNNX_REQUIRE(ncodes >= 1 && ncodes <= 65535, "code_segment_sum_sorted: num_codes must be in [1, 65535]");
    GNNX_REQUIRE(H % 4 == 0 && ((uintptr_t)dh & 15) == 0, "code_segment_sum_sorted: hidden must be a multiple of 4 and dh 16-byte aligned");
    GNNX_REQUIRE(ws_bytes >= gnnx_code_segment_sum_sorted_workspace_bytes(rows, ncodes, H), "code_segment_sum_sorted: workspace too small");
    cudaStream_t st = as_stream(stream);
    if (rows == 0) { GNNX_CUDA(cudaMemsetAsync(dtable, 0, (size_t)ncodes * H * sizeof(float), st)); return GNNX_OK; }
    const int max_chunks = (int)(rows / kCsRowsPerChunk + ncodes);          // sum_k ceil(n_k / R) <= rows / R + ncodes
    const int slabs = ceil_div(H, 128);
    launch(k_code_sum_sorted, max_chunks * slabs, 256, 0, st, dh, order, code_ptr, chunk_ptr, ncodes, H, max_chunks, (float*)ws);
    launch(k_code_sum_fold, dim3(slabs, ncodes), 256, 0, st, (const float*)ws, chunk_ptr, ncodes, H, dtable);
    return check_launch("code_segment_sum_sorted", 2);
}

int gnnx_pna_aggregate_fwd(const float* m, const int32_t* rowptr, int64_t N, int32_t T, int32_t F, float* agg, void* stream) {
    GNNX_REQUIRE(T >= 1 && F >= 4 && F % 4 == 0, "pna_aggregate_fwd: F must be a positive multiple of 4");
    if (N == 0) return GNNX_OK;
    long long nwork = (long long)N * ((T * F + 127) / 128);
    launch(k_pna_aggregate_fwd, warp_grid(nwork), 256, 0, as_stream(stream), m, rowptr, (int)N, T, F, agg, 4ll * T * F, nullptr, 0, nullptr);
    return check_launch("pna_aggregate_fwd");
}

int gnnx_pna_aggregate_rows_fwd(const float* m, const int32_t* rowptr, int64_t N, int32_t T, int32_t F, const int32_t* row_node,
                                int64_t rows, const float* x, float* out, int64_t ld_out, void* stream) {
    GNNX_REQUIRE(T >= 1 && F >= 4 && F % 4 == 0, "pna_aggregate_rows_fwd: F must be a positive multiple of 4");
    GNNX_REQUIRE(row_node && ld_out % 4 == 0 && ld_out >= (x ? 5 : 4) * (int64_t)T * F, "pna_aggregate_rows_fwd: row map / stride");
    if (rows == 0) return GNNX_OK;
    long long nwork = (long long)rows * ((T * F + 127) / 128);
    launch(k_pna_aggregate_fwd, warp_grid(nwork), 256, 0, as_stream(stream), m, rowptr, (int)N, T, F, out, ld_out, row_node, (int)rows, x);
    return check_launch("pna_aggregate_rows_fwd");
}

int gnnx_pna_aggregate_bwd(const float* dagg, const float* m, const float* agg, const int32_t* rowptr, int64_t N,
                           int32_t T, int32_t F, float* dm, void* stream) {
    GNNX_REQUIRE(T >= 1 && F >= 4 && F % 4 == 0, "pna_aggregate_bwd: F must be a positive multiple of 4");
    if (N == 0) return GNNX_OK;
    long long nwork = (long long)N * ((T * F + 127) / 128);
    launch(k_pna_aggregate_bwd, warp_grid(nwork), 256, 0, as_stream(stream), dagg, m, agg, rowptr, (int)N, T, F, dm, 4ll * T * F, nullptr, nullptr);
    return check_launch("pna_aggregate_bwd");
}

int gnnx_pna_aggregate_rows_bwd(const float* dagg, const float* m, const float* agg, int64_t ld, const int32_t* node_pos,
                                const int32_t* rowptr, int64_t N, int32_t T, int32_t F, float* dm, float* dx, void* stream) {
    GNNX_REQUIRE(T >= 1 && F >= 4 && F % 4 == 0, "pna_aggregate_rows_bwd: F must be a positive multiple of 4");
    GNNX_REQUIRE(node_pos && ld % 4 == 0 && ld >= (dx ? 5 : 4) * (int64_t)T * F, "pna_aggregate_rows_bwd: row map / stride");
    if (N == 0) return GNNX_OK;
    long long nwork = (long long)N * ((T * F + 127) / 128);
    launch(k_pna_aggregate_bwd, warp_grid(nwork), 256, 0, as_stream(stream), dagg, m, agg, rowptr, (int)N, T, F, dm, ld, node_pos, dx);
    return check_launch("pna_aggregate_rows_bwd");
}

int gnnx_pna_post_combine_fwd(const float* z, const float* bias, const int32_t* rowptr, float avg_deg_log, int64_t N,
                              int32_t H, int32_t relu, float* out, void* stream) {
    GNNX_REQUIRE(H % 4 == 0, "pna_post_combine_fwd: h_out must be a multiple of 4");
    if (N == 0) return GNNX_OK;
    launch(k_pna_post_combine_fwd, flat_grid(N * (H / 4)), 256, 0, as_stream(stream), z, bias, rowptr, avg_deg_log, N, H, relu, out);
    return check_launch("pna_post_combine_fwd");
}

int gnnx_pna_post_combine_bwd(const float* dout, const float* out, const int32_t* rowptr, float avg_deg_log, int64_t N,
                              int32_t H, int32_t relu, float* dz, void* stream) {
    GNNX_REQUIRE(H % 4 == 0, "pna_post_combine_bwd: h_out must be a multiple of 4");
    if (N == 0) return GNNX_OK;
    launch(k_pna_post_combine_bwd, flat_grid(N * (H / 4)), 256, 0, as_stream(stream), dout, out, rowptr, avg_deg_log, N, H, relu, dz);
    return check_launch("pna_post_combine_bwd");
}

int gnnx_relu_bwd(const float* dy, const float* y, int64_t count, float* dx, void* stream) {
    if (count == 0) return GNNX_OK;
    GNNX_REQUIRE(((uintptr_t)dy % 16 == 0) && ((uintptr_t)y % 16 == 0) && ((uintptr_t)dx % 16 == 0), "relu_bwd: pointers must be 16-byte aligned");
    launch(k_relu_bwd, flat_grid(count / 4 + 1), 256, 0, as_stream(stream), dy, y, count, dx);
    return check_launch("relu_bwd");
}

}  // extern "C"
