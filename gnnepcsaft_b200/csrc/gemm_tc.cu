// tcgen05 + TMEM + TMA GEMMs (kind::tf32, fp32 in HBM, fp32 accumulate) for the dense transforms of the
// hot path: the Linear layers inside PNAConv / GATv2Conv / the MLP head and their dX / dW products.
//
// Two kernel families, both warp-specialised (warp 0 = TMA producer, warp 1 = MMA issuer + TMEM owner, warps 2-5 =
// epilogue, one per TMEM lane quadrant), fed by 128-byte-swizzled TMA boxes straight from the fp32
// tensors (no conversion pass: the tensor map's TFLOAT32 type rounds on load) through an mbarrier ring:
//
//   k_gemm_tf32_rows2  C[M,N] = A[M,K] * op(B)      A K-major.  op(B) = B^T with B [N,K] (Linear forward,
//                      K-major) or B [K,N] (dX = dY W, MN-major B descriptor).  One CTA owns a 128-row tile and ONE
//                      128 x BN accumulator, two accumulators ping-pong in TMEM; 2-CTA clusters multicast the weights.
//   k_gemm_tf32_red    C[Mo,No] = A[R,Mo]^T * B[R,No]  (dW = dY^T X): both operands MN-major, the reduction
//                      runs over the R rows; grid = output tiles x row splits, partials reduced by a
//                      deterministic second pass (no atomics).
//
// At the hot path's shapes (K = 256..1024, N = 256..768, M = 1e5..2e5 rows of fp32) these products are
// HBM/L2-bound, not tensor-bound: see DESIGN.md for the roofline of each.
#include <cudaTypedefs.h>

#include <string.h>

#include <atomic>
#include <map>
#include <mutex>
#include <unordered_map>
#include <utility>
#include <type_traits>

#include "common.cuh"
#include "tc_ptx.cuh"

namespace gnnx {

int splitk_reduce_launch(const float* partial, int splits, int M, int N, float* C, long long ldc, const float* bias,
                         int flags, const float* mask, long long ldm, cudaStream_t st);
int pack_relu_bits_launch(const float* C, long long ldc, int M, int N, uint32_t* bits, cudaStream_t st);
int colsum_final_launch(const float* partial, int nchunks, int cols, float* out, cudaStream_t st);

namespace tc {

constexpr int kStages = 3;
constexpr int kThreads = 192;
constexpr int kBK = 32;                         // fp32 per k-block = one 128-byte swizzle row
constexpr int kAStage = 256 * 128;              // 32 KB: 256 rows x 128 B (rows kernel) / 8 chunks x 32 k x 128 B (red kernel)
constexpr int kBStage = 256 * 128;              // 32 KB
constexpr int kStageBytes = kAStage + kBStage;  // 64 KB
constexpr int kEpiPitch = 36;                   // floats; 144-byte rows keep float4 staging conflict-free (register-store path)
constexpr int kEpiBytes = 4 * 2 * 4096;         // per epilogue warp: two 32x32 fp32 TMA-store staging tiles (1024-B aligned)
constexpr int kSmemBytes = 1024 + kStages * kStageBytes + kEpiBytes + 128;
constexpr int kTmemCols = 512;

struct Smem {
    static constexpr int kBufs = 2;                 // TMA-store staging tiles per epilogue warp
    uint8_t* base;      // 1024-aligned
    __device__ uint32_t a(int s) const { return smem_u32(base + s * kStageBytes); }
    __device__ uint32_t b(int s) const { return smem_u32(base + s * kStageBytes + kAStage); }
    __device__ float* epi(int q) const { return reinterpret_cast<float*>(base + kStages * kStageBytes + q * 8192); }
    __device__ uint64_t* bars() const { return reinterpret_cast<uint64_t*>(base + kStages * kStageBytes + kEpiBytes); }
    __device__ uint32_t full(int s) const { return smem_u32(bars() + s); }
    __device__ uint32_t empty(int s) const { return smem_u32(bars() + kStages + s); }
    __device__ uint32_t tmem_full() const { return smem_u32(bars() + 2 * kStages); }
    __device__ uint32_t tmem_empty() const { return smem_u32(bars() + 2 * kStages + 1); }
    __device__ uint32_t* tmem_ptr() const { return reinterpret_cast<uint32_t*>(bars() + 2 * kStages + 2); }
};

__device__ __forceinline__ Smem carve_smem(uint8_t* raw) {
    Smem s;
    s.base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~uintptr_t(1023));
    return s;
}

// Common prologue: barriers, TMEM allocation (all 512 columns: one CTA per SM by shared-memory footprint).
__device__ __forceinline__ uint32_t prologue(const Smem& sm, int warp) {
    if (threadIdx.x == 0) {
        for (int s = 0; s < kStages; ++s) {
            mbar_init(sm.full(s), 1);
            mbar_init(sm.empty(s), 1);
        }
        mbar_init(sm.tmem_full(), 1);
        mbar_init(sm.tmem_empty(), 4);
        mbar_fence_init();
    }
    if (warp == 1) tmem_alloc(smem_u32(sm.tmem_ptr()), kTmemCols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    return *sm.tmem_ptr();
}

// Epilogue of one 128-row accumulator: TMEM -> registers -> per-warp smem transpose -> coalesced float4 rows.
// Thread t of quadrant q holds accumulator row q*32+t; after the transpose 8 lanes cover one 128-byte row segment.
template <class SM>
__device__ __forceinline__ void epilogue_acc(const SM& sm, uint32_t tmem_acc, int q, int lane, float* __restrict__ C,
                                            long long ldc, long long row0, long long row_end, int col0, int col_end,
                                            int ncols, const float* __restrict__ bias, int flags,
                                            const float* __restrict__ mask, long long ldm) {
    float* st = sm.epi(q);
    const int cc = (lane & 7) * 4;
    const bool acc = (flags & GNNX_GEMM_ACCUMULATE) != 0;
    for (int c = 0; c < ncols; c += 32) {
        const int gcol = col0 + c + cc;
        const bool col_ok = gcol < col_end;
        // previous C (accumulate) and the ReLU mask rows are fetched BEFORE the TMEM load so that the 8 independent
        // global loads of this chunk overlap it (a load issued after a store to C cannot be hoisted by the compiler).
        float4 prev[8], mk[8];
        if (acc && col_ok) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const long long grow = row0 + q * 32 + i * 4 + (lane >> 3);
                prev[i] = grow < row_end ? *reinterpret_cast<const float4*>(C + grow * ldc + gcol) : make_float4(0.f, 0.f, 0.f, 0.f);
            }
        }
        if (mask && col_ok) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const long long grow = row0 + q * 32 + i * 4 + (lane >> 3);
                if (grow >= row_end) { mk[i] = make_float4(0.f, 0.f, 0.f, 0.f); continue; }
                if (flags & GNNX_GEMM_MASK_BITS) {       // expand this lane's 4 bits to the fp32 form used below
                    const uint32_t w = __ldg(reinterpret_cast<const uint32_t*>(mask) + grow * ldm + (gcol >> 5)) >> (gcol & 31);
                    mk[i] = make_float4((float)(w & 1u), (float)((w >> 1) & 1u), (float)((w >> 2) & 1u), (float)((w >> 3) & 1u));
                } else {
                    mk[i] = __ldcs(reinterpret_cast<const float4*>(mask + grow * ldm + gcol));
                }
            }
        }
        float v[32];
        tmem_ld32(tmem_acc + ((uint32_t)(q * 32) << 16) + (uint32_t)c, v);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 8; ++j)
            *reinterpret_cast<float4*>(st + lane * kEpiPitch + 4 * j) = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
        __syncwarp();
        if (col_ok) {
            float4 bv = make_float4(0.f, 0.f, 0.f, 0.f);
            if (bias) bv = __ldg(reinterpret_cast<const float4*>(bias + gcol));
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int r = i * 4 + (lane >> 3);
                const long long grow = row0 + q * 32 + r;
                if (grow < row_end) {
                    float4 o = *reinterpret_cast<const float4*>(st + r * kEpiPitch + cc);
                    o.x += bv.x; o.y += bv.y; o.z += bv.z; o.w += bv.w;
                    if (acc) { o.x += prev[i].x; o.y += prev[i].y; o.z += prev[i].z; o.w += prev[i].w; }
                    if (mask) {
                        o.x = mk[i].x > 0.f ? o.x : 0.f; o.y = mk[i].y > 0.f ? o.y : 0.f;
                        o.z = mk[i].z > 0.f ? o.z : 0.f; o.w = mk[i].w > 0.f ? o.w : 0.f;
                    }
                    if (flags & GNNX_GEMM_RELU) {
                        o.x = fmaxf(o.x, 0.f); o.y = fmaxf(o.y, 0.f); o.z = fmaxf(o.z, 0.f); o.w = fmaxf(o.w, 0.f);
                    }
                    *reinterpret_cast<float4*>(C + grow * ldc + gcol) = o;
                }
            }
        }
        __syncwarp();
    }
}

// Epilogue of one 128-row accumulator through TMA stores: TMEM -> registers (+bias, ReLU) -> 128B-swizzled 32x32 smem
// tile -> cp.async.bulk.tensor store (or reduce-add when accumulating), two tiles in flight per warp.  The bulk
// copy engine moves whole 128-byte rows, so four epilogue warps are enough to keep the SM's store path full; rows
// and columns beyond the tensor bounds are clipped by the tensor map.  `it` counts this warp's staged tiles.
template <class SM>
__device__ __forceinline__ void epilogue_tma(const SM& sm, uint32_t tmem_acc, int q, int lane, const CUtensorMap* tmC,
                                            int row0, int row_end, int col0, int ncols, const float* __restrict__ bias,
                                            int col_end, int flags, uint32_t& it, const float* __restrict__ mask = nullptr,
                                            long long ldm = 0, uint32_t* __restrict__ bits_out = nullptr, long long ldbits = 0,
                                            int c_first = 0, int c_stride = 32, uint8_t* stage1 = nullptr) {
    // stage1 != null: this warp shares its TMEM lane quadrant with a second epilogue warp (each takes every other
    // 32-column chunk: c_first / c_stride) and has ONE 4 KB staging tile of its own instead of two
    uint8_t* stage = stage1 ? stage1 : reinterpret_cast<uint8_t*>(sm.epi(q));
    const int mr = lane >> 3, mj = lane & 7;            // coalesced mask mapping: 8 lanes cover one 128-byte row segment
    const bool mbits = mask && (flags & GNNX_GEMM_MASK_BITS);
    const long long myrow = (long long)row0 + q * 32 + lane;   // this thread's accumulator row
    for (int c = c_first; c < ncols; c += c_stride, ++it) {
        // ReLU mask of the layer below, requested before the TMEM load.  Packed: ONE word per thread covers its row's
        // 32 columns of this chunk.  fp32: 4 full 128-byte rows per load instruction, applied on the staged tile.
        float4 mk[8];
        uint32_t mw = 0xffffffffu;
        if (mbits) {
            if (myrow < row_end && col0 + c < col_end)
                mw = __ldg(reinterpret_cast<const uint32_t*>(mask) + myrow * ldm + ((col0 + c) >> 5));
        } else if (mask) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const long long gr = (long long)row0 + q * 32 + i * 4 + mr;
                mk[i] = (gr < row_end && col0 + c + 4 * mj < col_end)
                            ? __ldcs(reinterpret_cast<const float4*>(mask + gr * ldm + col0 + c + 4 * mj))
                            : make_float4(0.f, 0.f, 0.f, 0.f);
            }
        }
        float4 bv[8];
        if (bias) {
#pragma unroll
            for (int j = 0; j < 8; ++j)
                bv[j] = (col0 + c + 4 * j < col_end) ? __ldg(reinterpret_cast<const float4*>(bias + col0 + c + 4 * j))
                                                     : make_float4(0.f, 0.f, 0.f, 0.f);
        }
        float v[32];
        tmem_ld32(tmem_acc + ((uint32_t)(q * 32) << 16) + (uint32_t)c, v);
        uint8_t* buf = stage1 ? stage : stage + (it % SM::kBufs) * 4096;
        if (lane == 0) { if (stage1) tma_store_wait_read<0>(); else tma_store_wait_read<SM::kBufs - 1>(); }   // the store that last read this buffer has drained it
        __syncwarp();
        tmem_ld_wait();
        if (bias) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                v[4 * j] += bv[j].x; v[4 * j + 1] += bv[j].y; v[4 * j + 2] += bv[j].z; v[4 * j + 3] += bv[j].w;
            }
        }
        if (flags & GNNX_GEMM_RELU) {
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j], 0.f);
        }
        if (mbits) {
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = ((mw >> j) & 1u) ? v[j] : 0.f;
        }
        if (bits_out) {                                  // packed (C > 0) of this row's 32 columns, for the backward
            uint32_t w = 0;
#pragma unroll
            for (int j = 0; j < 32; ++j) w |= (v[j] > 0.f ? 1u : 0u) << j;
            if (myrow < row_end && col0 + c < col_end) bits_out[myrow * ldbits + ((col0 + c) >> 5)] = w;
        }
#pragma unroll
        for (int j = 0; j < 8; ++j)
            *reinterpret_cast<float4*>(buf + lane * 128 + ((j ^ (lane & 7)) << 4)) = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
        if (mask && !mbits) {                            // second pass over the staged tile in the mask's (coalesced) mapping
            __syncwarp();
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int r = i * 4 + mr;
                float4* p = reinterpret_cast<float4*>(buf + r * 128 + ((mj ^ (r & 7)) << 4));
                float4 o = *p;
                o.x = mk[i].x > 0.f ? o.x : 0.f; o.y = mk[i].y > 0.f ? o.y : 0.f;
                o.z = mk[i].z > 0.f ? o.z : 0.f; o.w = mk[i].w > 0.f ? o.w : 0.f;
                *p = o;
            }
        }
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) {
            if (row0 + q * 32 < row_end && col0 + c < col_end)
                tma_store_2d(tmC, smem_u32(buf), col0 + c, row0 + q * 32, (flags & GNNX_GEMM_ACCUMULATE) != 0);
            tma_store_commit();
        }
    }
}

// ------------------------------------------------------------------ C[M,N] = A[M,K] * op(B)
// 128 rows per CTA and tile (ONE 128 x BN accumulator = 256 TMEM columns), so the 512 columns hold two accumulators
// and the epilogue of tile i overlaps the loads + MMAs of tile i+1; 4 stages x 48 KB.  With kCluster == 2 two CTAs
// (a thread-block cluster, adjacent row tiles, same N tile) each load HALF of every weight stage and multicast it to
// both, so the L2 traffic per output element equals that of a 256-row tile.  A stage is released to the producers of
// BOTH CTAs by a multicast tcgen05.commit (empty barriers count kCluster arrivals).
#ifndef GNNX_ROWS2_STAGES
#define GNNX_ROWS2_STAGES 4
#endif
#ifndef GNNX_ROWS2_BUFS
#define GNNX_ROWS2_BUFS 2
#endif
constexpr int kStages2 = GNNX_ROWS2_STAGES;
constexpr int kEpiBufs2 = GNNX_ROWS2_BUFS;      // TMA-store staging tiles (4 KB) per epilogue warp
constexpr int kEpiBytes2 = 4 * kEpiBufs2 * 4096;
constexpr int kA2 = 128 * 128;                  // 16 KB
constexpr int kB2 = 256 * 128;                  // 32 KB
constexpr int kStage2 = kA2 + kB2;              // 48 KB
constexpr int kSmem2 = 1024 + kStages2 * kStage2 + kEpiBytes2 + 256;
static_assert(kSmem2 <= 232448, "rows2: shared memory");
constexpr int kThreads2 = 256;                  // + warps 6, 7: column sums of the A tiles (the bias gradient riding a dX product)
constexpr int kCsWarps = 1;                     // column-sum warps per CTA (two, 64 rows each, were measured: no faster -- reading a stage is ~1 us of a 42 us product)
constexpr int kBackoffFlag = 1 << 16;             // internal launch flag: waiting roles sleep between polls
constexpr int kColsumMaxK = 256;                // columns the fused sum covers (8 k-blocks of accumulators in registers)

struct Smem2 {
    static constexpr int kBufs = kEpiBufs2;
    uint8_t* base;
    __device__ uint32_t a(int s) const { return smem_u32(base + s * kStage2); }
    __device__ uint32_t b(int s) const { return smem_u32(base + s * kStage2 + kA2); }
    __device__ float* epi(int q) const { return reinterpret_cast<float*>(base + kStages2 * kStage2 + q * kEpiBufs2 * 4096); }
    __device__ uint64_t* bars() const { return reinterpret_cast<uint64_t*>(base + kStages2 * kStage2 + kEpiBytes2); }
    __device__ uint32_t full(int s) const { return smem_u32(bars() + s); }
    __device__ uint32_t empty(int s) const { return smem_u32(bars() + kStages2 + s); }
    __device__ uint32_t tmem_full(int b) const { return smem_u32(bars() + 2 * kStages2 + b); }
    __device__ uint32_t tmem_empty(int b) const { return smem_u32(bars() + 2 * kStages2 + 2 + b); }
    __device__ uint32_t* tmem_ptr() const { return reinterpret_cast<uint32_t*>(bars() + 2 * kStages2 + 4); }
};

// The column-sum warp's tail: fold the four row lanes and write this warp's partial row.
__device__ __forceinline__ void colsum_store_partial(float4 (&acc)[kColsumMaxK / kBK], int lane, int cw, int KB, int K,
                                                    float* __restrict__ colsum_partial) {
    const int cj = lane & 7, rr = lane >> 3;
    const int my_row = (int)blockIdx.x * kCsWarps + cw;
#pragma unroll
    for (int kb = 0; kb < kColsumMaxK / kBK; ++kb) {
        float4 v = acc[kb];                                   // fold the four row lanes (fixed order)
#pragma unroll
        for (int o = 8; o <= 16; o <<= 1) {
            v.x += __shfl_xor_sync(0xffffffffu, v.x, o); v.y += __shfl_xor_sync(0xffffffffu, v.y, o);
            v.z += __shfl_xor_sync(0xffffffffu, v.z, o); v.w += __shfl_xor_sync(0xffffffffu, v.w, o);
        }
        const int col = kb * kBK + cj * 4;
        if (kb < KB && rr == 0 && col < K)                    // K % 4 == 0 (tc_supported: lda % 4, TMA box)
            *reinterpret_cast<float4*>(colsum_partial + (long long)my_row * K + col) = v;
    }
    __threadfence();
}

// After the kernel's final CTA-wide sync: the CTA that arrives LAST at the ticket folds all partial rows with all of its
// threads (no second launch for the bias gradient).  One warp doing this alone took 13 us of a 55 us product: ~150 KB
// through 16 loads in flight.  Here thread t owns the float4 column t % (K / 4) and the row group t / (K / 4); a group
// walks its rows g, g + G, ... sixteen at a time, the groups are then summed in group order through `scratch` (shared
// memory, >= 4 KB, free by now).  Which CTA folds varies from run to run, the order of the sum does not.  The ticket is
// left at 0.  Every thread of the CTA must call this.
__device__ __forceinline__ void colsum_cta_fold(int K, const float* __restrict__ colsum_partial, float* __restrict__ colsum_out,
                                               int* __restrict__ ticket, float* scratch, int* flag) {
    if (threadIdx.x == 0) {
        __threadfence();
        *flag = atomicAdd(ticket, 1) == (int)gridDim.x - 1;
    }
    __syncthreads();
    if (!*flag) return;
    __threadfence();
    const int nrows = (int)gridDim.x * kCsWarps;
    const int nc4 = K >> 2;                                   // float4 columns (<= 64)
    const int G = (int)blockDim.x / nc4;                      // row groups (>= 4 at 256 threads)
    const int c4 = (int)threadIdx.x % nc4, g = (int)threadIdx.x / nc4;
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    if (g < G) {
        const float4* base = reinterpret_cast<const float4*>(colsum_partial) + c4;
        // sixteen rows in flight; past the end the loads are clamped to the last row and their values dropped
        // (unconditional loads: a load under a branch would serialise the round trips)
        for (int r = g; r < nrows; r += 16 * G) {
            float4 t[16];
#pragma unroll
            for (int j = 0; j < 16; ++j) t[j] = __ldcg(base + (long long)min(r + j * G, nrows - 1) * nc4);
#pragma unroll
            for (int j = 0; j < 16; ++j) {
                const bool ok = r + j * G < nrows;
                s.x += ok ? t[j].x : 0.f; s.y += ok ? t[j].y : 0.f; s.z += ok ? t[j].z : 0.f; s.w += ok ? t[j].w : 0.f;
            }
        }
        reinterpret_cast<float4*>(scratch)[g * nc4 + c4] = s;
    }
    __syncthreads();
    if ((int)threadIdx.x < nc4) {
        float4 o = reinterpret_cast<const float4*>(scratch)[threadIdx.x];
        for (int j = 1; j < G; ++j) {
            const float4 t = reinterpret_cast<const float4*>(scratch)[j * nc4 + threadIdx.x];
            o.x += t.x; o.y += t.y; o.z += t.z; o.w += t.w;
        }
        reinterpret_cast<float4*>(colsum_out)[threadIdx.x] = o;
    }
    if (threadIdx.x == 0) *ticket = 0;
}

// (An eight-epilogue-warp variant was measured and removed: dX products 84.0 us against 81.9, forward 72.2 against 71.7 --
// the un-overlapped last epilogue is not what limits a launch.  DESIGN.md 8.1.)
template <bool kBMN, int kCluster>
__global__ void __launch_bounds__(kThreads2, 1)
k_gemm_tf32_rows2(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                  const __grid_constant__ CUtensorMap tmC, float* __restrict__ C, long long ldc, int M, int N, int K, int BN,
                  const float* __restrict__ bias, int flags, const float* __restrict__ mask, long long ldm,
                  uint32_t* __restrict__ bits_out, long long ldbits, float* __restrict__ colsum_partial,
                  const int* __restrict__ bucket_ptr, int n_groups, int group_rows, float* __restrict__ colsum_out,
                  int* __restrict__ ticket) {
    extern __shared__ uint8_t smem_raw[];
    Smem2 sm;
    sm.base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int rank = kCluster > 1 ? (int)cluster_ctarank() : 0;
    if (threadIdx.x == 0) {
        for (int s = 0; s < kStages2; ++s) {
            mbar_init(sm.full(s), 1);
            mbar_init(sm.empty(s), kCluster + (colsum_partial ? kCsWarps : 0));     // + this CTA's column-sum warps
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(sm.tmem_full(b), 1);
            mbar_init(sm.tmem_empty(b), 4);
        }
        mbar_fence_init();
    }
    if (warp == 1) tmem_alloc(smem_u32(sm.tmem_ptr()), kTmemCols);
    tc_fence_before();
    if (kCluster > 1) cluster_sync_all(); else __syncthreads();      // peers must see initialised barriers before any multicast
    tc_fence_after();
    const uint32_t tmem_base = *sm.tmem_ptr();
    pdl_enter();           // barriers, TMEM and the cluster handshake overlap the previous kernel's tail; global memory only from here
    // Optional (GNNX_GEMM_BACKOFF=1): waiting roles sleep between polls.  In k_pna_post_fused spinning waiters cost the
    // producer warps a third of the kernel; here the epilogue warps do not mind, and it is 1-2 % slower.
    const bool nap = (flags & kBackoffFlag) != 0;
    flags &= ~kBackoffFlag;
    auto wait = [&](uint32_t bar, uint32_t parity, unsigned ns) { if (nap) mbar_wait_sleep(bar, parity, ns); else mbar_wait(bar, parity); };

    const int n_tiles = (N + BN - 1) / BN;
    const int m_groups = ((M + 127) / 128 + kCluster - 1) / kCluster;
    const int work = n_tiles * m_groups;
    const int cid = blockIdx.x / kCluster, ncl = gridDim.x / kCluster;
    const int KB = (K + kBK - 1) / kBK;

    if (warp == 0) {
        if (lane == 0) {
            tma_prefetch_desc(&tmA);
            tma_prefetch_desc(&tmB);
            int stage = 0;
            uint32_t phase = 0;
            const uint32_t tx = kA2 + (uint32_t)BN * 128u;
            const int half = BN / kCluster;                      // rows (K-major) / columns (MN-major) of B this CTA fetches for the cluster
            constexpr uint16_t kAll = (uint16_t)((1u << kCluster) - 1u);
            for (int w = cid; w < work; w += ncl) {
                const int m0 = ((w / n_tiles) * kCluster + rank) * 128, n0 = (w % n_tiles) * BN;
                // grouped product: the row block's group (bucket boundaries are multiples of 128 * kCluster rows, so both
                // CTAs of a cluster agree) selects a block of B rows: weights [G * N, K] (K-major) or [G * K, N] (MN-major)
                int goff = 0;
                if (bucket_ptr) {
                    const int mbase = (w / n_tiles) * kCluster * 128;
                    for (int d = 1; d < n_groups; ++d) goff += (mbase >= __ldg(bucket_ptr + d)) ? group_rows : 0;
                }
                for (int kb = 0; kb < KB; ++kb) {
                    wait(sm.empty(stage), phase ^ 1, 64);
                    mbar_expect_tx(sm.full(stage), tx);
                    tma_load_2d(sm.a(stage), &tmA, sm.full(stage), kb * kBK, m0);
                    if (kCluster == 1) {
                        if (kBMN) tma_load_3d(sm.b(stage), &tmB, sm.full(stage), 0, kb * kBK + goff, n0 / 32);
                        else tma_load_2d(sm.b(stage), &tmB, sm.full(stage), kb * kBK, n0 + goff);
                    } else {
                        if (kBMN) tma_load_3d_mc(sm.b(stage) + rank * (half / 32) * 4096, &tmB, sm.full(stage), 0, kb * kBK + goff, n0 / 32 + rank * (half / 32), kAll);
                        else tma_load_2d_mc(sm.b(stage) + rank * half * 128, &tmB, sm.full(stage), kb * kBK, n0 + rank * half + goff, kAll);
                    }
                    if (++stage == kStages2) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            const uint32_t idesc = make_idesc_tf32(128, BN, 0, kBMN ? 1 : 0);
            int stage = 0;
            uint32_t phase = 0, i = 0;
            for (int w = cid; w < work; w += ncl, ++i) {
                const uint32_t buf = i & 1;
                wait(sm.tmem_empty(buf), ((i >> 1) & 1) ^ 1, 64);
                tc_fence_after();
                for (int kb = 0; kb < KB; ++kb) {
                    wait(sm.full(stage), phase, 20);
                    tc_fence_after();
#pragma unroll
                    for (int k = 0; k < kBK / 8; ++k) {
                        const uint64_t bdesc = kBMN ? make_smem_desc(sm.b(stage) + k * 1024, 32 * 128, 512, kLayoutSw128B32)
                                                    : make_smem_desc(sm.b(stage) + k * 32, 16, 1024, kLayoutSw128);
                        const uint64_t adesc = make_smem_desc(sm.a(stage) + k * 32, 16, 1024, kLayoutSw128);
                        mma_tf32(tmem_base + buf * 256, adesc, bdesc, idesc, (kb | k) ? 1u : 0u);
                    }
                    if (kCluster > 1) mma_commit_mc(sm.empty(stage), (uint16_t)((1u << kCluster) - 1u)); else mma_commit(sm.empty(stage));
                    if (kb == KB - 1) mma_commit(sm.tmem_full(buf));
                    if (++stage == kStages2) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp >= 6 && warp < 6 + kCsWarps) {
        // Column sums of A over this CTA's row tiles: out[k] = sum_m A[m, k] (rows past M arrive as TMA zero fill).
        // Lane l reads the 16-byte chunk (l & 7) -- columns 4(l&7)..+3 of the k-block -- of rows (l >> 3), +4, +8, ...;
        // a stage is released to the producer only after this warp has read it.
        if (colsum_partial) {
            float4 acc[kColsumMaxK / kBK];
#pragma unroll
            for (int j = 0; j < kColsumMaxK / kBK; ++j) acc[j] = make_float4(0.f, 0.f, 0.f, 0.f);
            const int cj = lane & 7, rr = (lane >> 3) + (warp - 6) * (128 / kCsWarps);
            int stage = 0;
            uint32_t phase = 0;
            for (int w = cid; w < work; w += ncl) {
                const bool first_n = (w % n_tiles) == 0;              // every A tile is loaded once per N tile: count it once
#pragma unroll
                for (int kb = 0; kb < kColsumMaxK / kBK; ++kb) {
                    if (kb < KB) {
                        wait(sm.full(stage), phase, 64);
                        if (first_n) {
                            const uint8_t* a = sm.base + stage * kStage2;
                            float4 s0 = make_float4(0.f, 0.f, 0.f, 0.f), s1 = s0;
#pragma unroll
                            for (int r = rr; r < rr + 128 / kCsWarps; r += 8) {       // K-major SW128: 16-byte chunk index ^ (row & 7)
                                const float4 v0 = *reinterpret_cast<const float4*>(a + r * 128 + ((cj ^ (r & 7)) << 4));
                                const float4 v1 = *reinterpret_cast<const float4*>(a + (r + 4) * 128 + ((cj ^ ((r + 4) & 7)) << 4));
                                s0.x += v0.x; s0.y += v0.y; s0.z += v0.z; s0.w += v0.w;
                                s1.x += v1.x; s1.y += v1.y; s1.z += v1.z; s1.w += v1.w;
                            }
                            acc[kb].x += s0.x + s1.x; acc[kb].y += s0.y + s1.y; acc[kb].z += s0.z + s1.z; acc[kb].w += s0.w + s1.w;
                        }
                        __syncwarp();
                        if (lane == 0) mbar_arrive(sm.empty(stage));
                        if (++stage == kStages2) { stage = 0; phase ^= 1; }
                    }
                }
            }
            colsum_store_partial(acc, lane, warp - 6, KB, K, colsum_partial);
        }
    } else if (warp < 6) {
        const int q = warp & 3;
        uint32_t it = 0, i = 0;
        const bool use_tma = !((flags & GNNX_GEMM_ACCUMULATE) && ((flags & GNNX_GEMM_RELU) || mask));
        if (use_tma && lane == 0) tma_prefetch_desc(&tmC);
        for (int w = cid; w < work; w += ncl, ++i) {
            const int m0 = ((w / n_tiles) * kCluster + rank) * 128, n0 = (w % n_tiles) * BN;
            const uint32_t buf = i & 1;
            const int ncols = min(BN, ((N - n0) + 31) / 32 * 32);
            wait(sm.tmem_full(buf), (i >> 1) & 1, 128);
            tc_fence_after();
            if (m0 < M) {
                if (use_tma)
                    epilogue_tma(sm, tmem_base + buf * 256, q, lane, &tmC, m0, M, n0, ncols, bias, N, flags, it, mask, ldm, bits_out, ldbits);
                else
                    epilogue_acc(sm, tmem_base + buf * 256, q, lane, C, ldc, (long long)m0, M, n0, N, ncols, bias, flags, mask, ldm);
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(sm.tmem_empty(buf));
        }
        if (use_tma && lane == 0) tma_store_wait_all();
    }
    tc_fence_before();
    __syncwarp();                                                    // re-converge the single-lane role warps
    if (kCluster > 1) cluster_sync_all(); else __syncthreads();      // no CTA may exit while its peer can still multicast into it
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, kTmemCols);
    }
    if (colsum_partial)                                              // staging tiles and the TMEM-pointer slot are free by now
        colsum_cta_fold(K, colsum_partial, colsum_out, ticket, sm.epi(0), reinterpret_cast<int*>(sm.tmem_ptr()) + 1);
}

// ------------------------------------------------------------------ the same product on CTA PAIRS (cta_group::2)
// k_gemm_tf32_rows2 stages the WHOLE weight tile in each CTA (48 KB stages: 16 KB of A rows, 32 KB of weights, 4 stages).
// Measured: these products are bound by bytes in flight -- 64 KB of A per SM against a ~3.8 us
// release -> refill loop is 17 KB/us, a 128-row tile per 7.7 us where HBM would allow 5.8.  A CTA pair issues ONE
// tcgen05.mma.cta_group::2 (M = 256: 128 rows from each SM) that reads half of the weight tile from each CTA's shared
// memory, so a stage is 16 KB + 16 KB and the same 192 KB hold SIX stages: 96 KB of A in flight per SM.  Each CTA loads
// its own A rows and its own half of the weights (no multicast: half the L2 -> SM weight traffic of rows2<.., 2>); both
// signal the LEADER's full barrier, the leader's MMA thread releases the stage in both CTAs with a multicast commit.
// Same accumulation order over K as rows2, same epilogue: results are bit-identical (tests/test_gpu_gemm_pair.py).
#ifndef GNNX_PAIR_STAGES
#define GNNX_PAIR_STAGES 6
#endif
#ifndef GNNX_PAIR_BUFS
#define GNNX_PAIR_BUFS 2
#endif
constexpr int kStagesP = GNNX_PAIR_STAGES;
constexpr int kEpiBufsP = GNNX_PAIR_BUFS;
constexpr int kEpiBytesP = 4 * kEpiBufsP * 4096;
constexpr int kBP = 128 * 128;                  // this CTA's half of a weight stage (<= 128 rows / 4 chunks of 32 columns)
constexpr int kStageP = kA2 + kBP;              // 32 KB
constexpr int kSmemP = 1024 + kStagesP * kStageP + kEpiBytesP + 256;
static_assert(kSmemP <= 232448, "pair: shared memory");

struct SmemP {
    static constexpr int kBufs = kEpiBufsP;
    uint8_t* base;
    __device__ uint32_t a(int s) const { return smem_u32(base + s * kStageP); }
    __device__ uint32_t b(int s) const { return smem_u32(base + s * kStageP + kA2); }
    __device__ float* epi(int q) const { return reinterpret_cast<float*>(base + kStagesP * kStageP + q * kEpiBufsP * 4096); }
    __device__ uint64_t* bars() const { return reinterpret_cast<uint64_t*>(base + kStagesP * kStageP + kEpiBytesP); }
    __device__ uint32_t full(int s) const { return smem_u32(bars() + s); }                       // used in the leader only
    __device__ uint32_t empty(int s) const { return smem_u32(bars() + kStagesP + s); }
    __device__ uint32_t afull(int s) const { return smem_u32(bars() + 2 * kStagesP + s); }       // "stage landed", forwarded to the column-sum warps
    __device__ uint32_t tmem_full(int b) const { return smem_u32(bars() + 3 * kStagesP + b); }
    __device__ uint32_t tmem_empty(int b) const { return smem_u32(bars() + 3 * kStagesP + 2 + b); }   // used in the leader only
    __device__ uint32_t* tmem_ptr() const { return reinterpret_cast<uint32_t*>(bars() + 3 * kStagesP + 4); }
};

template <bool kBMN>
__global__ void __launch_bounds__(kThreads2, 1)
k_gemm_tf32_pair(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                 const __grid_constant__ CUtensorMap tmC, float* __restrict__ C, long long ldc, int M, int N, int K, int BN,
                 const float* __restrict__ bias, int flags, const float* __restrict__ mask, long long ldm,
                 uint32_t* __restrict__ bits_out, long long ldbits, float* __restrict__ colsum_partial,
                 const int* __restrict__ bucket_ptr, int n_groups, int group_rows, float* __restrict__ colsum_out,
                 int* __restrict__ ticket) {
    extern __shared__ uint8_t smem_raw[];
    SmemP sm;
    sm.base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int rank = (int)cluster_ctarank();                       // 0: the leader (issues the MMAs)
    if (threadIdx.x == 0) {
        for (int s = 0; s < kStagesP; ++s) {
            mbar_init(sm.full(s), 1);                                // the leader's expect_tx arrival; bytes of BOTH CTAs
            mbar_init(sm.empty(s), 1 + (colsum_partial ? kCsWarps : 0));   // the pair's commit (+ this CTA's column-sum warps)
            mbar_init(sm.afull(s), 1);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(sm.tmem_full(b), 1);
            mbar_init(sm.tmem_empty(b), 8);                          // four epilogue warps of each CTA
        }
        mbar_fence_init();
    }
    if (warp == 1) tmem_alloc_pair(smem_u32(sm.tmem_ptr()), kTmemCols);
    tc_fence_before();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem_base = *sm.tmem_ptr();
    pdl_enter();

    const int n_tiles = (N + BN - 1) / BN;
    const int m_groups = ((M + 127) / 128 + 1) / 2;
    const int work = n_tiles * m_groups;
    const int cid = blockIdx.x / 2, ncl = gridDim.x / 2;
    const int KB = (K + kBK - 1) / kBK;
    const int half = BN / 2;                                         // weight rows (K-major) / columns (MN-major) held by each CTA

    if (warp == 0) {
        if (lane == 0) {
            tma_prefetch_desc(&tmA);
            tma_prefetch_desc(&tmB);
            int stage = 0;
            uint32_t phase = 0;
            const uint32_t tx = 2u * (kA2 + (uint32_t)half * 128u);
            for (int w = cid; w < work; w += ncl) {
                const int m0 = ((w / n_tiles) * 2 + rank) * 128, n0 = (w % n_tiles) * BN;
                int goff = 0;
                if (bucket_ptr) {
                    const int mbase = (w / n_tiles) * 256;
                    for (int d = 1; d < n_groups; ++d) goff += (mbase >= __ldg(bucket_ptr + d)) ? group_rows : 0;
                }
                for (int kb = 0; kb < KB; ++kb) {
                    mbar_wait(sm.empty(stage), phase ^ 1);
                    const uint32_t lead_full = mapa_cluster(sm.full(stage), 0);
                    if (rank == 0) mbar_expect_tx(sm.full(stage), tx);
                    tma_load_2d_pair(sm.a(stage), &tmA, lead_full, kb * kBK, m0);
                    if (kBMN) tma_load_3d_pair(sm.b(stage), &tmB, lead_full, 0, kb * kBK + goff, n0 / 32 + rank * (half / 32));
                    else tma_load_2d_pair(sm.b(stage), &tmB, lead_full, kb * kBK, n0 + rank * half + goff);
                    if (++stage == kStagesP) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0 && rank == 0) {
            const uint32_t idesc = make_idesc_tf32(256, BN, 0, kBMN ? 1 : 0);
            int stage = 0;
            uint32_t phase = 0, i = 0;
            for (int w = cid; w < work; w += ncl, ++i) {
                const uint32_t buf = i & 1;
                mbar_wait_cluster(sm.tmem_empty(buf), ((i >> 1) & 1) ^ 1);
                tc_fence_after();
                for (int kb = 0; kb < KB; ++kb) {
                    mbar_wait(sm.full(stage), phase);
                    tc_fence_after();
                    if (colsum_partial) {                            // the column-sum warps of both CTAs may read the stage now
                        mbar_arrive(sm.afull(stage));
                        mbar_arrive_cluster(mapa_cluster(sm.afull(stage), 1));
                    }
#pragma unroll
                    for (int k = 0; k < kBK / 8; ++k) {
                        const uint64_t bdesc = kBMN ? make_smem_desc(sm.b(stage) + k * 1024, 32 * 128, 512, kLayoutSw128B32)
                                                    : make_smem_desc(sm.b(stage) + k * 32, 16, 1024, kLayoutSw128);
                        const uint64_t adesc = make_smem_desc(sm.a(stage) + k * 32, 16, 1024, kLayoutSw128);
                        mma_tf32_pair(tmem_base + buf * 256, adesc, bdesc, idesc, (kb | k) ? 1u : 0u);
                    }
                    mma_commit_pair(sm.empty(stage), 3);
                    if (kb == KB - 1) mma_commit_pair(sm.tmem_full(buf), 3);
                    if (++stage == kStagesP) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp >= 6 && warp < 6 + kCsWarps) {
        // column sums of this CTA's A tiles (see k_gemm_tf32_rows2); the "stage landed" signal comes from the leader's MMA thread
        if (colsum_partial) {
            float4 acc[kColsumMaxK / kBK];
#pragma unroll
            for (int j = 0; j < kColsumMaxK / kBK; ++j) acc[j] = make_float4(0.f, 0.f, 0.f, 0.f);
            const int cj = lane & 7, rr = (lane >> 3) + (warp - 6) * (128 / kCsWarps);
            int stage = 0;
            uint32_t phase = 0;
            for (int w = cid; w < work; w += ncl) {
                const bool first_n = (w % n_tiles) == 0;
#pragma unroll
                for (int kb = 0; kb < kColsumMaxK / kBK; ++kb) {
                    if (kb < KB) {
                        mbar_wait_cluster(sm.afull(stage), phase);
                        if (first_n) {
                            const uint8_t* a = sm.base + stage * kStageP;
                            float4 s0 = make_float4(0.f, 0.f, 0.f, 0.f), s1 = s0;
#pragma unroll
                            for (int r = rr; r < rr + 128 / kCsWarps; r += 8) {
                                const float4 v0 = *reinterpret_cast<const float4*>(a + r * 128 + ((cj ^ (r & 7)) << 4));
                                const float4 v1 = *reinterpret_cast<const float4*>(a + (r + 4) * 128 + ((cj ^ ((r + 4) & 7)) << 4));
                                s0.x += v0.x; s0.y += v0.y; s0.z += v0.z; s0.w += v0.w;
                                s1.x += v1.x; s1.y += v1.y; s1.z += v1.z; s1.w += v1.w;
                            }
                            acc[kb].x += s0.x + s1.x; acc[kb].y += s0.y + s1.y; acc[kb].z += s0.z + s1.z; acc[kb].w += s0.w + s1.w;
                        }
                        __syncwarp();
                        if (lane == 0) mbar_arrive(sm.empty(stage));
                        if (++stage == kStagesP) { stage = 0; phase ^= 1; }
                    }
                }
            }
            colsum_store_partial(acc, lane, warp - 6, KB, K, colsum_partial);
        }
    } else if (warp < 6) {
        const int q = warp & 3;
        uint32_t it = 0, i = 0;
        const bool use_tma = !((flags & GNNX_GEMM_ACCUMULATE) && ((flags & GNNX_GEMM_RELU) || mask));
        if (use_tma && lane == 0) tma_prefetch_desc(&tmC);
        for (int w = cid; w < work; w += ncl, ++i) {
            const int m0 = ((w / n_tiles) * 2 + rank) * 128, n0 = (w % n_tiles) * BN;
            const uint32_t buf = i & 1;
            const int ncols = min(BN, ((N - n0) + 31) / 32 * 32);
            mbar_wait_cluster(sm.tmem_full(buf), (i >> 1) & 1);
            tc_fence_after();
            if (m0 < M) {
                if (use_tma)
                    epilogue_tma(sm, tmem_base + buf * 256, q, lane, &tmC, m0, M, n0, ncols, bias, N, flags, it, mask, ldm, bits_out, ldbits);
                else
                    epilogue_acc(sm, tmem_base + buf * 256, q, lane, C, ldc, (long long)m0, M, n0, N, ncols, bias, flags, mask, ldm);
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(mapa_cluster(sm.tmem_empty(buf), 0));   // both CTAs' epilogues free the pair's accumulator
        }
        if (use_tma && lane == 0) tma_store_wait_all();
    }
    tc_fence_before();
    __syncwarp();
    cluster_sync_all();                                              // no CTA may exit (or free TMEM) while the pair's MMAs / commits can reach it
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc_pair(tmem_base, kTmemCols);
    }
    if (colsum_partial)
        colsum_cta_fold(K, colsum_partial, colsum_out, ticket, sm.epi(0), reinterpret_cast<int*>(sm.tmem_ptr()) + 1);
}

// ------------------------------------------------------------------ Linear -> ReLU -> Linear (-> ...) chains, ONE launch
// Up to kChainMax same-width products C_p = act(A_p W_p^T + b_p) with A_{p+1} = C_p, row-local: tile t of product p + 1
// needs only tile t of product p.  Separate launches pay the kernel boundary per product (drain + refill, ~12 us of a
// ~45 us product at 100 k rows) and read every intermediate back from HBM.  Here a cluster walks its row tiles in groups
// of kChainGroup: all tiles of the group through product 0, then through product 1, ... -- so the store of (tile, p) has
// a group's worth of other work items to complete before (tile, p + 1) loads it back, from L2.  The epilogue warps
// signal "the stores of item i are complete" on a small ring of mbarriers (cp.async.bulk.wait_group 0, then arrive),
// and the TMA thread waits for the item it depends on before issuing that item's A loads.  Everything else is
// k_gemm_tf32_rows2: multicast weights, two accumulators ping-ponging in TMEM, TMA-store epilogue, packed ReLU masks.
constexpr int kChainMax = 4;
constexpr int kChainGroupMax = 8;                   // tiles per group (run-time G <= this)
constexpr int kChainSlots = 16;                     // ring of "stores complete" barriers: > G + 2
struct ChainArgs {
    CUtensorMap a[kChainMax], b[kChainMax], c[kChainMax];
    const float* bias[kChainMax];
    uint32_t* bits[kChainMax];
    int flags[kChainMax];
    int n;
};

struct ChainSmem : Smem2 {
    __device__ uint32_t cdone(int s) const { return smem_u32(bars() + 2 * kStages2 + 6 + s); }
};
constexpr int kSmemChain = kSmem2 + 64;

// item i of a cluster's schedule -> (row-tile index within the cluster's list, product); g = tiles in the item's group
struct ChainItem { int tile, p, g; };
__device__ __forceinline__ ChainItem chain_item(int i, int ntiles, int n, int G) {
    const int per_full = G * n;
    const int full = ntiles / G;                           // full groups
    int grp = i / per_full, r = i - grp * per_full, g = G;
    if (grp >= full) {                                     // the last, partial group
        grp = full;
        r = i - full * per_full;
        g = ntiles - full * G;
    }
    ChainItem it;
    it.g = g;
    it.p = r / g;
    it.tile = grp * G + (r - it.p * g);
    return it;
}

template <int kCluster>
__global__ void __launch_bounds__(kThreads2, 1)
k_gemm_tf32_chain(const __grid_constant__ ChainArgs args, int M, int W, int G) {
    extern __shared__ uint8_t smem_raw[];
    ChainSmem sm;
    sm.base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int rank = kCluster > 1 ? (int)cluster_ctarank() : 0;
    if (threadIdx.x == 0) {
        for (int s = 0; s < kStages2; ++s) {
            mbar_init(sm.full(s), 1);
            mbar_init(sm.empty(s), kCluster);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(sm.tmem_full(b), 1);
            mbar_init(sm.tmem_empty(b), 4);
        }
        for (int s = 0; s < kChainSlots; ++s) mbar_init(sm.cdone(s), 4);
        mbar_fence_init();
    }
    if (warp == 1) tmem_alloc(smem_u32(sm.tmem_ptr()), kTmemCols);
    tc_fence_before();
    if (kCluster > 1) cluster_sync_all(); else __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *sm.tmem_ptr();
    pdl_enter();

    const int n = args.n;
    const int m_groups = ((M + 127) / 128 + kCluster - 1) / kCluster;      // row blocks of kCluster tiles
    const int cid = blockIdx.x / kCluster, ncl = gridDim.x / kCluster;
    const int ntiles = cid < m_groups ? (m_groups - cid + ncl - 1) / ncl : 0;   // this cluster's row blocks: cid, cid + ncl, ...
    const int items = ntiles * n;
    const int KB = (W + kBK - 1) / kBK;
    const int BN = W;                                                      // one N tile: W <= 256

    if (warp == 0) {
        if (lane == 0) {
            for (int p = 0; p < n; ++p) { tma_prefetch_desc(&args.a[p]); tma_prefetch_desc(&args.b[p]); }
            int stage = 0;
            uint32_t phase = 0;
            const uint32_t tx = kA2 + (uint32_t)BN * 128u;
            const int half = BN / kCluster;
            constexpr uint16_t kAll = (uint16_t)((1u << kCluster) - 1u);
            for (int i = 0; i < items; ++i) {
                const ChainItem it = chain_item(i, ntiles, n, G);
                const int m0 = ((cid + it.tile * ncl) * kCluster + rank) * 128;
                if (it.p > 0) {                                            // A of this item = C of item i - g: wait for its stores
                    const int d = i - it.g;
                    mbar_wait(sm.cdone(d % kChainSlots), (uint32_t)((d / kChainSlots) & 1));
                    asm volatile("fence.proxy.async;" ::: "memory");
                }
                for (int kb = 0; kb < KB; ++kb) {
                    mbar_wait(sm.empty(stage), phase ^ 1);
                    mbar_expect_tx(sm.full(stage), tx);
                    tma_load_2d(sm.a(stage), &args.a[it.p], sm.full(stage), kb * kBK, m0);
                    if (kCluster == 1) tma_load_2d(sm.b(stage), &args.b[it.p], sm.full(stage), kb * kBK, 0);
                    else tma_load_2d_mc(sm.b(stage) + rank * half * 128, &args.b[it.p], sm.full(stage), kb * kBK, rank * half, kAll);
                    if (++stage == kStages2) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            const uint32_t idesc = make_idesc_tf32(128, BN, 0, 0);
            int stage = 0;
            uint32_t phase = 0;
            for (int i = 0; i < items; ++i) {
                const uint32_t buf = i & 1;
                mbar_wait(sm.tmem_empty(buf), ((i >> 1) & 1) ^ 1);
                tc_fence_after();
                for (int kb = 0; kb < KB; ++kb) {
                    mbar_wait(sm.full(stage), phase);
                    tc_fence_after();
#pragma unroll
                    for (int k = 0; k < kBK / 8; ++k) {
                        const uint64_t bdesc = make_smem_desc(sm.b(stage) + k * 32, 16, 1024, kLayoutSw128);
                        const uint64_t adesc = make_smem_desc(sm.a(stage) + k * 32, 16, 1024, kLayoutSw128);
                        mma_tf32(tmem_base + buf * 256, adesc, bdesc, idesc, (kb | k) ? 1u : 0u);
                    }
                    if (kCluster > 1) mma_commit_mc(sm.empty(stage), (uint16_t)((1u << kCluster) - 1u)); else mma_commit(sm.empty(stage));
                    if (kb == KB - 1) mma_commit(sm.tmem_full(buf));
                    if (++stage == kStages2) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp >= 2 && warp < 6) {
        const int q = warp & 3;
        uint32_t sit = 0;
        int signalled = 0;                                                 // items whose completion this warp has announced
        for (int i = 0; i < items; ++i) {
            const ChainItem it = chain_item(i, ntiles, n, G);
            const int m0 = ((cid + it.tile * ncl) * kCluster + rank) * 128;
            const uint32_t buf = i & 1;
            mbar_wait(sm.tmem_full(buf), (i >> 1) & 1);
            tc_fence_after();
            if (m0 < M)
                epilogue_tma(sm, tmem_base + buf * 256, q, lane, &args.c[it.p], m0, M, 0, (BN + 31) / 32 * 32, args.bias[it.p], W,
                             args.flags[it.p], sit, nullptr, 0, args.bits[it.p], (W + 31) / 32);
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {
                mbar_arrive(sm.tmem_empty(buf));
                // "stores complete" signals.  Waiting for item i's own stores here would idle this warp for the store
                // latency on every item; instead item i - 1 is signalled now -- after item i's groups were committed,
                // wait_group <groups of one item> leaves only those pending -- unless a nearby item depends on item i
                // (groups of one or two tiles), which gets the full wait.
                const bool urgent = it.g <= 2 && it.p + 1 < n;
                if (urgent || i == items - 1) {
                    tma_store_wait_all();
                    for (; signalled <= i; ++signalled) mbar_arrive(sm.cdone(signalled % kChainSlots));
                } else {
                    tma_store_wait_pending((BN + 31) / 32);
                    for (; signalled < i; ++signalled) mbar_arrive(sm.cdone(signalled % kChainSlots));
                }
            }
            __syncwarp();
        }
    }
    tc_fence_before();
    __syncwarp();
    if (kCluster > 1) cluster_sync_all(); else __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, kTmemCols);
    }
}

// ------------------------------------------------------------------ PNAConv: aggregate -> first post Linear, ONE kernel
// out[r, :] = act([mean | min | max | std | x](row r) W_eff(d_r)^T + b) for the degree-bucketed rows (degree_groups.cu).
// Same pipeline as k_gemm_tf32_rows2 (weights by multicast TMA, tcgen05 accumulators ping-ponging in TMEM, TMA-store
// epilogue), but the A operand never exists in HBM on the way in: eight PRODUCER warps build every 128-row x 32-column
// A stage in shared memory -- the swizzled K-major layout TMA would have written -- straight from the dst-sorted
// messages m[E, H] (mean / min / max / std of each row's incoming run, k_pna_aggregate_fwd's operation sequence) and from
// x[row_node[r]].  The K order is permuted so that the four aggregates of one 32-column chunk are consecutive k-blocks
// (the weight box is fetched at the matching K coordinate, the sum over K does not care): the chunk's messages are read
// from HBM once and served to the other three from L1 / L2.  The product is tensor-bound (K = 5H), so the producers'
// loads hide under the MMAs; the separate aggregate pass (write + re-read of [rows, 5H]) is gone.  For the backward's
// weight gradient the producers can also store what they computed (`agg_out`, fp32, unrounded).
// Warp roles: warps 0-3 = TMA, MMA + two idle warps, warps 4-7 = epilogue (TMEM lane quadrant = warp % 4), warps 8-15 =
// producers.  512 threads x 128 registers; setmaxnreg re-balancing was tried and made ptxas spill more, not less.
constexpr int kProdWarps = 8;                       // two groups of four: alternate units of k-blocks
constexpr int kFirstEpiWarp = 4;
constexpr int kFirstProdWarp = 8;
constexpr int kThreadsF = (kFirstProdWarp + kProdWarps) * 32;
constexpr float kStdMaskF = 0.003162277629598975f;  // sqrt(1e-5f) rounded as in pna.cu (GNNX_STD_MASK)

__device__ __forceinline__ float to_tf32(float v, int mode) {
    uint32_t r;
    if (mode == 0) asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(v));
    else if (mode == 1) asm("cvt.rn.tf32.f32 %0, %1;" : "=r"(r) : "f"(v));
    else return v;
    return __uint_as_float(r);
}

// One output float4 of aggregate `typ` (0 mean, 1 min, 2 max, 3 std) of a row with `deg` messages, the first U of them in
// registers (entries past deg are ignored), the tail beyond 4 read from memory (hypervalent rows only).  The operation
// sequence is agg_acc / k_pna_aggregate_fwd's of pna.cu, so the results are bit-identical with that kernel's.
template <int U>
__device__ __forceinline__ float4 agg_reduce4(const float4 (&v)[U], int deg, int typ, const float* __restrict__ tail, int H) {
    float s[4] = {0.f, 0.f, 0.f, 0.f}, s2[4] = {0.f, 0.f, 0.f, 0.f};
    float mn[4] = {INFINITY, INFINITY, INFINITY, INFINITY}, mx[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
    for (int u = 0; u < U; ++u) {
        const bool on = u < deg;
        const float e[4] = {v[u].x, v[u].y, v[u].z, v[u].w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            s[j] = on ? __fadd_rn(s[j], e[j]) : s[j];
            s2[j] = on ? __fadd_rn(s2[j], __fmul_rn(e[j], e[j])) : s2[j];
            mn[j] = on ? fminf(mn[j], e[j]) : mn[j];
            mx[j] = on ? fmaxf(mx[j], e[j]) : mx[j];
        }
    }
    if (U >= 4)
        for (int u = 4; u < deg; ++u) {
            const float4 q = __ldg(reinterpret_cast<const float4*>(tail + (long long)u * H));
            const float e[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                s[j] = __fadd_rn(s[j], e[j]);
                s2[j] = __fadd_rn(s2[j], __fmul_rn(e[j], e[j]));
                mn[j] = fminf(mn[j], e[j]);
                mx[j] = fmaxf(mx[j], e[j]);
            }
        }
    const float cnt = fmaxf((float)deg, 1.f);
    float o[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const float mu = __fdiv_rn(s[j], cnt);
        if (typ == 0) o[j] = mu;
        else if (typ == 1) o[j] = deg ? mn[j] : 0.f;
        else if (typ == 2) o[j] = deg ? mx[j] : 0.f;
        else {
            const float m2 = __fdiv_rn(s2[j], cnt);
            const float var = __fsub_rn(m2, __fmul_rn(mu, mu));
            const float sd = __fsqrt_rn(fmaxf(var, 1e-5f));
            o[j] = (sd <= kStdMaskF) ? 0.f : sd;
        }
    }
    return make_float4(o[0], o[1], o[2], o[3]);
}

template <int kCluster>
__global__ void __launch_bounds__(kThreadsF, 1)
k_pna_post_fused(const __grid_constant__ CUtensorMap tmB, const __grid_constant__ CUtensorMap tmC,
                 const float* __restrict__ msg, const int* __restrict__ rowptr, const int* __restrict__ row_node,
                 const float* __restrict__ x, int M, int N, int H, int F, int BN, const float* __restrict__ bias, int flags,
                 uint32_t* __restrict__ bits_out, long long ldbits, const int* __restrict__ bucket_ptr, int n_groups,
                 float* __restrict__ agg_out, long long ld_agg, int round_mode) {
    extern __shared__ uint8_t smem_raw[];
    Smem2 sm;
    sm.base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int rank = kCluster > 1 ? (int)cluster_ctarank() : 0;
    if (threadIdx.x == 0) {
        for (int s = 0; s < kStages2; ++s) {
            mbar_init(sm.full(s), (round_mode & 512) ? 1 : (round_mode & 4096) ? 2 : 1 + kProdWarps);   // the weight TMA's expect_tx + the eight producer warps
            mbar_init(sm.empty(s), kCluster);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(sm.tmem_full(b), 1);
            mbar_init(sm.tmem_empty(b), 4);
        }
        mbar_fence_init();
    }
    if (warp == 1) tmem_alloc(smem_u32(sm.tmem_ptr()), kTmemCols);
    tc_fence_before();
    if (kCluster > 1) cluster_sync_all(); else __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *sm.tmem_ptr();
    pdl_enter();

    const int n_tiles = (N + BN - 1) / BN;
    const int m_groups = ((M + 127) / 128 + kCluster - 1) / kCluster;
    const int work = n_tiles * m_groups;
    const int cid = blockIdx.x / kCluster, ncl = gridDim.x / kCluster;
    const int HB = H / kBK;                          // 32-column chunks of the hidden width
    const int KB = 5 * HB;                           // k-blocks: (chunk, aggregate) pairs, then the x chunks
    // K coordinate (column of W_eff / of [agg | x]) of k-block kb: agg layout is [tower][mean|min|max|std][F]
    auto kcoord = [&](int kb) -> int {
        if (kb >= 4 * HB) return 4 * H + (kb - 4 * HB) * kBK;
        const int c = (kb >> 2) * kBK, typ = kb & 3;
        return (c / F) * 4 * F + typ * F + (c % F);
    };

    if (warp == 0) {
        if (lane == 0) {
            tma_prefetch_desc(&tmB);
            int stage = 0;
            uint32_t phase = 0;
            const uint32_t tx = (uint32_t)BN * 128u;
            const int half = BN / 2;
            for (int w = cid; w < work; w += ncl) {
                const int n0 = (w % n_tiles) * BN;
                int goff = 0;
                const int mbase = (w / n_tiles) * kCluster * 128;
                for (int d = 1; d < n_groups; ++d) goff += (mbase >= __ldg(bucket_ptr + d)) ? N : 0;
                for (int kb = 0; kb < KB; ++kb) {
                    mbar_wait_sleep(sm.empty(stage), phase ^ 1, 64);
                    mbar_expect_tx(sm.full(stage), tx);
                    if (kCluster == 1) tma_load_2d(sm.b(stage), &tmB, sm.full(stage), kcoord(kb), n0 + goff);
                    else tma_load_2d_mc(sm.b(stage) + rank * half * 128, &tmB, sm.full(stage), kcoord(kb), n0 + rank * half + goff, 3);
                    if (++stage == kStages2) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            const uint32_t idesc = make_idesc_tf32(128, BN, 0, 0);
            int stage = 0;
            uint32_t phase = 0, i = 0;
            for (int w = cid; w < work; w += ncl, ++i) {
                const uint32_t buf = i & 1;
                mbar_wait_sleep(sm.tmem_empty(buf), ((i >> 1) & 1) ^ 1, 64);
                tc_fence_after();
                for (int kb = 0; kb < KB; ++kb) {
                    mbar_wait_sleep(sm.full(stage), phase, 32);
                    tc_fence_after();
#pragma unroll
                    for (int k = 0; k < kBK / 8; ++k) {
                        const uint64_t bdesc = make_smem_desc(sm.b(stage) + k * 32, 16, 1024, kLayoutSw128);
                        const uint64_t adesc = make_smem_desc(sm.a(stage) + k * 32, 16, 1024, kLayoutSw128);
                        mma_tf32(tmem_base + buf * 256, adesc, bdesc, idesc, (kb | k) ? 1u : 0u);
                    }
                    if (kCluster > 1) mma_commit_mc(sm.empty(stage), 3); else mma_commit(sm.empty(stage));
                    if (kb == KB - 1) mma_commit(sm.tmem_full(buf));
                    if (++stage == kStages2) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp >= kFirstProdWarp && !(round_mode & 512) && (!(round_mode & 4096) || warp == kFirstProdWarp)) {
        // ---- producers: all eight warps build every k-block.  Warp pw owns rows pw * 16 .. +15 of the tile; lane l covers
        // the 16-byte chunk (l & 7) of rows (l >> 3) + 4 i, i < 4: 8 lanes read one 128-byte row segment of a message,
        // 4 rows per load instruction.  The k-blocks come in UNITS -- the four aggregates of one 32-column chunk from ONE
        // set of message loads held in registers, then the x chunks two at a time -- and with in-degree <= 2 the loads of
        // unit u + 1 are issued before unit u is computed.  Loads are unconditional (clamped row indices, selects in the
        // arithmetic) so that all loads of a unit are in flight together.
        const int pw = warp - kFirstProdWarp;
        const int cj = lane & 7, rr = lane >> 3;
        const float* dummy = x + cj * 4;                      // any mapped address: deg-0 / padding rows load it and discard it
        const int NU = HB + (HB + 1) / 2;                     // units per tile
        uint32_t it = 0;                                      // k-blocks of this CTA so far (all work items)
        for (int w = cid; w < work; w += ncl) {
            const int m0 = ((w / n_tiles) * kCluster + rank) * 128;
            const bool store_agg = agg_out != nullptr && (w % n_tiles) == 0;
            // per row group: node id, and first edge slot + in-degree in one register (e0 << 4 | min(deg, 15): the bucket
            // layout caps the number of degree classes at GNNX_MAX_GROUPS = 16)
            uint32_t ed[4];
            int nd[4];
            int dmax = 0;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int r = m0 + pw * 16 + i * 4 + rr;
                const int node = (r < M && !(round_mode & 2048)) ? __ldg(row_node + r) : -1;
                const int b = node >= 0 ? __ldg(rowptr + node) : 0;
                const int d = node >= 0 ? __ldg(rowptr + node + 1) - b : 0;
                nd[i] = node;
                ed[i] = ((uint32_t)b << 4) | (uint32_t)min(d, 15);
                dmax = max(dmax, d);
            }
            dmax = __reduce_max_sync(0xffffffffu, dmax);

            // value -> A stage (tf32-rounded, swizzled K-major row) and, for the backward, the fp32 operand copy
            auto emit = [&](uint32_t k, int i, int kc, const float4& o) {
                if (round_mode & 128) return;
                const int r = pw * 16 + i * 4 + rr;
                if (store_agg && !(round_mode & 64) && m0 + r < M) __stcs(reinterpret_cast<float4*>(agg_out + (long long)(m0 + r) * ld_agg + kc), o);
                uint8_t* a = sm.base + (int)(k % kStages2) * kStage2;
                *reinterpret_cast<float4*>(a + r * 128 + ((cj ^ (r & 7)) << 4)) =
                    make_float4(to_tf32(o.x, round_mode & 3), to_tf32(o.y, round_mode & 3), to_tf32(o.z, round_mode & 3), to_tf32(o.w, round_mode & 3));
            };
            // one poller per warp: 256 threads spinning on try_wait keep the barrier's shared-memory word busy and delay
            // the very arrivals (tcgen05.commit, TMA complete_tx) they are waiting for
            auto wait_empty = [&](uint32_t k) {
                if (lane == 0 && !(round_mode & 1024)) mbar_wait_sleep(sm.empty((int)(k % kStages2)), ((k / kStages2) & 1u) ^ 1u, 20);
                __syncwarp();
            };
            auto publish = [&](uint32_t k) {
                if (!(round_mode & 128)) fence_proxy_async_smem();   // generic-proxy writes -> visible to the tensor core's async proxy
                __syncwarp();
                if (lane == 0) mbar_arrive(sm.full((int)(k % kStages2)));
            };
            // loads of unit u into v[row group][slot]: U message rows of chunk u, or two x chunks
            auto load_agg = [&](auto& v, int u) {
                constexpr int U = sizeof(v[0]) / sizeof(float4);
                const int col = u * kBK + cj * 4;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int d = (int)(ed[i] & 15u), b = (int)(ed[i] >> 4);
#pragma unroll
                    for (int k = 0; k < U; ++k)
                        v[i][k] = (round_mode & 16) ? make_float4(1.f, 2.f, 3.f, 4.f)
                                                    : __ldg(reinterpret_cast<const float4*>(d > 0 ? msg + (long long)(b + min(k, d - 1)) * H + col : dummy));
                }
            };
            auto load_x = [&](auto& v, int u) {
                const int xc = 2 * (u - HB), n = min(2, HB - xc);
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int k = 0; k < 2; ++k)
                        v[i][k] = __ldg(reinterpret_cast<const float4*>((nd[i] >= 0 && k < n) ? x + (long long)nd[i] * H + (xc + k) * kBK + cj * 4 : dummy));
            };
            auto produce_agg = [&](auto& v, int u) {
                constexpr int U = sizeof(v[0]) / sizeof(float4);
                const int col = u * kBK + cj * 4;
#pragma unroll
                for (int typ = 0; typ < 4; ++typ) {
                    wait_empty(it + typ);
                    const int kc = kcoord(4 * u + typ) + cj * 4;
#pragma unroll
                    for (int i = 0; i < 4; ++i)
                        emit(it + typ, i, kc, (round_mode & 32) ? v[i][0] : agg_reduce4<U>(v[i], (int)(ed[i] & 15u), typ, msg + (long long)(ed[i] >> 4) * H + col, H));
                    publish(it + typ);
                }
                it += 4;
            };
            auto produce_x = [&](auto& v, int u) {
                const int xc = 2 * (u - HB), n = min(2, HB - xc);
#pragma unroll
                for (int k = 0; k < 2; ++k)
                    if (k < n) {
                        wait_empty(it + k);
                        const int kc = 4 * H + (xc + k) * kBK + cj * 4;
#pragma unroll
                        for (int i = 0; i < 4; ++i) emit(it + k, i, kc, nd[i] >= 0 ? v[i][k] : make_float4(0.f, 0.f, 0.f, 0.f));
                        publish(it + k);
                    }
                it += n;
            };

            if (dmax <= 2) {
                // (a double-buffered variant that issued unit u + 1's loads before reducing unit u spilled ~45 registers
                // to local memory, whose lines the streaming loads keep evicting from the 28 KB L1: 90 % local-load misses)
#pragma unroll 1
                for (int u = 0; u < NU; ++u) {
                    float4 v[4][2];
                    if (u < HB) { load_agg(v, u); produce_agg(v, u); }
                    else { load_x(v, u); produce_x(v, u); }
                }
            } else {
#pragma unroll 1
                for (int u = 0; u < HB; ++u) {
                    float4 v[4][4];
                    load_agg(v, u);
                    produce_agg(v, u);
                }
#pragma unroll 1
                for (int u = HB; u < NU; ++u) {
                    float4 v[4][2];
                    load_x(v, u);
                    produce_x(v, u);
                }
            }
        }
    } else if (warp >= kFirstEpiWarp && warp < kFirstProdWarp) {
        const int q = warp & 3;
        uint32_t it = 0, i = 0;
        if (lane == 0) tma_prefetch_desc(&tmC);
        for (int w = cid; w < work; w += ncl, ++i) {
            const int m0 = ((w / n_tiles) * kCluster + rank) * 128, n0 = (w % n_tiles) * BN;
            const uint32_t buf = i & 1;
            mbar_wait_sleep(sm.tmem_full(buf), (i >> 1) & 1, 256);
            tc_fence_after();
            const int ncols = min(BN, ((N - n0) + 31) / 32 * 32);
            if (m0 < M && !(round_mode & 256))
                epilogue_tma(sm, tmem_base + buf * 256, q, lane, &tmC, m0, M, n0, ncols, bias, N, flags, it, nullptr, 0, bits_out, ldbits);
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(sm.tmem_empty(buf));
        }
        if (lane == 0) tma_store_wait_all();
    }
    tc_fence_before();
    __syncwarp();
    if (kCluster > 1) cluster_sync_all(); else __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, kTmemCols);
    }
}

// ------------------------------------------------------------------ partial[z][Mo,No] = A[rows_z, Mo]^T * B[rows_z, No]
// Fold of the row-split partials inside the product's own launch (C != null): after a grid barrier every CTA of a
// problem sums a slice of the output over the splits, in split order, exactly as k_splitk_reduce does (bit-identical),
// and applies bias / accumulate / ReLU.  The host enables it only when the whole grid is resident (<= one CTA per SM);
// counters[0] / [1] are the barrier's arrival and exit counts (stream_ticket() + 16), left at 0.
struct RedFold {
    float* C;
    long long ldc;
    const float* bias;
    int flags;
    int* counters;
};

__device__ __forceinline__ void red_body(const CUtensorMap* tmA, const CUtensorMap* tmB, const CUtensorMap* tmP,
                                         float* __restrict__ partial, int Mo, int No, int R, int rows_per_split, int BM,
                                         int BN, const int* __restrict__ split_table, const int split, const RedFold fold) {
    extern __shared__ uint8_t smem_raw[];
    const Smem sm = carve_smem(smem_raw);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t tmem_base = prologue(sm, warp);
    pdl_enter();           // barriers, TMEM and descriptors are set up while the previous kernel drains; global memory only from here

    const int n_tiles = (No + BN - 1) / BN;
    const int m0 = (blockIdx.x / n_tiles) * BM, n0 = (blockIdx.x % n_tiles) * BN;
    // row range of this split: uniform, or (grouped dW) read from the per-batch split table {begin, end, group}
    const int r_begin = split_table ? split_table[3 * split] : split * rows_per_split;
    const int r_end = split_table ? split_table[3 * split + 1] : min(R, r_begin + rows_per_split);
    const int RB = (r_end - r_begin + kBK - 1) / kBK;
    const int halves = BM / 128;

    if (warp == 0) {
        if (lane == 0) {
            tma_prefetch_desc(tmA);
            tma_prefetch_desc(tmB);
            int stage = 0;
            uint32_t phase = 0;
            const uint32_t tx = (uint32_t)(BM + BN) * 128u;
            for (int rb = 0; rb < RB; ++rb) {
                mbar_wait(sm.empty(stage), phase ^ 1);
                mbar_expect_tx(sm.full(stage), tx);
                tma_load_3d(sm.a(stage), tmA, sm.full(stage), 0, r_begin + rb * kBK, m0 / 32);
                tma_load_3d(sm.b(stage), tmB, sm.full(stage), 0, r_begin + rb * kBK, n0 / 32);
                if (++stage == kStages) { stage = 0; phase ^= 1; }
            }
        }
    } else if (warp == 1) {
        if (lane == 0 && RB > 0) {
            const uint32_t idesc = make_idesc_tf32(128, BN, 1, 1);
            int stage = 0;
            uint32_t phase = 0;
            for (int rb = 0; rb < RB; ++rb) {
                mbar_wait(sm.full(stage), phase);
                tc_fence_after();
#pragma unroll
                for (int k = 0; k < kBK / 8; ++k) {
                    const uint64_t bdesc = make_smem_desc(sm.b(stage) + k * 1024, 32 * 128, 512, kLayoutSw128B32);
                    for (int h = 0; h < halves; ++h) {
                        const uint64_t adesc = make_smem_desc(sm.a(stage) + h * (4 * 32 * 128) + k * 1024, 32 * 128, 512, kLayoutSw128B32);
                        mma_tf32(tmem_base + h * 256, adesc, bdesc, idesc, (rb | k) ? 1u : 0u);
                    }
                }
                mma_commit(sm.empty(stage));
                if (rb == RB - 1) mma_commit(sm.tmem_full());
                if (++stage == kStages) { stage = 0; phase ^= 1; }
            }
        }
    } else {
        const int q = warp & 3;
        float* out = partial + (long long)split * Mo * No;
        if (RB > 0) {
            mbar_wait(sm.tmem_full(), 0);
            tc_fence_after();
            const int ncols = min(BN, No - n0);
            uint32_t it = 0;
            // tmP views the partials as [splits * Mo, No]; Mo and No are multiples of 32, so a 32x32 staging tile is
            // either wholly inside this split's [Mo, No] block or skipped (row_end / col_end guards).
            const int prow = split * Mo;
            for (int h = 0; h < halves; ++h)
                epilogue_tma(sm, tmem_base + h * 256, q, lane, tmP, prow + m0 + h * 128, prow + Mo, n0, ncols, nullptr, No, 0, it);
            if (lane == 0) {
                tma_store_wait_all();
                if (fold.C) asm volatile("fence.proxy.async;" ::: "memory");     // bulk-store writes before the generic-proxy release below
            }
        } else if (!split_table) {                 // empty row range: this split contributes zeros (the grouped reduce skips it)
            for (int h = 0; h < halves; ++h)
                for (int r = q * 32 + lane; r < 128; r += 128) {
                    const int grow = m0 + h * 128 + r;
                    if (grow < Mo)
                        for (int c = n0; c < min(No, n0 + BN); ++c) out[(long long)grow * No + c] = 0.f;
                }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, kTmemCols);
    }
    if (fold.C) {
        const int nctas = (int)(gridDim.x * gridDim.y * gridDim.z);
        grid_barrier(fold.counters, nctas);
        asm volatile("fence.proxy.async;" ::: "memory");
        const int ci = (int)(blockIdx.x + gridDim.x * blockIdx.y), nci = (int)(gridDim.x * gridDim.y);   // the CTAs of this problem
        const int splits = (int)gridDim.y;
        const long long total = (long long)Mo * No, total4 = total >> 2;                                 // Mo, No multiples of 32
        const float4* p4 = reinterpret_cast<const float4*>(partial);
        for (long long i = (long long)ci * kThreads + threadIdx.x; i < total4; i += (long long)nci * kThreads) {
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            int sp = 0;
            for (; sp + 8 <= splits; sp += 8) {                 // 8 independent loads in flight, summed in split order
                float4 t[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) t[j] = __ldcg(p4 + (long long)(sp + j) * total4 + i);
#pragma unroll
                for (int j = 0; j < 8; ++j) { v.x += t[j].x; v.y += t[j].y; v.z += t[j].z; v.w += t[j].w; }
            }
            for (; sp < splits; ++sp) {
                const float4 t = __ldcg(p4 + (long long)sp * total4 + i);
                v.x += t.x; v.y += t.y; v.z += t.z; v.w += t.w;
            }
            const long long e = i << 2;
            const int m = (int)(e / No), n = (int)(e % No);
            float* dst = fold.C + (long long)m * fold.ldc + n;
            if (fold.bias) { const float4 b = __ldg(reinterpret_cast<const float4*>(fold.bias + n)); v.x += b.x; v.y += b.y; v.z += b.z; v.w += b.w; }
            if (fold.flags & GNNX_GEMM_ACCUMULATE) { const float4 c = *reinterpret_cast<const float4*>(dst); v.x += c.x; v.y += c.y; v.z += c.z; v.w += c.w; }
            if (fold.flags & GNNX_GEMM_RELU) { v.x = fmaxf(v.x, 0.f); v.y = fmaxf(v.y, 0.f); v.z = fmaxf(v.z, 0.f); v.w = fmaxf(v.w, 0.f); }
            *reinterpret_cast<float4*>(dst) = v;
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            if (atomicAdd(fold.counters + 1, 1) == nctas - 1) { fold.counters[0] = 0; fold.counters[1] = 0; __threadfence(); }
        }
    }
}


__global__ void __launch_bounds__(kThreads, 1)
k_gemm_tf32_red(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                const __grid_constant__ CUtensorMap tmP, float* __restrict__ partial, int Mo, int No, int R,
                int rows_per_split, int BM, int BN, const int* __restrict__ split_table, const RedFold fold) {
    red_body(&tmA, &tmB, &tmP, partial, Mo, No, R, rows_per_split, BM, BN, split_table, (int)blockIdx.y, fold);
}

// Up to kRedBatch weight gradients of one shape in ONE launch (blockIdx.z = problem): the three post_nn products of a
// PNAConv layer share the row count, so each gets a third of the CTAs, three times the rows per CTA (the prologue, the
// un-overlapped 256 KB epilogue and the launch are paid once instead of three times) and a third of the partials.
constexpr int kRedBatch = 4;
struct RedBatch { CUtensorMap a[kRedBatch], b[kRedBatch], p[kRedBatch]; float* c[kRedBatch]; long long ldc[kRedBatch]; int* counters; };
__global__ void __launch_bounds__(kThreads, 1)
k_gemm_tf32_red_batched(const __grid_constant__ RedBatch bt, float* __restrict__ partial, long long partial_stride, int Mo,
                        int No, int R, int rows_per_split, int BM, int BN) {
    const int z = blockIdx.z;
    const RedFold fold = {bt.counters ? bt.c[z] : nullptr, bt.ldc[z], nullptr, 0, bt.counters};
    red_body(&bt.a[z], &bt.b[z], &bt.p[z], partial + z * partial_stride, Mo, No, R, rows_per_split, BM, BN, nullptr,
             (int)blockIdx.y, fold);
}

// ------------------------------------------------------------------ host side
static PFN_cuTensorMapEncodeTiled_v12000 encode_fn() {
    static std::once_flag once;
    static PFN_cuTensorMapEncodeTiled_v12000 fn = nullptr;
    std::call_once(once, [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(p);
    });
    return fn;
}

// cudaFuncAttributeMaxDynamicSharedMemorySize is a per-device property of a kernel: set it once per (kernel set, device).
template <class SetFn>
static int ensure_attrs(std::atomic<uint64_t>& done, SetFn set) {
    int dev = 0;
    GNNX_CUDA(cudaGetDevice(&dev));
    const uint64_t bit = 1ull << (dev & 63);
    if (done.load(std::memory_order_acquire) & bit) return GNNX_OK;
    int rc = set();                                       // idempotent: two threads racing here both set the same value
    if (rc == GNNX_OK) done.fetch_or(bit, std::memory_order_release);
    return rc;
}

static CUtensorMapDataType tma_dtype() {
    static std::once_flag once;
    static int f32 = 0;
    std::call_once(once, [] {
        const char* e = getenv("GNNX_TMA_F32");      // diagnostic: raw fp32 bits (MMA truncates) instead of round-to-tf32
        f32 = (e && e[0] == '1') ? 1 : 0;
    });
    return f32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_TFLOAT32;
}

// ---- tensor-map cache.  A CUtensorMap is a pure function of (address, dims, row stride, box, kind): the allocator
// hands the same blocks back step after step, so the ~520 cuTensorMapEncodeTiled calls of a cfg2 step become lookups.
struct MapKey {
    const void* p; int64_t rows, cols, ld; int32_t box, kind, dtype, pad;
    bool operator==(const MapKey& o) const { return memcmp(this, &o, sizeof(MapKey)) == 0; }
};
struct MapKeyHash {
    size_t operator()(const MapKey& k) const {
        const uint64_t* w = reinterpret_cast<const uint64_t*>(&k);
        uint64_t h = 1469598103934665603ull;
        for (size_t i = 0; i < sizeof(MapKey) / 8; ++i) { h ^= w[i]; h *= 1099511628211ull; }
        return (size_t)h;
    }
};
enum { kMapKMajor = 0, kMapMNMajor = 1, kMapStore = 2 };

static int encode_map(CUtensorMap* m, const MapKey& k) {
    const float* p = static_cast<const float*>(k.p);
    CUresult r;
    if (k.kind == kMapKMajor) {
        // K-major view of a row-major [rows, cols] fp32 matrix: box = 32 floats x box rows.
        cuuint64_t dims[2] = {(cuuint64_t)k.cols, (cuuint64_t)k.rows};
        cuuint64_t strides[1] = {(cuuint64_t)k.ld * 4};
        cuuint32_t box[2] = {32, (cuuint32_t)k.box};
        cuuint32_t es[2] = {1, 1};
        r = encode_fn()(m, (CUtensorMapDataType)k.dtype, 2, const_cast<float*>(p), dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    } else if (k.kind == kMapMNMajor) {
        // MN-major view of a row-major [rows, cols] fp32 matrix (cols % 32 == 0): dims (32, rows, cols/32) so that one
        // box lands as [chunk][k row][32 floats] = the canonical MN-major operand layout of tf32 (128B swizzle, 32B atoms).
        cuuint64_t dims[3] = {32, (cuuint64_t)k.rows, (cuuint64_t)(k.cols / 32)};
        cuuint64_t strides[2] = {(cuuint64_t)k.ld * 4, 128};
        cuuint32_t box[3] = {32, 32, (cuuint32_t)k.box};
        cuuint32_t es[3] = {1, 1, 1};
        r = encode_fn()(m, (CUtensorMapDataType)k.dtype, 3, const_cast<float*>(p), dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    } else {
        // Store view of a row-major [rows, cols] fp32 matrix: 32x32 tiles, 128B swizzle (matches epilogue_tma's staging).
        cuuint64_t dims[2] = {(cuuint64_t)k.cols, (cuuint64_t)k.rows};
        cuuint64_t strides[1] = {(cuuint64_t)k.ld * 4};
        cuuint32_t box[2] = {32, 32};
        cuuint32_t es[2] = {1, 1};
        r = encode_fn()(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(p), dims, strides, box, es,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    }
    return r == CUDA_SUCCESS ? 0 : set_error(GNNX_ERR_INVALID, "cuTensorMapEncodeTiled (kind %d) failed: %d", k.kind, (int)r);
}

static int cached_map(CUtensorMap* m, const void* p, int64_t rows, int64_t cols, int64_t ld, int box, int kind) {
    static std::mutex mu;
    static std::unordered_map<MapKey, CUtensorMap, MapKeyHash> cache;
    MapKey k;
    memset(&k, 0, sizeof k);
    k.p = p; k.rows = rows; k.cols = cols; k.ld = ld; k.box = box; k.kind = kind;
    k.dtype = kind == kMapStore ? (int)CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : (int)tma_dtype();
    {
        std::lock_guard<std::mutex> lock(mu);
        auto it = cache.find(k);
        if (it != cache.end()) { *m = it->second; return 0; }
    }
    if (encode_map(m, k)) return GNNX_ERR_INVALID;
    std::lock_guard<std::mutex> lock(mu);
    if (cache.size() >= 8192) cache.clear();               // bounded: a long run with ever-new shapes starts over
    cache.emplace(k, *m);
    return 0;
}
static int make_map_kmajor(CUtensorMap* m, const float* p, int64_t rows, int64_t cols, int64_t ld, int box_rows) {
    return cached_map(m, p, rows, cols, ld, box_rows, kMapKMajor);
}
static int make_map_mnmajor(CUtensorMap* m, const float* p, int64_t rows, int64_t cols, int64_t ld, int box_chunks) {
    return cached_map(m, p, rows, cols, ld, box_chunks, kMapMNMajor);
}
static int make_map_store(CUtensorMap* m, const float* p, int64_t rows, int64_t cols, int64_t ld) {
    return cached_map(m, p, rows, cols, ld, 32, kMapStore);
}

static inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

struct RedPlan {
    int BM, BN, tiles, splits, rows_per_split;
};
static RedPlan red_plan(int64_t Mo, int64_t No, int64_t R) {
    RedPlan p;
    p.BM = Mo > 128 ? 256 : 128;
    p.BN = (int)(No < 256 ? No : 256);
    p.tiles = ceil_div(Mo, p.BM) * ceil_div(No, p.BN);
    int want = num_sms() / p.tiles;
    if (want < 1) want = 1;
    int64_t max_splits = (R + 255) / 256;                 // at least 256 rows per split
    if (max_splits < 1) max_splits = 1;
    int splits = (int)(want < max_splits ? want : max_splits);
    int rps = ceil_div(ceil_div(R, splits), kBK) * kBK;
    if (rps < kBK) rps = kBK;
    p.rows_per_split = rps;
    p.splits = R > 0 ? ceil_div(R, rps) : 1;
    return p;
}

}  // namespace tc

using namespace tc;

// One zero-initialised arrival counter per (device, stream): kernels of a stream run in order and each leaves it at 0.
// The counters of a device live in ONE block allocated (and zeroed) on first use, so that handing a counter to a new
// stream -- e.g. the capture stream of a CUDA graph -- involves no CUDA call at all (cudaMalloc / cudaMemset are not
// allowed while a stream is being captured; the first tensor-core product of a process must therefore run eagerly,
// which the usual warm-up step before a capture does anyway).
int* stream_ticket(cudaStream_t stream) {   // 64 ints per stream: [0] the GEMM column-sum ticket, [8..9] BatchNorm's grid barrier
    constexpr int kSlots = 1024, kStride = 64;             // ints: 256 B apart
    struct Pool { int* base = nullptr; int next = 0; std::map<cudaStream_t, int> slot; };
    static std::mutex mu;
    static std::map<int, Pool> pools;
    std::lock_guard<std::mutex> lock(mu);
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return nullptr;
    Pool& p = pools[dev];
    if (!p.base) {
        int* q = nullptr;
        if (cudaMalloc(&q, (size_t)kSlots * kStride * sizeof(int)) != cudaSuccess) { cudaGetLastError(); return nullptr; }
        if (cudaMemset(q, 0, (size_t)kSlots * kStride * sizeof(int)) != cudaSuccess) { cudaGetLastError(); cudaFree(q); return nullptr; }
        p.base = q;
    }
    auto it = p.slot.find(stream);
    if (it == p.slot.end()) it = p.slot.emplace(stream, p.next++ % kSlots).first;     // > 1024 live streams would share
    return p.base + (size_t)it->second * kStride;
}

static bool red_fold_enabled() {
    const char* e = getenv("GNNX_RED_FOLD");        // read per call: tests compare both paths in one process
    return !(e && e[0] == '0');
}

bool tc_supported(int64_t M, int64_t N, int64_t K, int ta, int tb, int64_t lda, int64_t ldb, int64_t ldc, const void* A,
                  const void* B, const void* C, int precision) {
    (void)precision;
    if (!encode_fn()) return false;
    if (!aligned16(A) || !aligned16(B) || !aligned16(C)) return false;
    if ((lda & 3) || (ldb & 3) || (ldc & 3)) return false;
    if (M > 0x7fffffff || N > 0x7fffffff || K > 0x7fffffff) return false;
    if (!ta) {
        if (M < 1 || N < 16 || K < 8 || (N & 3)) return false;
        if (tb) return true;                               // C = A B^T, both K-major
        return (N % 32) == 0;                              // C = A B, B [K,N] as an MN-major operand
    }
    if (tb) return false;                                  // A^T B^T: not on the hot path
    return M >= 32 && N >= 32 && (M % 32) == 0 && (N % 32) == 0 && K >= 32;
}

size_t tc_workspace_bytes(int64_t M, int64_t N, int64_t K, int ta, int tb, int precision) {
    (void)precision;
    if (ta && !tb) {
        RedPlan p = red_plan(M, N, K);
        return (size_t)p.splits * M * N * sizeof(float);
    }
    return 0;
}

int gemm_tc(const float* A, int64_t lda, int ta, const float* B, int64_t ldb, int tb, float* C, int64_t ldc, int64_t M,
            int64_t N, int64_t K, const float* bias, int flags, const float* mask, int64_t ldm, uint32_t* bits_out,
            float* colsum_a_out, int precision, void* ws, size_t ws_bytes, cudaStream_t st, const int* bucket_ptr,
            int n_groups) {
    (void)precision;
    if (!bucket_ptr) n_groups = 1;
    static std::atomic<uint64_t> attrs_done{0};
    {
        int rc = ensure_attrs(attrs_done, []() -> int {
            GNNX_CUDA(cudaFuncSetAttribute(k_gemm_tf32_red, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
            GNNX_CUDA(cudaFuncSetAttribute(k_gemm_tf32_rows2<false, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem2));
            GNNX_CUDA(cudaFuncSetAttribute(k_gemm_tf32_rows2<true, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem2));
            GNNX_CUDA(cudaFuncSetAttribute(k_gemm_tf32_rows2<false, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem2));
            GNNX_CUDA(cudaFuncSetAttribute(k_gemm_tf32_rows2<true, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem2));
            GNNX_CUDA(cudaFuncSetAttribute(k_gemm_tf32_rows2<false, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem2));
            GNNX_CUDA(cudaFuncSetAttribute(k_gemm_tf32_rows2<true, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem2));
            GNNX_CUDA(cudaFuncSetAttribute(k_gemm_tf32_pair<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemP));
            GNNX_CUDA(cudaFuncSetAttribute(k_gemm_tf32_pair<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemP));
            return GNNX_OK;
        });
        if (rc) return rc;
    }
    CUtensorMap mA, mB, mC;
    if (!ta) {
        const int BN = (int)(N >= 256 ? 256 : (tb ? (N + 15) / 16 * 16 : N));
        const int m_tiles = ceil_div(M, 128);
        // 2-CTA clusters halve the weight traffic from L2 per output tile.  4-CTA clusters (GNNX_GEMM_CLUSTER=4) halve it
        // again but were measured SLOWER on a B200 (cfg2 step: forward products 62.8 us against 58.9, dX 68.9 against 65.1):
        // these products are not bound by L2 -> SM bandwidth, and fewer 4-CTA clusters fit the GPCs.  Default: pairs.
        static std::once_flag cl_once;
        static int max_cluster = 2;
        std::call_once(cl_once, [] { const char* e = getenv("GNNX_GEMM_CLUSTER"); if (e && (e[0] == '1' || e[0] == '2' || e[0] == '4')) max_cluster = e[0] - '0'; });
        int cluster = (m_tiles >= 2 && (tb ? (BN % 16 == 0) : (BN % 64 == 0))) ? 2 : 1;
        if (cluster == 2 && max_cluster >= 4 && !bucket_ptr && m_tiles >= 16 && (tb ? (BN % 32 == 0) : (BN % 128 == 0))) cluster = 4;
        if (cluster > max_cluster) cluster = max_cluster;
        if (make_map_kmajor(&mA, A, M, K, lda, 128)) return GNNX_ERR_INVALID;
        if (tb) {                                        // grouped: B stacks one [N, K] / [K, N] weight block per group
            if (make_map_kmajor(&mB, B, N * n_groups, K, ldb, BN / cluster)) return GNNX_ERR_INVALID;
        } else {
            if (make_map_mnmajor(&mB, B, K * n_groups, N, ldb, BN / 32 / cluster)) return GNNX_ERR_INVALID;
        }
        const int group_rows = (int)(tb ? N : K);
        if (make_map_store(&mC, C, M, N, ldc)) return GNNX_ERR_INVALID;
        const int work = ceil_div(m_tiles, cluster) * ceil_div(N, BN);
        int max_cl = num_sms() / cluster;
        if (cluster == 4) {
            // clusters are placed inside one GPC: how many 4-CTA clusters fit at once is a property of the die, not SMs / 4
            static std::atomic<int> fit4{0};
            int f = fit4.load();
            if (f == 0) {
                cudaLaunchConfig_t q = {};
                q.gridDim = dim3(num_sms() / 4 * 4); q.blockDim = dim3(kThreads2); q.dynamicSmemBytes = kSmem2;
                cudaLaunchAttribute qa[1];
                qa[0].id = cudaLaunchAttributeClusterDimension;
                qa[0].val.clusterDim.x = 4; qa[0].val.clusterDim.y = 1; qa[0].val.clusterDim.z = 1;
                q.attrs = qa; q.numAttrs = 1;
                int n = 0;
                if (cudaOccupancyMaxActiveClusters(&n, k_gemm_tf32_rows2<false, 4>, &q) != cudaSuccess || n < 1) { cudaGetLastError(); n = -1; }
                fit4.store(n); f = n;
            }
            if (f < 0) {                                         // the query failed: stay with pairs
                cluster = 2;
                if (tb) { if (make_map_kmajor(&mB, B, N * n_groups, K, ldb, BN / cluster)) return GNNX_ERR_INVALID; }
                else if (make_map_mnmajor(&mB, B, K * n_groups, N, ldb, BN / 32 / cluster)) return GNNX_ERR_INVALID;
                max_cl = num_sms() / cluster;
            } else if (f < max_cl) {
                max_cl = f;
            }
        }
        const int work_c = ceil_div(m_tiles, cluster) * ceil_div(N, BN);
        // Balanced rounds: 400 work items on 74 clusters are 5 full rounds plus one that keeps 30 clusters busy at their
        // per-SM rate; ceil(400 / 6) = 67 clusters all run 6 rounds at the chip's HBM rate instead (GNNX_GEMM_BALANCE=0: off)
        static std::once_flag bal_once;
        static int balance = 1;
        std::call_once(bal_once, [] { const char* e = getenv("GNNX_GEMM_BALANCE"); balance = (e && e[0] == '0') ? 0 : 1; });
        int ncl = work_c < max_cl ? work_c : max_cl;
        if (balance && work_c > max_cl) ncl = ceil_div(work_c, ceil_div(work_c, max_cl));
        (void)work;
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(ncl * cluster);
        cfg.blockDim = dim3(kThreads2);
        cfg.dynamicSmemBytes = kSmem2;
        cfg.stream = st;
        cudaLaunchAttribute at[2];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = cluster; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;      // see common.cuh: pdl_enter()
        at[1].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = at; cfg.numAttrs = pdl_enabled() ? 2 : 1;
        static std::once_flag nap_once;
        static int nap = 0;          // measured on a B200 (cfg2): 83.8 / 73.5 us per dX / forward product with back-off, 82.5 / 72.0 without
        std::call_once(nap_once, [] { const char* e = getenv("GNNX_GEMM_BACKOFF"); nap = (e && e[0] == '1') ? 1 : 0; });
        if (nap) flags |= kBackoffFlag;
        const int Mi = (int)M, Ni = (int)N, Ki = (int)K;
        const long long ldc_ = ldc, ldm_ = ldm;
        const long long ldbits = (N + 31) / 32;          // packed masks are row-major: [M, ceil(N/32)] words
        float* csp = nullptr;                            // per-CTA partial column sums of A (fused bias gradient)
        int* tk = nullptr;                               // arrival counter: the last CTA folds the partials in-kernel
        if (colsum_a_out) {
            GNNX_REQUIRE(K <= kColsumMaxK, "gemm_tc: fused column sums cover K <= %d", kColsumMaxK);
            GNNX_REQUIRE(ws && ws_bytes >= (size_t)ncl * cluster * kCsWarps * K * sizeof(float), "gemm_tc: workspace too small for the column-sum partials");
            csp = (float*)ws;
            GNNX_REQUIRE(((uintptr_t)colsum_a_out & 15) == 0, "gemm_tc: colsum_a_out must be 16-byte aligned");
            tk = stream_ticket(st);
            if (!tk) return set_error(GNNX_ERR_INVALID, "gemm_tc: could not allocate the column-sum ticket");
        }
        cudaError_t e;
        // CTA pairs (cta_group::2): read per call so that a test can compare both kernels in one process
        const char* pe = getenv("GNNX_GEMM_PAIR");
        const bool pair = cluster == 2 && BN % 64 == 0 && pe && pe[0] == '1';
        if (trace_on())
            trace_before(pair ? (tb ? (const void*)k_gemm_tf32_pair<false> : (const void*)k_gemm_tf32_pair<true>)
                         : cluster == 4 ? (tb ? (const void*)k_gemm_tf32_rows2<false, 4> : (const void*)k_gemm_tf32_rows2<true, 4>)
                         : cluster == 2 ? (tb ? (const void*)k_gemm_tf32_rows2<false, 2> : (const void*)k_gemm_tf32_rows2<true, 2>)
                                      : (tb ? (const void*)k_gemm_tf32_rows2<false, 1> : (const void*)k_gemm_tf32_rows2<true, 1>), st);
        if (pair) {
            cfg.dynamicSmemBytes = kSmemP;
            e = tb ? cudaLaunchKernelEx(&cfg, k_gemm_tf32_pair<false>, mA, mB, mC, C, ldc_, Mi, Ni, Ki, BN, bias, flags, mask, ldm_, bits_out, ldbits, csp, bucket_ptr, n_groups, group_rows, colsum_a_out, tk)
                   : cudaLaunchKernelEx(&cfg, k_gemm_tf32_pair<true>, mA, mB, mC, C, ldc_, Mi, Ni, Ki, BN, bias, flags, mask, ldm_, bits_out, ldbits, csp, bucket_ptr, n_groups, group_rows, colsum_a_out, tk);
        } else if (cluster == 4)
            e = tb ? cudaLaunchKernelEx(&cfg, k_gemm_tf32_rows2<false, 4>, mA, mB, mC, C, ldc_, Mi, Ni, Ki, BN, bias, flags, mask, ldm_, bits_out, ldbits, csp, bucket_ptr, n_groups, group_rows, colsum_a_out, tk)
                   : cudaLaunchKernelEx(&cfg, k_gemm_tf32_rows2<true, 4>, mA, mB, mC, C, ldc_, Mi, Ni, Ki, BN, bias, flags, mask, ldm_, bits_out, ldbits, csp, bucket_ptr, n_groups, group_rows, colsum_a_out, tk);
        else if (cluster == 2)
            e = tb ? cudaLaunchKernelEx(&cfg, k_gemm_tf32_rows2<false, 2>, mA, mB, mC, C, ldc_, Mi, Ni, Ki, BN, bias, flags, mask, ldm_, bits_out, ldbits, csp, bucket_ptr, n_groups, group_rows, colsum_a_out, tk)
                   : cudaLaunchKernelEx(&cfg, k_gemm_tf32_rows2<true, 2>, mA, mB, mC, C, ldc_, Mi, Ni, Ki, BN, bias, flags, mask, ldm_, bits_out, ldbits, csp, bucket_ptr, n_groups, group_rows, colsum_a_out, tk);
        else
            e = tb ? cudaLaunchKernelEx(&cfg, k_gemm_tf32_rows2<false, 1>, mA, mB, mC, C, ldc_, Mi, Ni, Ki, BN, bias, flags, mask, ldm_, bits_out, ldbits, csp, bucket_ptr, n_groups, group_rows, colsum_a_out, tk)
                   : cudaLaunchKernelEx(&cfg, k_gemm_tf32_rows2<true, 1>, mA, mB, mC, C, ldc_, Mi, Ni, Ki, BN, bias, flags, mask, ldm_, bits_out, ldbits, csp, bucket_ptr, n_groups, group_rows, colsum_a_out, tk);
        if (trace_on()) trace_after(st);
        if (e != cudaSuccess) return set_error((int)e, "gemm_tf32_rows2 launch: %s", cudaGetErrorString(e));
        count_tc_launches(1);
        return check_launch("gemm_tf32_rows2");
    }
    // C[M,N] = A[K,M]^T B[K,N]: reduction over the K rows
    GNNX_REQUIRE(!bucket_ptr, "gemm_tc: grouped dW goes through gemm_tc_red_grouped");
    RedPlan p = red_plan(M, N, K);
    GNNX_REQUIRE(ws && ws_bytes >= (size_t)p.splits * M * N * sizeof(float), "gemm_tc: workspace too small for the row-split partials");
    if (make_map_mnmajor(&mA, A, K, M, lda, p.BM / 32)) return GNNX_ERR_INVALID;
    if (make_map_mnmajor(&mB, B, K, N, ldb, p.BN / 32)) return GNNX_ERR_INVALID;
    if (make_map_store(&mC, (const float*)ws, (int64_t)p.splits * M, N, N)) return GNNX_ERR_INVALID;
    // the ordered fold of the partials runs inside the same launch when the whole grid is resident (one CTA per SM) and
    // no mask is applied; GNNX_RED_FOLD=0: always the separate k_splitk_reduce launch (same sums, same order)
    int* ctr = nullptr;
    if (!mask && red_fold_enabled() && (long long)p.tiles * p.splits <= num_sms()) ctr = stream_ticket(st);
    const RedFold fold = {ctr ? C : nullptr, (long long)ldc, bias, flags, ctr ? ctr + 16 : nullptr};
    if (ctr) launch_coop(k_gemm_tf32_red, dim3(p.tiles, p.splits), kThreads, kSmemBytes, st, mA, mB, mC, (float*)ws, (int)M, (int)N, (int)K,
                         p.rows_per_split, p.BM, p.BN, nullptr, fold);
    else launch(k_gemm_tf32_red, dim3(p.tiles, p.splits), kThreads, kSmemBytes, st, mA, mB, mC, (float*)ws, (int)M, (int)N, (int)K,
                p.rows_per_split, p.BM, p.BN, nullptr, fold);
    count_tc_launches(1);
    int rc = check_launch("gemm_tf32_red");
    if (rc) return rc;
    if (!ctr) rc = splitk_reduce_launch((const float*)ws, p.splits, (int)M, (int)N, C, ldc, bias, flags, mask, ldm, st);
    if (rc == GNNX_OK && bits_out) rc = pack_relu_bits_launch(C, ldc, (int)M, (int)N, bits_out, st);
    return rc;
}


// Grouped dW: partial[z][Mo,No] = A[rows_z, Mo]^T B[rows_z, No] for the S row ranges of `split_table` ({begin, end,
// group} per split; every range lies inside one group and is a multiple of 32 rows; begin == end marks an unused
// split, whose partial is left unwritten).  The caller folds the partials with per-group coefficients.
int tc_red_grouped_splits(int64_t Mo, int64_t No, int64_t R) { return red_plan(Mo, No, R).splits; }

int gemm_tc_red_grouped(const float* A, int64_t lda, const float* B, int64_t ldb, int64_t Mo, int64_t No, int64_t R,
                        const int* split_table, int S, float* partial, cudaStream_t st) {
    static std::atomic<uint64_t> attrs_done{0};
    {
        int rc = ensure_attrs(attrs_done, []() -> int {
            GNNX_CUDA(cudaFuncSetAttribute(k_gemm_tf32_red, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
            return GNNX_OK;
        });
        if (rc) return rc;
    }
    RedPlan p = red_plan(Mo, No, R);
    CUtensorMap mA, mB, mC;
    if (make_map_mnmajor(&mA, A, R, Mo, lda, p.BM / 32)) return GNNX_ERR_INVALID;
    if (make_map_mnmajor(&mB, B, R, No, ldb, p.BN / 32)) return GNNX_ERR_INVALID;
    if (make_map_store(&mC, partial, (int64_t)S * Mo, No, No)) return GNNX_ERR_INVALID;
    const RedFold none = {nullptr, 0, nullptr, 0, nullptr};             // folded by k_group_reduce with per-group coefficients
    launch(k_gemm_tf32_red, dim3(p.tiles, S), kThreads, kSmemBytes, st, mA, mB, mC, partial, (int)Mo, (int)No, (int)R, 0, p.BM, p.BN, split_table, none);
    count_tc_launches(1);
    return check_launch("gemm_tf32_red(grouped)");
}

// n weight gradients C_z[Mo,No] = A_z[R,Mo]^T B_z[R,No] of one shape in one launch + one ordered reduce per problem.
bool tc_red_batched_ok(int n, int64_t Mo, int64_t No, int64_t R) {
    return n >= 2 && n <= kRedBatch && ceil_div(Mo, Mo > 128 ? 256 : 128) * ceil_div(No, 256) * n <= num_sms() && R >= 256;
}
static RedPlan red_plan_batched(int n, int64_t Mo, int64_t No, int64_t R) {
    RedPlan p = red_plan(Mo, No, R);
    int want = num_sms() / (p.tiles * n);
    if (want < 1) want = 1;
    int64_t max_splits = (R + 255) / 256;
    int splits = (int)(want < max_splits ? want : max_splits);
    int rps = ceil_div(ceil_div(R, splits), kBK) * kBK;
    if (rps < kBK) rps = kBK;
    p.rows_per_split = rps;
    p.splits = ceil_div(R, rps);
    return p;
}
size_t tc_red_batched_workspace_bytes(int n, int64_t Mo, int64_t No, int64_t R) {
    return (size_t)n * red_plan_batched(n, Mo, No, R).splits * Mo * No * sizeof(float);
}
int gemm_tc_red_batched(int n, const float* const* A, const int64_t* lda, const float* const* B, const int64_t* ldb,
                        float* const* C, const int64_t* ldc, int64_t Mo, int64_t No, int64_t R, void* ws, size_t ws_bytes,
                        cudaStream_t st) {
    static std::atomic<uint64_t> attrs_done{0};
    {
        int rc = ensure_attrs(attrs_done, []() -> int {
            GNNX_CUDA(cudaFuncSetAttribute(k_gemm_tf32_red_batched, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
            return GNNX_OK;
        });
        if (rc) return rc;
    }
    const RedPlan p = red_plan_batched(n, Mo, No, R);
    const long long stride = (long long)p.splits * Mo * No;
    GNNX_REQUIRE(ws && ws_bytes >= (size_t)n * stride * sizeof(float), "gemm_dw_batched: workspace too small");
    RedBatch bt;
    for (int z = 0; z < n; ++z) {
        if (make_map_mnmajor(&bt.a[z], A[z], R, Mo, lda[z], p.BM / 32)) return GNNX_ERR_INVALID;
        if (make_map_mnmajor(&bt.b[z], B[z], R, No, ldb[z], p.BN / 32)) return GNNX_ERR_INVALID;
        if (make_map_store(&bt.p[z], (const float*)ws + z * stride, (int64_t)p.splits * Mo, No, No)) return GNNX_ERR_INVALID;
    }
    for (int z = n; z < kRedBatch; ++z) { bt.a[z] = bt.a[0]; bt.b[z] = bt.b[0]; bt.p[z] = bt.p[0]; }
    int* ctr = nullptr;
    if (red_fold_enabled() && (long long)p.tiles * p.splits * n <= num_sms()) ctr = stream_ticket(st);
    for (int z = 0; z < kRedBatch; ++z) { bt.c[z] = C[z < n ? z : 0]; bt.ldc[z] = ldc[z < n ? z : 0]; }
    bt.counters = ctr ? ctr + 16 : nullptr;
    if (ctr) launch_coop(k_gemm_tf32_red_batched, dim3(p.tiles, p.splits, n), kThreads, kSmemBytes, st, bt, (float*)ws, stride, (int)Mo,
                         (int)No, (int)R, p.rows_per_split, p.BM, p.BN);
    else launch(k_gemm_tf32_red_batched, dim3(p.tiles, p.splits, n), kThreads, kSmemBytes, st, bt, (float*)ws, stride, (int)Mo,
                (int)No, (int)R, p.rows_per_split, p.BM, p.BN);
    count_tc_launches(1);
    int rc = check_launch("gemm_tf32_red(batched)");
    for (int z = 0; z < n && rc == GNNX_OK && !ctr; ++z)
        rc = splitk_reduce_launch((const float*)ws + z * stride, p.splits, (int)Mo, (int)No, C[z], ldc[z], nullptr, 0, nullptr, 0, st);
    return rc;
}

// ------------------------------------------------------------------ Linear chains (forward)
bool tc_chain_ok(int n, int64_t M, int64_t W, const float* A, int64_t lda, const float* const* Wt, const int64_t* ldw,
                 float* const* C, const int64_t* ldc) {
    if (!encode_fn() || n < 2 || n > kChainMax) return false;
    if (W < 32 || W > 256 || W % 32 || M < 768) return false;
    if (!aligned16(A) || (lda & 3)) return false;
    for (int p = 0; p < n; ++p)
        if (!aligned16(Wt[p]) || !aligned16(C[p]) || (ldw[p] & 3) || (ldc[p] & 3)) return false;
    return true;
}

int gemm_tc_chain(int n, const float* A, int64_t lda, const float* const* Wt, const int64_t* ldw, const float* const* bias,
                  const int* flags, float* const* C, const int64_t* ldc, uint32_t* const* bits, int64_t M, int64_t W,
                  cudaStream_t st) {
    static std::atomic<uint64_t> attrs_done{0};
    {
        int rc = ensure_attrs(attrs_done, []() -> int {
            GNNX_CUDA(cudaFuncSetAttribute(k_gemm_tf32_chain<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemChain));
            return GNNX_OK;
        });
        if (rc) return rc;
    }
    constexpr int cluster = 2;
    ChainArgs args;
    memset(&args, 0, sizeof args);
    args.n = n;
    for (int p = 0; p < n; ++p) {
        const float* a = p == 0 ? A : C[p - 1];
        const int64_t la = p == 0 ? lda : ldc[p - 1];
        if (make_map_kmajor(&args.a[p], a, M, W, la, 128)) return GNNX_ERR_INVALID;
        if (make_map_kmajor(&args.b[p], Wt[p], W, W, ldw[p], (int)W / cluster)) return GNNX_ERR_INVALID;
        if (make_map_store(&args.c[p], C[p], M, W, ldc[p])) return GNNX_ERR_INVALID;
        args.bias[p] = bias ? bias[p] : nullptr;
        args.bits[p] = bits ? bits[p] : nullptr;
        args.flags[p] = flags ? flags[p] : 0;
    }
    for (int p = n; p < kChainMax; ++p) { args.a[p] = args.a[0]; args.b[p] = args.b[0]; args.c[p] = args.c[0]; }
    const int m_groups = ceil_div(ceil_div(M, 128), cluster);
    const int max_cl = num_sms() / cluster;
    int ncl = m_groups < max_cl ? m_groups : max_cl;
    if (m_groups > max_cl) ncl = ceil_div(m_groups, ceil_div(m_groups, max_cl));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(ncl * cluster);
    cfg.blockDim = dim3(kThreads2);
    cfg.dynamicSmemBytes = kSmemChain;
    cfg.stream = st;
    cudaLaunchAttribute at[2];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = cluster; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = pdl_enabled() ? 2 : 1;
    static int group = 0;
    // tiles per group: 6 = all row tiles of a cluster at cfg2 (138 us per 3-product chain against 142 at 3 or 4; three separate
    // launches: 157).  GNNX_CHAIN_GROUP overrides.
    if (group == 0) { const char* e = getenv("GNNX_CHAIN_GROUP"); group = e ? atoi(e) : 6; if (group < 1 || group > kChainGroupMax) group = 6; }
    if (trace_on()) trace_before((const void*)k_gemm_tf32_chain<2>, st);
    cudaError_t e = cudaLaunchKernelEx(&cfg, k_gemm_tf32_chain<2>, args, (int)M, (int)W, group);
    if (trace_on()) trace_after(st);
    if (e != cudaSuccess) return set_error((int)e, "gemm_tf32_chain launch: %s", cudaGetErrorString(e));
    count_tc_launches(1);
    return check_launch("gemm_tf32_chain");
}

// ------------------------------------------------------------------ fused aggregate -> grouped post Linear (forward)
bool pna_post_fused_ok(int64_t rows, int64_t Ho, int64_t H, int64_t F, int64_t ldw, int64_t ld_out, int64_t ld_agg,
                       const void* weff, const void* out, const void* agg_out, const void* msg, const void* x) {
    if (!encode_fn()) return false;
    if (rows < 256 || rows % 256 || rows > 0x7fffffff) return false;        // buckets are padded to 2-CTA row blocks
    if (H < 32 || H % 32 || F < 32 || F % 32 || H % F) return false;        // 32-column chunks never straddle a tower block
    if (Ho < 16 || Ho % 16 || (Ho > 256 && Ho % 256)) return false;
    if ((ldw & 3) || (ld_out & 3) || (ld_agg & 3)) return false;
    return aligned16(weff) && aligned16(out) && aligned16(agg_out) && aligned16(msg) && aligned16(x);
}

int pna_post_fused(const float* msg, const int* rowptr, const int* row_node, int64_t rows, const float* x, int64_t H, int64_t F,
                   const float* weff, int64_t ldw, const int* bucket_ptr, int n_groups, const float* bias, int flags,
                   float* out, int64_t ld_out, int64_t Ho, uint32_t* bits_out, float* agg_out, int64_t ld_agg, cudaStream_t st) {
    static std::atomic<uint64_t> attrs_done{0};
    {
        int rc = ensure_attrs(attrs_done, []() -> int {
            GNNX_CUDA(cudaFuncSetAttribute(k_pna_post_fused<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem2));
            return GNNX_OK;
        });
        if (rc) return rc;
    }
    static std::once_flag rm_once;
    // how the producers round fp32 -> tf32: 1 = cvt.rn (what the TMA's TFLOAT32 tensor maps do: measured, the fused
    // product then equals the two-kernel product up to the K order of the accumulation), 0 = cvt.rna, 2 = leave the low
    // mantissa bits for the tensor core to drop (GNNX_FUSED_ROUND: diagnostic)
    static int round_mode = 1;
    std::call_once(rm_once, [] {
        const char* e = getenv("GNNX_FUSED_ROUND");
        if (e && e[0] >= '0' && e[0] <= '2') round_mode = e[0] - '0';
        const char* d = getenv("GNNX_FUSED_DBG");           // diagnostic (wrong results): 16 no message loads, 32 no reduction, 64 no operand copy
        if (d) round_mode |= atoi(d) & 8176;
    });
    const int BN = (int)(Ho >= 256 ? 256 : Ho);
    constexpr int cluster = 2;
    CUtensorMap mB, mC;
    if (make_map_kmajor(&mB, weff, Ho * n_groups, 5 * H, ldw, BN / cluster)) return GNNX_ERR_INVALID;
    if (make_map_store(&mC, out, rows, Ho, ld_out)) return GNNX_ERR_INVALID;
    const int m_tiles = ceil_div(rows, 128);
    const int work = ceil_div(m_tiles, cluster) * ceil_div(Ho, BN);
    const int max_cl = num_sms() / cluster;
    int ncl = work < max_cl ? work : max_cl;
    if (work > max_cl) ncl = ceil_div(work, ceil_div(work, max_cl));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(ncl * cluster);
    cfg.blockDim = dim3(kThreadsF);
    cfg.dynamicSmemBytes = kSmem2;
    cfg.stream = st;
    cudaLaunchAttribute at[2];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = cluster; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = pdl_enabled() ? 2 : 1;
    const long long ldbits = (Ho + 31) / 32, ld_agg_ = ld_agg;
    if (trace_on()) trace_before((const void*)k_pna_post_fused<2>, st);
    cudaError_t e = cudaLaunchKernelEx(&cfg, k_pna_post_fused<2>, mB, mC, msg, rowptr, row_node, x, (int)rows, (int)Ho, (int)H, (int)F,
                                       BN, bias, flags, bits_out, ldbits, bucket_ptr, n_groups, agg_out, ld_agg_, round_mode);
    if (trace_on()) trace_after(st);
    if (e != cudaSuccess) return set_error((int)e, "pna_post_fused launch: %s", cudaGetErrorString(e));
    count_tc_launches(1);
    return check_launch("pna_post_fused");
}

}  // namespace gnnx
