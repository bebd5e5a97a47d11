// Shared pieces of the two segment-softmax-aggregate kernels (gatv2.cu, tconv.cu).
//
// Work decomposition: the H = heads*C feature columns of a destination row are cut into GROUPS that never split a
// head -- one 128-column chunk when C <= 128 (the chunk then holds 128/C whole heads), the C/128 chunks of ONE head
// when C > 128 -- and one warp owns one (row, group) pair: the softmax of a head needs nothing from other heads, so
// groups are independent work items.  A lane holds 4 consecutive columns of each of its NQ chunks.
#pragma once
#include "common.cuh"

namespace gnnx {

union F4 {
    float4 q;
    float a[4];
};
__device__ __forceinline__ F4 g4(const float* p) { F4 r; r.q = ldg4(p); return r; }
__device__ __forceinline__ F4 z4() { F4 r; r.q = make_float4(0.f, 0.f, 0.f, 0.f); return r; }

struct AttnShape {
    int heads, C, H;
    int nq;          // 128-column chunks per group (1 when C <= 128, else C / 128)
    int groups;      // groups per row
};
static inline AttnShape attn_shape(int heads, int C) {
    AttnShape s;
    s.heads = heads; s.C = C; s.H = heads * C;
    s.nq = C > 128 ? C / 128 : 1;
    s.groups = (s.H + 128 * s.nq - 1) / (128 * s.nq);
    return s;
}
static inline bool attn_shape_ok(int heads, int C) {
    return C >= 4 && (C & (C - 1)) == 0 && C <= 1024 && heads >= 1 && (long long)heads * C <= 4096;
}

// Per-lane partial dot product -> the full dot product of the lane's head, on every lane of that head.
// NQ > 1: the whole warp (all chunks) is one head.  NQ == 1: a head is C/4 consecutive lanes.
template <int NQ>
__device__ __forceinline__ float head_sum(float part, int C) {
    if (NQ > 1 || C >= 128) return warp_sum(part);
    for (int o = C >> 3; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
    return part;
}

// Attention dropout (F.dropout on alpha, training mode): one counter-based draw per (slot, head), a pure function of
// (seed, slot * heads + head) so that forward and backward agree without storing a mask -- the forward still records
// the outcome in the SIGN of the saved alpha (alpha >= 0; dropped entries are stored negated, -0.0 included).
__device__ __forceinline__ bool attn_keep(unsigned long long seed, long long index, unsigned int threshold) {
    return dropout_keep(seed, index, threshold);          // common.cuh (shared with the node-feature dropout)
}
static inline unsigned int attn_drop_threshold(float p) { return dropout_threshold(p); }
__device__ __forceinline__ bool attn_dropped(float stored_alpha) { return (__float_as_uint(stored_alpha) >> 31) != 0u; }

// Grid for the destination-side kernels: every warp keeps ONE head group for its whole life and strides over rows
// (no index division in the row loop), so the number of warps must be a multiple of `groups`.
static inline int attn_row_grid(long long N, int groups, int blocks_per_sm) {
    long long blocks = (N * groups + 7) / 8;
    long long cap = (long long)num_sms() * blocks_per_sm;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return (int)((blocks + groups - 1) / groups * groups);
}
static inline int ilog2(int v) { int l = 0; while ((1 << l) < v) ++l; return l; }

// grid for a warp-per-work-item grid-stride kernel with 8 warps per block
static inline int attn_grid(long long work) {
    long long blocks = (work + 7) / 8;
    long long cap = (long long)num_sms() * 16;
    return (int)(blocks < cap ? (blocks > 0 ? blocks : 1) : cap);
}

}  // namespace gnnx
