// TransformerConv attention over the dst-sorted CSR (reference models.py:497-510):
//   logit_e = <q_i, k_j + e_e>_h / sqrt(C),  alpha = segment softmax over the incoming edges of i (R8),
//   agg_i = sum_e alpha_e (v_j + e_e),  out_i = agg_i + lin_skip(x)_i.
// One warp owns one (destination row, head group) pair (attn.cuh); two edges are gathered per iteration so that
// six independent 16-byte loads per lane are in flight.  Deterministic and atomic-free: every output element is
// written by exactly one thread, sums run in CSR order.
#include "attn.cuh"

namespace gnnx {

template <int NQ, typename EIdx>
__global__ void __launch_bounds__(256, NQ == 1 ? 4 : (NQ == 2 ? 2 : 1))
k_tconv_fwd(const float* __restrict__ qkv, const float* __restrict__ ee, long long ld_e, const EIdx* __restrict__ eidx,
            const int* __restrict__ rowptr, const int* __restrict__ csr_src, int N, int heads, int C, int log2c, int groups,
            float scale, unsigned int drop_thr, float keep_scale, unsigned long long seed, const float* skip,
            float* __restrict__ agg, float* out, float* alpha) {
    pdl_enter();
    const int H = heads * C;
    const int lane = threadIdx.x & 31;
    const int wpb = blockDim.x >> 5;
    const long long ld = 3ll * H;
    const int hpw = (NQ > 1 || C >= 128) ? 1 : 128 / C;          // heads per warp
    const int lph = 32 / hpw;                                    // lanes per head
    const int lih = lane & (lph - 1);                            // lane index inside its head
    const int gw = blockIdx.x * wpb + (threadIdx.x >> 5);        // this warp's head group never changes: no index
    const int row_stride = (gridDim.x * wpb) / groups;           // division in the row loop (host: warps % groups == 0)
    const int cb = (gw % groups) * 128 * NQ;
    const int c0 = cb + lane * 4;
    const bool act = c0 < H;
    const int head = act ? c0 >> log2c : 0;
    const bool writer = act && (NQ > 1 ? lane == 0 : (c0 & (C - 1)) == 0);
    int i = gw / groups;
    int nb = 0, ne = 0;
    if (i < N) { nb = rowptr[i]; ne = rowptr[i + 1]; }
    for (; i < N; i += row_stride) {
        const int b = nb, e = ne;
        if (i + row_stride < N) {                                // next row's bounds: latency hidden behind this row
            nb = rowptr[i + row_stride]; ne = rowptr[i + row_stride + 1];
        }
        const int nslots = e - b;
        const bool spill = nslots > lph;                         // logits stay in registers (lane t of a head <- slot t)
        float mylg = 0.f;                                        // unless the row has more slots than a head has lanes
        F4 qv[NQ], acc[NQ];
#pragma unroll
        for (int q = 0; q < NQ; ++q) {
            qv[q] = act ? g4(qkv + (long long)i * ld + c0 + 128 * q) : z4();
#pragma unroll
            for (int j = 0; j < 4; ++j) qv[q].a[j] *= scale;
            acc[q] = z4();
        }
        float mx = -INFINITY, den = 0.f;
        auto step = [&](const F4 (&k)[NQ], const F4 (&v)[NQ], const F4 (&ev)[NQ], int slot) {
            float part = 0.f;
#pragma unroll
            for (int q = 0; q < NQ; ++q)
#pragma unroll
                for (int j = 0; j < 4; ++j) part = fmaf(qv[q].a[j], k[q].a[j] + ev[q].a[j], part);
            const float lg = head_sum<NQ>(part, C);
            if (spill) { if (writer) alpha[(long long)slot * heads + head] = lg; }      // raw logit, normalised below
            else if (lih == slot - b) mylg = lg;
            const float dl = lg - mx;                                    // one of the two rescale factors is exp(0):
            const float ex = expf(-fabsf(dl));                           // a single exp (dl = +inf on the first edge -> 0)
            const float nm = fmaxf(mx, lg);
            const float sc = dl > 0.f ? ex : 1.f;
            const float wgt = dl > 0.f ? 1.f : ex;
            den = den * sc + wgt;
            const float wk = (drop_thr == 0u || attn_keep(seed, (long long)slot * heads + head, drop_thr)) ? wgt : 0.f;
#pragma unroll
            for (int q = 0; q < NQ; ++q)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[q].a[j] = acc[q].a[j] * sc + wk * (v[q].a[j] + ev[q].a[j]);
            mx = nm;
        };
        for (int t = b; t < e; t += 2) {
            const bool two = t + 1 < e;
            const int s0 = csr_src[t], s1 = two ? csr_src[t + 1] : s0;
            const long long r0 = (long long)eidx[t] * ld_e, r1 = two ? (long long)eidx[t + 1] * ld_e : r0;
            F4 k0[NQ], v0[NQ], e0[NQ], k1[NQ], v1[NQ], e1[NQ];
#pragma unroll
            for (int q = 0; q < NQ; ++q) {
                const int c = c0 + 128 * q;
                if (act) {
                    k0[q] = g4(qkv + (long long)s0 * ld + H + c);
                    v0[q] = g4(qkv + (long long)s0 * ld + 2 * H + c);
                    e0[q] = g4(ee + r0 + c);
                    k1[q] = g4(qkv + (long long)s1 * ld + H + c);
                    v1[q] = g4(qkv + (long long)s1 * ld + 2 * H + c);
                    e1[q] = g4(ee + r1 + c);
                } else {
                    k0[q] = v0[q] = e0[q] = k1[q] = v1[q] = e1[q] = z4();
                }
            }
            step(k0, v0, e0, t);
            if (two) step(k1, v1, e1, t + 1);
        }
        const float inv = 1.f / (den + 1e-16f);
#pragma unroll
        for (int q = 0; q < NQ; ++q) {
            const int c = c0 + 128 * q;
            if (act) {
                F4 a;
#pragma unroll
                for (int j = 0; j < 4; ++j) a.a[j] = acc[q].a[j] * (inv * keep_scale);
                stg4(agg + (long long)i * H + c, a.q);
                if (out) {
                    if (skip) {
                        const float4 s = *reinterpret_cast<const float4*>(skip + (long long)i * H + c);   // may alias out
                        a.a[0] += s.x; a.a[1] += s.y; a.a[2] += s.z; a.a[3] += s.w;
                    }
                    stg4(out + (long long)i * H + c, a.q);
                }
            }
        }
        // alpha = exp(logit - max) / (sum + 1e-16), negated where attention dropout removed the entry
        if (!spill) {
            if (act && lih < nslots) {
                const int slot = b + lih;
                float a = expf(mylg - mx) * inv;
                if (!(drop_thr == 0u || attn_keep(seed, (long long)slot * heads + head, drop_thr))) a = -a;
                alpha[(long long)slot * heads + head] = a;
            }
            continue;
        }
        __syncwarp();                                            // long rows: lanes spread over the (slot, head) pairs
        const int npairs = nslots * hpw;
        for (int base = 0; base < npairs; base += 32) {
            const int idx = base + lane;
            const int hl = idx % hpw, t = idx / hpw;
            const int from = (hl * (C >> 2)) & 31;
            const float hmx = __shfl_sync(0xffffffffu, mx, from), hinv = __shfl_sync(0xffffffffu, inv, from);
            const int hd = (cb >> log2c) + hl;
            if (idx < npairs && hd < heads) {
                float* ap = alpha + (long long)(b + t) * heads + hd;
                const float a = expf(*ap - hmx) * hinv;
                const bool keep = drop_thr == 0u || attn_keep(seed, (long long)(b + t) * heads + hd, drop_thr);
                *ap = keep ? a : -a;
            }
        }
    }
}

// Backward, destination side: dq rows, and per slot dk (= d logit * q / sqrt(C)) and the edge-term gradient
// de = dk + alpha * dout_i.  S_h = sum_e alpha_e <dout_i, v_j + e_e>_h equals <dout_i, agg_i>_h, so no extra pass.
template <int NQ, typename EIdx>
__global__ void __launch_bounds__(256, NQ == 1 ? 4 : (NQ == 2 ? 2 : 1))
k_tconv_bwd_dst(const float* __restrict__ dout, const float* __restrict__ qkv, const float* __restrict__ ee, long long ld_e,
                const EIdx* __restrict__ eidx, const float* __restrict__ alpha, const float* __restrict__ agg,
                const int* __restrict__ rowptr, const int* __restrict__ csr_src, const int* __restrict__ csr_eid, int N,
                int heads, int C, int log2c, int groups, float scale, float keep_scale, float* __restrict__ dqkv,
                float* __restrict__ ds, float* __restrict__ de) {
    pdl_enter();
    const int H = heads * C;
    const int lane = threadIdx.x & 31;
    const int wpb = blockDim.x >> 5;
    const long long ld = 3ll * H;
    const int gw = blockIdx.x * wpb + (threadIdx.x >> 5);
    const int row_stride = (gridDim.x * wpb) / groups;           // host guarantees divisibility
    const int cb = (gw % groups) * 128 * NQ;
    const int c0 = cb + lane * 4;
    const bool act = c0 < H;
    const int head = act ? c0 >> log2c : 0;
    int i = gw / groups;
    int nb = 0, ne = 0;
    if (i < N) { nb = rowptr[i]; ne = rowptr[i + 1]; }
    for (; i < N; i += row_stride) {
        const int b = nb, e = ne;
        if (i + row_stride < N) { nb = rowptr[i + row_stride]; ne = rowptr[i + row_stride + 1]; }
        F4 go[NQ], qv[NQ], dq[NQ];
        float sp = 0.f;
#pragma unroll
        for (int q = 0; q < NQ; ++q) {
            const int c = c0 + 128 * q;
            go[q] = act ? g4(dout + (long long)i * H + c) : z4();
            qv[q] = act ? g4(qkv + (long long)i * ld + c) : z4();
            F4 ag = act ? g4(agg + (long long)i * H + c) : z4();
            dq[q] = z4();
#pragma unroll
            for (int j = 0; j < 4; ++j) sp = fmaf(go[q].a[j], ag.a[j], sp);
        }
        const float S = head_sum<NQ>(sp, C);
        auto step = [&](const F4 (&k)[NQ], const F4 (&v)[NQ], const F4 (&ev)[NQ], float stored, int slot) {
            float part = 0.f;
#pragma unroll
            for (int q = 0; q < NQ; ++q)
#pragma unroll
                for (int j = 0; j < 4; ++j) part = fmaf(go[q].a[j], v[q].a[j] + ev[q].a[j], part);
            const float kf = attn_dropped(stored) ? 0.f : keep_scale;     // d(alpha after dropout) / d alpha
            const float asm_ = fabsf(stored);                             // softmax output
            const float a = asm_ * kf;                                    // weight the forward used
            const float dlg = asm_ * (kf * head_sum<NQ>(part, C) - S) * scale;
            const long long drow = csr_eid ? csr_eid[slot] : slot;
#pragma unroll
            for (int q = 0; q < NQ; ++q) {
                const int c = c0 + 128 * q;
                F4 dk, dd;
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    dq[q].a[j] = fmaf(dlg, k[q].a[j] + ev[q].a[j], dq[q].a[j]);
                    dk.a[j] = dlg * qv[q].a[j];
                    dd.a[j] = fmaf(a, go[q].a[j], dk.a[j]);
                }
                if (act) {
                    stg4(ds + (long long)slot * H + c, dk.q);
                    stg4(de + drow * H + c, dd.q);
                }
            }
        };
        for (int t = b; t < e; t += 2) {
            const bool two = t + 1 < e;
            const int s0 = csr_src[t], s1 = two ? csr_src[t + 1] : s0;
            const long long r0 = (long long)eidx[t] * ld_e, r1 = two ? (long long)eidx[t + 1] * ld_e : r0;
            const float a0 = alpha[(long long)t * heads + head];
            const float a1 = two ? alpha[(long long)(t + 1) * heads + head] : 0.f;
            F4 k0[NQ], v0[NQ], e0[NQ], k1[NQ], v1[NQ], e1[NQ];
#pragma unroll
            for (int q = 0; q < NQ; ++q) {
                const int c = c0 + 128 * q;
                if (act) {
                    k0[q] = g4(qkv + (long long)s0 * ld + H + c);
                    v0[q] = g4(qkv + (long long)s0 * ld + 2 * H + c);
                    e0[q] = g4(ee + r0 + c);
                    k1[q] = g4(qkv + (long long)s1 * ld + H + c);
                    v1[q] = g4(qkv + (long long)s1 * ld + 2 * H + c);
                    e1[q] = g4(ee + r1 + c);
                } else {
                    k0[q] = v0[q] = e0[q] = k1[q] = v1[q] = e1[q] = z4();
                }
            }
            step(k0, v0, e0, a0, t);
            if (two) step(k1, v1, e1, a1, t + 1);
        }
#pragma unroll
        for (int q = 0; q < NQ; ++q) {
            const int c = c0 + 128 * q;
            if (act) stg4(dqkv + (long long)i * ld + c, dq[q].q);
        }
    }
}

// Backward, source side: dk_j = sum over out-edges of ds[slot]; dv_j = sum alpha[slot] * dout[dst].  Warp per
// (source row, 128-column chunk); no head coupling here.
__global__ void __launch_bounds__(256)
k_tconv_bwd_src(const float* __restrict__ dout, const float* __restrict__ ds, const float* __restrict__ alpha,
                const int* __restrict__ colptr, const int* __restrict__ csc_pos, const int* __restrict__ csc_dst, int N,
                int heads, int C, int log2c, float keep_scale, float* __restrict__ dqkv) {
    pdl_enter();
    const int H = heads * C;
    const int lane = threadIdx.x & 31;
    const int wpb = blockDim.x >> 5;
    const int chunks = (H + 127) / 128;
    const long long ld = 3ll * H;
    const int gw = blockIdx.x * wpb + (threadIdx.x >> 5);        // fixed column chunk per warp, rows strided
    const int row_stride = (gridDim.x * wpb) / chunks;           // host guarantees divisibility
    const int c = (gw % chunks) * 128 + lane * 4;
    if (c >= H) return;
    const int head = c >> log2c;
    for (int j = gw / chunks; j < N; j += row_stride) {
        const int b = colptr[j], e = colptr[j + 1];
        F4 dk = z4(), dv = z4();
        for (int p = b; p < e; p += 2) {
            const bool two = p + 1 < e;
            const int p0 = csc_pos[p], d0 = csc_dst[p];
            const int p1 = two ? csc_pos[p + 1] : p0, d1 = two ? csc_dst[p + 1] : d0;
            const F4 x0 = g4(ds + (long long)p0 * H + c), g0 = g4(dout + (long long)d0 * H + c);
            const F4 x1 = g4(ds + (long long)p1 * H + c), g1 = g4(dout + (long long)d1 * H + c);
            const float r0 = alpha[(long long)p0 * heads + head];
            const float r1 = two ? alpha[(long long)p1 * heads + head] : 0.f;
            const float a0 = attn_dropped(r0) ? 0.f : r0 * keep_scale, a1 = attn_dropped(r1) ? 0.f : r1 * keep_scale;
#pragma unroll
            for (int jj = 0; jj < 4; ++jj) {
                dk.a[jj] += x0.a[jj];
                dv.a[jj] = fmaf(a0, g0.a[jj], dv.a[jj]);
            }
            if (two) {
#pragma unroll
                for (int jj = 0; jj < 4; ++jj) {
                    dk.a[jj] += x1.a[jj];
                    dv.a[jj] = fmaf(a1, g1.a[jj], dv.a[jj]);
                }
            }
        }
        stg4(dqkv + (long long)j * ld + H + c, dk.q);
        stg4(dqkv + (long long)j * ld + 2 * H + c, dv.q);
    }
}

}  // namespace gnnx

using namespace gnnx;

#define TCONV_DISPATCH(S, CALL)                                              \
    do {                                                                     \
        if ((S).nq == 1) { CALL(1); }                                        \
        else if ((S).nq == 2) { CALL(2); }                                   \
        else if ((S).nq == 4) { CALL(4); }                                   \
        else { CALL(8); }                                                    \
    } while (0)

extern "C" {

int gnnx_tconv_fwd(const float* qkv, const float* e, int64_t ld_e, const int32_t* eidx_i32, const uint8_t* eidx_u8,
                   const int32_t* rowptr, const int32_t* csr_src, int64_t N, int64_t E, int32_t heads, int32_t C,
                   float dropout_p, int64_t seed, const float* skip, float* agg, float* out, float* alpha, void* stream) {
    (void)E;
    GNNX_REQUIRE(dropout_p >= 0.f && dropout_p < 1.f, "tconv_fwd: dropout_p must be in [0, 1)");
    GNNX_REQUIRE(attn_shape_ok(heads, C), "tconv_fwd: need C a power of two in [4, 1024] and heads*C <= 4096");
    GNNX_REQUIRE(E == 0 || (eidx_i32 != nullptr) != (eidx_u8 != nullptr), "tconv_fwd: exactly one of eidx_i32 / eidx_u8");
    GNNX_REQUIRE(ld_e % 4 == 0, "tconv_fwd: ld_e must be a multiple of 4");
    if (N == 0) return GNNX_OK;
    const AttnShape s = attn_shape(heads, C);
    const float scale = 1.0f / sqrtf((float)C);
    cudaStream_t st = as_stream(stream);
    const int grid = attn_row_grid(N, s.groups, 16);
    const unsigned int thr = attn_drop_threshold(dropout_p);
    const float ks = 1.0f / (1.0f - dropout_p);
    const unsigned long long sd = (unsigned long long)seed;
#define CALL(NQ)                                                                                                          \
    if (eidx_i32) launch(k_tconv_fwd<NQ, int>, grid, 256, 0, st, qkv, e, ld_e, eidx_i32, rowptr, csr_src, (int)N, heads, C, ilog2(C), s.groups, scale, thr, ks, sd, skip, agg, out, alpha); \
    else launch(k_tconv_fwd<NQ, uint8_t>, grid, 256, 0, st, qkv, e, ld_e, eidx_u8, rowptr, csr_src, (int)N, heads, C, ilog2(C), s.groups, scale, thr, ks, sd, skip, agg, out, alpha)
    TCONV_DISPATCH(s, CALL);
#undef CALL
    return check_launch("tconv_fwd");
}

int gnnx_tconv_bwd(const float* dout, const float* qkv, const float* e, int64_t ld_e, const int32_t* eidx_i32,
                   const uint8_t* eidx_u8, const float* alpha, const float* agg, const int32_t* rowptr,
                   const int32_t* csr_src, const int32_t* csr_eid, const int32_t* colptr, const int32_t* csc_pos,
                   const int32_t* csc_dst, int64_t N, int64_t E, int32_t heads, int32_t C, float dropout_p, float* dqkv,
                   float* ds, float* de, void* stream) {
    (void)E;
    GNNX_REQUIRE(dropout_p >= 0.f && dropout_p < 1.f, "tconv_bwd: dropout_p must be in [0, 1)");
    GNNX_REQUIRE(attn_shape_ok(heads, C), "tconv_bwd: need C a power of two in [4, 1024] and heads*C <= 4096");
    GNNX_REQUIRE(E == 0 || (eidx_i32 != nullptr) != (eidx_u8 != nullptr), "tconv_bwd: exactly one of eidx_i32 / eidx_u8");
    GNNX_REQUIRE(ld_e % 4 == 0, "tconv_bwd: ld_e must be a multiple of 4");
    if (N == 0) return GNNX_OK;
    const AttnShape s = attn_shape(heads, C);
    const float scale = 1.0f / sqrtf((float)C);
    cudaStream_t st = as_stream(stream);
    const int grid = attn_row_grid(N, s.groups, 16);
    const float ks = 1.0f / (1.0f - dropout_p);
#define CALL(NQ)                                                                                                          \
    if (eidx_i32) launch(k_tconv_bwd_dst<NQ, int>, grid, 256, 0, st, dout, qkv, e, ld_e, eidx_i32, alpha, agg, rowptr, csr_src, csr_eid, (int)N, heads, C, ilog2(C), s.groups, scale, ks, dqkv, ds, de); \
    else launch(k_tconv_bwd_dst<NQ, uint8_t>, grid, 256, 0, st, dout, qkv, e, ld_e, eidx_u8, alpha, agg, rowptr, csr_src, csr_eid, (int)N, heads, C, ilog2(C), s.groups, scale, ks, dqkv, ds, de)
    TCONV_DISPATCH(s, CALL);
#undef CALL
    launch(k_tconv_bwd_src, attn_row_grid(N, (s.H + 127) / 128, 16), 256, 0, st, dout, ds, alpha, colptr, csc_pos, csc_dst, (int)N, heads, C, ilog2(C), ks, dqkv);
    return check_launch("tconv_bwd", 2);
}

}  // extern "C"
