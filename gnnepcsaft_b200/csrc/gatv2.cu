// GATv2 attention over the dst-sorted CSR: additive attention logits, segment
// softmax over each destination (online, single gather of x_l[src] per edge), weighted aggregation,
// self loops synthesised in-kernel (loop edge term = mean of the incoming edge terms, which equals
// lin_edge(mean(edge_attr)) because lin_edge has no bias), attention dropout, and the matching backward.
// One warp owns one (destination row, head group) pair (attn.cuh) and gathers two edges per iteration.
// Deterministic and atomic-free: the only cross-row reduction (d att) goes through ordered partials.
#include "attn.cuh"
#include <cstdlib>

namespace gnnx {

template <int NQ, typename EIdx>
__global__ void __launch_bounds__(256, NQ == 1 ? 4 : (NQ == 2 ? 2 : 1))
k_gatv2_fwd(const float* __restrict__ xlr, const float* __restrict__ ee, long long ld_e, const EIdx* __restrict__ eidx,
            const float* __restrict__ att, const float* __restrict__ bias, const int* __restrict__ rowptr,
            const int* __restrict__ csr_src, int N, int E, int heads, int C, int log2c, int groups, float slope,
            int self_loops, unsigned int drop_thr, float keep_scale, unsigned long long seed, float* __restrict__ out, float* alpha) {
    pdl_enter();
    const int H = heads * C;
    const int lane = threadIdx.x & 31;
    const int wpb = blockDim.x >> 5;
    const long long ld = 2ll * H;
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
        const int nreal = e - b;
        const int nslots = nreal + (self_loops ? 1 : 0);
        const bool spill = nslots > lph;                         // logits kept in registers (lane t of a head <- slot t) unless
        float mylg = 0.f;                                        // the row has more slots than a head has lanes
        F4 xr[NQ], av[NQ], acc[NQ], esum[NQ], xself[NQ];
#pragma unroll
        for (int q = 0; q < NQ; ++q) {
            const int c = c0 + 128 * q;
            xr[q] = act ? g4(xlr + (long long)i * ld + H + c) : z4();
            xself[q] = (act && self_loops) ? g4(xlr + (long long)i * ld + c) : z4();
            av[q] = act ? g4(att + c) : z4();
            acc[q] = z4(); esum[q] = z4();
        }
        float mx = -INFINITY, den = 0.f;
        int cnt = 0;
        auto step = [&](const F4 (&xl)[NQ], const F4 (&ev)[NQ], int slot, int tt) {
            float part = 0.f;
#pragma unroll
            for (int q = 0; q < NQ; ++q)
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    float s = xr[q].a[j] + xl[q].a[j] + ev[q].a[j];
                    s = s > 0.f ? s : s * slope;
                    part = fmaf(s, av[q].a[j], part);
                }
            const float lg = head_sum<NQ>(part, C);
            if (spill) { if (writer) alpha[(long long)slot * heads + head] = lg; }      // raw logit, normalised below
            else if (lih == tt) mylg = lg;
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
                for (int j = 0; j < 4; ++j) acc[q].a[j] = acc[q].a[j] * sc + wk * xl[q].a[j];
            mx = nm;
        };
        for (int t = b; t < e; t += 2) {
            const bool two = t + 1 < e;
            const int s0 = csr_src[t], s1 = two ? csr_src[t + 1] : s0;
            const long long r0 = (long long)eidx[t] * ld_e, r1 = two ? (long long)eidx[t + 1] * ld_e : r0;
            F4 x0[NQ], e0[NQ], x1[NQ], e1[NQ];
#pragma unroll
            for (int q = 0; q < NQ; ++q) {
                const int c = c0 + 128 * q;
                if (act) {
                    x0[q] = g4(xlr + (long long)s0 * ld + c);
                    e0[q] = g4(ee + r0 + c);
                    x1[q] = g4(xlr + (long long)s1 * ld + c);
                    e1[q] = g4(ee + r1 + c);
                } else {
                    x0[q] = e0[q] = x1[q] = e1[q] = z4();
                }
            }
            if (!(self_loops && s0 == i)) {                              // remove_self_loops
#pragma unroll
                for (int q = 0; q < NQ; ++q)
#pragma unroll
                    for (int j = 0; j < 4; ++j) esum[q].a[j] += e0[q].a[j];
                ++cnt;
                step(x0, e0, t, t - b);
            } else if (!spill && lih == t - b) {
                mylg = INFINITY;                                         // marks a removed edge: alpha = 0
            }
            if (two) {
                if (!(self_loops && s1 == i)) {
#pragma unroll
                    for (int q = 0; q < NQ; ++q)
#pragma unroll
                        for (int j = 0; j < 4; ++j) esum[q].a[j] += e1[q].a[j];
                    ++cnt;
                    step(x1, e1, t + 1, t + 1 - b);
                } else if (!spill && lih == t + 1 - b) {
                    mylg = INFINITY;
                }
            }
        }
        if (self_loops) {                                                // the i -> i edge, edge term = mean of the row's
            const float inv_cnt = 1.f / (float)max(cnt, 1);
            F4 ev[NQ];
#pragma unroll
            for (int q = 0; q < NQ; ++q)
#pragma unroll
                for (int j = 0; j < 4; ++j) ev[q].a[j] = esum[q].a[j] * inv_cnt;
            step(xself, ev, E + i, nreal);
        }
        // out = acc / (den + 1e-16) + bias
        const float inv = 1.f / (den + 1e-16f);
#pragma unroll
        for (int q = 0; q < NQ; ++q) {
            const int c = c0 + 128 * q;
            if (act) {
                F4 o, bv = bias ? g4(bias + c) : z4();
#pragma unroll
                for (int j = 0; j < 4; ++j) o.a[j] = acc[q].a[j] * (inv * keep_scale) + bv.a[j];
                stg4(out + (long long)i * H + c, o.q);
            }
        }
        // alpha = exp(logit - max) / (sum + 1e-16), negated where attention dropout removed the entry
        if (!spill) {
            if (act && lih < nslots) {
                const int slot = lih < nreal ? b + lih : E + i;
                float a = 0.f;
                if (mylg != INFINITY) {
                    a = expf(mylg - mx) * inv;
                    if (!(drop_thr == 0u || attn_keep(seed, (long long)slot * heads + head, drop_thr))) a = -a;
                }
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
                const int slot = t < nreal ? b + t : E + i;
                float* ap = alpha + (long long)slot * heads + hd;
                if (t < nreal && self_loops && csr_src[slot] == i) { *ap = 0.f; continue; }     // removed self loop
                const float a = expf(*ap - hmx) * hinv;
                const bool keep = drop_thr == 0u || attn_keep(seed, (long long)slot * heads + hd, drop_thr);
                *ap = keep ? a : -a;
            }
        }
    }
}

// The same forward for C == 128 with an even number of heads: a HALF-warp owns a head (16 lanes x 8 columns), so a
// warp advances two heads of a destination at once.  k_gatv2_fwd<1> at heads = 2 x C = 128 is instruction-bound (72 %
// issue-active at 0.3 of the DRAM peak): with 8 columns per lane the per-edge index loads, the online-softmax scalar
// work (one exp, compares, rescale factors) and the reduction shuffles (4 levels inside 16 lanes instead of 5 inside
// 32) are issued once per PAIR of heads.  The reduction tree differs from k_gatv2_fwd<1>, so results agree to fp32
// rounding, not bit for bit.  GNNX_GATV2_H2=0 selects the one-head-per-warp kernel.
template <typename EIdx>
__global__ void __launch_bounds__(256, 2)
k_gatv2_fwd_h2(const float* __restrict__ xlr, const float* __restrict__ ee, long long ld_e, const EIdx* __restrict__ eidx,
               const float* __restrict__ att, const float* __restrict__ bias, const int* __restrict__ rowptr,
               const int* __restrict__ csr_src, int N, int E, int heads, int groups, float slope, int self_loops,
               unsigned int drop_thr, float keep_scale, unsigned long long seed, float* __restrict__ out, float* alpha) {
    pdl_enter();
    const int H = heads * 128;
    const int lane = threadIdx.x & 31;
    const int wpb = blockDim.x >> 5;
    const long long ld = 2ll * H;
    const int gw = blockIdx.x * wpb + (threadIdx.x >> 5);
    const int row_stride = (gridDim.x * wpb) / groups;
    const int li = lane & 15;                                    // lane inside its head
    const int head = (gw % groups) * 2 + (lane >> 4);
    const int c0 = head * 128 + li * 4;                          // this lane's columns: c0 .. +3 and c0 + 64 .. +3
    int i = gw / groups;
    int nb = 0, ne = 0;
    if (i < N) { nb = rowptr[i]; ne = rowptr[i + 1]; }
    for (; i < N; i += row_stride) {
        const int b = nb, e = ne;
        if (i + row_stride < N) { nb = rowptr[i + row_stride]; ne = rowptr[i + row_stride + 1]; }
        const int nreal = e - b;
        const int nslots = nreal + (self_loops ? 1 : 0);
        const bool spill = nslots > 16;                          // logits in registers (lane t of a head <- slot t) unless the row is longer
        float mylg = 0.f;
        F4 xr[2], av[2], acc[2], esum[2], xself[2];
#pragma unroll
        for (int v = 0; v < 2; ++v) {
            const int c = c0 + 64 * v;
            xr[v] = g4(xlr + (long long)i * ld + H + c);
            xself[v] = self_loops ? g4(xlr + (long long)i * ld + c) : z4();
            av[v] = g4(att + c);
            acc[v] = z4(); esum[v] = z4();
        }
        float mx = -INFINITY, den = 0.f;
        int cnt = 0;
        auto step = [&](const F4 (&xl)[2], const F4 (&ev)[2], int slot, int tt) {
            float part = 0.f;
#pragma unroll
            for (int v = 0; v < 2; ++v)
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    float s = xr[v].a[j] + xl[v].a[j] + ev[v].a[j];
                    s = s > 0.f ? s : s * slope;
                    part = fmaf(s, av[v].a[j], part);
                }
#pragma unroll
            for (int o = 8; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);      // inside the 16-lane half
            const float lg = part;
            if (spill) { if (li == 0) alpha[(long long)slot * heads + head] = lg; }
            else if (li == tt) mylg = lg;
            const float dl = lg - mx;
            const float ex = expf(-fabsf(dl));
            const float nm = fmaxf(mx, lg);
            const float sc = dl > 0.f ? ex : 1.f;
            const float wgt = dl > 0.f ? 1.f : ex;
            den = den * sc + wgt;
            const float wk = (drop_thr == 0u || attn_keep(seed, (long long)slot * heads + head, drop_thr)) ? wgt : 0.f;
#pragma unroll
            for (int v = 0; v < 2; ++v)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[v].a[j] = acc[v].a[j] * sc + wk * xl[v].a[j];
            mx = nm;
        };
        for (int t = b; t < e; t += 2) {
            const bool two = t + 1 < e;
            const int s0 = csr_src[t], s1 = two ? csr_src[t + 1] : s0;
            const long long r0 = (long long)eidx[t] * ld_e, r1 = two ? (long long)eidx[t + 1] * ld_e : r0;
            F4 x0[2], e0[2], x1[2], e1[2];
#pragma unroll
            for (int v = 0; v < 2; ++v) {
                const int c = c0 + 64 * v;
                x0[v] = g4(xlr + (long long)s0 * ld + c);
                e0[v] = g4(ee + r0 + c);
                x1[v] = g4(xlr + (long long)s1 * ld + c);
                e1[v] = g4(ee + r1 + c);
            }
            if (!(self_loops && s0 == i)) {
#pragma unroll
                for (int v = 0; v < 2; ++v)
#pragma unroll
                    for (int j = 0; j < 4; ++j) esum[v].a[j] += e0[v].a[j];
                ++cnt;
                step(x0, e0, t, t - b);
            } else if (!spill && li == t - b) {
                mylg = INFINITY;                                 // marks a removed edge: alpha = 0
            }
            if (two) {
                if (!(self_loops && s1 == i)) {
#pragma unroll
                    for (int v = 0; v < 2; ++v)
#pragma unroll
                        for (int j = 0; j < 4; ++j) esum[v].a[j] += e1[v].a[j];
                    ++cnt;
                    step(x1, e1, t + 1, t + 1 - b);
                } else if (!spill && li == t + 1 - b) {
                    mylg = INFINITY;
                }
            }
        }
        if (self_loops) {
            const float inv_cnt = 1.f / (float)max(cnt, 1);
            F4 ev[2];
#pragma unroll
            for (int v = 0; v < 2; ++v)
#pragma unroll
                for (int j = 0; j < 4; ++j) ev[v].a[j] = esum[v].a[j] * inv_cnt;
            step(xself, ev, E + i, nreal);
        }
        const float inv = 1.f / (den + 1e-16f);
#pragma unroll
        for (int v = 0; v < 2; ++v) {
            const int c = c0 + 64 * v;
            F4 o, bv = bias ? g4(bias + c) : z4();
#pragma unroll
            for (int j = 0; j < 4; ++j) o.a[j] = acc[v].a[j] * (inv * keep_scale) + bv.a[j];
            stg4(out + (long long)i * H + c, o.q);
        }
        if (!spill) {
            if (li < nslots) {
                const int slot = li < nreal ? b + li : E + i;
                float a = 0.f;
                if (mylg != INFINITY) {
                    a = expf(mylg - mx) * inv;
                    if (!(drop_thr == 0u || attn_keep(seed, (long long)slot * heads + head, drop_thr))) a = -a;
                }
                alpha[(long long)slot * heads + head] = a;
            }
            continue;
        }
        __syncwarp();                                            // long rows: the 16 lanes of a head walk its slots
        for (int t = li; t < nslots; t += 16) {
            const int slot = t < nreal ? b + t : E + i;
            float* ap = alpha + (long long)slot * heads + head;
            if (t < nreal && self_loops && csr_src[slot] == i) { *ap = 0.f; continue; }     // removed self loop
            const float a = expf(*ap - mx) * inv;
            const bool keep = drop_thr == 0u || attn_keep(seed, (long long)slot * heads + head, drop_thr);
            *ap = keep ? a : -a;
        }
    }
}

// Backward, destination side.  Writes ds rows (gradient at the pre-activation s) for every slot, dxlr[:, H:2H]
// (= d x_r), and per-warp partial sums of d att.  S_h = sum_e alpha'_e <dout_i, x_l[src_e]>_h is <dout_i, out_i - bias>_h.
// A warp keeps one head group for its whole life (rows advance by total_warps / groups), so d att accumulates in registers.
template <int NQ, typename EIdx>
__global__ void __launch_bounds__(256, NQ == 1 ? 4 : (NQ == 2 ? 2 : 1))
k_gatv2_bwd_dst(const float* __restrict__ dout, const float* __restrict__ out, const float* __restrict__ bias,
                const float* __restrict__ xlr, const float* __restrict__ ee, long long ld_e, const EIdx* __restrict__ eidx,
                const float* __restrict__ att, const float* __restrict__ alpha, const int* __restrict__ rowptr,
                const int* __restrict__ csr_src, int N, int E, int heads, int C, int log2c, int groups, float slope,
                int self_loops, float keep_scale, float* __restrict__ dxlr, float* __restrict__ ds, float* __restrict__ datt_partial,
                float* __restrict__ de, const int* __restrict__ de_row) {
    // de != null: also emit the gradient at the edge term of every slot, de[row(k)] = ds[k] + ds[loop(dst)] / cnt(dst)
    // (row(k) = de_row ? de_row[k] : k) -- the mean-filled self loop passes its gradient back to the incoming edges.
    pdl_enter();
    const int H = heads * C;
    const int lane = threadIdx.x & 31;
    const int wpb = blockDim.x >> 5;
    const long long ld = 2ll * H;
    const int gw = blockIdx.x * wpb + (threadIdx.x >> 5);
    const int row_stride = (gridDim.x * wpb) / groups;           // host guarantees divisibility
    const int cb = (gw % groups) * 128 * NQ;
    const int c0 = cb + lane * 4;
    const bool act = c0 < H;
    const int head = act ? c0 >> log2c : 0;
    F4 av[NQ], datt[NQ], bv[NQ];
#pragma unroll
    for (int q = 0; q < NQ; ++q) {
        const int c = c0 + 128 * q;
        av[q] = act ? g4(att + c) : z4();
        bv[q] = (act && bias) ? g4(bias + c) : z4();
        datt[q] = z4();
    }
    for (int i = gw / groups; i < N; i += row_stride) {
        const int b = rowptr[i], e = rowptr[i + 1];
        F4 xr[NQ], go[NQ], dxr[NQ], esum[NQ];
        float sp = 0.f;
#pragma unroll
        for (int q = 0; q < NQ; ++q) {
            const int c = c0 + 128 * q;
            xr[q] = act ? g4(xlr + (long long)i * ld + H + c) : z4();
            go[q] = act ? g4(dout + (long long)i * H + c) : z4();
            const F4 o = act ? g4(out + (long long)i * H + c) : z4();
            dxr[q] = z4(); esum[q] = z4();
#pragma unroll
            for (int j = 0; j < 4; ++j) sp = fmaf(go[q].a[j], o.a[j] - bv[q].a[j], sp);
        }
        const float S = head_sum<NQ>(sp, C);
        int cnt = 0;
        F4 dl[NQ];                                   // the loop slot's ds / cnt, added to every incoming slot's de
#pragma unroll
        for (int q = 0; q < NQ; ++q) dl[q] = z4();
        auto step = [&](const F4 (&xl)[NQ], const F4 (&ev)[NQ], float stored, int slot, bool is_loop, float loop_scale) {
            float part = 0.f;
#pragma unroll
            for (int q = 0; q < NQ; ++q)
#pragma unroll
                for (int j = 0; j < 4; ++j) part = fmaf(go[q].a[j], xl[q].a[j], part);
            const float kf = attn_dropped(stored) ? 0.f : keep_scale;
            const float dlg = fabsf(stored) * (kf * head_sum<NQ>(part, C) - S);
#pragma unroll
            for (int q = 0; q < NQ; ++q) {
                F4 d;
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float s = xr[q].a[j] + xl[q].a[j] + ev[q].a[j];
                    const float a = s > 0.f ? s : s * slope;
                    datt[q].a[j] = fmaf(dlg, a, datt[q].a[j]);
                    d.a[j] = dlg * av[q].a[j] * (s > 0.f ? 1.f : slope);
                    dxr[q].a[j] += d.a[j];
                }
                if (act) stg4(ds + (long long)slot * H + c0 + 128 * q, d.q);
                if (is_loop) {
#pragma unroll
                    for (int j = 0; j < 4; ++j) dl[q].a[j] = d.a[j] * loop_scale;
                } else if (de && act) {
#pragma unroll
                    for (int j = 0; j < 4; ++j) d.a[j] += dl[q].a[j];
                    const long long row = de_row ? de_row[slot] : slot;
                    stg4(de + row * H + c0 + 128 * q, d.q);
                }
            }
        };
        auto zero_slot = [&](int slot) {
#pragma unroll
            for (int q = 0; q < NQ; ++q)
                if (act) {
                    stg4(ds + (long long)slot * H + c0 + 128 * q, make_float4(0.f, 0.f, 0.f, 0.f));
                    if (de) stg4(de + (de_row ? (long long)de_row[slot] : (long long)slot) * H + c0 + 128 * q, make_float4(0.f, 0.f, 0.f, 0.f));
                }
        };
        if (self_loops) {
            // the loop slot first: its edge term is the mean of the incoming edge terms (a pass over the 60-row table /
            // the slots' e rows, L1-resident for the main pass below), and its ds feeds every incoming slot's de
            for (int t = b; t < e; ++t) {
                if (csr_src[t] == i) continue;
                const long long r0 = (long long)eidx[t] * ld_e;
#pragma unroll
                for (int q = 0; q < NQ; ++q)
                    if (act) {
                        const F4 ev = g4(ee + r0 + c0 + 128 * q);
#pragma unroll
                        for (int j = 0; j < 4; ++j) esum[q].a[j] += ev.a[j];
                    }
                ++cnt;
            }
            const float inv_cnt = 1.f / (float)max(cnt, 1);
            const float al = alpha[(long long)(E + i) * heads + head];
            F4 xl[NQ], ev[NQ];
#pragma unroll
            for (int q = 0; q < NQ; ++q) {
                xl[q] = act ? g4(xlr + (long long)i * ld + c0 + 128 * q) : z4();
#pragma unroll
                for (int j = 0; j < 4; ++j) ev[q].a[j] = esum[q].a[j] * inv_cnt;
            }
            step(xl, ev, al, E + i, true, inv_cnt);
        }
        for (int t = b; t < e; t += 2) {
            const bool two = t + 1 < e;
            const int s0 = csr_src[t], s1 = two ? csr_src[t + 1] : s0;
            const long long r0 = (long long)eidx[t] * ld_e, r1 = two ? (long long)eidx[t + 1] * ld_e : r0;
            const float a0 = alpha[(long long)t * heads + head];
            const float a1 = two ? alpha[(long long)(t + 1) * heads + head] : 0.f;
            F4 x0[NQ], e0[NQ], x1[NQ], e1[NQ];
#pragma unroll
            for (int q = 0; q < NQ; ++q) {
                const int c = c0 + 128 * q;
                if (act) {
                    x0[q] = g4(xlr + (long long)s0 * ld + c);
                    e0[q] = g4(ee + r0 + c);
                    x1[q] = g4(xlr + (long long)s1 * ld + c);
                    e1[q] = g4(ee + r1 + c);
                } else {
                    x0[q] = e0[q] = x1[q] = e1[q] = z4();
                }
            }
            if (!(self_loops && s0 == i)) step(x0, e0, a0, t, false, 0.f);
            else zero_slot(t);
            if (two) {
                if (!(self_loops && s1 == i)) step(x1, e1, a1, t + 1, false, 0.f);
                else zero_slot(t + 1);
            }
        }
#pragma unroll
        for (int q = 0; q < NQ; ++q)
            if (act) stg4(dxlr + (long long)i * ld + H + c0 + 128 * q, dxr[q].q);
    }
#pragma unroll
    for (int q = 0; q < NQ; ++q)
        stg4(datt_partial + (long long)gw * (128 * NQ) + q * 128 + lane * 4, datt[q].q);
}

// k_gatv2_bwd_dst with a HALF-warp per head (C == 128, even head count; see k_gatv2_fwd_h2): 8 columns per lane, the
// index / alpha loads, the per-slot scalar arithmetic and a 4-level reduction are issued once per pair of heads.  Warp gw
// owns heads 2 (gw % groups), + 1 = 256 consecutive columns; its d att partial row is [256] (datt_partial + gw * 256).
template <typename EIdx>
__global__ void __launch_bounds__(256, 2)
k_gatv2_bwd_dst_h2(const float* __restrict__ dout, const float* __restrict__ out, const float* __restrict__ bias,
                   const float* __restrict__ xlr, const float* __restrict__ ee, long long ld_e, const EIdx* __restrict__ eidx,
                   const float* __restrict__ att, const float* __restrict__ alpha, const int* __restrict__ rowptr,
                   const int* __restrict__ csr_src, int N, int E, int heads, int groups, float slope, int self_loops,
                   float keep_scale, float* __restrict__ dxlr, float* __restrict__ ds, float* __restrict__ datt_partial,
                   float* __restrict__ de, const int* __restrict__ de_row) {
    pdl_enter();
    const int H = heads * 128;
    const int lane = threadIdx.x & 31;
    const int wpb = blockDim.x >> 5;
    const long long ld = 2ll * H;
    const int gw = blockIdx.x * wpb + (threadIdx.x >> 5);
    const int row_stride = (gridDim.x * wpb) / groups;
    const int li = lane & 15;
    const int cb = (gw % groups) * 256;
    const int head = (cb >> 7) + (lane >> 4);
    const int c0 = head * 128 + li * 4;                          // this lane's columns: c0 .. +3 and c0 + 64 .. +3
    auto hsum = [](float v) {
#pragma unroll
        for (int o = 8; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        return v;
    };
    F4 av[2], datt[2], bv[2];
#pragma unroll
    for (int v = 0; v < 2; ++v) {
        const int c = c0 + 64 * v;
        av[v] = g4(att + c);
        bv[v] = bias ? g4(bias + c) : z4();
        datt[v] = z4();
    }
    for (int i = gw / groups; i < N; i += row_stride) {
        const int b = rowptr[i], e = rowptr[i + 1];
        F4 xr[2], go[2], dxr[2], esum[2];
        float sp = 0.f;
#pragma unroll
        for (int v = 0; v < 2; ++v) {
            const int c = c0 + 64 * v;
            xr[v] = g4(xlr + (long long)i * ld + H + c);
            go[v] = g4(dout + (long long)i * H + c);
            const F4 o = g4(out + (long long)i * H + c);
            dxr[v] = z4(); esum[v] = z4();
#pragma unroll
            for (int j = 0; j < 4; ++j) sp = fmaf(go[v].a[j], o.a[j] - bv[v].a[j], sp);
        }
        const float S = hsum(sp);
        int cnt = 0;
        F4 dl[2];
        dl[0] = z4(); dl[1] = z4();
        auto step = [&](const F4 (&xl)[2], const F4 (&ev)[2], float stored, int slot, bool is_loop, float loop_scale) {
            float part = 0.f;
#pragma unroll
            for (int v = 0; v < 2; ++v)
#pragma unroll
                for (int j = 0; j < 4; ++j) part = fmaf(go[v].a[j], xl[v].a[j], part);
            const float kf = attn_dropped(stored) ? 0.f : keep_scale;
            const float dlg = fabsf(stored) * (kf * hsum(part) - S);
#pragma unroll
            for (int v = 0; v < 2; ++v) {
                F4 d;
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float s = xr[v].a[j] + xl[v].a[j] + ev[v].a[j];
                    const float a = s > 0.f ? s : s * slope;
                    datt[v].a[j] = fmaf(dlg, a, datt[v].a[j]);
                    d.a[j] = dlg * av[v].a[j] * (s > 0.f ? 1.f : slope);
                    dxr[v].a[j] += d.a[j];
                }
                stg4(ds + (long long)slot * H + c0 + 64 * v, d.q);
                if (is_loop) {
#pragma unroll
                    for (int j = 0; j < 4; ++j) dl[v].a[j] = d.a[j] * loop_scale;
                } else if (de) {
#pragma unroll
                    for (int j = 0; j < 4; ++j) d.a[j] += dl[v].a[j];
                    const long long row = de_row ? de_row[slot] : slot;
                    stg4(de + row * H + c0 + 64 * v, d.q);
                }
            }
        };
        auto zero_slot = [&](int slot) {
#pragma unroll
            for (int v = 0; v < 2; ++v) {
                stg4(ds + (long long)slot * H + c0 + 64 * v, make_float4(0.f, 0.f, 0.f, 0.f));
                if (de) stg4(de + (de_row ? (long long)de_row[slot] : (long long)slot) * H + c0 + 64 * v, make_float4(0.f, 0.f, 0.f, 0.f));
            }
        };
        if (self_loops) {
            for (int t = b; t < e; ++t) {
                if (csr_src[t] == i) continue;
                const long long r0 = (long long)eidx[t] * ld_e;
#pragma unroll
                for (int v = 0; v < 2; ++v) {
                    const F4 ev = g4(ee + r0 + c0 + 64 * v);
#pragma unroll
                    for (int j = 0; j < 4; ++j) esum[v].a[j] += ev.a[j];
                }
                ++cnt;
            }
            const float inv_cnt = 1.f / (float)max(cnt, 1);
            const float al = alpha[(long long)(E + i) * heads + head];
            F4 xl[2], ev[2];
#pragma unroll
            for (int v = 0; v < 2; ++v) {
                xl[v] = g4(xlr + (long long)i * ld + c0 + 64 * v);
#pragma unroll
                for (int j = 0; j < 4; ++j) ev[v].a[j] = esum[v].a[j] * inv_cnt;
            }
            step(xl, ev, al, E + i, true, inv_cnt);
        }
        for (int t = b; t < e; t += 2) {
            const bool two = t + 1 < e;
            const int s0 = csr_src[t], s1 = two ? csr_src[t + 1] : s0;
            const long long r0 = (long long)eidx[t] * ld_e, r1 = two ? (long long)eidx[t + 1] * ld_e : r0;
            const float a0 = alpha[(long long)t * heads + head];
            const float a1 = two ? alpha[(long long)(t + 1) * heads + head] : 0.f;
            F4 x0[2], e0[2], x1[2], e1[2];
#pragma unroll
            for (int v = 0; v < 2; ++v) {
                const int c = c0 + 64 * v;
                x0[v] = g4(xlr + (long long)s0 * ld + c);
                e0[v] = g4(ee + r0 + c);
                x1[v] = g4(xlr + (long long)s1 * ld + c);
                e1[v] = g4(ee + r1 + c);
            }
            if (!(self_loops && s0 == i)) step(x0, e0, a0, t, false, 0.f);
            else zero_slot(t);
            if (two) {
                if (!(self_loops && s1 == i)) step(x1, e1, a1, t + 1, false, 0.f);
                else zero_slot(t + 1);
            }
        }
#pragma unroll
        for (int v = 0; v < 2; ++v) stg4(dxlr + (long long)i * ld + H + c0 + 64 * v, dxr[v].q);
    }
#pragma unroll
    for (int v = 0; v < 2; ++v)
        stg4(datt_partial + (long long)gw * 256 + (c0 - cb) + 64 * v, datt[v].q);
}

// Backward, source side: dxlr[j, 0:H] = sum_{edges j->i} (ds[pos] + alpha'[pos] * dout[i]) + loop term.
// Warp per (source row, 128-column chunk), two out-edges per iteration.
__global__ void __launch_bounds__(256)
k_gatv2_bwd_src(const float* __restrict__ dout, const float* __restrict__ ds, const float* __restrict__ alpha,
                const int* __restrict__ colptr, const int* __restrict__ csc_pos, const int* __restrict__ csc_dst, int N,
                int E, int heads, int C, int log2c, int self_loops, float keep_scale, float* __restrict__ dxlr) {
    pdl_enter();
    const int H = heads * C;
    const int lane = threadIdx.x & 31;
    const int wpb = blockDim.x >> 5;
    const int chunks = (H + 127) / 128;
    const long long ld = 2ll * H;
    const int gw = blockIdx.x * wpb + (threadIdx.x >> 5);        // fixed column chunk per warp, rows strided
    const int row_stride = (gridDim.x * wpb) / chunks;           // host guarantees divisibility
    const int c = (gw % chunks) * 128 + lane * 4;
    if (c >= H) return;
    const int head = c >> log2c;
    for (int j = gw / chunks; j < N; j += row_stride) {
        const int b = colptr[j], e = colptr[j + 1];
        F4 acc = z4();
        auto add = [&](int pos, int dst) {
            const F4 d = g4(ds + (long long)pos * H + c), g = g4(dout + (long long)dst * H + c);
            const float r = alpha[(long long)pos * heads + head];
            const float a = attn_dropped(r) ? 0.f : r * keep_scale;
#pragma unroll
            for (int jj = 0; jj < 4; ++jj) acc.a[jj] += fmaf(a, g.a[jj], d.a[jj]);
        };
        for (int p = b; p < e; p += 2) {
            const bool two = p + 1 < e;
            const int p0 = csc_pos[p], d0 = csc_dst[p];
            const int p1 = two ? csc_pos[p + 1] : p0, d1 = two ? csc_dst[p + 1] : d0;
            const F4 x0 = g4(ds + (long long)p0 * H + c), g0 = g4(dout + (long long)d0 * H + c);
            const F4 x1 = g4(ds + (long long)p1 * H + c), g1 = g4(dout + (long long)d1 * H + c);
            const float r0 = alpha[(long long)p0 * heads + head], r1 = alpha[(long long)p1 * heads + head];
            const float a0 = attn_dropped(r0) ? 0.f : r0 * keep_scale, a1 = attn_dropped(r1) ? 0.f : r1 * keep_scale;
            if (!(self_loops && d0 == j)) {
#pragma unroll
                for (int jj = 0; jj < 4; ++jj) acc.a[jj] += fmaf(a0, g0.a[jj], x0.a[jj]);
            }
            if (two && !(self_loops && d1 == j)) {
#pragma unroll
                for (int jj = 0; jj < 4; ++jj) acc.a[jj] += fmaf(a1, g1.a[jj], x1.a[jj]);
            }
        }
        if (self_loops) add(E + j, j);
        stg4(dxlr + (long long)j * ld + c, acc.q);
    }
}

// de[slot or eid] = ds[slot] + ds[loop(dst)] / deg(dst); warp per destination row
__global__ void __launch_bounds__(256)
k_gatv2_edge_grad(const float* __restrict__ ds, const int* __restrict__ rowptr, const int* __restrict__ csr_src,
                  const int* __restrict__ csr_eid, int N, int E, int H, int self_loops, float* __restrict__ de) {
    pdl_enter();
    const int lane = threadIdx.x & 31;
    const int chunks = (H + 127) / 128;
    const long long nwork = (long long)N * chunks;
    const int wpb = blockDim.x >> 5;
    for (long long w = blockIdx.x * (long long)wpb + (threadIdx.x >> 5); w < nwork; w += (long long)gridDim.x * wpb) {
        int i = (int)(w / chunks);
        int c = (int)(w % chunks) * 128 + lane * 4;
        if (c >= H) continue;
        int b = rowptr[i], e = rowptr[i + 1];
        if (b == e) continue;
        F4 lp = z4();
        if (self_loops) {
            int cnt = 0;
            for (int k = b; k < e; ++k) cnt += (csr_src[k] != i);
            lp = g4(ds + (long long)(E + i) * H + c);
            float inv = 1.f / (float)max(cnt, 1);
#pragma unroll
            for (int j = 0; j < 4; ++j) lp.a[j] *= inv;
        }
        for (int k = b; k < e; ++k) {
            F4 d = g4(ds + (long long)k * H + c);
            if (self_loops && csr_src[k] == i) d = z4();       // removed self loop: no gradient
            else {
#pragma unroll
                for (int j = 0; j < 4; ++j) d.a[j] += lp.a[j];
            }
            long long row = csr_eid ? csr_eid[k] : k;
            stg4(de + row * H + c, d.q);
        }
    }
}

}  // namespace gnnx

using namespace gnnx;

#define GAT_DISPATCH(S, CALL)                                                \
    do {                                                                     \
        if ((S).nq == 1) { CALL(1); }                                        \
        else if ((S).nq == 2) { CALL(2); }                                   \
        else if ((S).nq == 4) { CALL(4); }                                   \
        else { CALL(8); }                                                    \
    } while (0)

static int gat_bwd_grid(int64_t N, int groups) { return attn_row_grid(N, groups, 8); }
// half-warp-per-head kernels (C == 128, even head count); GNNX_GATV2_H2=0 selects the warp-per-head ones (read per call)
static bool gat_h2(int heads, int C) {
    const char* pe = getenv("GNNX_GATV2_H2");
    return C == 128 && heads % 2 == 0 && !(pe && pe[0] == '0');
}

extern "C" {

int gnnx_gatv2_fwd(const float* xlr, const float* e, int64_t ld_e, const int32_t* eidx_i32, const uint8_t* eidx_u8,
                   const float* att, const float* bias, const int32_t* rowptr, const int32_t* csr_src, int64_t N, int64_t E,
                   int32_t heads, int32_t C, float slope, int32_t self_loops, float dropout_p, int64_t seed, float* out,
                   float* alpha, void* stream) {
    GNNX_REQUIRE(attn_shape_ok(heads, C), "gatv2_fwd: need C a power of two in [4, 1024] and heads*C <= 4096");
    GNNX_REQUIRE(dropout_p >= 0.f && dropout_p < 1.f, "gatv2_fwd: dropout_p must be in [0, 1)");
    GNNX_REQUIRE(E == 0 || (eidx_i32 != nullptr) != (eidx_u8 != nullptr), "gatv2_fwd: exactly one of eidx_i32 / eidx_u8");
    GNNX_REQUIRE(ld_e % 4 == 0, "gatv2_fwd: ld_e must be a multiple of 4");
    if (N == 0) return GNNX_OK;
    const AttnShape s = attn_shape(heads, C);
    cudaStream_t st = as_stream(stream);
    const int grid = attn_row_grid(N, s.groups, 16);
    const unsigned int thr = attn_drop_threshold(dropout_p);
    const float ks = 1.0f / (1.0f - dropout_p);
    const unsigned long long sd = (unsigned long long)seed;
    {   // a half-warp per head (GNNX_GATV2_H2=0: a warp per head); read per call so that a test can compare both
        if (gat_h2(heads, C)) {
            const int groups = heads / 2;
            const int grid2 = attn_row_grid(N, groups, 8);
            if (eidx_i32) launch(k_gatv2_fwd_h2<int>, grid2, 256, 0, st, xlr, e, ld_e, eidx_i32, att, bias, rowptr, csr_src, (int)N, (int)E, heads, groups, slope, self_loops, thr, ks, sd, out, alpha);
            else launch(k_gatv2_fwd_h2<uint8_t>, grid2, 256, 0, st, xlr, e, ld_e, eidx_u8, att, bias, rowptr, csr_src, (int)N, (int)E, heads, groups, slope, self_loops, thr, ks, sd, out, alpha);
            return check_launch("gatv2_fwd");
        }
    }
#define CALL(NQ)                                                                                                              \
    if (eidx_i32) launch(k_gatv2_fwd<NQ, int>, grid, 256, 0, st, xlr, e, ld_e, eidx_i32, att, bias, rowptr, csr_src, (int)N, (int)E, heads, C, ilog2(C), s.groups, slope, self_loops, thr, ks, sd, out, alpha); \
    else launch(k_gatv2_fwd<NQ, uint8_t>, grid, 256, 0, st, xlr, e, ld_e, eidx_u8, att, bias, rowptr, csr_src, (int)N, (int)E, heads, C, ilog2(C), s.groups, slope, self_loops, thr, ks, sd, out, alpha)
    GAT_DISPATCH(s, CALL);
#undef CALL
    return check_launch("gatv2_fwd");
}

size_t gnnx_gatv2_bwd_workspace_bytes(int64_t N, int64_t E, int32_t heads, int32_t C) {
    (void)E;
    if (!attn_shape_ok(heads, C)) return 0;
    const AttnShape s = attn_shape(heads, C);
    // per-warp partials of d att, viewed as a row-major [warps / groups, groups * width] matrix, + the workspace of the
    // column sum that folds them
    const size_t part = (size_t)gat_bwd_grid(N, s.groups) * 8 * 128 * s.nq * sizeof(float);
    const int64_t rows = (int64_t)gat_bwd_grid(N, s.groups) * 8 / s.groups;
    size_t need = part + gnnx_colsum_workspace_bytes(rows, (int64_t)heads * C);
    if (C == 128 && heads % 2 == 0) {                     // the half-warp-per-head layout: [256] per warp, heads / 2 groups
        const int g2 = heads / 2;
        const size_t part2 = (size_t)gat_bwd_grid(N, g2) * 8 * 256 * sizeof(float);
        const size_t need2 = part2 + gnnx_colsum_workspace_bytes((int64_t)gat_bwd_grid(N, g2) * 8 / g2, (int64_t)heads * C);
        if (need2 > need) need = need2;
    }
    return need;
}

int gnnx_gatv2_bwd(const float* dout, const float* out, const float* bias, const float* xlr, const float* e, int64_t ld_e,
                   const int32_t* eidx_i32, const uint8_t* eidx_u8, const float* att, const float* alpha,
                   const int32_t* rowptr, const int32_t* csr_src, const int32_t* colptr, const int32_t* csc_pos,
                   const int32_t* csc_dst, int64_t N, int64_t E, int32_t heads, int32_t C, float slope, int32_t self_loops,
                   float dropout_p, float* dxlr, float* datt, float* ds, float* de, const int32_t* de_row, void* ws,
                   size_t ws_bytes, void* stream) {
    const int H = heads * C;
    GNNX_REQUIRE(attn_shape_ok(heads, C), "gatv2_bwd: need C a power of two in [4, 1024] and heads*C <= 4096");
    GNNX_REQUIRE(dropout_p >= 0.f && dropout_p < 1.f, "gatv2_bwd: dropout_p must be in [0, 1)");
    GNNX_REQUIRE(E == 0 || (eidx_i32 != nullptr) != (eidx_u8 != nullptr), "gatv2_bwd: exactly one of eidx_i32 / eidx_u8");
    GNNX_REQUIRE(ws_bytes >= gnnx_gatv2_bwd_workspace_bytes(N, E, heads, C), "gatv2_bwd: workspace too small");
    cudaStream_t st = as_stream(stream);
    if (N == 0) { GNNX_CUDA(cudaMemsetAsync(datt, 0, H * sizeof(float), st)); return GNNX_OK; }
    const AttnShape s = attn_shape(heads, C);
    const float ks = 1.0f / (1.0f - dropout_p);
    if (gat_h2(heads, C)) {
        const int g2 = heads / 2, grid2 = gat_bwd_grid(N, g2);
        if (eidx_i32) launch(k_gatv2_bwd_dst_h2<int>, grid2, 256, 0, st, dout, out, bias, xlr, e, ld_e, eidx_i32, att, alpha, rowptr, csr_src, (int)N, (int)E, heads, g2, slope, self_loops, ks, dxlr, ds, (float*)ws, de, de_row);
        else launch(k_gatv2_bwd_dst_h2<uint8_t>, grid2, 256, 0, st, dout, out, bias, xlr, e, ld_e, eidx_u8, att, alpha, rowptr, csr_src, (int)N, (int)E, heads, g2, slope, self_loops, ks, dxlr, ds, (float*)ws, de, de_row);
        launch(k_gatv2_bwd_src, attn_row_grid(N, (H + 127) / 128, 16), 256, 0, st, dout, ds, alpha, colptr, csc_pos, csc_dst, (int)N, (int)E, heads, C, ilog2(C), self_loops, ks, dxlr);
        int rc2 = check_launch("gatv2_bwd", 2);
        if (rc2) return rc2;
        const size_t part2 = (size_t)grid2 * 8 * 256 * sizeof(float);
        return gnnx_colsum((const float*)ws, (int64_t)g2 * 256, (int64_t)grid2 * 8 / g2, H, datt, (char*)ws + part2, ws_bytes - part2, stream);
    }
    const int grid = gat_bwd_grid(N, s.groups);
#define CALL(NQ)                                                                                                              \
    if (eidx_i32) launch(k_gatv2_bwd_dst<NQ, int>, grid, 256, 0, st, dout, out, bias, xlr, e, ld_e, eidx_i32, att, alpha, rowptr, csr_src, (int)N, (int)E, heads, C, ilog2(C), s.groups, slope, self_loops, ks, dxlr, ds, (float*)ws, de, de_row); \
    else launch(k_gatv2_bwd_dst<NQ, uint8_t>, grid, 256, 0, st, dout, out, bias, xlr, e, ld_e, eidx_u8, att, alpha, rowptr, csr_src, (int)N, (int)E, heads, C, ilog2(C), s.groups, slope, self_loops, ks, dxlr, ds, (float*)ws, de, de_row)
    GAT_DISPATCH(s, CALL);
#undef CALL
    launch(k_gatv2_bwd_src, attn_row_grid(N, (H + 127) / 128, 16), 256, 0, st, dout, ds, alpha, colptr, csc_pos, csc_dst, (int)N, (int)E, heads, C, ilog2(C), self_loops, ks, dxlr);
    int rc = check_launch("gatv2_bwd", 2);
    if (rc) return rc;
    // d att = column sums of the per-warp partials: warp gw owns head group gw % groups, so `groups` consecutive partial
    // rows are one row of all H columns.  (A one-warp-per-column fold walked 300 strided rows per lane: 35 us for 5 MB;
    // the two-pass column sum reads them coalesced.)
    const size_t part = (size_t)grid * 8 * 128 * s.nq * sizeof(float);
    return gnnx_colsum((const float*)ws, (int64_t)s.groups * 128 * s.nq, (int64_t)grid * 8 / s.groups, H, datt,
                       (char*)ws + part, ws_bytes - part, stream);
}

int gnnx_gatv2_edge_grad(const float* ds, const int32_t* rowptr, const int32_t* csr_src, const int32_t* csr_eid, int64_t N,
                         int64_t E, int32_t H, int32_t self_loops, float* de, void* stream) {
    GNNX_REQUIRE(H % 4 == 0, "gatv2_edge_grad: hidden must be a multiple of 4");
    if (N == 0 || E == 0) return GNNX_OK;
    long long nwork = (long long)N * ((H + 127) / 128);
    long long blocks = (nwork + 7) / 8, cap = (long long)num_sms() * 32;
    launch(k_gatv2_edge_grad, (int)(blocks < cap ? blocks : cap), 256, 0, as_stream(stream), ds, rowptr, csr_src, csr_eid, (int)N, (int)E, H, self_loops, de);
    return check_launch("gatv2_edge_grad");
}

}  // extern "C"
