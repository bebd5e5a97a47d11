// GATv2 attention over the dst-sorted CSR: additive attention logits, segment
// softmax over each destination (online, single gather of x_l[src] per edge), weighted aggregation,
// self loops synthesised in-kernel (loop edge term = mean of the incoming edge terms, which equals
// lin_edge(mean(edge_attr)) because lin_edge has no bias), and the matching backward.
// One warp owns one destination row and all H = heads*C columns (NQ chunks of 128); C is a power of 2.
// Deterministic and atomic-free: the only cross-row reduction (d att) goes through ordered partials.
#include "common.cuh"

namespace gnnx {

union F4 {
    float4 q;
    float a[4];
};
__device__ __forceinline__ F4 g4(const float* p) { F4 r; r.q = ldg4(p); return r; }
__device__ __forceinline__ F4 z4() { F4 r; r.q = make_float4(0.f, 0.f, 0.f, 0.f); return r; }

// part[q] (per-lane partial dot products) -> logit of the head that owns columns (q*128 + lane*4 ..)
template <int NQ>
__device__ __forceinline__ void head_reduce(float (&part)[NQ], int C) {
    if (C >= 128) {
        const int qph = C >> 7;
#pragma unroll
        for (int q = 0; q < NQ; ++q) part[q] = warp_sum(part[q]);
        if (qph > 1) {
            float tot[NQ];
#pragma unroll
            for (int q = 0; q < NQ; ++q) {
                float t = 0.f;
#pragma unroll
                for (int p = 0; p < NQ; ++p) t += (p / qph == q / qph) ? part[p] : 0.f;
                tot[q] = t;
            }
#pragma unroll
            for (int q = 0; q < NQ; ++q) part[q] = tot[q];
        }
    } else {
#pragma unroll
        for (int q = 0; q < NQ; ++q) {
            float v = part[q];
            for (int o = C >> 3; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
            part[q] = v;
        }
    }
}

// is this lane the writer for the head owning (q, lane)?  (first lane of the head's first chunk)
__device__ __forceinline__ bool head_writer(int q, int lane, int C) {
    int col = q * 128 + lane * 4;
    return col % C == 0;
}

template <int NQ, typename EIdx>
__global__ void __launch_bounds__(256)
k_gatv2_fwd(const float* __restrict__ xlr, const float* __restrict__ ee, long long ld_e, const EIdx* __restrict__ eidx,
            const float* __restrict__ att, const float* __restrict__ bias, const int* __restrict__ rowptr,
            const int* __restrict__ csr_src, int N, int E, int heads, int C, float slope, int self_loops,
            float* __restrict__ out, float* alpha) {
    const int H = heads * C;
    const int lane = threadIdx.x & 31;
    const int wpb = blockDim.x >> 5;
    const long long ld = 2ll * H;
    F4 av[NQ];
#pragma unroll
    for (int q = 0; q < NQ; ++q) {
        int c = q * 128 + lane * 4;
        av[q] = c < H ? g4(att + c) : z4();
    }
    for (int i = blockIdx.x * wpb + (threadIdx.x >> 5); i < N; i += gridDim.x * wpb) {
        const int b = rowptr[i], e = rowptr[i + 1];
        F4 xr[NQ], acc[NQ], esum[NQ];
        float mx[NQ], den[NQ];
#pragma unroll
        for (int q = 0; q < NQ; ++q) {
            int c = q * 128 + lane * 4;
            xr[q] = c < H ? g4(xlr + (long long)i * ld + H + c) : z4();
            acc[q] = z4(); esum[q] = z4();
            mx[q] = -INFINITY; den[q] = 0.f;
        }
        int cnt = 0;
        const int nslots = (e - b) + (self_loops ? 1 : 0);
        for (int t = 0; t < nslots; ++t) {
            const bool is_loop = t == (e - b);
            int src, slot;
            if (is_loop) { src = i; slot = E + i; }
            else {
                slot = b + t; src = csr_src[slot];
                if (self_loops && src == i) {            // remove_self_loops
                    if (lane < heads) alpha[(long long)slot * heads + lane] = 0.f;
                    continue;
                }
            }
            F4 xl[NQ];
            float part[NQ];
            const float inv_cnt = 1.f / (float)max(cnt, 1);
#pragma unroll
            for (int q = 0; q < NQ; ++q) {
                int c = q * 128 + lane * 4;
                part[q] = 0.f;
                xl[q] = z4();
                if (c < H) {
                    xl[q] = g4(xlr + (long long)src * ld + c);
                    F4 ev;
                    if (is_loop) {
#pragma unroll
                        for (int j = 0; j < 4; ++j) ev.a[j] = esum[q].a[j] * inv_cnt;
                    } else {
                        ev = g4(ee + (long long)eidx[slot] * ld_e + c);
#pragma unroll
                        for (int j = 0; j < 4; ++j) esum[q].a[j] += ev.a[j];
                    }
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        float s = xr[q].a[j] + xl[q].a[j] + ev.a[j];
                        s = s > 0.f ? s : s * slope;
                        part[q] = fmaf(s, av[q].a[j], part[q]);
                    }
                }
            }
            if (!is_loop) ++cnt;
            head_reduce<NQ>(part, C);
#pragma unroll
            for (int q = 0; q < NQ; ++q) {
                int c = q * 128 + lane * 4;
                if (c < H) {
                    float lg = part[q];
                    if (head_writer(q, lane, C)) alpha[(long long)slot * heads + c / C] = lg;   // raw logit for now
                    float nm = fmaxf(mx[q], lg);
                    float sc = expf(mx[q] - nm);        // exp(-inf) = 0 on the first edge
                    float w = expf(lg - nm);
                    den[q] = den[q] * sc + w;
#pragma unroll
                    for (int j = 0; j < 4; ++j) acc[q].a[j] = acc[q].a[j] * sc + w * xl[q].a[j];
                    mx[q] = nm;
                }
            }
        }
        // normalise: out = acc / (den + 1e-16) + bias ; alpha = exp(lg - mx) / (den + 1e-16)
#pragma unroll
        for (int q = 0; q < NQ; ++q) {
            int c = q * 128 + lane * 4;
            if (c < H) {
                float inv = 1.f / (den[q] + 1e-16f);
                F4 o, bv = bias ? g4(bias + c) : z4();
#pragma unroll
                for (int j = 0; j < 4; ++j) o.a[j] = acc[q].a[j] * inv + bv.a[j];
                stg4(out + (long long)i * H + c, o.q);
                if (head_writer(q, lane, C)) {
                    int hd = c / C;
                    for (int t = 0; t < nslots; ++t) {
                        int slot = t == (e - b) ? E + i : b + t;
                        if (t < (e - b) && self_loops && csr_src[slot] == i) continue;
                        float* ap = alpha + (long long)slot * heads + hd;
                        *ap = expf(*ap - mx[q]) * inv;
                    }
                }
            }
        }
    }
}

// Backward, destination side.  Writes ds rows (gradient at the pre-activation s) for every slot,
// dxlr[:, H:2H] (= d x_r), and per-block partial sums of d att.
template <int NQ, typename EIdx>
__global__ void __launch_bounds__(256)
k_gatv2_bwd_dst(const float* __restrict__ dout, const float* __restrict__ xlr, const float* __restrict__ ee, long long ld_e,
                const EIdx* __restrict__ eidx, const float* __restrict__ att, const float* __restrict__ alpha,
                const int* __restrict__ rowptr, const int* __restrict__ csr_src, int N, int E, int heads, int C,
                float slope, int self_loops, float* __restrict__ dxlr, float* __restrict__ ds,
                float* __restrict__ datt_partial) {
    extern __shared__ float sh_datt[];        // [warps][H]
    const int H = heads * C;
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int wpb = blockDim.x >> 5;
    const long long ld = 2ll * H;
    F4 av[NQ], datt[NQ];
#pragma unroll
    for (int q = 0; q < NQ; ++q) {
        int c = q * 128 + lane * 4;
        av[q] = c < H ? g4(att + c) : z4();
        datt[q] = z4();
    }
    for (int i = blockIdx.x * wpb + warp; i < N; i += gridDim.x * wpb) {
        const int b = rowptr[i], e = rowptr[i + 1];
        const int nslots = (e - b) + (self_loops ? 1 : 0);
        F4 xr[NQ], go[NQ], dxr[NQ], esum[NQ];
        float S[NQ];
#pragma unroll
        for (int q = 0; q < NQ; ++q) {
            int c = q * 128 + lane * 4;
            xr[q] = c < H ? g4(xlr + (long long)i * ld + H + c) : z4();
            go[q] = c < H ? g4(dout + (long long)i * H + c) : z4();
            dxr[q] = z4(); esum[q] = z4();
            S[q] = 0.f;
        }
        // pass 1: S_h = sum_e alpha_e * <dout_i, x_l[src_e]>_h
        for (int t = 0; t < nslots; ++t) {
            const bool is_loop = t == (e - b);
            int slot = is_loop ? E + i : b + t;
            int src = is_loop ? i : csr_src[slot];
            if (!is_loop && self_loops && src == i) continue;
            float part[NQ];
#pragma unroll
            for (int q = 0; q < NQ; ++q) {
                int c = q * 128 + lane * 4;
                part[q] = 0.f;
                if (c < H) {
                    F4 xl = g4(xlr + (long long)src * ld + c);
#pragma unroll
                    for (int j = 0; j < 4; ++j) part[q] = fmaf(go[q].a[j], xl.a[j], part[q]);
                }
            }
            head_reduce<NQ>(part, C);
#pragma unroll
            for (int q = 0; q < NQ; ++q) {
                int c = q * 128 + lane * 4;
                if (c < H) S[q] += alpha[(long long)slot * heads + c / C] * part[q];
            }
        }
        // pass 2: d logit, d s, accumulate d x_r / d att; loop slot last (needs the mean edge term)
        int cnt = 0;
        for (int t = 0; t < nslots; ++t) {
            const bool is_loop = t == (e - b);
            int slot = is_loop ? E + i : b + t;
            int src = is_loop ? i : csr_src[slot];
            if (!is_loop && self_loops && src == i) {
#pragma unroll
                for (int q = 0; q < NQ; ++q) {
                    int c = q * 128 + lane * 4;
                    if (c < H) stg4(ds + (long long)slot * H + c, make_float4(0.f, 0.f, 0.f, 0.f));
                }
                continue;
            }
            const float inv_cnt = 1.f / (float)max(cnt, 1);
            F4 sv[NQ];
            float part[NQ];
#pragma unroll
            for (int q = 0; q < NQ; ++q) {
                int c = q * 128 + lane * 4;
                part[q] = 0.f;
                sv[q] = z4();
                if (c < H) {
                    F4 xl = g4(xlr + (long long)src * ld + c);
                    F4 ev;
                    if (is_loop) {
#pragma unroll
                        for (int j = 0; j < 4; ++j) ev.a[j] = esum[q].a[j] * inv_cnt;
                    } else {
                        ev = g4(ee + (long long)eidx[slot] * ld_e + c);
#pragma unroll
                        for (int j = 0; j < 4; ++j) esum[q].a[j] += ev.a[j];
                    }
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        sv[q].a[j] = xr[q].a[j] + xl.a[j] + ev.a[j];
                        part[q] = fmaf(go[q].a[j], xl.a[j], part[q]);
                    }
                }
            }
            if (!is_loop) ++cnt;
            head_reduce<NQ>(part, C);     // d alpha for this lane's heads
#pragma unroll
            for (int q = 0; q < NQ; ++q) {
                int c = q * 128 + lane * 4;
                if (c < H) {
                    float a = alpha[(long long)slot * heads + c / C];
                    float dlg = a * (part[q] - S[q]);
                    F4 d;
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        float s = sv[q].a[j];
                        float act = s > 0.f ? s : s * slope;
                        datt[q].a[j] = fmaf(dlg, act, datt[q].a[j]);
                        d.a[j] = dlg * av[q].a[j] * (s > 0.f ? 1.f : slope);
                        dxr[q].a[j] += d.a[j];
                    }
                    stg4(ds + (long long)slot * H + c, d.q);
                }
            }
        }
#pragma unroll
        for (int q = 0; q < NQ; ++q) {
            int c = q * 128 + lane * 4;
            if (c < H) stg4(dxlr + (long long)i * ld + H + c, dxr[q].q);
        }
    }
    // ordered block reduction of d att
#pragma unroll
    for (int q = 0; q < NQ; ++q) {
        int c = q * 128 + lane * 4;
        if (c < H) stg4(sh_datt + warp * H + c, datt[q].q);
    }
    __syncthreads();
    for (int c = threadIdx.x; c < H; c += blockDim.x) {
        float s = 0.f;
        for (int w = 0; w < wpb; ++w) s += sh_datt[w * H + c];
        datt_partial[(long long)blockIdx.x * H + c] = s;
    }
}

// Backward, source side: dxlr[j, 0:H] = sum_{edges j->i} (ds[pos] + alpha[pos] * dout[i]) + loop term
template <int NQ>
__global__ void __launch_bounds__(256)
k_gatv2_bwd_src(const float* __restrict__ dout, const float* __restrict__ ds, const float* __restrict__ alpha,
                const int* __restrict__ colptr, const int* __restrict__ csc_pos, const int* __restrict__ csr_dst_of_pos,
                int N, int E, int heads, int C, int self_loops, float* __restrict__ dxlr) {
    const int H = heads * C;
    const int lane = threadIdx.x & 31;
    const int wpb = blockDim.x >> 5;
    const long long ld = 2ll * H;
    for (int j = blockIdx.x * wpb + (threadIdx.x >> 5); j < N; j += gridDim.x * wpb) {
        F4 acc[NQ];
#pragma unroll
        for (int q = 0; q < NQ; ++q) acc[q] = z4();
        const int b = colptr[j], e = colptr[j + 1];
        const int nslots = (e - b) + (self_loops ? 1 : 0);
        for (int t = 0; t < nslots; ++t) {
            const bool is_loop = t == (e - b);
            int pos, dst;
            if (is_loop) { pos = E + j; dst = j; }
            else {
                pos = csc_pos[b + t]; dst = csr_dst_of_pos[b + t];
                if (self_loops && dst == j) continue;
            }
#pragma unroll
            for (int q = 0; q < NQ; ++q) {
                int c = q * 128 + lane * 4;
                if (c < H) {
                    F4 d = g4(ds + (long long)pos * H + c);
                    F4 g = g4(dout + (long long)dst * H + c);
                    float a = alpha[(long long)pos * heads + c / C];
#pragma unroll
                    for (int jj = 0; jj < 4; ++jj) acc[q].a[jj] += d.a[jj] + a * g.a[jj];
                }
            }
        }
#pragma unroll
        for (int q = 0; q < NQ; ++q) {
            int c = q * 128 + lane * 4;
            if (c < H) stg4(dxlr + (long long)j * ld + c, acc[q].q);
        }
    }
}

__global__ void k_sum_rows(const float* __restrict__ partial, int nrows, int H, float* __restrict__ out) {
    int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= H) return;
    float s = 0.f;
    for (int r = 0; r < nrows; ++r) s += partial[(long long)r * H + c];
    out[c] = s;
}

// de[slot or eid] = ds[slot] + ds[loop(dst)] / deg(dst); warp per destination row
__global__ void __launch_bounds__(256)
k_gatv2_edge_grad(const float* __restrict__ ds, const int* __restrict__ rowptr, const int* __restrict__ csr_src,
                  const int* __restrict__ csr_eid, int N, int E, int H, int self_loops, float* __restrict__ de) {
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

static int gat_grid(int64_t N) {
    long long blocks = (N + 7) / 8;
    long long cap = (long long)kNumSMs * 8;
    return (int)(blocks < cap ? (blocks > 0 ? blocks : 1) : cap);
}

static bool pow2(int v) { return v > 0 && (v & (v - 1)) == 0; }

}  // namespace gnnx

using namespace gnnx;

#define GAT_DISPATCH_NQ(H, CALL)                                              \
    do {                                                                      \
        int _nq = ((H) + 127) / 128;                                          \
        if (_nq <= 1) { CALL(1); }                                            \
        else if (_nq <= 2) { CALL(2); }                                       \
        else if (_nq <= 4) { CALL(4); }                                       \
        else { CALL(8); }                                                     \
    } while (0)

extern "C" {

int gnnx_gatv2_fwd(const float* xlr, const float* e, int64_t ld_e, const int32_t* eidx_i32, const uint8_t* eidx_u8,
                   const float* att, const float* bias, const int32_t* rowptr, const int32_t* csr_src, int64_t N, int64_t E,
                   int32_t heads, int32_t C, float slope, int32_t self_loops, float* out, float* alpha, void* stream) {
    const int H = heads * C;
    GNNX_REQUIRE(pow2(C) && C >= 4 && heads >= 1 && heads <= 32 && H <= 1024, "gatv2_fwd: need C a power of two >= 4, heads <= 32, heads*C <= 1024");
    GNNX_REQUIRE((eidx_i32 != nullptr) != (eidx_u8 != nullptr), "gatv2_fwd: exactly one of eidx_i32 / eidx_u8");
    GNNX_REQUIRE(ld_e % 4 == 0, "gatv2_fwd: ld_e must be a multiple of 4");
    if (N == 0) return GNNX_OK;
    cudaStream_t st = as_stream(stream);
    int grid = gat_grid(N);
#define CALL(NQ)                                                                                                              \
    if (eidx_i32) k_gatv2_fwd<NQ, int><<<grid, 256, 0, st>>>(xlr, e, ld_e, eidx_i32, att, bias, rowptr, csr_src, (int)N, (int)E, heads, C, slope, self_loops, out, alpha); \
    else k_gatv2_fwd<NQ, uint8_t><<<grid, 256, 0, st>>>(xlr, e, ld_e, eidx_u8, att, bias, rowptr, csr_src, (int)N, (int)E, heads, C, slope, self_loops, out, alpha)
    GAT_DISPATCH_NQ(H, CALL);
#undef CALL
    return check_launch("gatv2_fwd");
}

size_t gnnx_gatv2_bwd_workspace_bytes(int64_t N, int64_t E, int32_t heads, int32_t C) {
    (void)E;
    return (size_t)gat_grid(N) * heads * C * sizeof(float);
}

int gnnx_gatv2_bwd(const float* dout, const float* xlr, const float* e, int64_t ld_e, const int32_t* eidx_i32,
                   const uint8_t* eidx_u8, const float* att, const float* alpha, const int32_t* rowptr, const int32_t* csr_src,
                   const int32_t* colptr, const int32_t* csc_pos, const int32_t* csc_dst, int64_t N, int64_t E, int32_t heads,
                   int32_t C, float slope, int32_t self_loops, float* dxlr, float* datt, float* ds, void* ws, size_t ws_bytes,
                   void* stream) {
    const int H = heads * C;
    GNNX_REQUIRE(pow2(C) && C >= 4 && heads >= 1 && heads <= 32 && H <= 1024, "gatv2_bwd: need C a power of two >= 4, heads <= 32, heads*C <= 1024");
    GNNX_REQUIRE((eidx_i32 != nullptr) != (eidx_u8 != nullptr), "gatv2_bwd: exactly one of eidx_i32 / eidx_u8");
    GNNX_REQUIRE(ws_bytes >= gnnx_gatv2_bwd_workspace_bytes(N, E, heads, C), "gatv2_bwd: workspace too small");
    cudaStream_t st = as_stream(stream);
    if (N == 0) { GNNX_CUDA(cudaMemsetAsync(datt, 0, H * sizeof(float), st)); return GNNX_OK; }
    int grid = gat_grid(N);
    size_t smem = (size_t)8 * H * sizeof(float);
#define CALL(NQ)                                                                                                              \
    if (eidx_i32) k_gatv2_bwd_dst<NQ, int><<<grid, 256, smem, st>>>(dout, xlr, e, ld_e, eidx_i32, att, alpha, rowptr, csr_src, (int)N, (int)E, heads, C, slope, self_loops, dxlr, ds, (float*)ws); \
    else k_gatv2_bwd_dst<NQ, uint8_t><<<grid, 256, smem, st>>>(dout, xlr, e, ld_e, eidx_u8, att, alpha, rowptr, csr_src, (int)N, (int)E, heads, C, slope, self_loops, dxlr, ds, (float*)ws); \
    k_gatv2_bwd_src<NQ><<<grid, 256, 0, st>>>(dout, ds, alpha, colptr, csc_pos, csc_dst, (int)N, (int)E, heads, C, self_loops, dxlr)
    GAT_DISPATCH_NQ(H, CALL);
#undef CALL
    k_sum_rows<<<ceil_div(H, 128), 128, 0, st>>>((const float*)ws, grid, H, datt);
    return check_launch("gatv2_bwd", 3);
}

int gnnx_gatv2_edge_grad(const float* ds, const int32_t* rowptr, const int32_t* csr_src, const int32_t* csr_eid, int64_t N,
                         int64_t E, int32_t H, int32_t self_loops, float* de, void* stream) {
    GNNX_REQUIRE(H % 4 == 0, "gatv2_edge_grad: hidden must be a multiple of 4");
    if (N == 0 || E == 0) return GNNX_OK;
    long long nwork = (long long)N * ((H + 127) / 128);
    long long blocks = (nwork + 7) / 8, cap = (long long)kNumSMs * 32;
    k_gatv2_edge_grad<<<(int)(blocks < cap ? blocks : cap), 256, 0, as_stream(stream)>>>(ds, rowptr, csr_src, csr_eid, (int)N, (int)E, H, self_loops, de);
    return check_launch("gatv2_edge_grad");
}

}  // extern "C"
