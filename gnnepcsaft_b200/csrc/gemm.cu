// gnnx_gemm: precision dispatch for the dense transforms (Linear layers of PNAConv / GATv2Conv / head).
//   GNNX_PREC_FP32 -> CUDA-core fp32 kernel (gemm_simt.cu), the rel-1e-5 parity tier.
//   GNNX_PREC_TF32 -> tcgen05 + TMEM kernels (gemm_tc.cu); shapes the tensor-core
//   kernel does not cover (tiny or unaligned: the 60-row bond tables, the num_para-wide head output)
//   run on the fp32 kernel, which is at least as accurate.
#include <stdlib.h>

#include "common.cuh"

namespace gnnx {
size_t simt_workspace_bytes(int64_t M, int64_t N, int64_t K);
int gemm_simt(const float* A, int64_t lda, int ta, const float* B, int64_t ldb, int tb, float* C, int64_t ldc, int64_t M,
              int64_t N, int64_t K, const float* bias, int flags, const float* mask, int64_t ldm, uint32_t* bits_out,
              void* ws, size_t ws_bytes, cudaStream_t st, const int* bucket_ptr = nullptr, int n_groups = 1);
int gemm_simt_red_grouped(const float* A, int64_t lda, const float* B, int64_t ldb, int64_t Mo, int64_t No, int64_t R,
                          const int* split_table, int S, float* partial, cudaStream_t st);
bool tc_supported(int64_t M, int64_t N, int64_t K, int ta, int tb, int64_t lda, int64_t ldb, int64_t ldc, const void* A,
                  const void* B, const void* C, int precision);
size_t tc_workspace_bytes(int64_t M, int64_t N, int64_t K, int ta, int tb, int precision);
int gemm_tc(const float* A, int64_t lda, int ta, const float* B, int64_t ldb, int tb, float* C, int64_t ldc, int64_t M,
            int64_t N, int64_t K, const float* bias, int flags, const float* mask, int64_t ldm, uint32_t* bits_out,
            float* colsum_a_out, int precision, void* ws, size_t ws_bytes, cudaStream_t st,
            const int* bucket_ptr = nullptr, int n_groups = 1);
int tc_red_grouped_splits(int64_t Mo, int64_t No, int64_t R);
int gemm_tc_red_grouped(const float* A, int64_t lda, const float* B, int64_t ldb, int64_t Mo, int64_t No, int64_t R,
                        const int* split_table, int S, float* partial, cudaStream_t st);
bool tc_red_batched_ok(int n, int64_t Mo, int64_t No, int64_t R);
size_t tc_red_batched_workspace_bytes(int n, int64_t Mo, int64_t No, int64_t R);
int gemm_tc_red_batched(int n, const float* const* A, const int64_t* lda, const float* const* B, const int64_t* ldb,
                        float* const* C, const int64_t* ldc, int64_t Mo, int64_t No, int64_t R, void* ws, size_t ws_bytes,
                        cudaStream_t st);
int group_reduce_launch(const float* partial, const int* split_table, int S, const float* coef, int J, long long total,
                        float* out, cudaStream_t st);
bool tc_chain_ok(int n, int64_t M, int64_t W, const float* A, int64_t lda, const float* const* Wt, const int64_t* ldw,
                 float* const* C, const int64_t* ldc);
int gemm_tc_chain(int n, const float* A, int64_t lda, const float* const* Wt, const int64_t* ldw, const float* const* bias,
                  const int* flags, float* const* C, const int64_t* ldc, uint32_t* const* bits, int64_t M, int64_t W,
                  cudaStream_t st);
bool pna_post_fused_ok(int64_t rows, int64_t Ho, int64_t H, int64_t F, int64_t ldw, int64_t ld_out, int64_t ld_agg,
                       const void* weff, const void* out, const void* agg_out, const void* msg, const void* x);
int pna_post_fused(const float* msg, const int* rowptr, const int* row_node, int64_t rows, const float* x, int64_t H, int64_t F,
                   const float* weff, int64_t ldw, const int* bucket_ptr, int n_groups, const float* bias, int flags,
                   float* out, int64_t ld_out, int64_t Ho, uint32_t* bits_out, float* agg_out, int64_t ld_agg, cudaStream_t st);
}  // namespace gnnx

using namespace gnnx;

extern "C" {

size_t gnnx_gemm_workspace_bytes(int64_t M, int64_t N, int64_t K, int32_t ta, int32_t tb, int32_t precision) {
    size_t s = simt_workspace_bytes(M, N, K);
    if (!ta) {                                             // room for colsum_a_out: per-CTA partials or the two-pass kernel's
        size_t c = gnnx_colsum_workspace_bytes(M, K);
        size_t f = (size_t)num_sms() * 2 * K * sizeof(float);    // two column-sum warps per CTA (gemm_tc.cu kCsWarps)
        if (f > c) c = f;
        if (c > s) s = c;
    }
    if (precision != GNNX_PREC_FP32) {
        size_t t = tc_workspace_bytes(M, N, K, ta, tb, precision);
        if (t > s) s = t;
    }
    return s;
}

static int gemm_dispatch(const float* A, int64_t lda, int32_t ta, const float* B, int64_t ldb, int32_t tb, float* C, int64_t ldc,
              int64_t M, int64_t N, int64_t K, const float* bias, int32_t flags, const void* relu_mask_v, int64_t ld_mask,
              uint32_t* relu_bits_out, float* colsum_a_out, int32_t precision, void* ws, size_t ws_bytes, void* stream,
              const int32_t* bucket_ptr, int32_t n_groups) {
    const float* relu_mask = static_cast<const float*>(relu_mask_v);      // uint32 words under GNNX_GEMM_MASK_BITS
    GNNX_REQUIRE(M >= 0 && N >= 0 && K >= 0, "gemm: negative dimension");
    if (precision == GNNX_PREC_BF16) return set_error(GNNX_ERR_UNSUPPORTED, "gemm: no bf16 kernels are built (use GNNX_PREC_TF32)");
    GNNX_REQUIRE(precision == GNNX_PREC_FP32 || precision == GNNX_PREC_TF32, "gemm: unknown precision %d", precision);
    cudaStream_t st = as_stream(stream);
    if (M == 0 || N == 0) {                                // empty result; the column sum of a zero-row A is zero
        if (colsum_a_out && K > 0) GNNX_CUDA(cudaMemsetAsync(colsum_a_out, 0, (size_t)K * sizeof(float), st));
        return GNNX_OK;
    }
    // K == 0 (a weight gradient over zero rows: a batch without edges) is an empty sum: the fp32 kernel writes the
    // epilogue of a zero accumulator and never reads A / B, which may then be null (empty device tensors)
    GNNX_REQUIRE(C && (K == 0 || (A && B)), "gemm: null operand");
    const bool bits = (flags & GNNX_GEMM_MASK_BITS) != 0;
    GNNX_REQUIRE(!bits || relu_mask, "gemm: GNNX_GEMM_MASK_BITS without a mask");
    GNNX_REQUIRE(!bits || ld_mask >= (N + 31) / 32, "gemm: packed mask rows need at least ceil(N/32) words");
    GNNX_REQUIRE(!(relu_bits_out && (flags & GNNX_GEMM_ACCUMULATE)), "gemm: relu_bits_out is not available with ACCUMULATE");
    const bool mask_ok = !relu_mask || bits || (((uintptr_t)relu_mask & 15) == 0 && (ld_mask & 3) == 0);
    GNNX_REQUIRE(!(colsum_a_out && ta), "gemm: colsum_a_out needs trans_a = 0");
    // tiny products (60-row bond tables, weight-space 256^3) are launch-latency-bound: one CTA of the tensor-core
    // kernel would walk them alone, the 32x32-tile fp32 kernel spreads them over the SMs (and is exact)
    const bool force_tc = (flags & GNNX_GEMM_FORCE_TC) != 0;
    flags &= ~GNNX_GEMM_FORCE_TC;
    const bool tiny = !force_tc && (double)M * (double)N * (double)K < 33554432.0;
    // grouped tensor-core products: the row blocks of a 2-CTA cluster share one weight block (256-row buckets), and an
    // MN-major weight stack [G*K, N] must not let a k-block straddle two groups
    const bool group_ok = !bucket_ptr || (tb || K % 32 == 0);
    const bool tc = precision != GNNX_PREC_FP32 && !tiny && mask_ok && group_ok &&
                    tc_supported(M, N, K, ta, tb, lda, ldb, ldc, A, B, C, precision);
    if (force_tc && !tc)
        return set_error(GNNX_ERR_UNSUPPORTED, "gemm: GNNX_GEMM_FORCE_TC, but the tensor-core kernels do not take this product "
                         "(precision %d, M %lld N %lld K %lld, trans %d/%d, alignment / mask / group constraints)", precision,
                         (long long)M, (long long)N, (long long)K, ta, tb);
    const bool fuse_colsum = colsum_a_out && tc && K <= 256 && (K & 3) == 0;   // rides the tensor-core kernel's A tiles
    if (colsum_a_out && !fuse_colsum) {                    // separate two-pass kernel (same result, one more read of A)
        int rc = gnnx_colsum(A, lda, M, K, colsum_a_out, ws, ws_bytes, stream);
        if (rc) return rc;
    }
    if (tc)
        return gemm_tc(A, lda, ta, B, ldb, tb, C, ldc, M, N, K, bias, flags, relu_mask, ld_mask, relu_bits_out,
                       fuse_colsum ? colsum_a_out : nullptr, precision, ws, ws_bytes, st, bucket_ptr, n_groups);
    return gemm_simt(A, lda, ta, B, ldb, tb, C, ldc, M, N, K, bias, flags, relu_mask, ld_mask, relu_bits_out, ws, ws_bytes, st,
                     bucket_ptr, n_groups);
}

int gnnx_gemm(const float* A, int64_t lda, int32_t ta, const float* B, int64_t ldb, int32_t tb, float* C, int64_t ldc,
              int64_t M, int64_t N, int64_t K, const float* bias, int32_t flags, const void* relu_mask, int64_t ld_mask,
              uint32_t* relu_bits_out, float* colsum_a_out, int32_t precision, void* ws, size_t ws_bytes, void* stream) {
    return gemm_dispatch(A, lda, ta, B, ldb, tb, C, ldc, M, N, K, bias, flags, relu_mask, ld_mask, relu_bits_out, colsum_a_out,
                         precision, ws, ws_bytes, stream, nullptr, 1);
}

int gnnx_gemm_grouped(const float* A, int64_t lda, const float* B, int64_t ldb, int32_t tb, float* C, int64_t ldc,
                      int64_t M, int64_t N, int64_t K, const int32_t* bucket_ptr, int32_t n_groups, const float* bias,
                      int32_t flags, const void* relu_mask, int64_t ld_mask, uint32_t* relu_bits_out, float* colsum_a_out,
                      int32_t precision, void* ws, size_t ws_bytes, void* stream) {
    GNNX_REQUIRE(bucket_ptr && n_groups >= 1 && n_groups <= GNNX_MAX_GROUPS, "gemm_grouped: 1..%d groups and a bucket_ptr", GNNX_MAX_GROUPS);
    return gemm_dispatch(A, lda, 0, B, ldb, tb, C, ldc, M, N, K, bias, flags, relu_mask, ld_mask, relu_bits_out, colsum_a_out,
                         precision, ws, ws_bytes, stream, bucket_ptr, n_groups);
}

int gnnx_gemm_chain(int32_t n, const float* A, int64_t lda, const float* const* W, const int64_t* ld_w,
                    const float* const* bias, const int32_t* flags, float* const* C, const int64_t* ld_c,
                    uint32_t* const* relu_bits_out, int64_t M, int64_t width, int32_t precision, void* stream) {
    GNNX_REQUIRE(n >= 1 && A && W && ld_w && C && ld_c, "gemm_chain: null argument");
    GNNX_REQUIRE(M >= 0 && width >= 1, "gemm_chain: bad shape");
    if (M == 0) return GNNX_OK;
    static int chain = -1;
    if (chain < 0) { const char* e = getenv("GNNX_GEMM_CHAIN"); chain = (e && e[0] == '0') ? 0 : 1; }
    if (chain && precision != GNNX_PREC_FP32 && tc_chain_ok(n, M, width, A, lda, W, ld_w, C, ld_c))
        return gemm_tc_chain(n, A, lda, W, ld_w, bias, flags, C, ld_c, relu_bits_out, M, width, as_stream(stream));
    for (int p = 0; p < n; ++p) {                          // shapes / tiers the chain kernel does not take: one product at a time
        int rc = gnnx_gemm(p == 0 ? A : C[p - 1], p == 0 ? lda : ld_c[p - 1], 0, W[p], ld_w[p], 1, C[p], ld_c[p], M, width, width,
                           bias ? bias[p] : nullptr, flags ? flags[p] : 0, nullptr, 0, relu_bits_out ? relu_bits_out[p] : nullptr,
                           nullptr, precision, nullptr, 0, stream);
        if (rc) return rc;
    }
    return GNNX_OK;
}

int32_t gnnx_pna_post_fused_supported(int64_t rows, int32_t towers, int32_t f, int32_t h_out, int32_t precision) {
    const int64_t H = (int64_t)towers * f;
    if (precision == GNNX_PREC_FP32 || towers < 1 ||
        !pna_post_fused_ok(rows, h_out, H, f, 5 * H, h_out, 5 * H, nullptr, nullptr, nullptr, nullptr, nullptr))
        return 0;
    // 1: the kernel takes the shape; 2: and the product is large enough that gnnx_gemm would send it to the tensor cores
    return (double)rows * (double)h_out * (double)(5 * H) < 33554432.0 ? 1 : 2;
}

int gnnx_pna_post_fused_fwd(const float* m, const int32_t* rowptr, const int32_t* row_node, int64_t rows, const float* x,
                            int32_t towers, int32_t f, const float* weff, int64_t ld_w, const int32_t* bucket_ptr,
                            int32_t n_groups, const float* bias, int32_t relu, float* out, int64_t ld_out, int32_t h_out,
                            uint32_t* relu_bits_out, float* agg_out, int64_t ld_agg, int32_t precision, void* stream) {
    GNNX_REQUIRE(precision != GNNX_PREC_FP32, "pna_post_fused_fwd: tensor-core tier only (the fp32 tier runs aggregate + grouped product)");
    GNNX_REQUIRE(rowptr && row_node && x && weff && bucket_ptr && out, "pna_post_fused_fwd: null operand");
    GNNX_REQUIRE(n_groups >= 1 && n_groups <= GNNX_MAX_GROUPS && towers >= 1, "pna_post_fused_fwd: 1..%d groups", GNNX_MAX_GROUPS);
    const int64_t H = (int64_t)towers * f;
    if (!agg_out) ld_agg = 5 * H;
    if (!pna_post_fused_ok(rows, h_out, H, f, ld_w, ld_out, ld_agg, weff, out, agg_out, m, x))
        return set_error(GNNX_ERR_UNSUPPORTED, "pna_post_fused_fwd: shape / alignment outside the fused kernel's envelope "
                         "(rows %% 256, F %% 32, h_out %% 16; see gnnx_pna_post_fused_supported)");
    GNNX_REQUIRE(ld_w >= 5 * H && ld_out >= h_out && ld_agg >= 5 * H, "pna_post_fused_fwd: row strides too small");
    return pna_post_fused(m, rowptr, row_node, rows, x, H, f, weff, ld_w, bucket_ptr, n_groups, bias,
                          relu ? GNNX_GEMM_RELU : 0, out, ld_out, h_out, relu_bits_out, agg_out, ld_agg, as_stream(stream));
}

static bool dw_batched_tc(int32_t n, const float* const* dY, const int64_t* ld_dy, const float* const* X, const int64_t* ld_x,
                          float* const* dW, const int64_t* ld_dw, int64_t Mo, int64_t No, int64_t R, int32_t precision) {
    if (precision == GNNX_PREC_FP32 || !tc_red_batched_ok(n, Mo, No, R)) return false;
    for (int z = 0; z < n; ++z)
        if (!tc_supported(Mo, No, R, 1, 0, ld_dy[z], ld_x[z], ld_dw[z], dY[z], X[z], dW[z], precision)) return false;
    return true;
}

size_t gnnx_gemm_dw_batched_workspace_bytes(int32_t n, int64_t Mo, int64_t No, int64_t R, int32_t precision) {
    size_t s = gnnx_gemm_workspace_bytes(Mo, No, R, 1, 0, precision);          // the one-by-one path
    if (precision != GNNX_PREC_FP32 && tc_red_batched_ok(n, Mo, No, R)) {
        size_t t = tc_red_batched_workspace_bytes(n, Mo, No, R);
        if (t > s) s = t;
    }
    return s;
}

int gnnx_gemm_dw_batched(int32_t n, const float* const* dY, const int64_t* ld_dy, const float* const* X, const int64_t* ld_x,
                         float* const* dW, const int64_t* ld_dw, int64_t Mo, int64_t No, int64_t R, int32_t precision,
                         void* ws, size_t ws_bytes, void* stream) {
    GNNX_REQUIRE(n >= 1 && dY && X && dW && ld_dy && ld_x && ld_dw, "gemm_dw_batched: null argument");
    if (dw_batched_tc(n, dY, ld_dy, X, ld_x, dW, ld_dw, Mo, No, R, precision))
        return gemm_tc_red_batched(n, dY, ld_dy, X, ld_x, dW, ld_dw, Mo, No, R, ws, ws_bytes, as_stream(stream));
    for (int z = 0; z < n; ++z) {                          // shapes the batched kernel does not take: one product at a time
        int rc = gnnx_gemm(dY[z], ld_dy[z], 1, X[z], ld_x[z], 0, dW[z], ld_dw[z], Mo, No, R, nullptr, 0, nullptr, 0, nullptr,
                           nullptr, precision, ws, ws_bytes, stream);
        if (rc) return rc;
    }
    return GNNX_OK;
}

static bool grouped_dw_tc(int64_t Mo, int64_t No, int64_t R, int64_t lda, int64_t ldb, const void* A, const void* B, int precision) {
    return precision != GNNX_PREC_FP32 && tc_supported(Mo, No, R, 1, 0, lda, ldb, No, A, B, A, precision);
}

int32_t gnnx_gemm_grouped_dw_splits(int64_t Mo, int64_t No, int64_t R, int32_t n_groups, int32_t precision) {
    int s = precision != GNNX_PREC_FP32 ? tc_red_grouped_splits(Mo, No, R) : 2 * num_sms() / (ceil_div(Mo, 128) * ceil_div(No, 128));
    if (s < n_groups) s = n_groups;                       // every non-empty group needs a split of its own
    if (s < 1) s = 1;
    return s;
}

size_t gnnx_gemm_grouped_dw_workspace_bytes(int64_t Mo, int64_t No, int32_t splits) {
    return (size_t)splits * Mo * No * sizeof(float);
}

int gnnx_gemm_grouped_dw(const float* A, int64_t lda, const float* B, int64_t ldb, int64_t R, int64_t Mo, int64_t No,
                         const int32_t* split_table, int32_t splits, const float* coef, int32_t J, float* out,
                         int32_t precision, void* ws, size_t ws_bytes, void* stream) {
    GNNX_REQUIRE(A && B && out && split_table && coef, "gemm_grouped_dw: null operand");
    GNNX_REQUIRE(J >= 1 && J <= 4 && splits >= 1, "gemm_grouped_dw: 1..4 coefficient columns, at least one split");
    GNNX_REQUIRE((Mo * No) % 4 == 0, "gemm_grouped_dw: Mo * No must be a multiple of 4");
    GNNX_REQUIRE(ws && ws_bytes >= gnnx_gemm_grouped_dw_workspace_bytes(Mo, No, splits), "gemm_grouped_dw: workspace too small");
    cudaStream_t st = as_stream(stream);
    int rc;
    if (grouped_dw_tc(Mo, No, R, lda, ldb, A, B, precision))
        rc = gemm_tc_red_grouped(A, lda, B, ldb, Mo, No, R, split_table, splits, (float*)ws, st);
    else
        rc = gemm_simt_red_grouped(A, lda, B, ldb, Mo, No, R, split_table, splits, (float*)ws, st);
    if (rc) return rc;
    return group_reduce_launch((const float*)ws, split_table, splits, coef, J, (long long)Mo * No, out, st);
}

}  // extern "C"
