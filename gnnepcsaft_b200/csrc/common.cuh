// Shared helpers for libgnnx_sm100.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/gnnx.h"

namespace gnnx {

// Thread-local last-error message, read through gnnx_last_error().
int set_error(int code, const char* fmt, ...);

// Number of kernels this library has launched (read through gnnx_launch_count()), and how many of them were
// tcgen05 tensor-core kernels (gnnx_tc_launch_count(): lets a test tell the tensor path from the fp32 fallback).
void count_launches(int n);
void count_tc_launches(int n);

inline int check_launch(const char* what, int kernels = 1) {
    count_launches(kernels);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return set_error((int)e, "%s: %s", what, cudaGetErrorString(e));
    return GNNX_OK;
}

#define GNNX_CUDA(call)                                                                   \
    do {                                                                                  \
        cudaError_t _e = (call);                                                          \
        if (_e != cudaSuccess)                                                            \
            return gnnx::set_error((int)_e, "%s:%d %s", __FILE__, __LINE__, cudaGetErrorString(_e)); \
    } while (0)

#define GNNX_REQUIRE(cond, ...)                                                           \
    do {                                                                                  \
        if (!(cond)) return gnnx::set_error(GNNX_ERR_INVALID, __VA_ARGS__);               \
    } while (0)

// ---- programmatic dependent launch (PDL) ------------------------------------------------------------------------
// Every gnnx kernel is launched through launch() with cudaLaunchAttributeProgrammaticStreamSerialization and starts
// with pdl_enter(): `griddepcontrol.wait` blocks until the preceding kernel of the stream has completed and its
// writes are visible (so NO global memory may be touched before it), `griddepcontrol.launch_dependents` then lets the
// next kernel's CTAs become resident as this kernel's CTAs retire: its launch latency, prologue and ramp-up overlap
// this kernel's tail instead of following it (measured: 2 us per kernel boundary, ~400 boundaries per cfg2 step).
// GNNX_PDL=0 in the environment launches everything fully serialised (the wait / trigger are then no-ops).
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_enter() { pdl_wait(); pdl_trigger(); }

// Grid-wide barrier for kernels whose whole grid is resident (the host sizes the grid: SMs x co-resident CTAs): an
// arrival counter in global memory that thread 0 of each CTA bumps and then polls with acquire loads.  `target` =
// arrivals expected so far (grid size x barrier ordinal).  A counter that never fills traps after ~2 s of SM clocks
// instead of hanging the GPU.  Used by the cooperative BatchNorm kernels (norm.cu) and the split fold of the weight-
// gradient products (gemm_tc.cu); counters come from stream_ticket() and are left at 0 by the last CTA out.
__device__ __forceinline__ void grid_barrier(int* counter, int target) {
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        atomicAdd(counter, 1);
        long long t0 = 0;
        for (unsigned it = 0;; ++it) {
            int v;
            asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(counter) : "memory");
            if (v >= target) break;
            __nanosleep(40);
            if ((it & 4095u) == 4095u) {
                long long now = clock64();
                if (t0 == 0) t0 = now;
                else if (now - t0 > 4000000000ll) __trap();
            }
        }
        __threadfence();
    }
    __syncthreads();
}

bool pdl_enabled();
// Per-kernel device timeline (gnnx_trace_begin / gnnx_trace_end): while a trace is open every launch() is bracketed by
// two CUDA events on its stream, so the report gives each kernel's in-situ duration (warm caches, real predecessors)
// and the time the stream spent between gnnx kernels (other libraries' kernels, launch gaps).
int* stream_ticket(cudaStream_t stream);   // gemm_tc.cu: 64 zero-initialised ints per (device, stream); every user leaves its ints at 0
bool trace_on();
void trace_before(const void* func, cudaStream_t st);
void trace_after(cudaStream_t st);

template <typename... KArgs, typename... Args>
inline cudaError_t launch(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    if (pdl_enabled()) { cfg.attrs = attr; cfg.numAttrs = 1; }
    if (!trace_on()) return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
    trace_before(reinterpret_cast<const void*>(kernel), st);
    cudaError_t e = cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
    trace_after(st);
    return e;
}

// Launch of a kernel that contains a grid barrier.  Default: like launch() -- the host has sized the grid so that every
// CTA fits on the GPU at once, and a grid is dispatched whole before a later one starts.  With gang launch ON
// (gnnx_set_gang_launch(1); train.Trainer turns it on for data-parallel runs, GNNX_GANG_LAUNCH=1 forces it) the kernel is
// launched with cudaLaunchAttributeCooperative instead of the programmatic-serialisation attribute: the driver then
// guarantees that all CTAs are co-resident before any runs, whatever another stream (NCCL's all-reduce kernel, launched
// from inside the backward) is doing to the SMs at that moment -- a partially resident grid spinning at its barrier next
// to a partially resident collective could otherwise wait for each other.  Costs the launch overlap (~2 us per kernel).
bool gang_launch();
void set_gang_launch(int on);
template <typename... KArgs, typename... Args>
inline cudaError_t launch_coop(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
    if (!gang_launch()) return launch(kernel, grid, block, smem, st, static_cast<Args&&>(args)...);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeCooperative;
    attr[0].val.cooperative = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    if (!trace_on()) return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
    trace_before(reinterpret_cast<const void*>(kernel), st);
    cudaError_t e = cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
    trace_after(st);
    return e;
}

static inline int ceil_div(long long a, long long b) { return (int)((a + b - 1) / b); }
static inline cudaStream_t as_stream(void* s) { return reinterpret_cast<cudaStream_t>(s); }

// SM count of the current device (148 on B200), queried once per device; grids of grid-stride kernels are sized in
// multiples of it.
int num_sms();

__device__ __forceinline__ float4 ldg4(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }
__device__ __forceinline__ void stg4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }
// streaming (evict-first) variants for data touched exactly once
__device__ __forceinline__ float4 ldg4_stream(const float* p) { return __ldcs(reinterpret_cast<const float4*>(p)); }
__device__ __forceinline__ void stg4_stream(float* p, float4 v) { __stcs(reinterpret_cast<float4*>(p), v); }

// Counter-based dropout draw: keep(index) is a pure function of (seed, index), so a forward kernel and its backward agree
// without a stored mask.  Used per (slot, head) by the attention kernels and per element by the node-feature dropout
// fused into the BatchNorm apply / embedding kernels.  P(drop) = threshold / 2^32.
__device__ __forceinline__ bool dropout_keep(unsigned long long seed, long long index, unsigned int threshold) {
    unsigned long long z = seed + 0x9E3779B97F4A7C15ull * (unsigned long long)(index + 1);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;          // splitmix64 finaliser
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    z ^= z >> 31;
    return (unsigned int)(z >> 32) >= threshold;
}
static inline unsigned int dropout_threshold(float p) {
    double t = (double)p * 4294967296.0;
    return t <= 0.0 ? 0u : (t >= 4294967295.0 ? 4294967295u : (unsigned int)t);
}
// Node-feature dropout applied in an epilogue: thr == 0 means "off"; kept values are scaled by 1 / (1 - p).
struct NodeDropout {
    unsigned long long seed;
    unsigned int thr;
    float scale;
};
static inline NodeDropout node_dropout(float p, long long seed) {
    NodeDropout d;
    d.seed = (unsigned long long)seed;
    d.thr = dropout_threshold(p);
    d.scale = p > 0.f && p < 1.f ? 1.f / (1.f - p) : (p >= 1.f ? 0.f : 1.f);
    return d;
}
__device__ __forceinline__ float drop_apply(const NodeDropout& d, long long index, float v) {
    return d.thr == 0u ? v : (dropout_keep(d.seed, index, d.thr) ? v * d.scale : 0.f);
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

}  // namespace gnnx
