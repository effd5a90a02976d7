// K0: ingest. fp32 -> bf16 cast fused with the per-row score multiplier (inverse L2 norm of the
// bf16-ROUNDED row, 0 for a zero row; 1 for inner product). One warp per row, 128-bit loads.
// HBM-bound: 4 B read + 2 B written per element (fp32 source) or 2 B read (bf16 source).
// Replaces the server side of collection.add (super_rag/vectorstore/seekdb_connector.py:68-85);
// zero-norm rule from graphiti_core/search/search_utils.py:63-64.
#include "common.cuh"

namespace {

constexpr int ING_THREADS = 256;  // 8 rows per CTA

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, o);
    return v;
}

__device__ __forceinline__ float scale_from_ss(float ss, int metric) {
    if (metric == SRAG_METRIC_IP) return 1.0f;
    return ss > 0.0f ? 1.0f / sqrtf(ss) : 0.0f;
}

__global__ void __launch_bounds__(ING_THREADS)
ingest_f32_kernel(const float* __restrict__ src, int64_t n, int dim, __nv_bfloat16* __restrict__ dst,
                  float* __restrict__ scale, int metric) {
    const int lane = threadIdx.x & 31;
    const int64_t warps_total = (int64_t)gridDim.x * (ING_THREADS / 32);
    for (int64_t row = (int64_t)blockIdx.x * (ING_THREADS / 32) + (threadIdx.x >> 5); row < n; row += warps_total) {
        const float* s = src + row * dim;
        __nv_bfloat16* d = dst + row * dim;
        float ss = 0.f;
        for (int i = lane * 4; i < dim; i += 128) {
            const float4 v = __ldg(reinterpret_cast<const float4*>(s + i));
            const __nv_bfloat16 b0 = __float2bfloat16_rn(v.x), b1 = __float2bfloat16_rn(v.y);
            const __nv_bfloat16 b2 = __float2bfloat16_rn(v.z), b3 = __float2bfloat16_rn(v.w);
            const float f0 = __bfloat162float(b0), f1 = __bfloat162float(b1);
            const float f2 = __bfloat162float(b2), f3 = __bfloat162float(b3);
            ss += f0 * f0 + f1 * f1 + f2 * f2 + f3 * f3;
            uint2 pk;
            pk.x = (uint32_t)__bfloat16_as_ushort(b0) | ((uint32_t)__bfloat16_as_ushort(b1) << 16);
            pk.y = (uint32_t)__bfloat16_as_ushort(b2) | ((uint32_t)__bfloat16_as_ushort(b3) << 16);
            *reinterpret_cast<uint2*>(d + i) = pk;
        }
        ss = warp_sum(ss);
        if (lane == 0) scale[row] = scale_from_ss(ss, metric);
    }
}

__global__ void __launch_bounds__(ING_THREADS)
scale_bf16_kernel(const __nv_bfloat16* __restrict__ rows, int64_t n, int dim, float* __restrict__ scale, int metric) {
    const int lane = threadIdx.x & 31;
    const int64_t warps_total = (int64_t)gridDim.x * (ING_THREADS / 32);
    for (int64_t row = (int64_t)blockIdx.x * (ING_THREADS / 32) + (threadIdx.x >> 5); row < n; row += warps_total) {
        const __nv_bfloat16* s = rows + row * dim;
        float ss = 0.f;
        for (int i = lane * 8; i < dim; i += 256) {
            const uint4 v = __ldg(reinterpret_cast<const uint4*>(s + i));
            const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float lo = __uint_as_float(w[j] << 16);
                const float hi = __uint_as_float(w[j] & 0xFFFF0000u);
                ss += lo * lo + hi * hi;
            }
        }
        ss = warp_sum(ss);
        if (lane == 0) scale[row] = scale_from_ss(ss, metric);
    }
}

int ingest_grid(int64_t n) {
    int64_t blocks = (n + (ING_THREADS / 32) - 1) / (ING_THREADS / 32);
    const int64_t cap = 148 * 8 * 4;  // persistent-ish: a few waves of 8 CTAs/SM
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return (int)blocks;
}

}  // namespace

extern "C" int srag_rows_ingest_f32(const float* src_f32, int64_t n, int dim, void* dst_bf16, float* scale,
                                    int metric, void* stream) {
    SRAG_CHECK_ARG(n >= 0 && dim >= 8 && dim % 8 == 0, "bad sizes n=%lld dim=%d", (long long)n, dim);
    SRAG_CHECK_ARG(metric == SRAG_METRIC_COSINE || metric == SRAG_METRIC_IP, "bad metric %d", metric);
    if (n == 0) return SRAG_OK;
    SRAG_CHECK_ARG(src_f32 && dst_bf16 && scale, "null pointer argument");
    SRAG_CHECK_ARG(((uintptr_t)src_f32 & 15) == 0 && ((uintptr_t)dst_bf16 & 15) == 0, "rows must be 16-byte aligned");
    ingest_f32_kernel<<<ingest_grid(n), ING_THREADS, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
        src_f32, n, dim, reinterpret_cast<__nv_bfloat16*>(dst_bf16), scale, metric);
    SRAG_CUDA(cudaGetLastError());
    return SRAG_OK;
}

extern "C" int srag_rows_scale_bf16(const void* rows_bf16, int64_t n, int dim, float* scale, int metric,
                                    void* stream) {
    SRAG_CHECK_ARG(n >= 0 && dim >= 8 && dim % 8 == 0, "bad sizes n=%lld dim=%d", (long long)n, dim);
    SRAG_CHECK_ARG(metric == SRAG_METRIC_COSINE || metric == SRAG_METRIC_IP, "bad metric %d", metric);
    if (n == 0) return SRAG_OK;
    SRAG_CHECK_ARG(rows_bf16 && scale, "null pointer argument");
    SRAG_CHECK_ARG(((uintptr_t)rows_bf16 & 15) == 0, "rows must be 16-byte aligned");
    scale_bf16_kernel<<<ingest_grid(n), ING_THREADS, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
        reinterpret_cast<const __nv_bfloat16*>(rows_bf16), n, dim, scale, metric);
    SRAG_CUDA(cudaGetLastError());
    return SRAG_OK;
}
