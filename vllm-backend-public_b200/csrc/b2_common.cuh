// b2_common.cuh -- shared helpers for libb200attn.so (sm_100a only).
#pragma once

#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_fp8.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "b200attn.h"

// ---------------------------------------------------------------------------------------------
// host: thread-local error text (never throws across the C ABI)
// ---------------------------------------------------------------------------------------------
void b2_set_error(const char* fmt, ...);

#define B2_REQUIRE(cond, code, ...)      \
    do {                                 \
        if (!(cond)) {                   \
            b2_set_error(__VA_ARGS__);   \
            return (code);               \
        }                                \
    } while (0)

#define B2_CUDA_OK(expr)                                                              \
    do {                                                                              \
        cudaError_t _e = (expr);                                                      \
        if (_e != cudaSuccess) {                                                      \
            b2_set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e),      \
                         __FILE__, __LINE__);                                         \
            return B2_ECUDA;                                                          \
        }                                                                             \
    } while (0)

// TMA map over a paged KV cache: one box = one (block, kv head) tile (paged_decode.cu)
int b2_internal_make_cache_tmap(CUtensorMap* tm, const void* cache, int kv_dtype, int64_t num_blocks, int Hkv, int bs,
                                 int64_t block_stride, int64_t head_stride);

int b2_internal_encode_bf16_sw128(CUtensorMap* tm, const void* ptr, int rank, const uint64_t* dims,
                                  const uint64_t* strides_bytes, const uint32_t* box);
// tcgen05 prompt-chunk attention (prefill_tc5.cu); same arguments as b2_prefill_attention after validation
int b2_internal_prefill_tc5(void* out, const void* q, const void* k_cache, const void* v_cache,
                            const int32_t* block_tables, const int32_t* cu_q, const int32_t* ctx_start, int n_chunks,
                            int T, int Hq, int Hkv, int bs, int max_nb, int64_t num_blocks, int64_t q_stride,
                            int64_t out_stride, int64_t cache_block_stride, int64_t cache_head_stride, float scale,
                            const float* k_scale, const float* v_scale, void* stream);

static inline bool b2_aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

#ifdef __CUDACC__
// ---------------------------------------------------------------------------------------------
// device: PTX wrappers (mbarrier, TMA, ldmatrix, mma.sync)
// ---------------------------------------------------------------------------------------------
namespace b2 {

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void fence_mbar_init() {
    // make the inits visible to the async (TMA) proxy
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async() {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    return ok != 0;
}
// Bounded wait: try_wait sleeps in hardware between probes, so the guard costs nothing in the
// common case; a protocol bug traps instead of hanging the GPU box.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t spins = 0;
    while (!mbar_try_wait(bar, parity)) {
        if (++spins > (1u << 26)) __trap();
    }
}

// L2 eviction policy for streamed-once KV tiles.
__device__ __forceinline__ uint64_t policy_evict_first() {
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}

// 4-D tiled TMA load, global -> shared, completion on an mbarrier (bytes).
__device__ __forceinline__ void tma_load_4d(uint32_t dst, const CUtensorMap* tmap, uint32_t bar,
                                            int c0, int c1, int c2, int c3, uint64_t policy) {
    asm volatile(
        "cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes"
        ".L2::cache_hint [%0], [%1, {%3, %4, %5, %6}], [%2], %7;"
        ::"r"(dst), "l"(tmap), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "l"(policy)
        : "memory");
}
__device__ __forceinline__ void tma_load_5d(uint32_t dst, const CUtensorMap* tmap, uint32_t bar,
                                            int c0, int c1, int c2, int c3, int c4, uint64_t policy) {
    asm volatile(
        "cp.async.bulk.tensor.5d.shared::cluster.global.tile.mbarrier::complete_tx::bytes"
        ".L2::cache_hint [%0], [%1, {%3, %4, %5, %6, %7}], [%2], %8;"
        ::"r"(dst), "l"(tmap), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4), "l"(policy)
        : "memory");
}
// 1-D bulk copy global -> shared (16-byte aligned, size a multiple of 16), completion on an mbarrier (bytes).
__device__ __forceinline__ void bulk_load(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* tmap) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(tmap) : "memory");
}

__device__ __forceinline__ void ldmatrix_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2,
                                            uint32_t& r3) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
                 : "r"(addr));
}
__device__ __forceinline__ void ldmatrix_x4_trans(uint32_t addr, uint32_t& r0, uint32_t& r1,
                                                  uint32_t& r2, uint32_t& r3) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
                 : "r"(addr));
}
__device__ __forceinline__ uint32_t movmatrix_trans(uint32_t a) {
    uint32_t d;
    asm volatile("movmatrix.sync.aligned.m8n8.trans.b16 %0, %1;" : "=r"(d) : "r"(a));
    return d;
}
// D(16x8,f32) += A(16x16,bf16,row) * B(16x8,bf16,col)
__device__ __forceinline__ void mma_bf16_16816(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2,
                                               uint32_t a3, uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
        "{%0,%1,%2,%3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
    uint32_t r;
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
    return r;
}
__device__ __forceinline__ float fast_exp2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ int4 ld_nc_16(const void* p) {
    int4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    return r;
}

// Generic-proxy global stores of this thread become visible to later async-proxy (TMA) reads of global memory.
__device__ __forceinline__ void fence_proxy_async_global() {
    asm volatile("fence.proxy.async.global;" ::: "memory");
}
// 8 bf16 (one 16-byte vector) -> 8 e4m3 bytes: cvt.rn.satfinite.e4m3x2.f32(x / scale), IEEE division -- the KV-write
// encoding of b2_reshape_and_cache (bit-exact against the oracle), shared with the fused append of paged_decode.cu.
__device__ __forceinline__ uint2 pack_e4m3x8(const int4& val, float s) {
    const uint32_t w[4] = {static_cast<uint32_t>(val.x), static_cast<uint32_t>(val.y),
                           static_cast<uint32_t>(val.z), static_cast<uint32_t>(val.w)};
    uint32_t packed[2];
#pragma unroll
    for (int p = 0; p < 2; ++p) {
        uint32_t acc = 0;
#pragma unroll
        for (int q = 0; q < 2; ++q) {
            const uint32_t word = w[p * 2 + q];
            // bf16 -> f32 is a 16-bit shift; the division is IEEE (matches the oracle)
            const float lo = __fdiv_rn(__uint_as_float(word << 16), s);
            const float hi = __fdiv_rn(__uint_as_float(word & 0xffff0000u), s);
            const __nv_fp8x2_storage_t pr = __nv_cvt_float2_to_fp8x2(make_float2(lo, hi), __NV_SATFINITE, __NV_E4M3);
            acc |= static_cast<uint32_t>(pr) << (16 * q);
        }
        packed[p] = acc;
    }
    return make_uint2(packed[0], packed[1]);
}

}  // namespace b2
#endif  // __CUDACC__
