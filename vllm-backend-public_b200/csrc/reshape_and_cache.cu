// reshape_and_cache.cu -- KV write through slot_mapping.
//
// Contract: include/b200attn.h `b2_reshape_and_cache`.  Semantics follow vLLM's
// `reshape_and_cache` (the op the engine launched by reference
// benchmark-vllm/vllm_manager.py:371-374 runs after every QKV projection), on the HND cache
// layout [num_blocks, Hkv, bs, d].
//
// HBM-bound byte mover: one CTA per token, one 16-byte vector (8 bf16) per thread per trip,
// K and V handled by the same CTA.  Loads go through the non-allocating read-only path; the
// scattered 256-byte row stores are full 32-byte sectors.  bf16 is a bit-exact copy; e4m3 is
// `cvt.rn.satfinite.e4m3x2.f32(x / scale)`.
#include "b2_common.cuh"

namespace {

template <bool FP8>
__global__ void __launch_bounds__(256)
reshape_and_cache_kernel(const __nv_bfloat16* __restrict__ k_new,
                         const __nv_bfloat16* __restrict__ v_new,
                         void* __restrict__ k_cache, void* __restrict__ v_cache,
                         const int64_t* __restrict__ slot_mapping,
                         int Hkv, int d, int bs,
                         int64_t k_token_stride, int64_t v_token_stride,
                         int64_t cache_block_stride, int64_t cache_head_stride,
                         const float* __restrict__ k_scale, const float* __restrict__ v_scale,
                         int32_t* __restrict__ block_tables, const int32_t* __restrict__ seq_lens, int max_nb) {
    // the decode kernel that follows in the stream may be launched now (programmatic dependent launch): its prologue
    // runs under this kernel; it waits (griddepcontrol.wait) before it touches the cache or the block tables
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    const int t = blockIdx.x;
    const int64_t slot = slot_mapping[t];
    if (slot < 0) return;  // padding token
    const int64_t blk = slot / bs;
    const int off = static_cast<int>(slot - blk * bs);
    // decode step with device-resident block tables: token t is the newest token of batch row t, at position
    // seq_lens[t] - 1, so its slot names the block of that table column (idempotent: rewritten every step)
    if (block_tables && threadIdx.x == 0) {
        const int L = seq_lens[t];
        if (L > 0 && (L - 1) / bs < max_nb) block_tables[(int64_t)t * max_nb + (L - 1) / bs] = static_cast<int32_t>(blk);
    }
    const int vec_per_head = d >> 3;
    const int nvec = Hkv * vec_per_head;

    float ks = 1.f, vs = 1.f;
    if (FP8) {
        ks = *k_scale;
        vs = *v_scale;
    }

    for (int i = threadIdx.x; i < 2 * nvec; i += blockDim.x) {
        const bool is_v = i >= nvec;
        const int j = is_v ? i - nvec : i;
        const int h = j / vec_per_head;
        const int c = j - h * vec_per_head;
        const __nv_bfloat16* src =
            (is_v ? v_new + t * v_token_stride : k_new + t * k_token_stride) + h * d + c * 8;
        const int4 val = b2::ld_nc_16(src);
        const int64_t dst_elem =
            blk * cache_block_stride + h * cache_head_stride + static_cast<int64_t>(off) * d + c * 8;
        if (!FP8) {
            __nv_bfloat16* dst = static_cast<__nv_bfloat16*>(is_v ? v_cache : k_cache) + dst_elem;
            *reinterpret_cast<int4*>(dst) = val;
        } else {
            uint8_t* dst = static_cast<uint8_t*>(is_v ? v_cache : k_cache) + dst_elem;
            *reinterpret_cast<uint2*>(dst) = b2::pack_e4m3x8(val, is_v ? vs : ks);
        }
    }
}

// fp8 -> bf16 gather of whole cache blocks: out block i = float(e4m3 block src[i]), exact in bf16 (4 significant
// bits, exponent range inside bf16's); the per-tensor scales are applied by the consumer (b2_prefill_attention).
// grid (n blocks, K|V); a thread converts 16 cache bytes per trip (one 16-byte load, two 16-byte stores).
__global__ void __launch_bounds__(256)
dequant_blocks_kernel(__nv_bfloat16* __restrict__ k_out, __nv_bfloat16* __restrict__ v_out,
                      const uint8_t* __restrict__ k_cache, const uint8_t* __restrict__ v_cache,
                      const int32_t* __restrict__ src_blocks, int Hkv, int row_elems /* bs * d */,
                      int64_t cache_block_stride, int64_t cache_head_stride) {
    const int64_t i = blockIdx.x;
    const bool is_v = blockIdx.y != 0;
    const uint8_t* src = (is_v ? v_cache : k_cache) + (int64_t)src_blocks[i] * cache_block_stride;
    __nv_bfloat16* dst = (is_v ? v_out : k_out) + i * (int64_t)Hkv * row_elems;
    const int vec_per_head = row_elems >> 4;
    for (int v = threadIdx.x; v < Hkv * vec_per_head; v += blockDim.x) {
        const int h = v / vec_per_head, c = v - h * vec_per_head;
        const uint4 raw = *reinterpret_cast<const uint4*>(src + (int64_t)h * cache_head_stride + c * 16);
        const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
        uint32_t o[8];
#pragma unroll
        for (int q = 0; q < 4; ++q)
#pragma unroll
            for (int hh = 0; hh < 2; ++hh) {
                const __half2_raw h2 = __nv_cvt_fp8x2_to_halfraw2((__nv_fp8x2_storage_t)(w[q] >> (16 * hh)), __NV_E4M3);
                const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&h2));
                o[q * 2 + hh] = b2::pack_bf16x2(f.x, f.y);
            }
        uint4* d4 = reinterpret_cast<uint4*>(dst + (int64_t)h * row_elems + c * 16);
        d4[0] = make_uint4(o[0], o[1], o[2], o[3]);
        d4[1] = make_uint4(o[4], o[5], o[6], o[7]);
    }
}

int reshape_and_cache_impl(const void* k_new, const void* v_new, void* k_cache,
                           void* v_cache, const int64_t* slot_mapping, int T, int Hkv,
                           int d, int bs, int64_t k_token_stride, int64_t v_token_stride,
                           int64_t cache_block_stride, int64_t cache_head_stride,
                           int kv_dtype, const float* k_scale, const float* v_scale,
                           int32_t* block_tables, const int32_t* seq_lens, int max_nb, void* stream) {
    B2_REQUIRE(T >= 0 && Hkv > 0 && d > 0 && bs > 0, B2_EINVAL,
               "reshape_and_cache: bad sizes T=%d Hkv=%d d=%d bs=%d", T, Hkv, d, bs);
    if (T == 0) return B2_OK;
    B2_REQUIRE(k_new && v_new && k_cache && v_cache && slot_mapping, B2_EINVAL,
               "reshape_and_cache: null pointer");
    B2_REQUIRE(kv_dtype == B2_KV_BF16 || kv_dtype == B2_KV_FP8_E4M3, B2_EUNSUPPORTED,
               "reshape_and_cache: kv_dtype %d", kv_dtype);
    B2_REQUIRE(d % 8 == 0, B2_EUNSUPPORTED, "reshape_and_cache: d=%d must be a multiple of 8", d);
    B2_REQUIRE(b2_aligned16(k_new) && b2_aligned16(v_new) && b2_aligned16(k_cache) &&
                   b2_aligned16(v_cache),
               B2_EINVAL, "reshape_and_cache: pointers must be 16-byte aligned");
    B2_REQUIRE(k_token_stride % 8 == 0 && v_token_stride % 8 == 0 && cache_block_stride % 16 == 0 &&
                   cache_head_stride % 16 == 0,
               B2_EINVAL, "reshape_and_cache: strides must keep 16-byte alignment");
    B2_REQUIRE(k_token_stride >= (int64_t)Hkv * d && v_token_stride >= (int64_t)Hkv * d &&
                   cache_head_stride >= (int64_t)bs * d,
               B2_EINVAL, "reshape_and_cache: strides overlap");
    if (kv_dtype == B2_KV_FP8_E4M3)
        B2_REQUIRE(k_scale && v_scale, B2_EINVAL, "reshape_and_cache: fp8 needs k_scale/v_scale");

    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int nvec2 = 2 * Hkv * (d / 8);
    int threads = nvec2 < 256 ? ((nvec2 + 31) / 32) * 32 : 256;
    const __nv_bfloat16* kn = static_cast<const __nv_bfloat16*>(k_new);
    const __nv_bfloat16* vn = static_cast<const __nv_bfloat16*>(v_new);
    if (kv_dtype == B2_KV_BF16) {
        reshape_and_cache_kernel<false><<<T, threads, 0, st>>>(
            kn, vn, k_cache, v_cache, slot_mapping, Hkv, d, bs, k_token_stride, v_token_stride,
            cache_block_stride, cache_head_stride, nullptr, nullptr, block_tables, seq_lens, max_nb);
    } else {
        reshape_and_cache_kernel<true><<<T, threads, 0, st>>>(
            kn, vn, k_cache, v_cache, slot_mapping, Hkv, d, bs, k_token_stride, v_token_stride,
            cache_block_stride, cache_head_stride, k_scale, v_scale, block_tables, seq_lens, max_nb);
    }
    B2_CUDA_OK(cudaGetLastError());
    return B2_OK;
}

}  // namespace

extern "C" int b2_reshape_and_cache(const void* k_new, const void* v_new, void* k_cache,
                                    void* v_cache, const int64_t* slot_mapping, int T, int Hkv,
                                    int d, int bs, int64_t k_token_stride, int64_t v_token_stride,
                                    int64_t cache_block_stride, int64_t cache_head_stride,
                                    int kv_dtype, const float* k_scale, const float* v_scale,
                                    void* stream) {
    return reshape_and_cache_impl(k_new, v_new, k_cache, v_cache, slot_mapping, T, Hkv, d, bs, k_token_stride,
                                  v_token_stride, cache_block_stride, cache_head_stride, kv_dtype, k_scale, v_scale,
                                  nullptr, nullptr, 0, stream);
}

extern "C" int b2_reshape_and_cache_decode(const void* k_new, const void* v_new, void* k_cache, void* v_cache,
                                           const int64_t* slot_mapping, int32_t* block_tables, const int32_t* seq_lens,
                                           int B, int Hkv, int d, int bs, int max_nb, int64_t k_token_stride,
                                           int64_t v_token_stride, int64_t cache_block_stride, int64_t cache_head_stride,
                                           int kv_dtype, const float* k_scale, const float* v_scale, void* stream) {
    B2_REQUIRE(B == 0 || (block_tables && seq_lens && max_nb > 0), B2_EINVAL,
               "reshape_and_cache_decode: block_tables / seq_lens / max_nb");
    return reshape_and_cache_impl(k_new, v_new, k_cache, v_cache, slot_mapping, B, Hkv, d, bs, k_token_stride,
                                  v_token_stride, cache_block_stride, cache_head_stride, kv_dtype, k_scale, v_scale,
                                  block_tables, seq_lens, max_nb, stream);
}

extern "C" int b2_dequant_kv_blocks(void* k_out, void* v_out, const void* k_cache, const void* v_cache,
                                    const int32_t* src_blocks, int64_t n, int Hkv, int d, int bs,
                                    int64_t cache_block_stride, int64_t cache_head_stride, void* stream) {
    B2_REQUIRE(n >= 0 && Hkv > 0 && d > 0 && bs > 0, B2_EINVAL, "dequant_kv_blocks: bad sizes n=%lld Hkv=%d d=%d bs=%d",
               (long long)n, Hkv, d, bs);
    if (n == 0) return B2_OK;
    B2_REQUIRE(k_out && v_out && k_cache && v_cache && src_blocks, B2_EINVAL,
               "dequant_kv_blocks: null pointer");
    B2_REQUIRE(((int64_t)bs * d) % 16 == 0, B2_EUNSUPPORTED, "dequant_kv_blocks: bs * d = %d must be a multiple of 16", bs * d);
    B2_REQUIRE(b2_aligned16(k_out) && b2_aligned16(v_out) && b2_aligned16(k_cache) && b2_aligned16(v_cache), B2_EINVAL,
               "dequant_kv_blocks: pointers must be 16-byte aligned");
    B2_REQUIRE(cache_block_stride % 16 == 0 && cache_head_stride % 16 == 0 && cache_head_stride >= (int64_t)bs * d,
               B2_EINVAL, "dequant_kv_blocks: cache strides must keep 16-byte alignment");
    B2_REQUIRE(n <= 0x7fffffffLL, B2_EUNSUPPORTED, "dequant_kv_blocks: %lld blocks exceed the grid", (long long)n);
    dequant_blocks_kernel<<<dim3((unsigned)n, 2), 256, 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<__nv_bfloat16*>(k_out), static_cast<__nv_bfloat16*>(v_out), static_cast<const uint8_t*>(k_cache),
        static_cast<const uint8_t*>(v_cache), src_blocks, Hkv, bs * d, cache_block_stride, cache_head_stride);
    B2_CUDA_OK(cudaGetLastError());
    return B2_OK;
}
