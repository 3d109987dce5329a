// paged_decode_simt.cu -- plain CUDA-core paged-attention decode (no TMA, no tensor cores).
//
// Same contract as b2_paged_decode (include/b200attn.h).  This is NOT the product path: it is the
// "recompiled baseline" variant kept for the ncu A/B comparison in DESIGN.md / profiles/ and as an
// independent second implementation for the parity tests.  One CTA per (sequence, kv head); each
// of 8 warps walks tokens warp, warp+8, ...; a lane owns 4 of the 128 head_dim elements, so one
// warp-wide load is one contiguous token row.  Dot products are reduced with xor-shuffles.
#include "b2_common.cuh"

namespace {

constexpr int D = 128;
constexpr int NW = 8;
constexpr float LOG2E = 1.4426950408889634f;

template <bool FP8>
__device__ __forceinline__ void load_row4(const void* cache, int64_t elem, float (&x)[4]) {
    if (!FP8) {
        const uint2 w = *reinterpret_cast<const uint2*>(static_cast<const __nv_bfloat16*>(cache) + elem);
        x[0] = __uint_as_float(w.x << 16);
        x[1] = __uint_as_float(w.x & 0xffff0000u);
        x[2] = __uint_as_float(w.y << 16);
        x[3] = __uint_as_float(w.y & 0xffff0000u);
    } else {
        const uint32_t w = *reinterpret_cast<const uint32_t*>(static_cast<const uint8_t*>(cache) + elem);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const __half_raw hr = __nv_cvt_fp8_to_halfraw(static_cast<__nv_fp8_storage_t>((w >> (8 * i)) & 0xff), __NV_E4M3);
            x[i] = __half2float(__half(hr));
        }
    }
}

template <bool FP8, int G>
__global__ void __launch_bounds__(NW * 32)
paged_decode_simt_kernel(__nv_bfloat16* __restrict__ out, const __nv_bfloat16* __restrict__ q,
                         const void* __restrict__ k_cache, const void* __restrict__ v_cache,
                         const int32_t* __restrict__ block_tables, const int32_t* __restrict__ seq_lens,
                         int bs, int max_nb, int g0, int Gtotal, int64_t q_stride, int64_t out_stride,
                         int64_t block_stride, int64_t head_stride, float scale_log2,
                         const float* __restrict__ k_scale, const float* __restrict__ v_scale) {
    const int kvh = blockIdx.x, b = blockIdx.y;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int seq_len = seq_lens[b];
    const int head0 = kvh * Gtotal + g0;
    float sl2 = scale_log2;
    float vs = 1.f;
    if (FP8) {
        sl2 *= *k_scale;
        vs = *v_scale;
    }

    float qr[G][4];
#pragma unroll
    for (int h = 0; h < G; ++h) {
        const uint2 w = *reinterpret_cast<const uint2*>(q + b * q_stride + (int64_t)(head0 + h) * D + lane * 4);
        qr[h][0] = __uint_as_float(w.x << 16);
        qr[h][1] = __uint_as_float(w.x & 0xffff0000u);
        qr[h][2] = __uint_as_float(w.y << 16);
        qr[h][3] = __uint_as_float(w.y & 0xffff0000u);
    }
    float m[G], l[G], acc[G][4];
#pragma unroll
    for (int h = 0; h < G; ++h) {
        m[h] = -INFINITY;
        l[h] = 0.f;
        acc[h][0] = acc[h][1] = acc[h][2] = acc[h][3] = 0.f;
    }

    const int32_t* bt = block_tables + (int64_t)b * max_nb;
    for (int t = warp; t < seq_len; t += NW) {
        const int blk = bt[t / bs];
        const int64_t elem = (int64_t)blk * block_stride + (int64_t)kvh * head_stride + (int64_t)(t % bs) * D + lane * 4;
        float kx[4], vx[4];
        load_row4<FP8>(k_cache, elem, kx);
        load_row4<FP8>(v_cache, elem, vx);
#pragma unroll
        for (int h = 0; h < G; ++h) {
            float s = qr[h][0] * kx[0] + qr[h][1] * kx[1] + qr[h][2] * kx[2] + qr[h][3] * kx[3];
#pragma unroll
            for (int off = 16; off > 0; off >>= 1) s += __shfl_xor_sync(0xffffffffu, s, off);
            const float m_new = fmaxf(m[h], s);
            const float alpha = exp2f((m[h] - m_new) * sl2);
            const float pj = exp2f((s - m_new) * sl2);
            m[h] = m_new;
            l[h] = l[h] * alpha + pj;
#pragma unroll
            for (int i = 0; i < 4; ++i) acc[h][i] = acc[h][i] * alpha + pj * vx[i];
        }
    }

    __shared__ float sm_m[NW][G], sm_l[NW][G];
    __shared__ float sm_o[NW][G][D];
#pragma unroll
    for (int h = 0; h < G; ++h) {
        if (lane == 0) {
            sm_m[warp][h] = m[h];
            sm_l[warp][h] = l[h];
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) sm_o[warp][h][lane * 4 + i] = acc[h][i];
    }
    __syncthreads();
    for (int idx = threadIdx.x; idx < G * D; idx += NW * 32) {
        const int h = idx / D, d = idx % D;
        float M = -INFINITY;
        for (int w = 0; w < NW; ++w) M = fmaxf(M, sm_m[w][h]);
        float num = 0.f, den = 0.f;
        for (int w = 0; w < NW; ++w) {
            const float wt = (sm_m[w][h] == -INFINITY) ? 0.f : exp2f((sm_m[w][h] - M) * sl2);
            num += wt * sm_o[w][h][d];
            den += wt * sm_l[w][h];
        }
        const float r = den > 0.f ? vs * num / den : 0.f;
        out[b * out_stride + (int64_t)(head0 + h) * D + d] = __float2bfloat16(r);
    }
}

template <bool FP8>
int launch_simt(int G, dim3 grid, cudaStream_t st, __nv_bfloat16* out, const __nv_bfloat16* q, const void* kc,
                const void* vc, const int32_t* bt, const int32_t* sl, int bs, int max_nb, int64_t q_stride,
                int64_t out_stride, int64_t block_stride, int64_t head_stride, float scale_log2,
                const float* k_scale, const float* v_scale) {
    // heads are processed in groups of up to 8 (register budget); each group re-reads the KV rows
    for (int g0 = 0; g0 < G;) {
        const int rem = G - g0;
#define B2_SIMT_LAUNCH(N)                                                                                   \
    paged_decode_simt_kernel<FP8, N><<<grid, NW * 32, 0, st>>>(out, q, kc, vc, bt, sl, bs, max_nb, g0, G,    \
                                                              q_stride, out_stride, block_stride,          \
                                                              head_stride, scale_log2, k_scale, v_scale);  \
    g0 += N
        if (rem >= 8) { B2_SIMT_LAUNCH(8); }
        else if (rem >= 4) { B2_SIMT_LAUNCH(4); }
        else if (rem >= 2) { B2_SIMT_LAUNCH(2); }
        else { B2_SIMT_LAUNCH(1); }
#undef B2_SIMT_LAUNCH
        B2_CUDA_OK(cudaGetLastError());
    }
    return B2_OK;
}

}  // namespace

extern "C" int b2_paged_decode_simt(void* out, const void* q, const void* k_cache, const void* v_cache,
                                    const int32_t* block_tables, const int32_t* seq_lens, int B, int Hq,
                                    int Hkv, int d, int bs, int max_nb, int64_t q_stride, int64_t out_stride,
                                    int64_t cache_block_stride, int64_t cache_head_stride, float scale,
                                    int kv_dtype, const float* k_scale, const float* v_scale, void* stream) {
    B2_REQUIRE(B >= 0 && Hq > 0 && Hkv > 0 && bs > 0 && max_nb > 0, B2_EINVAL, "paged_decode_simt: bad sizes");
    if (B == 0) return B2_OK;
    B2_REQUIRE(out && q && k_cache && v_cache && block_tables && seq_lens, B2_EINVAL, "paged_decode_simt: null pointer");
    B2_REQUIRE(d == D, B2_EUNSUPPORTED, "paged_decode_simt: head_dim %d unsupported", d);
    B2_REQUIRE(Hq % Hkv == 0, B2_EINVAL, "paged_decode_simt: Hq %% Hkv != 0");
    B2_REQUIRE(kv_dtype == B2_KV_BF16 || kv_dtype == B2_KV_FP8_E4M3, B2_EUNSUPPORTED, "paged_decode_simt: kv_dtype");
    B2_REQUIRE(Hkv <= 65535 && B <= 65535, B2_EUNSUPPORTED, "paged_decode_simt: grid too large");
    if (kv_dtype == B2_KV_FP8_E4M3) B2_REQUIRE(k_scale && v_scale, B2_EINVAL, "paged_decode_simt: fp8 needs scales");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    dim3 grid(Hkv, B);
    const float sl2 = scale * LOG2E;
    if (kv_dtype == B2_KV_BF16)
        return launch_simt<false>(Hq / Hkv, grid, st, static_cast<__nv_bfloat16*>(out), static_cast<const __nv_bfloat16*>(q),
                                  k_cache, v_cache, block_tables, seq_lens, bs, max_nb, q_stride, out_stride,
                                  cache_block_stride, cache_head_stride, sl2, k_scale, v_scale);
    return launch_simt<true>(Hq / Hkv, grid, st, static_cast<__nv_bfloat16*>(out), static_cast<const __nv_bfloat16*>(q),
                             k_cache, v_cache, block_tables, seq_lens, bs, max_nb, q_stride, out_stride,
                             cache_block_stride, cache_head_stride, sl2, k_scale, v_scale);
}
