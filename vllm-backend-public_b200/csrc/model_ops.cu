// model_ops.cu -- the HBM-bound element-wise ops either side of the attention path in one decoder layer
// (the random-init Llama forward that calls the hot path every step).
//
//   b2_rms_norm        (optional fused residual add)           vLLM: rms_norm / fused_add_rms_norm
//   b2_rope_and_cache  neox rotary on q and k in place + KV write through slot_mapping in the same pass
//                                                             vLLM: rotary_embedding + reshape_and_cache
//   b2_silu_and_mul    SwiGLU gate                             vLLM: silu_and_mul
//   b2_sample          greedy / temperature (Gumbel-max) token choice from bf16 logits
//
// The engine the reference launches (reference benchmark-vllm/vllm_manager.py:371-374) runs these between the
// cuBLAS GEMMs; activation dtype is VLLMConfig.dtype (reference benchmark-vllm/models.py:14) = bf16.
// All are byte movers: 16-byte vector loads/stores, fp32 math, one rounding to bf16 on the way out.
#include "b2_common.cuh"

#include <stdlib.h>

namespace {

__device__ __forceinline__ float bf_lo(uint32_t w) { return __uint_as_float(w << 16); }
__device__ __forceinline__ float bf_hi(uint32_t w) { return __uint_as_float(w & 0xffff0000u); }

__device__ __forceinline__ void unpack8(const int4& v, float (&f)[8]) {
    const uint32_t w[4] = {(uint32_t)v.x, (uint32_t)v.y, (uint32_t)v.z, (uint32_t)v.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        f[2 * i] = bf_lo(w[i]);
        f[2 * i + 1] = bf_hi(w[i]);
    }
}
__device__ __forceinline__ int4 pack8(const float (&f)[8]) {
    int4 r;
    r.x = (int)b2::pack_bf16x2(f[0], f[1]);
    r.y = (int)b2::pack_bf16x2(f[2], f[3]);
    r.z = (int)b2::pack_bf16x2(f[4], f[5]);
    r.w = (int)b2::pack_bf16x2(f[6], f[7]);
    return r;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// ---------------------------------------------------------------------------------------------
// RMSNorm: one CTA per row.  residual != nullptr: h = bf16(x + residual) is written back to residual
// and normalised; out = bf16(w * h * rsqrt(mean(h^2) + eps)).  VPT vectors of 8 per thread stay in registers.
// ---------------------------------------------------------------------------------------------
template <int VPT>
__global__ void __launch_bounds__(256)
rms_norm_kernel(__nv_bfloat16* out, const __nv_bfloat16* x,   // out may alias x (in-place norm): no __restrict__ on these two
                __nv_bfloat16* __restrict__ residual, const __nv_bfloat16* __restrict__ weight,
                int H, int64_t x_stride, int64_t out_stride, int64_t res_stride, float eps) {
    const int row = blockIdx.x;
    const int nvec = H >> 3;
    const __nv_bfloat16* xr = x + row * x_stride;
    __nv_bfloat16* rr = residual ? residual + row * res_stride : nullptr;
    float h[VPT][8];
    float ss = 0.f;
    // every load of the row (x, residual, weight) is issued before the first use: one DRAM round trip per CTA
    int4 xv[VPT], rv[VPT], wv[VPT];
#pragma unroll
    for (int k = 0; k < VPT; ++k) {
        const int v = threadIdx.x + k * blockDim.x;
        xv[k] = rv[k] = wv[k] = make_int4(0, 0, 0, 0);
        if (v < nvec) {
            xv[k] = *reinterpret_cast<const int4*>(xr + v * 8);
            if (rr) rv[k] = *reinterpret_cast<const int4*>(rr + v * 8);
            wv[k] = *reinterpret_cast<const int4*>(weight + v * 8);
        }
    }
#pragma unroll
    for (int k = 0; k < VPT; ++k) {
        const int v = threadIdx.x + k * blockDim.x;
        if (v < nvec) {
            unpack8(xv[k], h[k]);
            if (rr) {
                float r[8];
                unpack8(rv[k], r);
#pragma unroll
                for (int i = 0; i < 8; ++i) h[k][i] += r[i];
                const int4 packed = pack8(h[k]);          // the residual stream is bf16: round once, then norm
                *reinterpret_cast<int4*>(rr + v * 8) = packed;
                unpack8(packed, h[k]);
            }
#pragma unroll
            for (int i = 0; i < 8; ++i) ss += h[k][i] * h[k][i];
        }
    }
    __shared__ float red[8];
    ss = warp_sum(ss);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
    __syncthreads();
    float tot = 0.f;
    const int nw = blockDim.x >> 5;
    for (int i = 0; i < nw; ++i) tot += red[i];
    const float inv = rsqrtf(tot / (float)H + eps);
    __nv_bfloat16* orow = out + row * out_stride;
#pragma unroll
    for (int k = 0; k < VPT; ++k) {
        const int v = threadIdx.x + k * blockDim.x;
        if (v < nvec) {
            float w[8], o[8];
            unpack8(wv[k], w);
#pragma unroll
            for (int i = 0; i < 8; ++i) o[i] = w[i] * (h[k][i] * inv);
            *reinterpret_cast<int4*>(orow + v * 8) = pack8(o);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// SwiGLU: out[t, i] = silu(gate[t, i]) * up[t, i], gate_up = [T, 2I] (gate | up)
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
silu_and_mul_kernel(__nv_bfloat16* __restrict__ out, const __nv_bfloat16* __restrict__ gate_up, int I,
                    int64_t in_stride, int64_t out_stride) {
    const int t = blockIdx.y;
    const int v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= (I >> 3)) return;
    const __nv_bfloat16* g = gate_up + t * in_stride + v * 8;
    float a[8], b[8], o[8];
    unpack8(b2::ld_nc_16(g), a);
    unpack8(b2::ld_nc_16(g + I), b);
#pragma unroll
    for (int i = 0; i < 8; ++i) o[i] = a[i] / (1.f + __expf(-a[i])) * b[i];
    *reinterpret_cast<int4*>(out + t * out_stride + v * 8) = pack8(o);
}

// ---------------------------------------------------------------------------------------------
// rotary (neox pairing i <-> i + d/2) on q and k in place, then K (rotated) and V into the paged cache.
// One CTA per token.  Work items: (Hq + Hkv) heads x (d/16) vector pairs, then Hkv x (d/8) V vectors.
// cos_sin: fp32 [max_pos, d] = cos[0..d/2) | sin[0..d/2) per position.
// ---------------------------------------------------------------------------------------------
template <bool FP8>
__device__ __forceinline__ void store_cache8(void* cache, int64_t elem, const float (&f)[8], float scale) {
    if (!FP8) {
        *reinterpret_cast<int4*>(static_cast<__nv_bfloat16*>(cache) + elem) = pack8(f);
    } else {
        // quantise what the bf16 tensor holds: round to bf16 first (same bytes b2_reshape_and_cache would read)
        const int4 p = pack8(f);
        float g[8];
        unpack8(p, g);
        uint32_t w[2];
#pragma unroll
        for (int q = 0; q < 2; ++q) {
            uint32_t acc = 0;
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                const __nv_fp8x2_storage_t pr = __nv_cvt_float2_to_fp8x2(
                    make_float2(__fdiv_rn(g[q * 4 + j * 2], scale), __fdiv_rn(g[q * 4 + j * 2 + 1], scale)),
                    __NV_SATFINITE, __NV_E4M3);
                acc |= static_cast<uint32_t>(pr) << (16 * j);
            }
            w[q] = acc;
        }
        *reinterpret_cast<uint2*>(static_cast<uint8_t*>(cache) + elem) = make_uint2(w[0], w[1]);
    }
}

template <bool FP8>
__global__ void __launch_bounds__(256, 5)      // <= 48 registers: five CTAs (40 warps) per SM keep the loads in flight
rope_and_cache_kernel(__nv_bfloat16* __restrict__ qkv, const int32_t* __restrict__ positions,
                      const float* __restrict__ cos_sin, int max_pos, void* __restrict__ k_cache,
                      void* __restrict__ v_cache, const int64_t* __restrict__ slot_mapping, int Hq, int Hkv, int d, int bs,
                      int64_t token_stride, int64_t cache_block_stride, int64_t cache_head_stride,
                      const float* __restrict__ k_scale, const float* __restrict__ v_scale) {
    // the attention kernel that follows may be launched now (programmatic dependent launch): it waits
    // (griddepcontrol.wait) before it reads q, the cache or the block tables
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    const int t = blockIdx.x;
    const int half = d >> 1;
    const int pairs_per_head = half >> 3;              // vector pairs of 8 elements
    const int vec_per_head = d >> 3;
    const int n_rot = (Hq + Hkv) * pairs_per_head;
    const int n_v = Hkv * vec_per_head;
    __nv_bfloat16* row = qkv + t * token_stride;

    // A thread owns up to ITEMS work items (i = tid + k * blockDim): a rotation pair (two 16-byte vectors of one head)
    // or one 16-byte vector of V.  All of a thread's q/k/v loads are issued before anything depends on them; the
    // position / slot / cos-sin chain runs underneath.  (One item per loop trip left a CTA with one DRAM round trip
    // in flight per thread: 3.4 TB/s at 8k tokens.)
    constexpr int ITEMS = 2;
    int4 xa[ITEMS], xb[ITEMS];
#pragma unroll
    for (int k = 0; k < ITEMS; ++k) {
        const int i = threadIdx.x + k * blockDim.x;
        xa[k] = xb[k] = make_int4(0, 0, 0, 0);
        if (i < n_rot) {
            const int h = i / pairs_per_head, c = i - h * pairs_per_head;
            xa[k] = *reinterpret_cast<const int4*>(row + h * d + c * 8);
            xb[k] = *reinterpret_cast<const int4*>(row + h * d + c * 8 + half);
        } else if (i < n_rot + n_v) {
            const int j = i - n_rot;
            xa[k] = *reinterpret_cast<const int4*>(row + (int64_t)(Hq + Hkv) * d + j * 8);
        }
    }
    const int pos = min(max(positions[t], 0), max_pos - 1);   // table bound; the host validates positions
    const int64_t slot = slot_mapping[t];
    int64_t blk = 0;
    int off = 0;
    if (slot >= 0) {
        if (slot < (int64_t)0x7fffffff) {                     // 32-bit divide (the 64-bit one is a subroutine call)
            const uint32_t s32 = (uint32_t)slot;
            blk = s32 / (uint32_t)bs;
            off = (int)(s32 - (uint32_t)blk * (uint32_t)bs);
        } else {
            blk = slot / bs;
            off = (int)(slot - blk * bs);
        }
    }
    const float* cs = cos_sin + (int64_t)pos * d;
    float ks = 1.f, vs = 1.f;
    if (FP8) { ks = *k_scale; vs = *v_scale; }

    for (int base = 0; base < n_rot + n_v; base += ITEMS * blockDim.x) {
        if (base > 0) {                                       // more than ITEMS items per thread (wide models)
#pragma unroll
            for (int k = 0; k < ITEMS; ++k) {
                const int i = base + threadIdx.x + k * blockDim.x;
                if (i < n_rot) {
                    const int h = i / pairs_per_head, c = i - h * pairs_per_head;
                    xa[k] = *reinterpret_cast<const int4*>(row + h * d + c * 8);
                    xb[k] = *reinterpret_cast<const int4*>(row + h * d + c * 8 + half);
                } else if (i < n_rot + n_v) {
                    xa[k] = *reinterpret_cast<const int4*>(row + (int64_t)(Hq + Hkv) * d + (i - n_rot) * 8);
                }
            }
        }
#pragma unroll
        for (int k = 0; k < ITEMS; ++k) {
            const int i = base + threadIdx.x + k * blockDim.x;
            if (i < n_rot) {
                const int h = i / pairs_per_head;
                const int c = i - h * pairs_per_head;
                __nv_bfloat16* p = row + h * d + c * 8;
                float x1[8], x2[8], co[8], si[8], o1[8], o2[8];
                unpack8(xa[k], x1);
                unpack8(xb[k], x2);
                *reinterpret_cast<float4*>(co) = *reinterpret_cast<const float4*>(cs + c * 8);
                *reinterpret_cast<float4*>(co + 4) = *reinterpret_cast<const float4*>(cs + c * 8 + 4);
                *reinterpret_cast<float4*>(si) = *reinterpret_cast<const float4*>(cs + half + c * 8);
                *reinterpret_cast<float4*>(si + 4) = *reinterpret_cast<const float4*>(cs + half + c * 8 + 4);
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    o1[j] = __fsub_rn(__fmul_rn(x1[j], co[j]), __fmul_rn(x2[j], si[j]));
                    o2[j] = __fadd_rn(__fmul_rn(x2[j], co[j]), __fmul_rn(x1[j], si[j]));
                }
                *reinterpret_cast<int4*>(p) = pack8(o1);
                *reinterpret_cast<int4*>(p + half) = pack8(o2);
                if (h >= Hq && slot >= 0) {
                    const int64_t e = blk * cache_block_stride + (int64_t)(h - Hq) * cache_head_stride + (int64_t)off * d + c * 8;
                    store_cache8<FP8>(k_cache, e, o1, ks);
                    store_cache8<FP8>(k_cache, e + half, o2, ks);
                }
            } else if (i < n_rot + n_v && slot >= 0) {
                const int j = i - n_rot;
                const int h = j / vec_per_head;
                const int c = j - h * vec_per_head;
                float f[8];
                unpack8(xa[k], f);
                const int64_t e = blk * cache_block_stride + (int64_t)h * cache_head_stride + (int64_t)off * d + c * 8;
                store_cache8<FP8>(v_cache, e, f, vs);
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------
// Streaming form of the same op (the default when there are tokens enough to pipeline): a persistent grid; each CTA
// walks tokens t, t + grid, ... through a ring of RING stages in shared memory.  A producer warp fetches the whole
// q|k|v row of a token (12 KiB at Llama-3-8B widths) and its cos|sin row with two bulk copies (cp.async.bulk, completion
// counted on an mbarrier) up to RING tokens ahead; the 8 consumer warps rotate out of shared memory and store.
// Why: with one token per CTA (the kernel above) a CTA has its 12 KiB of loads in flight for only the first third of
// its life, and ncu shows the loads taking ~3 us under the mixed read/write traffic -- 48 KiB in flight per SM is
// 2.5 TB/s of reads.  The ring keeps up to 4 CTAs x RING x 12 KiB in flight per SM without a register per byte.
// write_k = 0 skips the in-place store of the rotated K rows (an engine that reads K only from the cache does not need
// them: 2 of 26 KiB per token).
// ---------------------------------------------------------------------------------------------
constexpr int ROPE_RING = 4;
constexpr int ROPE_CONSUMERS = 256;
constexpr int ROPE_MAX_ITEMS = 3;        // work items per consumer thread: rows of up to 80 heads of 128

template <bool FP8, int MAXI>          // MAXI work items per consumer thread: 2 covers rows of up to 512 items (Llama-3-8B: 448)
__global__ void __launch_bounds__(ROPE_CONSUMERS + 32, 4)     // <= 56 registers: four CTAs (36 warps, 16 stages) per SM
rope_and_cache_ring_kernel(__nv_bfloat16* __restrict__ qkv, const int32_t* __restrict__ positions,
                           const float* __restrict__ cos_sin, int max_pos, void* __restrict__ k_cache,
                           void* __restrict__ v_cache, const int64_t* __restrict__ slot_mapping, int T, int Hq, int Hkv,
                           int d, int bs, int64_t token_stride, int64_t cache_block_stride, int64_t cache_head_stride,
                           const float* __restrict__ k_scale, const float* __restrict__ v_scale, int write_k,
                           int stage_bytes) {
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ __align__(8) uint64_t bars[2 * ROPE_RING];          // full[RING] | empty[RING]
    const int half = d >> 1;
    const int pairs_per_head = half >> 3;
    const int vec_per_head = d >> 3;
    const int n_rot = (Hq + Hkv) * pairs_per_head;
    const int n_items = n_rot + Hkv * vec_per_head;
    const uint32_t row_bytes = (uint32_t)(Hq + 2 * Hkv) * d * sizeof(__nv_bfloat16);
    const uint32_t cs_bytes = (uint32_t)d * sizeof(float);
    // stage layout: row | cos|sin row | slot (int64)
    const uint32_t full0 = b2::smem_u32(&bars[0]), empty0 = b2::smem_u32(&bars[ROPE_RING]);
    if (threadIdx.x == 0) {
        for (int s = 0; s < ROPE_RING; ++s) {
            b2::mbar_init(full0 + 8 * s, 1);
            b2::mbar_init(empty0 + 8 * s, ROPE_CONSUMERS / 32);
        }
        b2::fence_mbar_init();
    }
    __syncthreads();
    const int first = blockIdx.x, step = gridDim.x;
    if (first >= T) return;

    if (threadIdx.x >= ROPE_CONSUMERS) {
        // ---------------- producer warp ----------------
        // The 32 lanes fetch position and slot of the next 32 tokens of this CTA at once (one exposed load latency per
        // 32 tokens, and the batch after is already in flight); lane 0 issues the bulk copies.
        const int lane = threadIdx.x & 31;
        auto fetch = [&](int j0, int& pos_l, int64_t& slot_l) {
            const int64_t tl = (int64_t)first + (int64_t)(j0 + lane) * step;
            pos_l = 0;
            slot_l = -1;
            if (tl < T) { pos_l = positions[tl]; slot_l = slot_mapping[tl]; }
        };
        int pos_l, npos_l;
        int64_t slot_l, nslot_l;
        fetch(0, pos_l, slot_l);
        for (int j0 = 0; (int64_t)first + (int64_t)j0 * step < T; j0 += 32) {
            fetch(j0 + 32, npos_l, nslot_l);
            for (int jj = 0; jj < 32; ++jj) {
                const int j = j0 + jj;
                const int64_t t = (int64_t)first + (int64_t)j * step;
                if (t >= T) break;
                const int pos = __shfl_sync(0xffffffffu, pos_l, jj);
                const int64_t slot = __shfl_sync(0xffffffffu, slot_l, jj);
                if (lane == 0) {
                    const int s = j % ROPE_RING;
                    if (j >= ROPE_RING) b2::mbar_wait(empty0 + 8 * s, ((j / ROPE_RING) - 1) & 1);
                    unsigned char* st = smem + (size_t)s * stage_bytes;
                    *reinterpret_cast<int64_t*>(st + row_bytes + cs_bytes) = slot;
                    const int p = min(max(pos, 0), max_pos - 1);
                    b2::mbar_arrive_expect_tx(full0 + 8 * s, row_bytes + cs_bytes);  // release: the slot store is visible
                    b2::bulk_load(b2::smem_u32(st), qkv + t * token_stride, row_bytes, full0 + 8 * s);
                    b2::bulk_load(b2::smem_u32(st + row_bytes), cos_sin + (int64_t)p * d, cs_bytes, full0 + 8 * s);
                }
                __syncwarp();
            }
            pos_l = npos_l;
            slot_l = nslot_l;
        }
        return;
    }
    // ---------------- consumers ----------------
    // What a thread's (up to MAXI) work items are does not depend on the token: decode them once -- the kernel is
    // issue-bound (ncu: 2.7 warp instructions per cycle per SM), and the per-token integer divides were a fifth of it.
    float ks = 1.f, vs = 1.f;
    if (FP8) { ks = *k_scale; vs = *v_scale; }
    int so[MAXI], kind[MAXI], col[MAXI];      // row offset (elements); 0 none 1 q 2 k 3 v; c * 8
    int64_t coff[MAXI];                                           // offset inside a cache block
#pragma unroll
    for (int k = 0; k < MAXI; ++k) {
        const int i = threadIdx.x + k * ROPE_CONSUMERS;
        so[k] = 0; kind[k] = 0; col[k] = (threadIdx.x % pairs_per_head) * 8; coff[k] = 0;
        if (i < n_rot) {
            const int h = i / pairs_per_head, c = i - h * pairs_per_head;
            so[k] = h * d + c * 8;
            col[k] = c * 8;
            kind[k] = h < Hq ? 1 : 2;
            coff[k] = (int64_t)(h - Hq) * cache_head_stride + c * 8;
        } else if (i < n_items) {
            const int jv = i - n_rot;
            const int h = jv / vec_per_head, c = jv - h * vec_per_head;
            so[k] = (Hq + Hkv) * d + jv * 8;
            kind[k] = 3;
            coff[k] = (int64_t)h * cache_head_stride + c * 8;
        }
    }
    const int bs_shift = __ffs(bs) - 1;                                    // bs is a power of two (host checks)
    int t = first;
    for (int j = 0; t < T; ++j, t += step) {
        const int s = j % ROPE_RING;
        b2::mbar_wait(full0 + 8 * s, (j / ROPE_RING) & 1);
        const unsigned char* st = smem + (size_t)s * stage_bytes;
        const __nv_bfloat16* srow = reinterpret_cast<const __nv_bfloat16*>(st);
        const float* cs = reinterpret_cast<const float*>(st + row_bytes);
        const int64_t slot = *reinterpret_cast<const int64_t*>(st + row_bytes + cs_bytes);
        const int64_t tok_off = slot >= 0 ? (slot >> bs_shift) * cache_block_stride + (slot & (bs - 1)) * (int64_t)d : 0;
        __nv_bfloat16* row = qkv + t * token_stride;
        // every rotation item of a thread has the same column (the pairs per head divide the consumer count: host
        // check), so the token's cos / sin vectors are read from shared memory once
        float co[8], si[8];
        *reinterpret_cast<float4*>(co) = *reinterpret_cast<const float4*>(cs + col[0]);
        *reinterpret_cast<float4*>(co + 4) = *reinterpret_cast<const float4*>(cs + col[0] + 4);
        *reinterpret_cast<float4*>(si) = *reinterpret_cast<const float4*>(cs + half + col[0]);
        *reinterpret_cast<float4*>(si + 4) = *reinterpret_cast<const float4*>(cs + half + col[0] + 4);
#pragma unroll
        for (int k = 0; k < MAXI; ++k) {
            if (kind[k] == 1 || kind[k] == 2) {
                float x1[8], x2[8], o1[8], o2[8];
                unpack8(*reinterpret_cast<const int4*>(srow + so[k]), x1);
                unpack8(*reinterpret_cast<const int4*>(srow + so[k] + half), x2);
#pragma unroll
                for (int q = 0; q < 8; ++q) {
                    o1[q] = __fsub_rn(__fmul_rn(x1[q], co[q]), __fmul_rn(x2[q], si[q]));
                    o2[q] = __fadd_rn(__fmul_rn(x2[q], co[q]), __fmul_rn(x1[q], si[q]));
                }
                if (kind[k] == 1 || write_k) {
                    *reinterpret_cast<int4*>(row + so[k]) = pack8(o1);
                    *reinterpret_cast<int4*>(row + so[k] + half) = pack8(o2);
                }
                if (kind[k] == 2 && slot >= 0) {
                    store_cache8<FP8>(k_cache, tok_off + coff[k], o1, ks);
                    store_cache8<FP8>(k_cache, tok_off + coff[k] + half, o2, ks);
                }
            } else if (kind[k] == 3 && slot >= 0) {
                float f[8];
                unpack8(*reinterpret_cast<const int4*>(srow + so[k]), f);
                store_cache8<FP8>(v_cache, tok_off + coff[k], f, vs);
            }
        }
        __syncwarp();
        if ((threadIdx.x & 31) == 0) b2::mbar_arrive(empty0 + 8 * s);           // this warp is done with the stage
    }
}

// ---------------------------------------------------------------------------------------------
// token choice: one CTA per row of bf16 logits [B, V].
//   temperature[b] <= 0: argmax (lowest index among equal maxima)
//   else: argmax of logit / T + g, g = -log(-log(u)), u from a counter hash of (seed[b], index)   (Gumbel-max)
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t hash_u32(uint64_t seed, uint32_t idx) {
    uint64_t z = seed + 0x9E3779B97F4A7C15ull * (uint64_t)(idx + 1);    // splitmix64 finaliser
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    z = z ^ (z >> 31);
    return (uint32_t)(z >> 40);                                          // 24 bits
}

// order-preserving 16-bit key of a bf16 bit pattern: key(a) < key(b)  <=>  a < b  (-0 < +0; NaN sorts outside)
__device__ __forceinline__ uint32_t bf16_key(uint32_t u) { return (u & 0x8000u) ? (~u & 0xffffu) : (u | 0x8000u); }

// argmax over the row of (logit / T + Gumbel noise) [T > 0] or of the logit [T <= 0], restricted to keys >= min_key
// (0 = every token); lowest index among equal maxima.  Every thread returns; thread 0 holds the result.
__device__ __forceinline__ int sample_row(const __nv_bfloat16* __restrict__ row, int V, float T, uint64_t seed,
                                          uint32_t min_key) {
    float best = -INFINITY;
    int best_i = 0x7fffffff;
    const float invT = T > 0.f ? 1.f / T : 1.f;
    for (int v = threadIdx.x; v < (V >> 3); v += blockDim.x) {
        const int4 raw = b2::ld_nc_16(row + v * 8);
        float f[8];
        unpack8(raw, f);
        const uint32_t w[4] = {(uint32_t)raw.x, (uint32_t)raw.y, (uint32_t)raw.z, (uint32_t)raw.w};
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int idx = v * 8 + j;
            if (min_key && bf16_key((j & 1) ? (w[j >> 1] >> 16) : (w[j >> 1] & 0xffffu)) < min_key) continue;
            float s = f[j];
            if (T > 0.f) {
                const float u = ((float)hash_u32(seed, (uint32_t)idx) + 0.5f) * (1.f / 16777216.f);
                s = s * invT - __logf(-__logf(u));
            }
            if (s > best || (s == best && idx < best_i)) { best = s; best_i = idx; }
        }
    }
    for (int idx = (V & ~7) + threadIdx.x; idx < V; idx += blockDim.x) {     // tail (V not a multiple of 8)
        if (min_key && bf16_key((uint32_t)__bfloat16_as_ushort(row[idx])) < min_key) continue;
        float s = __bfloat162float(row[idx]);
        if (T > 0.f) {
            const float u = ((float)hash_u32(seed, (uint32_t)idx) + 0.5f) * (1.f / 16777216.f);
            s = s * invT - __logf(-__logf(u));
        }
        if (s > best || (s == best && idx < best_i)) { best = s; best_i = idx; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float ob = __shfl_xor_sync(0xffffffffu, best, o);
        const int oi = __shfl_xor_sync(0xffffffffu, best_i, o);
        if (ob > best || (ob == best && oi < best_i)) { best = ob; best_i = oi; }
    }
    __shared__ float sb[32];
    __shared__ int si[32];
    __syncthreads();                                   // sb / si may still be read by an earlier call
    if ((threadIdx.x & 31) == 0) { sb[threadIdx.x >> 5] = best; si[threadIdx.x >> 5] = best_i; }
    __syncthreads();
    if (threadIdx.x < 32) {
        const int nw = blockDim.x >> 5;
        best = threadIdx.x < nw ? sb[threadIdx.x] : -INFINITY;
        best_i = threadIdx.x < nw ? si[threadIdx.x] : 0x7fffffff;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ob = __shfl_xor_sync(0xffffffffu, best, o);
            const int oi = __shfl_xor_sync(0xffffffffu, best_i, o);
            if (ob > best || (ob == best && oi < best_i)) { best = ob; best_i = oi; }
        }
    }
    return best_i == 0x7fffffff ? 0 : best_i;
}

// block-wide sums of (count, mass) over the tokens with key in [lo_key, 65535] and key >= t_key;
// mass_i = exp2((logit_i - m) * c).  Fixed reduction order: the same bits on every launch.
__device__ __forceinline__ void measure_ge(const __nv_bfloat16* __restrict__ row, int V, uint32_t t_key, float m, float c,
                                           bool want_mass, int& count, float& mass) {
    int cnt = 0;
    float ms = 0.f;
    for (int v = threadIdx.x; v < (V >> 3); v += blockDim.x) {
        const int4 raw = b2::ld_nc_16(row + v * 8);
        const uint32_t w[4] = {(uint32_t)raw.x, (uint32_t)raw.y, (uint32_t)raw.z, (uint32_t)raw.w};
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const uint32_t u = (j & 1) ? (w[j >> 1] >> 16) : (w[j >> 1] & 0xffffu);
            if (bf16_key(u) >= t_key && (u & 0x7f80u) != 0x7f80u) {            // not inf / NaN
                ++cnt;
                if (want_mass) ms += b2::fast_exp2((__uint_as_float(u << 16) - m) * c);
            }
        }
    }
    for (int idx = (V & ~7) + threadIdx.x; idx < V; idx += blockDim.x) {
        const uint32_t u = __bfloat16_as_ushort(row[idx]);
        if (bf16_key(u) >= t_key && (u & 0x7f80u) != 0x7f80u) {
            ++cnt;
            if (want_mass) ms += b2::fast_exp2((__uint_as_float(u << 16) - m) * c);
        }
    }
    __shared__ int rc[32];
    __shared__ float rm[32];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
        ms += __shfl_xor_sync(0xffffffffu, ms, o);
    }
    __syncthreads();
    if ((threadIdx.x & 31) == 0) { rc[threadIdx.x >> 5] = cnt; rm[threadIdx.x >> 5] = ms; }
    __syncthreads();
    cnt = 0;
    ms = 0.f;
    const int nw = blockDim.x >> 5;
    for (int i = 0; i < nw; ++i) { cnt += rc[i]; ms += rm[i]; }
    count = cnt;
    mass = ms;
}

__global__ void __launch_bounds__(1024)
sample_kernel(int64_t* __restrict__ tokens, const __nv_bfloat16* __restrict__ logits, int V, int64_t stride,
              const float* __restrict__ temperature, const int64_t* __restrict__ seeds,
              const int32_t* __restrict__ top_k, const float* __restrict__ top_p) {
    const int b = blockIdx.x;
    const float T = temperature ? temperature[b] : 0.f;
    const uint64_t seed = seeds ? (uint64_t)seeds[b] : 0ull;
    const __nv_bfloat16* row = logits + b * stride;
    const int k = top_k ? top_k[b] : 0;
    const float p = top_p ? top_p[b] : 1.f;
    uint32_t min_key = 0;
    if (T > 0.f && ((k > 0 && k < V) || p < 1.f)) {
        // the candidate set is {key >= t}: t = the k-th largest logit (ties kept), then raised to the smallest set of
        // top logits whose softmax mass reaches p of the top-k mass (ties kept).  Both by bisection over the 16-bit key
        // space: 16 passes over the row (L2 resident) each.
        uint32_t lo = 0, hi = 65536;
        int cnt;
        float mass;
        if (k > 0 && k < V) {
            while (hi - lo > 1) {
                const uint32_t mid = (lo + hi) >> 1;
                measure_ge(row, V, mid, 0.f, 0.f, false, cnt, mass);
                if (cnt >= k) lo = mid; else hi = mid;
            }
        }
        min_key = lo;
        if (p < 1.f) {
            // row maximum (finite values) for a stable exponent
            float mx = -INFINITY;
            for (int idx = threadIdx.x; idx < V; idx += blockDim.x) {
                const float x = __bfloat162float(row[idx]);
                if (x > mx && x < INFINITY) mx = x;
            }
            __shared__ float smx[32];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
            if ((threadIdx.x & 31) == 0) smx[threadIdx.x >> 5] = mx;
            __syncthreads();
            mx = smx[0];
            for (int i = 1; i < (int)(blockDim.x >> 5); ++i) mx = fmaxf(mx, smx[i]);
            const float c = 1.4426950408889634f / T;
            measure_ge(row, V, min_key, mx, c, true, cnt, mass);
            const float target = p * mass;
            lo = min_key;
            hi = 65536;
            while (hi - lo > 1) {
                const uint32_t mid = (lo + hi) >> 1;
                measure_ge(row, V, mid, mx, c, true, cnt, mass);
                if (mass >= target && cnt > 0) lo = mid; else hi = mid;
            }
            min_key = lo;
        }
    }
    const int tok = sample_row(row, V, T, seed, min_key);
    if (threadIdx.x == 0) tokens[b] = tok;
}

}  // namespace

// =============================================================================================
// C ABI
// =============================================================================================
extern "C" int b2_rms_norm(void* out, const void* x, void* residual, const void* weight, int T, int H,
                           int64_t x_stride, int64_t out_stride, int64_t res_stride, float eps, void* stream) {
    B2_REQUIRE(T >= 0 && H > 0, B2_EINVAL, "rms_norm: bad sizes T=%d H=%d", T, H);
    if (T == 0) return B2_OK;
    B2_REQUIRE(out && x && weight, B2_EINVAL, "rms_norm: null pointer");
    B2_REQUIRE(H % 8 == 0 && H <= 16384, B2_EUNSUPPORTED, "rms_norm: H=%d must be a multiple of 8 and <= 16384", H);
    B2_REQUIRE(b2_aligned16(out) && b2_aligned16(x) && b2_aligned16(weight) && b2_aligned16(residual), B2_EINVAL,
               "rms_norm: pointers must be 16-byte aligned");
    B2_REQUIRE(x_stride % 8 == 0 && out_stride % 8 == 0 && res_stride % 8 == 0, B2_EINVAL,
               "rms_norm: row strides must keep 16-byte alignment");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int nvec = H / 8;
    auto* o = static_cast<__nv_bfloat16*>(out);
    auto* xi = static_cast<const __nv_bfloat16*>(x);
    auto* r = static_cast<__nv_bfloat16*>(residual);
    auto* w = static_cast<const __nv_bfloat16*>(weight);
    if (nvec <= 256) {
        const int threads = ((nvec + 31) / 32) * 32;
        rms_norm_kernel<1><<<T, threads, 0, st>>>(o, xi, r, w, H, x_stride, out_stride, res_stride, eps);
    } else if (nvec <= 512) {
        rms_norm_kernel<2><<<T, 256, 0, st>>>(o, xi, r, w, H, x_stride, out_stride, res_stride, eps);
    } else if (nvec <= 1024) {
        rms_norm_kernel<4><<<T, 256, 0, st>>>(o, xi, r, w, H, x_stride, out_stride, res_stride, eps);
    } else {
        rms_norm_kernel<8><<<T, 256, 0, st>>>(o, xi, r, w, H, x_stride, out_stride, res_stride, eps);
    }
    B2_CUDA_OK(cudaGetLastError());
    return B2_OK;
}

extern "C" int b2_silu_and_mul(void* out, const void* gate_up, int T, int I, int64_t in_stride, int64_t out_stride,
                               void* stream) {
    B2_REQUIRE(T >= 0 && I > 0, B2_EINVAL, "silu_and_mul: bad sizes T=%d I=%d", T, I);
    if (T == 0) return B2_OK;
    B2_REQUIRE(out && gate_up, B2_EINVAL, "silu_and_mul: null pointer");
    B2_REQUIRE(I % 8 == 0, B2_EUNSUPPORTED, "silu_and_mul: I=%d must be a multiple of 8", I);
    B2_REQUIRE(b2_aligned16(out) && b2_aligned16(gate_up) && in_stride % 8 == 0 && out_stride % 8 == 0, B2_EINVAL,
               "silu_and_mul: pointers / strides must keep 16-byte alignment");
    B2_REQUIRE(T <= 65535 * 64, B2_EUNSUPPORTED, "silu_and_mul: T=%d too large", T);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int nvec = I / 8;
    auto* o = static_cast<__nv_bfloat16*>(out);
    auto* g = static_cast<const __nv_bfloat16*>(gate_up);
    for (int t0 = 0; t0 < T; t0 += 65535) {       // grid.y limit
        const int n = T - t0 < 65535 ? T - t0 : 65535;
        silu_and_mul_kernel<<<dim3((nvec + 255) / 256, n), 256, 0, st>>>(o + t0 * out_stride, g + t0 * in_stride, I,
                                                                        in_stride, out_stride);
    }
    B2_CUDA_OK(cudaGetLastError());
    return B2_OK;
}

extern "C" int b2_rope_and_cache(void* qkv, const int32_t* positions, const float* cos_sin, int max_pos,
                                 void* k_cache, void* v_cache, const int64_t* slot_mapping, int T, int Hq, int Hkv,
                                 int d, int bs, int64_t token_stride, int64_t cache_block_stride,
                                 int64_t cache_head_stride, int kv_dtype, const float* k_scale, const float* v_scale,
                                 void* stream) {
    return b2_rope_and_cache_ex(qkv, positions, cos_sin, max_pos, k_cache, v_cache, slot_mapping, T, Hq, Hkv, d, bs,
                                token_stride, cache_block_stride, cache_head_stride, kv_dtype, k_scale, v_scale, 1, stream);
}

extern "C" int b2_rope_and_cache_ex(void* qkv, const int32_t* positions, const float* cos_sin, int max_pos,
                                 void* k_cache, void* v_cache, const int64_t* slot_mapping, int T, int Hq, int Hkv,
                                 int d, int bs, int64_t token_stride, int64_t cache_block_stride,
                                 int64_t cache_head_stride, int kv_dtype, const float* k_scale, const float* v_scale,
                                 int write_k, void* stream) {
    B2_REQUIRE(T >= 0 && Hq > 0 && Hkv > 0 && d > 0 && bs > 0 && max_pos > 0, B2_EINVAL,
               "rope_and_cache: bad sizes T=%d Hq=%d Hkv=%d d=%d bs=%d", T, Hq, Hkv, d, bs);
    if (T == 0) return B2_OK;
    B2_REQUIRE(qkv && positions && cos_sin && k_cache && v_cache && slot_mapping, B2_EINVAL,
               "rope_and_cache: null pointer");
    B2_REQUIRE(kv_dtype == B2_KV_BF16 || kv_dtype == B2_KV_FP8_E4M3, B2_EUNSUPPORTED, "rope_and_cache: kv_dtype %d",
               kv_dtype);
    B2_REQUIRE(d % 16 == 0, B2_EUNSUPPORTED, "rope_and_cache: d=%d must be a multiple of 16", d);
    B2_REQUIRE(b2_aligned16(qkv) && b2_aligned16(cos_sin) && b2_aligned16(k_cache) && b2_aligned16(v_cache), B2_EINVAL,
               "rope_and_cache: pointers must be 16-byte aligned");
    B2_REQUIRE(token_stride % 8 == 0 && token_stride >= (int64_t)(Hq + 2 * Hkv) * d && cache_block_stride % 16 == 0 &&
                   cache_head_stride % 16 == 0 && cache_head_stride >= (int64_t)bs * d,
               B2_EINVAL, "rope_and_cache: bad strides");
    if (kv_dtype == B2_KV_FP8_E4M3) B2_REQUIRE(k_scale && v_scale, B2_EINVAL, "rope_and_cache: fp8 needs scales");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    auto* p = static_cast<__nv_bfloat16*>(qkv);
    // streaming kernel: a persistent grid, each CTA a ring of ROPE_RING token rows in shared memory fed by bulk copies;
    // taken when there are tokens enough to pipeline.  B2_ROPE_IMPL=token forces the one-token-per-CTA kernel (A/B)
    static const int sm_count = [] {
        int dev = 0, n = 148;
        if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
        return n;
    }();
    static const bool force_token = [] { const char* e = getenv("B2_ROPE_IMPL"); return e && e[0] == 't'; }();
    const size_t row_bytes = (size_t)(Hq + 2 * Hkv) * d * 2;
    const int stage_bytes = (int)((row_bytes + (size_t)d * 4 + 16 + 127) / 128 * 128);
    const size_t smem_bytes = (size_t)stage_bytes * ROPE_RING;
    int per_sm = (int)((220u * 1024u) / (smem_bytes + 1024));
    if (per_sm > 4) per_sm = 4;
    const int grid = sm_count * (per_sm > 0 ? per_sm : 1);
    const int n_items = (Hq + Hkv) * (d / 16) + Hkv * (d / 8);
    if (!force_token && per_sm >= 1 && smem_bytes <= 200u * 1024u && row_bytes % 16 == 0 && d % 16 == 0 &&
        (bs & (bs - 1)) == 0 && n_items <= ROPE_MAX_ITEMS * ROPE_CONSUMERS && ROPE_CONSUMERS % (d / 16) == 0 &&
        T >= 2 * grid) {
        const bool two = n_items <= 2 * ROPE_CONSUMERS;
        auto kern = kv_dtype == B2_KV_BF16
                        ? (two ? rope_and_cache_ring_kernel<false, 2> : rope_and_cache_ring_kernel<false, ROPE_MAX_ITEMS>)
                        : (two ? rope_and_cache_ring_kernel<true, 2> : rope_and_cache_ring_kernel<true, ROPE_MAX_ITEMS>);
        static size_t attr_set[4] = {0, 0, 0, 0};
        size_t& cur = attr_set[(kv_dtype == B2_KV_BF16 ? 0 : 1) + (two ? 0 : 2)];
        if (smem_bytes > cur) {
            B2_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes));
            cur = smem_bytes;
        }
        kern<<<grid, ROPE_CONSUMERS + 32, smem_bytes, st>>>(p, positions, cos_sin, max_pos, k_cache, v_cache, slot_mapping,
                                                            T, Hq, Hkv, d, bs, token_stride, cache_block_stride,
                                                            cache_head_stride, k_scale, v_scale, write_k, stage_bytes);
        B2_CUDA_OK(cudaGetLastError());
        return B2_OK;
    }
    if (kv_dtype == B2_KV_BF16)
        rope_and_cache_kernel<false><<<T, 256, 0, st>>>(p, positions, cos_sin, max_pos, k_cache, v_cache, slot_mapping, Hq, Hkv, d,
                                                        bs, token_stride, cache_block_stride, cache_head_stride,
                                                        nullptr, nullptr);
    else
        rope_and_cache_kernel<true><<<T, 256, 0, st>>>(p, positions, cos_sin, max_pos, k_cache, v_cache, slot_mapping, Hq, Hkv, d,
                                                       bs, token_stride, cache_block_stride, cache_head_stride, k_scale,
                                                       v_scale);
    B2_CUDA_OK(cudaGetLastError());
    return B2_OK;
}

extern "C" int b2_sample(int64_t* tokens, const void* logits, int B, int V, int64_t stride, const float* temperature,
                         const int64_t* seeds, void* stream) {
    return b2_sample_filtered(tokens, logits, B, V, stride, temperature, seeds, nullptr, nullptr, stream);
}

extern "C" int b2_sample_filtered(int64_t* tokens, const void* logits, int B, int V, int64_t stride,
                                  const float* temperature, const int64_t* seeds, const int32_t* top_k,
                                  const float* top_p, void* stream) {
    B2_REQUIRE(B >= 0 && V > 0, B2_EINVAL, "sample: bad sizes B=%d V=%d", B, V);
    if (B == 0) return B2_OK;
    B2_REQUIRE(tokens && logits, B2_EINVAL, "sample: null pointer");
    B2_REQUIRE(b2_aligned16(logits) && stride % 8 == 0, B2_EINVAL, "sample: logits rows must be 16-byte aligned");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    sample_kernel<<<B, 1024, 0, st>>>(tokens, static_cast<const __nv_bfloat16*>(logits), V, stride, temperature, seeds,
                                      top_k, top_p);
    B2_CUDA_OK(cudaGetLastError());
    return B2_OK;
}
