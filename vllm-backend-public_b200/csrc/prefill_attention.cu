// prefill_attention.cu -- causal attention of prompt chunks over the paged KV cache (sm_100a).
// the step on the other side of the cache (it produces, through
// b2_rope_and_cache / b2_reshape_and_cache, the KV the decode kernel reads).
//
// Contract: include/b200attn.h `b2_prefill_attention`.  Semantics follow vLLM's chunked-prefill /
// prefix attention (flash_attn_varlen_func with a block table): chunk c holds len_c new tokens whose
// positions in their sequence are ctx_start[c] .. ctx_start[c] + len_c - 1; token i attends the cache
// positions j <= ctx_start[c] + i of its own sequence.  K/V of the chunk itself are read from the cache
// too (they were written just before), so a chunk of a long prompt and a whole short prompt are the
// same case.  Engine behind reference benchmark-vllm/vllm_manager.py:371-374; max_model_len /
// block_size from reference benchmark-vllm/models.py:15,10.
//
// Kernel `prefill_mma_kernel` (FlashAttention-2 dataflow, mma.sync m16n8k16 bf16):
//   * CTA = 128 query tokens of one q head (8 consumer warps x 16 rows) + 1 producer warp.
//   * producer: walks the chunk's block table, one TMA box per (block, kv head) tile and tensor into a
//     4-stage ring of 64-token K|V stages (128-byte swizzle), byte-counted mbarriers -- the same tile
//     format the decode kernel streams.
//   * consumers: S = Q.K^T (Q fragments live in registers), fp32 online softmax (exp2, scale folded),
//     O += P.V with P rounded to bf16; the causal diagonal and the tail beyond the sequence are masked,
//     V rows beyond the sequence are zeroed so uninitialised cache bytes cannot reach the MMA.
//   * grid = (Hq, q tiles); heads of one kv group are adjacent so their KV tiles meet in L2; q tiles
//     are issued heaviest (latest in the sequence) first.
// Tensor-bound (4*d flop per (q, kv) pair against 2*d*2 B of KV per kv token per 128 queries).
#include "b2_common.cuh"

#include <stdlib.h>

namespace {

constexpr int D = 128;
constexpr int BM = 128;          // query rows per CTA
constexpr int NCW = BM / 16;     // consumer warps
constexpr int BN = 64;           // kv tokens per stage
constexpr int NSTAGE = 4;
constexpr int NTHREADS = (NCW + 1) * 32;
constexpr int TILE_BYTES = BN * D * 2;           // one K (or V) stage tile
constexpr int STAGE_BYTES = 2 * TILE_BYTES;
constexpr int RING_BYTES = NSTAGE * STAGE_BYTES;  // 128 KiB
constexpr int SMEM_BYTES = RING_BYTES + 2 * NSTAGE * 8 + 64;
constexpr float LOG2E = 1.4426950408889634f;

struct PrefillParams {
    __nv_bfloat16* out;
    const __nv_bfloat16* q;
    const int32_t* block_tables;   // [n_chunks, max_nb]
    const int32_t* cu_q;           // [n_chunks + 1] first token row of each chunk
    const int32_t* ctx_start;      // [n_chunks] tokens of the sequence already in the cache before the chunk
    int n_chunks, Hq, G, bs, max_nb;
    int64_t q_stride, out_stride;
    float scale_log2;
    const float* k_scale;          // optional device scalars: the cache holds K / k_scale and V / v_scale
    const float* v_scale;          // (an fp8 cache dequantised into bf16 scratch blocks, exactly)
};

__device__ __forceinline__ uint32_t tile_addr(uint32_t base, int row, int chunk) {
    const int line = 2 * row + (chunk >> 3);           // 128-byte swizzle line (bf16 row = 2 lines)
    return base + line * 128 + (((chunk & 7) ^ (line & 7)) << 4);
}

__global__ void __launch_bounds__(NTHREADS, 1)
prefill_mma_kernel(const __grid_constant__ CUtensorMap tmK, const __grid_constant__ CUtensorMap tmV,
                   const PrefillParams p) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    __shared__ int s_tile[4];   // chunk, q0 (row of the tile inside the chunk), len, start

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int hq = blockIdx.x;
    const int kvh = hq / p.G;

    // q tile -> (chunk, offset): tiles are numbered chunk after chunk; heaviest (last) tiles of a chunk first is
    // approximated by walking the grid backwards.
    if (tid == 0) {
        int want = (int)(gridDim.y - 1 - blockIdx.y), c = 0, found = -1, q0 = 0;
        for (; c < p.n_chunks; ++c) {
            const int len = p.cu_q[c + 1] - p.cu_q[c];
            const int nt = (len + BM - 1) / BM;
            if (want < nt) { found = c; q0 = want * BM; break; }
            want -= nt;
        }
        s_tile[0] = found;
        s_tile[1] = q0;
        s_tile[2] = found >= 0 ? p.cu_q[found + 1] - p.cu_q[found] : 0;
        s_tile[3] = found >= 0 ? p.ctx_start[found] : 0;
    }
    const uint32_t ring = b2::smem_u32(smem_raw);
    const uint32_t bar_full = ring + RING_BYTES;
    const uint32_t bar_empty = bar_full + NSTAGE * 8;
    if (tid == 0) {
        if (ring & 1023u) __trap();
        for (int s = 0; s < NSTAGE; ++s) {
            b2::mbar_init(bar_full + s * 8, 1);
            b2::mbar_init(bar_empty + s * 8, NCW);
        }
        b2::fence_mbar_init();
    }
    __syncthreads();
    const int chunk = s_tile[0];
    if (chunk < 0) return;                       // grid upper bound: no such tile
    const int q0 = s_tile[1], len = s_tile[2], start = s_tile[3];
    const int rows = min(BM, len - q0);          // valid query rows of this tile
    const int kv_hi = start + q0 + rows;         // keys [0, kv_hi) are visible to the tile's last row
    const int n_tiles = (kv_hi + BN - 1) / BN;
    const int row0 = p.cu_q[chunk] + q0;         // first token row in q / out

    if (warp == NCW) {
        // ---------------------------------- producer ----------------------------------
        if (lane == 0) {
            b2::tma_prefetch_desc(&tmK);
            b2::tma_prefetch_desc(&tmV);
        }
        uint64_t pol;   // KV tiles are re-read by the G heads of the group and by later q tiles: keep them in L2
        asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
        const int32_t* bt = p.block_tables + (int64_t)chunk * p.max_nb;
        const int bps = BN / p.bs;                       // blocks per stage
        const uint32_t box = p.bs * D * 2;
        const int n_blk = (kv_hi + p.bs - 1) / p.bs;
        int cur = lane < n_blk ? bt[lane] : 0;           // 32 table entries per batch, next batch prefetched
        int nxt = 32 + lane < n_blk ? bt[32 + lane] : 0;
        for (int it = 0; it < n_tiles; ++it) {
            const uint32_t s = it % NSTAGE, ph = (it / NSTAGE) & 1;
            b2::mbar_wait(bar_empty + s * 8, ph ^ 1);
            const int nb = min(bps, n_blk - it * bps);
            if (lane == 0) b2::mbar_arrive_expect_tx(bar_full + s * 8, nb * 2 * box);
            const uint32_t kdst = ring + s * STAGE_BYTES, vdst = kdst + TILE_BYTES;
            for (int j = 0; j < bps; ++j) {
                const int gi = it * bps + j;
                const int blk = __shfl_sync(0xffffffffu, cur, gi & 31);
                if (j < nb && lane == 0) {
                    b2::tma_load_5d(kdst + j * box, &tmK, bar_full + s * 8, 0, 0, 0, kvh, blk, pol);
                    b2::tma_load_5d(vdst + j * box, &tmV, bar_full + s * 8, 0, 0, 0, kvh, blk, pol);
                }
                if ((gi & 31) == 31) {
                    cur = nxt;
                    nxt = gi + 33 + lane < n_blk ? bt[gi + 33 + lane] : 0;
                }
            }
        }
        return;
    }

    // ---------------------------------- consumers ----------------------------------
    const int g = lane >> 2, t4 = lane & 3;
    const int mi = lane >> 3, ri = lane & 7;
    const int wr = warp * 16;                          // first row of this warp inside the tile
    // Q fragments: rows wr+g and wr+g+8, all 8 k-steps of d
    uint32_t qf[8][4];
    {
        const __nv_bfloat16* qa = p.q + (int64_t)(row0 + wr + g) * p.q_stride + (int64_t)hq * D;
        const __nv_bfloat16* qb = qa + 8 * p.q_stride;
        const bool va = wr + g < rows, vb = wr + g + 8 < rows;
#pragma unroll
        for (int kk = 0; kk < 8; ++kk) {
            const int c = 16 * kk + 2 * t4;
            qf[kk][0] = va ? *reinterpret_cast<const uint32_t*>(qa + c) : 0u;
            qf[kk][1] = vb ? *reinterpret_cast<const uint32_t*>(qb + c) : 0u;
            qf[kk][2] = va ? *reinterpret_cast<const uint32_t*>(qa + c + 8) : 0u;
            qf[kk][3] = vb ? *reinterpret_cast<const uint32_t*>(qb + c + 8) : 0u;
        }
    }
    float o[16][4];
#pragma unroll
    for (int i = 0; i < 16; ++i) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;
    float m_a = -INFINITY, m_b = -INFINITY, l_a = 0.f, l_b = 0.f;
    const float sl2 = p.scale_log2 * (p.k_scale ? *p.k_scale : 1.f);
    // last visible key of rows a / b (padding rows behave like the tile's last valid row)
    const int lim_a = start + q0 + min(wr + g, rows - 1);
    const int lim_b = start + q0 + min(wr + g + 8, rows - 1);
    const int warp_hi = start + q0 + min(wr + 15, rows - 1);   // last key any row of this warp sees
    const int warp_lo = start + q0 + min(wr, rows - 1);        // keys <= warp_lo are visible to every row

    for (int it = 0; it < n_tiles; ++it) {
        const uint32_t s = it % NSTAGE, ph = (it / NSTAGE) & 1;
        b2::mbar_wait(bar_full + s * 8, ph);
        const int j0 = it * BN;
        if (j0 <= warp_hi && wr < rows) {
            const uint32_t kbase = ring + s * STAGE_BYTES, vbase = kbase + TILE_BYTES;
            // ------------------------------ S = Q . K^T ------------------------------
            float sc[8][4];
#pragma unroll
            for (int nt = 0; nt < 8; ++nt) sc[nt][0] = sc[nt][1] = sc[nt][2] = sc[nt][3] = 0.f;
#pragma unroll
            for (int np = 0; np < 4; ++np) {
                const int tok = 16 * np + (mi >> 1) * 8 + ri;
#pragma unroll
                for (int kk = 0; kk < 8; ++kk) {
                    uint32_t b0, b1, b2_, b3;
                    b2::ldmatrix_x4(tile_addr(kbase, tok, 2 * kk + (mi & 1)), b0, b1, b2_, b3);
                    b2::mma_bf16_16816(sc[2 * np], qf[kk][0], qf[kk][1], qf[kk][2], qf[kk][3], b0, b1);
                    b2::mma_bf16_16816(sc[2 * np + 1], qf[kk][0], qf[kk][1], qf[kk][2], qf[kk][3], b2_, b3);
                }
            }
            // ------------------------------ mask + online softmax ------------------------------
            const bool need_mask = j0 + BN - 1 > warp_lo;
            if (need_mask) {
#pragma unroll
                for (int nt = 0; nt < 8; ++nt) {
                    const int j = j0 + 8 * nt + 2 * t4;
                    if (j > lim_a) sc[nt][0] = -INFINITY;
                    if (j + 1 > lim_a) sc[nt][1] = -INFINITY;
                    if (j > lim_b) sc[nt][2] = -INFINITY;
                    if (j + 1 > lim_b) sc[nt][3] = -INFINITY;
                }
            }
            float mx_a = -INFINITY, mx_b = -INFINITY;
#pragma unroll
            for (int nt = 0; nt < 8; ++nt) {
                mx_a = fmaxf(mx_a, fmaxf(sc[nt][0], sc[nt][1]));
                mx_b = fmaxf(mx_b, fmaxf(sc[nt][2], sc[nt][3]));
            }
            mx_a = fmaxf(mx_a, __shfl_xor_sync(0xffffffffu, mx_a, 1));
            mx_a = fmaxf(mx_a, __shfl_xor_sync(0xffffffffu, mx_a, 2));
            mx_b = fmaxf(mx_b, __shfl_xor_sync(0xffffffffu, mx_b, 1));
            mx_b = fmaxf(mx_b, __shfl_xor_sync(0xffffffffu, mx_b, 2));
            // key 0 is visible to every row and tiles arrive in order, so the running max is finite from tile 0 on
            const float mn_a = fmaxf(m_a, mx_a), mn_b = fmaxf(m_b, mx_b);
            const float al_a = b2::fast_exp2((m_a - mn_a) * sl2), al_b = b2::fast_exp2((m_b - mn_b) * sl2);
            m_a = mn_a;
            m_b = mn_b;
            const float ms_a = mn_a * sl2, ms_b = mn_b * sl2;
            uint32_t pf[8][2];
            float rs_a = 0.f, rs_b = 0.f;
#pragma unroll
            for (int nt = 0; nt < 8; ++nt) {
                const float p0 = b2::fast_exp2(fmaf(sc[nt][0], sl2, -ms_a));
                const float p1 = b2::fast_exp2(fmaf(sc[nt][1], sl2, -ms_a));
                const float p2 = b2::fast_exp2(fmaf(sc[nt][2], sl2, -ms_b));
                const float p3 = b2::fast_exp2(fmaf(sc[nt][3], sl2, -ms_b));
                pf[nt][0] = b2::pack_bf16x2(p0, p1);
                pf[nt][1] = b2::pack_bf16x2(p2, p3);
                // the denominator sums the ROUNDED weights: the output stays a convex combination of V rows
                rs_a += __uint_as_float(pf[nt][0] << 16) + __uint_as_float(pf[nt][0] & 0xffff0000u);
                rs_b += __uint_as_float(pf[nt][1] << 16) + __uint_as_float(pf[nt][1] & 0xffff0000u);
            }
            l_a = fmaf(l_a, al_a, rs_a);
            l_b = fmaf(l_b, al_b, rs_b);
#pragma unroll
            for (int dt = 0; dt < 16; ++dt) {
                o[dt][0] *= al_a;
                o[dt][1] *= al_a;
                o[dt][2] *= al_b;
                o[dt][3] *= al_b;
            }
            // ------------------------------ O += P . V ------------------------------
            const int n_valid = kv_hi - j0;                 // keys of this stage that exist in the cache
#pragma unroll
            for (int np = 0; np < 4; ++np) {
                if (16 * np >= n_valid) break;              // warp-uniform: nothing valid from here on
                uint32_t mk_lo = 0xffffffffu, mk_hi = 0xffffffffu;
                if (16 * np + 16 > n_valid) {
                    const int r = n_valid - 16 * np;
                    mk_lo = (2 * t4 < r ? 0x0000ffffu : 0u) | (2 * t4 + 1 < r ? 0xffff0000u : 0u);
                    mk_hi = (8 + 2 * t4 < r ? 0x0000ffffu : 0u) | (9 + 2 * t4 < r ? 0xffff0000u : 0u);
                }
                const int tok = 16 * np + (mi & 1) * 8 + ri;
#pragma unroll
                for (int dp = 0; dp < 8; ++dp) {
                    uint32_t b0, b1, b2_, b3;
                    b2::ldmatrix_x4_trans(tile_addr(vbase, tok, 2 * dp + (mi >> 1)), b0, b1, b2_, b3);
                    b0 &= mk_lo; b1 &= mk_hi; b2_ &= mk_lo; b3 &= mk_hi;
                    b2::mma_bf16_16816(o[2 * dp], pf[2 * np][0], pf[2 * np][1], pf[2 * np + 1][0], pf[2 * np + 1][1], b0, b1);
                    b2::mma_bf16_16816(o[2 * dp + 1], pf[2 * np][0], pf[2 * np][1], pf[2 * np + 1][0], pf[2 * np + 1][1], b2_, b3);
                }
            }
        }
        __syncwarp();
        if (lane == 0) b2::mbar_arrive(bar_empty + s * 8);
    }

    // ------------------------------ epilogue ------------------------------
    l_a += __shfl_xor_sync(0xffffffffu, l_a, 1);
    l_a += __shfl_xor_sync(0xffffffffu, l_a, 2);
    l_b += __shfl_xor_sync(0xffffffffu, l_b, 1);
    l_b += __shfl_xor_sync(0xffffffffu, l_b, 2);
    const float vsc = p.v_scale ? *p.v_scale : 1.f;
    const float inv_a = vsc / l_a, inv_b = vsc / l_b;
    __nv_bfloat16* oa = p.out + (int64_t)(row0 + wr + g) * p.out_stride + (int64_t)hq * D;
    __nv_bfloat16* ob = oa + 8 * p.out_stride;
    const bool va = wr + g < rows, vb = wr + g + 8 < rows;
#pragma unroll
    for (int dt = 0; dt < 16; ++dt) {
        const int c = 8 * dt + 2 * t4;
        if (va) *reinterpret_cast<uint32_t*>(oa + c) = b2::pack_bf16x2(o[dt][0] * inv_a, o[dt][1] * inv_a);
        if (vb) *reinterpret_cast<uint32_t*>(ob + c) = b2::pack_bf16x2(o[dt][2] * inv_b, o[dt][3] * inv_b);
    }
}

}  // namespace

// =============================================================================================
// C ABI
// =============================================================================================
extern "C" int b2_prefill_attention(void* out, const void* q, const void* k_cache, const void* v_cache,
                                    const int32_t* block_tables, const int32_t* cu_q, const int32_t* ctx_start,
                                    int n_chunks, int T, int Hq, int Hkv, int d, int bs, int max_nb,
                                    int64_t num_blocks, int64_t q_stride, int64_t out_stride,
                                    int64_t cache_block_stride, int64_t cache_head_stride, float scale, int kv_dtype,
                                    const float* k_scale, const float* v_scale, void* stream) {
    B2_REQUIRE(n_chunks >= 0 && T >= 0 && Hq > 0 && Hkv > 0 && max_nb > 0 && num_blocks > 0, B2_EINVAL,
               "prefill_attention: bad sizes n_chunks=%d T=%d Hq=%d Hkv=%d max_nb=%d", n_chunks, T, Hq, Hkv, max_nb);
    if (n_chunks == 0 || T == 0) return B2_OK;
    B2_REQUIRE(out && q && k_cache && v_cache && block_tables && cu_q && ctx_start, B2_EINVAL,
               "prefill_attention: null pointer");
    B2_REQUIRE(d == D, B2_EUNSUPPORTED, "prefill_attention: head_dim %d (only 128 is built)", d);
    B2_REQUIRE(bs == 8 || bs == 16 || bs == 32 || bs == 64, B2_EUNSUPPORTED, "prefill_attention: block_size %d", bs);
    B2_REQUIRE(kv_dtype == B2_KV_BF16, B2_EUNSUPPORTED,
               "prefill_attention: kv_dtype %d (prompt chunks attend a bf16 cache; fp8 KV is decode-only)", kv_dtype);
    B2_REQUIRE(Hq % Hkv == 0, B2_EINVAL, "prefill_attention: Hq=%d not a multiple of Hkv=%d", Hq, Hkv);
    B2_REQUIRE(b2_aligned16(out) && b2_aligned16(q) && b2_aligned16(k_cache) && b2_aligned16(v_cache), B2_EINVAL,
               "prefill_attention: pointers must be 16-byte aligned");
    B2_REQUIRE(q_stride % 8 == 0 && out_stride % 8 == 0 && q_stride >= (int64_t)Hq * d && out_stride >= (int64_t)Hq * d,
               B2_EINVAL, "prefill_attention: bad q/out strides");
    B2_REQUIRE(cache_block_stride % 8 == 0 && cache_head_stride % 8 == 0 && cache_head_stride >= (int64_t)bs * d,
               B2_EINVAL, "prefill_attention: bad cache strides");
    const int64_t n_qtiles = (int64_t)T / BM + n_chunks;       // upper bound of sum ceil(len_c / BM)
    B2_REQUIRE(n_qtiles <= 65535, B2_EUNSUPPORTED, "prefill_attention: %lld q tiles exceed the grid", (long long)n_qtiles);

    // default: the tcgen05 kernel (prefill_tc5.cu); B2_PREFILL_IMPL=mma keeps this file's mma.sync kernel as the A/B
    static const bool use_mma = []() { const char* e = getenv("B2_PREFILL_IMPL"); return e && e[0] == 'm'; }();
    if (!use_mma)
        return b2_internal_prefill_tc5(out, q, k_cache, v_cache, block_tables, cu_q, ctx_start, n_chunks, T, Hq, Hkv, bs,
                                       max_nb, num_blocks, q_stride, out_stride, cache_block_stride, cache_head_stride,
                                       scale, k_scale, v_scale, stream);

    CUtensorMap tmK, tmV;
    int rc = b2_internal_make_cache_tmap(&tmK, k_cache, kv_dtype, num_blocks, Hkv, bs, cache_block_stride, cache_head_stride);
    if (rc != B2_OK) return rc;
    rc = b2_internal_make_cache_tmap(&tmV, v_cache, kv_dtype, num_blocks, Hkv, bs, cache_block_stride, cache_head_stride);
    if (rc != B2_OK) return rc;

    PrefillParams p;
    p.out = static_cast<__nv_bfloat16*>(out);
    p.q = static_cast<const __nv_bfloat16*>(q);
    p.block_tables = block_tables;
    p.cu_q = cu_q;
    p.ctx_start = ctx_start;
    p.n_chunks = n_chunks;
    p.Hq = Hq;
    p.G = Hq / Hkv;
    p.bs = bs;
    p.max_nb = max_nb;
    p.q_stride = q_stride;
    p.out_stride = out_stride;
    p.scale_log2 = scale * LOG2E;
    p.k_scale = k_scale;
    p.v_scale = v_scale;

    static bool attr_set = false;
    if (!attr_set) {
        B2_CUDA_OK(cudaFuncSetAttribute(prefill_mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
        attr_set = true;
    }
    prefill_mma_kernel<<<dim3(Hq, (unsigned)n_qtiles), NTHREADS, SMEM_BYTES, static_cast<cudaStream_t>(stream)>>>(tmK, tmV, p);
    B2_CUDA_OK(cudaGetLastError());
    return B2_OK;
}
