// paged_decode.cu -- paged-attention decode for B200 (sm_100a).
// SURVEY.md section 8(a) rows a6 (single pass), a7 (split-K + merge), a8 (fp8 dequant in load).
//
// Contract: include/b200attn.h `b2_paged_decode`.  Semantics follow vLLM paged_attention_v1/v2,
// the op run each decode step by the engine the reference launches
// (reference benchmark-vllm/vllm_manager.py:371-374); block_size / max_num_seqs / max_model_len
// come from reference benchmark-vllm/models.py:9,10,15.
//
// Design (HBM-bound gather):
//   * grid = (split, kv_head x head_group, seq).  One CTA streams the K and V tiles of its
//     (sequence, kv head, KV-length split) exactly once and serves all G = Hq/Hkv query heads
//     that share that kv head, so KV bytes are read once per kv head, not once per q head.
//   * warp 4 is the producer: it walks the block table and issues one TMA tiled copy
//     (cp.async.bulk.tensor.4d, 128-byte swizzle, L2 evict-first) per 16 x 64-element half tile
//     into a ring of NSTAGE x 64-token stages; completion is counted in bytes on an mbarrier.
//   * warps 0-3 are consumers: each owns 16 tokens of every stage.  S^T = K.Q^T and
//     O^T += V^T.P^T run on mma.sync m16n8k16 (tokens / head_dim on the M side, the <=8 grouped
//     query heads on the N side, so no MMA row is wasted on padding tokens); ldmatrix(.trans)
//     feeds the A operand straight from the swizzled stage; softmax is fp32 with warp shuffles;
//     the P^T B-operand is produced from the S^T accumulator by movmatrix.
//   * the four consumers merge their (m, l, O) in shared memory; with num_splits > 1 the CTA
//     writes a normalised partial + log2-sum-exp and `merge_splits_kernel` combines them.
#include "b2_common.cuh"

namespace {

constexpr int D = 128;        // head_dim (Llama-3 geometry)
constexpr int NCW = 4;        // consumer warps
constexpr int TOK = NCW * 16; // tokens per pipeline stage
constexpr int NTHREADS = (NCW + 1) * 32;
constexpr float LOG2E = 1.4426950408889634f;

struct DecodeParams {
    __nv_bfloat16* out;
    const __nv_bfloat16* q;
    const int32_t* block_tables;
    const int32_t* seq_lens;
    float* ws_o;     // [B, Hq, num_splits, D] normalised partial outputs
    float* ws_lse;   // [B, Hq, num_splits]    log2-sum-exp of each partial
    const float* k_scale;
    const float* v_scale;
    int Hq, Hkv, G, HG, max_nb, num_splits, split_tokens;
    int64_t q_stride, out_stride;
    float scale_log2;  // softmax scale * log2(e)
};

template <int BS, bool FP8, int NT, int NSTAGE>
struct Cfg {
    static constexpr int ESZ = FP8 ? 1 : 2;
    static constexpr int ROWB = D * ESZ;           // bytes per token row
    static constexpr int NHALF = ROWB / 128;       // 128-byte swizzle spans per row
    static constexpr int TILE_BYTES = TOK * ROWB;  // one K (or V) stage tile
    static constexpr int STAGE_BYTES = 2 * TILE_BYTES;
    static constexpr int BLK_PER_STAGE = TOK / BS;
    static constexpr int BOX_BYTES = BS * 128;     // one TMA box
    static constexpr int HEADS = 8 * NT;           // q heads served per CTA
    static constexpr int RING_BYTES = NSTAGE * STAGE_BYTES;
    static constexpr int MERGE_BYTES = NCW * (D * HEADS + 2 * HEADS) * 4;
    static constexpr int SMEM_BYTES = 1024 + RING_BYTES + 2 * NSTAGE * 8;
    static_assert(TOK % BS == 0 && BS % 8 == 0, "block size must divide the stage and keep the swizzle phase");
    static_assert(MERGE_BYTES <= RING_BYTES, "merge scratch reuses the ring");
};

// fp8 (e4m3 x2, low 16 bits of `v`) -> f16x2, exact
__device__ __forceinline__ uint32_t e4m3x2_to_f16x2(uint32_t v) {
    uint32_t r;
    asm("{\n\t.reg .b16 lo;\n\tcvt.u16.u32 lo, %1;\n\tcvt.rn.f16x2.e4m3x2 %0, lo;\n\t}" : "=r"(r) : "r"(v));
    return r;
}
__device__ __forceinline__ uint32_t pack_f16x2(float lo, float hi) {
    uint32_t r;
    asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
    return r;
}
__device__ __forceinline__ void mma_f16_16816(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2,
                                              uint32_t a3, uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
        "{%0,%1,%2,%3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

template <bool FP8>
__device__ __forceinline__ void mma_kv(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2,
                                       uint32_t a3, uint32_t b0, uint32_t b1) {
    if (FP8) mma_f16_16816(c, a0, a1, a2, a3, b0, b1);
    else b2::mma_bf16_16816(c, a0, a1, a2, a3, b0, b1);
}

template <int BS, bool FP8, int NT, int NSTAGE, int MINB>
__global__ void __launch_bounds__(NTHREADS, MINB)
paged_decode_kernel(const __grid_constant__ CUtensorMap tmK, const __grid_constant__ CUtensorMap tmV,
                    const DecodeParams p) {
    using C = Cfg<BS, FP8, NT, NSTAGE>;
    extern __shared__ uint8_t smem_raw[];

    const int split = blockIdx.x;
    const int kvh = blockIdx.y / p.HG;
    const int hg = blockIdx.y - kvh * p.HG;
    const int b = blockIdx.z;
    const int tid = threadIdx.x;
    const int warp = tid >> 5;
    const int lane = tid & 31;

    const int seq_len = p.seq_lens[b];
    const int t_begin = split * p.split_tokens;
    const int head0 = kvh * p.G + hg * C::HEADS;                 // first q head of this CTA
    const int nheads = min(C::HEADS, p.G - hg * C::HEADS);       // valid q heads of this CTA

    if (t_begin >= seq_len) {
        // empty split: nothing to read.  Single-pass mode must still define the output row.
        if (p.num_splits == 1) {
            for (int i = tid; i < nheads * D; i += NTHREADS)
                p.out[b * p.out_stride + (int64_t)(head0 + i / D) * D + (i % D)] = __float2bfloat16(0.f);
        }
        return;
    }
    const int t_end = min(seq_len, t_begin + p.split_tokens);
    const int n_tok = t_end - t_begin;
    const int n_stage = (n_tok + TOK - 1) / TOK;
    const int n_blk = (n_tok + BS - 1) / BS;
    const int blk0 = t_begin / BS;

    const uint32_t base = (b2::smem_u32(smem_raw) + 1023u) & ~1023u;   // swizzle phase = address bits
    const uint32_t bar_full = base + C::RING_BYTES;
    const uint32_t bar_empty = bar_full + NSTAGE * 8;

    if (tid == 0) {
        for (int s = 0; s < NSTAGE; ++s) {
            b2::mbar_init(bar_full + s * 8, 1);       // producer's arrive.expect_tx
            b2::mbar_init(bar_empty + s * 8, NCW);    // one arrive per consumer warp
        }
        b2::fence_mbar_init();
    }
    __syncthreads();

    if (warp == NCW) {
        // ================================ producer warp =====================================
        if (lane == 0) {
            b2::tma_prefetch_desc(&tmK);
            b2::tma_prefetch_desc(&tmV);
        }
        const uint64_t pol = b2::policy_evict_first();
        const int32_t* bt = p.block_tables + (int64_t)b * p.max_nb + blk0;
        int cur = (lane < n_blk) ? bt[lane] : 0;            // 32 table entries per batch
        int nxt = (32 + lane < n_blk) ? bt[32 + lane] : 0;  // next batch, prefetched
        for (int it = 0; it < n_stage; ++it) {
            const int s = it % NSTAGE;
            const uint32_t ph = (it / NSTAGE) & 1;
            b2::mbar_wait(bar_empty + s * 8, ph ^ 1);
            const int nb_stage = min(C::BLK_PER_STAGE, n_blk - it * C::BLK_PER_STAGE);
            const uint32_t kdst = base + s * C::STAGE_BYTES;
            const uint32_t vdst = kdst + C::TILE_BYTES;
            if (lane == 0)
                b2::mbar_arrive_expect_tx(bar_full + s * 8, nb_stage * 2 * C::NHALF * C::BOX_BYTES);
#pragma unroll
            for (int j = 0; j < C::BLK_PER_STAGE; ++j) {
                const int gi = it * C::BLK_PER_STAGE + j;
                const int src = gi & 31;
                const int blk = __shfl_sync(0xffffffffu, cur, src);
                if (j < nb_stage && lane == 0) {
#pragma unroll
                    for (int h = 0; h < C::NHALF; ++h) {
                        const uint32_t off = h * (TOK * 128) + j * C::BOX_BYTES;
                        b2::tma_load_4d(kdst + off, &tmK, bar_full + s * 8, h * 64, 0, kvh, blk, pol);
                        b2::tma_load_4d(vdst + off, &tmV, bar_full + s * 8, h * 64, 0, kvh, blk, pol);
                    }
                }
                if (src == 31) {  // warp-uniform: this batch of 32 entries is used up
                    cur = nxt;
                    nxt = (gi + 33 + lane < n_blk) ? bt[gi + 33 + lane] : 0;
                }
            }
        }
    } else {
        // ================================ consumer warps ====================================
        const int g = lane >> 2;   // mma "group id": row of A/C, column of B
        const int t4 = lane & 3;   // thread in group
        float sl2 = p.scale_log2;
        if (FP8) sl2 *= *p.k_scale;   // K dequant scale folds into the softmax scale

        // ---- Q^T as the B operand of S^T = K.Q^T: b0 = Q[head g][k-pair], b1 = the +8 pair.
        // bf16: logical k == physical d.  fp8: ldmatrix hands each thread 4 consecutive fp8 of a
        // 16-element chunk, so logical k = 2*t4+{0,1} <-> d = 16kc + 4*t4 + {0,1} and the +8 pair
        // <-> d = 16kc + 4*t4 + {2,3}; the contraction is permutation invariant.
        uint32_t qf[NT][8][2];
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
            const int hl = nt * 8 + g;
            const bool valid = hl < nheads;
            const uint32_t* qrow =
                reinterpret_cast<const uint32_t*>(p.q + b * p.q_stride + (int64_t)(head0 + (valid ? hl : 0)) * D);
#pragma unroll
            for (int kc = 0; kc < 8; ++kc) {
                uint32_t w0, w1;
                if (!FP8) {
                    w0 = qrow[kc * 8 + t4];
                    w1 = qrow[kc * 8 + 4 + t4];
                } else {
                    w0 = qrow[kc * 8 + 2 * t4];
                    w1 = qrow[kc * 8 + 2 * t4 + 1];
                    // bf16x2 -> f16x2 (the fp8 path runs the f16 MMA; see DESIGN.md on range)
                    w0 = pack_f16x2(__uint_as_float(w0 << 16), __uint_as_float(w0 & 0xffff0000u));
                    w1 = pack_f16x2(__uint_as_float(w1 << 16), __uint_as_float(w1 & 0xffff0000u));
                }
                qf[nt][kc][0] = valid ? w0 : 0u;
                qf[nt][kc][1] = valid ? w1 : 0u;
            }
        }

        float o[NT][8][4];
        float m_run[NT][2], l_run[NT][2];
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
            m_run[nt][0] = m_run[nt][1] = -INFINITY;
            l_run[nt][0] = l_run[nt][1] = 0.f;
#pragma unroll
            for (int dt = 0; dt < 8; ++dt) o[nt][dt][0] = o[nt][dt][1] = o[nt][dt][2] = o[nt][dt][3] = 0.f;
        }

        // per-lane ldmatrix geometry: lane -> (matrix mi, row ri)
        const int mi = lane >> 3, ri = lane & 7;
        const int row_base = warp * 16;

        for (int it = 0; it < n_stage; ++it) {
            const int s = it % NSTAGE;
            const uint32_t ph = (it / NSTAGE) & 1;
            b2::mbar_wait(bar_full + s * 8, ph);
            const int nvalid = n_tok - (it * TOK + row_base);   // valid tokens in this warp's slab
            if (nvalid > 0) {
                const uint32_t kbase = base + s * C::STAGE_BYTES;
                const uint32_t vbase = kbase + C::TILE_BYTES;

                // ------------------------------ S^T = K . Q^T ------------------------------
                float sacc[NT][4];
#pragma unroll
                for (int nt = 0; nt < NT; ++nt) sacc[nt][0] = sacc[nt][1] = sacc[nt][2] = sacc[nt][3] = 0.f;
                if (!FP8) {
                    // matrices: 0 = rows 0-7 / chunk lo, 1 = rows 8-15 / lo, 2 = rows 0-7 / hi, 3 = rows 8-15 / hi
                    const int row = row_base + (mi & 1) * 8 + ri;
#pragma unroll
                    for (int kc = 0; kc < 8; ++kc) {
                        const int chunk = 2 * kc + (mi >> 1);        // 16-byte chunk of the 256-byte row
                        const uint32_t addr = kbase + (chunk >> 3) * (TOK * 128) + row * 128 +
                                              (((chunk & 7) ^ (row & 7)) << 4);
                        uint32_t a0, a1, a2, a3;
                        b2::ldmatrix_x4(addr, a0, a1, a2, a3);
#pragma unroll
                        for (int nt = 0; nt < NT; ++nt)
                            mma_kv<FP8>(sacc[nt], a0, a1, a2, a3, qf[nt][kc][0], qf[nt][kc][1]);
                    }
                } else {
                    // fp8 rows are 128 bytes = 8 chunks of 16 elements; one ldmatrix.x4 covers two k-chunks:
                    // matrices: 0 = rows 0-7 / chunk c, 1 = rows 8-15 / c, 2 = rows 0-7 / c+1, 3 = rows 8-15 / c+1
                    const int row = row_base + (mi & 1) * 8 + ri;
#pragma unroll
                    for (int kc2 = 0; kc2 < 4; ++kc2) {
                        const int chunk = 2 * kc2 + (mi >> 1);
                        const uint32_t addr = kbase + row * 128 + (((chunk & 7) ^ (row & 7)) << 4);
                        uint32_t r0, r1, r2, r3;
                        b2::ldmatrix_x4(addr, r0, r1, r2, r3);
#pragma unroll
                        for (int nt = 0; nt < NT; ++nt) {
                            mma_kv<FP8>(sacc[nt], e4m3x2_to_f16x2(r0), e4m3x2_to_f16x2(r1),
                                        e4m3x2_to_f16x2(r0 >> 16), e4m3x2_to_f16x2(r1 >> 16),
                                        qf[nt][2 * kc2][0], qf[nt][2 * kc2][1]);
                            mma_kv<FP8>(sacc[nt], e4m3x2_to_f16x2(r2), e4m3x2_to_f16x2(r3),
                                        e4m3x2_to_f16x2(r2 >> 16), e4m3x2_to_f16x2(r3 >> 16),
                                        qf[nt][2 * kc2 + 1][0], qf[nt][2 * kc2 + 1][1]);
                        }
                    }
                }
                // sacc[nt] = {S[tok g][head 2t4], S[tok g][2t4+1], S[tok g+8][2t4], S[tok g+8][2t4+1]}
                if (nvalid < 16) {
#pragma unroll
                    for (int nt = 0; nt < NT; ++nt) {
                        if (g >= nvalid) sacc[nt][0] = sacc[nt][1] = -INFINITY;
                        if (g + 8 >= nvalid) sacc[nt][2] = sacc[nt][3] = -INFINITY;
                    }
                }

                // ------------------------------ online softmax ------------------------------
                uint32_t pb[NT][2];
#pragma unroll
                for (int nt = 0; nt < NT; ++nt) {
                    float pr[4];
#pragma unroll
                    for (int j = 0; j < 2; ++j) {
                        float mx = fmaxf(sacc[nt][j], sacc[nt][j + 2]);
                        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 4));
                        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 8));
                        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 16));
                        const float m_new = fmaxf(m_run[nt][j], mx);       // finite: slab has a valid token
                        const float alpha = b2::fast_exp2((m_run[nt][j] - m_new) * sl2);
                        const float ms = m_new * sl2;
                        pr[j] = b2::fast_exp2(fmaf(sacc[nt][j], sl2, -ms));
                        pr[j + 2] = b2::fast_exp2(fmaf(sacc[nt][j + 2], sl2, -ms));
                        m_run[nt][j] = m_new;
                        l_run[nt][j] = fmaf(l_run[nt][j], alpha, pr[j] + pr[j + 2]);
#pragma unroll
                        for (int dt = 0; dt < 8; ++dt) {
                            o[nt][dt][j] *= alpha;
                            o[nt][dt][j + 2] *= alpha;
                        }
                    }
                    // P^T as the B operand of O^T += V^T.P^T: transpose the 8x8 (token x head) tiles
                    const uint32_t x0 = FP8 ? pack_f16x2(pr[0], pr[1]) : b2::pack_bf16x2(pr[0], pr[1]);
                    const uint32_t x1 = FP8 ? pack_f16x2(pr[2], pr[3]) : b2::pack_bf16x2(pr[2], pr[3]);
                    pb[nt][0] = b2::movmatrix_trans(x0);   // (tokens 2t4,2t4+1 ; head g)
                    pb[nt][1] = b2::movmatrix_trans(x1);   // (tokens 8+2t4,.. ; head g)
                }

                // ------------------------------ O^T += V^T . P^T -----------------------------
                // V garbage beyond seq_len must not reach the MMA (0 * NaN): zero those token columns.
                uint32_t mk_lo = 0xffffffffu, mk_hi = 0xffffffffu;
                if (nvalid < 16) {
                    mk_lo = (2 * t4 < nvalid ? 0x0000ffffu : 0u) | (2 * t4 + 1 < nvalid ? 0xffff0000u : 0u);
                    mk_hi = (8 + 2 * t4 < nvalid ? 0x0000ffffu : 0u) | (9 + 2 * t4 < nvalid ? 0xffff0000u : 0u);
                }
                if (!FP8) {
                    // .trans matrices: 0 = tokens 0-7 / d lo, 1 = tokens 0-7 / d hi, 2 = tokens 8-15 / lo, 3 = tokens 8-15 / hi
                    const int row = row_base + (mi >> 1) * 8 + ri;
#pragma unroll
                    for (int dt = 0; dt < 8; ++dt) {
                        const int chunk = 2 * dt + (mi & 1);
                        const uint32_t addr = vbase + (chunk >> 3) * (TOK * 128) + row * 128 +
                                              (((chunk & 7) ^ (row & 7)) << 4);
                        uint32_t a0, a1, a2, a3;
                        b2::ldmatrix_x4_trans(addr, a0, a1, a2, a3);
                        a0 &= mk_lo; a1 &= mk_lo; a2 &= mk_hi; a3 &= mk_hi;
#pragma unroll
                        for (int nt = 0; nt < NT; ++nt)
                            mma_kv<FP8>(o[nt][dt], a0, a1, a2, a3, pb[nt][0], pb[nt][1]);
                    }
                } else {
                    // A 16-byte fp8 chunk holds 16 d values, so the b16 transpose moves (d 2g, d 2g+1) pairs:
                    // a thread receives {tok 2t4: d 2g, d 2g+1 ; tok 2t4+1: d 2g, d 2g+1}.  PRMT regroups it to
                    // {d 2g: tok pair ; d 2g+1: tok pair}; MMA row g <-> d = 16c + 2g, row g+8 <-> d = 16c + 2g + 1.
                    // .trans matrices: 0 = tokens 0-7 / chunk c, 1 = tokens 8-15 / c, 2 = tokens 0-7 / c+1, 3 = tokens 8-15 / c+1
                    const int row = row_base + (mi & 1) * 8 + ri;
#pragma unroll
                    for (int dt2 = 0; dt2 < 4; ++dt2) {
                        const int chunk = 2 * dt2 + (mi >> 1);
                        const uint32_t addr = vbase + row * 128 + (((chunk & 7) ^ (row & 7)) << 4);
                        uint32_t r0, r1, r2, r3;
                        b2::ldmatrix_x4_trans(addr, r0, r1, r2, r3);
                        r0 = __byte_perm(r0, 0, 0x3120); r1 = __byte_perm(r1, 0, 0x3120);
                        r2 = __byte_perm(r2, 0, 0x3120); r3 = __byte_perm(r3, 0, 0x3120);
#pragma unroll
                        for (int nt = 0; nt < NT; ++nt) {
                            mma_kv<FP8>(o[nt][2 * dt2], e4m3x2_to_f16x2(r0) & mk_lo, e4m3x2_to_f16x2(r0 >> 16) & mk_lo,
                                        e4m3x2_to_f16x2(r1) & mk_hi, e4m3x2_to_f16x2(r1 >> 16) & mk_hi,
                                        pb[nt][0], pb[nt][1]);
                            mma_kv<FP8>(o[nt][2 * dt2 + 1], e4m3x2_to_f16x2(r2) & mk_lo, e4m3x2_to_f16x2(r2 >> 16) & mk_lo,
                                        e4m3x2_to_f16x2(r3) & mk_hi, e4m3x2_to_f16x2(r3 >> 16) & mk_hi,
                                        pb[nt][0], pb[nt][1]);
                        }
                    }
                }
            }
            __syncwarp();
            if (lane == 0) b2::mbar_arrive(bar_empty + s * 8);   // stage may be overwritten
        }

        // lane-local row sums -> full sums over the slab's 16 token rows (lanes differing in g)
#pragma unroll
        for (int nt = 0; nt < NT; ++nt)
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                float l = l_run[nt][j];
                l += __shfl_xor_sync(0xffffffffu, l, 4);
                l += __shfl_xor_sync(0xffffffffu, l, 8);
                l += __shfl_xor_sync(0xffffffffu, l, 16);
                l_run[nt][j] = l;
            }

        // Every stage this CTA armed has been consumed by all four warps once they pass this
        // named barrier, so the ring can be reused as merge scratch.
        asm volatile("bar.sync 1, %0;" ::"n"(NCW * 32));

        // ---- publish (m, l, O^T) of this warp: sc_o[warp][d][head], sc_m/sc_l[warp][head]
        float* sc = reinterpret_cast<float*>(smem_raw + (base - b2::smem_u32(smem_raw)));
        float* sc_o = sc + warp * (D * C::HEADS);
        float* sc_m = sc + NCW * (D * C::HEADS) + warp * C::HEADS;
        float* sc_l = sc_m + NCW * C::HEADS;
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
#pragma unroll
            for (int dt = 0; dt < 8; ++dt) {
                // bf16: MMA row r <-> d = 16dt + r.  fp8: row g <-> d = 16dt + 2g, row g+8 <-> d = 16dt + 2g + 1.
                const int d_lo = FP8 ? 16 * dt + 2 * g : 16 * dt + g;
                const int d_hi = FP8 ? d_lo + 1 : d_lo + 8;
                *reinterpret_cast<float2*>(sc_o + d_lo * C::HEADS + nt * 8 + 2 * t4) =
                    make_float2(o[nt][dt][0], o[nt][dt][1]);
                *reinterpret_cast<float2*>(sc_o + d_hi * C::HEADS + nt * 8 + 2 * t4) =
                    make_float2(o[nt][dt][2], o[nt][dt][3]);
            }
            if (g == 0) {
                sc_m[nt * 8 + 2 * t4] = m_run[nt][0];
                sc_m[nt * 8 + 2 * t4 + 1] = m_run[nt][1];
                sc_l[nt * 8 + 2 * t4] = l_run[nt][0];
                sc_l[nt * 8 + 2 * t4 + 1] = l_run[nt][1];
            }
        }
        asm volatile("bar.sync 1, %0;" ::"n"(NCW * 32));

        // ---- merge the four slabs; thread -> (head h, 8*NT... d values), bank-conflict-free reads
        float vs = 1.f;
        if (FP8) vs = *p.v_scale;
        const float* sc_all = sc;
        const float* sm_all = sc + NCW * (D * C::HEADS);
        const float* sl_all = sm_all + NCW * C::HEADS;
        constexpr int DG = (NCW * 32) / C::HEADS;        // d-groups covered per pass
        const int h = tid % C::HEADS;
        const int dg = tid / C::HEADS;
        if (h < nheads) {
            float mw[NCW], M = -INFINITY;
#pragma unroll
            for (int w = 0; w < NCW; ++w) {
                mw[w] = sm_all[w * C::HEADS + h];
                M = fmaxf(M, mw[w]);
            }
            float wgt[NCW], L = 0.f;
#pragma unroll
            for (int w = 0; w < NCW; ++w) {
                wgt[w] = (mw[w] == -INFINITY) ? 0.f : b2::fast_exp2((mw[w] - M) * sl2);
                L = fmaf(wgt[w], sl_all[w * C::HEADS + h], L);
            }
            const float inv = vs / L;   // L > 0: the split holds at least one valid token
            const int hq = head0 + h;
            if (p.num_splits == 1) {
                __nv_bfloat16* orow = p.out + b * p.out_stride + (int64_t)hq * D;
                for (int d = dg; d < D; d += DG) {
                    float acc = 0.f;
#pragma unroll
                    for (int w = 0; w < NCW; ++w) acc = fmaf(wgt[w], sc_all[w * (D * C::HEADS) + d * C::HEADS + h], acc);
                    orow[d] = __float2bfloat16(acc * inv);
                }
            } else {
                const int64_t pi = ((int64_t)b * p.Hq + hq) * p.num_splits + split;
                float* orow = p.ws_o + pi * D;
                for (int d = dg; d < D; d += DG) {
                    float acc = 0.f;
#pragma unroll
                    for (int w = 0; w < NCW; ++w) acc = fmaf(wgt[w], sc_all[w * (D * C::HEADS) + d * C::HEADS + h], acc);
                    orow[d] = acc * inv;
                }
                if (dg == 0) p.ws_lse[pi] = fmaf(M, sl2, log2f(L));
            }
        }
    }
}

// a7: out = sum_c 2^(lse_c - max) o_c / sum_c 2^(lse_c - max) over the non-empty splits.
__global__ void __launch_bounds__(D)
merge_splits_kernel(__nv_bfloat16* __restrict__ out, const float* __restrict__ ws_o,
                    const float* __restrict__ ws_lse, const int32_t* __restrict__ seq_lens,
                    int Hq, int num_splits, int split_tokens, int64_t out_stride) {
    const int hq = blockIdx.x, b = blockIdx.y, d = threadIdx.x;
    const int seq_len = seq_lens[b];
    int nsp = (seq_len + split_tokens - 1) / split_tokens;
    nsp = nsp < num_splits ? nsp : num_splits;
    __nv_bfloat16* o = out + b * out_stride + (int64_t)hq * D + d;
    if (nsp <= 0) {
        *o = __float2bfloat16(0.f);
        return;
    }
    const int64_t pi = ((int64_t)b * Hq + hq) * num_splits;
    float mx = -INFINITY;
    for (int c = 0; c < nsp; ++c) mx = fmaxf(mx, ws_lse[pi + c]);
    float num = 0.f, den = 0.f;
    for (int c = 0; c < nsp; ++c) {
        const float w = exp2f(ws_lse[pi + c] - mx);
        num = fmaf(w, ws_o[(pi + c) * D + d], num);
        den += w;
    }
    *o = __float2bfloat16(num / den);
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode_tiled() {
    static EncodeTiledFn fn = []() -> EncodeTiledFn {
        void* sym = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &qres) != cudaSuccess ||
            qres != cudaDriverEntryPointSuccess)
            return nullptr;
        return reinterpret_cast<EncodeTiledFn>(sym);
    }();
    return fn;
}

// cache viewed as [num_blocks][Hkv][bs][128 elements]; one box = bs rows x 128 bytes
int make_cache_tmap(CUtensorMap* tm, const void* cache, int kv_dtype, int64_t num_blocks, int Hkv,
                    int bs, int64_t block_stride, int64_t head_stride) {
    EncodeTiledFn enc = get_encode_tiled();
    B2_REQUIRE(enc != nullptr, B2_ECUDA, "cuTensorMapEncodeTiled entry point not available");
    const int esz = kv_dtype == B2_KV_FP8_E4M3 ? 1 : 2;
    cuuint64_t dims[4] = {(cuuint64_t)D, (cuuint64_t)bs, (cuuint64_t)Hkv, (cuuint64_t)num_blocks};
    cuuint64_t strides[3] = {(cuuint64_t)D * esz, (cuuint64_t)head_stride * esz, (cuuint64_t)block_stride * esz};
    cuuint32_t box[4] = {(cuuint32_t)(128 / esz), (cuuint32_t)bs, 1, 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    const CUresult r = enc(tm, esz == 1 ? CU_TENSOR_MAP_DATA_TYPE_UINT8 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4,
                           const_cast<void*>(cache), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                           CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    B2_REQUIRE(r == CUDA_SUCCESS, B2_ECUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
    return B2_OK;
}

constexpr int kNumSMs = 148;

struct Plan {
    int nt;            // n-tiles of 8 q heads per CTA
    int hg;            // head groups per kv head
    int num_splits;
    int split_tokens;
};

int heads_per_cta(int G) { return G <= 8 ? 8 : 16; }

// Split heuristic: enough CTAs for >= 8 waves over 148 SMs x 2 resident CTAs so the tail wave is
// small, but never splits shorter than 256 tokens (merge traffic stays < 2 % of KV bytes).
int auto_splits(int B, int Hkv, int G, int bs, int max_nb) {
    const int hpc = heads_per_cta(G);
    const int64_t items = (int64_t)B * Hkv * ((G + hpc - 1) / hpc);
    const int64_t max_ctx = (int64_t)max_nb * bs;
    const int64_t target = 8LL * kNumSMs * 2;
    int64_t want = (target + items - 1) / items;
    int64_t cap = max_ctx / 256;
    if (cap < 1) cap = 1;
    if (want > cap) want = cap;
    if (want < 1) want = 1;
    if (want > 64) want = 64;
    return (int)want;
}

int make_plan(Plan* pl, int B, int Hkv, int G, int bs, int max_nb, int num_splits) {
    pl->nt = G <= 8 ? 1 : 2;
    const int hpc = 8 * pl->nt;
    pl->hg = (G + hpc - 1) / hpc;
    if (num_splits <= 0) num_splits = auto_splits(B, Hkv, G, bs, max_nb);
    const int64_t max_ctx = (int64_t)max_nb * bs;
    int64_t st = (max_ctx + num_splits - 1) / num_splits;
    st = ((st + TOK - 1) / TOK) * TOK;   // whole stages, hence whole blocks
    pl->split_tokens = (int)st;
    pl->num_splits = (int)((max_ctx + st - 1) / st);
    if (pl->num_splits < 1) pl->num_splits = 1;
    return B2_OK;
}

template <int BS, bool FP8, int NT>
int launch_decode(const CUtensorMap& tmK, const CUtensorMap& tmV, const DecodeParams& p, int B, cudaStream_t st) {
    constexpr int NSTAGE = FP8 ? 4 : 3;
    constexpr int MINB = 2;
    using C = Cfg<BS, FP8, NT, NSTAGE>;
    auto kern = paged_decode_kernel<BS, FP8, NT, NSTAGE, MINB>;
    static bool attr_set = false;   // per instantiation; idempotent, so a race is harmless
    if (!attr_set) {
        B2_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM_BYTES));
        attr_set = true;
    }
    dim3 grid(p.num_splits, p.Hkv * p.HG, B);
    kern<<<grid, NTHREADS, C::SMEM_BYTES, st>>>(tmK, tmV, p);
    B2_CUDA_OK(cudaGetLastError());
    return B2_OK;
}

template <int BS>
int dispatch_decode(const CUtensorMap& tmK, const CUtensorMap& tmV, const DecodeParams& p, int B, int nt,
                    bool fp8, cudaStream_t st) {
    if (!fp8) return nt == 1 ? launch_decode<BS, false, 1>(tmK, tmV, p, B, st) : launch_decode<BS, false, 2>(tmK, tmV, p, B, st);
    return nt == 1 ? launch_decode<BS, true, 1>(tmK, tmV, p, B, st) : launch_decode<BS, true, 2>(tmK, tmV, p, B, st);
}

}  // namespace

extern "C" int b2_paged_decode_auto_splits(int B, int Hq, int Hkv, int d, int bs, int max_nb, int kv_dtype) {
    (void)d; (void)kv_dtype;
    if (B <= 0 || Hq <= 0 || Hkv <= 0 || Hq % Hkv != 0 || bs <= 0 || max_nb <= 0) return B2_EINVAL;
    Plan pl;
    make_plan(&pl, B, Hkv, Hq / Hkv, bs, max_nb, 0);
    return pl.num_splits;
}

extern "C" size_t b2_paged_decode_workspace_bytes(int B, int Hq, int Hkv, int d, int bs, int max_nb,
                                                  int num_splits) {
    if (B <= 0 || Hq <= 0 || Hkv <= 0 || Hq % Hkv != 0 || bs <= 0 || max_nb <= 0 || d != D) return 0;
    Plan pl;
    make_plan(&pl, B, Hkv, Hq / Hkv, bs, max_nb, num_splits);
    if (pl.num_splits == 1) return 0;
    return (size_t)B * Hq * pl.num_splits * (D + 1) * sizeof(float);
}

extern "C" int b2_paged_decode(void* out, const void* q, const void* k_cache, const void* v_cache,
                               const int32_t* block_tables, const int32_t* seq_lens, int B, int Hq,
                               int Hkv, int d, int bs, int max_nb, int64_t num_blocks,
                               int64_t q_stride, int64_t out_stride, int64_t cache_block_stride,
                               int64_t cache_head_stride, float scale, int kv_dtype,
                               const float* k_scale, const float* v_scale, void* workspace,
                               size_t workspace_bytes, int num_splits, void* stream) {
    B2_REQUIRE(B >= 0 && Hq > 0 && Hkv > 0 && bs > 0 && max_nb > 0 && num_blocks > 0, B2_EINVAL,
               "paged_decode: bad sizes B=%d Hq=%d Hkv=%d bs=%d max_nb=%d num_blocks=%lld", B, Hq, Hkv, bs,
               max_nb, (long long)num_blocks);
    if (B == 0) return B2_OK;
    B2_REQUIRE(out && q && k_cache && v_cache && block_tables && seq_lens, B2_EINVAL, "paged_decode: null pointer");
    B2_REQUIRE(d == D, B2_EUNSUPPORTED, "paged_decode: head_dim %d unsupported (kernels are built for 128)", d);
    B2_REQUIRE(Hq % Hkv == 0, B2_EINVAL, "paged_decode: Hq=%d not a multiple of Hkv=%d", Hq, Hkv);
    B2_REQUIRE(bs == 8 || bs == 16 || bs == 32, B2_EUNSUPPORTED, "paged_decode: block_size %d unsupported (8/16/32)", bs);
    B2_REQUIRE(kv_dtype == B2_KV_BF16 || kv_dtype == B2_KV_FP8_E4M3, B2_EUNSUPPORTED, "paged_decode: kv_dtype %d", kv_dtype);
    B2_REQUIRE(b2_aligned16(out) && b2_aligned16(q) && b2_aligned16(k_cache) && b2_aligned16(v_cache), B2_EINVAL,
               "paged_decode: pointers must be 16-byte aligned");
    B2_REQUIRE(q_stride % 8 == 0 && out_stride % 8 == 0 && q_stride >= (int64_t)Hq * D && out_stride >= (int64_t)Hq * D,
               B2_EINVAL, "paged_decode: q/out strides");
    B2_REQUIRE(cache_block_stride % 16 == 0 && cache_head_stride % 16 == 0 && cache_head_stride >= (int64_t)bs * D,
               B2_EINVAL, "paged_decode: cache strides must keep 16-byte alignment");
    B2_REQUIRE(scale > 0.f, B2_EINVAL, "paged_decode: scale must be positive");
    if (kv_dtype == B2_KV_FP8_E4M3) B2_REQUIRE(k_scale && v_scale, B2_EINVAL, "paged_decode: fp8 needs k_scale/v_scale");
    B2_REQUIRE((int64_t)Hkv * ((Hq / Hkv + 15) / 16) <= 65535 && B <= 65535, B2_EUNSUPPORTED, "paged_decode: grid too large");

    const int G = Hq / Hkv;
    Plan pl;
    make_plan(&pl, B, Hkv, G, bs, max_nb, num_splits);

    DecodeParams p;
    p.out = static_cast<__nv_bfloat16*>(out);
    p.q = static_cast<const __nv_bfloat16*>(q);
    p.block_tables = block_tables;
    p.seq_lens = seq_lens;
    p.k_scale = k_scale;
    p.v_scale = v_scale;
    p.Hq = Hq; p.Hkv = Hkv; p.G = G; p.HG = pl.hg; p.max_nb = max_nb;
    p.num_splits = pl.num_splits; p.split_tokens = pl.split_tokens;
    p.q_stride = q_stride; p.out_stride = out_stride;
    p.scale_log2 = scale * LOG2E;
    p.ws_o = nullptr; p.ws_lse = nullptr;
    if (pl.num_splits > 1) {
        const size_t need = (size_t)B * Hq * pl.num_splits * (D + 1) * sizeof(float);
        B2_REQUIRE(workspace && workspace_bytes >= need && b2_aligned16(workspace), B2_ENOMEM,
                   "paged_decode: workspace %zu bytes < required %zu", workspace_bytes, need);
        p.ws_o = static_cast<float*>(workspace);
        p.ws_lse = p.ws_o + (size_t)B * Hq * pl.num_splits * D;
    }

    CUtensorMap tmK, tmV;
    int rc = make_cache_tmap(&tmK, k_cache, kv_dtype, num_blocks, Hkv, bs, cache_block_stride, cache_head_stride);
    if (rc != B2_OK) return rc;
    rc = make_cache_tmap(&tmV, v_cache, kv_dtype, num_blocks, Hkv, bs, cache_block_stride, cache_head_stride);
    if (rc != B2_OK) return rc;

    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const bool fp8 = kv_dtype == B2_KV_FP8_E4M3;
    if (bs == 8) rc = dispatch_decode<8>(tmK, tmV, p, B, pl.nt, fp8, st);
    else if (bs == 16) rc = dispatch_decode<16>(tmK, tmV, p, B, pl.nt, fp8, st);
    else rc = dispatch_decode<32>(tmK, tmV, p, B, pl.nt, fp8, st);
    if (rc != B2_OK) return rc;

    if (pl.num_splits > 1) {
        merge_splits_kernel<<<dim3(Hq, B), D, 0, st>>>(p.out, p.ws_o, p.ws_lse, seq_lens, Hq, pl.num_splits,
                                                       pl.split_tokens, out_stride);
        B2_CUDA_OK(cudaGetLastError());
    }
    return B2_OK;
}
