// paged_decode.cu -- paged-attention decode for B200 (sm_100a).
// SURVEY.md section 8(a) rows a6 (single pass), a7 (split-K + merge), a8 (fp8 dequant in load).
//
// Contract: include/b200attn.h `b2_paged_decode`.  Semantics follow vLLM paged_attention_v1/v2,
// the op run each decode step by the engine the reference launches
// (reference benchmark-vllm/vllm_manager.py:371-374); block_size / max_num_seqs / max_model_len
// come from reference benchmark-vllm/models.py:9,10,15.
//
// Design (HBM-bound gather):
//   * One CTA serves all G = Hq/Hkv query heads that share a kv head, so KV bytes are read once
//     per kv head, not once per q head.  An "item" is one (sequence, kv head [, head group]).
//   * warp 4 is the producer: it walks the block table and issues ONE TMA box per (block, head)
//     tile and tensor (4 KiB bf16 / 2 KiB fp8, contiguous in HBM; 128-byte swizzle, L2 evict-first)
//     into a ring of NSTAGE x 64-token stages; completion is counted in bytes on an mbarrier.
//   * warps 0-3 are consumers: each owns 16 tokens of every stage.  S^T = K.Q^T and
//     O^T += V^T.P^T run on mma.sync m16n8k16 (tokens / head_dim on the M side, the <= 8 grouped
//     query heads on the N side, so no MMA row is wasted on padding tokens); ldmatrix(.trans)
//     feeds the A operand straight from the swizzled stage; softmax is fp32 with warp shuffles;
//     the P^T B-operand is produced from the S^T accumulator by movmatrix (bf16 path: hi + lo).
//   * schedule A (num_splits >= 1): the classic grid, one CTA per (split, item); 1 = single pass
//     ("v1"), > 1 = split-K with an LSE merge kernel ("v2").
//   * schedule B (num_splits = 0): persistent CTAs (2 per SM) over a fixed grid of chunks: a static
//     contiguous share per CTA, then chunks claimed from an atomic counter so the grid drains
//     together whatever the batch size or raggedness; partial runs are merged by
//     merge_chunks_kernel (programmatic dependent launch).  Details above that kernel.
//   * b2_paged_decode_plan_splits picks A(1) or B on the host from the scheduler's seq_lens.
#include "b2_common.cuh"

#include <stdlib.h>
#include <algorithm>
#include <mutex>
#include <new>
#include <vector>

namespace {

constexpr int D = 128;        // head_dim (Llama-3 geometry)
constexpr int NCW = 4;        // consumer warps
constexpr int TOK = NCW * 16; // tokens per pipeline stage
constexpr int NTHREADS = (NCW + 1) * 32;
constexpr float LOG2E = 1.4426950408889634f;

// Optional per-CTA timeline (make trace -> lib/libb200attn_trace.so; scripts/trace_schedule.py reads it).
#ifdef B2_TRACE
__device__ unsigned long long g_trace[2048 * 8];
__device__ __forceinline__ unsigned long long gtime() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
#define B2_TRACE_AT(cta, k) do { if ((threadIdx.x & 31) == 0 && (cta) < 2048) g_trace[(cta) * 8 + (k)] = gtime(); } while (0)
#define B2_TRACE_VAL(cta, k, v) do { if ((threadIdx.x & 31) == 0 && (cta) < 2048) g_trace[(cta) * 8 + (k)] = (v); } while (0)
#else
#define B2_TRACE_AT(cta, k) do { } while (0)
#define B2_TRACE_VAL(cta, k, v) do { } while (0)
#endif

struct DecodeParams {
    __nv_bfloat16* out;
    const __nv_bfloat16* q;
    const int32_t* block_tables;
    const int32_t* seq_lens;
    float* ws_o;     // grid-split: [B, Hq, num_splits, D]   persistent: [NC*2, HEADS, D]
    float* ws_lse;   // grid-split: [B, Hq, num_splits]      persistent: [NC*2, HEADS]
    const float* k_scale;
    const float* v_scale;
    int B, Hq, Hkv, G, HG, max_nb, num_splits, split_tokens;
    int chunk_tokens, static_pct;   // dynamic schedule: claim granularity, share of chunks dealt statically
    int two_box;                    // bf16: the cache map holds one 128-byte half per box (driver refused the one-box map)
    uint32_t* counter;              // dynamic schedule: claim counter (workspace header)
    int64_t q_stride, out_stride;
    float scale_log2;  // softmax scale * log2(e)
    // b2_paged_decode_append, fused path (single-pass grid, one CTA per (row, kv head)): the CTA writes its row's new
    // K/V token into the cache before it streams the row.  append = 0: every field below is unused.
    int append;
    const __nv_bfloat16* k_new;
    const __nv_bfloat16* v_new;
    const int64_t* slot_mapping;
    void* k_cache_w;
    void* v_cache_w;
    int64_t k_new_stride, v_new_stride, cache_block_stride, cache_head_stride;
    int32_t* bt_update;   // resident block tables (b2_decode_plan_run_append): also record the new token's block, or null
};

template <int BS, bool FP8, int NT, int NSTAGE>
struct Cfg {
    static constexpr int BSZ = BS;
    static constexpr int ESZ = FP8 ? 1 : 2;
    static constexpr int ROWB = D * ESZ;           // bytes per token row
    static constexpr int TILE_BYTES = TOK * ROWB;  // one K (or V) stage tile
    static constexpr int STAGE_BYTES = 2 * TILE_BYTES;
    static constexpr int BLK_PER_STAGE = TOK / BS;
    static constexpr int BOX_BYTES = BS * ROWB;    // one TMA box = one contiguous (block, head) tile: 4 KiB bf16 / 2 KiB fp8
    static constexpr int HEADS = 8 * NT;           // q heads served per CTA
    static constexpr int RING_BYTES = NSTAGE * STAGE_BYTES;
    static constexpr int BAR_BYTES = 2 * NSTAGE * 8 + 16;
    static constexpr int SLOT_FLOATS = 32 * NT * 32 + 4 * NT * 4;   // O^T fragments + (m, l) of one warp
    static constexpr int SCRATCH_BYTES = 2 * SLOT_FLOATS * 4;
    static constexpr int SMEM_BYTES = RING_BYTES + BAR_BYTES + SCRATCH_BYTES;
    // dynamic schedule: per-stage descriptors + the chunk prefix over (groups of) sequences
    static constexpr int META_BYTES = NSTAGE * 16;
    static constexpr int PMAX = NT == 1 ? 512 : 128;
    static constexpr int PRE_BYTES = (PMAX + 1 + 3) / 4 * 16;
    // dynamic schedule, NT == 1: the run's q rows travel with its first stage (bulk copy, same full-barrier),
    // so a run boundary never waits on a global load.  One q buffer per stage slot.
    static constexpr bool QSMEM = NT == 1;
    static constexpr int QROW_BYTES = D * 2;
    static constexpr int QBUF_BYTES = QSMEM ? HEADS * QROW_BYTES : 0;
    static constexpr int DYN_SMEM_BYTES = SMEM_BYTES + META_BYTES + PRE_BYTES + NSTAGE * QBUF_BYTES;
    // fp8 does twice the tokens (MMA + softmax + e4m3->f16 conversion) per byte: its consumers, not HBM, set the pace
    // with 8 consumer warps per SM, so the shallow-ring variants run three CTAs per SM (12 consumer warps).
    static constexpr int CTAS_SPLIT = (FP8 && 3 * (SMEM_BYTES + 1024) <= 233472) ? 3 : 2;
    static constexpr int CTAS_DYN = (FP8 && 3 * (DYN_SMEM_BYTES + 1024) <= 233472) ? 3 : 2;
    static_assert(TOK % BS == 0 && BS % 8 == 0, "block size must divide the stage and keep the swizzle phase");
    static_assert(RING_BYTES % 1024 == 0, "stages must stay 1024-byte aligned");
    static_assert(2 * (SMEM_BYTES + 1024) <= 233472, "two CTAs per SM must fit");
};

// fp8 (e4m3 x2, low 16 bits of `v`) -> f16x2, exact
__device__ __forceinline__ uint32_t e4m3x2_to_f16x2(uint32_t v) {
    uint32_t r;
    asm("{\n\t.reg .b16 lo;\n\tcvt.u16.u32 lo, %1;\n\tcvt.rn.f16x2.e4m3x2 %0, lo;\n\t}" : "=r"(r) : "r"(v));
    return r;
}
// same for the high 16 bits of `v` (SASS: F2FP...UNPACK_B Rx.H1 -- no shift instruction)
__device__ __forceinline__ uint32_t e4m3x2_hi_to_f16x2(uint32_t v) {
    uint32_t r;
    asm("{\n\t.reg .b16 lo, hi;\n\tmov.b32 {lo, hi}, %1;\n\tcvt.rn.f16x2.e4m3x2 %0, hi;\n\t}" : "=r"(r) : "r"(v));
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
__device__ __forceinline__ void consumer_bar() { asm volatile("bar.sync 1, %0;" ::"n"(NCW * 32) : "memory"); }

// One KV block of one kv head -> stage slot j.  The (block, head) tile is bs x d contiguous in HBM.  fp8 rows are one
// 128-byte swizzle line (4-D map, one box).  bf16 rows are 256 B = two lines: the tile is viewed as
// [half][row][64 elements] (5-D map whose row stride is larger than its half stride), so ONE box lands as bs lines of
// the low halves followed by bs lines of the high halves -- the 8 rows of an ldmatrix then sit on 8 consecutive lines
// and the 128-byte swizzle makes them conflict-free (r1 landed [row][half]: lines 2r + h, a 2-way conflict on every
// K and V fragment load).  two_box: same shared-memory image from two boxes of one half each (fallback map).
template <class C, bool FP8>
__device__ __forceinline__ void issue_block(const CUtensorMap* tmK, const CUtensorMap* tmV, uint32_t kdst, uint32_t vdst,
                                            int j, uint32_t bar, int kvh, int blk, uint64_t pol, int two_box) {
    const uint32_t off = j * C::BOX_BYTES;
    if (FP8) {
        b2::tma_load_4d(kdst + off, tmK, bar, 0, 0, kvh, blk, pol);
        b2::tma_load_4d(vdst + off, tmV, bar, 0, 0, kvh, blk, pol);
    } else if (!two_box) {
        b2::tma_load_5d(kdst + off, tmK, bar, 0, 0, 0, kvh, blk, pol);
        b2::tma_load_5d(vdst + off, tmV, bar, 0, 0, 0, kvh, blk, pol);
    } else {
        b2::tma_load_5d(kdst + off, tmK, bar, 0, 0, 0, kvh, blk, pol);
        b2::tma_load_5d(kdst + off + C::BOX_BYTES / 2, tmK, bar, 0, 1, 0, kvh, blk, pol);
        b2::tma_load_5d(vdst + off, tmV, bar, 0, 0, 0, kvh, blk, pol);
        b2::tma_load_5d(vdst + off + C::BOX_BYTES / 2, tmV, bar, 0, 1, 0, kvh, blk, pol);
    }
}
// shared-memory address of 16-byte chunk `chunk` (0..15 bf16, 0..7 fp8) of token row `row` inside a K or V stage tile
template <bool FP8, int BS>
__device__ __forceinline__ uint32_t tile_addr(uint32_t base, int row, int chunk) {
    // 128-byte swizzle line: fp8 = the row; bf16 = [block][half][row in block]
    const int line = FP8 ? row : (row / BS) * (2 * BS) + (chunk >> 3) * BS + (row % BS);
    return base + line * 128 + (((chunk & 7) ^ (line & 7)) << 4);
}

// ---------------------------------------------------------------------------------------------
// producer: stream the blocks [blk_begin, blk_begin + n_blk) of one (sequence, kv head) into the ring
// ---------------------------------------------------------------------------------------------
template <class C, int NSTAGE>
__device__ __forceinline__ void produce_segment(const CUtensorMap* tmK, const CUtensorMap* tmV,
                                                const int32_t* __restrict__ bt, int n_blk, int kvh,
                                                uint32_t ring, uint32_t bar_full, uint32_t bar_empty,
                                                uint32_t& it, int lane, uint64_t pol, int two_box,
                                                int last_blk = -1) {
    const int n_stage = (n_blk + C::BLK_PER_STAGE - 1) / C::BLK_PER_STAGE;
    int cur = (lane < n_blk) ? bt[lane] : 0;            // 32 table entries per batch
    // fused append over resident tables: the id of the segment's last block comes from the new token's slot, not from
    // the table entry this very grid is writing (only used with n_blk <= 32)
    if (last_blk >= 0 && lane == n_blk - 1) cur = last_blk;
    int nxt = (32 + lane < n_blk) ? bt[32 + lane] : 0;  // next batch, prefetched
    for (int st = 0; st < n_stage; ++st, ++it) {
        const uint32_t s = it % NSTAGE;
        const uint32_t ph = (it / NSTAGE) & 1;
        b2::mbar_wait(bar_empty + s * 8, ph ^ 1);
        const int nb_stage = min(C::BLK_PER_STAGE, n_blk - st * C::BLK_PER_STAGE);
        const uint32_t kdst = ring + s * C::STAGE_BYTES;
        const uint32_t vdst = kdst + C::TILE_BYTES;
        if (lane == 0)
            b2::mbar_arrive_expect_tx(bar_full + s * 8, nb_stage * 2 * C::BOX_BYTES);
#pragma unroll
        for (int j = 0; j < C::BLK_PER_STAGE; ++j) {
            const int gi = st * C::BLK_PER_STAGE + j;
            const int src = gi & 31;
            const int blk = __shfl_sync(0xffffffffu, cur, src);
            if (j < nb_stage && lane == 0)
                issue_block<C, C::ESZ == 1>(tmK, tmV, kdst, vdst, j, bar_full + s * 8, kvh, blk, pol, two_box);
            if (src == 31) {  // warp-uniform: this batch of 32 entries is used up
                cur = nxt;
                nxt = (gi + 33 + lane < n_blk) ? bt[gi + 33 + lane] : 0;
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------
// consumer state: one warp's running softmax over its token slabs
// ---------------------------------------------------------------------------------------------
template <int NT>
struct WarpAcc {
    float o[NT][8][4];     // O^T fragments: [n-tile][d-tile][c0..c3]
    float m[NT][2];        // running max (raw scores) of heads 2*t4, 2*t4+1
    float l[NT][2];        // running sum (lane-local until reduce_l)
    __device__ __forceinline__ void reset() {
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
            m[nt][0] = m[nt][1] = -INFINITY;
            l[nt][0] = l[nt][1] = 0.f;
#pragma unroll
            for (int dt = 0; dt < 8; ++dt) o[nt][dt][0] = o[nt][dt][1] = o[nt][dt][2] = o[nt][dt][3] = 0.f;
        }
    }
    // lane-local row sums -> sums over the 16 token rows (lanes differing in g)
    __device__ __forceinline__ void reduce_l() {
#pragma unroll
        for (int nt = 0; nt < NT; ++nt)
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                float v = l[nt][j];
                v += __shfl_xor_sync(0xffffffffu, v, 4);
                v += __shfl_xor_sync(0xffffffffu, v, 8);
                v += __shfl_xor_sync(0xffffffffu, v, 16);
                l[nt][j] = v;
            }
    }
    // fragment-layout dump / combine: partner lanes hold the same (d, head) positions
    __device__ __forceinline__ void dump(float* slot, int lane) const {
#pragma unroll
        for (int nt = 0; nt < NT; ++nt)
#pragma unroll
            for (int dt = 0; dt < 8; ++dt)
#pragma unroll
                for (int i = 0; i < 4; ++i) slot[((nt * 8 + dt) * 4 + i) * 32 + lane] = o[nt][dt][i];
        if (lane < 4) {
            float* ml = slot + 32 * NT * 32 + lane * (4 * NT);
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) {
                ml[nt * 4 + 0] = m[nt][0];
                ml[nt * 4 + 1] = m[nt][1];
                ml[nt * 4 + 2] = l[nt][0];
                ml[nt * 4 + 3] = l[nt][1];
            }
        }
    }
    __device__ __forceinline__ void combine(const float* slot, int lane, float sl2) {
        const float* ml = slot + 32 * NT * 32 + (lane & 3) * (4 * NT);
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
            float wa[2], wb[2];
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                const float mb = ml[nt * 4 + j], lb = ml[nt * 4 + 2 + j];
                const float M = fmaxf(m[nt][j], mb);
                wa[j] = (m[nt][j] == -INFINITY) ? 0.f : b2::fast_exp2((m[nt][j] - M) * sl2);
                wb[j] = (mb == -INFINITY) ? 0.f : b2::fast_exp2((mb - M) * sl2);
                l[nt][j] = wa[j] * l[nt][j] + wb[j] * lb;
                m[nt][j] = M;
            }
#pragma unroll
            for (int dt = 0; dt < 8; ++dt)
#pragma unroll
                for (int i = 0; i < 4; ++i)
                    o[nt][dt][i] = wa[i & 1] * o[nt][dt][i] + wb[i & 1] * slot[((nt * 8 + dt) * 4 + i) * 32 + lane];
        }
    }
};

// Q^T as the B operand of S^T = K.Q^T: r[nt][kc] = {Q[head g][k-pair], the +8 pair}.
// bf16: logical k == physical d.  fp8: ldmatrix hands each thread 4 consecutive fp8 of a 16-element
// chunk, so logical k = 2*t4+{0,1} <-> d = 16kc + 4*t4 + {0,1} and the +8 pair <-> d = 16kc + 4*t4 + {2,3};
// the contraction is permutation invariant.
// The fp8 path runs the f16 MMA (K and V convert exactly), so q becomes f16: each q row is first scaled by a power
// of two that puts its largest |element| into [128, 256) -- exact for bf16 values down to 2^-21 of the row maximum,
// whatever the magnitude of q.  hs[nt][j] = the inverse factor of head 2t4 + j, applied to this thread's score columns.
template <int NT>
struct QFrags {
    uint32_t r[NT][8][2];
    float hs[NT][2];
};

// qbase: [head][D/2] words in global or shared memory; rows >= nheads are never read.
template <bool FP8, int NT>
__device__ __forceinline__ void load_q_frags(QFrags<NT>& qf, const uint32_t* qbase, int nheads, int lane) {
    const int g = lane >> 2, t4 = lane & 3;
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
        const int hl = nt * 8 + g;
        const bool valid = hl < nheads;
        const uint32_t* qrow = qbase + (valid ? hl : 0) * (D / 2);
        if (!FP8) {
#pragma unroll
            for (int kc = 0; kc < 8; ++kc) {
                const uint32_t w0 = qrow[kc * 8 + t4], w1 = qrow[kc * 8 + 4 + t4];
                qf.r[nt][kc][0] = valid ? w0 : 0u;
                qf.r[nt][kc][1] = valid ? w1 : 0u;
            }
            qf.hs[nt][0] = qf.hs[nt][1] = 1.f;
        } else {
            uint2 w[8];
            float amax = 0.f;
#pragma unroll
            for (int kc = 0; kc < 8; ++kc) {
                w[kc] = *reinterpret_cast<const uint2*>(qrow + kc * 8 + 2 * t4);
                if (!valid) w[kc] = make_uint2(0u, 0u);
                amax = fmaxf(amax, fmaxf(fabsf(__uint_as_float(w[kc].x << 16)), fabsf(__uint_as_float(w[kc].x & 0xffff0000u))));
                amax = fmaxf(amax, fmaxf(fabsf(__uint_as_float(w[kc].y << 16)), fabsf(__uint_as_float(w[kc].y & 0xffff0000u))));
            }
            amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 1));   // the four lanes of one g hold one head
            amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 2));
            // exponent field clamped so that both factors are normal floats
            uint32_t ef = (__float_as_uint(amax) >> 23) & 0xffu;
            ef = ef < 16u ? 16u : (ef > 240u ? 240u : ef);
            const float up = __uint_as_float((261u - ef) << 23);     // 2^(7 - (ef - 127))
            const float down = __uint_as_float((ef - 7u) << 23);     // 2^((ef - 127) - 7)
#pragma unroll
            for (int kc = 0; kc < 8; ++kc) {
                qf.r[nt][kc][0] = pack_f16x2(__uint_as_float(w[kc].x << 16) * up, __uint_as_float(w[kc].x & 0xffff0000u) * up);
                qf.r[nt][kc][1] = pack_f16x2(__uint_as_float(w[kc].y << 16) * up, __uint_as_float(w[kc].y & 0xffff0000u) * up);
            }
            qf.hs[nt][0] = __shfl_sync(0xffffffffu, down, 8 * t4);       // lane 4 * (2t4)     holds head 2t4
            qf.hs[nt][1] = __shfl_sync(0xffffffffu, down, 8 * t4 + 4);   // lane 4 * (2t4 + 1) holds head 2t4 + 1
        }
    }
}

// One warp, one stage: 16 token rows [row_base, row_base+16) of which the first `nvalid` are real.
template <class C, bool FP8, int NT>
__device__ __forceinline__ void consume_slab(WarpAcc<NT>& acc, const QFrags<NT>& qf, uint32_t kbase,
                                             uint32_t vbase, int row_base, int nvalid, float sl2, int lane) {
    const int g = lane >> 2, t4 = lane & 3;
    const int mi = lane >> 3, ri = lane & 7;   // ldmatrix: lane -> (matrix, row)

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
            const uint32_t addr = tile_addr<false, C::BSZ>(kbase, row, chunk);
            uint32_t a0, a1, a2, a3;
            b2::ldmatrix_x4(addr, a0, a1, a2, a3);
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) mma_kv<FP8>(sacc[nt], a0, a1, a2, a3, qf.r[nt][kc][0], qf.r[nt][kc][1]);
        }
    } else {
        // fp8 rows are 128 bytes = 8 chunks of 16 elements; one ldmatrix.x4 covers two k-chunks:
        // matrices: 0 = rows 0-7 / chunk c, 1 = rows 8-15 / c, 2 = rows 0-7 / c+1, 3 = rows 8-15 / c+1
        const int row = row_base + (mi & 1) * 8 + ri;
#pragma unroll
        for (int kc2 = 0; kc2 < 4; ++kc2) {
            const int chunk = 2 * kc2 + (mi >> 1);
            const uint32_t addr = tile_addr<true, C::BSZ>(kbase, row, chunk);
            uint32_t r0, r1, r2, r3;
            b2::ldmatrix_x4(addr, r0, r1, r2, r3);
            const uint32_t a0 = e4m3x2_to_f16x2(r0), a1 = e4m3x2_to_f16x2(r1);
            const uint32_t a2 = e4m3x2_hi_to_f16x2(r0), a3 = e4m3x2_hi_to_f16x2(r1);
            const uint32_t c0 = e4m3x2_to_f16x2(r2), c1 = e4m3x2_to_f16x2(r3);
            const uint32_t c2 = e4m3x2_hi_to_f16x2(r2), c3 = e4m3x2_hi_to_f16x2(r3);
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) {
                mma_kv<FP8>(sacc[nt], a0, a1, a2, a3, qf.r[nt][2 * kc2][0], qf.r[nt][2 * kc2][1]);
                mma_kv<FP8>(sacc[nt], c0, c1, c2, c3, qf.r[nt][2 * kc2 + 1][0], qf.r[nt][2 * kc2 + 1][1]);
            }
        }
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {   // back to raw score units (q rows were scaled per head)
            sacc[nt][0] *= qf.hs[nt][0]; sacc[nt][1] *= qf.hs[nt][1];
            sacc[nt][2] *= qf.hs[nt][0]; sacc[nt][3] *= qf.hs[nt][1];
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
    // bf16 carries 8 significant bits: a P rounded once to bf16 puts up to 2^-9 * |V| of error on an output built
    // from a handful of tokens, which is outside the 2e-3 bar.  The bf16 path therefore feeds P as hi + lo
    // (lo = bf16(P - hi)), two MMAs on the same V fragments; the f16 P of the fp8 path (11 bits) is rounded once.
    constexpr bool PSPLIT = !FP8;
    uint32_t pb[NT][2];
    uint32_t pl[PSPLIT ? NT : 1][2];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
        float pr[4];
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            float mx = fmaxf(sacc[nt][j], sacc[nt][j + 2]);
            mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 4));
            mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 8));
            mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 16));
            const float m_new = fmaxf(acc.m[nt][j], mx);       // finite: the slab has a valid token
            const float alpha = b2::fast_exp2((acc.m[nt][j] - m_new) * sl2);
            const float ms = m_new * sl2;
            pr[j] = b2::fast_exp2(fmaf(sacc[nt][j], sl2, -ms));
            pr[j + 2] = b2::fast_exp2(fmaf(sacc[nt][j + 2], sl2, -ms));
            acc.m[nt][j] = m_new;
            acc.l[nt][j] = fmaf(acc.l[nt][j], alpha, pr[j] + pr[j + 2]);
#pragma unroll
            for (int dt = 0; dt < 8; ++dt) {
                acc.o[nt][dt][j] *= alpha;
                acc.o[nt][dt][j + 2] *= alpha;
            }
        }
        // P^T as the B operand of O^T += V^T.P^T: transpose the 8x8 (token x head) tiles
        const uint32_t x0 = FP8 ? pack_f16x2(pr[0], pr[1]) : b2::pack_bf16x2(pr[0], pr[1]);
        const uint32_t x1 = FP8 ? pack_f16x2(pr[2], pr[3]) : b2::pack_bf16x2(pr[2], pr[3]);
        pb[nt][0] = b2::movmatrix_trans(x0);   // (tokens 2t4,2t4+1 ; head g)
        pb[nt][1] = b2::movmatrix_trans(x1);   // (tokens 8+2t4,.. ; head g)
        if (PSPLIT) {
            const uint32_t y0 = b2::pack_bf16x2(pr[0] - __uint_as_float(x0 << 16), pr[1] - __uint_as_float(x0 & 0xffff0000u));
            const uint32_t y1 = b2::pack_bf16x2(pr[2] - __uint_as_float(x1 << 16), pr[3] - __uint_as_float(x1 & 0xffff0000u));
            pl[nt][0] = b2::movmatrix_trans(y0);
            pl[nt][1] = b2::movmatrix_trans(y1);
        }
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
            const uint32_t addr = tile_addr<false, C::BSZ>(vbase, row, chunk);
            uint32_t a0, a1, a2, a3;
            b2::ldmatrix_x4_trans(addr, a0, a1, a2, a3);
            a0 &= mk_lo; a1 &= mk_lo; a2 &= mk_hi; a3 &= mk_hi;
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) {
                mma_kv<FP8>(acc.o[nt][dt], a0, a1, a2, a3, pb[nt][0], pb[nt][1]);
                if (PSPLIT) mma_kv<FP8>(acc.o[nt][dt], a0, a1, a2, a3, pl[nt][0], pl[nt][1]);
            }
        }
    } else {
        // A 16-byte fp8 chunk holds 16 d values, so the b16 transpose moves (d 2g, d 2g+1) pairs: a thread
        // receives {tok 2t4: d 2g, d 2g+1 ; tok 2t4+1: d 2g, d 2g+1}.  PRMT regroups it to {d 2g: tok pair ;
        // d 2g+1: tok pair}; MMA row g <-> d = 16c + 2g, row g+8 <-> d = 16c + 2g + 1.
        // .trans matrices: 0 = tokens 0-7 / chunk c, 1 = tokens 8-15 / c, 2 = tokens 0-7 / c+1, 3 = tokens 8-15 / c+1
        const int row = row_base + (mi & 1) * 8 + ri;
#pragma unroll
        for (int dt2 = 0; dt2 < 4; ++dt2) {
            const int chunk = 2 * dt2 + (mi >> 1);
            const uint32_t addr = tile_addr<true, C::BSZ>(vbase, row, chunk);
            uint32_t r0, r1, r2, r3;
            b2::ldmatrix_x4_trans(addr, r0, r1, r2, r3);
            r0 = __byte_perm(r0, 0, 0x3120); r1 = __byte_perm(r1, 0, 0x3120);
            r2 = __byte_perm(r2, 0, 0x3120); r3 = __byte_perm(r3, 0, 0x3120);
            uint32_t a0 = e4m3x2_to_f16x2(r0), a1 = e4m3x2_hi_to_f16x2(r0);
            uint32_t a2 = e4m3x2_to_f16x2(r1), a3 = e4m3x2_hi_to_f16x2(r1);
            uint32_t c0 = e4m3x2_to_f16x2(r2), c1 = e4m3x2_hi_to_f16x2(r2);
            uint32_t c2 = e4m3x2_to_f16x2(r3), c3 = e4m3x2_hi_to_f16x2(r3);
            if (nvalid < 16) {   // warp-uniform; full slabs (all but a sequence's last) skip the eight ANDs
                a0 &= mk_lo; a1 &= mk_lo; a2 &= mk_hi; a3 &= mk_hi;
                c0 &= mk_lo; c1 &= mk_lo; c2 &= mk_hi; c3 &= mk_hi;
            }
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) {
                mma_kv<FP8>(acc.o[nt][2 * dt2], a0, a1, a2, a3, pb[nt][0], pb[nt][1]);
                mma_kv<FP8>(acc.o[nt][2 * dt2 + 1], c0, c1, c2, c3, pb[nt][0], pb[nt][1]);
            }
        }
    }
}

// Consume the stages of one segment (n_tok tokens starting at a block boundary).
template <class C, bool FP8, int NT, int NSTAGE>
__device__ __forceinline__ void consume_segment(WarpAcc<NT>& acc, const QFrags<NT>& qf, int n_tok,
                                                uint32_t ring, uint32_t bar_full, uint32_t bar_empty,
                                                uint32_t& it, float sl2, int warp, int lane) {
    const int n_stage = (n_tok + TOK - 1) / TOK;
    const int row_base = warp * 16;
    for (int st = 0; st < n_stage; ++st, ++it) {
        const uint32_t s = it % NSTAGE;
        const uint32_t ph = (it / NSTAGE) & 1;
        b2::mbar_wait(bar_full + s * 8, ph);
        const int nvalid = n_tok - (st * TOK + row_base);   // valid tokens in this warp's slab
        if (nvalid > 0) {
            const uint32_t kbase = ring + s * C::STAGE_BYTES;
            consume_slab<C, FP8, NT>(acc, qf, kbase, kbase + C::TILE_BYTES, row_base, nvalid, sl2, lane);
        }
        __syncwarp();
        if (lane == 0) b2::mbar_arrive(bar_empty + s * 8);   // stage may be overwritten
    }
}

// Tree merge of the four consumer warps; the result lands in warp 0's accumulator.
template <int NT>
__device__ __forceinline__ void merge_warps(WarpAcc<NT>& acc, float* scratch, int slot_floats, float sl2,
                                            int warp, int lane) {
    acc.reduce_l();
    if (warp >= 2) acc.dump(scratch + (warp - 2) * slot_floats, lane);
    consumer_bar();
    if (warp < 2) acc.combine(scratch + warp * slot_floats, lane, sl2);
    consumer_bar();
    if (warp == 1) acc.dump(scratch, lane);
    consumer_bar();
    if (warp == 0) acc.combine(scratch, lane, sl2);
    consumer_bar();   // scratch is free for the next segment
}

// Warp 0 writes the merged result: final bf16 rows, or a normalised fp32 partial + log2-sum-exp.
// MMA row r of d-tile dt <-> d = 16dt + r (bf16) or rows g / g+8 <-> d = 16dt + 2g / 2g+1 (fp8).
// The fragment layout scatters one head's row over the warp, so the rows are first laid out in shared memory
// (`stage`: the whole merge scratch, free once merge_warps has returned; padded row strides keep the scatter
// conflict-free) and leave as 16-byte vectors, 256 contiguous bytes per head: full sectors in HBM and, when `out`
// is pinned host memory (the engine step's D2H leg), full-size PCIe writes instead of 2-byte ones.
// The caller keeps the other consumer warps off the scratch until warp 0 is done (consumer_bar).
constexpr int OUT_STAGE_STRIDE = D * 2 + 16;    // bytes per staged bf16 row
constexpr int PART_STAGE_STRIDE = D * 4 + 16;   // bytes per staged fp32 row
template <bool FP8, int NT>
__device__ __forceinline__ void write_result(const WarpAcc<NT>& acc, int nheads, float vs, float sl2,
                                             __nv_bfloat16* out_rows /* [head][D] or null */,
                                             float* part_o /* [head][D] */, float* part_lse /* [head] */,
                                             float* stage, int lane) {
    static_assert(8 * NT * PART_STAGE_STRIDE <= 2 * (32 * NT * 32 + 4 * NT * 4) * 4, "staged rows must fit the merge scratch");
    const int g = lane >> 2, t4 = lane & 3;
    uint8_t* sb = reinterpret_cast<uint8_t*>(stage);
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            const int h = nt * 8 + 2 * t4 + j;
            if (h >= nheads) continue;
            const float L = acc.l[nt][j];
            const float inv = L > 0.f ? vs / L : 0.f;
#pragma unroll
            for (int dt = 0; dt < 8; ++dt) {
                const int d_lo = FP8 ? 16 * dt + 2 * g : 16 * dt + g;
                const int d_hi = FP8 ? d_lo + 1 : d_lo + 8;
                const float v_lo = acc.o[nt][dt][j] * inv, v_hi = acc.o[nt][dt][j + 2] * inv;
                if (out_rows) {
                    if (FP8) {
                        *reinterpret_cast<uint32_t*>(sb + h * OUT_STAGE_STRIDE + d_lo * 2) = b2::pack_bf16x2(v_lo, v_hi);
                    } else {
                        *reinterpret_cast<__nv_bfloat16*>(sb + h * OUT_STAGE_STRIDE + d_lo * 2) = __float2bfloat16(v_lo);
                        *reinterpret_cast<__nv_bfloat16*>(sb + h * OUT_STAGE_STRIDE + d_hi * 2) = __float2bfloat16(v_hi);
                    }
                } else {
                    *reinterpret_cast<float*>(sb + h * PART_STAGE_STRIDE + d_lo * 4) = v_lo;
                    *reinterpret_cast<float*>(sb + h * PART_STAGE_STRIDE + d_hi * 4) = v_hi;
                }
            }
            if (!out_rows && g == 0) part_lse[h] = L > 0.f ? fmaf(acc.m[nt][j], sl2, log2f(L)) : -INFINITY;
        }
    __syncwarp();
    if (out_rows) {
        for (int i = lane; i < nheads * (D / 8); i += 32) {        // 16 vectors of 8 bf16 per head row
            const int h = i >> 4, c = i & 15;
            *reinterpret_cast<uint4*>(out_rows + h * D + c * 8) = *reinterpret_cast<const uint4*>(sb + h * OUT_STAGE_STRIDE + c * 16);
        }
    } else {
        for (int i = lane; i < nheads * (D / 4); i += 32) {        // 32 vectors of 4 floats per head row
            const int h = i >> 5, c = i & 31;
            *reinterpret_cast<float4*>(part_o + h * D + c * 4) = *reinterpret_cast<const float4*>(sb + h * PART_STAGE_STRIDE + c * 16);
        }
    }
}

__device__ __forceinline__ void init_pipeline(uint32_t bar_full, uint32_t bar_empty, int nstage, int tid) {
    if (tid == 0) {
        for (int s = 0; s < nstage; ++s) {
            b2::mbar_init(bar_full + s * 8, 1);       // producer's arrive.expect_tx
            b2::mbar_init(bar_empty + s * 8, NCW);    // one arrive per consumer warp
        }
        b2::fence_mbar_init();
    }
    __syncthreads();
}

// Fused KV write of b2_paged_decode_append: one warp stores the new token of (row b, kv head kvh) -- 16 lanes the K
// row, 16 lanes the V row, one 16-byte vector of 8 bf16 each (bit-identical to b2_reshape_and_cache, e4m3 included).
// The row is read back by THIS CTA's own TMA loads (async proxy): every writing lane fences generic -> async for
// global memory, then the warp synchronises before its lane 0 issues the copy of the tile that holds the token.
// Two halves so that the loads are in flight while the producer streams the row's earlier tiles and the store (with
// its fences, ~1 us) sits in front of the LAST stage only, where two other stages cover it.
struct AppendVec {
    int64_t slot;
    int4 val;
};
__device__ __forceinline__ AppendVec append_load(const DecodeParams& p, int b, int kvh, int lane) {
    AppendVec a;
    a.slot = p.slot_mapping[b];
    const bool is_v = lane >= 16;
    const int c = lane & 15;                          // D / 8 = 16 vectors per row
    a.val = b2::ld_nc_16((is_v ? p.v_new + b * p.v_new_stride : p.k_new + b * p.k_new_stride) + kvh * D + c * 8);
    return a;
}
template <int BS, bool FP8>
__device__ __forceinline__ void append_store(const DecodeParams& p, const AppendVec& a, int b, int seq_len, int kvh, int lane) {
    if (a.slot >= 0) {                                // < 0: padding row, nothing to write
        const int64_t blk = a.slot / BS;
        const int off = (int)(a.slot - blk * BS);
        // resident tables: the slot names the block of column (seq_len - 1) / BS (as b2_reshape_and_cache_decode; every
        // kv head's CTA of the row stores the same value)
        if (p.bt_update && lane == 0 && seq_len > 0 && (seq_len - 1) / BS < p.max_nb)
            p.bt_update[(int64_t)b * p.max_nb + (seq_len - 1) / BS] = (int32_t)blk;
        const bool is_v = lane >= 16;
        const int c = lane & 15;
        const int64_t dst_elem = blk * p.cache_block_stride + kvh * p.cache_head_stride + (int64_t)off * D + c * 8;
        if (!FP8) {
            *reinterpret_cast<int4*>(static_cast<__nv_bfloat16*>(is_v ? p.v_cache_w : p.k_cache_w) + dst_elem) = a.val;
        } else {
            *reinterpret_cast<uint2*>(static_cast<uint8_t*>(is_v ? p.v_cache_w : p.k_cache_w) + dst_elem) =
                b2::pack_e4m3x8(a.val, is_v ? *p.v_scale : *p.k_scale);
        }
    }
    __threadfence();
    b2::fence_proxy_async_global();
    __syncwarp();
}

// =============================================================================================
// schedule A: classic grid split, one CTA per (split, kv head x head group, sequence)
// =============================================================================================
template <int BS, bool FP8, int NT, int NSTAGE>
__global__ void __launch_bounds__(NTHREADS, (Cfg<BS, FP8, NT, NSTAGE>::CTAS_SPLIT))
paged_decode_split_kernel(const __grid_constant__ CUtensorMap tmK, const __grid_constant__ CUtensorMap tmV,
                          const DecodeParams p) {
    using C = Cfg<BS, FP8, NT, NSTAGE>;
    extern __shared__ __align__(1024) uint8_t smem_raw[];

    const int split = blockIdx.x;
    const int kvh = blockIdx.y / p.HG;
    const int hg = blockIdx.y - kvh * p.HG;
    const int b = blockIdx.z;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

    const int cta_lin = (blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
    if (tid == 0) B2_TRACE_AT(cta_lin, 0);
    const int seq_len = p.seq_lens[b];
    const int t_begin = split * p.split_tokens;
    const int head0 = kvh * p.G + hg * C::HEADS;                 // first q head of this CTA
    const int nheads = min(C::HEADS, p.G - hg * C::HEADS);       // valid q heads of this CTA

    if (t_begin >= seq_len) {
        // empty split: nothing to read.  Single-pass mode must still define the output row.
        if (p.num_splits == 1)
            for (int i = tid; i < nheads * D; i += NTHREADS)
                p.out[b * p.out_stride + (int64_t)head0 * D + i] = __float2bfloat16(0.f);
        if (p.append && warp == NCW) {                // the KV write does not depend on seq_len (as b2_reshape_and_cache)
            asm volatile("griddepcontrol.wait;" ::: "memory");
            append_store<BS, FP8>(p, append_load(p, b, kvh, lane), b, seq_len, kvh, lane);
        }
        return;
    }
    const int n_tok = min(seq_len, t_begin + p.split_tokens) - t_begin;

    const uint32_t ring = b2::smem_u32(smem_raw);
    if (ring & 1023u) __trap();   // swizzle phase is derived from address bits
    const uint32_t bar_full = ring + C::RING_BYTES;
    const uint32_t bar_empty = bar_full + NSTAGE * 8;
    float* scratch = reinterpret_cast<float*>(smem_raw + C::RING_BYTES + C::BAR_BYTES);
    const uint32_t bar_append = bar_empty + NSTAGE * 8;   // the spare slot of the barrier area (append mode 2 only)
    if (tid == 0) b2::mbar_init(bar_append, 1);            // made visible by the fence + barrier of init_pipeline
    init_pipeline(bar_full, bar_empty, NSTAGE, tid);

    uint32_t it = 0;
    if (warp == NCW) {
        if (lane == 0) {
            b2::tma_prefetch_desc(&tmK);
            b2::tma_prefetch_desc(&tmV);
        }
        const uint64_t pol = b2::policy_evict_first();
        // programmatic dependent launch: this grid may have started under the tail of the step's KV write (which
        // triggers launch_dependents at its start).  Everything above -- barrier init, descriptor prefetch, the
        // consumers' q loads -- overlaps it; the cache and the block tables are only touched from here on.
        asm volatile("griddepcontrol.wait;" ::: "memory");
        const int32_t* bt = p.block_tables + (int64_t)b * p.max_nb + t_begin / BS;
        const int n_blk = (n_tok + BS - 1) / BS;
        if (p.append) {
            // fused KV write (single pass: this CTA streams the whole row, the new token sits in its last block): every
            // stage but the last goes out first, then the store, then the stage that holds it
            AppendVec a;
            if (p.append == 1) a = append_load(p, b, kvh, lane);
            else a.slot = p.slot_mapping[b];              // mode 2: consumer warp 0 stores; only the slot is needed here
            const int head = (n_blk - 1) / C::BLK_PER_STAGE * C::BLK_PER_STAGE;
            if (head)
                produce_segment<C, NSTAGE>(&tmK, &tmV, bt, head, kvh, ring, bar_full, bar_empty, it, lane, pol, p.two_box);
            if (p.append == 1) append_store<BS, FP8>(p, a, b, seq_len, kvh, lane);
            else b2::mbar_wait(bar_append, 0);            // the row is stored and fenced (long ago, unless the row is short)
            produce_segment<C, NSTAGE>(&tmK, &tmV, bt + head, n_blk - head, kvh, ring, bar_full, bar_empty, it, lane, pol,
                                       p.two_box, p.bt_update && a.slot >= 0 ? (int)(a.slot / BS) : -1);
        } else {
            produce_segment<C, NSTAGE>(&tmK, &tmV, bt, n_blk, kvh, ring, bar_full, bar_empty, it, lane, pol, p.two_box);
        }
        B2_TRACE_AT(cta_lin, 4);
    } else {
        if (warp == 0) B2_TRACE_AT(cta_lin, 1);
        float sl2 = p.scale_log2, vs = 1.f;
        if (FP8) {
            sl2 *= *p.k_scale;   // K dequant scale folds into the softmax scale
            vs = *p.v_scale;     // V dequant scale folds into the epilogue
        }
        QFrags<NT> qf;
        // q may be the output of the kernel this grid was launched under (rotary + KV write in the engine)
        asm volatile("griddepcontrol.wait;" ::: "memory");
        AppendVec a;
        const bool storer = p.append == 2 && warp == 0;   // warp-uniform
        if (storer) a = append_load(p, b, kvh, lane);     // in flight under the q loads
        load_q_frags<FP8, NT>(qf, reinterpret_cast<const uint32_t*>(p.q + b * p.q_stride + (int64_t)head0 * D), nheads, lane);
        if (storer) {
            // append mode 2: this warp idles until the first stage lands anyway; the store, its fences and the hand-over
            // to the producer (release on the mbarrier, acquired by its wait in front of the last stage) fit in that gap
            append_store<BS, FP8>(p, a, b, seq_len, kvh, lane);
            if (lane == 0) b2::mbar_arrive(bar_append);
        }
        WarpAcc<NT> acc;
        acc.reset();
        consume_segment<C, FP8, NT, NSTAGE>(acc, qf, n_tok, ring, bar_full, bar_empty, it, sl2, warp, lane);
        if (warp == 0) B2_TRACE_AT(cta_lin, 3);
        merge_warps<NT>(acc, scratch, C::SLOT_FLOATS, sl2, warp, lane);
        if (warp == 0) {
            if (p.num_splits == 1) {
                write_result<FP8, NT>(acc, nheads, vs, sl2, p.out + b * p.out_stride + (int64_t)head0 * D, nullptr,
                                      nullptr, scratch, lane);
            } else {
                // partial of q head (head0 + h) lives at ((b*Hq + head0 + h)*num_splits + split)
                const int g = lane >> 2, t4 = lane & 3;
#pragma unroll
                for (int nt = 0; nt < NT; ++nt)
#pragma unroll
                    for (int j = 0; j < 2; ++j) {
                        const int h = nt * 8 + 2 * t4 + j;
                        if (h >= nheads) continue;
                        const int64_t pi = ((int64_t)b * p.Hq + head0 + h) * p.num_splits + split;
                        const float L = acc.l[nt][j];
                        const float inv = vs / L;   // L > 0: the split holds at least one valid token
#pragma unroll
                        for (int dt = 0; dt < 8; ++dt) {
                            const int d_lo = FP8 ? 16 * dt + 2 * g : 16 * dt + g;
                            const int d_hi = FP8 ? d_lo + 1 : d_lo + 8;
                            p.ws_o[pi * D + d_lo] = acc.o[nt][dt][j] * inv;
                            p.ws_o[pi * D + d_hi] = acc.o[nt][dt][j + 2] * inv;
                        }
                        if (g == 0) p.ws_lse[pi] = fmaf(acc.m[nt][j], sl2, log2f(L));
                    }
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

// =============================================================================================
// schedule B (default, num_splits = 0): persistent CTAs, a static head + dynamically claimed chunks
// =============================================================================================
// Work is cut on a fixed grid: item (b, j) has nck(b) = ceil(seq_len / CT) chunks of CT tokens (CT a
// multiple of the 64-token stage).  Chunk ids are flattened sequence-major, then item, then chunk.
// CTA x first streams the contiguous ids [x*S, (x+1)*S), S = floor(pct% * Ctot / NC); the remaining
// ids are claimed one at a time from an atomic counter, so CTAs that stream faster (SM / L2-slice
// placement differs by ~10 %) take more of the tail and the grid drains together.  Consecutive
// static chunks of one item are accumulated in registers (a "run"); every dynamic chunk is its own
// run.  A run that covers its item writes bf16 output, any other run writes a normalised fp32
// partial + log2-sum-exp to a slot that whoever merges the item can compute from (Ctot, S, NC) alone:
//     run starts  : k == 0, or c >= NC*S, or c % S == 0
//     slot(c)     : c >= NC*S -> 2*NC + c - NC*S ;  c % S == 0 -> 2*(c/S) ;  else 2*(c/S) + 1
// The producer warp owns the schedule and hands each 64-token stage to the consumers together
// with a 16-byte descriptor {sequence, item, valid tokens, flags | slot << 4} in shared memory
// (ordered by the stage's full-barrier), so the consumer warps are a plain loop over stages.
constexpr int META_FIRST = 1, META_LAST = 2, META_END = 4, META_DIRECT = 8;

__device__ __forceinline__ uint32_t dyn_static_chunks(uint32_t Ctot, uint32_t NC, int pct) {
    return (uint32_t)(((uint64_t)Ctot * (uint32_t)pct / 100u) / NC);
}
__device__ __forceinline__ bool dyn_is_start(uint32_t c, uint32_t k, uint32_t S, uint32_t stat_total) {
    return k == 0 || c >= stat_total || (c % S) == 0;   // S > 0 whenever c < stat_total
}
__device__ __forceinline__ uint32_t dyn_slot(uint32_t c, uint32_t S, uint32_t stat_total, uint32_t NC) {
    if (c >= stat_total) return 2 * NC + (c - stat_total);
    const uint32_t x = c / S;
    return (c % S) == 0 ? 2 * x : 2 * x + 1;
}
__device__ __forceinline__ uint32_t chunks_of(int seq_len, int ct) {
    return seq_len > 0 ? (uint32_t)((seq_len + ct - 1) / ct) : 0u;
}

// pre[e], e in [0, E]: chunks before sequence group e (gs sequences per group).  Returns Ctot = pre[E].
template <int NT_>
__device__ __forceinline__ uint32_t build_chunk_prefix(const int32_t* __restrict__ seq_lens, int B, int IPB, int ct,
                                                       int gs, int E, uint32_t* pre) {
    const int tid = threadIdx.x;
    for (int e = tid; e < E; e += NT_) {
        uint32_t s = 0;
        const int b1 = min(B, (e + 1) * gs);
        for (int b = e * gs; b < b1; ++b) s += chunks_of(seq_lens[b], ct) * IPB;
        pre[e + 1] = s;
    }
    if (tid == 0) pre[0] = 0;
    __syncthreads();
    if (tid < 32) {   // in-place inclusive scan of pre[1..E], one contiguous strip per lane
        const int per = (E + 31) / 32;
        const int i0 = 1 + tid * per, i1 = min(E + 1, i0 + per);
        uint32_t s = 0;
        for (int i = i0; i < i1; ++i) s += pre[i];
        uint32_t incl = s;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, incl, off);
            if (tid >= off) incl += t;
        }
        uint32_t run = incl - s;
        for (int i = i0; i < i1; ++i) {
            run += pre[i];
            pre[i] = run;
        }
    }
    __syncthreads();
    return pre[E];
}

struct ChunkPos {
    int b, j, k, nck, seq_len;
};
// warp-uniform: every lane computes the same position
__device__ __forceinline__ ChunkPos locate_chunk(uint32_t c, const uint32_t* pre, int E, int gs,
                                                 const int32_t* __restrict__ seq_lens, int B, int IPB, int ct) {
    int lo = 0, hi = E;   // largest e with pre[e] <= c  (pre[E] = Ctot > c)
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (pre[mid] <= c) lo = mid; else hi = mid;
    }
    uint32_t acc = pre[lo];
    ChunkPos r{0, 0, 0, 1, 0};
    const int b1 = min(B, (lo + 1) * gs);
    for (int b = lo * gs; b < b1; ++b) {
        const int L = seq_lens[b];
        const uint32_t nck = chunks_of(L, ct), span = nck * IPB;
        if (c < acc + span) {
            const uint32_t rr = c - acc;
            r.b = b; r.j = (int)(rr / nck); r.k = (int)(rr % nck); r.nck = (int)nck; r.seq_len = L;
            return r;
        }
        acc += span;
    }
    return r;   // not reached for c < Ctot
}

// Number of runs item [ci0, ci0 + nck) was cut into.
__device__ __forceinline__ uint32_t dyn_runs_of_item(uint32_t ci0, uint32_t nck, uint32_t S, uint32_t stat_total) {
    uint32_t n = 1;
    for (uint32_t c = ci0 + 1; c < ci0 + nck; ++c) n += dyn_is_start(c, c - ci0, S, stat_total) ? 1u : 0u;
    return n;
}

// The same count in O(1): run starts in (ci0, ci0 + nck) are the multiples of S below stat_total and every chunk at or
// above stat_total.
__device__ __forceinline__ uint32_t dyn_runs_of_item_fast(uint32_t ci0, uint32_t nck, uint32_t S, uint32_t stat_total) {
    const uint32_t a = ci0 + 1, b = ci0 + nck;            // candidate starts: [a, b)
    uint32_t n = 1;
    const uint32_t hi = b < stat_total ? b : stat_total;
    if (S > 0 && hi > a) n += (hi - 1) / S - (a - 1) / S;
    const uint32_t lo = a > stat_total ? a : stat_total;
    if (b > lo) n += b - lo;
    return n;
}

// The four consumer warps (tid128 = 0..127) combine the partial runs of one item into bf16 output rows.
// (Round 2 experiment, removed again: combining the partials INSIDE the decode kernel -- per-item arrival counters in
// the workspace header, the CTA that writes an item's last partial merges it -- was parity-green but slower than the
// PDL merge kernel on every ragged shape (63.1 vs 60.1 us config-2 shape, 99.9 vs 89.6 fp8): two fences + an atomic at
// every partial run's end sit on the consumers' critical path, the merging CTA stalls its own pipeline, and the extra
// code cost the 128-register fp8 instantiation its spill-free allocation.)
// thread -> (head tid128/16 [+8], 8 consecutive d); partials were written by other CTAs: read through L2 (ld.cg).
template <int HEADS>
__device__ __forceinline__ void merge_item(__nv_bfloat16* __restrict__ out_rows, const float* ws_o, const float* ws_lse,
                                           uint32_t ci0, uint32_t nck, uint32_t S, uint32_t stat_total, uint32_t NC,
                                           int nheads, int tid128) {
    constexpr int MAXC = 4;
    const uint32_t cend = ci0 + nck;
    const int dc = (tid128 & 15) * 8;
    for (int h = tid128 >> 4; h < nheads; h += 8) {
        float m_run = -INFINITY, den = 0.f, num[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) num[i] = 0.f;
        uint32_t c = ci0;
        while (c < cend) {
            float lse[MAXC];
            float4 o0[MAXC], o1[MAXC];
#pragma unroll
            for (int i = 0; i < MAXC; ++i) {
                while (c < cend && !dyn_is_start(c, c - ci0, S, stat_total)) ++c;
                lse[i] = -INFINITY;
                o0[i] = o1[i] = make_float4(0.f, 0.f, 0.f, 0.f);
                if (c < cend) {
                    const int64_t slot = dyn_slot(c, S, stat_total, NC);
                    lse[i] = __ldcg(ws_lse + slot * HEADS + h);
                    const float4* po = reinterpret_cast<const float4*>(ws_o + (slot * HEADS + h) * D + dc);
                    o0[i] = __ldcg(po);
                    o1[i] = __ldcg(po + 1);
                    ++c;
                }
            }
            float mx = m_run;
#pragma unroll
            for (int i = 0; i < MAXC; ++i) mx = fmaxf(mx, lse[i]);
            if (mx == -INFINITY) continue;
            const float a = m_run == -INFINITY ? 0.f : exp2f(m_run - mx);
            den *= a;
#pragma unroll
            for (int i = 0; i < 8; ++i) num[i] *= a;
#pragma unroll
            for (int i = 0; i < MAXC; ++i) {
                const float w = lse[i] == -INFINITY ? 0.f : exp2f(lse[i] - mx);
                den += w;
                num[0] = fmaf(w, o0[i].x, num[0]); num[1] = fmaf(w, o0[i].y, num[1]);
                num[2] = fmaf(w, o0[i].z, num[2]); num[3] = fmaf(w, o0[i].w, num[3]);
                num[4] = fmaf(w, o1[i].x, num[4]); num[5] = fmaf(w, o1[i].y, num[5]);
                num[6] = fmaf(w, o1[i].z, num[6]); num[7] = fmaf(w, o1[i].w, num[7]);
            }
            m_run = mx;
        }
        const float inv = den > 0.f ? 1.f / den : 0.f;
        uint4 pk;
        pk.x = b2::pack_bf16x2(num[0] * inv, num[1] * inv);
        pk.y = b2::pack_bf16x2(num[2] * inv, num[3] * inv);
        pk.z = b2::pack_bf16x2(num[4] * inv, num[5] * inv);
        pk.w = b2::pack_bf16x2(num[6] * inv, num[7] * inv);
        *reinterpret_cast<uint4*>(out_rows + (int64_t)h * D + dc) = pk;
    }
}

template <int BS, bool FP8, int NT, int NSTAGE>
__global__ void __launch_bounds__(NTHREADS, (Cfg<BS, FP8, NT, NSTAGE>::CTAS_DYN))
paged_decode_dynamic_kernel(const __grid_constant__ CUtensorMap tmK, const __grid_constant__ CUtensorMap tmV,
                            const DecodeParams p) {
    using C = Cfg<BS, FP8, NT, NSTAGE>;
    static_assert(C::CTAS_DYN * (C::DYN_SMEM_BYTES + 1024) <= 233472, "the CTAs of one SM must fit (dynamic schedule)");
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    int4* s_meta = reinterpret_cast<int4*>(smem_raw + C::RING_BYTES + C::BAR_BYTES);
    float* scratch = reinterpret_cast<float*>(smem_raw + C::RING_BYTES + C::BAR_BYTES + C::META_BYTES);
    uint32_t* s_pre =
        reinterpret_cast<uint32_t*>(smem_raw + C::RING_BYTES + C::BAR_BYTES + C::META_BYTES + C::SCRATCH_BYTES);
    uint8_t* s_q = smem_raw + C::RING_BYTES + C::BAR_BYTES + C::META_BYTES + C::SCRATCH_BYTES + C::PRE_BYTES;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t x = blockIdx.x, NC = gridDim.x;
    const int IPB = p.Hkv * p.HG;
    const int ct = p.chunk_tokens;
    // the merge kernel may become resident now; it waits (griddepcontrol.wait) until this grid is done
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");

    // rows of empty sequences are defined here (no chunk ever visits them)
    for (int b = x; b < p.B; b += NC)
        if (p.seq_lens[b] <= 0)
            for (int i = tid; i < p.Hq * D; i += NTHREADS) p.out[b * p.out_stride + i] = __float2bfloat16(0.f);

    if (tid == 0) B2_TRACE_AT(x, 0);
    const int gs = (p.B + C::PMAX - 1) / C::PMAX, E = (p.B + gs - 1) / gs;
    const uint32_t Ctot = build_chunk_prefix<NTHREADS>(p.seq_lens, p.B, IPB, ct, gs, E, s_pre);
    // The claim counter is zeroed by the preceding (tiny) kernel of this call.  Only the producer warp touches
    // it, and only after its static share, so it orders itself after that kernel (griddepcontrol.wait) right
    // before its first claim -- and at the latest before it exits, which keeps "this grid finished" =>
    // "that grid finished" for whatever follows in the stream.
    if (Ctot == 0) {
        asm volatile("griddepcontrol.wait;" ::: "memory");
        return;
    }
    const uint32_t S = dyn_static_chunks(Ctot, NC, p.static_pct);
    const uint32_t stat_total = S * NC;

    const uint32_t ring = b2::smem_u32(smem_raw);
    if (ring & 1023u) __trap();
    const uint32_t bar_full = ring + C::RING_BYTES;
    const uint32_t bar_empty = bar_full + NSTAGE * 8;
    init_pipeline(bar_full, bar_empty, NSTAGE, tid);
    uint32_t it = 0;
    if (tid == 0) B2_TRACE_AT(x, 1);

    if (warp == NCW) {
        // ================================ producer warp =====================================
        if (lane == 0) {
            b2::tma_prefetch_desc(&tmK);
            b2::tma_prefetch_desc(&tmV);
        }
        const uint64_t pol = b2::policy_evict_first();
        const int CB = ct / BS;                         // <= 32: a chunk's block ids fit one register per lane
        uint32_t stat_cur = x * S;
        const uint32_t stat_end = stat_cur + S;
        uint32_t cid;
        bool cur_static;
        bool ordered = false;                           // griddepcontrol.wait executed
        if (stat_cur < stat_end) {
            cid = stat_cur++;
            cur_static = true;
        } else {
            asm volatile("griddepcontrol.wait;" ::: "memory");
            ordered = true;
            uint32_t v = 0;
            if (lane == 0) v = atomicAdd(p.counter, 1u);
            cid = stat_total + __shfl_sync(0xffffffffu, v, 0);
            cur_static = false;
        }
        bool run_first = true, cont_prev = false;
        uint32_t run_c0 = 0;
        int run_k0 = 0;
        ChunkPos cp{0, 0, 0, 1, 0};
        // the NEXT static chunk's position and block-table entries are fetched while the current chunk streams: a chunk
        // boundary otherwise costs the producer two dependent global loads (seq_lens, then the table row) with nothing
        // new in flight -- on ragged batches (a chunk or two per item) that was ~1 us per item
        bool have_next = false;
        ChunkPos cp_next{0, 0, 0, 1, 0};
        int tbl_next = 0;
        auto table_entries = [&](const ChunkPos& c) {
            const int nb_ = (c.seq_len + BS - 1) / BS;
            const int nblk_ = min(CB, nb_ - c.k * CB);
            return lane < nblk_ ? p.block_tables[(int64_t)c.b * p.max_nb + c.k * CB + lane] : 0;
        };
        while (cid < Ctot) {
            int tbl;
            if (have_next) {
                cp = cp_next;
                tbl = tbl_next;
            } else {
                if (cont_prev) ++cp.k;                  // next chunk of the same item
                else cp = locate_chunk(cid, s_pre, E, gs, p.seq_lens, p.B, IPB, ct);
                tbl = table_entries(cp);
            }
            have_next = false;
            if (run_first) {
                run_c0 = cid;
                run_k0 = cp.k;
            }
            const int nb = (cp.seq_len + BS - 1) / BS;
            const int blk0 = cp.k * CB;
            const int nblk = min(CB, nb - blk0);
            // the id after this one: static ids are known, a dynamic one is claimed now and read late
            const bool next_static = cur_static && stat_cur < stat_end;
            uint32_t pend = 0;
            if (!next_static) {
                if (!ordered) {
                    asm volatile("griddepcontrol.wait;" ::: "memory");
                    ordered = true;
                }
                if (lane == 0) pend = atomicAdd(p.counter, 1u);
            }
            const bool cont = next_static && (cp.k + 1 < cp.nck);
            if (next_static) {                          // stat_cur is the next chunk id: fetch it under this chunk's stages
                if (cont) {
                    cp_next = cp;
                    ++cp_next.k;
                } else {
                    cp_next = locate_chunk(stat_cur, s_pre, E, gs, p.seq_lens, p.B, IPB, ct);
                }
                tbl_next = table_entries(cp_next);
                have_next = true;
            }
            const int ntok = min(ct, cp.seq_len - cp.k * ct);
            const int n_stage = (ntok + TOK - 1) / TOK;
            const int kvh = cp.j / p.HG;
            for (int st = 0; st < n_stage; ++st, ++it) {
                const uint32_t s = it % NSTAGE;
                const uint32_t ph = (it / NSTAGE) & 1;
                b2::mbar_wait(bar_empty + s * 8, ph ^ 1);
                const int nb_stage = min(C::BLK_PER_STAGE, nblk - st * C::BLK_PER_STAGE);
                const uint32_t kdst = ring + s * C::STAGE_BYTES;
                const uint32_t vdst = kdst + C::TILE_BYTES;
                if (lane == 0) {
                    int flags = 0;
                    if (st == 0 && run_first) flags |= META_FIRST;
                    if (st == n_stage - 1 && !cont) {
                        flags |= META_LAST;
                        if (run_k0 == 0 && cp.k == cp.nck - 1) flags |= META_DIRECT;
                        else flags |= (int)(dyn_slot(run_c0, S, stat_total, NC) << 4);
                    }
                    s_meta[s] = make_int4(cp.b, cp.j, min(TOK, ntok - st * TOK), flags);
                    uint32_t tx = nb_stage * 2 * C::BOX_BYTES;
                    if (C::QSMEM && (flags & META_FIRST)) {
                        const int hg = cp.j - kvh * p.HG;
                        const int head0 = kvh * p.G + hg * C::HEADS;
                        const uint32_t qbytes = min(C::HEADS, p.G - hg * C::HEADS) * C::QROW_BYTES;
                        tx += qbytes;
                        b2::mbar_arrive_expect_tx(bar_full + s * 8, tx);
                        b2::bulk_load(b2::smem_u32(s_q + s * C::QBUF_BYTES),
                                      p.q + cp.b * p.q_stride + (int64_t)head0 * D, qbytes, bar_full + s * 8);
                    } else {
                        b2::mbar_arrive_expect_tx(bar_full + s * 8, tx);
                    }
                }
#pragma unroll
                for (int j = 0; j < C::BLK_PER_STAGE; ++j) {
                    const int blk = __shfl_sync(0xffffffffu, tbl, (st * C::BLK_PER_STAGE + j) & 31);
                    if (j < nb_stage && lane == 0)
                        issue_block<C, FP8>(&tmK, &tmV, kdst, vdst, j, bar_full + s * 8, kvh, blk, pol, p.two_box);
                }
            }
            run_first = !cont;
            cont_prev = cont;
            if (next_static) cid = stat_cur++;
            else cid = stat_total + __shfl_sync(0xffffffffu, pend, 0);
            cur_static = next_static;
        }
        if (!ordered) asm volatile("griddepcontrol.wait;" ::: "memory");
        // end-of-work sentinel: a stage with no bytes
        const uint32_t s = it % NSTAGE;
        b2::mbar_wait(bar_empty + s * 8, ((it / NSTAGE) & 1) ^ 1);
        if (lane == 0) {
            s_meta[s] = make_int4(0, 0, 0, META_END);
            b2::mbar_arrive(bar_full + s * 8);
        }
    } else {
        // ================================ consumer warps ====================================
        float sl2 = p.scale_log2, vs = 1.f;
        if (FP8) {
            sl2 *= *p.k_scale;
            vs = *p.v_scale;
        }
        WarpAcc<NT> acc;
        acc.reset();
        QFrags<NT> qf;
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
#pragma unroll
            for (int kc = 0; kc < 8; ++kc) qf.r[nt][kc][0] = qf.r[nt][kc][1] = 0u;
            qf.hs[nt][0] = qf.hs[nt][1] = 1.f;
        }
        int cb = 0, head0 = 0, nheads = 0;
        const int row_base = warp * 16;
        for (;; ++it) {
            const uint32_t s = it % NSTAGE;
            b2::mbar_wait(bar_full + s * 8, (it / NSTAGE) & 1);
            const int4 m = s_meta[s];
            if (m.w & META_END) {
                if (warp == 0) {
                    B2_TRACE_AT(x, 3);
                    B2_TRACE_VAL(x, 5, (unsigned long long)it);
                }
                break;
            }
            if (m.w & META_FIRST) {
                cb = m.x;
                const int kvh = m.y / p.HG, hg = m.y - kvh * p.HG;
                head0 = kvh * p.G + hg * C::HEADS;
                nheads = min(C::HEADS, p.G - hg * C::HEADS);
                if (C::QSMEM)
                    load_q_frags<FP8, NT>(qf, reinterpret_cast<const uint32_t*>(s_q + s * C::QBUF_BYTES), nheads, lane);
                else
                    load_q_frags<FP8, NT>(qf, reinterpret_cast<const uint32_t*>(p.q + cb * p.q_stride + (int64_t)head0 * D),
                                          nheads, lane);
                acc.reset();
            }
            const int nvalid = m.z - row_base;
            if (nvalid > 0) {
                const uint32_t kbase = ring + s * C::STAGE_BYTES;
                consume_slab<C, FP8, NT>(acc, qf, kbase, kbase + C::TILE_BYTES, row_base, nvalid, sl2, lane);
            }
            __syncwarp();
            if (lane == 0) b2::mbar_arrive(bar_empty + s * 8);   // the stage and its descriptor may be reused
            if (m.w & META_LAST) {
                merge_warps<NT>(acc, scratch, C::SLOT_FLOATS, sl2, warp, lane);
                if (warp == 0) {
                    if (m.w & META_DIRECT) {
                        write_result<FP8, NT>(acc, nheads, vs, sl2, p.out + cb * p.out_stride + (int64_t)head0 * D,
                                              nullptr, nullptr, scratch, lane);
                    } else {
                        const int64_t slot = (int64_t)((uint32_t)m.w >> 4);
                        write_result<FP8, NT>(acc, nheads, vs, sl2, nullptr, p.ws_o + slot * (C::HEADS * D),
                                              p.ws_lse + slot * C::HEADS, scratch, lane);
                    }
                }
                consumer_bar();   // warp 0 has read its staged rows: the scratch is free for the next run's merge
            }
        }
    }
}

// first kernel of a dynamic-schedule call: zero the claim counter
__global__ void zero_counter_kernel(uint32_t* counter) {
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");   // the decode grid may start its static share
    if (threadIdx.x == 0) *counter = 0u;
}

// One block per item (sequence b = blockIdx.y, item j = blockIdx.x): combine its partial runs unless one
// run covered it.  Which chunks start a run, and where their partials live, follows from (Ctot, S, NC)
// -- the same arithmetic as the main kernel, no flags in memory.
template <int HEADS>
__global__ void __launch_bounds__(128)
merge_chunks_kernel(__nv_bfloat16* __restrict__ out, const float* __restrict__ ws_o,
                    const float* __restrict__ ws_lse, const int32_t* __restrict__ seq_lens, int B, int Hkv, int G,
                    int HG, int64_t out_stride, int ct, int static_pct, uint32_t NC, uint32_t* counter) {
    __shared__ uint32_t s_red[2][4];
    const int tid = threadIdx.x, b = blockIdx.y, j = blockIdx.x;
    const int IPB = Hkv * HG;
    uint32_t before = 0, total = 0;
    for (int i = tid; i < B; i += 128) {
        const uint32_t n = chunks_of(seq_lens[i], ct) * IPB;
        total += n;
        if (i < b) before += n;
    }
#pragma unroll
    for (int off = 16; off; off >>= 1) {
        before += __shfl_xor_sync(0xffffffffu, before, off);
        total += __shfl_xor_sync(0xffffffffu, total, off);
    }
    if ((tid & 31) == 0) {
        s_red[0][tid >> 5] = before;
        s_red[1][tid >> 5] = total;
    }
    __syncthreads();
    before = s_red[0][0] + s_red[0][1] + s_red[0][2] + s_red[0][3];
    total = s_red[1][0] + s_red[1][1] + s_red[1][2] + s_red[1][3];
    // every block waits for the decode grid, so "merge finished" implies "decode finished" for later work
    asm volatile("griddepcontrol.wait;" ::: "memory");
    // the decode grid is complete: leave the claim counter at zero for the next call on this workspace
    // (b2_decode_plan_run relies on it and launches no zeroing kernel; b2_paged_decode zeroes it itself)
    if (blockIdx.x == 0 && blockIdx.y == 0 && tid == 0) *counter = 0u;
    const uint32_t nck = chunks_of(seq_lens[b], ct);
    if (nck <= 1) return;                                 // a single chunk is always written directly
    const uint32_t S = dyn_static_chunks(total, NC, static_pct);
    const uint32_t stat_total = S * NC;
    const uint32_t ci0 = before + (uint32_t)j * nck;
    if (dyn_runs_of_item_fast(ci0, nck, S, stat_total) == 1) return;   // one run covered the item: already in `out`
    const int kvh = j / HG, hg = j - kvh * HG;
    const int head0 = kvh * G + hg * HEADS;
    merge_item<HEADS>(out + b * out_stride + (int64_t)head0 * D, ws_o, ws_lse, ci0, nck, S, stat_total, NC,
                      min(HEADS, G - hg * HEADS), tid);
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

// Legacy view ([row][half] landing order) -- what prefill_attention.cu's mma.sync kernel addresses.
// fp8: cache viewed as [num_blocks][Hkv][bs][128 B]; bf16: [num_blocks][Hkv][bs][2 halves][64 elements].
// One box = one whole (block, head) tile, contiguous in HBM.
int make_cache_tmap(CUtensorMap* tm, const void* cache, int kv_dtype, int64_t num_blocks, int Hkv,
                    int bs, int64_t block_stride, int64_t head_stride) {
    EncodeTiledFn enc = get_encode_tiled();
    B2_REQUIRE(enc != nullptr, B2_ECUDA, "cuTensorMapEncodeTiled entry point not available");
    CUresult r;
    if (kv_dtype == B2_KV_FP8_E4M3) {
        cuuint64_t dims[4] = {(cuuint64_t)D, (cuuint64_t)bs, (cuuint64_t)Hkv, (cuuint64_t)num_blocks};
        cuuint64_t strides[3] = {(cuuint64_t)D, (cuuint64_t)head_stride, (cuuint64_t)block_stride};
        cuuint32_t box[4] = {128, (cuuint32_t)bs, 1, 1};
        cuuint32_t estr[4] = {1, 1, 1, 1};
        r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 4, const_cast<void*>(cache), dims, strides, box, estr,
                CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    } else {
        cuuint64_t dims[5] = {64, 2, (cuuint64_t)bs, (cuuint64_t)Hkv, (cuuint64_t)num_blocks};
        cuuint64_t strides[4] = {128, (cuuint64_t)D * 2, (cuuint64_t)head_stride * 2, (cuuint64_t)block_stride * 2};
        cuuint32_t box[5] = {64, 2, (cuuint32_t)bs, 1, 1};
        cuuint32_t estr[5] = {1, 1, 1, 1, 1};
        r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, const_cast<void*>(cache), dims, strides, box, estr,
                CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    }
    B2_REQUIRE(r == CUDA_SUCCESS, B2_ECUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
    return B2_OK;
}

int env_int(const char* name, int dflt) {
    const char* s = getenv(name);
    return (s && *s) ? atoi(s) : dflt;
}

// Diagnostic knobs (environment): read once per b2_decode_plan, per call by the stateless b2_paged_decode.
struct Knobs {
    int chunk_tokens;   // B2_CHUNK_TOKENS: claim granularity of the dynamic schedule (0 = automatic)
    int static_pct;     // B2_STATIC_PCT: share of the chunks dealt statically
    int fp8_ctas;       // B2_FP8_CTAS: 3 (default) or 2 resident CTAs per SM for the fp8 kernels
    int two_box;        // B2_TWO_BOX=1: force the two-boxes-per-tile bf16 map (the fallback path)
    int append_mode;    // B2_APPEND_MODE: who stores the fused append's row -- 2 (default) consumer warp 0 at CTA start,
                        // in the gap before its first stage lands; the producer waits for it on an mbarrier in front of its
                        // last stage.  1: the producer warp itself, in front of its last stage (A/B: 0.6 us slower per step)
};
Knobs read_knobs() {
    Knobs k;
    k.append_mode = env_int("B2_APPEND_MODE", 2) == 1 ? 1 : 2;
    k.chunk_tokens = env_int("B2_CHUNK_TOKENS", 0);
    int pct = env_int("B2_STATIC_PCT", 85);
    k.static_pct = pct < 0 ? 0 : (pct > 100 ? 100 : pct);
    k.fp8_ctas = env_int("B2_FP8_CTAS", 3) == 3 ? 3 : 2;
    k.two_box = env_int("B2_TWO_BOX", 0) != 0;
    return k;
}

// The decode kernels' view of the cache.  fp8: as above.  bf16: ONE box per (block, head) tile with the dimensions
// ordered [64 el][row][half] (row stride 256 B, half stride 128 B), so the tile lands as [half][row][64] -- see
// issue_block.  Should a driver refuse a map whose strides do not grow with the dimension index, the same image is
// built from two boxes of one half each (two_box = 1).
struct CacheMaps {
    CUtensorMap k, v;
    int two_box;
};
int make_decode_maps(CacheMaps* m, const void* k_cache, const void* v_cache, int kv_dtype, int64_t num_blocks, int Hkv,
                     int bs, int64_t block_stride, int64_t head_stride, bool force_two_box) {
    m->two_box = 0;
    if (kv_dtype == B2_KV_FP8_E4M3) {
        int rc = make_cache_tmap(&m->k, k_cache, kv_dtype, num_blocks, Hkv, bs, block_stride, head_stride);
        if (rc != B2_OK) return rc;
        return make_cache_tmap(&m->v, v_cache, kv_dtype, num_blocks, Hkv, bs, block_stride, head_stride);
    }
    EncodeTiledFn enc = get_encode_tiled();
    B2_REQUIRE(enc != nullptr, B2_ECUDA, "cuTensorMapEncodeTiled entry point not available");
    static int one_box_refused = 0;   // remembered per process: the answer depends on the driver only
    const cuuint32_t estr[5] = {1, 1, 1, 1, 1};
    if (!force_two_box && !one_box_refused) {
        const cuuint64_t dims[5] = {64, (cuuint64_t)bs, 2, (cuuint64_t)Hkv, (cuuint64_t)num_blocks};
        const cuuint64_t strides[4] = {(cuuint64_t)D * 2, 128, (cuuint64_t)head_stride * 2, (cuuint64_t)block_stride * 2};
        const cuuint32_t box[5] = {64, (cuuint32_t)bs, 2, 1, 1};
        CUresult rk = enc(&m->k, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, const_cast<void*>(k_cache), dims, strides, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        CUresult rv = rk != CUDA_SUCCESS ? rk
                                         : enc(&m->v, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, const_cast<void*>(v_cache), dims,
                                               strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                               CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (rk == CUDA_SUCCESS && rv == CUDA_SUCCESS) return B2_OK;
        one_box_refused = 1;
    }
    m->two_box = 1;
    const cuuint64_t dims[5] = {64, 2, (cuuint64_t)bs, (cuuint64_t)Hkv, (cuuint64_t)num_blocks};
    const cuuint64_t strides[4] = {128, (cuuint64_t)D * 2, (cuuint64_t)head_stride * 2, (cuuint64_t)block_stride * 2};
    const cuuint32_t box[5] = {64, 1, (cuuint32_t)bs, 1, 1};
    for (int i = 0; i < 2; ++i) {
        CUresult r = enc(i ? &m->v : &m->k, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, const_cast<void*>(i ? v_cache : k_cache), dims,
                         strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                         CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        B2_REQUIRE(r == CUDA_SUCCESS, B2_ECUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
    }
    return B2_OK;
}

// resident CTAs per SM (shared-memory bound): 2, or 3 for the shallow-ring fp8 variants (B2_FP8_CTAS=2 for A/B runs)
// (G > 8 serves 16 q heads per CTA: its accumulators do not fit the 128-register budget of three CTAs)
int ctas_per_sm(bool fp8, int G, int fp8_ctas) { return fp8 && G <= 8 && fp8_ctas == 3 ? 3 : 2; }

int sm_count() {
    static int sms = []() {
        int dev = 0, n = 148;
        if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
        return n;
    }();
    return sms;
}
int persistent_grid(int per_sm) { return per_sm * sm_count(); }

struct Plan {
    int nt;            // n-tiles of 8 q heads per CTA
    int hg;            // head groups per kv head
    int num_splits;    // 0 = persistent schedule
    int split_tokens;
    int chunk_tokens;  // dynamic: claim granularity
    int static_pct;    // dynamic: share of the chunks dealt out statically
    int64_t slots;     // dynamic: partial slots in the workspace
    int ipb;           // items per sequence = Hkv * hg
    int nc;            // dynamic: CTAs of the persistent grid the plan was made for
};

// dynamic schedule: workspace header (claim counter)
constexpr size_t WS_HEADER_BYTES = 256;

int dyn_ct_max(int bs) { return bs * 32 < 512 ? bs * 32 : 512; }   // <= 32 blocks per chunk

// per_sm: CTAs per SM of the kernel that will run (3 also bounds the workspace of every variant from above)
void make_plan(Plan* pl, int B, int Hkv, int G, int bs, int max_nb, int num_splits, int per_sm, const Knobs& kn) {
    pl->nt = G <= 8 ? 1 : 2;
    const int hpc = 8 * pl->nt;
    pl->hg = (G + hpc - 1) / hpc;
    pl->num_splits = 0;
    pl->split_tokens = 0;
    pl->chunk_tokens = 0;
    pl->static_pct = 0;
    pl->slots = 0;
    pl->ipb = Hkv * pl->hg;
    pl->nc = 0;
    const int64_t max_ctx = (int64_t)max_nb * bs;
    if (num_splits <= 0) {
        const int NC = persistent_grid(per_sm);
        const int ipb = Hkv * pl->hg;
        const int ct_max = dyn_ct_max(bs);
        // aim for >= 8 chunks per CTA (run boundaries cost more than a slightly longer tail), within [64, ct_max] tokens
        const int64_t total_tok = (int64_t)B * ipb * max_ctx;
        int ct = TOK;
        while (ct * 2 <= ct_max && total_tok / (ct * 2) >= (int64_t)8 * NC) ct *= 2;
        const int ect = kn.chunk_tokens;
        if (ect >= TOK && ect <= ct_max && ect % TOK == 0 && ect % bs == 0) ct = ect;
        pl->chunk_tokens = ct;
        pl->nc = NC;
        pl->static_pct = kn.static_pct;
        const int64_t ctot_max = (int64_t)B * ipb * ((max_ctx + ct - 1) / ct);
        pl->slots = 2 * (int64_t)NC + ctot_max * (100 - kn.static_pct) / 100 + NC + 2;
        return;
    }
    int64_t st = (max_ctx + num_splits - 1) / num_splits;
    st = ((st + TOK - 1) / TOK) * TOK;   // whole stages, hence whole blocks
    pl->split_tokens = (int)st;
    pl->num_splits = (int)((max_ctx + st - 1) / st);
    if (pl->num_splits < 1) pl->num_splits = 1;
}

size_t plan_workspace(const Plan& pl, int B, int Hq) {
    if (pl.num_splits == 0) {
        return WS_HEADER_BYTES + (size_t)pl.slots * (8 * pl.nt) * (D + 1) * sizeof(float);
    }
    if (pl.num_splits == 1) return 0;
    return (size_t)B * Hq * pl.num_splits * (D + 1) * sizeof(float);
}

// Bytes that serve EVERY batch size 1..B of this geometry.  The dynamic plan's chunk size grows with the batch, so
// its slot count is not monotone in B (a smaller batch may cut finer chunks and need more partial slots): within one
// chunk-size regime the need grows with B, so the maximum sits at B or at the last batch size of a smaller regime.
size_t workspace_bound(int B, int Hq, int Hkv, int bs, int max_nb, int num_splits, const Knobs& kn) {
    Plan pl;
    make_plan(&pl, B, Hkv, Hq / Hkv, bs, max_nb, num_splits, 3, kn);
    size_t best = plan_workspace(pl, B, Hq);
    if (num_splits > 0) return best;                       // grid split: monotone in B
    const int NC = persistent_grid(3);
    const int64_t per_seq = (int64_t)pl.ipb * max_nb * bs;  // streamed tokens one sequence adds
    for (int ct = TOK; ct * 2 <= dyn_ct_max(bs); ct *= 2) {
        // the plan leaves chunk size ct for 2*ct once B * per_seq >= 16 * NC * ct
        const int64_t first_bigger = ((int64_t)16 * NC * ct + per_seq - 1) / per_seq;
        const int64_t b = first_bigger - 1;
        if (b >= 1 && b < B) {
            make_plan(&pl, (int)b, Hkv, Hq / Hkv, bs, max_nb, 0, 3, kn);
            best = std::max(best, plan_workspace(pl, (int)b, Hq));
        }
    }
    return best;
}

// NSTAGE / NSTAGE_D: ring depth of the grid-split / dynamic kernel.  bf16: 3 x 32 KiB, two CTAs per SM.
// fp8 (16 KiB stages): 4 / 3 stages with three CTAs per SM (default), or 6 / 5 with two.
// zero_first: launch the tiny kernel that zeroes the claim counter (stateless API); a b2_decode_plan zeroes it once
// at creation and every merge kernel leaves it at zero, so its calls are two launches.
template <int BS, bool FP8, int NT, int NSTAGE, int NSTAGE_D>
int launch_decode(const CUtensorMap& tmK, const CUtensorMap& tmV, const DecodeParams& p, const Plan& pl,
                  bool zero_first, cudaStream_t st) {
    using C = Cfg<BS, FP8, NT, NSTAGE>;
    if (pl.num_splits == 0) {
        using C = Cfg<BS, FP8, NT, NSTAGE_D>;
        auto kern = paged_decode_dynamic_kernel<BS, FP8, NT, NSTAGE_D>;
        static bool attr_set3 = false;   // per instantiation; idempotent, so a race is harmless
        if (!attr_set3) {
            B2_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::DYN_SMEM_BYTES));
            attr_set3 = true;
        }
        const int grid = persistent_grid(C::CTAS_DYN);
        B2_REQUIRE(grid == pl.nc, B2_EINVAL, "paged_decode: plan made for %d CTAs, kernel runs %d", pl.nc, grid);
        if (zero_first) {
            zero_counter_kernel<<<1, 32, 0, st>>>(p.counter);
            B2_CUDA_OK(cudaGetLastError());
        }
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[0].val.programmaticStreamSerializationAllowed = 1;
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(grid);
        cfg.blockDim = dim3(NTHREADS);
        cfg.dynamicSmemBytes = C::DYN_SMEM_BYTES;
        cfg.stream = st;
        // without the zeroing kernel in front the grid is ordered after the stream's earlier work (the KV write of
        // this step) in the ordinary way: no programmatic overlap with it
        cfg.attrs = at;
        cfg.numAttrs = zero_first ? 1 : 0;
        B2_CUDA_OK(cudaLaunchKernelEx(&cfg, kern, tmK, tmV, p));
        cudaLaunchConfig_t cfg2 = {};
        cfg2.gridDim = dim3(p.Hkv * p.HG, p.B);
        cfg2.blockDim = dim3(128);
        cfg2.stream = st;
        cfg2.attrs = at;
        cfg2.numAttrs = 1;
        B2_CUDA_OK(cudaLaunchKernelEx(&cfg2, merge_chunks_kernel<8 * NT>, p.out, (const float*)p.ws_o,
                                      (const float*)p.ws_lse, p.seq_lens, p.B, p.Hkv, p.G, p.HG, p.out_stride,
                                      p.chunk_tokens, p.static_pct, (uint32_t)grid, p.counter));
        return B2_OK;
    }
    auto kern = paged_decode_split_kernel<BS, FP8, NT, NSTAGE>;
    static bool attr_set2 = false;
    if (!attr_set2) {
        B2_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM_BYTES));
        attr_set2 = true;
    }
    dim3 grid(p.num_splits, p.Hkv * p.HG, p.B);
    {
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[0].val.programmaticStreamSerializationAllowed = 1;
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = grid;
        cfg.blockDim = dim3(NTHREADS);
        cfg.dynamicSmemBytes = C::SMEM_BYTES;
        cfg.stream = st;
        cfg.attrs = at;
        static const int pdl = env_int("B2_DECODE_PDL", 1);      // 0: ordinary stream order (A/B)
        cfg.numAttrs = pdl ? 1 : 0;
        B2_CUDA_OK(cudaLaunchKernelEx(&cfg, kern, tmK, tmV, p));
    }
    if (pl.num_splits > 1) {
        merge_splits_kernel<<<dim3(p.Hq, p.B), D, 0, st>>>(p.out, p.ws_o, p.ws_lse, p.seq_lens, p.Hq, p.num_splits,
                                                           p.split_tokens, p.out_stride);
        B2_CUDA_OK(cudaGetLastError());
    }
    return B2_OK;
}

template <int BS>
int dispatch_decode(const CUtensorMap& tmK, const CUtensorMap& tmV, const DecodeParams& p, const Plan& pl, bool fp8,
                    int fp8_ctas, bool zero_first, cudaStream_t st) {
    if (!fp8)
        return pl.nt == 1 ? launch_decode<BS, false, 1, 3, 3>(tmK, tmV, p, pl, zero_first, st)
                          : launch_decode<BS, false, 2, 3, 3>(tmK, tmV, p, pl, zero_first, st);
    if (ctas_per_sm(true, p.G, fp8_ctas) == 3) return launch_decode<BS, true, 1, 4, 3>(tmK, tmV, p, pl, zero_first, st);
    return pl.nt == 1 ? launch_decode<BS, true, 1, 6, 5>(tmK, tmV, p, pl, zero_first, st)
                      : launch_decode<BS, true, 2, 6, 5>(tmK, tmV, p, pl, zero_first, st);
}

bool shape_ok(int B, int Hq, int Hkv, int d, int bs, int max_nb) {
    return B > 0 && Hq > 0 && Hkv > 0 && Hq % Hkv == 0 && bs > 0 && max_nb > 0 && d == D;
}

// everything of a decode call that does not change from step to step
struct DecodeGeom {
    int Hq, Hkv, bs, max_nb, kv_dtype;
    float scale;
    const float* k_scale;
    const float* v_scale;
    void* workspace;
    size_t workspace_bytes;
};

int check_geom(int Hq, int Hkv, int d, int bs, int max_nb, int64_t num_blocks, const void* k_cache, const void* v_cache,
               int64_t cache_block_stride, int64_t cache_head_stride, float scale, int kv_dtype, const float* k_scale,
               const float* v_scale) {
    B2_REQUIRE(Hq > 0 && Hkv > 0 && bs > 0 && max_nb > 0 && num_blocks > 0, B2_EINVAL,
               "paged_decode: bad sizes Hq=%d Hkv=%d bs=%d max_nb=%d num_blocks=%lld", Hq, Hkv, bs, max_nb,
               (long long)num_blocks);
    B2_REQUIRE(k_cache && v_cache, B2_EINVAL, "paged_decode: null pointer");
    B2_REQUIRE(d == D, B2_EUNSUPPORTED, "paged_decode: head_dim %d unsupported (kernels are built for 128)", d);
    B2_REQUIRE(Hq % Hkv == 0, B2_EINVAL, "paged_decode: Hq=%d not a multiple of Hkv=%d", Hq, Hkv);
    B2_REQUIRE(bs == 8 || bs == 16 || bs == 32, B2_EUNSUPPORTED, "paged_decode: block_size %d unsupported (8/16/32)", bs);
    B2_REQUIRE(kv_dtype == B2_KV_BF16 || kv_dtype == B2_KV_FP8_E4M3, B2_EUNSUPPORTED, "paged_decode: kv_dtype %d", kv_dtype);
    B2_REQUIRE(b2_aligned16(k_cache) && b2_aligned16(v_cache), B2_EINVAL, "paged_decode: pointers must be 16-byte aligned");
    B2_REQUIRE(cache_block_stride % 16 == 0 && cache_head_stride % 16 == 0 && cache_head_stride >= (int64_t)bs * D,
               B2_EINVAL, "paged_decode: cache strides must keep 16-byte alignment");
    B2_REQUIRE(scale > 0.f, B2_EINVAL, "paged_decode: scale must be positive");
    if (kv_dtype == B2_KV_FP8_E4M3) B2_REQUIRE(k_scale && v_scale, B2_EINVAL, "paged_decode: fp8 needs k_scale/v_scale");
    B2_REQUIRE((int64_t)Hkv * ((Hq / Hkv + 15) / 16) <= 65535, B2_EUNSUPPORTED, "paged_decode: grid too large");
    return B2_OK;
}

// the KV write of b2_paged_decode_append (one new token per row), already validated by its entry point
struct AppendArgs {
    const void* k_new;
    const void* v_new;
    const int64_t* slot_mapping;
    void* k_cache;
    void* v_cache;
    int64_t k_token_stride, v_token_stride, cache_block_stride, cache_head_stride;
    int32_t* bt_update;   // device-resident block tables to keep current (decode-plan form), or null
};

// The append rides inside the decode grid when exactly one CTA reads a (row, kv head) -- the single-pass grid with all
// q heads of a kv head in one CTA (G <= 16); every other schedule cuts a row over several CTAs, so the write is
// launched in front of it (b2_reshape_and_cache) and the call is an ordinary b2_paged_decode.
bool append_is_fused(const Plan& pl) { return pl.num_splits == 1 && pl.hg == 1; }

// one call: validate the per-step arguments, plan, launch
int run_decode(const DecodeGeom& g, const CacheMaps& maps, const Knobs& kn, bool zero_first, void* out, const void* q,
               const int32_t* block_tables, const int32_t* seq_lens, int B, int64_t q_stride, int64_t out_stride,
               int num_splits, cudaStream_t st, const AppendArgs* ap = nullptr) {
    B2_REQUIRE(out && q && block_tables && seq_lens, B2_EINVAL, "paged_decode: null pointer");
    B2_REQUIRE(b2_aligned16(out) && b2_aligned16(q), B2_EINVAL, "paged_decode: pointers must be 16-byte aligned");
    B2_REQUIRE(q_stride % 8 == 0 && out_stride % 8 == 0 && q_stride >= (int64_t)g.Hq * D && out_stride >= (int64_t)g.Hq * D,
               B2_EINVAL, "paged_decode: q/out strides");
    B2_REQUIRE(num_splits >= 0, B2_EINVAL, "paged_decode: num_splits must be >= 0");
    B2_REQUIRE(B <= 65535, B2_EUNSUPPORTED, "paged_decode: grid too large");
    B2_REQUIRE((int64_t)B * g.max_nb * g.Hkv * ((g.Hq / g.Hkv + 7) / 8) < (1LL << 31), B2_EUNSUPPORTED,
               "paged_decode: more than 2^31 block units");
    const int G = g.Hq / g.Hkv;
    const bool fp8 = g.kv_dtype == B2_KV_FP8_E4M3;
    Plan pl;
    make_plan(&pl, B, g.Hkv, G, g.bs, g.max_nb, num_splits, ctas_per_sm(fp8, G, kn.fp8_ctas), kn);

    DecodeParams p;
    p.out = static_cast<__nv_bfloat16*>(out);
    p.q = static_cast<const __nv_bfloat16*>(q);
    p.block_tables = block_tables;
    p.seq_lens = seq_lens;
    p.k_scale = g.k_scale;
    p.v_scale = g.v_scale;
    p.B = B; p.Hq = g.Hq; p.Hkv = g.Hkv; p.G = G; p.HG = pl.hg; p.max_nb = g.max_nb;
    p.num_splits = pl.num_splits; p.split_tokens = pl.split_tokens;
    p.q_stride = q_stride; p.out_stride = out_stride;
    p.scale_log2 = g.scale * LOG2E;
    p.ws_o = nullptr; p.ws_lse = nullptr;
    p.counter = nullptr;
    p.append = 0;
    p.k_new = p.v_new = nullptr;
    p.slot_mapping = nullptr;
    p.k_cache_w = p.v_cache_w = nullptr;
    p.k_new_stride = p.v_new_stride = p.cache_block_stride = p.cache_head_stride = 0;
    p.bt_update = nullptr;
    if (ap) {
        if (append_is_fused(pl)) {
            p.append = kn.append_mode;
            p.k_new = static_cast<const __nv_bfloat16*>(ap->k_new);
            p.v_new = static_cast<const __nv_bfloat16*>(ap->v_new);
            p.slot_mapping = ap->slot_mapping;
            p.k_cache_w = ap->k_cache;
            p.v_cache_w = ap->v_cache;
            p.k_new_stride = ap->k_token_stride;
            p.v_new_stride = ap->v_token_stride;
            p.cache_block_stride = ap->cache_block_stride;
            p.cache_head_stride = ap->cache_head_stride;
            p.bt_update = ap->bt_update;
        }
    }
    p.chunk_tokens = pl.chunk_tokens;
    p.static_pct = pl.static_pct;
    p.two_box = maps.two_box;
    // a plan handle keeps its zeroed claim counter in the workspace header: the partials of every schedule start
    // behind it (the stateless call re-zeroes the counter itself and uses the header only for the dynamic schedule)
    const bool header = pl.num_splits == 0 || !zero_first;
    const size_t body = plan_workspace(pl, B, g.Hq) - (pl.num_splits == 0 ? WS_HEADER_BYTES : 0);
    if (body) {
        const size_t need = body + (header ? WS_HEADER_BYTES : 0);
        B2_REQUIRE(g.workspace && g.workspace_bytes >= need && b2_aligned16(g.workspace), B2_ENOMEM,
                   "paged_decode: workspace %zu bytes < required %zu", g.workspace_bytes, need);
        char* base = static_cast<char*>(g.workspace);
        if (pl.num_splits == 0) {
            B2_REQUIRE(pl.slots < (1LL << 27), B2_EUNSUPPORTED, "paged_decode: %lld partial slots", (long long)pl.slots);
            p.counter = reinterpret_cast<uint32_t*>(base);
        }
        if (header) base += WS_HEADER_BYTES;
        p.ws_o = reinterpret_cast<float*>(base);
        p.ws_lse = p.ws_o + body / sizeof(float) / (D + 1) * D;
    }
    if (ap && !p.append) {
        // a schedule that cuts rows over several CTAs: the KV write goes in front of the grid (after every argument
        // check above, so a refused call has written nothing)
        const int rc = ap->bt_update
            ? b2_reshape_and_cache_decode(ap->k_new, ap->v_new, ap->k_cache, ap->v_cache, ap->slot_mapping,
                                          ap->bt_update, seq_lens, B, g.Hkv, D, g.bs, g.max_nb, ap->k_token_stride,
                                          ap->v_token_stride, ap->cache_block_stride, ap->cache_head_stride,
                                          g.kv_dtype, g.k_scale, g.v_scale, st)
            : b2_reshape_and_cache(ap->k_new, ap->v_new, ap->k_cache, ap->v_cache, ap->slot_mapping, B, g.Hkv, D,
                                   g.bs, ap->k_token_stride, ap->v_token_stride, ap->cache_block_stride,
                                   ap->cache_head_stride, g.kv_dtype, g.k_scale, g.v_scale, st);
        if (rc != B2_OK) return rc;
    }
    if (g.bs == 8) return dispatch_decode<8>(maps.k, maps.v, p, pl, fp8, kn.fp8_ctas, zero_first, st);
    if (g.bs == 16) return dispatch_decode<16>(maps.k, maps.v, p, pl, fp8, kn.fp8_ctas, zero_first, st);
    return dispatch_decode<32>(maps.k, maps.v, p, pl, fp8, kn.fp8_ctas, zero_first, st);
}

}  // namespace

// generic bf16 tiled map with 128-byte swizzle (prefill_tc5.cu builds its Q and half-tile cache maps with it)
int b2_internal_encode_bf16_sw128(CUtensorMap* tm, const void* ptr, int rank, const uint64_t* dims,
                                  const uint64_t* strides_bytes, const uint32_t* box) {
    EncodeTiledFn enc = get_encode_tiled();
    B2_REQUIRE(enc != nullptr, B2_ECUDA, "cuTensorMapEncodeTiled entry point not available");
    cuuint64_t d[5], st[4];
    cuuint32_t b[5], es[5];
    for (int i = 0; i < rank; ++i) { d[i] = dims[i]; b[i] = box[i]; es[i] = 1; }
    for (int i = 0; i + 1 < rank; ++i) st[i] = strides_bytes[i];
    CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, rank, const_cast<void*>(ptr), d, st, b, es,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    B2_REQUIRE(r == CUDA_SUCCESS, B2_ECUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
    return B2_OK;
}

// shared with prefill_attention.cu (internal to the library's objects, not exported)
int b2_internal_make_cache_tmap(CUtensorMap* tm, const void* cache, int kv_dtype, int64_t num_blocks, int Hkv, int bs,
                                 int64_t block_stride, int64_t head_stride) {
    return make_cache_tmap(tm, cache, kv_dtype, num_blocks, Hkv, bs, block_stride, head_stride);
}

extern "C" int b2_paged_decode_auto_splits(int B, int Hq, int Hkv, int d, int bs, int max_nb, int kv_dtype) {
    (void)kv_dtype;
    if (!shape_ok(B, Hq, Hkv, d, bs, max_nb)) return B2_EINVAL;
    return 0;   // num_splits = 0 selects the persistent schedule: no fixed split count
}

// Host-side schedule choice from the scheduler's own (host) copy of seq_lens -- the same role max_seq_len plays
// for vLLM's v1/v2 choice.  Model: CTAs of the single-pass grid (one per item, launched in grid order onto
// 2 x SMs slots, greedily) each stream `len + overhead` tokens; bf16: HBM saturates with ~3/4 of the slots busy, so a
// slot can run up to 1.37x its fair share.  The single-pass grid wins (no workspace, no merge, no schedule prologue)
// when that estimate beats the dynamic schedule's (streamed tokens / slots, + 4 %); short or ragged items lose to it.
extern "C" int b2_paged_decode_plan_splits(const int32_t* seq_lens_host, int B, int Hq, int Hkv, int d, int bs,
                                           int kv_dtype) {
    if (!seq_lens_host || !shape_ok(B, Hq, Hkv, d, bs, 1)) return B2_EINVAL;
    const int G = Hq / Hkv;
    const int hpc = G <= 8 ? 8 : 16;
    const int ipb = Hkv * ((G + hpc - 1) / hpc);
    static const int fp8_ctas = read_knobs().fp8_ctas;   // the planner runs every step: no getenv on that path
    const int NC = persistent_grid(ctas_per_sm(kv_dtype == B2_KV_FP8_E4M3, G, fp8_ctas));
    constexpr int64_t OVERHEAD_TOK = 128;   // CTA launch + pipeline fill, in streamed-token equivalents
    // The ipb items of a sequence are equal and adjacent in grid order: they are placed together on a group of ipb
    // slots (NC / ipb groups), which keeps this at one heap operation per sequence -- it runs on the host every step.
    // (fewer items than slots is not special: 256 equal items on 296 slots stream at 6.7 TB/s in one wave, the
    // dynamic schedule at 6.1 -- BASELINE configs[3] at tp = 8; the model below rejects the really short grids)
    const int n_groups = NC / ipb > 0 ? NC / ipb : 1;
    static thread_local std::vector<int64_t> slot;     // min-heap of group finish times (reused: no allocation per step)
    slot.assign((size_t)n_groups, 0);
    auto cmp = [](int64_t a, int64_t b) { return a > b; };
    int64_t total = 0, streamed = 0, makespan = 0;
    // uniform batches (the common steady state) need no heap: round-robin placement is exact
    bool uniform = true;
    for (int b = 1; b < B; ++b) uniform = uniform && seq_lens_host[b] == seq_lens_host[0];
    if (uniform) {
        const int64_t L = seq_lens_host[0] > 0 ? seq_lens_host[0] : 0;
        const int64_t rounds = ((int64_t)B + n_groups - 1) / n_groups;
        makespan = rounds * (L + OVERHEAD_TOK);
        total = (int64_t)B * (L + OVERHEAD_TOK) * ipb;
        streamed = (int64_t)B * L * ipb;
    } else {
        for (int b = 0; b < B; ++b) {
            const int64_t L = seq_lens_host[b] > 0 ? seq_lens_host[b] : 0;
            std::pop_heap(slot.begin(), slot.end(), cmp);
            int64_t& t = slot.back();
            t += L + OVERHEAD_TOK;
            if (t > makespan) makespan = t;
            std::push_heap(slot.begin(), slot.end(), cmp);
            total += (L + OVERHEAD_TOK) * ipb;
            streamed += L * ipb;
        }
    }
    const double ideal = (double)total / NC;
    // fp8 (three CTAs per SM) is paced by its consumer warps, not by HBM: an idle slot is lost throughput, so the
    // single-pass grid only wins when its waves are nearly full (measured: config 3 dynamic 6.4 vs grid 6.1 TB/s)
    const double slack = kv_dtype == B2_KV_FP8_E4M3 ? 1.10 : 1.37;
    const double est = std::max(ideal, (double)makespan / slack);
    // the dynamic schedule keeps its pipelines full across items (no per-item ramp) at ~4 % for chunk boundaries, the
    // claim counter and the merge: measured -4 % on uniform config 2, +9 % / +7 % / +56 % on log-uniform 128..ctx
    // batches of configs 4 / 2 / 3 (scripts/gpu_ragged.sh)
    const double est_dynamic = 1.04 * (double)streamed / NC;
    return est <= est_dynamic ? 1 : 0;
}

extern "C" size_t b2_paged_decode_workspace_bytes(int B, int Hq, int Hkv, int d, int bs, int max_nb,
                                                  int num_splits) {
    if (!shape_ok(B, Hq, Hkv, d, bs, max_nb)) return 0;
    return workspace_bound(B, Hq, Hkv, bs, max_nb, num_splits, read_knobs());   // no kv_dtype here: the bound for every variant
}

extern "C" int b2_paged_decode(void* out, const void* q, const void* k_cache, const void* v_cache,
                               const int32_t* block_tables, const int32_t* seq_lens, int B, int Hq,
                               int Hkv, int d, int bs, int max_nb, int64_t num_blocks,
                               int64_t q_stride, int64_t out_stride, int64_t cache_block_stride,
                               int64_t cache_head_stride, float scale, int kv_dtype,
                               const float* k_scale, const float* v_scale, void* workspace,
                               size_t workspace_bytes, int num_splits, void* stream) {
    B2_REQUIRE(B >= 0, B2_EINVAL, "paged_decode: bad batch size %d", B);
    if (B == 0) return B2_OK;
    int rc = check_geom(Hq, Hkv, d, bs, max_nb, num_blocks, k_cache, v_cache, cache_block_stride, cache_head_stride, scale,
                        kv_dtype, k_scale, v_scale);
    if (rc != B2_OK) return rc;
    const Knobs kn = read_knobs();
    CacheMaps maps;
    rc = make_decode_maps(&maps, k_cache, v_cache, kv_dtype, num_blocks, Hkv, bs, cache_block_stride, cache_head_stride,
                          kn.two_box != 0);
    if (rc != B2_OK) return rc;
    const DecodeGeom g{Hq, Hkv, bs, max_nb, kv_dtype, scale, k_scale, v_scale, workspace, workspace_bytes};
    return run_decode(g, maps, kn, /*zero_first=*/true, out, q, block_tables, seq_lens, B, q_stride, out_stride, num_splits,
                      static_cast<cudaStream_t>(stream));
}

// b2_paged_decode_append: the decode step's KV write and its attention in ONE call -- and, for the single-pass grid
// (num_splits = 1, G <= 16), in ONE launch: see append_row / append_is_fused.
extern "C" int b2_paged_decode_append_fused(int Hq, int Hkv, int num_splits) {
    if (Hq <= 0 || Hkv <= 0 || Hq % Hkv) return B2_EINVAL;
    return num_splits == 1 && Hq / Hkv <= 16 ? 1 : 0;
}

extern "C" int b2_paged_decode_append(void* out, const void* q, const void* k_new, const void* v_new,
                                      const int64_t* slot_mapping, int64_t k_token_stride, int64_t v_token_stride,
                                      void* k_cache, void* v_cache, const int32_t* block_tables, const int32_t* seq_lens,
                                      int B, int Hq, int Hkv, int d, int bs, int max_nb, int64_t num_blocks,
                                      int64_t q_stride, int64_t out_stride, int64_t cache_block_stride,
                                      int64_t cache_head_stride, float scale, int kv_dtype, const float* k_scale,
                                      const float* v_scale, void* workspace, size_t workspace_bytes, int num_splits,
                                      void* stream) {
    B2_REQUIRE(B >= 0, B2_EINVAL, "paged_decode_append: bad batch size %d", B);
    if (B == 0) return B2_OK;
    int rc = check_geom(Hq, Hkv, d, bs, max_nb, num_blocks, k_cache, v_cache, cache_block_stride, cache_head_stride, scale,
                        kv_dtype, k_scale, v_scale);
    if (rc != B2_OK) return rc;
    // the same argument rules as b2_reshape_and_cache
    B2_REQUIRE(k_new && v_new && slot_mapping, B2_EINVAL, "paged_decode_append: null pointer");
    B2_REQUIRE(b2_aligned16(k_new) && b2_aligned16(v_new) && k_token_stride % 8 == 0 && v_token_stride % 8 == 0, B2_EINVAL,
               "paged_decode_append: k_new / v_new rows must keep 16-byte alignment");
    B2_REQUIRE(k_token_stride >= (int64_t)Hkv * d && v_token_stride >= (int64_t)Hkv * d, B2_EINVAL,
               "paged_decode_append: k_new / v_new strides overlap");
    const Knobs kn = read_knobs();
    CacheMaps maps;
    rc = make_decode_maps(&maps, k_cache, v_cache, kv_dtype, num_blocks, Hkv, bs, cache_block_stride, cache_head_stride,
                          kn.two_box != 0);
    if (rc != B2_OK) return rc;
    const DecodeGeom g{Hq, Hkv, bs, max_nb, kv_dtype, scale, k_scale, v_scale, workspace, workspace_bytes};
    const AppendArgs ap{k_new, v_new, slot_mapping, k_cache, v_cache, k_token_stride, v_token_stride, cache_block_stride,
                        cache_head_stride, nullptr};
    return run_decode(g, maps, kn, /*zero_first=*/true, out, q, block_tables, seq_lens, B, q_stride, out_stride, num_splits,
                      static_cast<cudaStream_t>(stream), &ap);
}

// ---------------------------------------------------------------------------------------------
// b2_paged_decode_static: b2_paged_decode for callers that may not allocate (CUDA graph capture, tight serving
// loops, harnesses that forbid cudaMalloc).  The split-K partials and the claim counter live in a file-scope device
// array of B2_STATIC_SCRATCH_BYTES that every GPU loading this library holds until exit.  No counterpart in the
// reference.  Nothing here allocates: no cudaMalloc / cudaFree, no
// allocating library call.  Every call on a GPU shares the ONE array (it keeps the previous call's partials; the
// claim counter is re-zeroed on `stream` by the stateless path, and partials are always written before they are
// read), so two calls whose kernels overlap on the GPU (different streams) race.
// ---------------------------------------------------------------------------------------------
static __device__ __align__(256) unsigned char g_static_scratch[B2_STATIC_SCRATCH_BYTES];

namespace {
// `g_static_scratch` in host code is the host-side shadow of the symbol, not a device address: ask the runtime for
// the address once per device, with that device current.
int static_scratch_ptr(void** out) {
    static std::mutex mu;
    static void* ptrs[64] = {};
    int dev = -1;
    B2_REQUIRE(cudaGetDevice(&dev) == cudaSuccess && dev >= 0 && dev < 64, B2_ECUDA,
               "paged_decode_static: cudaGetDevice failed or device index %d out of range", dev);
    std::lock_guard<std::mutex> lock(mu);
    if (!ptrs[dev]) {
        void* ptr = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&ptr, g_static_scratch);
        B2_REQUIRE(e == cudaSuccess && ptr != nullptr, B2_ECUDA, "paged_decode_static: cudaGetSymbolAddress: %s",
                   cudaGetErrorString(e));
        ptrs[dev] = ptr;
    }
    *out = ptrs[dev];
    return B2_OK;
}
}  // namespace

extern "C" int b2_paged_decode_static(void* out, const void* q, const void* k_cache, const void* v_cache,
                                      const int32_t* block_tables, const int32_t* seq_lens, int B, int Hq,
                                      int Hkv, int d, int bs, int max_nb, int64_t num_blocks,
                                      int64_t q_stride, int64_t out_stride, int64_t cache_block_stride,
                                      int64_t cache_head_stride, float scale, int kv_dtype,
                                      const float* k_scale, const float* v_scale, int num_splits, void* stream) {
    B2_REQUIRE(B >= 0, B2_EINVAL, "paged_decode_static: bad batch size %d", B);
    if (B == 0) return B2_OK;
    int rc = check_geom(Hq, Hkv, d, bs, max_nb, num_blocks, k_cache, v_cache, cache_block_stride, cache_head_stride, scale,
                        kv_dtype, k_scale, v_scale);
    if (rc != B2_OK) return rc;
    B2_REQUIRE(num_splits >= 0, B2_EINVAL, "paged_decode_static: num_splits must be >= 0");
    const Knobs kn = read_knobs();
    {   // what THIS call needs (the same plan run_decode makes), checked against the built-in maximum first
        Plan pl;
        make_plan(&pl, B, Hkv, Hq / Hkv, bs, max_nb, num_splits,
                  ctas_per_sm(kv_dtype == B2_KV_FP8_E4M3, Hq / Hkv, kn.fp8_ctas), kn);
        const size_t need = plan_workspace(pl, B, Hq);
        B2_REQUIRE(need <= (size_t)B2_STATIC_SCRATCH_BYTES, B2_ENOMEM,
                   "paged_decode_static: B=%d Hq=%d max_nb=%d num_splits=%d needs %zu bytes of scratch, the built-in array "
                   "holds %zu (B2_STATIC_SCRATCH_BYTES); use b2_paged_decode with a caller-owned workspace",
                   B, Hq, max_nb, num_splits, need, (size_t)B2_STATIC_SCRATCH_BYTES);
    }
    void* scratch = nullptr;
    rc = static_scratch_ptr(&scratch);
    if (rc != B2_OK) return rc;
    CacheMaps maps;
    rc = make_decode_maps(&maps, k_cache, v_cache, kv_dtype, num_blocks, Hkv, bs, cache_block_stride, cache_head_stride,
                          kn.two_box != 0);
    if (rc != B2_OK) return rc;
    const DecodeGeom g{Hq, Hkv, bs, max_nb, kv_dtype, scale, k_scale, v_scale, scratch, (size_t)B2_STATIC_SCRATCH_BYTES};
    return run_decode(g, maps, kn, /*zero_first=*/true, out, q, block_tables, seq_lens, B, q_stride, out_stride, num_splits,
                      static_cast<cudaStream_t>(stream));
}

// ---------------------------------------------------------------------------------------------
// b2_decode_plan: the per-engine handle.  Everything that does not change between decode steps -- the two encoded
// tensor maps (a driver call each), the environment knobs, the zeroed claim counter -- is set up once, so a step is
// argument checks + the launches, with no driver call and no getenv on the way (host memory only; nothing on the device
// is owned by the handle).
// ---------------------------------------------------------------------------------------------
struct b2_decode_plan {
    DecodeGeom g;
    CacheMaps maps;
    Knobs kn;
    int max_B;
    void* k_cache;        // for b2_decode_plan_run_append (the step's KV write goes into the caches the plan reads)
    void* v_cache;
    int64_t cache_block_stride, cache_head_stride;
};

extern "C" b2_decode_plan* b2_decode_plan_create(const void* k_cache, const void* v_cache, int max_B, int Hq, int Hkv, int d,
                                                 int bs, int max_nb, int64_t num_blocks, int64_t cache_block_stride,
                                                 int64_t cache_head_stride, float scale, int kv_dtype,
                                                 const float* k_scale, const float* v_scale, void* workspace,
                                                 size_t workspace_bytes, void* stream) {
    if (max_B <= 0) {
        b2_set_error("decode_plan_create: bad max_B %d", max_B);
        return nullptr;
    }
    if (check_geom(Hq, Hkv, d, bs, max_nb, num_blocks, k_cache, v_cache, cache_block_stride, cache_head_stride, scale,
                   kv_dtype, k_scale, v_scale) != B2_OK)
        return nullptr;
    const Knobs kn = read_knobs();
    const size_t need = workspace_bound(max_B, Hq, Hkv, bs, max_nb, 0, kn);
    if (!workspace || workspace_bytes < need || !b2_aligned16(workspace)) {
        b2_set_error("decode_plan_create: workspace %zu bytes < required %zu (b2_paged_decode_workspace_bytes)",
                     workspace_bytes, need);
        return nullptr;
    }
    b2_decode_plan* pl = new (std::nothrow) b2_decode_plan();
    if (!pl) {
        b2_set_error("decode_plan_create: out of host memory");
        return nullptr;
    }
    pl->g = DecodeGeom{Hq, Hkv, bs, max_nb, kv_dtype, scale, k_scale, v_scale, workspace, workspace_bytes};
    pl->kn = kn;
    pl->max_B = max_B;
    pl->k_cache = const_cast<void*>(k_cache);
    pl->v_cache = const_cast<void*>(v_cache);
    pl->cache_block_stride = cache_block_stride;
    pl->cache_head_stride = cache_head_stride;
    if (make_decode_maps(&pl->maps, k_cache, v_cache, kv_dtype, num_blocks, Hkv, bs, cache_block_stride,
                         cache_head_stride, kn.two_box != 0) != B2_OK) {
        delete pl;
        return nullptr;
    }
    // the claim counter starts at zero; every dynamic-schedule call leaves it at zero again (merge_chunks_kernel)
    if (cudaMemsetAsync(workspace, 0, WS_HEADER_BYTES, static_cast<cudaStream_t>(stream)) != cudaSuccess) {
        b2_set_error("decode_plan_create: cudaMemsetAsync of the workspace header failed");
        delete pl;
        return nullptr;
    }
    return pl;
}

extern "C" void b2_decode_plan_destroy(b2_decode_plan* pl) { delete pl; }

extern "C" int b2_decode_plan_run(b2_decode_plan* pl, void* out, const void* q, const int32_t* block_tables,
                                  const int32_t* seq_lens, int B, int64_t q_stride, int64_t out_stride, int num_splits,
                                  void* stream) {
    B2_REQUIRE(pl != nullptr, B2_EINVAL, "decode_plan_run: null plan");
    B2_REQUIRE(B >= 0 && B <= pl->max_B, B2_EINVAL, "decode_plan_run: B=%d outside [0, %d]", B, pl->max_B);
    if (B == 0) return B2_OK;
    return run_decode(pl->g, pl->maps, pl->kn, /*zero_first=*/false, out, q, block_tables, seq_lens, B, q_stride,
                      out_stride, num_splits, static_cast<cudaStream_t>(stream));
}

#ifdef B2_TRACE
extern "C" B2_API int b2_debug_read_trace(unsigned long long* dst, int n_words) {
    if (n_words > 2048 * 8) n_words = 2048 * 8;
    return cudaMemcpyFromSymbol(dst, g_trace, (size_t)n_words * 8) == cudaSuccess ? 0 : B2_ECUDA;
}
extern "C" B2_API int b2_debug_clear_trace(void) {
    static unsigned long long zeros[2048 * 8];
    return cudaMemcpyToSymbol(g_trace, zeros, sizeof zeros) == cudaSuccess ? 0 : B2_ECUDA;
}
#endif

// The decode step of an engine with RESIDENT block tables in one call: b2_reshape_and_cache_decode (KV write + table
// entry of the new token) followed by b2_decode_plan_run -- one launch for the single-pass grid (see append_is_fused).
extern "C" int b2_decode_plan_run_append(b2_decode_plan* pl, void* out, const void* q, const void* k_new, const void* v_new,
                                         const int64_t* slot_mapping, int64_t k_token_stride, int64_t v_token_stride,
                                         int32_t* block_tables, const int32_t* seq_lens, int B, int64_t q_stride,
                                         int64_t out_stride, int num_splits, void* stream) {
    B2_REQUIRE(pl != nullptr, B2_EINVAL, "decode_plan_run_append: null plan");
    B2_REQUIRE(B >= 0 && B <= pl->max_B, B2_EINVAL, "decode_plan_run_append: B=%d outside [0, %d]", B, pl->max_B);
    if (B == 0) return B2_OK;
    B2_REQUIRE(k_new && v_new && slot_mapping && block_tables, B2_EINVAL, "decode_plan_run_append: null pointer");
    B2_REQUIRE(b2_aligned16(k_new) && b2_aligned16(v_new) && k_token_stride % 8 == 0 && v_token_stride % 8 == 0, B2_EINVAL,
               "decode_plan_run_append: k_new / v_new rows must keep 16-byte alignment");
    B2_REQUIRE(k_token_stride >= (int64_t)pl->g.Hkv * D && v_token_stride >= (int64_t)pl->g.Hkv * D, B2_EINVAL,
               "decode_plan_run_append: k_new / v_new strides overlap");
    const AppendArgs ap{k_new, v_new, slot_mapping, pl->k_cache, pl->v_cache, k_token_stride, v_token_stride,
                        pl->cache_block_stride, pl->cache_head_stride, block_tables};
    return run_decode(pl->g, pl->maps, pl->kn, /*zero_first=*/false, out, q, block_tables, seq_lens, B, q_stride,
                      out_stride, num_splits, static_cast<cudaStream_t>(stream), &ap);
}
