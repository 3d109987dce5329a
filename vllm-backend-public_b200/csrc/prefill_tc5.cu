// prefill_tc5.cu -- prompt-chunk attention over the paged KV cache on the 5th-gen tensor cores (sm_100a):
// tcgen05.mma with TMEM accumulators, TMA-staged paged K/V, warp-specialised (FlashAttention-4 style dataflow).
// Same contract as prefill_attention.cu (include/b200attn.h `b2_prefill_attention`); this is the default kernel,
// the mma.sync one is kept as the A/B (B2_PREFILL_IMPL=mma).
//
// One CTA = 256 query tokens of one q head = two 128-row tiles A and B, 11 warps:
//   warps 0-3  softmax of tile A   (one thread = one query row = one TMEM lane)
//   warps 4-7  softmax of tile B
//   warp  8    TMA producer: Q (two 128x128 tiles) once, then the chunk's block table -> K tiles of 128 tokens
//   warp  10   TMA producer: V tiles                            (one box of bs x 128 B per lane in both)
//   warp  9    MMA issuer (warp-uniform loop, one elected lane per instruction group): S = Q.K^T and O += P.V as
//              tcgen05.mma 128x128x16, bf16 -> fp32 in TMEM
// TMEM (512 columns x 128 lanes): S_A [0,128) S_B [128,256) O_A [256,384) O_B [384,512); P (bf16 pairs) overwrites the
// first 64 columns of its S tile and is fed back as the A operand straight from TMEM, so P never touches shared memory.
// Issue order  QK_A(0) QK_B(0) | PV_A(j) QK_A(j+1) PV_B(j) QK_B(j+1) ...  lets the softmax of one tile run under the
// MMAs of the other; P.V(j) is issued in two halves as the softmax hands over keys 0-63 and 64-127.  Rescaling of O is lazy: a row keeps its stale maximum until the true one exceeds it by 2^8.
//
// Round 2, what bounded the kernel and what did not (8k chunk over an 8k prefix, Llama-3-8B heads, block_size 16;
// clock64 timeline from `make trace` + scripts/trace_prefill.py, profiles/r2_trace_prefill_issuer*.txt):
//   1055 TF/s  round-1 kernel.  ncu: tensor pipe 48 %, XU pipe 48 %, softmax warps waiting for S in 56 % of samples.
//   not it     softmax latency hiding: TWO threads per query row (scripts/experiments_prefill_two_threads_per_row.cu.txt,
//              profiles/r2_ncu_prefill_two_threads_per_row.md) 1038; maximum pass off the chain (speculative maximum) 947;
//              polynomial exp2 for 1-3 of every 4 weights (B2_PREFILL_POLY) slower then and now (1371 / 1266 / 1164).
//   not it     shared-memory operand bandwidth of cta_group::1 (an earlier reading of the timeline, WRONG):
//              scripts/umma_rate.cu shows SS, TS, MN-major B, the kernel's own P.V | Q.K^T mix and all of them under
//              concurrent TMEM traffic at 64.1 cycles per 128x128x16 = full rate (profiles/r2_umma_rate.txt).
//   +8 %       the producer: ONE lane issued the 32 TMA boxes of a step (block_size 16: 2 KiB boxes) one after the other,
//              ~70 cycles each.  Now warp 8 streams K and warp 10 streams V, one box per lane, and the V stage is released
//              before the next Q.K^T is issued.  Tiles land ~1500 cycles after their stage frees (1140 TF/s).
//   +21 %      the issuer: the loop ran inside `if (lane == 0)`, so every operand lived in a per-thread register and ptxas
//              wrapped EVERY tcgen05.mma in an ELECT / R2UR.BROADCAST / BRA.U.ANY loop, ~12 dependent instructions that
//              the issuing thread -- sharing its scheduler with two softmax warps -- needed 80-105 cycles for, where the
//              pipe needs 64.  The whole warp now runs the loop with warp-uniform operands and elect.sync picks the lane
//              per instruction group (what CUTLASS does): 1-3 SASS instructions between UTCHMMAs (1375 TF/s).
//   +2.5 %     P is handed over in two halves, so P.V of keys 0-63 runs under the exponentials of keys 64-127 (1409 TF/s,
//              0.845 of the measured cuBLAS bf16 burst).  Quarters would cost the softmax as much as they save the chain.
//   small      launches of less than one wave at 256 rows per CTA (a 512-token chunk over a long prefix) run 128 rows per
//              CTA; with a single tile the tensor pipe idled during the tile's own softmax.  The SPLIT kernel (below)
//              gives both slots the same 128 rows and interleaves the key tiles between them: 512 / 3.5k 55.7 -> 36-41 us
//              (578 -> 760-880 TF/s), and at 8k / 8k it is within 6 % of the 256-row kernel (B2_PREFILL_ROWS=128).
// What is left is the dependency chain of a tile, softmax(j) -> P.V(j) -> Q.K^T(j+1) -> softmax(j+1): ~1680 cycles of
// softmax (TMEM load 30, maximum 300, 16 384 exp2 at 16 per clock + pack 1310, hand-over 40) + ~1170 until the next S
// (4 P.V + 8 Q.K^T instructions = 768 pipe cycles, the rest barrier and commit latency) = the measured period of
// 2850 cycles per 128-key step, in which the tensor pipe and the XU pipe are each busy 2048 (72 %).  P overwrites S, so
// a tile cannot overlap its own chain; TMEM (2 x S + 2 x O = 512 columns) has no room for a second S per tile.
//
// Shared memory (1024-byte aligned, 128-byte swizzle, the canonical UMMA layouts):
//   Q_A, Q_B     2 x [2 halves][128 rows][64 el]           K-major operand A           64 KiB
//   stage s      K: [2 halves][128 tok][64 el]  K-major operand B (QK^T)
//                V: [2 halves][128 tok][64 el]  MN-major operand B (P.V)               2 x 64 KiB
// One TMA box = one 64-element half of one (block, kv head) tile: bs rows x 128 B, contiguous in HBM.
#include "b2_common.cuh"

#ifdef B2_TRACE
// diagnostic build (make trace): clock64 stamps of the softmax warps 0 (tile A) and 4 (tile B) of CTA (0, 0), lane 0,
// for the first 32 steps: [tile][step][S ready | S in registers | maximum known | P computed | P handed to the MMA warp]
__device__ unsigned long long g_pf_trace[2][32][16];  // 5..8: the MMA issuer -- P seen | P.V issued | Q.K^T issued | V, K waits done; 9..11: P handed over by warps 1..3 of the tile; 12, 13: producer of K (x = 0) / V (x = 1) -- stage free | boxes issued; 14: K(j+1) (x = 0) / V(j) (x = 1) seen full by the issuer; 15: ... landed (watcher warp 11)
#define PF_STAMP(k) do { if (trace_on && j < 32) g_pf_trace[x][j][k] = clock64(); } while (0)
#define PF_ISSUE_STAMP(k) do { if ((threadIdx.x & 31) == 0 && blockIdx.x == 0 && blockIdx.y == 1 && j < 32) g_pf_trace[x][j][k] = clock64(); } while (0)
#else
#define PF_STAMP(k) do { } while (0)
#define PF_ISSUE_STAMP(k) do { } while (0)
#endif

namespace {

constexpr int D = 128;
constexpr int BM = 128;             // rows per q tile
constexpr int BN = 128;             // kv tokens per stage
constexpr int NSTAGE = 2;
constexpr int NTHREADS = 384;       // 3 warpgroups: softmax A, softmax B, {TMA producer, MMA issuer, 2 idle warps}
constexpr int SUB_BYTES = 128 * 128;               // one [128 rows][64 el] sub-tile
constexpr int TILE_BYTES = 2 * SUB_BYTES;          // 32 KiB: a 128 x 128 bf16 tile
constexpr int Q_BYTES = 2 * TILE_BYTES;
constexpr int STAGE_BYTES = 2 * TILE_BYTES;        // K | V
constexpr int SMEM_BYTES = Q_BYTES + NSTAGE * STAGE_BYTES + 256;
constexpr float LOG2E = 1.4426950408889634f;
constexpr float RESCALE_THRESHOLD = 8.f;           // log2 units

// barrier slots (8 bytes each) after the tiles
enum { BAR_Q = 0, BAR_K = 1, BAR_V = BAR_K + NSTAGE, BAR_KE = BAR_V + NSTAGE, BAR_VE = BAR_KE + NSTAGE,
       BAR_S = BAR_VE + NSTAGE, BAR_P = BAR_S + 2, BAR_P2 = BAR_P + 2, BAR_PV = BAR_P2 + 2, BAR_COUNT = BAR_PV + 2 };
static_assert(BAR_COUNT * 8 + 16 <= 256, "barrier area");

struct Tc5Params {
    __nv_bfloat16* out;
    const int32_t* block_tables;
    const int32_t* cu_q;
    const int32_t* ctx_start;
    int n_chunks, Hq, G, bs, max_nb;
    int rows_per_cta;      // 256 (tiles A and B) or 128 (tile A only: small batches fill the 148 SMs better)
    int64_t out_stride;
    float scale_log2;
    const float* k_scale;  // optional device scalars: the cache holds K / k_scale and V / v_scale (dequantised fp8)
    const float* v_scale;
};

// ---- tcgen05 wrappers --------------------------------------------------------------------------
__device__ __forceinline__ uint64_t smem_desc(uint32_t addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    return (uint64_t)((addr & 0x3FFFFu) >> 4) | ((uint64_t)(lbo_bytes >> 4) << 16) | ((uint64_t)(sbo_bytes >> 4) << 32) |
           (1ull << 46) | (2ull << 61);             // version 1 (Blackwell), SWIZZLE_128B
}
__host__ __device__ constexpr uint32_t idesc_bf16(int M, int N, int b_mn_major) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)b_mn_major << 16) | ((uint32_t)(N >> 3) << 17) |
           ((uint32_t)(M >> 4) << 24);              // D f32, A bf16, B bf16, A K-major
}
__device__ __forceinline__ void mma_ss(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, bool acc) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"((uint32_t)acc) : "memory");
}
__device__ __forceinline__ void mma_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, bool acc) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
                 ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"((uint32_t)acc) : "memory");
}
// The descriptor of a tile is constant except for the 14-bit start address (bytes >> 4) in its low word: the issuing
// warp keeps one low word per operand and advances it by a constant per k-step (uniform-register arithmetic).
__device__ __forceinline__ uint32_t desc_lo(uint32_t addr, uint32_t lbo_bytes) {
    return ((addr & 0x3FFFFu) >> 4) | ((lbo_bytes >> 4) << 16);
}
__device__ __forceinline__ uint64_t desc_join(uint32_t lo) {           // SBO = 1024 B, version 1, SWIZZLE_128B
    uint64_t d;
    asm("mov.b64 %0, {%1, %2};" : "=l"(d) : "r"(lo), "r"((1024u >> 4) | (1u << 14) | (2u << 29)));
    return d;
}
__device__ __forceinline__ void mma_commit(uint32_t bar) {       // arrives on `bar` when all MMAs issued so far are done
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ bool elect_one() {                   // true in exactly one lane of the (converged) warp
    uint32_t pred;
    asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
    return pred != 0;
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// 32 lanes x 32 columns of 32 bits: thread i of the warp gets lane (base + i), 32 consecutive columns
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
          "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
          "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr));
}
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x32.b32 [%32], "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31};"
        ::"r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]),
          "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]),
          "r"(r[16]), "r"(r[17]), "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]),
          "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31]),
          "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }


// 2^x for x <= 0 on the FMA/ALU pipes (the MUFU unit does 16 ex2 per clock and SM: with every weight on it the two
// tiles' softmax needs as many cycles as their MMAs, profiles/r1_ncu_prefill_tc5.md).  Round x to the nearest integer
// n with the 1.5 * 2^23 trick, 2^f on [-0.5, 0.5] by a cubic (relative error 7.5e-5, far below the bf16 rounding of
// P), n added to the exponent field.
__device__ __forceinline__ float poly_exp2(float x) {
    x = fmaxf(x, -126.f);
    const float t = x + 12582912.f;
    const float f = x - (t - 12582912.f);
    float p = fmaf(f, 0.05517157f, 0.24261111f);
    p = fmaf(p, f, 0.69326103f);
    p = fmaf(p, f, 0.99992806f);
    return __uint_as_float(__float_as_uint(p) + (__float_as_uint(t) << 23));
}

// Blackwell's packed fp32 pipe: two FMAs / adds per instruction (the softmax warps are issue- and latency-bound).
__device__ __forceinline__ void fma_x2(float& d0, float& d1, float a0, float a1, float b, float c) {
    asm("{\n\t.reg .b64 ra, rb, rc, rd;\n\t"
        "mov.b64 ra, {%2, %3};\n\tmov.b64 rb, {%4, %4};\n\tmov.b64 rc, {%5, %5};\n\t"
        "fma.rn.f32x2 rd, ra, rb, rc;\n\tmov.b64 {%0, %1}, rd;\n\t}"
        : "=f"(d0), "=f"(d1) : "f"(a0), "f"(a1), "f"(b), "f"(c));
}
__device__ __forceinline__ void add_x2(float& d0, float& d1, float a0, float a1) {   // (d0, d1) += (a0, a1)
    asm("{\n\t.reg .b64 ra, rd;\n\t"
        "mov.b64 ra, {%2, %3};\n\tmov.b64 rd, {%0, %1};\n\t"
        "add.rn.f32x2 rd, rd, ra;\n\tmov.b64 {%0, %1}, rd;\n\t}"
        : "+f"(d0), "+f"(d1) : "f"(a0), "f"(a1));
}

// =============================================================================================
// the kernel.  POLY of every 4 consecutive softmax weights take poly_exp2, the rest MUFU.EX2.
// =============================================================================================
// SPLIT (rows_per_cta = 128: launches that would leave SMs idle at 256 rows per CTA, e.g. a 512-token chunk over a long
// prefix): both slots work on the SAME 128 query rows and share the keys between them -- slot x takes the 128-key
// tiles x, x + 2, ... (always ring stage x), so the two slots still cover each other's softmax -> P.V -> Q.K^T chain
// instead of one tile leaving the tensor pipe idle during its own softmax.  Their (m, l, O) are merged by the warps
// of slot 0 at the end: a thread reads O of both slots from TMEM (same lane, other columns), m and l of slot 1 come
// through shared memory.
template <int POLY, bool SPLIT>
__global__ void __launch_bounds__(NTHREADS, 1)
prefill_tc5_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                   const __grid_constant__ CUtensorMap tmV, const Tc5Params p) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    __shared__ int s_tile[4];
    __shared__ uint32_t s_tmem;
    __shared__ float s_ml[SPLIT ? 2 * BM : 2];    // SPLIT: (m, l) of slot 1 per query row

    const int tid = threadIdx.x, lane = tid & 31;
    const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);      // warp-uniform for the compiler too (see the MMA issuer)
    const int hq = blockIdx.x, kvh = hq / p.G;
    const uint32_t sbase = b2::smem_u32(smem_raw);
    const uint32_t q_smem = sbase;
    const uint32_t ring = sbase + Q_BYTES;
    const uint32_t bars = ring + NSTAGE * STAGE_BYTES;
    auto bar = [&](int i) { return bars + 8u * i; };

    if (tid == 0) {
        if (sbase & 1023u) __trap();
        int want = (int)(gridDim.y - 1 - blockIdx.y), found = -1, q0 = 0;
        for (int c = 0; c < p.n_chunks; ++c) {
            const int len = p.cu_q[c + 1] - p.cu_q[c];
            const int nt = (len + p.rows_per_cta - 1) / p.rows_per_cta;
            if (want < nt) { found = c; q0 = want * p.rows_per_cta; break; }
            want -= nt;
        }
        s_tile[0] = found;
        s_tile[1] = q0;
        s_tile[2] = found >= 0 ? p.cu_q[found + 1] - p.cu_q[found] : 0;
        s_tile[3] = found >= 0 ? p.ctx_start[found] : 0;
        b2::mbar_init(bar(BAR_Q), 1);
        for (int s = 0; s < NSTAGE; ++s) {
            b2::mbar_init(bar(BAR_K + s), 1);
            b2::mbar_init(bar(BAR_V + s), 1);
            b2::mbar_init(bar(BAR_KE + s), 1);
            b2::mbar_init(bar(BAR_VE + s), 1);
        }
        for (int x = 0; x < 2; ++x) {
            b2::mbar_init(bar(BAR_S + x), 1);
            b2::mbar_init(bar(BAR_P + x), 4);      // one arrive per softmax warp of the tile: keys 0-63 of P are in TMEM
            b2::mbar_init(bar(BAR_P2 + x), 4);     // ... keys 64-127
            b2::mbar_init(bar(BAR_PV + x), 1);
        }
        b2::fence_mbar_init();
    }
    if (warp == 9) {                               // TMEM: all 512 columns (one CTA per SM)
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(b2::smem_u32(&s_tmem)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = s_tmem;
    // register file: 384 threads start with 168 each (where the 128 live S values of a row + packed P spill); the
    // third warpgroup keeps 64, the two softmax warpgroups take 216
    // (setmaxnreg opens each role's branch below)
    const int chunk = s_tile[0];
    const int q0 = s_tile[1], len = s_tile[2], start = s_tile[3];

    const bool active = chunk >= 0;               // CTA-uniform
    {
        const int rows = min(p.rows_per_cta, len - q0);         // valid query rows of the CTA
        const int rowsX[2] = {min(BM, rows), SPLIT ? min(BM, rows) : max(0, rows - BM)};
        // keys [0, hiX) are visible to the last row of tile X; tiles of 128 keys
        const int hiX[2] = {start + q0 + rowsX[0],
                            SPLIT ? start + q0 + rowsX[0] : rowsX[1] ? start + q0 + BM + rowsX[1] : 0};
        const int n_tiles = (max(hiX[0], hiX[1]) + BN - 1) / BN;
        // steps of slot x; its step i works on key tile SPLIT ? 2 i + x : i
        const int nX[2] = {SPLIT ? (n_tiles + 1) / 2 : (hiX[0] + BN - 1) / BN, SPLIT ? n_tiles / 2 : (hiX[1] + BN - 1) / BN};
        const int kv_hi = max(hiX[0], hiX[1]);
        const int row0 = active ? p.cu_q[chunk] + q0 : 0;       // first token row in q / out

        if (warp >= 8) {
        asm volatile("setmaxnreg.dec.sync.aligned.u32 64;");
        if (active && (warp == 8 || warp == 10)) {
            // ------------------------------------------ TMA producers: warp 8 streams K (and Q), warp 10 streams V ------------------------------------------
            // One box = one 64-element half of one (block, kv head) tile = bs x 128 B; at block_size 16 a 128-token tile
            // is 16 boxes.  Each box is issued by its own lane (a lone lane issuing all 32 boxes of a step one after the
            // other took longer than the step: the MMA warp waited ~800 cycles per step for V(j) / K(j+1), see
            // profiles/r2_trace_prefill_issuer.txt), and K and V wait for their own empty barriers in separate warps.
            const int kv = warp == 10;
            const CUtensorMap* tm = kv ? &tmV : &tmK;
            if (lane == 0) {
                if (!kv) b2::tma_prefetch_desc(&tmQ);
                b2::tma_prefetch_desc(tm);
            }
            uint64_t pol_keep, pol_once;
            asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol_keep));
            asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol_once));
            if (!kv && lane == 0) {
                b2::mbar_arrive_expect_tx(bar(BAR_Q), SPLIT ? TILE_BYTES : Q_BYTES);
                for (int x = 0; x < (SPLIT ? 1 : 2); ++x)       // SPLIT: one Q tile serves both slots
                    for (int h = 0; h < 2; ++h)                 // rows past T are zero-filled by the TMA unit
                        b2::tma_load_4d(q_smem + x * TILE_BYTES + h * SUB_BYTES, &tmQ, bar(BAR_Q), 0, h, hq,
                                        row0 + x * BM, pol_once);
            }
            const int32_t* bt = p.block_tables + (int64_t)chunk * p.max_nb;
            const int bps = BN / p.bs;
            const uint32_t box = p.bs * 128;                    // bytes of one half-tile box
            const int n_blk = (kv_hi + p.bs - 1) / p.bs;
            const int jb = lane >> 1, h = lane & 1;             // this lane's box of a tile: block jb, half h
            int win = 0;                                        // the 32 table entries that hold the current tiles
            for (int it = 0; it < n_tiles; ++it) {
                if (((it * bps) & 31) == 0) {                   // bps divides 32: a tile never straddles a window
                    const int first = it * bps;
                    win = first + lane < n_blk ? bt[first + lane] : 0;
                }
                const uint32_t s = it % NSTAGE, ph = (it / NSTAGE) & 1;
                b2::mbar_wait(bar((kv ? BAR_VE : BAR_KE) + s), ph ^ 1);
#ifdef B2_TRACE
                if (blockIdx.x == 0 && blockIdx.y == 1 && lane == 0 && it < 32) g_pf_trace[kv][it][12] = clock64();
#endif
                const int nb = min(bps, n_blk - it * bps);
                const uint32_t fb = bar((kv ? BAR_V : BAR_K) + s);
                if (lane == 0) b2::mbar_arrive_expect_tx(fb, nb * 2 * box);
                __syncwarp();
                const int blk = __shfl_sync(0xffffffffu, win, (it * bps + jb) & 31);
                if (jb < nb)
                    b2::tma_load_5d(ring + s * STAGE_BYTES + kv * TILE_BYTES + h * SUB_BYTES + jb * box, tm, fb, 0, h, 0, kvh,
                                    blk, pol_keep);
#ifdef B2_TRACE
                if (blockIdx.x == 0 && blockIdx.y == 1 && lane == 0 && it < 32) g_pf_trace[kv][it][13] = clock64();
#endif
            }
        }
#ifdef B2_TRACE
        else if (active && warp == 11) {                // trace build only: when do the K / V tiles actually land
            if (lane == 0 && blockIdx.x == 0 && blockIdx.y == 1)
                for (int it = 0; it < min(n_tiles, 32); ++it) {
                    const uint32_t s = it % NSTAGE, ph = (it / NSTAGE) & 1;
                    b2::mbar_wait(bar(BAR_K + s), ph);
                    g_pf_trace[0][it][15] = clock64();
                    b2::mbar_wait(bar(BAR_V + s), ph);
                    g_pf_trace[1][it][15] = clock64();
                }
        }
#endif
        else if (active && warp == 9) {
            // ------------------------------------------ MMA issuer ------------------------------------------
            // The WHOLE warp runs this loop with warp-uniform control flow and operands (`warp` comes from a shuffle,
            // every address below derives from kernel parameters and loop counters), and one elected lane issues each
            // group of instructions.  Round 2: with the loop inside `if (lane == 0)` the operands lived in per-thread
            // registers and ptxas wrapped EVERY tcgen05.mma in an ELECT / R2UR.BROADCAST / BRA.U.ANY loop of ~12
            // dependent instructions; sharing its scheduler with two softmax warps, the issuing thread then needed
            // 80-105 cycles per instruction where the pipe needs 64 (profiles/r2_trace_prefill_issuer.txt).
            {
                constexpr uint32_t ID_QK = idesc_bf16(128, 128, 0), ID_PV = idesc_bf16(128, 128, 1);
                // low descriptor words of the operands (see desc_lo): Q_x, the K half and the V half of stage 0
                const uint32_t q_lo0 = desc_lo(q_smem, 16);
                const uint32_t k_lo0 = desc_lo(ring, 16), v_lo0 = desc_lo(ring + TILE_BYTES, SUB_BYTES);
                auto issue_qk = [&](int x, int j) {        // S_x = Q_x . K(j)^T : 8 k-steps of 16 over d
                    const uint32_t kl = k_lo0 + (uint32_t)(j % NSTAGE) * (STAGE_BYTES >> 4);
                    const uint32_t ql = q_lo0 + (SPLIT ? 0u : (uint32_t)x * (TILE_BYTES >> 4));
                    if (elect_one()) {
#pragma unroll
                        for (int kk = 0; kk < 8; ++kk) {
                            const uint32_t off16 = (uint32_t)(((kk >> 2) * SUB_BYTES + (kk & 3) * 32) >> 4);
                            mma_ss(tmem + x * 128, desc_join(ql + off16), desc_join(kl + off16), ID_QK, kk > 0);
                        }
                        mma_commit(bar(BAR_S + x));
                    }
                    __syncwarp();
                };
                // rows of the last tile past the sequence hold whatever the cache (or an older stage) held: zero the
                // ones the last 16-token k-step still reads, so 0 * NaN cannot reach an accumulator
                auto zero_tail = [&](int jt) {
                    const uint32_t vb = ring + (jt % NSTAGE) * STAGE_BYTES + TILE_BYTES;
                    const int valid = kv_hi - jt * BN;
                    const int upto = min(BN, (valid + 15) & ~15);
                    if (lane < 16)                            // lane = (half, 16-byte column)
                        for (int r = valid; r < upto; ++r)
                            asm volatile("st.shared.v4.b32 [%0], {%1, %1, %1, %1};" ::"r"(vb + (lane >> 3) * SUB_BYTES + r * 128 + (lane & 7) * 16), "r"(0) : "memory");
                    b2::fence_proxy_async();
                    __syncwarp();
                };
                b2::mbar_wait(bar(BAR_Q), 0);
                if constexpr (SPLIT) {
                    // slot x owns ring stage x: its step i works on tile 2 i + x, barrier phases follow i
                    for (int x = 0; x < 2; ++x)
                        if (nX[x] > 0) {
                            b2::mbar_wait(bar(BAR_K + x), 0);
                            tc_fence_after();
                            issue_qk(x, x);
                            if (elect_one()) mma_commit(bar(BAR_KE + x));
                            __syncwarp();
                        }
                    for (int i = 0; i < nX[0]; ++i) {
                        for (int x = 0; x < 2; ++x) {
                            const int jt = 2 * i + x;
                            if (i < nX[x]) {
                                b2::mbar_wait(bar(BAR_V + x), i & 1);
                                if (jt == n_tiles - 1) zero_tail(jt);
                                b2::mbar_wait(bar(BAR_P + x), i & 1);      // keys 0-63 of P_x are in TMEM, O_x rescaled if needed
                                tc_fence_after();
                                const int ksteps = min(8, (kv_hi - jt * BN + 15) >> 4);
                                const uint32_t vl = v_lo0 + x * (STAGE_BYTES >> 4);
                                if (elect_one()) {
#pragma unroll
                                    for (int t = 0; t < 4; ++t)
                                        if (t < ksteps)
                                            mma_ts(tmem + 256 + x * 128, tmem + x * 128 + t * 8, desc_join(vl + t * (2048 >> 4)), ID_PV,
                                                   i > 0 || t > 0);
                                }
                                __syncwarp();
                                b2::mbar_wait(bar(BAR_P2 + x), i & 1);     // keys 64-127
                                tc_fence_after();
                                if (elect_one()) {
#pragma unroll
                                    for (int t = 4; t < 8; ++t)
                                        if (t < ksteps)
                                            mma_ts(tmem + 256 + x * 128, tmem + x * 128 + t * 8, desc_join(vl + t * (2048 >> 4)), ID_PV, true);
                                    mma_commit(bar(BAR_PV + x));
                                    mma_commit(bar(BAR_VE + x));
                                }
                                __syncwarp();
                            }
                            if (i + 1 < nX[x]) {
                                b2::mbar_wait(bar(BAR_K + x), (i + 1) & 1);
                                tc_fence_after();
                                issue_qk(x, jt + 2);
                                if (elect_one()) mma_commit(bar(BAR_KE + x));
                                __syncwarp();
                            }
                        }
                    }
                } else {
                b2::mbar_wait(bar(BAR_K + 0), 0);
                tc_fence_after();
                for (int x = 0; x < 2; ++x)
                    if (nX[x] > 0) issue_qk(x, 0);
                if (elect_one()) mma_commit(bar(BAR_KE + 0));                  // K(0) consumed
                __syncwarp();
                for (int j = 0; j < n_tiles; ++j) {
                    const uint32_t s = j % NSTAGE;
                    b2::mbar_wait(bar(BAR_V + s), (j / NSTAGE) & 1);
                    if (j == n_tiles - 1) zero_tail(j);
                    const bool more = j + 1 < n_tiles;
                    { const int x = 1; PF_ISSUE_STAMP(14); }
                    { const int x = 0; PF_ISSUE_STAMP(8); }
                    for (int x = 0; x < 2; ++x) {
                        if (j < nX[x]) {
                            b2::mbar_wait(bar(BAR_P + x), j & 1);          // keys 0-63 of P_x(j) are in TMEM, O_x rescaled if needed
                            tc_fence_after();
                            PF_ISSUE_STAMP(5);
                            const int ksteps = min(8, (hiX[x] - j * BN + 15) >> 4);
                            const uint32_t vl = v_lo0 + s * (STAGE_BYTES >> 4);
                            if (elect_one()) {                             // O_x += P_x . V(j) : k-steps of 16 tokens
#pragma unroll
                                for (int t = 0; t < 4; ++t)
                                    if (t < ksteps)
                                        mma_ts(tmem + 256 + x * 128, tmem + x * 128 + t * 8, desc_join(vl + t * (2048 >> 4)), ID_PV,
                                               j > 0 || t > 0);
                            }
                            __syncwarp();
                            b2::mbar_wait(bar(BAR_P2 + x), j & 1);         // keys 64-127
                            tc_fence_after();
                            if (elect_one()) {
#pragma unroll
                                for (int t = 4; t < 8; ++t)
                                    if (t < ksteps)
                                        mma_ts(tmem + 256 + x * 128, tmem + x * 128 + t * 8, desc_join(vl + t * (2048 >> 4)), ID_PV, true);
                                mma_commit(bar(BAR_PV + x));
                                if (x == 1) mma_commit(bar(BAR_VE + s));    // V(j) consumed: its stage refills under the Q.K^T below
                            }
                            __syncwarp();
                            PF_ISSUE_STAMP(6);
                        } else if (x == 1) {
                            if (elect_one()) mma_commit(bar(BAR_VE + s));
                            __syncwarp();
                        }
                        if (j + 1 < nX[x]) {
                            b2::mbar_wait(bar(BAR_K + (j + 1) % NSTAGE), ((j + 1) / NSTAGE) & 1);
#ifdef B2_TRACE
                            if (x == 0 && lane == 0 && blockIdx.x == 0 && blockIdx.y == 1 && j + 1 < 32) g_pf_trace[0][j + 1][14] = clock64();
#endif
                            issue_qk(x, j + 1);
                        }
                        PF_ISSUE_STAMP(7);
                    }
                    if (more) {
                        if (elect_one()) mma_commit(bar(BAR_KE + (j + 1) % NSTAGE));   // K(j+1) consumed (both tiles issued above)
                        __syncwarp();
                    }
                }
                }
            }
        }
        }
        else {
            asm volatile("setmaxnreg.inc.sync.aligned.u32 216;");
            if (active) {
            // ------------------------------------------ softmax warps: tile x = warp / 4 ------------------------------------------
            const int x = warp >> 2, wq = warp & 3;
            const int row = wq * 32 + lane;                       // query row in the tile = TMEM lane
            const uint32_t lane_off = (uint32_t)(wq * 32) << 16;
            const uint32_t s_addr = tmem + lane_off + x * 128, o_addr = tmem + lane_off + 256 + x * 128;
            const int n = nX[x];
            const float sl2 = p.scale_log2 * (p.k_scale ? *p.k_scale : 1.f);
            const int xrow = SPLIT ? 0 : x * BM;                  // first query row of the slot (SPLIT: the same rows in both)
            const int lim = start + q0 + xrow + min(row, max(rowsX[x] - 1, 0));     // last key this row may see
            const int lim_min = start + q0 + xrow;                                   // ... of the tile's first row
            float m = -INFINITY, l = 0.f;
#ifdef B2_TRACE
            const bool trace_on = blockIdx.x == 0 && blockIdx.y == 1 && wq == 0 && lane == 0;   // y = 0 is the spare (inactive) CTA of a single chunk
#endif
            for (int j = 0; j < n; ++j) {
                b2::mbar_wait(bar(BAR_S + x), j & 1);
                tc_fence_after();
                PF_STAMP(0);
                uint32_t sr[4][32];
#pragma unroll
                for (int c = 0; c < 4; ++c) tmem_ld32(s_addr + c * 32, sr[c]);
                tmem_wait_ld();
                PF_STAMP(1);
                const int j0 = (SPLIT ? 2 * j + x : j) * BN;      // first key of this step's tile
                if (j0 + BN - 1 > lim_min) {                      // tile touches the diagonal (or the end of the sequence)
#pragma unroll
                    for (int c = 0; c < 4; ++c)
#pragma unroll
                        for (int i = 0; i < 32; ++i)
                            if (j0 + c * 32 + i > lim) sr[c][i] = 0xff800000u;   // -inf
                }
                // one warp per tile and scheduler: nothing hides a 128-long dependent chain, so reduce in 8 short ones
                float mxp[8];
#pragma unroll
                for (int k = 0; k < 8; ++k) mxp[k] = -INFINITY;
#pragma unroll
                for (int c = 0; c < 4; ++c)
#pragma unroll
                    for (int i = 0; i < 32; ++i) mxp[i & 7] = fmaxf(mxp[i & 7], __uint_as_float(sr[c][i]));
                const float mx = fmaxf(fmaxf(fmaxf(mxp[0], mxp[1]), fmaxf(mxp[2], mxp[3])),
                                       fmaxf(fmaxf(mxp[4], mxp[5]), fmaxf(mxp[6], mxp[7])));
                // finite: key 0 is visible to every row, tiles in order.  SPLIT: slot 1 starts at key 128, which a row of a
                // chunk that starts inside the first tile may not see -- a finite floor keeps exp2(-inf - m) = 0, not NaN
                const float m_new = SPLIT ? fmaxf(fmaxf(m, mx), -1e30f) : fmaxf(m, mx);
                if (j == 0) {
                    m = m_new;
                } else {
                    const bool need = (m_new - m) * sl2 > RESCALE_THRESHOLD;
                    if (__any_sync(0xffffffffu, need)) {          // rare after the first tiles: bring O_x to the new maximum
                        b2::mbar_wait(bar(BAR_PV + x), (j - 1) & 1);
                        tc_fence_after();
                        const float alpha = need ? b2::fast_exp2((m - m_new) * sl2) : 1.f;
                        if (need) { m = m_new; l *= alpha; }
#pragma unroll
                        for (int c = 0; c < 4; ++c) {
                            uint32_t orr[32];
                            tmem_ld32(o_addr + c * 32, orr);
                            tmem_wait_ld();
#pragma unroll
                            for (int i = 0; i < 32; ++i) orr[i] = __float_as_uint(__uint_as_float(orr[i]) * alpha);
                            tmem_st32(o_addr + c * 32, orr);
                        }
                    }
                }
                PF_STAMP(2);
                const float ms = m * sl2;
                // l sums the fp32 weights (the bf16 rounding of P is zero-mean; FlashAttention convention), 4 chains
                float sum[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
                for (int c = 0; c < 2; ++c) {                     // 64 keys -> 32 packed bf16 pairs per store
                    uint32_t pk[32];
#pragma unroll
                    for (int i = 0; i < 32; ++i) {
                        const int cc = 2 * c + (i >> 4), ii = (2 * i) & 31;
                        float x0, x1;
                        fma_x2(x0, x1, __uint_as_float(sr[cc][ii]), __uint_as_float(sr[cc][ii + 1]), sl2, -ms);
                        // (i & 1, element) -> slot 0..3 of the group of four; slots < POLY go to the polynomial
                        const float p0 = (POLY >= 2 && (i & 1)) || POLY >= 3 ? poly_exp2(x0) : b2::fast_exp2(x0);
                        const float p1 = (POLY >= 1 && (i & 1)) || POLY >= 4 ? poly_exp2(x1) : b2::fast_exp2(x1);
                        pk[i] = b2::pack_bf16x2(p0, p1);
                        add_x2(sum[2 * (i & 1)], sum[2 * (i & 1) + 1], p0, p1);
                    }
                    tmem_st32(s_addr + c * 32, pk);               // P overwrites the first 64 columns of S_x
                    if (c == 0) {                                 // hand the first 64 keys over: their P.V runs under the
                        tmem_wait_st();                           // exponentials of the other 64
                        tc_fence_before();
                        __syncwarp();
                        if (lane == 0) b2::mbar_arrive(bar(BAR_P + x));
                    }
                }
                l += (sum[0] + sum[1]) + (sum[2] + sum[3]);
                PF_STAMP(3);
                tmem_wait_st();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) b2::mbar_arrive(bar(BAR_P2 + x));
                PF_STAMP(4);
#ifdef B2_TRACE
                if (blockIdx.x == 0 && blockIdx.y == 1 && wq > 0 && lane == 0 && j < 32) g_pf_trace[x][j][8 + wq] = clock64();
#endif
            }
            if constexpr (SPLIT) {
                // merge the two slots (same rows, interleaved key tiles): slot 1 publishes (m, l), the thread of slot 0
                // that owns the row reads O of both slots from TMEM (same lane, columns 256.. and 384..)
                if (x == 1) {
                    s_ml[row] = n > 0 ? m : -1e30f;
                    s_ml[BM + row] = n > 0 ? l : 0.f;
                    __threadfence_block();
                    asm volatile("bar.arrive 1, 256;" ::: "memory");
                } else {
                    b2::mbar_wait(bar(BAR_PV + 0), (n - 1) & 1);                  // n >= 1: slot 0 owns tile 0
                    asm volatile("bar.sync 1, 256;" ::: "memory");
                    const int n1 = nX[1];
                    // commits complete in issue order, so PV_1 lags PV_0 by less than one phase here (parity wait is safe)
                    if (n1 > 0) b2::mbar_wait(bar(BAR_PV + 1), (n1 - 1) & 1);
                    tc_fence_after();
                    const float m1 = s_ml[row], l1 = s_ml[BM + row];
                    const float mm = fmaxf(m, m1);
                    float a0 = b2::fast_exp2((m - mm) * sl2), a1 = n1 > 0 ? b2::fast_exp2((m1 - mm) * sl2) : 0.f;
                    const float inv = (p.v_scale ? *p.v_scale : 1.f) / (a0 * l + a1 * l1);
                    a0 *= inv;
                    a1 *= inv;
                    const bool valid = row < rowsX[0];
                    __nv_bfloat16* orow = p.out + (int64_t)(row0 + row) * p.out_stride + (int64_t)hq * D;
#pragma unroll
                    for (int c = 0; c < 4; ++c) {
                        uint32_t oa[32], ob[32];
                        tmem_ld32(o_addr + c * 32, oa);
                        if (n1 > 0) tmem_ld32(o_addr + 128 + c * 32, ob);
                        tmem_wait_ld();
                        if (valid) {
                            float r[32];
#pragma unroll
                            for (int i = 0; i < 32; ++i)
                                r[i] = n1 > 0 ? fmaf(__uint_as_float(ob[i]), a1, __uint_as_float(oa[i]) * a0) : __uint_as_float(oa[i]) * a0;
#pragma unroll
                            for (int i = 0; i < 32; i += 8) {
                                int4 v;
                                v.x = (int)b2::pack_bf16x2(r[i], r[i + 1]);
                                v.y = (int)b2::pack_bf16x2(r[i + 2], r[i + 3]);
                                v.z = (int)b2::pack_bf16x2(r[i + 4], r[i + 5]);
                                v.w = (int)b2::pack_bf16x2(r[i + 6], r[i + 7]);
                                *reinterpret_cast<int4*>(orow + c * 32 + i) = v;
                            }
                        }
                    }
                }
            } else if (n > 0) {
                b2::mbar_wait(bar(BAR_PV + x), (n - 1) & 1);
                tc_fence_after();
                const float inv = (p.v_scale ? *p.v_scale : 1.f) / l;
                const bool valid = row < rowsX[x];
                __nv_bfloat16* orow = p.out + (int64_t)(row0 + x * BM + row) * p.out_stride + (int64_t)hq * D;
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    uint32_t orr[32];
                    tmem_ld32(o_addr + c * 32, orr);
                    tmem_wait_ld();
                    if (valid) {
#pragma unroll
                        for (int i = 0; i < 32; i += 8) {
                            int4 v;
                            v.x = (int)b2::pack_bf16x2(__uint_as_float(orr[i]) * inv, __uint_as_float(orr[i + 1]) * inv);
                            v.y = (int)b2::pack_bf16x2(__uint_as_float(orr[i + 2]) * inv, __uint_as_float(orr[i + 3]) * inv);
                            v.z = (int)b2::pack_bf16x2(__uint_as_float(orr[i + 4]) * inv, __uint_as_float(orr[i + 5]) * inv);
                            v.w = (int)b2::pack_bf16x2(__uint_as_float(orr[i + 6]) * inv, __uint_as_float(orr[i + 7]) * inv);
                            *reinterpret_cast<int4*>(orow + c * 32 + i) = v;
                        }
                    }
                }
            }
            }
        }
    }

    // ------------------------------------------ teardown ------------------------------------------
    tc_fence_before();
    __syncthreads();
    if (warp == 9) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem) : "memory");
    }
}

}  // namespace

int b2_internal_prefill_tc5(void* out, const void* q, const void* k_cache, const void* v_cache,
                            const int32_t* block_tables, const int32_t* cu_q, const int32_t* ctx_start, int n_chunks,
                            int T, int Hq, int Hkv, int bs, int max_nb, int64_t num_blocks, int64_t q_stride,
                            int64_t out_stride, int64_t cache_block_stride, int64_t cache_head_stride, float scale,
                            const float* k_scale, const float* v_scale, void* stream) {
    // 256 query rows per CTA share each K/V stage between two tiles; when that leaves SMs idle, 128 rows per CTA
    static const int n_sm = []() { int dev = 0, n = 148; if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev); return n; }();
    static const int force_rows = []() { const char* e = getenv("B2_PREFILL_ROWS"); return e && *e ? atoi(e) : 0; }();   // A/B knob
    const int rows_per_cta = force_rows == BM || force_rows == 2 * BM ? force_rows
                             : ((int64_t)T / (2 * BM) + n_chunks) * Hq < n_sm ? BM : 2 * BM;    // less than one wave
    const int64_t n_qtiles = (int64_t)T / rows_per_cta + n_chunks;        // upper bound of sum ceil(len_c / rows_per_cta)
    B2_REQUIRE(n_qtiles <= 65535, B2_EUNSUPPORTED, "prefill_attention: %lld q tiles exceed the grid", (long long)n_qtiles);
    B2_REQUIRE(bs <= 64 && BN % bs == 0, B2_EUNSUPPORTED, "prefill_attention: block_size %d", bs);

    CUtensorMap tmQ, tmK, tmV;
    {   // q viewed as [T][Hq][2 halves][64]: one box = 128 rows x one 64-element half of one head
        const uint64_t dims[4] = {64, 2, (uint64_t)Hq, (uint64_t)T};
        const uint64_t strides[3] = {128, 256, (uint64_t)q_stride * 2};
        const uint32_t box[4] = {64, 1, 1, (uint32_t)BM};
        int rc = b2_internal_encode_bf16_sw128(&tmQ, q, 4, dims, strides, box);
        if (rc != B2_OK) return rc;
    }
    for (int kv = 0; kv < 2; ++kv) {   // cache viewed as [num_blocks][Hkv][bs][2 halves][64]: one box = bs rows x one half
        const uint64_t dims[5] = {64, 2, (uint64_t)bs, (uint64_t)Hkv, (uint64_t)num_blocks};
        const uint64_t strides[4] = {128, 256, (uint64_t)cache_head_stride * 2, (uint64_t)cache_block_stride * 2};
        const uint32_t box[5] = {64, 1, (uint32_t)bs, 1, 1};
        int rc = b2_internal_encode_bf16_sw128(kv ? &tmV : &tmK, kv ? v_cache : k_cache, 5, dims, strides, box);
        if (rc != B2_OK) return rc;
    }
    Tc5Params p;
    p.out = static_cast<__nv_bfloat16*>(out);
    p.block_tables = block_tables;
    p.cu_q = cu_q;
    p.ctx_start = ctx_start;
    p.n_chunks = n_chunks;
    p.Hq = Hq;
    p.G = Hq / Hkv;
    p.bs = bs;
    p.max_nb = max_nb;
    p.rows_per_cta = rows_per_cta;
    p.out_stride = out_stride;
    p.scale_log2 = scale * LOG2E;
    p.k_scale = k_scale;
    p.v_scale = v_scale;
    // share of the softmax weights computed by polynomial (B2_PREFILL_POLY = 0..3 of every 4; A/B knob)
    static const int poly = []() { const char* e = getenv("B2_PREFILL_POLY"); return e && *e ? atoi(e) : 0; }();
    // 128 rows per CTA run the SPLIT kernel (both slots on the same rows, key tiles interleaved); B2_PREFILL_SPLIT=0
    // keeps the single-tile kernel as the A/B
    static const bool split_on = []() { const char* e = getenv("B2_PREFILL_SPLIT"); return !(e && *e == '0'); }();
    using KernFn = void (*)(const CUtensorMap, const CUtensorMap, const CUtensorMap, const Tc5Params);
    KernFn kern = poly <= 0 ? prefill_tc5_kernel<0, false> : poly == 1 ? prefill_tc5_kernel<1, false>
                : poly == 2 ? prefill_tc5_kernel<2, false> : prefill_tc5_kernel<3, false>;
    if (rows_per_cta == BM && split_on && poly <= 0) kern = prefill_tc5_kernel<0, true>;
    B2_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
    kern<<<dim3(Hq, (unsigned)n_qtiles), NTHREADS, SMEM_BYTES, static_cast<cudaStream_t>(stream)>>>(tmQ, tmK, tmV, p);
    B2_CUDA_OK(cudaGetLastError());
    return B2_OK;
}

#ifdef B2_TRACE
extern "C" B2_API int b2_debug_read_prefill_trace(unsigned long long* dst) {
    return cudaMemcpyFromSymbol(dst, g_pf_trace, sizeof(g_pf_trace)) == cudaSuccess ? 0 : B2_ECUDA;
}
#endif
