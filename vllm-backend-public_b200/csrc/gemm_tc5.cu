// gemm_tc5.cu -- the projections of a DECODE step on the 5th-gen tensor cores: out[M, N] = x[M, K] . W[N, K]^T, bf16,
// M <= 256 rows (one per running sequence), W a model weight that is read exactly once per step.
//
// Status: an EXPERIMENT, parity-green (tests/test_gpu_gemm.py) but not on the engine's path.  SURVEY.md section 8(f)
// rank 3 lets the GEMMs be library calls; at M = 64 the op is a weight STREAM (64 flop per weight byte, HBM-bound), and
// in-graph cuBLAS reaches 53-88 % of the HBM bandwidth on the Llama-3-8B decode shapes and 13-49 % on the per-rank
// shapes of --tensor-parallel-size 8.  This kernel was written to beat that and does not yet: 0.47-0.81x of the
// library's speed (profiles/r2_decode_gemm.jsonl) -- one non-persistent CTA per (feature tile, K slice) pays TMEM
// allocation, pipeline ramp and two cluster barriers on <= 0.5 MB of weights.  DESIGN.md section 8 says what would.
//
// Shape of the kernel (swap-AB, split-K over a thread-block cluster):
//   * the WEIGHT tile is the M side of the MMA: one CTA owns 128 output features (rows of W), the activations are the N
//     side (N_mma = M rounded up to 16), so no MMA row is padding whatever the batch size:
//         D^T[128 features, M] += W[128, 64-chunk] . x[M, 64-chunk]^T      tcgen05.mma.cta_group::1.kind::f16, fp32 in TMEM
//   * K is cut over the `ksplit` CTAs of a cluster so that ~148 x k CTAs stream concurrently whatever N is
//     (N = 4096 is only 32 feature tiles); each CTA streams its [128 x K/ksplit] slab of W once through a ring of TMA
//     stages (16 KiB of W + M x 128 B of x per stage, SWIZZLE_128B, mbarrier full/empty, L2 evict-first for W);
//   * the partial sums meet in DISTRIBUTED SHARED MEMORY: every CTA parks its fp32 tile in its own shared memory
//     ([m][feature], so the TMEM-lane-per-feature reads land conflict-free), one cluster barrier, then CTA r of the
//     cluster sums feature slice r over all peers (ld.shared::cluster), rounds once to bf16 and stores 8-byte vectors.
//     No workspace, no second kernel, no atomics.
// Warps: 0-3 epilogue (one TMEM lane = one feature each), 4 TMA producer, 5 MMA issuer + TMEM allocation.
#include "b2_common.cuh"

#include <stdlib.h>

#include <mutex>
#include <unordered_map>

namespace {

constexpr int BF = 128;                 // features (rows of W) per CTA = MMA M
constexpr int KC = 64;                  // K elements per stage = one 128-byte swizzle row
constexpr int A_BYTES = BF * KC * 2;    // 16 KiB of W per stage
constexpr int NTHREADS = 192;
constexpr int MAX_M = 256;

struct GemmParams {
    __nv_bfloat16* out;
    int64_t ldo;
    int M, N, K;
    int n_mma;          // M rounded up to 16
    int ksplit;         // CTAs per cluster = K slices
    int nstage;
    int stage_bytes;    // A_BYTES + n_mma * 128
    int tmem_cols;      // power of two >= max(32, n_mma)
};

__device__ __forceinline__ uint64_t smem_desc(uint32_t addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    return (uint64_t)((addr & 0x3FFFFu) >> 4) | ((uint64_t)(lbo_bytes >> 4) << 16) | ((uint64_t)(sbo_bytes >> 4) << 32) |
           (1ull << 46) | (2ull << 61);             // version 1 (Blackwell), SWIZZLE_128B
}
__device__ __forceinline__ uint32_t idesc_bf16(int M, int N) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}                                                    // D f32, A bf16, B bf16, both K-major
__device__ __forceinline__ void mma_ss(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, bool acc) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"((uint32_t)acc) : "memory");
}
__device__ __forceinline__ void mma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {     // lane = thread, 16 consecutive columns
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr));
}
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* tmap, uint32_t bar, int c0, int c1,
                                            uint64_t policy) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes.L2::cache_hint"
        " [%0], [%1, {%3, %4}], [%2], %5;"
        ::"r"(dst), "l"(tmap), "r"(bar), "r"(c0), "r"(c1), "l"(policy) : "memory");
}
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t mapa(uint32_t addr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
    return r;
}
__device__ __forceinline__ float4 ld_dsmem_f4(uint32_t addr) {
    float4 v;
    asm volatile("ld.shared::cluster.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
    return v;
}

__global__ void __launch_bounds__(NTHREADS, 1)
skinny_gemm_kernel(const __grid_constant__ CUtensorMap tmW, const __grid_constant__ CUtensorMap tmX, const GemmParams p) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t s_bar[2 * 8 + 1];      // full[nstage] | empty[nstage] | done
    __shared__ uint32_t s_tmem;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t ks = (uint32_t)p.ksplit;
    const uint32_t split = ks > 1 ? cluster_ctarank() : 0u;
    const int tile = blockIdx.x / ks;
    const int n0 = tile * BF;
    const int nk = p.K / KC;                                 // 64-element chunks of K
    const int per = (nk + (int)ks - 1) / (int)ks;
    const int k_begin = (int)split * per, k_end = min(nk, k_begin + per);
    const int n_it = max(k_end - k_begin, 0);

    const uint32_t ring = b2::smem_u32(smem_raw);
    const uint32_t full0 = b2::smem_u32(&s_bar[0]), empty0 = b2::smem_u32(&s_bar[8]), done = b2::smem_u32(&s_bar[16]);
    if (tid == 0) {
        if (ring & 1023u) __trap();
        for (int s = 0; s < p.nstage; ++s) {
            b2::mbar_init(full0 + 8 * s, 1);
            b2::mbar_init(empty0 + 8 * s, 1);
        }
        b2::mbar_init(done, 1);
        b2::fence_mbar_init();
    }
    if (warp == 5) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(b2::smem_u32(&s_tmem)), "r"(p.tmem_cols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = s_tmem;

    if (warp == 4) {
        // ------------------------------ TMA producer ------------------------------
        if (lane == 0) {
            b2::tma_prefetch_desc(&tmW);
            b2::tma_prefetch_desc(&tmX);
            uint64_t pol_once, pol_keep;
            asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol_once));
            asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol_keep));
            for (int it = 0; it < n_it; ++it) {
                const int s = it % p.nstage;
                if (it >= p.nstage) b2::mbar_wait(empty0 + 8 * s, ((it / p.nstage) - 1) & 1);
                const uint32_t dst = ring + (uint32_t)s * p.stage_bytes;
                b2::mbar_arrive_expect_tx(full0 + 8 * s, (uint32_t)p.stage_bytes);
                const int kc = (k_begin + it) * KC;
                tma_load_2d(dst, &tmW, full0 + 8 * s, kc, n0, pol_once);                 // W[n0 .. n0+128, kc .. kc+64]
                tma_load_2d(dst + A_BYTES, &tmX, full0 + 8 * s, kc, 0, pol_keep);        // x[0 .. n_mma, kc .. kc+64]
            }
        }
        __syncwarp();
    } else if (warp == 5) {
        // ------------------------------ MMA issuer ------------------------------
        if (lane == 0) {
            const uint32_t idesc = idesc_bf16(BF, p.n_mma);
            for (int it = 0; it < n_it; ++it) {
                const int s = it % p.nstage;
                b2::mbar_wait(full0 + 8 * s, (it / p.nstage) & 1);
                tc_fence_after();
                const uint32_t a = ring + (uint32_t)s * p.stage_bytes, b = a + A_BYTES;
#pragma unroll
                for (int kk = 0; kk < KC / 16; ++kk)
                    mma_ss(tmem, smem_desc(a + kk * 32, 16, 1024), smem_desc(b + kk * 32, 16, 1024), idesc, it > 0 || kk > 0);
                mma_commit(empty0 + 8 * s);            // the stage is free once these MMAs have read it
            }
            mma_commit(done);                          // every MMA of this CTA has landed in TMEM
        }
        __syncwarp();
    }

    // ------------------------------ epilogue ------------------------------
    // park the fp32 tile as part[m][feature] in shared memory (the ring is free once `done` fires)
    float* part = reinterpret_cast<float*>(smem_raw);
    if (warp < 4) {
        b2::mbar_wait(done, 0);
        tc_fence_after();
        const int f = warp * 32 + lane;                              // feature = TMEM lane
        const uint32_t taddr = tmem + ((uint32_t)(warp * 32) << 16);
        for (int c = 0; c < p.n_mma; c += 16) {
            uint32_t r[16];
            if (n_it > 0) {
                tmem_ld16(taddr + c, r);
                tmem_wait_ld();
            } else {
#pragma unroll
                for (int i = 0; i < 16; ++i) r[i] = 0u;              // this split got no K chunk: contributes zero
            }
#pragma unroll
            for (int i = 0; i < 16; ++i) part[(c + i) * BF + f] = __uint_as_float(r[i]);
        }
        tc_fence_before();
    }
    __syncthreads();
    if (ks > 1) cluster_sync();                                       // every CTA's partial tile is parked
    {
        // CTA `split` of the cluster reduces features [split * F, (split + 1) * F) over all peers
        const int F = BF / (int)ks;
        const int vec_per_row = F / 4;
        const int total = p.M * vec_per_row;
        const uint32_t part_addr = b2::smem_u32(part);
        for (int idx = tid; idx < total; idx += NTHREADS) {
            const int m = idx / vec_per_row, v = idx - m * vec_per_row;
            const int f = (int)split * F + v * 4;
            const uint32_t off = (uint32_t)((m * BF + f) * 4);
            float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
            if (ks > 1) {
                for (uint32_t r = 0; r < ks; ++r) {                  // fixed order: the same bits on every launch
                    const float4 t = ld_dsmem_f4(mapa(part_addr + off, r));
                    acc.x += t.x; acc.y += t.y; acc.z += t.z; acc.w += t.w;
                }
            } else {
                acc = *reinterpret_cast<const float4*>(reinterpret_cast<const uint8_t*>(part) + off);
            }
            const int n = n0 + f;
            if (n + 3 < p.N) {
                uint2 pk;
                pk.x = b2::pack_bf16x2(acc.x, acc.y);
                pk.y = b2::pack_bf16x2(acc.z, acc.w);
                *reinterpret_cast<uint2*>(p.out + (int64_t)m * p.ldo + n) = pk;
            } else {
                const float a4[4] = {acc.x, acc.y, acc.z, acc.w};
                for (int i = 0; i < 4; ++i)
                    if (n + i < p.N) p.out[(int64_t)m * p.ldo + n + i] = __float2bfloat16(a4[i]);
            }
        }
    }
    if (ks > 1) cluster_sync();                                       // nobody leaves while a peer still reads its tile
    __syncthreads();
    if (warp == 5) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(p.tmem_cols) : "memory");
    }
}

// W's tensor map depends only on (pointer, N, K): encode it once per weight (a model has a few hundred)
struct WKey {
    const void* ptr;
    int N, K;
    bool operator==(const WKey& o) const { return ptr == o.ptr && N == o.N && K == o.K; }
};
struct WKeyHash {
    size_t operator()(const WKey& k) const {
        return std::hash<const void*>()(k.ptr) ^ (std::hash<int>()(k.N) * 31u) ^ (std::hash<int>()(k.K) * 131u);
    }
};
std::mutex g_map_mu;
std::unordered_map<WKey, CUtensorMap, WKeyHash> g_wmaps;

}  // namespace

extern "C" int b2_skinny_gemm(void* out, const void* x, const void* w, int M, int N, int K, int64_t ldx, int64_t ldo,
                              void* stream) {
    B2_REQUIRE(M >= 0 && N > 0 && K > 0, B2_EINVAL, "skinny_gemm: bad sizes M=%d N=%d K=%d", M, N, K);
    if (M == 0) return B2_OK;
    B2_REQUIRE(out && x && w, B2_EINVAL, "skinny_gemm: null pointer");
    B2_REQUIRE(M <= MAX_M, B2_EUNSUPPORTED, "skinny_gemm: M=%d > %d rows (use a library GEMM for prompt chunks)", M, MAX_M);
    B2_REQUIRE(K % KC == 0, B2_EUNSUPPORTED, "skinny_gemm: K=%d must be a multiple of %d", K, KC);
    B2_REQUIRE(N % 4 == 0 && ldo % 4 == 0 && ldx % 8 == 0 && ldx >= K && ldo >= N, B2_EINVAL,
               "skinny_gemm: N and ldo must be multiples of 4, ldx of 8, and cover the rows");
    B2_REQUIRE(b2_aligned16(x) && b2_aligned16(w) && (reinterpret_cast<uintptr_t>(out) & 7u) == 0, B2_EINVAL,
               "skinny_gemm: x / w must be 16-byte aligned, out 8-byte aligned");
    cudaStream_t st = static_cast<cudaStream_t>(stream);

    GemmParams p;
    p.out = static_cast<__nv_bfloat16*>(out);
    p.ldo = ldo;
    p.M = M; p.N = N; p.K = K;
    p.n_mma = (M + 15) / 16 * 16;
    p.tmem_cols = 32;
    while (p.tmem_cols < p.n_mma) p.tmem_cols *= 2;
    p.stage_bytes = A_BYTES + p.n_mma * 128;
    const int n_tiles = (N + BF - 1) / BF;
    const int nk = K / KC;
    // K slices per feature tile: fill the SMs (one CTA per SM: the ring takes most of the shared memory) and keep the
    // waves even; a slice costs one cluster barrier + a DSMEM pass (~2 chunks' worth), and F = 128 / ksplit must stay a
    // multiple of 4
    static const int sm_count = [] {
        int dev = 0, n = 148;
        if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
        return n;
    }();
    static const int fixed_cost = [] { const char* e = getenv("B2_GEMM_FIXED"); return e ? atoi(e) : 16; }();
    static const int force_ks = [] { const char* e = getenv("B2_GEMM_KSPLIT"); return e ? atoi(e) : 0; }();
    int best = 1;
    long best_cost = -1;
    for (int ksp = 1; ksp <= 8; ksp *= 2) {
        if (ksp > nk) break;
        const long waves = ((long)n_tiles * ksp + sm_count - 1) / sm_count;
        // a CTA's fixed cost (launch, TMEM allocation, pipeline ramp, epilogue; + the cluster barriers and the DSMEM
        // pass when K is split) is worth ~16-18 chunks of streaming: B2_GEMM_FIXED overrides (experiments)
        const long cost = waves * ((nk + ksp - 1) / ksp + (ksp > 1 ? fixed_cost + 2 : fixed_cost));
        if (best_cost < 0 || cost < best_cost) { best_cost = cost; best = ksp; }
    }
    if (force_ks == 1 || force_ks == 2 || force_ks == 4 || force_ks == 8) best = force_ks <= nk ? force_ks : best;
    p.ksplit = best;
    // the fp32 tile parked for the reduction shares the ring's memory: n_mma x 128 x 4 bytes
    const int part_bytes = p.n_mma * BF * 4;
    int nstage = (200 * 1024) / p.stage_bytes;
    if (nstage > 8) nstage = 8;
    const int need_it = (nk + p.ksplit - 1) / p.ksplit;
    if (nstage > need_it) nstage = need_it < 2 ? 2 : need_it;
    while (nstage * p.stage_bytes < part_bytes) ++nstage;
    B2_REQUIRE(nstage <= 8 && nstage * p.stage_bytes <= 220 * 1024, B2_EUNSUPPORTED, "skinny_gemm: M=%d does not fit shared memory", M);
    p.nstage = nstage;
    const size_t smem = (size_t)nstage * p.stage_bytes;

    CUtensorMap tmW, tmX;
    {
        std::lock_guard<std::mutex> lk(g_map_mu);
        const WKey key{w, N, K};
        auto it = g_wmaps.find(key);
        if (it == g_wmaps.end()) {
            const uint64_t dims[2] = {(uint64_t)K, (uint64_t)N};
            const uint64_t strides[1] = {(uint64_t)K * 2};
            const uint32_t box[2] = {KC, BF};
            CUtensorMap tm;
            int rc = b2_internal_encode_bf16_sw128(&tm, w, 2, dims, strides, box);
            if (rc != B2_OK) return rc;
            if (g_wmaps.size() > 4096) g_wmaps.clear();
            it = g_wmaps.emplace(key, tm).first;
        }
        tmW = it->second;
    }
    {
        const uint64_t dims[2] = {(uint64_t)K, (uint64_t)M};           // rows past M are zero-filled by the TMA unit
        const uint64_t strides[1] = {(uint64_t)ldx * 2};
        const uint32_t box[2] = {KC, (uint32_t)p.n_mma};
        int rc = b2_internal_encode_bf16_sw128(&tmX, x, 2, dims, strides, box);
        if (rc != B2_OK) return rc;
    }
    static size_t smem_set = 0;
    if (smem > smem_set) {
        B2_CUDA_OK(cudaFuncSetAttribute(skinny_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        smem_set = smem;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(n_tiles * p.ksplit);
    cfg.blockDim = dim3(NTHREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = p.ksplit;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    B2_CUDA_OK(cudaLaunchKernelEx(&cfg, skinny_gemm_kernel, tmW, tmX, p));
    return B2_OK;
}
