// Issue-rate probe for tcgen05.mma.cta_group::1.kind::f16 (bf16 in, fp32 in TMEM), one CTA per SM, one issuing thread:
// how many SM cycles does one 128 x N x 16 instruction cost when
//   (ss)  A and B come from shared memory (SWIZZLE_128B, K-major) -- the Q.K^T of csrc/prefill_tc5.cu
//   (ts)  A comes from TMEM, B from shared memory                  -- its P.V (V is MN-major there; K-major here)
// for N = 64 / 128 / 256.  The operands' VALUES do not matter (zeros); the shared-memory traffic does: an SS MMA at
// M = 128, N = 128 reads 4 KiB of A + 4 KiB of B per instruction.
// Build + run on the GPU box:  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/umma_rate scripts/umma_rate.cu && /tmp/umma_rate
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

constexpr int ITERS = 2048;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t smem_desc(uint32_t addr, uint32_t lbo, uint32_t sbo) {
    return (uint64_t)((addr & 0x3FFFFu) >> 4) | ((uint64_t)(lbo >> 4) << 16) | ((uint64_t)(sbo >> 4) << 32) | (1ull << 46) |
           (2ull << 61);
}
__device__ __forceinline__ uint32_t idesc_bf16(int M, int N) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

// MODE 0: SS   1: TS, B K-major   2: TS, B MN-major (the P.V of prefill_tc5: V tile [2 halves][128 tok][64 el])
//      3: the kernel's pair -- 8 x (TS, MN-major B) then 8 x SS, all 128 x 128 x 16
// LOAD > 0: warps 1..3 hammer TMEM with tcgen05.ld / tcgen05.st (32 columns each) while warp 0 issues, as softmax warps do
template <int N, int MODE, int LOAD>
__global__ void __launch_bounds__(128, 1) probe(long long* cycles) {
    constexpr bool TS = MODE == 1 || MODE == 2;
    extern __shared__ __align__(1024) uint8_t smem[];        // A: 128 x 64 bf16 (16 KiB) | B: up to 32 KiB
    __shared__ __align__(8) uint64_t bar;
    __shared__ uint32_t s_tmem;
    const int tid = threadIdx.x;
    for (int i = tid; i < (16384 + 32768) / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0u;
    __shared__ volatile int s_stop;
    if (tid == 0) s_stop = 0;
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (tid < 32) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&s_tmem)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = s_tmem;
    if (tid == 0) {
        const uint32_t a = smem_u32(smem), b = a + 16384;
        const uint32_t idesc = idesc_bf16(128, N) | ((MODE == 2 ? 1u : 0u) << 16);
        if (MODE >= 3) {
            // groups of GROUP instructions alternate between two kinds X / Y (descriptors precomputed, groups unrolled,
            // so the issuing thread executes almost nothing but the MMAs):
            //   3: X = TS, B MN-major (P.V)   Y = SS (Q.K^T)            -- the prefill kernel's pair, GROUP 8
            //   4: X = TS, B MN-major         Y = TS, B K-major         -- only the B layout (idesc) changes
            //   5: X = SS into D columns 0    Y = SS into D columns 128 -- only the accumulator changes
            //   6: X = TS, B K-major          Y = SS                    -- only the A source changes
            //   7: as 3 with GROUP 16         8: as 3 with GROUP 32
            constexpr int GROUP = MODE == 7 ? 16 : MODE == 8 ? 32 : 8;
            constexpr int KX = MODE == 4 ? 0 : MODE == 5 ? 2 : MODE == 6 ? 1 : 0;
            constexpr int KY = MODE == 4 ? 1 : MODE == 5 ? 3 : 2;
            uint64_t vd[8], kd[4], ad[4];
            for (int i = 0; i < 8; ++i) vd[i] = smem_desc(b + i * 2048, 16384, 1024);
            for (int i = 0; i < 4; ++i) { kd[i] = smem_desc(b + i * 32, 16, 1024); ad[i] = smem_desc(a + i * 32, 16, 1024); }
            auto issue = [&](int kind, int i) {
                if (kind == 0)          // TS, MN-major B, D at +128
                    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                                 "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
                                 ::"r"(tmem + 128), "r"(tmem + 256 + (i & 7) * 8), "l"(vd[i & 7]), "r"(idesc | (1u << 16)), "r"(1u) : "memory");
                else if (kind == 1)     // TS, K-major B, D at 0
                    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                                 "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
                                 ::"r"(tmem), "r"(tmem + 256 + (i & 7) * 8), "l"(kd[i & 3]), "r"(idesc), "r"(1u) : "memory");
                else                    // SS, D at 0 (kind 2) or +128 (kind 3)
                    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                                 "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                                 ::"r"(tmem + (kind == 3 ? 128 : 0)), "l"(ad[i & 3]), "l"(kd[i & 3]), "r"(idesc), "r"(1u) : "memory");
            };
            const long long t0 = clock64();
            for (int it = 0; it < ITERS; it += 2 * GROUP) {
#pragma unroll
                for (int i = 0; i < GROUP; ++i) issue(KX, i);
#pragma unroll
                for (int i = 0; i < GROUP; ++i) issue(KY, i);
            }
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
            uint32_t ok = 0;
            while (!ok)
                asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                             : "=r"(ok) : "r"(smem_u32(&bar)) : "memory");
            cycles[blockIdx.x] = clock64() - t0;
            s_stop = 1;
        } else {
        const long long t0 = clock64();
        for (int it = 0; it < ITERS; ++it) {
            const uint32_t off = (it & 3) * 32;                // the four 16-element k-steps of a 64-element swizzle row
            uint64_t bd = smem_desc(b + off, 16, 1024);
            if (MODE == 2) bd = smem_desc(b + (it & 7) * 2048, 16384, 1024);       // 16 tokens x 128 d, MN-major
            if (TS) {
                // A = 128 x 16 bf16 from TMEM columns [256, 264); D = columns [0, N)
                asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                             "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
                             ::"r"(tmem), "r"(tmem + 256 + (it & 3) * 8), "l"(bd), "r"(idesc), "r"(1u) : "memory");
            } else {
                const uint64_t ad = smem_desc(a + off, 16, 1024);
                asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                             "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                             ::"r"(tmem), "l"(ad), "l"(bd), "r"(idesc), "r"(1u) : "memory");
            }
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
        uint32_t ok = 0;
        while (!ok)
            asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                         : "=r"(ok) : "r"(smem_u32(&bar)) : "memory");
        cycles[blockIdx.x] = clock64() - t0;
        s_stop = 1;
        }
    } else if (LOAD > 0 && tid >= 32) {
        // TMEM traffic of "softmax" warps: load 32 columns, store them back, on columns [384, 512) (not used by the MMAs)
        const uint32_t taddr = tmem + ((uint32_t)((tid >> 5) * 32) << 16) + 384 + ((tid >> 5) - 1) * 32;
        while (!s_stop) {
            uint32_t r[32];
            asm volatile(
                "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                  "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
                  "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
                  "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
                : "r"(taddr));
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
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
            asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (tid < 32) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem) : "memory");
    }
}

template <int N, int MODE, int LOAD = 0>
void run(const char* name, int sms, long long* d, long long* h) {
    constexpr bool TS = MODE == 1 || MODE == 2;
    const int smem = 16384 + 32768;
    cudaFuncSetAttribute(probe<N, MODE, LOAD>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    for (int rep = 0; rep < 2; ++rep) probe<N, MODE, LOAD><<<sms, 128, smem>>>(d);
    if (cudaDeviceSynchronize() != cudaSuccess) { printf("%s: %s\n", name, cudaGetErrorString(cudaGetLastError())); return; }
    cudaMemcpy(h, d, sms * sizeof(long long), cudaMemcpyDeviceToHost);
    double mean = 0;
    long long mn = h[0], mx = h[0];
    for (int i = 0; i < sms; ++i) { mean += h[i]; mn = h[i] < mn ? h[i] : mn; mx = h[i] > mx ? h[i] : mx; }
    mean /= sms;
    const double per = mean / ITERS;
    const double flop = 2.0 * 128 * N * 16;
    printf("%-22s %7.1f cycles per MMA (min %.1f max %.1f over %d SMs)  = %6.0f flop/clk/SM; operand bytes from smem per MMA %d -> %5.1f B/clk\n",
           name, per, (double)mn / ITERS, (double)mx / ITERS, sms, flop / per, (TS ? 0 : 4096) + N * 32,
           ((TS ? 0 : 4096) + N * 32) / per);
}

int main() {
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    long long *d, *h = new long long[sms];
    cudaMalloc(&d, sms * sizeof(long long));
    printf("tcgen05.mma cta_group::1 kind::f16, M = 128, K = 16, %d back-to-back instructions per SM, all SMs busy\n", ITERS);
    run<64, 0>("SS  128 x  64 x 16", sms, d, h);
    run<128, 0>("SS  128 x 128 x 16", sms, d, h);
    run<256, 0>("SS  128 x 256 x 16", sms, d, h);
    run<64, 1>("TS  128 x  64 x 16", sms, d, h);
    run<128, 1>("TS  128 x 128 x 16", sms, d, h);
    run<256, 1>("TS  128 x 256 x 16", sms, d, h);
    run<128, 2>("TS  128x128x16 B MN", sms, d, h);
    run<128, 3>("8 TS-MN | 8 SS (pair)", sms, d, h);
    run<128, 4>("8 TS-MN | 8 TS-K", sms, d, h);
    run<128, 5>("8 SS D0 | 8 SS D128", sms, d, h);
    run<128, 6>("8 TS-K  | 8 SS", sms, d, h);
    run<128, 7>("16 TS-MN | 16 SS", sms, d, h);
    run<128, 8>("32 TS-MN | 32 SS", sms, d, h);
    run<128, 0, 1>("SS  128x128 +TMEM io", sms, d, h);
    run<128, 3, 1>("pair        +TMEM io", sms, d, h);
    return 0;
}
