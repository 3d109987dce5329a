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

template <int N, bool TS>
__global__ void __launch_bounds__(128, 1) probe(long long* cycles) {
    extern __shared__ __align__(1024) uint8_t smem[];        // A: 128 x 64 bf16 (16 KiB) | B: N x 64 bf16
    __shared__ __align__(8) uint64_t bar;
    __shared__ uint32_t s_tmem;
    const int tid = threadIdx.x;
    for (int i = tid; i < (16384 + N * 128) / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0u;
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
        const uint32_t idesc = idesc_bf16(128, N);
        const long long t0 = clock64();
        for (int it = 0; it < ITERS; ++it) {
            const uint32_t off = (it & 3) * 32;                // the four 16-element k-steps of a 64-element swizzle row
            const uint64_t bd = smem_desc(b + off, 16, 1024);
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
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (tid < 32) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem) : "memory");
    }
}

template <int N, bool TS>
void run(const char* name, int sms, long long* d, long long* h) {
    const int smem = 16384 + N * 128;
    cudaFuncSetAttribute(probe<N, TS>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    for (int rep = 0; rep < 2; ++rep) probe<N, TS><<<sms, 128, smem>>>(d);
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
    run<64, false>("SS  128 x  64 x 16", sms, d, h);
    run<128, false>("SS  128 x 128 x 16", sms, d, h);
    run<256, false>("SS  128 x 256 x 16", sms, d, h);
    run<64, true>("TS  128 x  64 x 16", sms, d, h);
    run<128, true>("TS  128 x 128 x 16", sms, d, h);
    run<256, true>("TS  128 x 256 x 16", sms, d, h);
    return 0;
}
