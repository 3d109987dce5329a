// read_bw.cu -- read-only HBM stream ceiling on this box (diagnostic; not part of the library).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o gpurun_out/read_bw scripts/read_bw.cu && gpurun_out/read_bw
// Two variants over a 2 GiB buffer (>> 126 MB L2):
//   ldg : 16-byte ld.global.nc loads, 8 in flight per thread, grid = 148 x 8 CTAs of 256 threads
//   bulk: cp.async.bulk (TMA, 1-D) 4 KiB copies into a 3/6-stage shared-memory ring, 2 CTAs per SM -- the same
//         access machinery as the decode kernel, with no consumer work
// Prints GB/s for each (best and median of 20 launches, CUDA events).
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <algorithm>
#include <vector>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

__global__ void __launch_bounds__(256) ldg_kernel(const int4* __restrict__ src, size_t n16, int* sink) {
    int acc = 0;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + 7 * stride < n16; i += 8 * stride) {
        int4 v[8];
#pragma unroll
        for (int u = 0; u < 8; ++u)
            asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0,%1,%2,%3}, [%4];"
                         : "=r"(v[u].x), "=r"(v[u].y), "=r"(v[u].z), "=r"(v[u].w) : "l"(src + i + u * stride));
#pragma unroll
        for (int u = 0; u < 8; ++u) acc ^= v[u].x ^ v[u].y ^ v[u].z ^ v[u].w;
    }
    if (acc == 0x12345678) *sink = acc;
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

template <int NSTAGE, int STAGE_BYTES, int COPY_BYTES>
__global__ void __launch_bounds__(64, 2) bulk_kernel(const uint8_t* __restrict__ src, size_t bytes, int* sink) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ __align__(8) uint64_t full[NSTAGE], empty[NSTAGE];
    const int tid = threadIdx.x;
    if (tid == 0) {
        for (int s = 0; s < NSTAGE; ++s) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&full[s])));
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&empty[s])));
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const size_t nstage_total = bytes / STAGE_BYTES;
    const size_t per = nstage_total / gridDim.x;
    const size_t s0 = per * blockIdx.x, s1 = s0 + per;
    if (tid < 32) {          // producer warp, lane 0 issues
        if (tid == 0) {
            uint32_t it = 0;
            for (size_t st = s0; st < s1; ++st, ++it) {
                const uint32_t s = it % NSTAGE, ph = (it / NSTAGE) & 1;
                uint32_t ok = 0;
                while (!ok) asm volatile("{.reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0,1,0,p;}"
                                         : "=r"(ok) : "r"(smem_u32(&empty[s])), "r"(ph ^ 1) : "memory");
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&full[s])), "r"(STAGE_BYTES) : "memory");
                for (int c = 0; c < STAGE_BYTES / COPY_BYTES; ++c)
                    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                                 ::"r"(smem_u32(smem + s * STAGE_BYTES + c * COPY_BYTES)),
                                   "l"(src + st * STAGE_BYTES + (size_t)c * COPY_BYTES), "r"(COPY_BYTES), "r"(smem_u32(&full[s])) : "memory");
            }
        }
    } else if (tid == 32) {  // consumer: just recycle the stage
        uint32_t it = 0;
        int acc = 0;
        for (size_t st = s0; st < s1; ++st, ++it) {
            const uint32_t s = it % NSTAGE, ph = (it / NSTAGE) & 1;
            uint32_t ok = 0;
            while (!ok) asm volatile("{.reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0,1,0,p;}"
                                     : "=r"(ok) : "r"(smem_u32(&full[s])), "r"(ph) : "memory");
            acc ^= *reinterpret_cast<volatile int*>(smem + s * STAGE_BYTES);
            asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&empty[s])) : "memory");
        }
        if (acc == 0x12345678) *sink = acc;
    }
}

template <class F>
void timeit(const char* name, size_t bytes, F launch) {
    cudaEvent_t a, b;
    CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
    for (int i = 0; i < 3; ++i) launch();
    CK(cudaDeviceSynchronize());
    std::vector<float> ms;
    for (int i = 0; i < 20; ++i) {
        CK(cudaEventRecord(a)); launch(); CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b));
        float t; CK(cudaEventElapsedTime(&t, a, b)); ms.push_back(t);
    }
    CK(cudaGetLastError());
    std::sort(ms.begin(), ms.end());
    printf("%-28s best %8.1f GB/s   median %8.1f GB/s   (%.1f us)\n", name, bytes / ms[0] / 1e6, bytes / ms[10] / 1e6, ms[10] * 1e3);
}

int main() {
    const size_t bytes = (size_t)2 << 30;
    uint8_t* buf; int* sink;
    CK(cudaMalloc(&buf, bytes)); CK(cudaMalloc(&sink, 4));
    CK(cudaMemset(buf, 1, bytes));
    int sms = 148; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    for (int mult : {4, 8, 16}) {
        char nm[64]; snprintf(nm, sizeof nm, "ldg x%d CTAs/SM", mult);
        timeit(nm, bytes, [&] { ldg_kernel<<<sms * mult, 256>>>((const int4*)buf, bytes / 16, sink); });
    }
    {
        auto k = bulk_kernel<3, 32768, 4096>;
        CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 3 * 32768));
        timeit("bulk 3x32K, 4K copies", bytes, [&] { k<<<sms * 2, 64, 3 * 32768>>>(buf, bytes, sink); });
    }
    {
        auto k = bulk_kernel<3, 32768, 2048>;
        CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 3 * 32768));
        timeit("bulk 3x32K, 2K copies", bytes, [&] { k<<<sms * 2, 64, 3 * 32768>>>(buf, bytes, sink); });
    }
    {
        auto k = bulk_kernel<6, 16384, 4096>;
        CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 6 * 16384));
        timeit("bulk 6x16K, 4K copies", bytes, [&] { k<<<sms * 2, 64, 6 * 16384>>>(buf, bytes, sink); });
    }
    {
        auto k = bulk_kernel<3, 65536, 8192>;
        CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 3 * 65536));
        timeit("bulk 3x64K 1CTA/SM, 8K", bytes, [&] { k<<<sms, 64, 3 * 65536>>>(buf, bytes, sink); });
    }
    return 0;
}
