// gather_bw.cu -- read-only HBM ceiling for the ACCESS PATTERN of the paged decode kernel (diagnostic; not part of
// the library).  scripts/read_bw.cu streams one sequential buffer; the decode kernel instead gathers 4 KiB
// (block, kv head) tiles from TWO allocations (K and V) through a randomly permuted block table.  This program
// reproduces exactly that traffic with no consumer work, so its number is the ceiling the kernel is measured against:
//   * caches [num_blocks][Hkv][4 KiB] x 2 (config 2: 16 448 blocks x 8 heads = 514 MiB each, >> 126 MB L2)
//   * block table [B][max_nb] = a random permutation of the pool (as bench.py builds it)
//   * one producer thread per CTA walks its table row and issues cp.async.bulk copies of one tile each into a ring of
//     NSTAGE x 32 KiB stages (4 blocks x (K tile + V tile)), L2 evict-first; one consumer thread recycles stages
//   * "grid": one CTA per (sequence, kv head) item, 2 CTAs resident per SM (the single-pass schedule of the kernel)
//     "persistent": 2 x SMs CTAs, items dealt round-robin (no per-item CTA launch, ring kept full across items)
//     "sequential": the same CTAs reading the same number of tiles from consecutive addresses
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o gpurun_out/gather_bw scripts/gather_bw.cu && gpurun_out/gather_bw
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <algorithm>
#include <numeric>
#include <random>
#include <vector>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

constexpr int TILE = 4096;          // one (block, kv head) tile of 16 tokens x 128 bf16
constexpr int BLK_PER_STAGE = 4;    // 64 tokens
constexpr int STAGE_BYTES = 2 * BLK_PER_STAGE * TILE;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t ph) {
    uint32_t ok = 0;
    while (!ok) asm volatile("{.reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0,1,0,p;}"
                             : "=r"(ok) : "r"(bar), "r"(ph) : "memory");
}

// mode 0: gather through the table; mode 1: consecutive tiles (same copy size and count)
template <int NSTAGE>
__global__ void __launch_bounds__(64, 2)
gather_kernel(const uint8_t* __restrict__ kc, const uint8_t* __restrict__ vc, const int32_t* __restrict__ bt,
              int n_items, int Hkv, int max_nb, int nb_per_item, int persistent, int mode, int* sink) {
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
    const int first = blockIdx.x, step = persistent ? gridDim.x : n_items;   // grid mode: exactly one item
    const int stages_per_item = nb_per_item / BLK_PER_STAGE;
    if (tid == 0) {
        uint64_t pol;
        asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
        uint32_t it = 0;
        for (int item = first; item < n_items; item += step) {
            const int b = item / Hkv, h = item % Hkv;
            const int32_t* row = bt + (size_t)b * max_nb;
            for (int st = 0; st < stages_per_item; ++st, ++it) {
                const uint32_t s = it % NSTAGE, ph = (it / NSTAGE) & 1;
                mbar_wait(smem_u32(&empty[s]), ph ^ 1);
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&full[s])), "r"(STAGE_BYTES) : "memory");
#pragma unroll
                for (int j = 0; j < BLK_PER_STAGE; ++j) {
                    size_t tile;
                    if (mode == 0) tile = (size_t)row[st * BLK_PER_STAGE + j] * Hkv + h;
                    else tile = ((size_t)item * nb_per_item + st * BLK_PER_STAGE + j) * 2;   // consecutive 8 KiB pieces
                    const uint8_t* ks = kc + tile * TILE;
                    const uint8_t* vs = mode == 0 ? vc + tile * TILE : ks + TILE;
                    const uint32_t dst = smem_u32(smem + s * STAGE_BYTES + j * TILE);
                    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
                                 ::"r"(dst), "l"(ks), "r"(TILE), "r"(smem_u32(&full[s])), "l"(pol) : "memory");
                    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
                                 ::"r"(dst + BLK_PER_STAGE * TILE), "l"(vs), "r"(TILE), "r"(smem_u32(&full[s])), "l"(pol) : "memory");
                }
            }
        }
    } else if (tid == 32) {
        uint32_t it = 0;
        int acc = 0;
        for (int item = first; item < n_items; item += step)
            for (int st = 0; st < stages_per_item; ++st, ++it) {
                const uint32_t s = it % NSTAGE, ph = (it / NSTAGE) & 1;
                mbar_wait(smem_u32(&full[s]), ph);
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
    for (int i = 0; i < 5; ++i) launch();
    CK(cudaDeviceSynchronize());
    std::vector<float> ms;
    for (int i = 0; i < 30; ++i) {
        CK(cudaEventRecord(a)); launch(); CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b));
        float t; CK(cudaEventElapsedTime(&t, a, b)); ms.push_back(t);
    }
    CK(cudaGetLastError());
    std::sort(ms.begin(), ms.end());
    printf("%-44s best %8.1f GB/s   median %8.1f GB/s   (%.1f us)\n", name, bytes / ms[0] / 1e6, bytes / ms[15] / 1e6, ms[15] * 1e3);
}

int main(int argc, char** argv) {
    // default = BASELINE configs[1]: B 64, Hkv 8, ctx 4096 (256 blocks of 16 tokens); argv: B Hkv ctx
    const int B = argc > 1 ? atoi(argv[1]) : 64, Hkv = argc > 2 ? atoi(argv[2]) : 8, ctx = argc > 3 ? atoi(argv[3]) : 4096;
    const int max_nb = ctx / 16, num_blocks = B * max_nb + 64;
    const size_t cache_bytes = (size_t)num_blocks * Hkv * TILE;
    uint8_t *kc, *vc; int32_t* bt; int* sink;
    // one allocation, K then V (the sequential sweep runs through both halves); the engine's pool is laid out the same
    CK(cudaMalloc(&kc, 2 * cache_bytes)); CK(cudaMalloc(&sink, 4));
    vc = kc + cache_bytes;
    CK(cudaMemset(kc, 1, 2 * cache_bytes));
    std::vector<int32_t> perm(num_blocks);
    std::iota(perm.begin(), perm.end(), 0);
    std::mt19937 rng(0);
    std::shuffle(perm.begin(), perm.end(), rng);
    CK(cudaMalloc(&bt, (size_t)B * max_nb * 4));
    CK(cudaMemcpy(bt, perm.data(), (size_t)B * max_nb * 4, cudaMemcpyHostToDevice));
    int sms = 148; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    const int n_items = B * Hkv;
    const size_t bytes = (size_t)n_items * max_nb * 2 * TILE;
    printf("gather ceiling: B %d, Hkv %d, ctx %d -> %d items x %d tiles x 2 tensors = %.1f MB per launch; %d SMs\n",
           B, Hkv, ctx, n_items, max_nb, bytes / 1e6, sms);
    auto run = [&](auto kern, int nstage, const char* tag) {
        CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, nstage * STAGE_BYTES));
        char nm[96];
        snprintf(nm, sizeof nm, "gather, grid (1 CTA/item), %s", tag);
        timeit(nm, bytes, [&] { kern<<<n_items, 64, nstage * STAGE_BYTES>>>(kc, vc, bt, n_items, Hkv, max_nb, max_nb, 0, 0, sink); });
        snprintf(nm, sizeof nm, "gather, persistent 2 CTAs/SM, %s", tag);
        timeit(nm, bytes, [&] { kern<<<2 * sms, 64, nstage * STAGE_BYTES>>>(kc, vc, bt, n_items, Hkv, max_nb, max_nb, 1, 0, sink); });
        snprintf(nm, sizeof nm, "sequential, grid (1 CTA/item), %s", tag);
        timeit(nm, bytes, [&] { kern<<<n_items, 64, nstage * STAGE_BYTES>>>(kc, vc, bt, n_items, Hkv, max_nb, max_nb, 0, 1, sink); });
    };
    run(gather_kernel<3>, 3, "ring 3 x 32 KiB");
    run(gather_kernel<2>, 2, "ring 2 x 32 KiB");
    return 0;
}
