// Instruction-rate probe for the fp8 decode design choice:
//   (a) mma.sync m16n8k16 f16      (what the fp8 path uses: K/V converted e4m3 -> f16 in registers)
//   (b) mma.sync m16n8k32 e4m3     (K/V fed to the MMA as stored, q and P split into e4m3 terms)
//   (c) cvt.rn.f16x2.e4m3x2        (the conversion (a) needs: one per two cache elements)
// Finding (B200, CUDA 12.9): sm_100a has no native legacy fp8 MMA.  `cuobjdump -sass` of this file shows (b) lowered
// to F2FP.F16.E4M3.UNPACK_B + two HMMA.16816.F32 per instruction; in this loop the operands are loop-invariant, so
// ptxas hoists the conversions and (b) *looks* 1.9x faster per instruction than (a) -- in a real kernel the
// conversion is paid per element either way.  (c) confirms F2FP runs on the half-rate ALU pipe (~3.9 G warp
// instructions/s/SM with its companions).  Native fp8 on Blackwell = tcgen05.mma.kind::f8f6f4.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scripts/bin/mma_rate scripts/mma_rate.cu ; run on the GPU box.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

constexpr int ITERS = 4096;
constexpr int ILP = 8;

__global__ void k_f16(float* out) {
    float c[ILP][4] = {};
    uint32_t a0 = threadIdx.x, a1 = a0 * 3, a2 = a0 * 5, a3 = a0 * 7, b0 = a0 * 11, b1 = a0 * 13;
    for (int i = 0; i < ITERS; ++i) {
#pragma unroll
        for (int j = 0; j < ILP; ++j)
            asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                         : "+f"(c[j][0]), "+f"(c[j][1]), "+f"(c[j][2]), "+f"(c[j][3])
                         : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
    }
    float s = 0;
    for (int j = 0; j < ILP; ++j) s += c[j][0] + c[j][1] + c[j][2] + c[j][3];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void k_e4m3(float* out) {
    float c[ILP][4] = {};
    uint32_t a0 = threadIdx.x, a1 = a0 * 3, a2 = a0 * 5, a3 = a0 * 7, b0 = a0 * 11, b1 = a0 * 13;
    for (int i = 0; i < ITERS; ++i) {
#pragma unroll
        for (int j = 0; j < ILP; ++j)
            asm volatile("mma.sync.aligned.m16n8k32.row.col.f32.e4m3.e4m3.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                         : "+f"(c[j][0]), "+f"(c[j][1]), "+f"(c[j][2]), "+f"(c[j][3])
                         : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
    }
    float s = 0;
    for (int j = 0; j < ILP; ++j) s += c[j][0] + c[j][1] + c[j][2] + c[j][3];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void k_cvt(uint32_t* out) {
    uint32_t v[ILP];
#pragma unroll
    for (int j = 0; j < ILP; ++j) v[j] = threadIdx.x * (2 * j + 1) + 0x3838;
    for (int i = 0; i < ITERS; ++i) {
#pragma unroll
        for (int j = 0; j < ILP; ++j) {
            uint32_t r;
            asm volatile("{\n\t.reg .b16 lo;\n\tcvt.u16.u32 lo, %1;\n\tcvt.rn.f16x2.e4m3x2 %0, lo;\n\t}" : "=r"(r) : "r"(v[j]));
            v[j] = r ^ (v[j] >> 16);
        }
    }
    uint32_t s = 0;
    for (int j = 0; j < ILP; ++j) s ^= v[j];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <class F>
float time_ms(F launch) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    launch();
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    launch();
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    return ms;
}

int main() {
    cudaDeviceProp prop;
    cudaGetDeviceProperties(&prop, 0);
    const int sms = prop.multiProcessorCount;
    float* out;
    cudaMalloc(&out, sizeof(float) * sms * 8 * 256);
    for (int warps : {4, 8, 16}) {
        const int threads = warps * 32, grid = sms;
        const double n_inst = (double)grid * warps * ITERS * ILP;   // warp instructions
        float a = time_ms([&] { k_f16<<<grid, threads>>>(out); });
        float b = time_ms([&] { k_e4m3<<<grid, threads>>>(out); });
        float c = time_ms([&] { k_cvt<<<grid, threads>>>((uint32_t*)out); });
        printf("warps/SM %2d: f16 k16 %.3f ms = %.1f TF/s (%.2f Ginst/s/SM) | e4m3 k32 %.3f ms = %.1f TF/s (%.2f Ginst/s/SM) | cvt+xor+shift %.3f ms (%.2f Ginst/s/SM of cvt)\n",
               warps, a, n_inst * 2 * 16 * 8 * 16 / a * 1e-9, n_inst / a * 1e-6 / sms, b, n_inst * 2 * 16 * 8 * 32 / b * 1e-9,
               n_inst / b * 1e-6 / sms, c, n_inst / c * 1e-6 / sms);
    }
    cudaError_t err = cudaGetLastError();
    printf("status: %s\n", cudaGetErrorString(err));
    return err != cudaSuccess;
}
