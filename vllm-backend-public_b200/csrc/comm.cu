// comm.cu -- tensor-parallel exchange of one decoder layer over NVLink peer memory (SURVEY.md section 8 rows a4 / e:
// `--tensor-parallel-size N`, reference benchmark-vllm/models.py:11 -> vllm_templates.py:55,226).
//
// With the heads (and the MLP columns) sharded over tp GPUs, `o_proj` and `down_proj` are row-parallel: every rank
// holds a PARTIAL [T, hidden] sum, and the layer needs  h = resid + sum_r partial_r ;  x = RMSNorm(h) * w  next.
// For a decode step that message is tiny (B x hidden bf16 = 512 KiB at B = 64) and the step has 2 x layers of them,
// so the cost is launch + synchronisation latency, not bytes.  One kernel does all of it:
//
//   b2_comm_allreduce_rms_norm   one-shot all-reduce over peer memory fused with the residual add and the RMSNorm:
//       every rank's GEMM writes its partial straight into its own peer-mapped buffer; the kernel tells the peers
//       (one 4-byte store per peer), waits for theirs, then each CTA reads its row from all tp buffers with 16-byte
//       peer loads over NVLink / NVSwitch, sums in rank order (bit-identical on every rank), adds the residual,
//       stores the new residual and the normalised row.  No NCCL call, no intermediate buffer, CUDA-graph capturable
//       (the barrier epoch lives in device memory).
//   b2_comm_allgather_cols       the vocab-parallel lm_head: every rank stores its [T, V/tp] logit columns into every
//       peer's [T, V] buffer (16-byte peer stores), then one barrier.
//
// Memory model: buffers are plain cudaMalloc regions exported with cudaIpcGetMemHandle (one process per GPU) -- the one
// place where the library allocates device memory, because the allocation must be exportable; the caller owns the
// pointer and frees it (b2_comm_free).  Synchronisation = one 32-bit arrival flag per (destination, source) pair,
// written with st.release.sys and polled with ld.acquire.sys; flags carry a monotonically increasing epoch, so
// nothing is ever reset and a graph replay needs no host-side state.
//
// Why reuse is safe with two partial-sum buffers (slot 0 after o_proj, slot 1 after down_proj): a rank can only leave
// barrier e+1 after every peer has entered it, i.e. after every peer finished the kernel of barrier e and with it every
// read of the slot used there; the slot is next written two barriers later.
#include "b2_common.cuh"

#include <stddef.h>
#include <stdlib.h>
#include <string.h>

#include <new>

namespace {

constexpr int MAX_TP = B2_COMM_MAX_RANKS;
constexpr size_t CTRL_BYTES = 4096;
constexpr int PUSH_MAX_ROWS = 256;   // <= 2 CTAs x 148 SMs (see push_allreduce_rms_norm_kernel)

struct Ctrl {                       // first CTRL_BYTES of every rank's region
    uint32_t epoch;                 // barriers this rank has completed (only this rank's kernels write it)
    uint32_t done;                  // CTAs of the running kernel that have finished
    uint32_t lam;                   // landing buffer of the next push exchange: 0, 1, 2, 0, ... (kept below 3, so no wrap)
    uint32_t pad[13];
    uint64_t dbg[8];                // B2_COMM_DEBUG=1: globaltimer stamps of CTA 0 / thread 0 of the last push exchange
    uint32_t flag[MAX_TP][32];      // flag[src][0] = last epoch rank `src` has reached (one 128-byte line per source)
};
static_assert(sizeof(Ctrl) <= CTRL_BYTES, "control block");

struct Peers {
    char* base[MAX_TP];
};

struct Comm {
    int rank, tp, max_rows, H, gather_rows, gather_cols, debug, weak_st;
    size_t region_bytes, off_in[2], off_gather, off_lam, lam_src_bytes, lam_buf_bytes;
    Peers peers;
};

__device__ __forceinline__ float bf_lo(uint32_t w) { return __uint_as_float(w << 16); }
__device__ __forceinline__ float bf_hi(uint32_t w) { return __uint_as_float(w & 0xffff0000u); }
__device__ __forceinline__ void unpack8(const int4& v, float (&f)[8]) {
    const uint32_t w[4] = {(uint32_t)v.x, (uint32_t)v.y, (uint32_t)v.z, (uint32_t)v.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        f[2 * i] = bf_lo(w[i]);
        f[2 * i + 1] = bf_hi(w[i]);
    }
}
__device__ __forceinline__ int4 pack8(const float (&f)[8]) {
    int4 r;
    r.x = (int)b2::pack_bf16x2(f[0], f[1]);
    r.y = (int)b2::pack_bf16x2(f[2], f[3]);
    r.z = (int)b2::pack_bf16x2(f[4], f[5]);
    r.w = (int)b2::pack_bf16x2(f[6], f[7]);
    return r;
}

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ uint32_t ld_volatile(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ uint64_t globaltimer_ns() {
    uint64_t t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
// Bounded wait for a peer's arrival flag: a peer that died traps this kernel after ~10 s (the engine exits non-zero
// and is restarted by its supervisor) instead of hanging the GPU.
__device__ __forceinline__ void wait_flag(const uint32_t* p, uint32_t epoch) {
    if ((int32_t)(ld_acquire_sys(p) - epoch) >= 0) return;
    const uint64_t t0 = globaltimer_ns();
    while ((int32_t)(ld_acquire_sys(p) - epoch) < 0) {
        if (globaltimer_ns() - t0 > 10000000000ull) __trap();
    }
}
// 16-byte load that bypasses L1 (peer data changes between launches; L1 is not coherent with remote writers)
__device__ __forceinline__ int4 ld_cg_16(const void* p) {
    int4 r;
    asm volatile("ld.global.cg.v4.s32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
// peers.base[i] for a runtime i without spilling the by-value parameter array to local memory
template <int TP>
__device__ __forceinline__ char* peer_base(const Peers& peers, int i) {
    char* b = peers.base[0];
#pragma unroll
    for (int p = 1; p < TP; ++p) b = (i == p) ? peers.base[p] : b;
    return b;
}

// ---------------------------------------------------------------------------------------------
// one-shot all-reduce + residual add + RMSNorm.  grid = T rows, block = nvec / VPT threads (nvec = H / 8).
//   partial_r = region_r + in_off, [T, H] bf16 contiguous, written by rank r's GEMM before this kernel on its stream
//   h    = bf16( sum_{r = 0..TP-1} partial_r  (+ residual) )          -> residual (when given)
//   out  = weight ? bf16(weight * h * rsqrt(mean(h^2) + eps)) : h
// ---------------------------------------------------------------------------------------------
template <int TP, int VPT>
__global__ void __launch_bounds__(512)
allreduce_rms_norm_kernel(const Peers peers, int rank, size_t in_off, __nv_bfloat16* __restrict__ out,
                          __nv_bfloat16* __restrict__ residual, const __nv_bfloat16* __restrict__ weight, int H,
                          int64_t out_stride, int64_t res_stride, float eps) {
    Ctrl* me = reinterpret_cast<Ctrl*>(peer_base<TP>(peers, rank));
    __shared__ uint32_t s_epoch;
    __shared__ float red[32];
    const int row = blockIdx.x;
    const int nvec = H >> 3;

    // loads that do not depend on the peers go out before the barrier: residual row and norm weight
    int4 rv[VPT], wv[VPT];
#pragma unroll
    for (int k = 0; k < VPT; ++k) {
        const int v = threadIdx.x + k * blockDim.x;
        rv[k] = wv[k] = make_int4(0, 0, 0, 0);
        if (v < nvec) {
            if (residual) rv[k] = *reinterpret_cast<const int4*>(residual + row * res_stride + v * 8);
            if (weight) wv[k] = *reinterpret_cast<const int4*>(weight + v * 8);
        }
    }

    if (threadIdx.x == 0) s_epoch = ld_volatile(&me->epoch) + 1;
    __syncthreads();
    const uint32_t e = s_epoch;
    if (TP > 1) {
        // CTA 0 announces this rank: its partial sums were written by the previous kernel on this stream
        if (blockIdx.x == 0 && threadIdx.x < TP) {
            __threadfence_system();
            st_release_sys(&reinterpret_cast<Ctrl*>(peer_base<TP>(peers, threadIdx.x))->flag[rank][0], e);
        }
        if (threadIdx.x < TP) wait_flag(&me->flag[threadIdx.x][0], e);
        __syncthreads();
    }

    float h[VPT][8];
    float ss = 0.f;
#pragma unroll
    for (int k = 0; k < VPT; ++k) {
        const int v = threadIdx.x + k * blockDim.x;
        if (v < nvec) {
            const size_t off = in_off + ((size_t)row * H + (size_t)v * 8) * sizeof(__nv_bfloat16);
            int4 pv[TP];
#pragma unroll
            for (int p = 0; p < TP; ++p) pv[p] = ld_cg_16(peers.base[p] + off);     // all peer loads in flight together
#pragma unroll
            for (int i = 0; i < 8; ++i) h[k][i] = 0.f;
#pragma unroll
            for (int p = 0; p < TP; ++p) {                                          // rank order: same bits on every rank
                float f[8];
                unpack8(pv[p], f);
#pragma unroll
                for (int i = 0; i < 8; ++i) h[k][i] += f[i];
            }
            if (residual) {
                float r[8];
                unpack8(rv[k], r);
#pragma unroll
                for (int i = 0; i < 8; ++i) h[k][i] += r[i];
            }
            const int4 packed = pack8(h[k]);               // the residual stream is bf16: round once, then norm
            if (residual) *reinterpret_cast<int4*>(residual + row * res_stride + v * 8) = packed;
            if (!weight) *reinterpret_cast<int4*>(out + row * out_stride + v * 8) = packed;
            unpack8(packed, h[k]);
#pragma unroll
            for (int i = 0; i < 8; ++i) ss += h[k][i] * h[k][i];
        }
    }
    if (weight) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
        if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
        __syncthreads();
        float tot = 0.f;
        const int nw = (blockDim.x + 31) >> 5;
        for (int i = 0; i < nw; ++i) tot += red[i];
        const float inv = rsqrtf(tot / (float)H + eps);
#pragma unroll
        for (int k = 0; k < VPT; ++k) {
            const int v = threadIdx.x + k * blockDim.x;
            if (v < nvec) {
                float w[8], o[8];
                unpack8(wv[k], w);
#pragma unroll
                for (int i = 0; i < 8; ++i) o[i] = w[i] * (h[k][i] * inv);
                *reinterpret_cast<int4*>(out + row * out_stride + v * 8) = pack8(o);
            }
        }
    }
    // the last CTA to finish publishes the epoch for the next kernel on this stream
    if (threadIdx.x == 0) {
        __threadfence();
        if (atomicAdd(&me->done, 1u) == gridDim.x - 1) {
            me->done = 0;
            me->epoch = e;
            __threadfence();
        }
    }
}

// ---------------------------------------------------------------------------------------------
// all-gather of column blocks by peer stores: gathered_p[t][rank*cols + c] = local[t][c] for every peer p, then a
// barrier (the last CTA to finish its stores signals every peer and waits for theirs).
// ---------------------------------------------------------------------------------------------
template <int TP>
__global__ void __launch_bounds__(256)
allgather_cols_kernel(const Peers peers, int rank, size_t gather_off, const __nv_bfloat16* __restrict__ local,
                      int64_t local_stride, int T, int cols) {
    Ctrl* me = reinterpret_cast<Ctrl*>(peer_base<TP>(peers, rank));
    __shared__ uint32_t s_last, s_epoch;
    const int cvec = cols >> 3;
    const int64_t total = (int64_t)T * cvec;
    const int64_t row_bytes = (int64_t)TP * cols * sizeof(__nv_bfloat16);
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int t = (int)(i / cvec), v = (int)(i - (int64_t)t * cvec);
        const int4 x = b2::ld_nc_16(local + t * local_stride + v * 8);
        const size_t off = gather_off + (size_t)t * row_bytes + ((size_t)rank * cols + (size_t)v * 8) * sizeof(__nv_bfloat16);
#pragma unroll
        for (int p = 0; p < TP; ++p) *reinterpret_cast<int4*>(peers.base[p] + off) = x;
    }
    __threadfence_system();                 // this thread's peer stores are performed before the CTA counts as done
    __syncthreads();
    if (threadIdx.x == 0) {
        s_last = atomicAdd(&me->done, 1u) == gridDim.x - 1;
        s_epoch = ld_volatile(&me->epoch) + 1;
    }
    __syncthreads();
    if (!s_last) return;
    const uint32_t e = s_epoch;
    if (TP > 1 && threadIdx.x < TP) {
        __threadfence_system();
        st_release_sys(&reinterpret_cast<Ctrl*>(peer_base<TP>(peers, threadIdx.x))->flag[rank][0], e);
        wait_flag(&me->flag[threadIdx.x][0], e);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        me->done = 0;
        me->epoch = e;
        __threadfence();
    }
}

// ---------------------------------------------------------------------------------------------
// PUSH variant of the same exchange (the default): no barrier, no flag, one NVLink crossing on the critical path.
// Every rank stores its partial row vectors straight into a landing buffer on every peer; a landing slot that still
// holds the sentinel (bf16 -0.0 in every lane) has not arrived yet, so the consumer polls the very 16 bytes it needs
// ("Lamport" style: the data is its own flag).  A partial sum that really is -0.0 is sent as +0.0 -- the same value.
//   landing area of rank d: lam[3][TP src][max_rows][H] bf16; call k (device counter, equal on all ranks) uses
//   buffer k % 3.  A thread re-arms exactly the slots it consumed (stores the sentinel back) right after reading
//   them.  Reuse is safe: a peer can run at most one call ahead (it cannot finish call k+1 without this rank's
//   call-k+1 data), so buffer k % 3 is next written in call k+3, long after this rank's call k has completed.
// Validity is checked on all 8 lanes of a vector, so no store atomicity beyond 2 bytes is assumed.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ bool has_sentinel(const int4& v) {
    const uint32_t w[4] = {(uint32_t)v.x, (uint32_t)v.y, (uint32_t)v.z, (uint32_t)v.w};
    bool s = false;
#pragma unroll
    for (int i = 0; i < 4; ++i) s |= ((w[i] & 0xffffu) == 0x8000u) | ((w[i] >> 16) == 0x8000u);
    return s;
}
__device__ __forceinline__ int4 no_neg_zero(int4 v) {
    uint32_t w[4] = {(uint32_t)v.x, (uint32_t)v.y, (uint32_t)v.z, (uint32_t)v.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        if ((w[i] & 0xffffu) == 0x8000u) w[i] &= 0xffff0000u;
        if ((w[i] >> 16) == 0x8000u) w[i] &= 0x0000ffffu;
    }
    return make_int4((int)w[0], (int)w[1], (int)w[2], (int)w[3]);
}
__device__ __forceinline__ int4 ld_volatile_16(const void* p) {
    int4 r;
    asm volatile("ld.volatile.global.v4.s32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ void st_volatile_16(void* p, const int4& v) {
    asm volatile("st.volatile.global.v4.s32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
                 : "memory");
}

__global__ void fill_u32_kernel(uint32_t* p, size_t n, uint32_t v) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) p[i] = v;
}

template <int TP, int VPT>
__global__ void __launch_bounds__(512, 2)      // two CTAs per SM: all PUSH_MAX_ROWS CTAs of a launch are resident, so
push_allreduce_rms_norm_kernel(                // no CTA can spin for data that a not-yet-scheduled peer CTA must push
const Peers peers, int rank, size_t lam_off, size_t lam_src_bytes, size_t lam_buf_bytes,
                               const __nv_bfloat16* __restrict__ partial, int64_t partial_stride,
                               __nv_bfloat16* __restrict__ out, __nv_bfloat16* __restrict__ residual,
                               const __nv_bfloat16* __restrict__ weight, int H, int64_t out_stride, int64_t res_stride,
                               float eps, int flags) {
    char* mine = peer_base<TP>(peers, rank);
    const bool dbg = (flags & 1) && blockIdx.x == 0 && threadIdx.x == 0;
    uint64_t ts[4];
    if (dbg) ts[0] = globaltimer_ns();
    Ctrl* me = reinterpret_cast<Ctrl*>(mine);
    __shared__ uint32_t s_lam;
    __shared__ float red[32];
    const int row = blockIdx.x;
    const int nvec = H >> 3;

    // this rank's partial row (written by the previous kernel on this stream), residual row and norm weight
    int4 xv[VPT], rv[VPT], wv[VPT];
#pragma unroll
    for (int k = 0; k < VPT; ++k) {
        const int v = threadIdx.x + k * blockDim.x;
        xv[k] = rv[k] = wv[k] = make_int4(0, 0, 0, 0);
        if (v < nvec) {
            xv[k] = *reinterpret_cast<const int4*>(partial + row * partial_stride + v * 8);
            if (residual) rv[k] = *reinterpret_cast<const int4*>(residual + row * res_stride + v * 8);
            if (weight) wv[k] = *reinterpret_cast<const int4*>(weight + v * 8);
        }
    }
    if (threadIdx.x == 0) s_lam = ld_volatile(&me->lam);
    __syncthreads();
    const uint32_t call = s_lam;                                  // 0, 1 or 2
    const size_t buf_off = lam_off + (size_t)call * lam_buf_bytes;

    // push: one 16-byte store per peer and vector, into the slot [buffer][src = rank][row][v] of every peer
#pragma unroll
    for (int k = 0; k < VPT; ++k) {
        const int v = threadIdx.x + k * blockDim.x;
        if (v < nvec) {
            xv[k] = no_neg_zero(xv[k]);
            const size_t off = buf_off + (size_t)rank * lam_src_bytes + ((size_t)row * H + (size_t)v * 8) * sizeof(__nv_bfloat16);
#pragma unroll
            for (int p = 1; p < TP; ++p) {                         // start with the next rank: spreads the switch ports
                const int dst = (rank + p) % TP;
                if (flags & 2) *reinterpret_cast<int4*>(peer_base<TP>(peers, dst) + off) = xv[k];
                else st_volatile_16(peer_base<TP>(peers, dst) + off, xv[k]);
            }
        }
    }
    if (dbg) ts[1] = globaltimer_ns();

    float h[VPT][8];
    float ss = 0.f;
    const int4 sentinel = make_int4((int)0x80008000u, (int)0x80008000u, (int)0x80008000u, (int)0x80008000u);
#pragma unroll
    for (int k = 0; k < VPT; ++k) {
        const int v = threadIdx.x + k * blockDim.x;
        if (v < nvec) {
            const size_t off = buf_off + ((size_t)row * H + (size_t)v * 8) * sizeof(__nv_bfloat16);
            int4 pv[TP];
            uint32_t pending = 0;
#pragma unroll
            for (int p = 0; p < TP; ++p) {
                if (p == rank) { pv[p] = xv[k]; continue; }
                pv[p] = ld_volatile_16(mine + off + (size_t)p * lam_src_bytes);
                pending |= has_sentinel(pv[p]) ? (1u << p) : 0u;
            }
            if (pending) {
                const uint64_t t0 = globaltimer_ns();
                while (pending) {
#pragma unroll
                    for (int p = 0; p < TP; ++p) {
                        if (pending & (1u << p)) {
                            pv[p] = ld_volatile_16(mine + off + (size_t)p * lam_src_bytes);
                            if (!has_sentinel(pv[p])) pending &= ~(1u << p);
                        }
                    }
                    if (pending && globaltimer_ns() - t0 > 10000000000ull) __trap();
                }
            }
            if (dbg && k == 0) ts[2] = globaltimer_ns();
#pragma unroll
            for (int p = 0; p < TP; ++p)                            // re-arm the consumed slots for call k + 3
                if (p != rank) st_volatile_16(mine + off + (size_t)p * lam_src_bytes, sentinel);
#pragma unroll
            for (int i = 0; i < 8; ++i) h[k][i] = 0.f;
#pragma unroll
            for (int p = 0; p < TP; ++p) {                          // rank order: same bits on every rank
                float f[8];
                unpack8(pv[p], f);
#pragma unroll
                for (int i = 0; i < 8; ++i) h[k][i] += f[i];
            }
            if (residual) {
                float r[8];
                unpack8(rv[k], r);
#pragma unroll
                for (int i = 0; i < 8; ++i) h[k][i] += r[i];
            }
            const int4 packed = pack8(h[k]);
            if (residual) *reinterpret_cast<int4*>(residual + row * res_stride + v * 8) = packed;
            if (!weight) *reinterpret_cast<int4*>(out + row * out_stride + v * 8) = packed;
            unpack8(packed, h[k]);
#pragma unroll
            for (int i = 0; i < 8; ++i) ss += h[k][i] * h[k][i];
        }
    }
    if (weight) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
        if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
        __syncthreads();
        float tot = 0.f;
        const int nw = (blockDim.x + 31) >> 5;
        for (int i = 0; i < nw; ++i) tot += red[i];
        const float inv = rsqrtf(tot / (float)H + eps);
#pragma unroll
        for (int k = 0; k < VPT; ++k) {
            const int v = threadIdx.x + k * blockDim.x;
            if (v < nvec) {
                float w[8], o[8];
                unpack8(wv[k], w);
#pragma unroll
                for (int i = 0; i < 8; ++i) o[i] = w[i] * (h[k][i] * inv);
                *reinterpret_cast<int4*>(out + row * out_stride + v * 8) = pack8(o);
            }
        }
    }
    if (dbg) {
        ts[3] = globaltimer_ns();
        for (int i = 0; i < 4; ++i) me->dbg[i] = ts[i];
    }
    if (threadIdx.x == 0) {
        __threadfence();
        if (atomicAdd(&me->done, 1u) == gridDim.x - 1) {
            me->done = 0;
            me->lam = call == 2u ? 0u : call + 1u;
            __threadfence();
        }
    }
}

template <int TP>
void launch_push(const Comm* c, const __nv_bfloat16* partial, int64_t partial_stride, __nv_bfloat16* out,
                 __nv_bfloat16* residual, const __nv_bfloat16* weight, int T, int64_t out_stride, int64_t res_stride,
                 float eps, cudaStream_t st) {
    const int nvec = c->H / 8;
#define B2_PUSH(VPT, THREADS)                                                                                         \
    push_allreduce_rms_norm_kernel<TP, VPT><<<T, THREADS, 0, st>>>(c->peers, c->rank, c->off_lam, c->lam_src_bytes,    \
                                                                   c->lam_buf_bytes, partial, partial_stride, out,     \
                                                                   residual, weight, c->H, out_stride, res_stride, eps, \
                                                                   c->debug | (c->weak_st << 1))
    if (nvec <= 512) {
        B2_PUSH(1, ((nvec + 31) / 32) * 32);
    } else if (nvec <= 1024) {
        B2_PUSH(2, 512);
    } else {
        B2_PUSH(4, 512);
    }
#undef B2_PUSH
}

template <int TP>
void launch_allreduce(const Comm* c, int slot, __nv_bfloat16* out, __nv_bfloat16* residual, const __nv_bfloat16* weight,
                      int T, int64_t out_stride, int64_t res_stride, float eps, cudaStream_t st) {
    const int nvec = c->H / 8;            // H <= 16384: at most 2048 vectors of 8 per row
    if (nvec <= 512) {
        const int threads = ((nvec + 31) / 32) * 32;
        allreduce_rms_norm_kernel<TP, 1><<<T, threads, 0, st>>>(c->peers, c->rank, c->off_in[slot], out, residual, weight,
                                                                c->H, out_stride, res_stride, eps);
    } else if (nvec <= 1024) {
        allreduce_rms_norm_kernel<TP, 2><<<T, 512, 0, st>>>(c->peers, c->rank, c->off_in[slot], out, residual, weight,
                                                            c->H, out_stride, res_stride, eps);
    } else {
        allreduce_rms_norm_kernel<TP, 4><<<T, 512, 0, st>>>(c->peers, c->rank, c->off_in[slot], out, residual, weight,
                                                            c->H, out_stride, res_stride, eps);
    }
}

template <int TP>
void launch_allgather(const Comm* c, const __nv_bfloat16* local, int64_t local_stride, int T, int cols, cudaStream_t st) {
    const int64_t total = (int64_t)T * (cols / 8);
    int grid = (int)((total + 255) / 256);
    if (grid > 296) grid = 296;            // two CTAs per SM keep enough peer stores in flight
    if (grid < 1) grid = 1;
    allgather_cols_kernel<TP><<<grid, 256, 0, st>>>(c->peers, c->rank, c->off_gather, local, local_stride, T, cols);
}

size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

}  // namespace

// =============================================================================================
// C ABI
// =============================================================================================
extern "C" int b2_comm_alloc(size_t bytes, void** ptr, void* handle) {
    B2_REQUIRE(bytes > 0 && ptr && handle, B2_EINVAL, "comm_alloc: bad arguments");
    static_assert(sizeof(cudaIpcMemHandle_t) == B2_COMM_HANDLE_BYTES, "IPC handle size");
    void* p = nullptr;
    B2_CUDA_OK(cudaMalloc(&p, bytes));
    cudaError_t e = cudaMemset(p, 0, bytes);
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    if (e == cudaSuccess) e = cudaIpcGetMemHandle(static_cast<cudaIpcMemHandle_t*>(handle), p);
    if (e != cudaSuccess) {
        cudaFree(p);
        b2_set_error("comm_alloc: %s", cudaGetErrorString(e));
        return B2_ECUDA;
    }
    *ptr = p;
    return B2_OK;
}

extern "C" int b2_comm_open(const void* handle, void** ptr) {
    B2_REQUIRE(handle && ptr, B2_EINVAL, "comm_open: null pointer");
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, sizeof(h));
    B2_CUDA_OK(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
    return B2_OK;
}

extern "C" int b2_comm_close(void* ptr) {
    if (ptr) B2_CUDA_OK(cudaIpcCloseMemHandle(ptr));
    return B2_OK;
}

extern "C" int b2_comm_free(void* ptr) {
    if (ptr) B2_CUDA_OK(cudaFree(ptr));
    return B2_OK;
}

extern "C" size_t b2_comm_region_bytes(int tp, int max_rows, int H, int gather_rows, int gather_cols) {
    if (tp < 1 || tp > MAX_TP || max_rows <= 0 || H <= 0 || gather_rows < 0 || gather_cols < 0) return 0;
    const size_t in = align_up((size_t)max_rows * H * sizeof(__nv_bfloat16), 256);
    // control | 2 partial-sum slots (pull variant) | gathered logits | 3 landing buffers x tp sources (push variant)
    return CTRL_BYTES + 2 * in + align_up((size_t)gather_rows * gather_cols * sizeof(__nv_bfloat16), 256) +
           3 * (size_t)tp * in;
}

extern "C" void* b2_comm_create(int rank, int tp, void* const* regions, size_t region_bytes, int max_rows, int H,
                                int gather_rows, int gather_cols) {
    if (!(tp == 1 || tp == 2 || tp == 4 || tp == 8) || rank < 0 || rank >= tp || !regions || max_rows <= 0 || H <= 0 ||
        H % 8 || H > 16384 || gather_rows < 0 || gather_cols < 0 || gather_cols % 8) {
        b2_set_error("comm_create: bad arguments (tp=%d rank=%d max_rows=%d H=%d gather=%dx%d; tp in {1,2,4,8}, "
                     "H %% 8 == 0, H <= 16384)", tp, rank, max_rows, H, gather_rows, gather_cols);
        return nullptr;
    }
    if (region_bytes < b2_comm_region_bytes(tp, max_rows, H, gather_rows, gather_cols)) {
        b2_set_error("comm_create: regions hold %zu bytes, %zu required", region_bytes,
                     b2_comm_region_bytes(tp, max_rows, H, gather_rows, gather_cols));
        return nullptr;
    }
    for (int r = 0; r < tp; ++r)
        if (!regions[r] || !b2_aligned16(regions[r])) {
            b2_set_error("comm_create: region %d is null or misaligned", r);
            return nullptr;
        }
    Comm* c = new (std::nothrow) Comm();
    if (!c) {
        b2_set_error("comm_create: out of host memory");
        return nullptr;
    }
    c->rank = rank; c->tp = tp; c->max_rows = max_rows; c->H = H;
    c->gather_rows = gather_rows; c->gather_cols = gather_cols; c->region_bytes = region_bytes;
    const char* env = getenv("B2_COMM_DEBUG");
    c->debug = env && env[0] == '1';
    env = getenv("B2_COMM_WEAK_ST");
    c->weak_st = env && env[0] == '1';
    const size_t in = align_up((size_t)max_rows * H * sizeof(__nv_bfloat16), 256);
    c->off_in[0] = CTRL_BYTES;
    c->off_in[1] = CTRL_BYTES + in;
    c->off_gather = CTRL_BYTES + 2 * in;
    c->off_lam = c->off_gather + align_up((size_t)gather_rows * gather_cols * sizeof(__nv_bfloat16), 256);
    c->lam_src_bytes = in;
    c->lam_buf_bytes = (size_t)tp * in;
    for (int r = 0; r < MAX_TP; ++r) c->peers.base[r] = r < tp ? static_cast<char*>(regions[r]) : nullptr;
    // arm this rank's landing buffers: every slot holds the sentinel until a peer's data lands in it
    fill_u32_kernel<<<592, 256>>>(reinterpret_cast<uint32_t*>(c->peers.base[rank] + c->off_lam),
                                  3 * c->lam_buf_bytes / sizeof(uint32_t), 0x80008000u);
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    if (e != cudaSuccess) {
        b2_set_error("comm_create: arming the landing buffers failed: %s", cudaGetErrorString(e));
        delete c;
        return nullptr;
    }
    return c;
}

extern "C" void b2_comm_destroy(void* comm) { delete static_cast<Comm*>(comm); }

extern "C" void* b2_comm_input(void* comm, int slot) {
    const Comm* c = static_cast<const Comm*>(comm);
    if (!c || slot < 0 || slot > 1) return nullptr;
    return c->peers.base[c->rank] + c->off_in[slot];
}

extern "C" int b2_comm_debug_stamps(void* comm, uint64_t* out4) {
    const Comm* c = static_cast<const Comm*>(comm);
    B2_REQUIRE(c && out4, B2_EINVAL, "comm_debug_stamps: null pointer");
    B2_CUDA_OK(cudaMemcpy(out4, c->peers.base[c->rank] + offsetof(Ctrl, dbg), 4 * sizeof(uint64_t), cudaMemcpyDeviceToHost));
    return B2_OK;
}

extern "C" void* b2_comm_gathered(void* comm) {
    const Comm* c = static_cast<const Comm*>(comm);
    if (!c) return nullptr;
    return c->peers.base[c->rank] + c->off_gather;
}

extern "C" int b2_comm_allreduce_rms_norm(void* comm, int slot, void* out, void* residual, const void* weight, int T,
                                          int64_t out_stride, int64_t res_stride, float eps, void* stream) {
    const Comm* c = static_cast<const Comm*>(comm);
    B2_REQUIRE(c && (slot == 0 || slot == 1), B2_EINVAL, "comm_allreduce_rms_norm: bad handle / slot");
    B2_REQUIRE(T >= 0 && T <= c->max_rows, B2_EINVAL, "comm_allreduce_rms_norm: T=%d outside [0, %d]", T, c->max_rows);
    if (T == 0) return B2_OK;
    B2_REQUIRE(out, B2_EINVAL, "comm_allreduce_rms_norm: null out");
    B2_REQUIRE(b2_aligned16(out) && b2_aligned16(residual) && b2_aligned16(weight) && out_stride % 8 == 0 &&
                   res_stride % 8 == 0, B2_EINVAL,
               "comm_allreduce_rms_norm: pointers / strides must keep 16-byte alignment");
    auto* o = static_cast<__nv_bfloat16*>(out);
    auto* r = static_cast<__nv_bfloat16*>(residual);
    auto* w = static_cast<const __nv_bfloat16*>(weight);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    switch (c->tp) {
        case 1: launch_allreduce<1>(c, slot, o, r, w, T, out_stride, res_stride, eps, st); break;
        case 2: launch_allreduce<2>(c, slot, o, r, w, T, out_stride, res_stride, eps, st); break;
        case 4: launch_allreduce<4>(c, slot, o, r, w, T, out_stride, res_stride, eps, st); break;
        default: launch_allreduce<8>(c, slot, o, r, w, T, out_stride, res_stride, eps, st); break;
    }
    B2_CUDA_OK(cudaGetLastError());
    return B2_OK;
}

extern "C" int b2_comm_push_allreduce_rms_norm(void* comm, const void* partial, int64_t partial_stride, void* out,
                                               void* residual, const void* weight, int T, int64_t out_stride,
                                               int64_t res_stride, float eps, void* stream) {
    const Comm* c = static_cast<const Comm*>(comm);
    B2_REQUIRE(c, B2_EINVAL, "comm_push_allreduce_rms_norm: null handle");
    B2_REQUIRE(T >= 0 && T <= c->max_rows && T <= PUSH_MAX_ROWS, B2_EINVAL,
               "comm_push_allreduce_rms_norm: T=%d outside [0, min(%d, %d)]", T, c->max_rows, PUSH_MAX_ROWS);
    if (T == 0) return B2_OK;
    B2_REQUIRE(out && partial, B2_EINVAL, "comm_push_allreduce_rms_norm: null pointer");
    B2_REQUIRE(b2_aligned16(out) && b2_aligned16(partial) && b2_aligned16(residual) && b2_aligned16(weight) &&
                   out_stride % 8 == 0 && res_stride % 8 == 0 && partial_stride % 8 == 0, B2_EINVAL,
               "comm_push_allreduce_rms_norm: pointers / strides must keep 16-byte alignment");
    auto* x = static_cast<const __nv_bfloat16*>(partial);
    auto* o = static_cast<__nv_bfloat16*>(out);
    auto* r = static_cast<__nv_bfloat16*>(residual);
    auto* w = static_cast<const __nv_bfloat16*>(weight);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    switch (c->tp) {
        case 1: launch_push<1>(c, x, partial_stride, o, r, w, T, out_stride, res_stride, eps, st); break;
        case 2: launch_push<2>(c, x, partial_stride, o, r, w, T, out_stride, res_stride, eps, st); break;
        case 4: launch_push<4>(c, x, partial_stride, o, r, w, T, out_stride, res_stride, eps, st); break;
        default: launch_push<8>(c, x, partial_stride, o, r, w, T, out_stride, res_stride, eps, st); break;
    }
    B2_CUDA_OK(cudaGetLastError());
    return B2_OK;
}

extern "C" int b2_comm_allgather_cols(void* comm, const void* local, int64_t local_stride, int T, int cols,
                                      void* stream) {
    const Comm* c = static_cast<const Comm*>(comm);
    B2_REQUIRE(c, B2_EINVAL, "comm_allgather_cols: null handle");
    B2_REQUIRE(T >= 0 && T <= c->gather_rows && cols > 0 && cols % 8 == 0 && cols * c->tp <= c->gather_cols, B2_EINVAL,
               "comm_allgather_cols: T=%d cols=%d x tp=%d outside the %d x %d gather buffer", T, cols, c->tp,
               c->gather_rows, c->gather_cols);
    if (T == 0) return B2_OK;
    B2_REQUIRE(local && b2_aligned16(local) && local_stride % 8 == 0, B2_EINVAL,
               "comm_allgather_cols: local rows must be 16-byte aligned");
    auto* l = static_cast<const __nv_bfloat16*>(local);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    switch (c->tp) {
        case 1: launch_allgather<1>(c, l, local_stride, T, cols, st); break;
        case 2: launch_allgather<2>(c, l, local_stride, T, cols, st); break;
        case 4: launch_allgather<4>(c, l, local_stride, T, cols, st); break;
        default: launch_allgather<8>(c, l, local_stride, T, cols, st); break;
    }
    B2_CUDA_OK(cudaGetLastError());
    return B2_OK;
}
