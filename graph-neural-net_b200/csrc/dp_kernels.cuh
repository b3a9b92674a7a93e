// dp_kernels.cuh -- data-parallel gradient exchange fused with the optimizer step, over NVLink peer memory.
//
// The serial reference sums per-sample gradients over the batch and applies one momentum update
// (src/SoftmaxCrossEntropyNeuralNet.java:305-342).  With the batch split over G GPUs the sum becomes a sum
// over GPUs.  Instead of all-reduce (2(G-1)/G x 4 B per parameter on the wire) + a separate update kernel,
// ONE kernel per layer does reduce-scatter + update + all-gather with plain loads / stores on the peers'
// memory (every GPU of an NVSwitch box reaches every peer at full NVLink bandwidth):
//   rank r owns the r-th slice of the layer's parameters (optimizer state is sharded: master w and v of a
//   slice live only on its owner);
//   1. handshake: tell every peer "my gradient of layer l is complete", wait for theirs;
//   2. for each owned float4: g = sum over ranks 0..G-1 (fixed order) of the peers' gradient, read over
//      NVLink; v = scale*g + mu*v; w -= v; the BF16 operand copy of w is stored into EVERY rank's wn;
//   3. handshake: "my slice is written everywhere"; the kernel ends when all peers said so, i.e. when this
//      rank's wn is complete and its gradient buffer is no longer being read.
// Wire traffic per parameter: (G-1)/G x (4 B in + 2 B out); the gradient never makes a second trip through HBM.
// TF32 nets have no BF16 operand copy -- their GEMMs read the FP32 master array itself -- so for them the owner
// replicates the updated FP32 weights (4 B out instead of 2); the velocity stays sharded either way.
// Every cross-GPU wait is bounded (Args::timeout_ns, %globaltimer): a peer that never arrives sets Flags::error,
// the kernel skips its work and ends, and the host turns the word into B2NN_ERR_CUDA at the next synchronising call.
#pragma once
#include <cuda_bf16.h>

#include <cstdint>

#include "common.hpp"
#include "gemm_tcgen05.cuh"

namespace b2nn {
namespace dp {

constexpr int MAX_LAYERS = DP_MAX_LAYERS;
// Small CTAs on purpose: the kernel runs BESIDE the persistent GEMM (320 threads x up to 128 registers = 41 K of
// the SM's 64 K registers); 128 threads x <= 96 registers fit next to it, a 256 x 172 CTA did not and the two
// kernels time-shared the SMs (measured at 8 GPUs: the next forward GEMM took 2.6x longer).
constexpr int DP_THREADS = 128;
// layout of the flag block at the head of every arena (DP_FLAG_BYTES)
struct Flags {
    unsigned int ready[MAX_LAYERS][DP_MAX_WORLD];  // [l][p]: rank p's gradient of layer l is complete (sequence number)
    unsigned int done[MAX_LAYERS][DP_MAX_WORLD];   // [l][p]: rank p has written its slice of layer l everywhere
    unsigned int go[MAX_LAYERS];                   // local: CTA 0 saw every peer's ready flag
    unsigned int finished[MAX_LAYERS];             // local: CTAs of this launch that finished their part
    unsigned int error;                            // local, sticky: 0, or DP_ERR_* | layer << 8 | peer of the wait that timed out
};
constexpr unsigned int DP_ERR_READY = 0x01000000u;  // a peer's gradient never became ready
constexpr unsigned int DP_ERR_DONE = 0x02000000u;   // a peer never reported its slice written
constexpr unsigned int DP_ERR_GO = 0x03000000u;     // CTA 0 of this launch never released the others
static_assert(sizeof(Flags) <= DP_FLAG_BYTES, "flag block does not fit");

struct Args {
    char *arena[DP_MAX_WORLD];  // every rank's arena (own entry included), same layout everywhere
    char *mc;                   // multicast mapping of all arenas (NVSwitch), or null: plain peer loads / stores
    int world, rank, layer;
    unsigned long long off_g, off_w, off_v, off_wn;  // byte offsets of this layer's arrays inside an arena
    long long n4;               // float4 elements of the layer (padded [in][ld_n] array)
    float scale, momentum;
    unsigned int seq;           // handshake sequence number, identical on all ranks for this launch
    int f32_operand;            // TF32 nets: replicate the FP32 weights themselves (no BF16 copy exists)
    int g_bf16;                 // B2NN_FLAG_BF16_WIRE: the per-GPU gradients are BF16 arrays at off_gb (2 B per parameter and peer)
    unsigned long long off_gb;
    unsigned long long timeout_ns;  // bound of every cross-GPU wait
};

__device__ __forceinline__ void st_release_sys(unsigned int *p, unsigned int v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned int ld_acquire_sys(const unsigned int *p) {
    unsigned int v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned int ld_acquire_gpu(const unsigned int *p) {
    unsigned int v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// 16 bytes summed over every GPU of the multicast team inside the switch
__device__ __forceinline__ float4 multimem_ld_reduce_add(const void *mc_addr) {
    float4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                 : "l"(mc_addr)
                 : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
// spin until *p == want (acquire at system scope); false when `timeout_ns` passed first
__device__ __forceinline__ bool wait_flag_sys(const unsigned int *p, unsigned int want, unsigned long long timeout_ns) {
    if (ld_acquire_sys(p) == want) return true;
    const unsigned long long t0 = globaltimer_ns();
    while (ld_acquire_sys(p) != want)
        if (globaltimer_ns() - t0 > timeout_ns) return false;
    return true;
}
// 8 bytes = four BF16 gradients summed over every GPU of the team inside the switch (FP32 accumulation, BF16 result)
__device__ __forceinline__ uint2 multimem_ld_reduce_add_bf16x4(const void *mc_addr) {
    uint2 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v2.bf16x2 {%0, %1}, [%2];"
                 : "=r"(v.x), "=r"(v.y)
                 : "l"(mc_addr)
                 : "memory");
    return v;
}
__device__ __forceinline__ float4 bf16x4_to_float4(uint2 u) {
    const float2 lo = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162 *>(&u.x));
    const float2 hi = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162 *>(&u.y));
    return make_float4(lo.x, lo.y, hi.x, hi.y);
}
// 16 bytes (four FP32) replicated into every GPU of the team by the switch
__device__ __forceinline__ void multimem_st_f32x4(void *mc_addr, float4 v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc_addr), "f"(v.x), "f"(v.y),
                 "f"(v.z), "f"(v.w)
                 : "memory");
}
// 8 bytes (four BF16) replicated into every GPU of the team by the switch
__device__ __forceinline__ void multimem_st_bf16x4(void *mc_addr, uint2 b) {
    asm volatile("multimem.st.relaxed.sys.global.v2.bf16x2 [%0], {%1, %2};" ::"l"(mc_addr), "r"(b.x), "r"(b.y) : "memory");
}

__global__ void __launch_bounds__(DP_THREADS, 4)
reduce_update_kernel(const Args a) {
    Flags *mine = reinterpret_cast<Flags *>(a.arena[a.rank]);
    const int l = a.layer;
    // ---- 1. every rank's gradient is complete ----
    if (blockIdx.x == 0) {
        if ((int)threadIdx.x < a.world) {
            Flags *peer = reinterpret_cast<Flags *>(a.arena[threadIdx.x]);
            __threadfence_system();
            st_release_sys(&peer->ready[l][a.rank], a.seq);
            if (!wait_flag_sys(&mine->ready[l][threadIdx.x], a.seq, a.timeout_ns))
                atomicCAS(&mine->error, 0u, DP_ERR_READY | ((unsigned)l << 8) | threadIdx.x);
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            __threadfence();
            asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(&mine->go[l]), "r"(a.seq) : "memory");
        }
    } else {
        if (threadIdx.x == 0 && ld_acquire_gpu(&mine->go[l]) != a.seq) {
            const unsigned long long t0 = globaltimer_ns();
            while (ld_acquire_gpu(&mine->go[l]) != a.seq)
                if (globaltimer_ns() - t0 > 2 * a.timeout_ns) {
                    atomicCAS(&mine->error, 0u, DP_ERR_GO | ((unsigned)l << 8));
                    break;
                }
        }
        __syncthreads();
    }
    // a wait above (in this or an earlier launch) timed out: a peer is gone.  Touch nothing, fall through to the end.
    const bool aborted = ld_acquire_gpu(&mine->error) != 0u;

    // ---- 2. owned slice: reduce over the peers, update, broadcast the BF16 copy ----
    const long long lo = (long long)a.rank * a.n4 / a.world, hi = (long long)(a.rank + 1) * a.n4 / a.world;
    float4 *w = reinterpret_cast<float4 *>(a.arena[a.rank] + a.off_w);
    float4 *v = reinterpret_cast<float4 *>(a.arena[a.rank] + a.off_v);
    const long long stride = (long long)gridDim.x * blockDim.x;
    // NVLink reads are latency-bound per request: keep UNROLL x world 16-byte loads in flight per thread
    constexpr int UNROLL = 4;
    for (long long q0 = lo + (long long)blockIdx.x * blockDim.x + threadIdx.x; q0 < hi && !aborted; q0 += stride * UNROLL) {
        float4 g[UNROLL], wv[UNROLL], vv[UNROLL];
#pragma unroll
        for (int u = 0; u < UNROLL; u++) g[u] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (a.g_bf16) {
            // BF16 wire: 8 bytes per owned float4 and rank (summed in FP32 here, or in the switch with a BF16 result)
            if (a.mc != nullptr) {
                const uint2 *gm = reinterpret_cast<const uint2 *>(a.mc + a.off_gb);
#pragma unroll
                for (int u = 0; u < UNROLL; u++) {
                    const long long q = q0 + u * stride;
                    if (q < hi) g[u] = bf16x4_to_float4(multimem_ld_reduce_add_bf16x4(gm + q));
                }
            } else {
#pragma unroll
                for (int p = 0; p < DP_MAX_WORLD; p++) {
                    if (p < a.world) {
                        const uint2 *gp = reinterpret_cast<const uint2 *>(a.arena[p] + a.off_gb);
                        uint2 t[UNROLL];
#pragma unroll
                        for (int u = 0; u < UNROLL; u++) {
                            const long long q = q0 + u * stride;
                            t[u] = q < hi ? __ldcg(gp + q) : make_uint2(0u, 0u);
                        }
#pragma unroll
                        for (int u = 0; u < UNROLL; u++) {
                            const float4 f = bf16x4_to_float4(t[u]);
                            g[u].x += f.x; g[u].y += f.y; g[u].z += f.z; g[u].w += f.w;
                        }
                    }
                }
            }
        } else
        if (a.mc != nullptr) {
            // in-switch reduction: ONE load per owned float4, the NVSwitch gathers and adds the G copies
            const float4 *gm = reinterpret_cast<const float4 *>(a.mc + a.off_g);
#pragma unroll
            for (int u = 0; u < UNROLL; u++) {
                const long long q = q0 + u * stride;
                if (q < hi) g[u] = multimem_ld_reduce_add(gm + q);
            }
        } else
#pragma unroll
        for (int p = 0; p < DP_MAX_WORLD; p++) {
            if (p < a.world) {
                const float4 *gp = reinterpret_cast<const float4 *>(a.arena[p] + a.off_g);
                float4 t[UNROLL];
#pragma unroll
                for (int u = 0; u < UNROLL; u++) {
                    const long long q = q0 + u * stride;
                    t[u] = q < hi ? __ldcg(gp + q) : make_float4(0.f, 0.f, 0.f, 0.f);
                }
#pragma unroll
                for (int u = 0; u < UNROLL; u++) { g[u].x += t[u].x; g[u].y += t[u].y; g[u].z += t[u].z; g[u].w += t[u].w; }
            }
        }
#pragma unroll
        for (int u = 0; u < UNROLL; u++) {
            const long long q = q0 + u * stride;
            if (q < hi) { wv[u] = w[q]; vv[u] = v[q]; }
        }
#pragma unroll
        for (int u = 0; u < UNROLL; u++) {
            const long long q = q0 + u * stride;
            if (q >= hi) continue;
            float4 &W = wv[u], &V = vv[u];
            V.x = a.scale * g[u].x + a.momentum * V.x; V.y = a.scale * g[u].y + a.momentum * V.y;
            V.z = a.scale * g[u].z + a.momentum * V.z; V.w = a.scale * g[u].w + a.momentum * V.w;
            W.x -= V.x; W.y -= V.y; W.z -= V.z; W.w -= V.w;
            v[q] = V;
            if (a.f32_operand) {
                // TF32 net: every rank's GEMMs read the FP32 array itself -- replicate it (own copy included)
                if (a.mc != nullptr) {
                    multimem_st_f32x4(reinterpret_cast<float4 *>(a.mc + a.off_w) + q, W);
                } else {
#pragma unroll
                    for (int p = 0; p < DP_MAX_WORLD; p++)
                        if (p < a.world) __stcg(reinterpret_cast<float4 *>(a.arena[p] + a.off_w) + q, W);
                }
                continue;
            }
            w[q] = W;
            const uint2 b = make_uint2(pack_bf16x2(W.x, W.y), pack_bf16x2(W.z, W.w));
            if (a.mc != nullptr) {
                multimem_st_bf16x4(reinterpret_cast<uint2 *>(a.mc + a.off_wn) + q, b);  // one store, G copies
            } else {
#pragma unroll
                for (int p = 0; p < DP_MAX_WORLD; p++)
                    if (p < a.world) __stcg(reinterpret_cast<uint2 *>(a.arena[p] + a.off_wn) + q, b);
            }
        }
    }

    // ---- 3. my slice is written everywhere; wait until every peer's is ----
    __threadfence_system();
    __syncthreads();
    __shared__ bool last;
    if (threadIdx.x == 0) last = atomicAdd(&mine->finished[l], 1u) == gridDim.x - 1;
    __syncthreads();
    if (last) {
        if (threadIdx.x == 0) mine->finished[l] = 0u;
        if ((int)threadIdx.x < a.world && ld_acquire_gpu(&mine->error) == 0u) {
            Flags *peer = reinterpret_cast<Flags *>(a.arena[threadIdx.x]);
            __threadfence_system();
            st_release_sys(&peer->done[l][a.rank], a.seq);
            if (!wait_flag_sys(&mine->done[l][threadIdx.x], a.seq, a.timeout_ns))
                atomicCAS(&mine->error, 0u, DP_ERR_DONE | ((unsigned)l << 8) | threadIdx.x);
        }
    }
}

}  // namespace dp
}  // namespace b2nn
