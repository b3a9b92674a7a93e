// gemm_tcgen05.cuh -- the wide-layer contraction of the MLP hot path on Blackwell tensor cores.
//
//   D[M,N] = A[M,K] * B[N,K]^T      A, B: BF16, K contiguous ("TN");  accumulate FP32 in TMEM
//
// One persistent, warp-specialised kernel (1 CTA per SM, 192 threads):
//   warp 0      TMA producer   : cp.async.bulk.tensor 128B-swizzled A (128x64) and B (256x64) tiles
//   warp 1      MMA issuer     : tcgen05.mma.cta_group::1.kind::f16, M=128 N=256 K=16, owns TMEM alloc
//   warps 2..5  epilogue       : tcgen05.ld 32x32b -> registers -> fused epilogue -> global
// Pipelines: smem full/empty ring (STAGES deep) and a double-buffered TMEM accumulator
// (2 x 256 columns) so the epilogue of tile i overlaps the main loop of tile i+1.
//
// The three contractions of a training step all take this one form (replaces the MAC loops
// at src/SoftmaxCrossEntropyNeuralNet.java:187-192, :272-278 and :279-283 + :309-317):
//   forward   Z   = A_l   * Wt_l^T        (M=batch, N=d_{l+1}, K=d_l)
//   delta     D_l = D_l+1 * Wn_l^T        (M=batch, N=d_l,     K=d_{l+1})
//   gradient  G_l = At_l  * Dt_l+1^T      (M=d_l,   N=d_{l+1}, K=batch)
// and differ only in the epilogue (EpiKind).
#pragma once
#include <cuda_bf16.h>

#include "sm100_ptx.cuh"

namespace b2nn {

constexpr int GEMM_BM = 128;
constexpr int GEMM_BN = 256;
constexpr int GEMM_BK = 64;  // 64 bf16 = 128 B = one swizzle row
constexpr int GEMM_STAGES = 4;
constexpr int GEMM_THREADS = 192;
constexpr int GEMM_A_BYTES = GEMM_BM * GEMM_BK * 2;  // 16 KB
constexpr int GEMM_B_BYTES = GEMM_BN * GEMM_BK * 2;  // 32 KB
constexpr int GEMM_STAGE_BYTES = GEMM_A_BYTES + GEMM_B_BYTES;
constexpr int GEMM_SMEM_BYTES = GEMM_STAGES * GEMM_STAGE_BYTES + 1024 /*align*/ + 256 /*barriers*/;
constexpr uint32_t GEMM_TMEM_COLS = 512;  // 2 accumulators x 256 fp32 columns

enum EpiKind {
    EPI_ACT = 0,        // a = f(acc): store BF16 row-major out0[M][ld0] and transposed out1[N][ld1]
    EPI_LOGITS = 1,     // store FP32 acc to out0[M][ld0]
    EPI_DX = 2,         // d = acc * f'(from stored activation aux[M][ldaux]); store like EPI_ACT
    EPI_DW_UPDATE = 3,  // v = s*acc + mu*v; w -= v (FP32 master w,v [M][ldw]); BF16 copies out0 / out1
    EPI_DW_STORE = 4    // store FP32 acc (the summed gradient) to w[M][ldw]
};

enum ActKind { ACT_LEAKY_RELU = 0, ACT_SIGMOID = 1, ACT_RELU = 2, ACT_TANH = 3, ACT_IDENTITY = 4 };

struct GemmArgs {
    int M, N, K;
    void *out0;   // row-major output (bf16, or fp32 for EPI_LOGITS)
    int ld0;
    __nv_bfloat16 *out1;  // transposed bf16 output [N][ld1] (may be null)
    int ld1;
    const __nv_bfloat16 *aux;  // stored activation for EPI_DX
    int ldaux;
    float *w;  // master weights / gradient out
    float *v;  // velocity
    int ldw;
    float scale;     // step / batch
    float momentum;  // mu
    int act;
    float act_param;
};

__device__ __forceinline__ float act_apply(int act, float p, float z) {
    switch (act) {
    case ACT_LEAKY_RELU: return z > 0.f ? z : p * z;
    case ACT_SIGMOID: return 1.f / (1.f + __expf(-z));
    case ACT_RELU: return z > 0.f ? z : 0.f;
    case ACT_TANH: return tanhf(z);
    default: return z;
    }
}
// derivative evaluated from the STORED activation a = f(z) (exact for these functions)
__device__ __forceinline__ float act_prime_from_a(int act, float p, float a) {
    switch (act) {
    case ACT_LEAKY_RELU: return a > 0.f ? 1.f : p;  // z<=0 <=> a<=0 for p>0 (MNISTTrainer.java:235)
    case ACT_SIGMOID: return a * (1.f - a);
    case ACT_RELU: return a > 0.f ? 1.f : 0.f;
    case ACT_TANH: return 1.f - a * a;
    default: return 1.f;
    }
}

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
    __nv_bfloat162 t = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t *>(&t);
}

// Store 32 consecutive values of row m as bf16 (row-major) and, optionally, transposed.
__device__ __forceinline__ void store_bf16_row_and_t(const GemmArgs &g, int m, int n0, const float (&x)[32]) {
    __nv_bfloat16 *o0 = reinterpret_cast<__nv_bfloat16 *>(g.out0);
    if (o0 != nullptr) {
        __nv_bfloat16 *row = o0 + (size_t)m * g.ld0 + n0;
        if (n0 + 32 <= g.N) {
#pragma unroll
            for (int q = 0; q < 4; q++) {
                uint4 u;
                u.x = pack_bf16x2(x[8 * q + 0], x[8 * q + 1]);
                u.y = pack_bf16x2(x[8 * q + 2], x[8 * q + 3]);
                u.z = pack_bf16x2(x[8 * q + 4], x[8 * q + 5]);
                u.w = pack_bf16x2(x[8 * q + 6], x[8 * q + 7]);
                reinterpret_cast<uint4 *>(row)[q] = u;
            }
        } else {
#pragma unroll
            for (int c = 0; c < 32; c++)
                if (n0 + c < g.N) row[c] = __float2bfloat16_rn(x[c]);
        }
    }
    if (g.out1 != nullptr) {
#pragma unroll
        for (int c = 0; c < 32; c++)
            if (n0 + c < g.N) g.out1[(size_t)(n0 + c) * g.ld1 + m] = __float2bfloat16_rn(x[c]);
    }
}

template <int EPI>
__device__ __forceinline__ void epilogue_chunk(const GemmArgs &g, int m, int n0, const uint32_t (&r)[32]) {
    if (m >= g.M || n0 >= g.N) return;
    if constexpr (EPI == EPI_ACT) {
        float x[32];
#pragma unroll
        for (int c = 0; c < 32; c++) x[c] = act_apply(g.act, g.act_param, __uint_as_float(r[c]));
        store_bf16_row_and_t(g, m, n0, x);
    } else if constexpr (EPI == EPI_LOGITS) {
        float *row = reinterpret_cast<float *>(g.out0) + (size_t)m * g.ld0 + n0;
        if (n0 + 32 <= g.N) {
#pragma unroll
            for (int q = 0; q < 8; q++)
                reinterpret_cast<float4 *>(row)[q] =
                    make_float4(__uint_as_float(r[4 * q]), __uint_as_float(r[4 * q + 1]),
                                __uint_as_float(r[4 * q + 2]), __uint_as_float(r[4 * q + 3]));
        } else {
#pragma unroll
            for (int c = 0; c < 32; c++)
                if (n0 + c < g.N) row[c] = __uint_as_float(r[c]);
        }
    } else if constexpr (EPI == EPI_DX) {
        const __nv_bfloat16 *arow = g.aux + (size_t)m * g.ldaux + n0;
        float x[32];
        if (n0 + 32 <= g.N) {
#pragma unroll
            for (int q = 0; q < 4; q++) {
                uint4 u = reinterpret_cast<const uint4 *>(arow)[q];
                const __nv_bfloat162 *h = reinterpret_cast<const __nv_bfloat162 *>(&u);
#pragma unroll
                for (int e = 0; e < 4; e++) {
                    float2 a = __bfloat1622float2(h[e]);
                    x[8 * q + 2 * e] = __uint_as_float(r[8 * q + 2 * e]) * act_prime_from_a(g.act, g.act_param, a.x);
                    x[8 * q + 2 * e + 1] =
                        __uint_as_float(r[8 * q + 2 * e + 1]) * act_prime_from_a(g.act, g.act_param, a.y);
                }
            }
        } else {
#pragma unroll
            for (int c = 0; c < 32; c++) {
                float a = (n0 + c < g.N) ? __bfloat162float(arow[c]) : 0.f;
                x[c] = __uint_as_float(r[c]) * act_prime_from_a(g.act, g.act_param, a);
            }
        }
        store_bf16_row_and_t(g, m, n0, x);
    } else if constexpr (EPI == EPI_DW_UPDATE) {
        // momentum SGD fused into the gradient GEMM (SoftmaxCrossEntropyNeuralNet.java:325-342):
        // adj = step*G/B + mu*prev ; W -= adj ; prev = adj.  G never reaches HBM.
        float *wrow = g.w + (size_t)m * g.ldw + n0;
        float *vrow = g.v + (size_t)m * g.ldw + n0;
        float x[32];
        if (n0 + 32 <= g.N) {
#pragma unroll
            for (int q = 0; q < 8; q++) {
                float4 w4 = reinterpret_cast<const float4 *>(wrow)[q];
                float4 v4 = reinterpret_cast<const float4 *>(vrow)[q];
                v4.x = g.scale * __uint_as_float(r[4 * q + 0]) + g.momentum * v4.x;
                v4.y = g.scale * __uint_as_float(r[4 * q + 1]) + g.momentum * v4.y;
                v4.z = g.scale * __uint_as_float(r[4 * q + 2]) + g.momentum * v4.z;
                v4.w = g.scale * __uint_as_float(r[4 * q + 3]) + g.momentum * v4.w;
                w4.x -= v4.x; w4.y -= v4.y; w4.z -= v4.z; w4.w -= v4.w;
                reinterpret_cast<float4 *>(wrow)[q] = w4;
                reinterpret_cast<float4 *>(vrow)[q] = v4;
                x[4 * q + 0] = w4.x; x[4 * q + 1] = w4.y; x[4 * q + 2] = w4.z; x[4 * q + 3] = w4.w;
            }
        } else {
#pragma unroll
            for (int c = 0; c < 32; c++) {
                x[c] = 0.f;
                if (n0 + c < g.N) {
                    float vn = g.scale * __uint_as_float(r[c]) + g.momentum * vrow[c];
                    float wn = wrow[c] - vn;
                    vrow[c] = vn; wrow[c] = wn; x[c] = wn;
                }
            }
        }
        store_bf16_row_and_t(g, m, n0, x);
    } else {  // EPI_DW_STORE
        float *grow = g.w + (size_t)m * g.ldw + n0;
        if (n0 + 32 <= g.N) {
#pragma unroll
            for (int q = 0; q < 8; q++)
                reinterpret_cast<float4 *>(grow)[q] =
                    make_float4(__uint_as_float(r[4 * q]), __uint_as_float(r[4 * q + 1]),
                                __uint_as_float(r[4 * q + 2]), __uint_as_float(r[4 * q + 3]));
        } else {
#pragma unroll
            for (int c = 0; c < 32; c++)
                if (n0 + c < g.N) grow[c] = __uint_as_float(r[c]);
        }
    }
}

template <int EPI>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_tn_bf16_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                    const GemmArgs g) {
    extern __shared__ uint8_t smem_raw[];
    // 128B swizzle needs 1024 B alignment of every tile
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t *smem_a = smem;
    uint8_t *smem_b = smem + GEMM_STAGES * GEMM_A_BYTES;
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem + GEMM_STAGES * GEMM_STAGE_BYTES);
    uint64_t *full_bar = bars;                       // [STAGES]  TMA -> MMA
    uint64_t *empty_bar = bars + GEMM_STAGES;        // [STAGES]  MMA -> TMA
    uint64_t *tfull_bar = bars + 2 * GEMM_STAGES;    // [2]       MMA -> epilogue
    uint64_t *tempty_bar = bars + 2 * GEMM_STAGES + 2;  // [2]    epilogue -> MMA
    uint32_t *tmem_base_slot = reinterpret_cast<uint32_t *>(bars + 2 * GEMM_STAGES + 4);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    const int m_tiles = (g.M + GEMM_BM - 1) / GEMM_BM;
    const int n_tiles = (g.N + GEMM_BN - 1) / GEMM_BN;
    const int num_tiles = m_tiles * n_tiles;
    const int k_blocks = (g.K + GEMM_BK - 1) / GEMM_BK;

    if (warp == 0 && lane == 0) {
        ptx::prefetch_tmap(&tmA);
        ptx::prefetch_tmap(&tmB);
        for (int s = 0; s < GEMM_STAGES; s++) {
            ptx::mbar_init(&full_bar[s], 1);
            ptx::mbar_init(&empty_bar[s], 1);
        }
        for (int a = 0; a < 2; a++) {
            ptx::mbar_init(&tfull_bar[a], 1);
            ptx::mbar_init(&tempty_bar[a], 4);  // one arrive per epilogue warp
        }
        ptx::fence_barrier_init();
    }
    if (warp == 1) {
        ptx::tmem_alloc(tmem_base_slot, GEMM_TMEM_COLS);
        ptx::tmem_relinquish();
    }
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_base_slot;

    if (warp == 0) {
        // ===================== TMA producer =====================
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
                const int m0 = (t / n_tiles) * GEMM_BM;
                const int n0 = (t % n_tiles) * GEMM_BN;
                for (int kb = 0; kb < k_blocks; kb++) {
                    ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
                    ptx::mbar_arrive_expect_tx(&full_bar[stage], GEMM_STAGE_BYTES);
                    ptx::tma_load_2d(smem_a + stage * GEMM_A_BYTES, &tmA, &full_bar[stage], kb * GEMM_BK, m0);
                    ptx::tma_load_2d(smem_b + stage * GEMM_B_BYTES, &tmB, &full_bar[stage], kb * GEMM_BK, n0);
                    if (++stage == GEMM_STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        if (lane == 0) {
            constexpr uint32_t idesc = ptx::make_idesc_bf16(GEMM_BM, GEMM_BN);
            int stage = 0;
            uint32_t phase = 0;
            int acc = 0;
            uint32_t acc_phase = 0;
            for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
                ptx::mbar_wait(&tempty_bar[acc], acc_phase ^ 1);  // epilogue has drained this accumulator
                ptx::tc_fence_after();
                const uint32_t d_tmem = tmem_base + (uint32_t)acc * GEMM_BN;
                for (int kb = 0; kb < k_blocks; kb++) {
                    ptx::mbar_wait(&full_bar[stage], phase);
                    ptx::tc_fence_after();
                    const uint64_t da = ptx::make_desc_kmajor_sw128(ptx::smem_u32(smem_a + stage * GEMM_A_BYTES));
                    const uint64_t db = ptx::make_desc_kmajor_sw128(ptx::smem_u32(smem_b + stage * GEMM_B_BYTES));
#pragma unroll
                    for (int k = 0; k < GEMM_BK / 16; k++) {
                        // advance 16 bf16 = 32 B along K inside the swizzle row: +2 in 16 B units
                        ptx::mma_f16_ss(d_tmem, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc,
                                        (kb | k) != 0 ? 1u : 0u);
                    }
                    ptx::mma_commit(&empty_bar[stage]);  // frees the smem slot when these MMAs retire
                    if (++stage == GEMM_STAGES) { stage = 0; phase ^= 1; }
                }
                ptx::mma_commit(&tfull_bar[acc]);  // accumulator complete -> epilogue
                if (++acc == 2) { acc = 0; acc_phase ^= 1; }
            }
        }
    } else {
        // ===================== epilogue warps (2..5) =====================
        const int q = warp & 3;  // TMEM lane quarter this warp may access
        int acc = 0;
        uint32_t acc_phase = 0;
        for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
            const int m0 = (t / n_tiles) * GEMM_BM;
            const int n0 = (t % n_tiles) * GEMM_BN;
            ptx::mbar_wait(&tfull_bar[acc], acc_phase);
            ptx::tc_fence_after();
            const int m = m0 + q * 32 + lane;
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)acc * GEMM_BN;
#pragma unroll 1
            for (int c0 = 0; c0 < GEMM_BN; c0 += 32) {
                uint32_t r[32];
                ptx::tmem_ld_32x32b_x32(taddr + (uint32_t)c0, r);
                ptx::tmem_ld_wait();
                epilogue_chunk<EPI>(g, m, n0 + c0, r);
            }
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(&tempty_bar[acc]);
            if (++acc == 2) { acc = 0; acc_phase ^= 1; }
        }
    }

    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc(tmem_base, GEMM_TMEM_COLS);
    }
}

// =====================================================================================================
// CTA-pair variant: the same contraction with tcgen05.mma.cta_group::2.  A cluster of two CTAs (the two
// SMs of a TPC) owns one 256 x BN tile: each CTA stages ITS 128 rows of A and ITS BN/2 rows of B, the
// leader's single MMA thread drives both tensor cores (M = 256), and each CTA's TMEM receives its 128
// accumulator rows.  Per-SM shared-memory read traffic per MMA drops from A+B to A+B/2 and the smem
// ring deepens (6-8 stages).  Barrier topology:
//   full[s]    leader only, 1 arrival (its expect_tx of BOTH CTAs' bytes); both CTAs' TMA credit it
//   empty[s]   one per CTA, signalled in both CTAs by the multicast tcgen05.commit
//   tfull[a]   one per CTA, multicast commit -> each CTA's epilogue warps
//   tempty[a]  leader only, 8 arrivals (4 epilogue warps of each CTA, the peer's arrive remotely)
template <int BN>
struct PairCfg {
    static constexpr int A_BYTES = GEMM_BM * GEMM_BK * 2;         // 16 KB: this CTA's 128 rows
    static constexpr int B_BYTES = (BN / 2) * GEMM_BK * 2;        // this CTA's half of the B tile
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int STAGES = (192 * 1024) / STAGE_BYTES;     // 6 (BN=256) or 8 (BN=128)
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 + 256;
    static constexpr uint32_t TMEM_COLS = 2 * BN;                 // double-buffered accumulator
};

template <int EPI, int BN>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(GEMM_THREADS, 1)
gemm_tn_bf16_pair_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                         const GemmArgs g) {
    using Cfg = PairCfg<BN>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t *smem_a = smem;
    uint8_t *smem_b = smem + Cfg::STAGES * Cfg::A_BYTES;
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem + Cfg::STAGES * Cfg::STAGE_BYTES);
    uint64_t *full_bar = bars;
    uint64_t *empty_bar = bars + Cfg::STAGES;
    uint64_t *tfull_bar = bars + 2 * Cfg::STAGES;
    uint64_t *tempty_bar = bars + 2 * Cfg::STAGES + 2;
    uint32_t *tmem_base_slot = reinterpret_cast<uint32_t *>(bars + 2 * Cfg::STAGES + 4);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const uint32_t rank = ptx::cluster_ctarank();
    const int pair = blockIdx.x >> 1, num_pairs = gridDim.x >> 1;

    const int m_tiles = (g.M + 2 * GEMM_BM - 1) / (2 * GEMM_BM);
    const int n_tiles = (g.N + BN - 1) / BN;
    const int num_tiles = m_tiles * n_tiles;
    const int k_blocks = (g.K + GEMM_BK - 1) / GEMM_BK;

    if (warp == 0 && lane == 0) {
        ptx::prefetch_tmap(&tmA);
        ptx::prefetch_tmap(&tmB);
        for (int s = 0; s < Cfg::STAGES; s++) {
            ptx::mbar_init(&full_bar[s], 1);
            ptx::mbar_init(&empty_bar[s], 1);
        }
        for (int a = 0; a < 2; a++) {
            ptx::mbar_init(&tfull_bar[a], 1);
            ptx::mbar_init(&tempty_bar[a], 8);
        }
        ptx::fence_barrier_init();
    }
    ptx::cluster_sync();  // both CTAs' barriers exist before anyone signals across the pair
    if (warp == 1) {
        ptx::tmem_alloc_pair(tmem_base_slot, Cfg::TMEM_COLS);
        ptx::tmem_relinquish_pair();
    }
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_base_slot;

    if (warp == 0) {
        // ===================== TMA producer (both CTAs) =====================
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int t = pair; t < num_tiles; t += num_pairs) {
                const int m0 = (t / n_tiles) * (2 * GEMM_BM) + (int)rank * GEMM_BM;
                const int n0 = (t % n_tiles) * BN + (int)rank * (BN / 2);
                for (int kb = 0; kb < k_blocks; kb++) {
                    ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
                    if (rank == 0) ptx::mbar_arrive_expect_tx(&full_bar[stage], 2 * Cfg::STAGE_BYTES);
                    ptx::tma_load_2d_pair(smem_a + stage * Cfg::A_BYTES, &tmA, &full_bar[stage], kb * GEMM_BK, m0);
                    ptx::tma_load_2d_pair(smem_b + stage * Cfg::B_BYTES, &tmB, &full_bar[stage], kb * GEMM_BK, n0);
                    if (++stage == Cfg::STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer (leader CTA only) =====================
        if (rank == 0 && lane == 0) {
            constexpr uint32_t idesc = ptx::make_idesc_bf16(2 * GEMM_BM, BN);
            int stage = 0;
            uint32_t phase = 0;
            int acc = 0;
            uint32_t acc_phase = 0;
            for (int t = pair; t < num_tiles; t += num_pairs) {
                ptx::mbar_wait(&tempty_bar[acc], acc_phase ^ 1);
                ptx::tc_fence_after();
                const uint32_t d_tmem = tmem_base + (uint32_t)acc * BN;
                for (int kb = 0; kb < k_blocks; kb++) {
                    ptx::mbar_wait(&full_bar[stage], phase);
                    ptx::tc_fence_after();
                    const uint64_t da = ptx::make_desc_kmajor_sw128(ptx::smem_u32(smem_a + stage * Cfg::A_BYTES));
                    const uint64_t db = ptx::make_desc_kmajor_sw128(ptx::smem_u32(smem_b + stage * Cfg::B_BYTES));
#pragma unroll
                    for (int k = 0; k < GEMM_BK / 16; k++)
                        ptx::mma_f16_ss_pair(d_tmem, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc,
                                             (kb | k) != 0 ? 1u : 0u);
                    ptx::mma_commit_pair(&empty_bar[stage]);
                    if (++stage == Cfg::STAGES) { stage = 0; phase ^= 1; }
                }
                ptx::mma_commit_pair(&tfull_bar[acc]);
                if (++acc == 2) { acc = 0; acc_phase ^= 1; }
            }
        }
    } else {
        // ===================== epilogue warps (2..5), each CTA drains its own 128 rows ==========
        const int q = warp & 3;
        int acc = 0;
        uint32_t acc_phase = 0;
        for (int t = pair; t < num_tiles; t += num_pairs) {
            const int m0 = (t / n_tiles) * (2 * GEMM_BM) + (int)rank * GEMM_BM;
            const int n0 = (t % n_tiles) * BN;
            ptx::mbar_wait(&tfull_bar[acc], acc_phase);
            ptx::tc_fence_after();
            const int m = m0 + q * 32 + lane;
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)acc * BN;
#pragma unroll 1
            for (int c0 = 0; c0 < BN; c0 += 32) {
                uint32_t r[32];
                ptx::tmem_ld_32x32b_x32(taddr + (uint32_t)c0, r);
                ptx::tmem_ld_wait();
                epilogue_chunk<EPI>(g, m, n0 + c0, r);
            }
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive_leader(&tempty_bar[acc]);
            if (++acc == 2) { acc = 0; acc_phase ^= 1; }
        }
    }

    ptx::tc_fence_before();
    ptx::cluster_sync();  // the peer may still read this CTA's smem / signal its barriers until here
    if (warp == 1) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc_pair(tmem_base, Cfg::TMEM_COLS);
    }
}

}  // namespace b2nn
