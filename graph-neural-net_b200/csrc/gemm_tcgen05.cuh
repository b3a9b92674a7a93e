// gemm_tcgen05.cuh -- the wide-layer contraction of the MLP hot path on Blackwell tensor cores.
//
//   D[M,N] = sum_k A(m,k) * B(n,k)        A, B: BF16 (kind::f16) or FP32 read as TF32 (kind::tf32) in HBM exactly
//                                         as the net stores them (row-major, no transposed copies);
//                                         accumulate FP32 in TMEM
//
// One persistent, warp-specialised kernel run by CTA PAIRS (cluster of 2 = the two SMs of a TPC,
// tcgen05.mma.cta_group::2): a pair owns a 256 x 256 output tile, each CTA stages ITS 128 rows of A and
// ITS 128 rows of B per 64-deep K block, the leader's single MMA thread drives both tensor cores, and each
// CTA's TMEM receives its 128 accumulator rows.
//   warp 0      TMA producer   : cp.async.bulk.tensor, 64x64 128B-swizzled boxes, 5-stage 32 KB ring
//   warp 1      MMA issuer     : leader CTA only; M=256, N=256 (or 128), K=16 per instruction
//   warps 2..9  epilogue       : tcgen05.ld 32x32b -> registers -> fused epilogue -> global (row-major)
// Pipelines: smem full/empty ring and a double-buffered TMEM accumulator (2 x 256 columns) so the epilogue
// of tile i overlaps the main loop of tile i+1.
//
// Operand major-ness.  An operand whose K index is contiguous in memory is "K-major", one whose M/N index
// is contiguous is "MN-major"; UMMA reads both (descriptor + instruction-descriptor bits), so the three
// contractions of a training step read the SAME row-major buffers (replaces the MAC loops at
// src/SoftmaxCrossEntropyNeuralNet.java:187-192, :272-278 and :279-283 + :309-317):
//   forward   Z   = A_l   . W_l        A = a[l] [B][d_l]     K-major,  B = W [d_l][d_l+1]  MN-major
//   delta     D_l = D_l+1 . W_l^T      A = d[l+1] [B][d_l+1] K-major,  B = W [d_l][d_l+1]  K-major
//   gradient  G_l = A_l^T . D_l+1      A = a[l] [B][d_l]     MN-major, B = d[l+1]          MN-major (K = batch)
// Every buffer has ONE tensor map (64 x 64 box); the kernel picks coordinates and descriptors by major-ness.
//
// TF32 operands (gemm_tf32_pair_kernel, B2NN_PREC_TF32).  Everything above is laid out in BYTES, so the same
// body serves 4-byte operands: a 128-byte swizzle row holds 32 elements instead of 64, hence 32 x 32 boxes
// (4 KB), a 32-deep K block per stage (still 16 KB of A and 16 KB of B per CTA), K = 8 per instruction (still
// four instructions per stage) and a 1024-byte instead of 2048-byte step along K in an MN-major tile.  One thing
// is NOT the same: kind::tf32 reads an MN-major operand only from the 128-byte swizzle with 32-byte atoms
// (descriptor layout type SWIZZLE_128B_BASE32B, TMA swizzle CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B), so a TF32 buffer
// has two tensor maps, one per major-ness it is read in (wide_path.cu).  The
// activations and deltas are then FP32 arrays and the B operand of the forward / delta GEMMs is the FP32
// master weight array itself (no operand copy); TMA hands the tensor core the FP32 words as TF32.
//
// Wave quantisation: with T tiles on P pairs the last, partial wave is split into 256 x 128 half tiles when
// that lets it finish in half a tile time (T mod P <= P/2), e.g. 4096^3: 256 tiles on 74 pairs = 3.5 waves
// instead of 4.
#pragma once
#include <cuda_bf16.h>

#include "sm100_ptx.cuh"

namespace b2nn {

constexpr int GEMM_BM = 128;        // rows of A per CTA (pair tile: 256)
constexpr int GEMM_BN = 256;        // pair tile width
constexpr int GEMM_BK = 64;         // 64 bf16 = 128 B = one swizzle row (BF16 operands; GemmCfg<true> for TF32)
#ifndef B2NN_GEMM_STAGES
#define B2NN_GEMM_STAGES 5
#endif
constexpr int GEMM_STAGES = B2NN_GEMM_STAGES;   // 5 x 32 KB operand stages + 32 KB of epilogue staging (fused-update epilogue) =
                                                // the 198 KB footprint of the 6-stage kernel of round 1: the kernels that run
                                                // beside the GEMM CTA (input prologue, exchange) still find shared memory on the SM
#ifndef B2NN_GEMM_EPI_WARPS
#define B2NN_GEMM_EPI_WARPS 8
#endif
constexpr int GEMM_EPI_WARPS = B2NN_GEMM_EPI_WARPS;   // GEMM_EPI_PARTS warps per TMEM lane quarter, each draining 1/PARTS of the columns
constexpr int GEMM_EPI_PARTS = GEMM_EPI_WARPS / 4;
#ifndef B2NN_GEMM_EPI_PIPE
#define B2NN_GEMM_EPI_PIPE 1
#endif
#ifndef B2NN_GEMM_BAND
#define B2NN_GEMM_BAND 8
#endif
// tile order: bands of GEMM_BAND tile rows, column by column inside a band (1 = plain row-major order)
constexpr int GEMM_BAND = B2NN_GEMM_BAND;
constexpr int GEMM_EPI_PIPE = B2NN_GEMM_EPI_PIPE;   // epilogue look-ahead: 0 none, 1 the aux loads of EPI_DX one chunk ahead (the TMEM reads one chunk
                                                    // ahead too cost 64 registers and the host-fed pipeline its co-residency: measured, rejected)
static_assert(GEMM_EPI_WARPS % 4 == 0 && (GEMM_EPI_PARTS == 1 || GEMM_EPI_PARTS == 2 || GEMM_EPI_PARTS == 4),
              "epilogue warps come in groups of four (one per TMEM lane quarter); 128-column half tiles split into <= 4 parts of 32");
constexpr int GEMM_THREADS = 64 + 32 * GEMM_EPI_WARPS;
constexpr int GEMM_CHUNK_BYTES = 64 * GEMM_BK * 2;      // one 64 x 64 TMA box: 8 KB
constexpr int GEMM_A_BYTES = 2 * GEMM_CHUNK_BYTES;      // this CTA's 128 rows of A
constexpr int GEMM_B_BYTES = 2 * GEMM_CHUNK_BYTES;      // this CTA's 128 rows of B (64 for a half tile)
constexpr int GEMM_STAGE_BYTES = GEMM_A_BYTES + GEMM_B_BYTES;
// staging of the fused-update epilogue: 4 KB per epilogue warp (a 32-row x 32-column FP32 accumulator chunk)
constexpr int GEMM_EPI_STAGE_BYTES = 32 * 128;
constexpr int GEMM_SMEM_BYTES = GEMM_STAGES * GEMM_STAGE_BYTES + 1024 /*align*/ + 512 /*barriers, item ring*/ +
                                GEMM_EPI_WARPS * GEMM_EPI_STAGE_BYTES;
static_assert(GEMM_SMEM_BYTES + 1024 <= 232448, "shared memory of one CTA (227 KB incl. the 1 KB the system reserves)");
constexpr uint32_t GEMM_TMEM_COLS = 512;  // 2 accumulators x 256 fp32 columns
constexpr int GEMM_SCHED_SLOTS = 4;       // depth of the work-item ring between the scheduler and the role warps
constexpr int GEMM_SCHED_READERS = 2 /*producers*/ + 1 /*MMA*/ + 2 * GEMM_EPI_WARPS;

// operand-type dependent tile geometry; the byte layout of a stage is the same for both
template <bool TF32>
struct GemmCfg {
    static constexpr int ESIZE = TF32 ? 4 : 2;
    static constexpr int CH = 128 / ESIZE;          // elements per 128-byte swizzle row = edge of a TMA box = BK
    static constexpr int BK = CH;                   // K depth of a stage
    static constexpr int CHUNK_BYTES = CH * 128;    // one CH x CH box: 8 KB (BF16) / 4 KB (TF32)
    static constexpr int CHUNKS_128 = 128 / CH;     // boxes per 128 operand rows: 2 / 4
    static constexpr int UMMA_K = 32 / ESIZE;       // K per tcgen05.mma: 16 / 8 (32 bytes of a K-major row)
    static constexpr int KSTEP_MN = UMMA_K * 128;   // bytes between K steps of an MN-major tile: 2048 / 1024
};
static_assert(GemmCfg<false>::BK == GEMM_BK && GemmCfg<false>::CHUNK_BYTES == GEMM_CHUNK_BYTES, "BF16 geometry");
static_assert(GemmCfg<true>::CHUNKS_128 * GemmCfg<true>::CHUNK_BYTES == GEMM_A_BYTES, "TF32 stage bytes");

enum EpiKind {
    EPI_ACT = 0,        // a = f(acc): store out0[M][ld0] in the operand type (BF16, or FP32 for TF32 nets)
    EPI_LOGITS = 1,     // store FP32 acc to out0[M][ld0]
    EPI_DX = 2,         // d = acc * f'(from stored activation aux[M][ldaux]): store out0 in the operand type
    EPI_DW_UPDATE = 3,  // v = s*acc + mu*v; w -= v (FP32 master w,v [M][ldw]); BF16 copy of w to out0 (BF16 nets)
    EPI_DW_STORE = 4,   // store FP32 acc (the summed gradient) to w[M][ldw]
    EPI_DW_STORE_BF16 = 5  // store the summed gradient rounded to BF16 to out0[M][ld0] (B2NN_FLAG_BF16_WIRE)
};

enum ActKind { ACT_LEAKY_RELU = 0, ACT_SIGMOID = 1, ACT_RELU = 2, ACT_TANH = 3, ACT_IDENTITY = 4 };

struct GemmArgs {
    int M, N, K;
    void *out0;   // row-major output (operand type, or fp32 for EPI_LOGITS)
    int ld0;
    const void *aux;  // stored activation for EPI_DX (operand type)
    int ldaux;
    float *w;  // master weights / gradient out
    float *v;  // velocity
    int ldw;
    float scale;     // step / batch
    float momentum;  // mu
    int act;
    float act_param;
    unsigned int *tile_counter;  // work-item counter of the dynamic scheduler (0 at launch, reset by the kernel)
    unsigned int *stalls;        // debug (B2NN_GEMM_STALLS=1): [cta][8] cycle counters of the role warps, or null
    unsigned long long *stamp;   // measurement aid (b2nn_profile_enable 3): [0] <- min %globaltimer at CTA start, [1] <- max at CTA end,
                                 // [2] / [3] <- CTA 0's lifetime in SM cycles / in nanoseconds; or null
    int dbg;                     // debug (B2NN_GEMM_DBG): bit 1 = release the dependent launch at the kernel start; bit 0 = override the TF32 kernel's source-format fields with bits 4-6
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

// Store 32 consecutive values of row m as bf16.
__device__ __forceinline__ void store_bf16_row(__nv_bfloat16 *o0, int ld0, int N, int m, int n0, const float (&x)[32]) {
    __nv_bfloat16 *row = o0 + (size_t)m * ld0 + n0;
    if (n0 + 32 <= N) {
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
            if (n0 + c < N) row[c] = __float2bfloat16_rn(x[c]);
    }
}

__device__ __forceinline__ void store_f32_row(float *base, int ld, int N, int m, int n0, const uint32_t (&r)[32]) {
    float *row = base + (size_t)m * ld + n0;
    if (n0 + 32 <= N) {
#pragma unroll
        for (int q = 0; q < 8; q++)
            reinterpret_cast<float4 *>(row)[q] = make_float4(__uint_as_float(r[4 * q]), __uint_as_float(r[4 * q + 1]),
                                                             __uint_as_float(r[4 * q + 2]), __uint_as_float(r[4 * q + 3]));
    } else {
#pragma unroll
        for (int c = 0; c < 32; c++)
            if (n0 + c < N) row[c] = __uint_as_float(r[c]);
    }
}

// Store 32 consecutive values of row m as fp32 (TF32 nets keep activations / deltas in FP32).
__device__ __forceinline__ void store_f32_row(float *base, int ld, int N, int m, int n0, const float (&x)[32]) {
    float *row = base + (size_t)m * ld + n0;
    if (n0 + 32 <= N) {
#pragma unroll
        for (int q = 0; q < 8; q++)
            reinterpret_cast<float4 *>(row)[q] = make_float4(x[4 * q], x[4 * q + 1], x[4 * q + 2], x[4 * q + 3]);
    } else {
#pragma unroll
        for (int c = 0; c < 32; c++)
            if (n0 + c < N) row[c] = x[c];
    }
}

// x[c] = acc[c] * fprime(a[c]) for 32 columns, a unpacked from BF16 on the fly
template <class F>
__device__ __forceinline__ void dx32(const uint4 (&u)[4], const uint32_t (&r)[32], float (&x)[32], F fprime) {
#pragma unroll
    for (int q = 0; q < 4; q++) {
        const __nv_bfloat162 *h = reinterpret_cast<const __nv_bfloat162 *>(&u[q]);
#pragma unroll
        for (int e = 0; e < 4; e++) {
            const float2 a = __bfloat1622float2(h[e]);
            x[8 * q + 2 * e] = __uint_as_float(r[8 * q + 2 * e]) * fprime(a.x);
            x[8 * q + 2 * e + 1] = __uint_as_float(r[8 * q + 2 * e + 1]) * fprime(a.y);
        }
    }
}

// the same for FP32 activations (TF32 nets)
template <class F>
__device__ __forceinline__ void dx32_f32(const float *arow, int n0, int N, const uint32_t (&r)[32], float (&x)[32], F fprime) {
    if (n0 + 32 <= N) {
#pragma unroll
        for (int q = 0; q < 8; q++) {
            const float4 a = reinterpret_cast<const float4 *>(arow)[q];
            x[4 * q + 0] = __uint_as_float(r[4 * q + 0]) * fprime(a.x);
            x[4 * q + 1] = __uint_as_float(r[4 * q + 1]) * fprime(a.y);
            x[4 * q + 2] = __uint_as_float(r[4 * q + 2]) * fprime(a.z);
            x[4 * q + 3] = __uint_as_float(r[4 * q + 3]) * fprime(a.w);
        }
    } else {
#pragma unroll
        for (int c = 0; c < 32; c++) x[c] = (n0 + c < N) ? __uint_as_float(r[c]) * fprime(arow[c]) : 0.f;
    }
}

// the 32 stored activations a = f(z) of row m, columns n0..n0+31, packed BF16 (zeros beyond the edges): the aux operand
// of the BF16 EPI_DX epilogue.  Does not depend on the accumulator, so the epilogue warps issue it one chunk AHEAD.
__device__ __forceinline__ void load_aux_bf16(const GemmArgs &g, int m, int n0, uint4 (&u)[4]) {
#pragma unroll
    for (int q = 0; q < 4; q++) u[q] = make_uint4(0u, 0u, 0u, 0u);
    if (m >= g.M || n0 >= g.N) return;
    const __nv_bfloat16 *arow = reinterpret_cast<const __nv_bfloat16 *>(g.aux) + (size_t)m * g.ldaux + n0;
    if (n0 + 32 <= g.N) {
#pragma unroll
        for (int q = 0; q < 4; q++) u[q] = reinterpret_cast<const uint4 *>(arow)[q];
    } else {
        uint32_t *uw = reinterpret_cast<uint32_t *>(u);
#pragma unroll
        for (int c = 0; c < 32; c += 2) {
            const float lo = (n0 + c < g.N) ? __bfloat162float(arow[c]) : 0.f;
            const float hi = (n0 + c + 1 < g.N) ? __bfloat162float(arow[c + 1]) : 0.f;
            uw[c >> 1] = pack_bf16x2(lo, hi);
        }
    }
}

// `u`: the chunk's aux values from load_aux_bf16 (BF16 EPI_DX only; ignored by every other epilogue)
template <int EPI, bool TF32>
__device__ __forceinline__ void epilogue_chunk(const GemmArgs &g, int m, int n0, const uint32_t (&r)[32], const uint4 (&u)[4]) {
    if (m >= g.M || n0 >= g.N) return;
    if constexpr (EPI == EPI_ACT) {
        float x[32];
#pragma unroll
        for (int c = 0; c < 32; c++) x[c] = __uint_as_float(r[c]);
        // branch on the activation once per 32 columns, not per element (fewer live registers)
        switch (g.act) {
        case ACT_LEAKY_RELU:
#pragma unroll
            for (int c = 0; c < 32; c++) x[c] = x[c] > 0.f ? x[c] : g.act_param * x[c];
            break;
        case ACT_RELU:
#pragma unroll
            for (int c = 0; c < 32; c++) x[c] = fmaxf(x[c], 0.f);
            break;
        case ACT_SIGMOID:
#pragma unroll
            for (int c = 0; c < 32; c++) x[c] = 1.f / (1.f + __expf(-x[c]));
            break;
        case ACT_TANH:
#pragma unroll
            for (int c = 0; c < 32; c++) x[c] = 1.f - 2.f / (1.f + __expf(2.f * x[c]));  // tanh, BF16-accurate
            break;
        default: break;  // identity
        }
        if constexpr (TF32) store_f32_row(reinterpret_cast<float *>(g.out0), g.ld0, g.N, m, n0, x);
        else store_bf16_row(reinterpret_cast<__nv_bfloat16 *>(g.out0), g.ld0, g.N, m, n0, x);
    } else if constexpr (EPI == EPI_LOGITS) {
        store_f32_row(reinterpret_cast<float *>(g.out0), g.ld0, g.N, m, n0, r);
    } else if constexpr (EPI == EPI_DX && TF32) {
        const float *arow = reinterpret_cast<const float *>(g.aux) + (size_t)m * g.ldaux + n0;
        float x[32];
        switch (g.act) {  // one branch per 32 columns; f' from the stored activation
        case ACT_LEAKY_RELU: { const float p = g.act_param; dx32_f32(arow, n0, g.N, r, x, [p](float a) { return a > 0.f ? 1.f : p; }); break; }
        case ACT_RELU: dx32_f32(arow, n0, g.N, r, x, [](float a) { return a > 0.f ? 1.f : 0.f; }); break;
        case ACT_SIGMOID: dx32_f32(arow, n0, g.N, r, x, [](float a) { return a * (1.f - a); }); break;
        case ACT_TANH: dx32_f32(arow, n0, g.N, r, x, [](float a) { return 1.f - a * a; }); break;
        default: dx32_f32(arow, n0, g.N, r, x, [](float) { return 1.f; });
        }
        store_f32_row(reinterpret_cast<float *>(g.out0), g.ld0, g.N, m, n0, x);
    } else if constexpr (EPI == EPI_DX) {
        float x[32];
        switch (g.act) {  // one branch per 32 columns; f' from the stored activation
        case ACT_LEAKY_RELU: { const float p = g.act_param; dx32(u, r, x, [p](float a) { return a > 0.f ? 1.f : p; }); break; }
        case ACT_RELU: dx32(u, r, x, [](float a) { return a > 0.f ? 1.f : 0.f; }); break;
        case ACT_SIGMOID: dx32(u, r, x, [](float a) { return a * (1.f - a); }); break;
        case ACT_TANH: dx32(u, r, x, [](float a) { return 1.f - a * a; }); break;
        default: dx32(u, r, x, [](float) { return 1.f; });
        }
        store_bf16_row(reinterpret_cast<__nv_bfloat16 *>(g.out0), g.ld0, g.N, m, n0, x);
    } else if constexpr (EPI == EPI_DW_STORE_BF16) {
        float x[32];
#pragma unroll
        for (int c = 0; c < 32; c++) x[c] = __uint_as_float(r[c]);
        store_bf16_row(reinterpret_cast<__nv_bfloat16 *>(g.out0), g.ld0, g.N, m, n0, x);
    } else {  // EPI_DW_STORE
        store_f32_row(g.w, g.ldw, g.N, m, n0, r);
    }
}

// ---- EPI_DW_UPDATE: momentum SGD fused into the gradient GEMM (SoftmaxCrossEntropyNeuralNet.java:325-342) ----------
// adj = step*G/B + mu*prev ; W -= adj ; prev = adj, on the FP32 master arrays, plus the BF16 operand copy (BF16 nets).  The
// summed gradient never reaches HBM and there is no separate update kernel streaming 22 B per parameter through the
// SMs' memory ports beside the contractions (13 % of the C4 step, profiles/r02_notes.md).
// The accumulator arrives one ROW per thread (tcgen05.ld 32x32b: lane = row); read-modify-write of w and v in that shape
// is 32 scattered 16-byte requests per warp instruction (round 1: 3.4 ms/step).  So the warp parks its 32 x 32 chunk in
// shared memory (16-byte slots XOR-swizzled by row: conflict-free both ways) and takes it back with EIGHT LANES PER ROW:
// every warp-wide load / store of w, v and the operand copy then covers four whole 128-byte row segments.
// one 32-row x 32-column accumulator chunk of this warp: r = row `lane`; m_base / n_base = its first row / column.
// Phase 1 issues ALL of the chunk's w / v loads (16 x 16 bytes per lane in flight: one memory round trip per chunk --
// interleaved with the stores the compiler must keep them in order, w and v being written through the same pointers);
// phase 2 takes the accumulators back from the staging rows, updates and stores.
template <bool TF32>
__device__ __forceinline__ void epilogue_update_chunk(const GemmArgs &g, float4 *stage, int lane, int m_base, int n_base,
                                                      const uint32_t (&r)[32]) {
#pragma unroll
    for (int q = 0; q < 8; q++)
        stage[lane * 8 + (q ^ (lane & 7))] = make_float4(__uint_as_float(r[4 * q]), __uint_as_float(r[4 * q + 1]),
                                                          __uint_as_float(r[4 * q + 2]), __uint_as_float(r[4 * q + 3]));
    __syncwarp();
    const int slot = lane & 7, n = n_base + 4 * slot;
    const int row0 = lane >> 3;                       // this lane's rows: row0 + 4 i
    const bool col_ok = n < g.N, full = n + 4 <= g.N;
    // two halves of four quads: 8 x 16 bytes per lane in flight keeps the kernel at ~110 registers, so that the input
    // prologue of the host-fed pipeline still finds room beside the GEMM CTA (with all 16 in flight: 162 registers and
    // the end-to-end step went from 0.85 to 1.11 ms)
#pragma unroll
    for (int h = 0; h < 2; h++) {
        float4 w4[4], v4[4];
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const int m = m_base + row0 + 4 * (4 * h + j);
            w4[j] = v4[j] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (col_ok && m < g.M) {
                const float *wp = g.w + (size_t)m * g.ldw + n, *vp = g.v + (size_t)m * g.ldw + n;
                if (full) {
                    w4[j] = *reinterpret_cast<const float4 *>(wp);
                    v4[j] = *reinterpret_cast<const float4 *>(vp);
                } else {   // ragged right edge: up to three valid columns
                    w4[j].x = wp[0]; v4[j].x = vp[0];
                    if (n + 1 < g.N) { w4[j].y = wp[1]; v4[j].y = vp[1]; }
                    if (n + 2 < g.N) { w4[j].z = wp[2]; v4[j].z = vp[2]; }
                }
            }
        }
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const int row = row0 + 4 * (4 * h + j), m = m_base + row;
            if (!col_ok || m >= g.M) continue;
            const float4 acc = stage[row * 8 + (slot ^ (row & 7))];
            float4 V = v4[j], W = w4[j];
            V.x = g.scale * acc.x + g.momentum * V.x; V.y = g.scale * acc.y + g.momentum * V.y;
            V.z = g.scale * acc.z + g.momentum * V.z; V.w = g.scale * acc.w + g.momentum * V.w;
            W.x -= V.x; W.y -= V.y; W.z -= V.z; W.w -= V.w;
            float *wp = g.w + (size_t)m * g.ldw + n, *vp = g.v + (size_t)m * g.ldw + n;
            if (full) {
                *reinterpret_cast<float4 *>(wp) = W;
                *reinterpret_cast<float4 *>(vp) = V;
                if constexpr (!TF32)   // TF32 nets: the FP32 master array IS the GEMM operand
                    *reinterpret_cast<uint2 *>(reinterpret_cast<__nv_bfloat16 *>(g.out0) + (size_t)m * g.ld0 + n) =
                        make_uint2(pack_bf16x2(W.x, W.y), pack_bf16x2(W.z, W.w));
            } else {
                const float wv[3] = {W.x, W.y, W.z}, vv[3] = {V.x, V.y, V.z};
                for (int c = 0; n + c < g.N; c++) {
                    wp[c] = wv[c]; vp[c] = vv[c];
                    if constexpr (!TF32) reinterpret_cast<__nv_bfloat16 *>(g.out0)[(size_t)m * g.ld0 + n + c] = __float2bfloat16_rn(wv[c]);
                }
            }
        }
    }
    __syncwarp();  // the chunk is out before the next one overwrites the staging rows
}
// pull the w / v lines a warp will update (32 rows x `cols` columns from (m_base, n_base)) into L2 while the tile's main
// loop is still running: the read-modify-write then starts from L2 hits instead of DRAM round trips
__device__ __forceinline__ void epilogue_update_prefetch(const GemmArgs &g, int lane, int m_base, int n_base, int cols) {
    const int lines = (cols * 4 + 127) / 128;         // 128-byte lines per row and array
    for (int e = lane; e < 32 * lines; e += 32) {
        const int m = m_base + e / lines, n = n_base + (e % lines) * 32;
        if (m < g.M && n < g.N) {
            asm volatile("prefetch.global.L2 [%0];" ::"l"(g.w + (size_t)m * g.ldw + n));
            asm volatile("prefetch.global.L2 [%0];" ::"l"(g.v + (size_t)m * g.ldw + n));
        }
    }
}

// Work list of one launch: items [0, full) are 256 x 256 tiles in row-major tile order, items
// [full, total) are the 256 x 128 halves of the remaining tiles (see "Wave quantisation" above).
// Tile order.  The pairs work on ~74 consecutive items at a time and march through K together, so what a wave pulls from
// HBM is (distinct tile rows + distinct tile columns among those items) x K.  In row-major order 74 items are 2-3 whole
// rows of an 8192-wide layer: 3 + 32 operand slabs per wave (C5: every forward GEMM streams the 134 MB weight matrix 55
// times, ~2.3 TB/s of HBM traffic beside the tensor cores).  In bands of GEMM_BAND = 8 tile rows walked column by column
// the same 74 items are an 8 x 9.25 patch: 8 + 10 slabs, and the band's 8 row slabs stay in L2 until the band is done.
struct TileSched {
    int n_tiles, m_tiles, full, total;
    __host__ __device__ __forceinline__ TileSched(int M, int N, int pairs) {
        m_tiles = (M + 2 * GEMM_BM - 1) / (2 * GEMM_BM);
        n_tiles = (N + GEMM_BN - 1) / GEMM_BN;
        const int T = m_tiles * n_tiles, rem = T % pairs;
        if (rem > 0 && 2 * rem <= pairs) { full = T - rem; total = T + rem; }
        else { full = T; total = T; }
    }
    // -> first row / column of the pair's tile and its width (256 or 128)
    __host__ __device__ __forceinline__ void decode(int item, int &m0, int &n0, int &bn) const {
        int t = item, half = 0;
        bn = GEMM_BN;
        if (item >= full) { t = full + ((item - full) >> 1); half = (item - full) & 1; bn = GEMM_BN / 2; }
        int mt, nt;
        if constexpr (GEMM_BAND <= 1) { mt = t / n_tiles; nt = t % n_tiles; }
        else {
            const int per_band = GEMM_BAND * n_tiles, band = t / per_band, idx = t - band * per_band;
            const int r0 = band * GEMM_BAND, left = m_tiles - r0, rows = left < GEMM_BAND ? left : GEMM_BAND;   // the last band may be shorter
            nt = idx / rows; mt = r0 + idx - nt * rows;
        }
        m0 = mt * (2 * GEMM_BM);
        n0 = nt * GEMM_BN + half * (GEMM_BN / 2);
    }
};

// Dynamic work distribution.  A persistent kernel with a static tile list stalls on its slowest CTA, and any
// kernel that shares SMs with it (the NCCL all-reduce, the side-stream update, the PCIe prologue) makes some
// CTAs slow or late.  So pairs DRAW items from a global counter: the otherwise idle MMA warp of the peer CTA
// (rank 1) is the pair's scheduler -- atomicAdd, publish the item into a 4-slot ring in BOTH CTAs' shared
// memory (st.shared::cluster + cluster-scope mbarrier release/acquire), every role warp of the pair reads it
// and releases the slot.  The pair that draws the last sentinel resets the counter for the next launch.
struct SchedRing {
    uint64_t *sfull, *sempty;      // [GEMM_SCHED_SLOTS]; sempty is only waited on in rank 1
    volatile uint32_t *sitem;      // [GEMM_SCHED_SLOTS]
    uint32_t sempty_rank1;         // shared::cluster address of rank 1's sempty[0]
    int slot = 0;
    uint32_t phase = 0;
    int static_next = -1, static_stride = 0;  // static round-robin list instead (tile_counter == nullptr)
    // warp-collective when whole_warp (epilogue warps), else called by one thread
    __device__ __forceinline__ int next(bool whole_warp, int lane) {
        if (static_stride) { const int it = static_next; static_next += static_stride; return it; }
        ptx::mbar_wait_cluster(&sfull[slot], phase);
        const int item = (int)sitem[slot];
        if (whole_warp) __syncwarp();
        if (!whole_warp || lane == 0) ptx::mbar_arrive_cluster(sempty_rank1 + 8u * (uint32_t)slot);
        if (++slot == GEMM_SCHED_SLOTS) { slot = 0; phase ^= 1; }
        return item;
    }
};

// Barrier topology of the pair:
//   full[s]    leader only, 1 arrival (its expect_tx of BOTH CTAs' bytes); both CTAs' TMA credit it
//   empty[s]   one per CTA, signalled in both CTAs by the multicast tcgen05.commit
//   tfull[a]   one per CTA, multicast commit -> each CTA's epilogue warps
//   tempty[a]  leader only, 2 x GEMM_EPI_WARPS arrivals (the peer's warps arrive remotely)
template <int EPI, bool A_MN, bool B_MN, bool TF32>
__device__ __forceinline__ void gemm_pair_body(const CUtensorMap &tmA, const CUtensorMap &tmB, const GemmArgs &g) {
    using Cfg = GemmCfg<TF32>;
    extern __shared__ uint8_t smem_raw[];
    // 128B swizzle needs 1024 B alignment of every box
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t *smem_a = smem;
    uint8_t *smem_b = smem + GEMM_STAGES * GEMM_A_BYTES;
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem + GEMM_STAGES * GEMM_STAGE_BYTES);
    uint64_t *full_bar = bars;
    uint64_t *empty_bar = bars + GEMM_STAGES;
    uint64_t *tfull_bar = bars + 2 * GEMM_STAGES;
    uint64_t *tempty_bar = bars + 2 * GEMM_STAGES + 2;
    uint32_t *tmem_base_slot = reinterpret_cast<uint32_t *>(bars + 2 * GEMM_STAGES + 4);
    SchedRing ring;
    ring.sfull = bars + 2 * GEMM_STAGES + 5;
    ring.sempty = ring.sfull + GEMM_SCHED_SLOTS;
    ring.sitem = reinterpret_cast<volatile uint32_t *>(ring.sempty + GEMM_SCHED_SLOTS);
    ring.sempty_rank1 = ptx::map_to_rank(ring.sempty, 1);
    float4 *epi_stage_base = reinterpret_cast<float4 *>(smem + GEMM_STAGES * GEMM_STAGE_BYTES + 512);
    if (g.tile_counter == nullptr) { ring.static_next = (int)(blockIdx.x >> 1); ring.static_stride = (int)(gridDim.x >> 1); }

    if (g.stamp != nullptr && threadIdx.x == 0) atomicMin(&g.stamp[0], ptx::globaltimer_ns());
    // ... and CTA 0 times itself in SM cycles and in nanoseconds: cycles / ns = the SM clock this launch really ran at
    long long stamp_c0 = 0;
    unsigned long long stamp_t0 = 0;
    if (g.stamp != nullptr && blockIdx.x == 0 && threadIdx.x == 0) { stamp_c0 = clock64(); stamp_t0 = ptx::globaltimer_ns(); }

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int rank = (int)ptx::cluster_ctarank();
    const int num_pairs = gridDim.x >> 1;
    const TileSched sched(g.M, g.N, num_pairs);
    const int k_blocks = (g.K + Cfg::BK - 1) / Cfg::BK;

    if (warp == 0 && lane == 0) {
        ptx::prefetch_tmap(&tmA);
        ptx::prefetch_tmap(&tmB);
        for (int s = 0; s < GEMM_STAGES; s++) {
            ptx::mbar_init(&full_bar[s], 1);
            ptx::mbar_init(&empty_bar[s], 1);
        }
        for (int a = 0; a < 2; a++) {
            ptx::mbar_init(&tfull_bar[a], 1);
            ptx::mbar_init(&tempty_bar[a], 2 * GEMM_EPI_WARPS);
        }
        for (int q = 0; q < GEMM_SCHED_SLOTS; q++) {
            ptx::mbar_init(&ring.sfull[q], 1);
            ptx::mbar_init(&ring.sempty[q], GEMM_SCHED_READERS);
        }
        ptx::fence_barrier_init();
    }
    ptx::cluster_sync();  // both CTAs' barriers exist before anyone signals across the pair
    if (warp == 1) {
        ptx::tmem_alloc_pair(tmem_base_slot, GEMM_TMEM_COLS);
        ptx::tmem_relinquish_pair();
    }
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_base_slot;
    // Programmatic dependent launch (wide_path.cu launches consecutive contractions with it): everything above touched
    // only this CTA's shared memory, TMEM and the kernel parameters, so it may run while the previous kernel of the
    // stream is still draining its last tiles on other SMs; from here on global memory is read and written.
    if (g.dbg & 2) ptx::grid_dep_launch_dependents();   // measurement aid: release the dependent grid at the START instead
    ptx::grid_dep_wait();

    if (warp == 0) {
        // ===================== TMA producer (both CTAs) =====================
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            uint32_t st_wait = 0;
            for (int item = ring.next(false, 0); item < sched.total; item = ring.next(false, 0)) {
                int m0, n0, bn;
                sched.decode(item, m0, n0, bn);
                const int am0 = m0 + rank * GEMM_BM;        // this CTA's rows of A
                const int b_chunks = (bn / 2) / Cfg::CH;    // boxes of B per CTA: 128 (full tile) or 64 (half) rows
                const int bn0 = n0 + rank * (bn / 2);       // this CTA's rows of B
                const uint32_t tx = 2u * (uint32_t)(GEMM_A_BYTES + b_chunks * Cfg::CHUNK_BYTES);
                for (int kb = 0; kb < k_blocks; kb++) {
                    const uint32_t c0 = g.stalls ? (uint32_t)clock64() : 0u;
                    ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
                    if (g.stalls) st_wait += (uint32_t)clock64() - c0;
                    if (rank == 0) ptx::mbar_arrive_expect_tx(&full_bar[stage], tx);
                    uint8_t *sa = smem_a + stage * GEMM_A_BYTES, *sb = smem_b + stage * GEMM_B_BYTES;
                    const int k0 = kb * Cfg::BK;
#pragma unroll
                    for (int c = 0; c < Cfg::CHUNKS_128; c++) {
                        if (A_MN) ptx::tma_load_2d_pair(sa + c * Cfg::CHUNK_BYTES, &tmA, &full_bar[stage], am0 + Cfg::CH * c, k0);
                        else ptx::tma_load_2d_pair(sa + c * Cfg::CHUNK_BYTES, &tmA, &full_bar[stage], k0, am0 + Cfg::CH * c);
                    }
                    for (int c = 0; c < b_chunks; c++) {
                        if (B_MN) ptx::tma_load_2d_pair(sb + c * Cfg::CHUNK_BYTES, &tmB, &full_bar[stage], bn0 + Cfg::CH * c, k0);
                        else ptx::tma_load_2d_pair(sb + c * Cfg::CHUNK_BYTES, &tmB, &full_bar[stage], k0, bn0 + Cfg::CH * c);
                    }
                    if (++stage == GEMM_STAGES) { stage = 0; phase ^= 1; }
                }
            }
            if (g.stalls) g.stalls[blockIdx.x * 8 + 0] = st_wait;
            // this CTA has requested its last operand block: the next kernel of the stream may be placed from now on (its
            // CTAs find room as ours exit).  Released at the start, the pending grid sits at the head of the hardware queue
            // for the whole kernel and can hold back the other streams' kernels (the host-fed input prologue: end to end
            // 1.53 instead of 0.80 ms/step in one run of two).
            ptx::grid_dep_launch_dependents();
        }
    } else if (warp == 1) {
        // ===================== MMA issuer (leader CTA only) =====================
        if (rank == 0 && lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            int acc = 0;
            uint32_t acc_phase = 0;
            uint32_t st_full = 0, st_tempty = 0, st_items = 0;
            const uint32_t st_begin = (uint32_t)clock64();
            // bytes to step one instruction (K = 16 BF16 / 8 TF32) along K inside a staged tile: 32 B in a K-major
            // row, UMMA_K rows of 128 B otherwise (descriptor address units of 16 B)
            constexpr uint64_t a_kstep = (A_MN ? Cfg::KSTEP_MN : 32) >> 4, b_kstep = (B_MN ? Cfg::KSTEP_MN : 32) >> 4;
            for (int item = ring.next(false, 0); item < sched.total; item = ring.next(false, 0)) {
                int m0, n0, bn;
                sched.decode(item, m0, n0, bn);
                const uint32_t idesc = ptx::make_idesc(2 * GEMM_BM, bn, A_MN, B_MN, TF32);
                uint32_t c0 = g.stalls ? (uint32_t)clock64() : 0u;
                ptx::mbar_wait(&tempty_bar[acc], acc_phase ^ 1);  // both CTAs' epilogues drained this accumulator
                if (g.stalls) { st_tempty += (uint32_t)clock64() - c0; st_items++; }
                ptx::tc_fence_after();
                const uint32_t d_tmem = tmem_base + (uint32_t)acc * GEMM_BN;
                for (int kb = 0; kb < k_blocks; kb++) {
                    c0 = g.stalls ? (uint32_t)clock64() : 0u;
                    ptx::mbar_wait(&full_bar[stage], phase);
                    if (g.stalls) st_full += (uint32_t)clock64() - c0;
                    ptx::tc_fence_after();
                    const uint32_t sa = ptx::smem_u32(smem_a + stage * GEMM_A_BYTES);
                    const uint32_t sb = ptx::smem_u32(smem_b + stage * GEMM_B_BYTES);
                    // MN-major TF32 tiles are staged in the 32-byte-atom swizzle (the only MN-major layout kind::tf32 reads)
                    const uint64_t da = !A_MN ? ptx::make_desc_kmajor_sw128(sa)
                                        : TF32 ? ptx::make_desc_mnmajor_sw128_base32b(sa, Cfg::CHUNK_BYTES)
                                               : ptx::make_desc_mnmajor_sw128(sa, Cfg::CHUNK_BYTES);
                    const uint64_t db = !B_MN ? ptx::make_desc_kmajor_sw128(sb)
                                        : TF32 ? ptx::make_desc_mnmajor_sw128_base32b(sb, Cfg::CHUNK_BYTES)
                                               : ptx::make_desc_mnmajor_sw128(sb, Cfg::CHUNK_BYTES);
#pragma unroll
                    for (int k = 0; k < Cfg::BK / Cfg::UMMA_K; k++) {
                        if constexpr (TF32) {
                            // debug (B2NN_GEMM_DBG = 1 | fmt << 4): override the two source-format fields
                            uint32_t id = idesc;
                            if (g.dbg & 1) id = (idesc & ~0x1F80u) | (((uint32_t)g.dbg >> 4 & 7u) << 7) | (((uint32_t)g.dbg >> 4 & 7u) << 10);
                            ptx::mma_tf32_ss_pair(d_tmem, da + a_kstep * k, db + b_kstep * k, id, (kb | k) != 0 ? 1u : 0u);
                        }
                        else
                            ptx::mma_f16_ss_pair(d_tmem, da + a_kstep * k, db + b_kstep * k, idesc, (kb | k) != 0 ? 1u : 0u);
                    }
                    ptx::mma_commit_pair(&empty_bar[stage]);  // frees the slot in both CTAs when these MMAs retire
                    if (++stage == GEMM_STAGES) { stage = 0; phase ^= 1; }
                }
                ptx::mma_commit_pair(&tfull_bar[acc]);
                if (++acc == 2) { acc = 0; acc_phase ^= 1; }
            }
            if (g.stalls) {
                unsigned int *o = g.stalls + blockIdx.x * 8;
                o[1] = st_full; o[2] = st_tempty; o[3] = (uint32_t)clock64() - st_begin; o[4] = st_items;
            }
        } else if (rank == 1 && lane == 0 && g.tile_counter != nullptr) {
            // ===================== scheduler of the pair (the peer's idle MMA warp) =====================
            const uint32_t sitem0 = ptx::map_to_rank(const_cast<uint32_t *>(ring.sitem), 0);
            const uint32_t sfull0 = ptx::map_to_rank(ring.sfull, 0);
            int slot = 0;
            uint32_t phase = 0;
            bool first = true;
            while (true) {
                ptx::mbar_wait_cluster(&ring.sempty[slot], phase ^ 1);  // every reader is done with this slot
                unsigned int item;
                if (first) {  // the pair's first item needs no round trip to L2: items [0, num_pairs) are pre-assigned
                    item = blockIdx.x >> 1;
                    first = false;
                } else {
                    const unsigned int raw = atomicAdd(g.tile_counter, 1u);
                    if (raw == (unsigned int)(sched.total - 1)) *g.tile_counter = 0u;  // last draw of all pairs
                    item = raw + (unsigned int)num_pairs;
                }
                ring.sitem[slot] = item;
                ptx::st_cluster_u32(sitem0 + 4u * (uint32_t)slot, item);
                ptx::mbar_arrive_cluster(ptx::map_to_rank(&ring.sfull[slot], 1));
                ptx::mbar_arrive_cluster(sfull0 + 8u * (uint32_t)slot);
                if (item >= (unsigned int)sched.total) break;
                if (++slot == GEMM_SCHED_SLOTS) { slot = 0; phase ^= 1; }
            }
        }
    } else {
        // ===================== epilogue warps: each CTA drains its own 128 rows =====================
        const int q = warp & 3;                 // TMEM lane quarter this warp may access
        const int part = (warp - 2) >> 2;       // which share of the tile's columns
        int acc = 0;
        uint32_t acc_phase = 0;
        uint32_t st_tfull = 0, st_busy = 0;
        for (int item = ring.next(true, lane); item < sched.total; item = ring.next(true, lane)) {
            int m0, n0, bn;
            sched.decode(item, m0, n0, bn);
            if constexpr (EPI == EPI_DW_UPDATE)   // while the tile's main loop runs
                epilogue_update_prefetch(g, lane, m0 + rank * GEMM_BM + q * 32, n0 + part * (bn / GEMM_EPI_PARTS),
                                         bn / GEMM_EPI_PARTS);
            const int m = m0 + rank * GEMM_BM + q * 32 + lane;
            const int cbeg = part * (bn / GEMM_EPI_PARTS), cend = cbeg + bn / GEMM_EPI_PARTS;
            // BF16 EPI_DX reads the stored activations of the tile (aux): they do not depend on the accumulator, so the
            // first chunk's are requested while the tile's main loop is still running and every later chunk's one chunk
            // ahead (ua / ub alternate; no copies)
            constexpr bool AUX = (EPI == EPI_DX) && !TF32 && GEMM_EPI_PIPE >= 1;
            uint4 ua[4], ub[4];
            if constexpr (AUX) load_aux_bf16(g, m, n0 + cbeg, ua);
            const uint32_t c0 = g.stalls ? (uint32_t)clock64() : 0u;
            ptx::mbar_wait(&tfull_bar[acc], acc_phase);
            const uint32_t c1 = g.stalls ? (uint32_t)clock64() : 0u;
            ptx::tc_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)acc * GEMM_BN;
            if constexpr (EPI == EPI_DW_UPDATE || GEMM_EPI_PIPE == 0) {
#pragma unroll 1
                for (int c0 = cbeg; c0 < cend; c0 += 32) {
                    uint32_t r[32];
                    ptx::tmem_ld_32x32b_x32(taddr + (uint32_t)c0, r);
                    ptx::tmem_ld_wait();
                    if constexpr (EPI == EPI_DW_UPDATE)
                        epilogue_update_chunk<TF32>(g, epi_stage_base + (size_t)(warp - 2) * (GEMM_EPI_STAGE_BYTES / 16), lane,
                                                    m0 + rank * GEMM_BM + q * 32, n0 + c0, r);
                    else {
                        if constexpr (EPI == EPI_DX && !TF32) load_aux_bf16(g, m, n0 + c0, ua);
                        epilogue_chunk<EPI, TF32>(g, m, n0 + c0, r, ua);
                    }
                }
            } else {
                // one accumulator chunk at a time; only the aux loads run ahead, so a chunk of EPI_DX no longer pays a TMEM
                // round trip AND a memory round trip before its first store (measured: 11.8 K -> 11.0 K busy cycles per
                // 256 x 256 tile; the K = 1000 delta GEMM, whose main loop is 8.2 K cycles per tile, 49.9 K -> 49.3 K)
#pragma unroll 1
                for (int c0 = cbeg; c0 < cend; c0 += 64) {
                    const bool second = c0 + 32 < cend, third = c0 + 64 < cend;
                    uint32_t r[32];
                    ptx::tmem_ld_32x32b_x32(taddr + (uint32_t)c0, r);
                    if constexpr (AUX) { if (second) load_aux_bf16(g, m, n0 + c0 + 32, ub); }
                    ptx::tmem_ld_wait();
                    epilogue_chunk<EPI, TF32>(g, m, n0 + c0, r, ua);
                    if (second) {
                        ptx::tmem_ld_32x32b_x32(taddr + (uint32_t)(c0 + 32), r);
                        if constexpr (AUX) { if (third) load_aux_bf16(g, m, n0 + c0 + 64, ua); }
                        ptx::tmem_ld_wait();
                        epilogue_chunk<EPI, TF32>(g, m, n0 + c0 + 32, r, ub);
                    }
                }
            }
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive_leader(&tempty_bar[acc]);
            if (g.stalls) { st_tfull += c1 - c0; st_busy += (uint32_t)clock64() - c1; }
            if (++acc == 2) { acc = 0; acc_phase ^= 1; }
        }
        if (g.stalls && warp == 2 && lane == 0) { g.stalls[blockIdx.x * 8 + 5] = st_tfull; g.stalls[blockIdx.x * 8 + 6] = st_busy; }
    }

    ptx::tc_fence_before();
    ptx::cluster_sync();  // the peer may still read this CTA's smem / signal its barriers until here
    if (warp == 1) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc_pair(tmem_base, GEMM_TMEM_COLS);
    }
    if (g.stamp != nullptr && threadIdx.x == 0) atomicMax(&g.stamp[1], ptx::globaltimer_ns());
    if (g.stamp != nullptr && blockIdx.x == 0 && threadIdx.x == 0) {
        g.stamp[2] = (unsigned long long)(clock64() - stamp_c0);
        g.stamp[3] = ptx::globaltimer_ns() - stamp_t0;
    }
}

template <int EPI, bool A_MN, bool B_MN>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(GEMM_THREADS, 1)
gemm_bf16_pair_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                      const GemmArgs g) {
    gemm_pair_body<EPI, A_MN, B_MN, false>(tmA, tmB, g);
}

// the same kernel over FP32 buffers read as TF32 (tcgen05.mma kind::tf32), see the header comment
template <int EPI, bool A_MN, bool B_MN>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(GEMM_THREADS, 1)
gemm_tf32_pair_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                      const GemmArgs g) {
    gemm_pair_body<EPI, A_MN, B_MN, true>(tmA, tmB, g);
}

}  // namespace b2nn
