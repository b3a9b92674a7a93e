// f64_fused.cuh -- K1: persistent fused FP64 training kernel for the reference's small nets
// (three weight matrices d0-d1-d2-d3, e.g. 784-300-100-10 batch 16).
//
// One cooperative launch runs `iters` consecutive mini-batch steps (forward, softmax/CE or general
// output delta, backprop, batch-summed gradient, momentum update -- SoftmaxCrossEntropyNeuralNet.java
// :164-198, :229-287, :297-346) with the weights resident in shared memory, distributed over all SMs:
//   * CTA c owns a slice of the COLUMNS of W0 (and the matching rows of W1 for backprop) and, if
//     c < d2-slices, a slice of the columns of W1; W2 (tiny) is replicated in every CTA.
//   * per step only TWO grid-wide barriers: after z1 (layer-1 pre-activations published to L2) and
//     after z2.  Everything after that -- z3, softmax, delta3, delta2 (replicated, tiny), delta1 for the
//     CTA's own columns, all gradients and updates -- is CTA-local, so the next step starts without
//     another barrier.  Replicated weights stay consistent because every replica applies the same
//     arithmetic to the same inputs in the same order.
//   * dot products are split over the lanes of a warp and combined with a fixed xor-shuffle tree
//     (deterministic; order differs from the reference's ascending-i loop => parity by tolerance);
//     batch sums and the update keep the reference's order and un-fused multiply/add.
#pragma once
#include <cstdint>

#include "f64_kernels.cuh"

namespace b2nn {
namespace f64 {

struct FusedParams {
    int d0, d1, d2, d3, B, iters;
    int kind, inner_act, last_act, loss_kind, stable;
    double inner_p, last_p, step, momentum;
    const double *X;        // [n][d0] rows (or null when Xu8 is set)
    const uint8_t *Xu8;     // [n][d0] rows as pixel bytes, value = byte/255.0 (exactly how the rows were made)
    const double *T;        // [n][d3] dense targets or null
    const int32_t *labels;  // [n] or null
    const int32_t *idx;     // [iters][B]
    double *W0, *W1, *W2, *V0, *V1, *V2;  // canonical [in][out] state (read at start, written at end)
    double *V0t;            // [d1][d0] scratch: velocity of W0, column-contiguous during the run
    double *V1b;            // [d1][d2] scratch: velocity replica for the row-slice copy of W1
    double *V2r;            // [grid][d2*d3] scratch: per-CTA velocity replica of W2
    double *z1buf;          // [2][B][d1] published pre-activations (double buffered by step parity)
    double *z2buf;          // [2][B][d2]
    unsigned long long *bar;  // [2]: grid barrier counter (never reset: bar_base = its value when this launch starts); sticky give-up word
    unsigned long long bar_base;
    double *loss_out;       // [iters] sum over the batch of per-sample losses, or null
    // shared-memory carve-up (in doubles), computed by the host with the same slice formulas
    int nc0max, nc1max;
    int o_a0, o_a1, o_a2, o_d2, o_w0, o_w1c, o_w1r, o_w2, o_z1o, o_d1o, o_z3, o_d3, o_t, o_misc;
    int o_a1o, o_part;  // own activations [B][nc0max]; chunk partial sums [KS][B][4]
    int o_w2t;  // [d3][d2] transposed copy of W2 for the delta2 pass
    int o_prof;                  // 16 cycle counters of the measurement aid
    int o_lut, o_stage, o_mbar;  // 256-entry f(byte/255) table, byte staging for one batch, mbarrier
    int bulk;                    // rows can be fetched with cp.async.bulk (16-byte aligned rows)
    unsigned long long *prof;    // optional [16] cycle sums per phase (CTA 0), measurement aid (B2NN_K1_PROF=1)
};

__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Wait (thread 0) until the barrier counter reaches `target`.  Bounded: a cooperative launch keeps every CTA resident, so
// the only way to wait forever is a CTA that died; after K1_MAX_POLLS polls (each an L2 round trip: seconds in total)
// the CTA records the fact in bar[1] (the host turns it into B2NN_ERR_CUDA), stops waiting for good and lets the launch end.
constexpr unsigned int K1_MAX_POLLS = 1u << 24;
__device__ __forceinline__ void grid_barrier_wait(unsigned long long *bar, unsigned long long target, bool &gave_up) {
    if (gave_up) return;
    unsigned long long v;
    unsigned int polls = 0;
    do {
        asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(bar) : "memory");
    } while (v < target && ++polls < K1_MAX_POLLS);
    if (v < target) {
        gave_up = true;
        atomicMax(bar + 1, (unsigned long long)polls);
    }
}

__device__ __forceinline__ void grid_barrier(unsigned long long *bar, unsigned long long target, bool &gave_up) {
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        atomicAdd(bar, 1ULL);
        grid_barrier_wait(bar, target, gave_up);
    }
    __syncthreads();
}

// the same barrier with `work()` run by the whole CTA between this CTA's arrival and the wait for the others: thread 0
// polls, every other warp is idle there anyway
template <class F>
__device__ __forceinline__ void grid_barrier_overlap(unsigned long long *bar, unsigned long long target, bool &gave_up,
                                                     F work) {
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        atomicAdd(bar, 1ULL);
    }
    work();
    if (threadIdx.x == 0) grid_barrier_wait(bar, target, gave_up);
    __syncthreads();
}

// ---- async bulk copy (TMA, 1-D) of one dataset row into shared memory, completion on an mbarrier ----
__device__ __forceinline__ uint32_t smem_addr(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void bulk_row_load(void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_addr(dst)),
                 "l"(src), "r"(bytes), "r"(smem_addr(bar))
                 : "memory");
}
__device__ __forceinline__ void mbar_expect(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait_parity(uint64_t *bar, uint32_t parity) {
    uint32_t ok = 0;
    while (!ok) {
        asm volatile(
            "{\n\t.reg .pred P;\n\tmbarrier.try_wait.parity.shared::cta.b64 P, [%1], %2;\n\tselp.b32 %0, 1, 0, P;\n\t}\n"
            : "=r"(ok)
            : "r"(smem_addr(bar)), "r"(parity)
            : "memory");
    }
}

__device__ __forceinline__ int slice_lo(int c, int n, int G) { return (int)(((long long)c * n) / G); }

// momentum update in the reference's evaluation order (S:333-339)
// (x / B is computed as x * (1/B) when B is a power of two: bit-identical, no divide)
__device__ __forceinline__ double sgd_adj(double step, double g, int B, double invB, double mom, double prev) {
    const double sg = __dmul_rn(step, g);
    const double q = (invB != 0.0) ? __dmul_rn(sg, invB) : __ddiv_rn(sg, (double)B);
    return __dadd_rn(q, __dmul_rn(mom, prev));
}

// Register-blocked partial dot products out of shared memory (the loops are shared-memory-bandwidth
// bound, so every loaded operand is reused): warp unit (rg, ks) handles rows 4*rg..4*rg+3, all nc <= 4
// columns, and the ks-th of KS contiguous chunks of K; lanes stride over k.  part[ks][b][c] receives the
// chunk sums (fixed xor-shuffle tree); the caller adds the KS chunks in ascending ks (deterministic).
constexpr int FUSED_CMAX = 4;
__device__ __forceinline__ int fused_ksplit(int B, int NW) {
    const int RG = (B + 3) / 4;
    return RG >= NW ? 1 : NW / RG;
}
__device__ __forceinline__ void blocked_dot(const double *A, const double *Wc, int B, int K, int nc, double *part,
                                            int warp, int lane, int NW) {
    const int RG = (B + 3) / 4, KS = fused_ksplit(B, NW);
    const int kchunk = (K + KS - 1) / KS;
    for (int u = warp; u < RG * KS; u += NW) {
        const int rg = u / KS, ks = u - rg * KS;
        const int kb = ks * kchunk, ke = min(K, kb + kchunk);
        double acc[4][FUSED_CMAX];
#pragma unroll
        for (int r = 0; r < 4; r++)
#pragma unroll
            for (int cc = 0; cc < FUSED_CMAX; cc++) acc[r][cc] = 0.0;
        const int b0 = rg * 4;
        for (int k = kb + lane; k < ke; k += 32) {
            double av[4], wv[FUSED_CMAX];
#pragma unroll
            for (int r = 0; r < 4; r++) av[r] = (b0 + r < B) ? A[(size_t)(b0 + r) * K + k] : 0.0;
#pragma unroll
            for (int cc = 0; cc < FUSED_CMAX; cc++) wv[cc] = (cc < nc) ? Wc[(size_t)cc * K + k] : 0.0;
#pragma unroll
            for (int r = 0; r < 4; r++)
#pragma unroll
                for (int cc = 0; cc < FUSED_CMAX; cc++) acc[r][cc] = fma(av[r], wv[cc], acc[r][cc]);
        }
        // 16 sums over the 32 lanes with ONE butterfly: at every step a lane keeps half of its values and hands the
        // other half to its partner, so 8 + 4 + 2 + 1 + 1 exchanges replace 16 x 5.  Each sum is formed by the same
        // xor tree (16, 8, 4, 2, 1) as warp_sum_d, hence the same bits.
        static_assert(FUSED_CMAX == 4, "the butterfly below reduces a 4 x 4 block");
        const bool u16 = (lane & 16) != 0, u8 = (lane & 8) != 0, u4 = (lane & 4) != 0, u2 = (lane & 2) != 0;
        double v8[8], v4[4], v2[2];
#pragma unroll
        for (int j = 0; j < 8; j++) {
            const double lo = acc[j >> 2][j & 3], hi = acc[2 + (j >> 2)][j & 3];
            v8[j] = (u16 ? hi : lo) + __shfl_xor_sync(0xffffffffu, u16 ? lo : hi, 16);
        }
#pragma unroll
        for (int j = 0; j < 4; j++) v4[j] = (u8 ? v8[4 + j] : v8[j]) + __shfl_xor_sync(0xffffffffu, u8 ? v8[j] : v8[4 + j], 8);
#pragma unroll
        for (int j = 0; j < 2; j++) v2[j] = (u4 ? v4[2 + j] : v4[j]) + __shfl_xor_sync(0xffffffffu, u4 ? v4[j] : v4[2 + j], 4);
        double v1 = (u2 ? v2[1] : v2[0]) + __shfl_xor_sync(0xffffffffu, u2 ? v2[0] : v2[1], 2);
        v1 += __shfl_xor_sync(0xffffffffu, v1, 1);
        const int r = (u16 ? 2 : 0) + (u8 ? 1 : 0), cc = (u4 ? 2 : 0) + (u2 ? 1 : 0);  // the element this lane ended up with
        if ((lane & 1) == 0 && cc < nc && b0 + r < B) part[((size_t)ks * B + b0 + r) * FUSED_CMAX + cc] = v1;
    }
}

// issue the asynchronous fetch of the batch of step `it` (rows named by idx) into `dst`; called by all lanes of ONE
// warp, one row per lane (a single thread issuing B copies one after the other, each behind a load of idx[b], cost
// ~4 K cycles per step on the critical path)
__device__ __forceinline__ void issue_batch_fetch(const FusedParams &p, int it, void *dst, uint64_t *mbar, int lane) {
    const int32_t *idx = p.idx + (size_t)it * p.B;
    const uint32_t row_bytes = p.Xu8 ? (uint32_t)p.d0 : (uint32_t)p.d0 * 8u;
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy accesses to dst are done
    if (lane == 0) mbar_expect(mbar, row_bytes * (uint32_t)p.B);
    __syncwarp();
    for (int b = lane; b < p.B; b += 32) {
        const size_t row = (size_t)idx[b];
        const void *src = p.Xu8 ? (const void *)(p.Xu8 + row * p.d0) : (const void *)(p.X + row * p.d0);
        bulk_row_load((char *)dst + (size_t)b * row_bytes, src, row_bytes, mbar);
    }
}

// cycle sums are kept in shared memory and written out once (a read-modify-write of global memory per mark cost
// ~700 cycles each and distorted the intervals)
#define K1_MARK(i)                                                      \
    if (p.prof != nullptr && c == 0 && tid == 0) {                      \
        const long long now_ = clock64();                               \
        profs[i] += (unsigned long long)(now_ - tprev_);                \
        tprev_ = now_;                                                  \
    }

template <int IACT>
__global__ void __launch_bounds__(512, 1) fused_train_kernel(const FusedParams p) {
    extern __shared__ double sm[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, NW = blockDim.x >> 5, NT = blockDim.x;
    const int c = blockIdx.x, G = gridDim.x;
    const int d0 = p.d0, d1 = p.d1, d2 = p.d2, d3 = p.d3, B = p.B;
    const int j0 = slice_lo(c, d1, G), j1 = slice_lo(c + 1, d1, G), nc0 = j1 - j0;  // own columns of W0 / rows of W1
    const int k0 = slice_lo(c, d2, G), k1 = slice_lo(c + 1, d2, G), nc1 = k1 - k0;  // own columns of W1
    const int ncm = p.nc0max;
    const double invB = ((B & (B - 1)) == 0) ? 1.0 / (double)B : 0.0;
    double *a0 = sm + p.o_a0;    // [B][d0]
    double *a1 = sm + p.o_a1;    // [B][d1]   (only staged by CTAs that own W1 columns)
    double *a2 = sm + p.o_a2;    // [B][d2]
    double *dl2 = sm + p.o_d2;   // [B][d2]
    double *w0 = sm + p.o_w0;    // [nc0][d0]   column-contiguous
    double *w1c = sm + p.o_w1c;  // [nc1][d1]   own columns of W1, column-contiguous
    double *w1r = sm + p.o_w1r;  // [nc0][d2]   own rows of W1
    double *w2 = sm + p.o_w2;    // [d2][d3]
    double *w2t = sm + p.o_w2t;  // [d3][d2]
    double *z1o = sm + p.o_z1o;  // [B][nc0max] own pre-activations
    double *d1o = sm + p.o_d1o;  // [B][nc0max] own deltas
    double *a1o_buf = sm + p.o_a1o;  // [2][B][nc0max] own activations f(z1), double buffered by step parity (the W1-row
                                     // update of step it runs under barrier 1 of step it + 1, after phase 1 has written the next)
    double *part = sm + p.o_part;  // [KS][B][4] chunk sums of the blocked dot products
    const int KS = fused_ksplit(B, NW);
    double *z3 = sm + p.o_z3;    // [B][d3]  (reused for exp terms / loss terms)
    double *dl3 = sm + p.o_d3;   // [B][d3]
    double *ttbuf = sm + p.o_t;  // [2][B][d3] targets, double buffered by step parity
    double *misc = sm + p.o_misc;  // [B] row maxima / sums
    double *lut = sm + p.o_lut;  // [256] f(byte / 255.0)
    uint8_t *stage = reinterpret_cast<uint8_t *>(sm + p.o_stage);  // [B][d0] bytes
    uint64_t *mbar = reinterpret_cast<uint64_t *>(sm + p.o_mbar);
    const bool u8 = p.Xu8 != nullptr;

    // ---- one-time setup: weight slices into shared memory, velocity slices into the scratch layouts ----
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_addr(mbar)) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    for (int jj = 0; jj < nc0; jj++)
        for (int k = tid; k < d0; k += NT) {
            w0[jj * d0 + k] = p.W0[(size_t)k * d1 + j0 + jj];
            p.V0t[(size_t)(j0 + jj) * d0 + k] = p.V0[(size_t)k * d1 + j0 + jj];
        }
    for (int kk = 0; kk < nc1; kk++)
        for (int j = tid; j < d1; j += NT) w1c[kk * d1 + j] = p.W1[(size_t)j * d2 + k0 + kk];
    for (int jj = 0; jj < nc0; jj++)
        for (int k = tid; k < d2; k += NT) {
            w1r[jj * d2 + k] = p.W1[(size_t)(j0 + jj) * d2 + k];
            p.V1b[(size_t)(j0 + jj) * d2 + k] = p.V1[(size_t)(j0 + jj) * d2 + k];
        }
    double *v2 = p.V2r + (size_t)c * d2 * d3;
    for (int e = tid; e < d2 * d3; e += NT) {
        w2[e] = p.W2[e]; v2[e] = p.V2[e];
        w2t[(e % d3) * d2 + e / d3] = p.W2[e];
    }
    if (u8 && tid < 256) lut[tid] = act(IACT, p.inner_p, (double)tid / 255.0);  // MNISTTrainer.java:97 + S:185
    __syncthreads();
    if (p.bulk && warp == 0) issue_batch_fetch(p, 0, u8 ? (void *)stage : (void *)a0, mbar, lane);

    unsigned long long bar_target = p.bar_base;
    bool gave_up = false;  // thread 0: a grid-barrier wait ran out of polls (see grid_barrier_wait)
    unsigned long long *profs = reinterpret_cast<unsigned long long *>(sm + p.o_prof);
    if (p.prof != nullptr && c == 0 && tid == 0)
        for (int i = 0; i < 16; i++) profs[i] = 0ULL;
    long long tprev_ = clock64();
    // ---- the W1 / W2 updates of a step (phase 5b) are deferred into the barrier waits of the NEXT step, where every CTA
    // idles anyway: the own-column copy of W1 (first needed by phase 2) under barrier 1, the own-row copy of W1 and W2
    // (first needed by phases 3 / 4) under barrier 2; after the loop for the last step.  Inputs: dl2, a1, dl3, a2 of the
    // step (untouched until phases 2 / 3 of the next one) and its a1o buffer.
    auto update_w1_columns = [&]() {
        // W1[:, own columns] (forward copy): g = sum_b delta2[b][k] * a1[b][j]
        for (int kk = 0; kk < nc1; kk++)
            for (int j = tid; j < d1; j += NT) {
                double *vp = p.V1 + (size_t)j * d2 + k0 + kk;
                const double vprev = *vp;
                double g = 0.0;
                for (int b = 0; b < B; b++) g = __dadd_rn(g, __dmul_rn(dl2[b * d2 + k0 + kk], a1[(size_t)b * d1 + j]));
                const double adj = sgd_adj(p.step, g, B, invB, p.momentum, vprev);
                w1c[kk * d1 + j] = __dsub_rn(w1c[kk * d1 + j], adj);
                *vp = adj;
            }
    };
    auto update_w1_rows_w2 = [&](const double *a1o) {
        // W1[own rows, :] (backprop copy): same arithmetic on the same inputs => identical values
        for (int e = tid; e < nc0 * d2; e += NT) {  // all (row, column) pairs at once: one L2 round trip, not nc0
            const int jj = e / d2, k = e - jj * d2;
            double *vp = p.V1b + (size_t)(j0 + jj) * d2 + k;
            const double vprev = *vp;
            double g = 0.0;
            for (int b = 0; b < B; b++) g = __dadd_rn(g, __dmul_rn(dl2[b * d2 + k], a1o[b * ncm + jj]));
            const double adj = sgd_adj(p.step, g, B, invB, p.momentum, vprev);
            w1r[jj * d2 + k] = __dsub_rn(w1r[jj * d2 + k], adj);
            *vp = adj;
        }
        // W2 (replicated): g = sum_b delta3[b][j] * a2[b][i]
        for (int e = tid; e < d2 * d3; e += NT) {
            const int i = e / d3, j = e - i * d3;
            const double vprev = v2[e];
            double g = 0.0;
            for (int b = 0; b < B; b++) g = __dadd_rn(g, __dmul_rn(dl3[b * d3 + j], a2[b * d2 + i]));
            const double adj = sgd_adj(p.step, g, B, invB, p.momentum, vprev);
            const double wn = __dsub_rn(w2[e], adj);
            w2[e] = wn;
            w2t[j * d2 + i] = wn;
            v2[e] = adj;
        }
    };

    for (int it = 0; it < p.iters; it++) {
        const int32_t *idx = p.idx + (size_t)it * B;
        double *a1o = a1o_buf + (size_t)(it & 1) * B * ncm;
        double *z1pub = p.z1buf + (size_t)(it & 1) * B * d1;
        double *z2pub = p.z2buf + (size_t)(it & 1) * B * d2;

        // ---- stage the batch: a0 = f(x) (the input layer is activated too, S:183-186) and the targets ----
        // targets are double buffered: this step's were fetched during the previous step (or below for
        // it == 0); the next step's are requested now and parked in a register until after phase 1
        double *tt = ttbuf + (size_t)(it & 1) * B * d3, *ttn = ttbuf + (size_t)((it + 1) & 1) * B * d3;
        if (it == 0) {
            for (int e = tid; e < B * d3; e += NT) {
                const int b = e / d3, j = e - b * d3;
                tt[e] = p.T ? p.T[(size_t)idx[b] * d3 + j] : ((p.labels[idx[b]] == j) ? 1.0 : 0.0);
            }
        }
        double tnext = 0.0;
        const bool have_next = it + 1 < p.iters;
        if (have_next && tid < B * d3) {
            const int b = tid / d3, j = tid - b * d3;
            const int32_t r = idx[B + b];
            tnext = p.T ? p.T[(size_t)r * d3 + j] : ((p.labels[r] == j) ? 1.0 : 0.0);
        }
        if (p.bulk) {
            mbar_wait_parity(mbar, (uint32_t)(it & 1));
            K1_MARK(15)  // targets + wait for the prefetched rows
            if (u8) {
                // d0 % 16 == 0 here: 4 bytes per lane per iteration, two 16-byte stores
                const uint32_t *st4 = reinterpret_cast<const uint32_t *>(stage);
                double2 *a02 = reinterpret_cast<double2 *>(a0);
                for (int q = tid; q < B * d0 / 4; q += NT) {
                    const uint32_t v = st4[q];
                    a02[2 * q] = make_double2(lut[v & 0xff], lut[(v >> 8) & 0xff]);
                    a02[2 * q + 1] = make_double2(lut[(v >> 16) & 0xff], lut[v >> 24]);
                }
            } else {
                for (int b = warp; b < B; b += NW)
                    for (int k = lane; k < d0; k += 32) a0[b * d0 + k] = act(IACT, p.inner_p, a0[b * d0 + k]);
            }
        } else {
            for (int b = warp; b < B; b += NW) {
                const size_t row = (size_t)idx[b] * d0;
                for (int k = lane; k < d0; k += 32)
                    a0[b * d0 + k] = u8 ? lut[p.Xu8[row + k]] : act(IACT, p.inner_p, p.X[row + k]);
            }
        }
        __syncthreads();

        K1_MARK(0)  // batch staged
        // ---- phase 1: z1[:, own columns] = a0 . W0[:, own]; publish ----
        blocked_dot(a0, w0, B, d0, nc0, part, warp, lane, NW);
        if (have_next) {
            if (tid < B * d3) ttn[tid] = tnext;
            for (int e = NT + tid; e < B * d3; e += NT) {  // only when B*d3 exceeds the block size
                const int b = e / d3, j = e - b * d3;
                const int32_t r = idx[B + b];
                ttn[e] = p.T ? p.T[(size_t)r * d3 + j] : ((p.labels[r] == j) ? 1.0 : 0.0);
            }
        }
        __syncthreads();
        for (int e = tid; e < B * nc0; e += NT) {
            const int b = e / nc0, jj = e - b * nc0;
            double z = 0.0;
            for (int ks = 0; ks < KS; ks++) z += part[((size_t)ks * B + b) * FUSED_CMAX + jj];
            z1o[b * ncm + jj] = z;
            a1o[b * ncm + jj] = act(IACT, p.inner_p, z);
            z1pub[(size_t)b * d1 + j0 + jj] = z;
        }
        K1_MARK(1)  // phase 1
        bar_target += G;
        // the byte staging buffer has been free since the batch was staged: warp 1 requests the NEXT batch while
        // thread 0 waits for the other CTAs
        grid_barrier_overlap(p.bar, bar_target, gave_up, [&] {
            if (p.bulk && u8 && warp == 1 && it + 1 < p.iters) issue_batch_fetch(p, it + 1, stage, mbar, lane);
            if (it > 0) update_w1_columns();   // phase 5b of the previous step, first part
        });
        K1_MARK(2)  // barrier 1

        // ---- phase 2: owners of W1 columns: a1 = f(z1) (all of it), z2[:, own] = a1 . W1[:, own]; publish ----
        if (nc1 > 0) {
            {
                // all of a thread's loads of the published z1 are issued before the first one is used: one L2 round
                // trip for up to 10 x 512 values instead of one per loop trip
                const int n = B * d1;
                constexpr int ZL = 10;
                for (int e0 = tid; e0 < n; e0 += ZL * NT) {
                    double q[ZL];
#pragma unroll
                    for (int u = 0; u < ZL; u++) q[u] = (e0 + u * NT < n) ? __ldcg(z1pub + e0 + u * NT) : 0.0;
#pragma unroll
                    for (int u = 0; u < ZL; u++)
                        if (e0 + u * NT < n) a1[e0 + u * NT] = act(IACT, p.inner_p, q[u]);
                }
            }
            __syncthreads();
            blocked_dot(a1, w1c, B, d1, nc1, part, warp, lane, NW);
            __syncthreads();
            for (int e = tid; e < B * nc1; e += NT) {
                const int b = e / nc1, kk = e - b * nc1;
                double z = 0.0;
                for (int ks = 0; ks < KS; ks++) z += part[((size_t)ks * B + b) * FUSED_CMAX + kk];
                z2pub[(size_t)b * d2 + k0 + kk] = z;
            }
        }
        K1_MARK(3)  // phase 2
        bar_target += G;
        grid_barrier_overlap(p.bar, bar_target, gave_up, [&] {
            if (it > 0) update_w1_rows_w2(a1o_buf + (size_t)((it - 1) & 1) * B * ncm);   // ... second part
        });
        K1_MARK(4)  // barrier 2

        // ---- phase 3 (replicated): a2 = f(z2), z3 = a2 . W2, output delta, delta2 ----
        {
            const int n = B * d2;
            constexpr int ZL = 4;
            for (int e0 = tid; e0 < n; e0 += ZL * NT) {
                double q[ZL];
#pragma unroll
                for (int u = 0; u < ZL; u++) q[u] = (e0 + u * NT < n) ? __ldcg(z2pub + e0 + u * NT) : 0.0;
#pragma unroll
                for (int u = 0; u < ZL; u++)
                    if (e0 + u * NT < n) {
                        dl2[e0 + u * NT] = act_prime(IACT, p.inner_p, q[u]);  // f'(z2), multiplied into delta2 below
                        a2[e0 + u * NT] = act(IACT, p.inner_p, q[u]);
                    }
            }
        }
        __syncthreads();
        K1_MARK(8)
        for (int e = tid; e < B * d3; e += NT) {
            const int b = e / d3, j = e - b * d3;
            const double *ar = a2 + (size_t)b * d2;
            double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
            int k = 0;
            for (; k + 3 < d2; k += 4) {
                s0 = fma(ar[k], w2[k * d3 + j], s0);
                s1 = fma(ar[k + 1], w2[(k + 1) * d3 + j], s1);
                s2 = fma(ar[k + 2], w2[(k + 2) * d3 + j], s2);
                s3 = fma(ar[k + 3], w2[(k + 3) * d3 + j], s3);
            }
            for (; k < d2; k++) s0 = fma(ar[k], w2[k * d3 + j], s0);
            z3[e] = (s0 + s1) + (s2 + s3);
        }
        __syncthreads();
        K1_MARK(9)
        if (p.kind == 0) {
            // softmax (S:357-376) + cross entropy (S:213-217) + delta (S:249-251)
            if (tid < B) {
                double mx = 0.0;
                if (p.stable) {
                    mx = z3[tid * d3];
                    for (int j = 1; j < d3; j++) mx = fmax(mx, z3[tid * d3 + j]);
                }
                misc[tid] = mx;
            }
            __syncthreads();
            for (int e = tid; e < B * d3; e += NT) z3[e] = exp(z3[e] - misc[e / d3]);
            __syncthreads();
            if (tid < B) {
                double sum = 0.0;
                for (int j = 0; j < d3; j++) sum = __dadd_rn(sum, z3[tid * d3 + j]);  // ascending i like the reference
                misc[tid] = sum;
            }
            __syncthreads();
            for (int e = tid; e < B * d3; e += NT) {
                const double pr = __ddiv_rn(z3[e], misc[e / d3]);
                dl3[e] = __dsub_rn(pr, tt[e]);
                if (p.loss_out) z3[e] = __dmul_rn(tt[e], log(pr));
            }
        } else {
            // general output layer (G:268-271, G:236-239)
            for (int e = tid; e < B * d3; e += NT) {
                const double z = z3[e], a = act(p.last_act, p.last_p, z);
                dl3[e] = __dmul_rn(loss_prime(p.loss_kind, a, tt[e]), act_prime(p.last_act, p.last_p, z));
                if (p.loss_out) z3[e] = -loss_fn(p.loss_kind, a, tt[e]);
            }
        }
        __syncthreads();
        K1_MARK(10)
        if (p.loss_out && c == 0) {
            // per-row loss in ascending j (one thread per row), then the batch sum in ascending b: the same
            // order of additions as a single serial loop
            if (tid < B) {
                double l = 0.0;
                for (int j = 0; j < d3; j++) l = __dsub_rn(l, z3[tid * d3 + j]);
                misc[tid] = l;
            }
            __syncthreads();
            if (tid == 0) {
                double tot = 0.0;
                for (int b = 0; b < B; b++) tot = __dadd_rn(tot, misc[b]);
                p.loss_out[it] = tot;
            }
        }
        K1_MARK(11)
        // delta2 = (W2 delta3) (.) f'(z2), ascending k like S:274-277
        for (int b = warp; b < B; b += NW)
            for (int jb = lane; jb < d2; jb += 128) {  // four columns per lane: four independent addition chains
                double s4[4] = {0.0, 0.0, 0.0, 0.0};
                for (int k = 0; k < d3; k++) {
                    const double dk = dl3[b * d3 + k];
#pragma unroll
                    for (int q = 0; q < 4; q++) {
                        const int j = jb + 32 * q;
                        if (j < d2) s4[q] = __dadd_rn(s4[q], __dmul_rn(w2t[k * d2 + j], dk));
                    }
                }
#pragma unroll
                for (int q = 0; q < 4; q++) {
                    const int j = jb + 32 * q;
                    if (j < d2) dl2[b * d2 + j] = __dmul_rn(s4[q], dl2[b * d2 + j]);
                }
            }
        __syncthreads();

        K1_MARK(5)  // phase 3
        // ---- phase 4: delta1 for the CTA's own columns: (W1[own rows] delta2) (.) f'(z1 own) ----
        for (int e = tid; e < B * nc0; e += NT) {
            const int b = e / nc0, jj = e - b * nc0;
            const double *wr = w1r + (size_t)jj * d2, *dr = dl2 + (size_t)b * d2;
            double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
            int k = 0;
            for (; k + 3 < d2; k += 4) {
                s0 = fma(wr[k], dr[k], s0); s1 = fma(wr[k + 1], dr[k + 1], s1);
                s2 = fma(wr[k + 2], dr[k + 2], s2); s3 = fma(wr[k + 3], dr[k + 3], s3);
            }
            for (; k < d2; k++) s0 = fma(wr[k], dr[k], s0);
            d1o[b * ncm + jj] = __dmul_rn((s0 + s1) + (s2 + s3), act_prime(IACT, p.inner_p, z1o[b * ncm + jj]));
        }
        __syncthreads();

        K1_MARK(6)  // phase 4
        // ---- phase 5: gradients (batch order, un-fused multiply/add like S:281 + S:309-317) and updates ----
        // W0[:, own]: g = sum_b delta1[b][jj] * a0[b][k]
        // two rows k per thread in flight (their velocity loads from L2 overlap); per element the order of the
        // additions is unchanged
        for (int kA = tid; kA < d0; kA += 2 * NT) {
            const int kB = kA + NT;
            const bool hasB = kB < d0;
            double vpA[FUSED_CMAX], vpB[FUSED_CMAX], gA[FUSED_CMAX], gB[FUSED_CMAX];
#pragma unroll
            for (int jj = 0; jj < FUSED_CMAX; jj++) {
                gA[jj] = 0.0; gB[jj] = 0.0;
                vpA[jj] = (jj < nc0) ? p.V0t[(size_t)(j0 + jj) * d0 + kA] : 0.0;
                vpB[jj] = (jj < nc0 && hasB) ? p.V0t[(size_t)(j0 + jj) * d0 + kB] : 0.0;
            }
            for (int b = 0; b < B; b++) {
                const double aA = a0[(size_t)b * d0 + kA];
                const double aB = hasB ? a0[(size_t)b * d0 + kB] : 0.0;
#pragma unroll
                for (int jj = 0; jj < FUSED_CMAX; jj++)
                    if (jj < nc0) {  // CTA-uniform: no FP64 work for columns this CTA does not own
                        const double dlt = d1o[b * ncm + jj];
                        gA[jj] = __dadd_rn(gA[jj], __dmul_rn(dlt, aA));
                        gB[jj] = __dadd_rn(gB[jj], __dmul_rn(dlt, aB));
                    }
            }
#pragma unroll
            for (int jj = 0; jj < FUSED_CMAX; jj++)
                if (jj < nc0) {
                    const double adjA = sgd_adj(p.step, gA[jj], B, invB, p.momentum, vpA[jj]);
                    w0[jj * d0 + kA] = __dsub_rn(w0[jj * d0 + kA], adjA);
                    p.V0t[(size_t)(j0 + jj) * d0 + kA] = adjA;
                    if (hasB) {
                        const double adjB = sgd_adj(p.step, gB[jj], B, invB, p.momentum, vpB[jj]);
                        w0[jj * d0 + kB] = __dsub_rn(w0[jj * d0 + kB], adjB);
                        p.V0t[(size_t)(j0 + jj) * d0 + kB] = adjB;
                    }
                }
        }
        K1_MARK(12)
        if (p.bulk && !u8) {
            // a0 is no longer needed: fetch the next batch's FP64 rows into it while W1 / W2 are updated
            __syncthreads();
            if (warp == 0 && it + 1 < p.iters) issue_batch_fetch(p, it + 1, a0, mbar, lane);
        }
        // (phase 5b, the W1 / W2 updates, runs under the barrier waits of the next step: see update_w1_columns)
        __syncthreads();
        K1_MARK(7)  // phase 5
    }

    if (p.iters > 0) {
        update_w1_columns();
        update_w1_rows_w2(a1o_buf + (size_t)((p.iters - 1) & 1) * B * ncm);
        __syncthreads();
    }
    if (p.prof != nullptr && c == 0 && tid == 0)
        for (int i = 0; i < 16; i++) p.prof[i] = profs[i];
    // ---- write the CTA's slices back to the canonical [in][out] state ----
    for (int jj = 0; jj < nc0; jj++)
        for (int k = tid; k < d0; k += NT) {
            p.W0[(size_t)k * d1 + j0 + jj] = w0[jj * d0 + k];
            p.V0[(size_t)k * d1 + j0 + jj] = p.V0t[(size_t)(j0 + jj) * d0 + k];
        }
    for (int kk = 0; kk < nc1; kk++)
        for (int j = tid; j < d1; j += NT) p.W1[(size_t)j * d2 + k0 + kk] = w1c[kk * d1 + j];  // V1 updated in place
    if (c == 0)
        for (int e = tid; e < d2 * d3; e += NT) { p.W2[e] = w2[e]; p.V2[e] = v2[e]; }
}

}  // namespace f64
}  // namespace b2nn
