// f64_kernels.cuh -- FP64 "parity mode" kernels (general shapes).
//
// Arithmetic follows the reference's evaluation order so that results track the JVM's to the
// last bits: every dot product accumulates in ascending k with a separate multiply and add
// (__dmul_rn / __dadd_rn: no FMA contraction, like Java), the batch sum runs in batch order, and
// the update is ((step*g)/B) + (mu*prev) left to right (src/SoftmaxCrossEntropyNeuralNet.java:333).
// Only exp()/log() differ from the JVM (each <= 1 ulp).
#pragma once
#include <cstdint>

namespace b2nn {
namespace f64 {

__device__ __forceinline__ double act(int kind, double p, double z) {
    switch (kind) {
    case 0: return (z > 0) ? z : __dmul_rn(p, z);                  // MNISTTrainer.java:234
    case 1: return __ddiv_rn(1.0, __dadd_rn(1.0, exp(-z)));        // sigmoid, doc/backprop.pdf
    case 2: return (z > 0) ? z : 0.0;
    case 3: return tanh(z);
    default: return z;
    }
}
__device__ __forceinline__ double act_prime(int kind, double p, double z) {
    switch (kind) {
    case 0: return (z <= 0.0) ? p : 1.0;                           // MNISTTrainer.java:235
    case 1: { double s = __ddiv_rn(1.0, __dadd_rn(1.0, exp(-z))); return __dmul_rn(s, __dsub_rn(1.0, s)); }
    case 2: return (z <= 0.0) ? 0.0 : 1.0;
    case 3: { double t = tanh(z); return __dsub_rn(1.0, __dmul_rn(t, t)); }
    default: return 1.0;
    }
}
__device__ __forceinline__ double loss_fn(int, double c, double e) {
    double d = __dsub_rn(c, e);
    return __dmul_rn(0.5, __dmul_rn(d, d));
}
__device__ __forceinline__ double loss_prime(int, double c, double e) { return __dsub_rn(c, e); }

enum { EPI_FWD = 0, EPI_DX = 1, EPI_DW_UPDATE = 2, EPI_STORE = 3 };

struct GemmArgs {
    int M, N, K;
    const double *A; long long sam, sak;  // A(m,k) = A[m*sam + k*sak]
    const double *B; long long sbk, sbn;  // B(k,n) = B[k*sbk + n*sbn]
    double *C; int ldc;                   // primary output [M][ldc]
    double *C2; int ldc2;                 // EPI_FWD: activation output (may be null)
    const double *Z; int ldz;             // EPI_DX: pre-activations for f'
    double *V;                            // EPI_DW_UPDATE: velocity (same layout as C = weights)
    int act; double act_param;
    double step, momentum; int batch;
};

// C = A*B with strict ascending-k accumulation.  Block 16x16 threads; thread (ty,tx) owns rows
// ty*TMT..+TMT and cols tx*4..+4 of a (16*TMT) x 64 tile.
template <int TMT, int EPI>
__global__ void __launch_bounds__(256) gemm_strict(const GemmArgs g) {
    constexpr int TM = 16 * TMT, TN = 64, BK = 16;
    __shared__ double As[BK][TM + 1];
    __shared__ double Bs[BK][TN + 1];
    const int tid = threadIdx.x;
    const int tx = tid & 15, ty = tid >> 4;
    const int m0 = blockIdx.y * TM, n0 = blockIdx.x * TN;
    double acc[TMT][4];
#pragma unroll
    for (int i = 0; i < TMT; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) acc[i][j] = 0.0;

    const bool a_kfast = (g.sak == 1);
    const bool b_nfast = (g.sbn == 1);
    for (int k0 = 0; k0 < g.K; k0 += BK) {
        for (int e = tid; e < TM * BK; e += 256) {
            int m, k;
            if (a_kfast) { k = e % BK; m = e / BK; } else { m = e % TM; k = e / TM; }
            int gm = m0 + m, gk = k0 + k;
            As[k][m] = (gm < g.M && gk < g.K) ? g.A[gm * g.sam + gk * g.sak] : 0.0;
        }
        for (int e = tid; e < TN * BK; e += 256) {
            int n, k;
            if (b_nfast) { n = e % TN; k = e / TN; } else { k = e % BK; n = e / BK; }
            int gn = n0 + n, gk = k0 + k;
            Bs[k][n] = (gn < g.N && gk < g.K) ? g.B[gk * g.sbk + gn * g.sbn] : 0.0;
        }
        __syncthreads();
        const int kmax = min(BK, g.K - k0);
        for (int k = 0; k < kmax; k++) {
            double b[4];
#pragma unroll
            for (int j = 0; j < 4; j++) b[j] = Bs[k][tx * 4 + j];
#pragma unroll
            for (int i = 0; i < TMT; i++) {
                double a = As[k][ty * TMT + i];
#pragma unroll
                for (int j = 0; j < 4; j++) acc[i][j] = __dadd_rn(acc[i][j], __dmul_rn(a, b[j]));
            }
        }
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < TMT; i++) {
        const int m = m0 + ty * TMT + i;
        if (m >= g.M) continue;
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const int n = n0 + tx * 4 + j;
            if (n >= g.N) continue;
            const double s = acc[i][j];
            if constexpr (EPI == EPI_FWD) {
                g.C[(size_t)m * g.ldc + n] = s;                                       // z_l
                if (g.C2) g.C2[(size_t)m * g.ldc2 + n] = act(g.act, g.act_param, s);  // a_l = f(z_l)
            } else if constexpr (EPI == EPI_DX) {
                g.C[(size_t)m * g.ldc + n] =
                    __dmul_rn(s, act_prime(g.act, g.act_param, g.Z[(size_t)m * g.ldz + n]));  // S:277
            } else if constexpr (EPI == EPI_DW_UPDATE) {
                const size_t o = (size_t)m * g.ldc + n;
                const double adj = __dadd_rn(__ddiv_rn(__dmul_rn(g.step, s), (double)g.batch),
                                             __dmul_rn(g.momentum, g.V[o]));          // S:333
                g.C[o] = __dsub_rn(g.C[o], adj);                                      // S:338
                g.V[o] = adj;                                                         // S:339,344
            } else {
                g.C[(size_t)m * g.ldc + n] = s;
            }
        }
    }
}

// a0 = f(x) for gathered rows (the reference applies the inner activation to the input layer
// too, S:183-186 with l-1 == 0), and the matching target rows.
static __global__ void gather_rows(const double *X, const double *T, const int32_t *labels, const int32_t *idx, int rows,
                            int d0, int dL, int act_kind, double act_param, double *a0, double *tout) {
    const int r = blockIdx.x;
    if (r >= rows) return;
    const long long src = idx ? idx[r] : r;
    for (int i = threadIdx.x; i < d0; i += blockDim.x)
        a0[(size_t)r * d0 + i] = act(act_kind, act_param, X[src * d0 + i]);
    if (tout) {
        for (int i = threadIdx.x; i < dL; i += blockDim.x)
            tout[(size_t)r * dL + i] = T ? T[src * dL + i] : ((labels[src] == i) ? 1.0 : 0.0);
    }
}

// Output layer, one thread per row (d_L is small on this path; order kept sequential):
//  softmax (S:357-376, no max subtraction unless `stable`), loss (S:213-217 / G:236-239),
//  delta (S:249-251 / G:268-271), `>=` argmax (MNISTTrainer.java:166-168).
static __global__ void output_rows(const double *Z, const double *T, int rows, int dL, int kind, int last_act,
                            double last_param, int loss_kind, int stable, double *P, double *loss, double *delta,
                            int32_t *pred) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= rows) return;
    const double *z = Z + (size_t)r * dL;
    const double *t = T ? T + (size_t)r * dL : nullptr;
    double l = 0.0;
    int best = 0;
    double bestv = 0.0;
    if (kind == 0) {  // softmax + cross entropy
        double mx = 0.0;
        if (stable) {
            mx = z[0];
            for (int i = 1; i < dL; i++) mx = fmax(mx, z[i]);
        }
        double sum = 0.0;
        for (int i = 0; i < dL; i++) sum = __dadd_rn(sum, exp(z[i] - mx));
        for (int i = 0; i < dL; i++) {
            const double p = __ddiv_rn(exp(z[i] - mx), sum);
            if (P) P[(size_t)r * dL + i] = p;
            if (t) {
                l = __dsub_rn(l, __dmul_rn(t[i], log(p)));
                if (delta) delta[(size_t)r * dL + i] = __dsub_rn(p, t[i]);
            }
            if (i == 0 || p >= bestv) { best = i; bestv = p; }
        }
    } else {
        for (int i = 0; i < dL; i++) {
            const double a = act(last_act, last_param, z[i]);
            if (P) P[(size_t)r * dL + i] = a;
            if (t) {
                l = __dadd_rn(l, loss_fn(loss_kind, a, t[i]));
                if (delta)
                    delta[(size_t)r * dL + i] =
                        __dmul_rn(loss_prime(loss_kind, a, t[i]), act_prime(last_act, last_param, z[i]));
            }
            if (i == 0 || a >= bestv) { best = i; bestv = a; }
        }
    }
    if (loss) loss[r] = l;
    if (pred) pred[r] = best;
}

static __global__ void iota_rows(int32_t *idx, int rows) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r < rows) idx[r] = r;
}

// Row-order sum of the per-row losses and the hit count (T:108-111 adds the losses in key order; the order is kept so the
// sum has the reference's bits).  One block: all threads fetch a slab of 256 losses into shared memory at once (one L2
// round trip per slab instead of one per row), thread 0 adds the slab in ascending row order; hits are counted by all
// threads (integer, order-free).
static __global__ void __launch_bounds__(256) reduce_rows(const double *loss, const int32_t *pred, const int32_t *labels,
                                                          int rows, double *loss_sum, long long *hits) {
    __shared__ double slab[256];
    __shared__ int s_hits[256];
    const int tid = threadIdx.x;
    double s = 0.0;
    int h = 0;
    for (int base = 0; base < rows; base += 256) {
        const int r = base + tid;
        if (loss) slab[tid] = r < rows ? loss[r] : 0.0;
        if (pred && labels && r < rows && pred[r] == labels[r]) h++;
        __syncthreads();
        if (loss && tid == 0) {
            const int n = min(256, rows - base);
            for (int i = 0; i < n; i++) s = __dadd_rn(s, slab[i]);
        }
        __syncthreads();
    }
    s_hits[tid] = h;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if (tid < o) s_hits[tid] += s_hits[tid + o];
        __syncthreads();
    }
    if (tid == 0) {
        if (loss_sum) *loss_sum = __dadd_rn(*loss_sum, s);
        if (hits) *hits += s_hits[0];
    }
}

}  // namespace f64
}  // namespace b2nn
