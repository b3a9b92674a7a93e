// wide_kernels.cuh -- the HBM-bound kernels around the tcgen05 GEMMs of the BF16 path:
// input prologue (K8), fused softmax-cross-entropy gradient (K5) / general output delta (K6),
// momentum-SGD update for the data-parallel case (K7) and the evaluation row pass.
// All of them read each input once, write row-major AND transposed BF16 copies through a
// 32x32 shared-memory tile (both global accesses coalesced), and are sized by the caller as
// plain grids over 32-row / 32x32 tiles.
#pragma once
#include <cuda_bf16.h>

#include <cstdint>

#include "gemm_tcgen05.cuh"

namespace b2nn {
namespace wide {

template <typename T>
__device__ __forceinline__ float load_as_float(const T *p);
template <>
__device__ __forceinline__ float load_as_float<double>(const double *p) { return (float)*p; }
template <>
__device__ __forceinline__ float load_as_float<float>(const float *p) { return *p; }
template <>
__device__ __forceinline__ float load_as_float<__nv_bfloat16>(const __nv_bfloat16 *p) { return __bfloat162float(*p); }

// a0 = f(x) (the reference activates the INPUT layer too, SoftmaxCrossEntropyNeuralNet.java:183-186)
// for the rows named by idx (or rows 0..B-1), cast to BF16, stored [B][ld_a] and transposed [d0][ldB].
// X may live in HBM or in mapped pinned host memory (then this kernel IS the host->device copy).
// grid (ceil(d0/32), ceil(B/32)), block (32, 8)
template <typename T>
__global__ void input_prologue(const T *__restrict__ X, const int32_t *__restrict__ idx, int B, int d0, int act,
                               float act_param, __nv_bfloat16 *__restrict__ a, int ld_a,
                               __nv_bfloat16 *__restrict__ at, int ldB) {
    __shared__ float tile[32][33];
    const int i0 = blockIdx.x * 32, b0 = blockIdx.y * 32;
    const int tx = threadIdx.x, ty = threadIdx.y;
#pragma unroll
    for (int r = ty; r < 32; r += 8) {
        const int b = b0 + r, i = i0 + tx;
        float v = 0.f;
        if (b < B && i < d0) {
            const long long src = idx ? idx[b] : b;
            v = act_apply(act, act_param, load_as_float<T>(X + src * d0 + i));
            a[(size_t)b * ld_a + i] = __float2bfloat16_rn(v);
        }
        tile[r][tx] = v;
    }
    __syncthreads();
#pragma unroll
    for (int r = ty; r < 32; r += 8) {
        const int i = i0 + r, b = b0 + tx;
        if (i < d0 && b < B) at[(size_t)i * ldB + b] = __float2bfloat16_rn(tile[tx][r]);
    }
}

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Output layer of the BF16 path.  One block = 32 batch rows, 256 threads (8 warps x 4 rows).
//  kind 0: p = softmax(z) (max-subtracted: identical maths, SoftmaxCrossEntropyNeuralNet.java:357-376),
//          loss = -sum t log p (:213-217), delta = p - t (:249-251); t one-hot from labels or dense.
//  kind 1: a = f_last(z); loss = sum L(a,t); delta = L'(a,t) f_last'(z) (GeneralNeuralNet.java:236-239,268-271).
// Writes delta as BF16 row-major d[B][ld_d] and transposed dt[dL][ldB]; row_loss[B] FP32.
__global__ void __launch_bounds__(256)
output_delta(const float *__restrict__ Z, int ld_z, const int32_t *__restrict__ labels,
             const float *__restrict__ tdense, int ld_t, int B, int dL, int kind, int last_act, float last_param,
             __nv_bfloat16 *__restrict__ d, int ld_d, __nv_bfloat16 *__restrict__ dt, int ldB,
             float *__restrict__ row_loss) {
    __shared__ float s_max[32], s_inv[32], s_lse[32];
    __shared__ float tile[32][33];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int b0 = blockIdx.x * 32;
    if (kind == 0) {
        for (int rr = 0; rr < 4; rr++) {
            const int r = warp * 4 + rr, b = b0 + r;
            float mx = -INFINITY, sum = 0.f;
            if (b < B) {
                const float *z = Z + (size_t)b * ld_z;
                for (int c = lane; c < dL; c += 32) mx = fmaxf(mx, z[c]);
                mx = warp_max(mx);
                for (int c = lane; c < dL; c += 32) sum += __expf(z[c] - mx);
                sum = warp_sum(sum);
            }
            if (lane == 0) { s_max[r] = mx; s_inv[r] = 1.f / sum; s_lse[r] = __logf(sum); }
        }
    }
    __syncthreads();
    float lacc[4] = {0.f, 0.f, 0.f, 0.f};
    for (int c0 = 0; c0 < dL; c0 += 32) {
        const int c = c0 + lane;
#pragma unroll
        for (int rr = 0; rr < 4; rr++) {
            const int r = warp * 4 + rr, b = b0 + r;
            float dv = 0.f;
            if (b < B && c < dL) {
                const float z = Z[(size_t)b * ld_z + c];
                float t;
                if (tdense) t = tdense[(size_t)b * ld_t + c];
                else t = (labels[b] == c) ? 1.f : 0.f;
                if (kind == 0) {
                    const float zs = z - s_max[r];
                    const float p = __expf(zs) * s_inv[r];
                    dv = p - t;
                    if (t != 0.f) lacc[rr] -= t * (zs - s_lse[r]);  // -t log p
                } else {
                    const float a = act_apply(last_act, last_param, z);
                    const float diff = a - t;  // half squared error: L = 0.5 diff^2, L' = diff
                    lacc[rr] += 0.5f * diff * diff;
                    dv = diff * act_prime_from_a(last_act, last_param, a);
                }
                d[(size_t)b * ld_d + c] = __float2bfloat16_rn(dv);
            }
            tile[r][lane] = dv;
        }
        __syncthreads();
#pragma unroll
        for (int rr = 0; rr < 4; rr++) {
            const int cc = c0 + warp * 4 + rr, b = b0 + lane;
            if (cc < dL && b < B) dt[(size_t)cc * ldB + b] = __float2bfloat16_rn(tile[lane][warp * 4 + rr]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int rr = 0; rr < 4; rr++) {
        const float l = warp_sum(lacc[rr]);
        const int b = b0 + warp * 4 + rr;
        if (lane == 0 && b < B && row_loss) row_loss[b] = l;
    }
}

// Evaluation row pass: one warp per row.  `>=` argmax so the LAST maximal index wins
// (MNISTTrainer.java:166-168) and the per-row loss.
__global__ void eval_rows(const float *__restrict__ Z, int ld_z, const int32_t *__restrict__ labels, int rows,
                          int dL, int kind, int last_act, float last_param, int32_t *__restrict__ pred,
                          float *__restrict__ row_loss) {
    const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (row >= rows) return;
    const float *z = Z + (size_t)row * ld_z;
    float best = -INFINITY;
    int besti = -1;
    for (int c = lane; c < dL; c += 32) {
        const float v = z[c];  // softmax / monotone last activations preserve the order of z
        if (v >= best) { best = v; besti = c; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float ob = __shfl_xor_sync(0xffffffffu, best, o);
        const int oi = __shfl_xor_sync(0xffffffffu, besti, o);
        if (ob > best || (ob == best && oi > besti)) { best = ob; besti = oi; }
    }
    float loss = 0.f;
    if (row_loss && labels) {
        const int lab = labels[row];
        if (kind == 0) {
            float sum = 0.f;
            for (int c = lane; c < dL; c += 32) sum += __expf(z[c] - best);
            sum = warp_sum(sum);
            loss = -(z[lab] - best - __logf(sum));
        } else {
            float acc = 0.f;
            for (int c = lane; c < dL; c += 32) {
                const float diff = act_apply(last_act, last_param, z[c]) - ((c == lab) ? 1.f : 0.f);
                acc += 0.5f * diff * diff;
            }
            loss = warp_sum(acc);
        }
    }
    if (lane == 0) {
        if (pred) pred[row] = besti;
        if (row_loss) row_loss[row] = loss;
    }
}

// Momentum-SGD update after the gradient all-reduce (SoftmaxCrossEntropyNeuralNet.java:325-342):
// v = scale*g + mu*v ; w -= v ; emit BF16 w row-major [in][ld_n] and transposed [out][ld_t].
// cast_only: just (re)build the BF16 copies from the master weights.
// grid (ceil(out/32), ceil(in/32)), block (32, 8)
__global__ void sgd_update(const float *__restrict__ g, float *__restrict__ w, float *__restrict__ v, int in,
                           int out, int ld_n, float scale, float momentum, int cast_only,
                           __nv_bfloat16 *__restrict__ wn, __nv_bfloat16 *__restrict__ wt, int ld_t) {
    __shared__ float tile[32][33];
    const int j0 = blockIdx.x * 32, i0 = blockIdx.y * 32;
    const int tx = threadIdx.x, ty = threadIdx.y;
#pragma unroll
    for (int r = ty; r < 32; r += 8) {
        const int i = i0 + r, j = j0 + tx;
        float wv = 0.f;
        if (i < in && j < out) {
            const size_t o = (size_t)i * ld_n + j;
            wv = w[o];
            if (!cast_only) {
                const float vn = scale * g[o] + momentum * v[o];
                wv -= vn;
                v[o] = vn;
                w[o] = wv;
            }
            wn[o] = __float2bfloat16_rn(wv);
        }
        tile[r][tx] = wv;
    }
    __syncthreads();
#pragma unroll
    for (int r = ty; r < 32; r += 8) {
        const int j = j0 + r, i = i0 + tx;
        if (j < out && i < in) wt[(size_t)j * ld_t + i] = __float2bfloat16_rn(tile[tx][r]);
    }
}

// row-order reduction of per-row FP32 losses / hit count into FP64 accumulators (one block)
__global__ void reduce_rows(const float *__restrict__ row_loss, const int32_t *__restrict__ pred,
                            const int32_t *__restrict__ labels, int rows, double *loss_sum, long long *hits) {
    __shared__ double s_l[256];
    __shared__ long long s_h[256];
    double l = 0.0;
    long long h = 0;
    for (int r = threadIdx.x; r < rows; r += blockDim.x) {
        if (row_loss) l += (double)row_loss[r];
        if (pred && labels && pred[r] == labels[r]) h++;
    }
    s_l[threadIdx.x] = l;
    s_h[threadIdx.x] = h;
    __syncthreads();
    for (int o = blockDim.x / 2; o > 0; o >>= 1) {
        if (threadIdx.x < o) { s_l[threadIdx.x] += s_l[threadIdx.x + o]; s_h[threadIdx.x] += s_h[threadIdx.x + o]; }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        if (loss_sum) *loss_sum += s_l[0];
        if (hits) *hits += s_h[0];
    }
}

__global__ void gather_labels(const int32_t *__restrict__ labels, const int32_t *__restrict__ idx, int B,
                              int32_t *__restrict__ out) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b < B) out[b] = labels[idx ? idx[b] : b];
}

// dense FP64 targets (rows picked by idx) -> FP32 [B][ld]
__global__ void gather_targets(const double *__restrict__ T, const int32_t *__restrict__ idx, int B, int dL,
                               float *__restrict__ out, int ld) {
    const int b = blockIdx.x;
    const long long src = idx ? idx[b] : b;
    for (int c = threadIdx.x; c < dL; c += blockDim.x) out[(size_t)b * ld + c] = (float)T[src * dL + c];
}

}  // namespace wide
}  // namespace b2nn
