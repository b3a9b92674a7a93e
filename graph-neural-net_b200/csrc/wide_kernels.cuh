// wide_kernels.cuh -- the HBM-bound kernels around the tcgen05 GEMMs of the BF16 path:
// input prologue (K8), fused softmax-cross-entropy gradient (K5) / general output delta (K6),
// momentum-SGD update for the data-parallel case (K7) and the evaluation row pass.
// All of them read each input once and write each output once, row-major, with 16-byte accesses
// where the shapes allow (the GEMMs read either major-ness, so no transposed copies are written).
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

// IDX pixel bytes: value = (raw & 0xff) / 255.0 (MNISTTrainer.java:97); the IEEE float quotient equals the double
// quotient rounded to float for all 256 bytes (checked in tests/test_capi_cpu.py)
template <>
__device__ __forceinline__ float load_as_float<uint8_t>(const uint8_t *p) { return __fdiv_rn((float)*p, 255.0f); }

// eight consecutive elements with the widest aligned load (one request per thread instead of eight: when X is
// pinned host memory every request is a multi-microsecond PCIe read that occupies the SM's load path)
template <typename T>
__device__ __forceinline__ void load8_as_float(const T *x, float (&v)[8]) {
#pragma unroll
    for (int e = 0; e < 8; e++) v[e] = load_as_float<T>(x + e);
}
template <>
__device__ __forceinline__ void load8_as_float<__nv_bfloat16>(const __nv_bfloat16 *x, float (&v)[8]) {
    if ((reinterpret_cast<uintptr_t>(x) & 15) == 0) {
        const uint4 u = __ldg(reinterpret_cast<const uint4 *>(x));
        const __nv_bfloat162 *h = reinterpret_cast<const __nv_bfloat162 *>(&u);
#pragma unroll
        for (int e = 0; e < 4; e++) {
            const float2 f = __bfloat1622float2(h[e]);
            v[2 * e] = f.x; v[2 * e + 1] = f.y;
        }
    } else {
#pragma unroll
        for (int e = 0; e < 8; e++) v[e] = __bfloat162float(x[e]);
    }
}
template <>
__device__ __forceinline__ void load8_as_float<float>(const float *x, float (&v)[8]) {
    if ((reinterpret_cast<uintptr_t>(x) & 15) == 0) {
        const float4 a = __ldg(reinterpret_cast<const float4 *>(x)), b = __ldg(reinterpret_cast<const float4 *>(x) + 1);
        v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
    } else {
#pragma unroll
        for (int e = 0; e < 8; e++) v[e] = x[e];
    }
}
template <>
__device__ __forceinline__ void load8_as_float<uint8_t>(const uint8_t *x, float (&v)[8]) {
    if ((reinterpret_cast<uintptr_t>(x) & 7) == 0) {
        const uint2 u = __ldg(reinterpret_cast<const uint2 *>(x));
#pragma unroll
        for (int e = 0; e < 4; e++) {
            v[e] = __fdiv_rn((float)((u.x >> (8 * e)) & 0xffu), 255.0f);
            v[4 + e] = __fdiv_rn((float)((u.y >> (8 * e)) & 0xffu), 255.0f);
        }
    } else {
#pragma unroll
        for (int e = 0; e < 8; e++) v[e] = load_as_float<uint8_t>(x + e);
    }
}

// operand-type stores: BF16 nets keep activations / deltas as BF16, TF32 nets as FP32
__device__ __forceinline__ void store8(__nv_bfloat16 *o, const float (&v)[8]) {
    uint4 u;
    u.x = pack_bf16x2(v[0], v[1]); u.y = pack_bf16x2(v[2], v[3]);
    u.z = pack_bf16x2(v[4], v[5]); u.w = pack_bf16x2(v[6], v[7]);
    *reinterpret_cast<uint4 *>(o) = u;
}
__device__ __forceinline__ void store8(float *o, const float (&v)[8]) {
    reinterpret_cast<float4 *>(o)[0] = make_float4(v[0], v[1], v[2], v[3]);
    reinterpret_cast<float4 *>(o)[1] = make_float4(v[4], v[5], v[6], v[7]);
}
__device__ __forceinline__ void store4(__nv_bfloat16 *o, float a, float b, float c, float d) {
    *reinterpret_cast<uint2 *>(o) = make_uint2(pack_bf16x2(a, b), pack_bf16x2(c, d));
}
__device__ __forceinline__ void store4(float *o, float a, float b, float c, float d) {
    *reinterpret_cast<float4 *>(o) = make_float4(a, b, c, d);
}
__device__ __forceinline__ void store2(__nv_bfloat16 *o, float a, float b) {
    *reinterpret_cast<uint32_t *>(o) = pack_bf16x2(a, b);
}
__device__ __forceinline__ void store2(float *o, float a, float b) { *reinterpret_cast<float2 *>(o) = make_float2(a, b); }
__device__ __forceinline__ void store1(__nv_bfloat16 *o, float a) { *o = __float2bfloat16_rn(a); }
__device__ __forceinline__ void store1(float *o, float a) { *o = a; }

// a0 = f(x) (the reference activates the INPUT layer too, SoftmaxCrossEntropyNeuralNet.java:183-186)
// for the rows named by idx (or rows 0..B-1), cast to the operand type O (BF16, or FP32 for TF32 nets), stored [B][ld_a].
// X may live in HBM or in mapped pinned host memory (then this kernel IS the host->device copy).
// One thread per 8 consecutive elements of a row (one 16-byte store); grid-stride over B * ceil(d0/8).
template <typename T, typename O = __nv_bfloat16>
__global__ void __launch_bounds__(256)
input_prologue(const T *__restrict__ X, const int32_t *__restrict__ idx, int B, int d0, int act, float act_param,
               O *__restrict__ a, int ld_a) {
    const int groups = (d0 + 7) >> 3;
    const long long total = (long long)B * groups;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total;
         t += (long long)gridDim.x * blockDim.x) {
        const int b = (int)(t / groups), i0 = (int)(t % groups) * 8;
        const long long src = idx ? idx[b] : b;
        const T *x = X + src * d0 + i0;
        O *o = a + (size_t)b * ld_a + i0;
        if (i0 + 8 <= d0) {
            float v[8];
            load8_as_float<T>(x, v);
#pragma unroll
            for (int e = 0; e < 8; e++) v[e] = act_apply(act, act_param, v[e]);
            store8(o, v);
        } else {
            for (int e = 0; i0 + e < d0; e++) store1(o + e, act_apply(act, act_param, load_as_float<T>(x + e)));
        }
    }
}

// Pixel-byte rows with d0 % 16 == 0: 16 bytes per thread per request.  Out of pinned host memory the gather is bound
// by the round trips each thread makes one after the other, not by bytes, so twice the elements per request halve its
// duration (and the time it spends beside the GEMMs).
__global__ void __launch_bounds__(256)
input_prologue_u8x16(const uint8_t *__restrict__ X, const int32_t *__restrict__ idx, int B, int d0, int act,
                     float act_param, __nv_bfloat16 *__restrict__ a, int ld_a) {
    const int groups = d0 >> 4;
    const long long total = (long long)B * groups;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total;
         t += (long long)gridDim.x * blockDim.x) {
        const int b = (int)(t / groups), i0 = (int)(t % groups) * 16;
        const long long src = idx ? idx[b] : b;
        const uint4 raw = __ldg(reinterpret_cast<const uint4 *>(X + src * d0 + i0));
        const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
        uint32_t o[8];
#pragma unroll
        for (int q = 0; q < 4; q++) {
            float v[4];
#pragma unroll
            for (int e = 0; e < 4; e++)
                v[e] = act_apply(act, act_param, __fdiv_rn((float)((w[q] >> (8 * e)) & 0xffu), 255.0f));
            o[2 * q] = pack_bf16x2(v[0], v[1]);
            o[2 * q + 1] = pack_bf16x2(v[2], v[3]);
        }
        uint4 *dst = reinterpret_cast<uint4 *>(a + (size_t)b * ld_a + i0);
        dst[0] = make_uint4(o[0], o[1], o[2], o[3]);
        dst[1] = make_uint4(o[4], o[5], o[6], o[7]);
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

// Output layer of the BF16 path.  One warp per batch row (8 rows per block); the row of logits is read
// from HBM once and re-read from L1.
//  kind 0: p = softmax(z) (max-subtracted: identical maths, SoftmaxCrossEntropyNeuralNet.java:357-376),
//          loss = -sum t log p (:213-217), delta = p - t (:249-251); t one-hot from labels or dense.
//  kind 1: a = f_last(z); loss = sum L(a,t); delta = L'(a,t) f_last'(z) (GeneralNeuralNet.java:236-239,268-271).
// Writes delta in the operand type O (BF16, or FP32 for TF32 nets) d[B][ld_d]; row_loss[B] FP32.
template <typename O>
__global__ void __launch_bounds__(256)
output_delta(const float *__restrict__ Z, int ld_z, const int32_t *__restrict__ labels,
             const float *__restrict__ tdense, int ld_t, int B, int dL, int kind, int last_act, float last_param,
             O *__restrict__ d, int ld_d, float *__restrict__ row_loss, double *loss_slot,
             double *__restrict__ block_partial, unsigned int *block_counter) {
    __shared__ float s_row[8];
    __shared__ bool s_last;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int b = blockIdx.x * (blockDim.x >> 5) + warp;
    float lacc = 0.f;
    if (b < B) {
    const float *z = Z + (size_t)b * ld_z;
    const float *t = tdense ? tdense + (size_t)b * ld_t : nullptr;
    const int lab = labels ? labels[b] : -1;
    O *drow = d + (size_t)b * ld_d;
    float mx = 0.f, inv = 1.f, lse = 0.f;
    if (kind == 0 && t == nullptr && dL <= 1024 && ((ld_z | ld_d) & 3) == 0) {
        // softmax / cross-entropy with labels, row held in registers: ONE pass over global memory, 16-byte loads
        // and 8-byte stores, all of a lane's loads in flight together (the three-pass loop below was latency-bound:
        // 16 us for 4096 x 1000 against 4 us of HBM time)
        float4 x[8];
#pragma unroll
        for (int i = 0; i < 8; i++) {
            const int c = 128 * i + 4 * lane;
            x[i] = make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY);
            if (c + 4 <= dL) x[i] = *reinterpret_cast<const float4 *>(z + c);
            else if (c < dL) {
                x[i].x = z[c];
                if (c + 1 < dL) x[i].y = z[c + 1];
                if (c + 2 < dL) x[i].z = z[c + 2];
            }
        }
        mx = -INFINITY;
#pragma unroll
        for (int i = 0; i < 8; i++) mx = fmaxf(mx, fmaxf(fmaxf(x[i].x, x[i].y), fmaxf(x[i].z, x[i].w)));
        mx = warp_max(mx);
        float sum = 0.f;
#pragma unroll
        for (int i = 0; i < 8; i++) {  // exp(-inf) = 0 for the masked columns
            x[i].x = __expf(x[i].x - mx); x[i].y = __expf(x[i].y - mx);
            x[i].z = __expf(x[i].z - mx); x[i].w = __expf(x[i].w - mx);
            sum += (x[i].x + x[i].y) + (x[i].z + x[i].w);
        }
        sum = warp_sum(sum);
        inv = 1.f / sum;
        lse = __logf(sum);
        if (lab >= 0 && lab < dL && (lab >> 2 & 31) == lane) lacc = lse - (z[lab] - mx);  // -log p[label]
#pragma unroll
        for (int i = 0; i < 8; i++) {
            const int c = 128 * i + 4 * lane;
            if (c >= dL) continue;
            float p0 = x[i].x * inv, p1 = x[i].y * inv, p2 = x[i].z * inv, p3 = x[i].w * inv;
            if (lab == c) p0 -= 1.f;
            if (lab == c + 1) p1 -= 1.f;
            if (lab == c + 2) p2 -= 1.f;
            if (lab == c + 3) p3 -= 1.f;
            if (c + 4 <= dL) store4(drow + c, p0, p1, p2, p3);
            else {
                store1(drow + c, p0);
                if (c + 1 < dL) store1(drow + c + 1, p1);
                if (c + 2 < dL) store1(drow + c + 2, p2);
            }
        }
        lacc = warp_sum(lacc);
        if (lane == 0 && row_loss) row_loss[b] = lacc;
    } else {
    if (kind == 0) {
        mx = -INFINITY;
        for (int c = lane; c < dL; c += 32) mx = fmaxf(mx, z[c]);
        mx = warp_max(mx);
        float sum = 0.f;
        for (int c = lane; c < dL; c += 32) sum += __expf(z[c] - mx);
        sum = warp_sum(sum);
        inv = 1.f / sum;
        lse = __logf(sum);
    }
    // two columns per lane per step -> 4-byte stores, 128 B per warp
    for (int c0 = 2 * lane; c0 < dL; c0 += 64) {
        float dv[2] = {0.f, 0.f};
#pragma unroll
        for (int e = 0; e < 2; e++) {
            const int c = c0 + e;
            if (c < dL) {
                const float zc = z[c];
                const float tc = t ? t[c] : ((c == lab) ? 1.f : 0.f);
                if (kind == 0) {
                    const float zs = zc - mx;
                    dv[e] = __expf(zs) * inv - tc;
                    if (tc != 0.f) lacc -= tc * (zs - lse);  // -t log p
                } else {
                    const float a = act_apply(last_act, last_param, zc);
                    const float diff = a - tc;  // half squared error: L = 0.5 diff^2, L' = diff
                    lacc += 0.5f * diff * diff;
                    dv[e] = diff * act_prime_from_a(last_act, last_param, a);
                }
            }
        }
        if (c0 + 1 < dL) store2(drow + c0, dv[0], dv[1]);
        else store1(drow + c0, dv[0]);
    }
    lacc = warp_sum(lacc);
    if (lane == 0 && row_loss) row_loss[b] = lacc;
    }
    }
    // optional batch loss sum with a fixed summation order (rows within a block, then blocks in index order by
    // whichever block finishes last), so a run is reproducible bit for bit and needs no extra launch
    if (loss_slot == nullptr) return;
    if (lane == 0) s_row[warp] = lacc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double s = 0.0;
        for (int w = 0; w < (int)(blockDim.x >> 5); w++) s += (double)s_row[w];
        block_partial[blockIdx.x] = s;
        __threadfence();
        s_last = atomicAdd(block_counter, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (s_last && warp == 0) {
        __threadfence();
        double s = 0.0;
        for (int i = lane; i < (int)gridDim.x; i += 32) s += __ldcg(block_partial + i);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if (lane == 0) { *loss_slot += s; *block_counter = 0u; }
    }
}

// Evaluation row pass: one warp per row.  The reference compares the ACTIVATED outputs with `>=` starting from
// index 0 (MNISTTrainer.java:164-168: `if (out[i] >= out[actual]) actual = i`), so the LAST maximal index wins, a NaN
// entry never wins, and a NaN at index 0 pins the answer to 0.  Softmax is strictly monotone in z, so softmax nets
// compare the logits; general nets compare f_last(z) (ReLU / saturating sigmoid and tanh outputs tie where z does not).
__global__ void eval_rows(const float *__restrict__ Z, int ld_z, const int32_t *__restrict__ labels, int rows,
                          int dL, int kind, int last_act, float last_param, int32_t *__restrict__ pred,
                          float *__restrict__ row_loss) {
    const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (row >= rows) return;
    const float *z = Z + (size_t)row * ld_z;
    float best = -INFINITY, zbest = -INFINITY;
    int besti = -1;
    for (int c = lane; c < dL; c += 32) {
        const float zc = z[c];
        const float v = kind == 0 ? zc : act_apply(last_act, last_param, zc);
        if (v >= best) { best = v; besti = c; }   // false for NaN
        zbest = fmaxf(zbest, zc);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float ob = __shfl_xor_sync(0xffffffffu, best, o);
        const int oi = __shfl_xor_sync(0xffffffffu, besti, o);
        if (oi >= 0 && (besti < 0 || ob > best || (ob == best && oi > besti))) { best = ob; besti = oi; }
        zbest = fmaxf(zbest, __shfl_xor_sync(0xffffffffu, zbest, o));
    }
    {
        const float z0 = z[0];
        const float v0 = kind == 0 ? z0 : act_apply(last_act, last_param, z0);
        if (besti < 0 || v0 != v0) besti = 0;
    }
    float loss = 0.f;
    if (row_loss && labels) {
        const int lab = labels[row];
        if (kind == 0) {
            float sum = 0.f;
            for (int c = lane; c < dL; c += 32) sum += __expf(z[c] - zbest);
            sum = warp_sum(sum);
            loss = -(z[lab] - zbest - __logf(sum));
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

// NeuralNet.propagate (src/NeuralNet.java:16) output pass of the tensor-core path: the FP32 logits of `rows` rows ->
// the net's output vector in double, on the device: softmax (S:357-376; max-subtracted, mathematically identical) or
// the last activation (G:214-218).  One warp per row; exp / sum / divide in double like the reference.
__global__ void propagate_rows(const float *__restrict__ Z, int ld_z, int rows, int dL, int kind, int last_act,
                               double last_param, double *__restrict__ P) {
    const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (row >= rows) return;
    const float *z = Z + (size_t)row * ld_z;
    double *p = P + (size_t)row * dL;
    if (kind == 0) {
        float mx = -INFINITY;
        for (int c = lane; c < dL; c += 32) mx = fmaxf(mx, z[c]);
        mx = warp_max(mx);
        double sum = 0.0;
        for (int c = lane; c < dL; c += 32) sum += exp((double)z[c] - (double)mx);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
        for (int c = lane; c < dL; c += 32) p[c] = exp((double)z[c] - (double)mx) / sum;
    } else {
        for (int c = lane; c < dL; c += 32) {
            const double v = (double)z[c];
            double a;
            switch (last_act) {
            case ACT_LEAKY_RELU: a = v > 0 ? v : last_param * v; break;
            case ACT_SIGMOID: a = 1.0 / (1.0 + exp(-v)); break;
            case ACT_RELU: a = v > 0 ? v : 0.0; break;
            case ACT_TANH: a = tanh(v); break;
            default: a = v;
            }
            p[c] = a;
        }
    }
}

// Momentum-SGD update (SoftmaxCrossEntropyNeuralNet.java:325-342) over the padded [in][ld_n] arrays as one
// flat stream of float4: v = scale*g + mu*v ; w -= v ; emit the BF16 operand copy wn.  Algorithmic traffic
// 12 B read + 10 B written per parameter.  cast_only: just (re)build wn from the master weights.
// wn == nullptr (TF32 nets: the FP32 master array IS the GEMM operand): no copy is written, 12 B + 8 B.
__global__ void __launch_bounds__(256)
sgd_update(const float4 *__restrict__ g, float4 *__restrict__ w, float4 *__restrict__ v, long long n4, float scale,
           float momentum, int cast_only, uint2 *__restrict__ wn) {
    for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < n4;
         q += (long long)gridDim.x * blockDim.x) {
        float4 wv = w[q];
        if (!cast_only) {
            const float4 gv = __ldcs(g + q);  // the gradient is dead after this read
            float4 vv = v[q];
            vv.x = scale * gv.x + momentum * vv.x; vv.y = scale * gv.y + momentum * vv.y;
            vv.z = scale * gv.z + momentum * vv.z; vv.w = scale * gv.w + momentum * vv.w;
            wv.x -= vv.x; wv.y -= vv.y; wv.z -= vv.z; wv.w -= vv.w;
            v[q] = vv;
            w[q] = wv;
        }
        if (wn) wn[q] = make_uint2(pack_bf16x2(wv.x, wv.y), pack_bf16x2(wv.z, wv.w));
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
