// f64_fused.cu -- host side of K1, the persistent fused FP64 training kernel (f64_fused.cuh).
#include "f64_fused.cuh"

#include <cstdio>
#include <cstdlib>

#include "f64_path.hpp"

namespace b2nn {

namespace {
constexpr int kMaxBatch = 64;

struct Layout {
    f64::FusedParams p;
    size_t smem_bytes;
};

// carve shared memory with the same slice formulas the kernel uses
void make_layout(const b2nn_net *net, int B, int G, Layout *out) {
    f64::FusedParams &p = out->p;
    const int d0 = net->dims[0], d1 = net->dims[1], d2 = net->dims[2], d3 = net->dims[3];
    p.d0 = d0; p.d1 = d1; p.d2 = d2; p.d3 = d3; p.B = B;
    p.nc0max = (d1 + G - 1) / G;
    p.nc1max = (d2 + G - 1) / G;
    int o = 0;
    auto take = [&](int n) { int r = o; o += (n + 1) & ~1; return r; };
    p.o_a0 = take(B * d0);
    p.o_a1 = take(B * d1);
    p.o_a2 = take(B * d2);
    p.o_d2 = take(B * d2);
    p.o_w0 = take(p.nc0max * d0);
    p.o_w1c = take(p.nc1max * d1);
    p.o_w1r = take(p.nc0max * d2);
    p.o_w2 = take(d2 * d3);
    p.o_z1o = take(B * p.nc0max);
    p.o_d1o = take(B * p.nc0max);
    p.o_z3 = take(B * d3);
    p.o_d3 = take(B * d3);
    p.o_t = take(2 * B * d3);
    p.o_misc = take(B);
    p.o_a1o = take(2 * B * p.nc0max);   // double buffered by step parity
    p.o_part = take(16 * 4 * 4 + B * 4);  // KS*B <= max(NW*4, B) rows of 4
    p.o_w2t = take(d2 * d3);
    p.o_lut = take(256);
    p.o_stage = take((B * d0 + 7) / 8);  // one batch of rows as bytes
    p.o_mbar = take(2);
    p.o_prof = take(16);
    out->smem_bytes = (size_t)o * sizeof(double);
}
const void *fused_kernel_for(int act) {
    switch (act) {
    case 0: return (const void *)f64::fused_train_kernel<0>;
    case 1: return (const void *)f64::fused_train_kernel<1>;
    case 2: return (const void *)f64::fused_train_kernel<2>;
    case 3: return (const void *)f64::fused_train_kernel<3>;
    default: return (const void *)f64::fused_train_kernel<4>;
    }
}
}  // namespace

bool f64_fused_supported(b2nn_net *net, int B) {
    if (net->fused_f64_disabled || net->precision != B2NN_PREC_F64) return false;
    if (net->n_dims != 4 || B < 1 || B > kMaxBatch) return false;
    if ((net->dims[1] + net->num_sms - 1) / net->num_sms > 4 || (net->dims[2] + net->num_sms - 1) / net->num_sms > 4)
        return false;  // at most FUSED_CMAX columns per CTA
    if (!net->coop_launch) return false;
    Layout lo;
    make_layout(net, B, net->num_sms, &lo);
    return lo.smem_bytes <= (size_t)net->max_smem_optin;
}

int f64_fused_train(b2nn_net *net, const double *X, const uint8_t *Xu8, const double *T, const int32_t *labels,
                    const int32_t *idx_dev, int iters, int B, double step, double momentum, double *loss_dev) {
    const int G = net->num_sms;
    Layout lo;
    make_layout(net, B, G, &lo);
    f64::FusedParams &p = lo.p;
    const int d0 = p.d0, d1 = p.d1, d2 = p.d2, d3 = p.d3;
    auto &fs = net->fused;
    if (fs.cap_B < B) {
        B2_CHECK_CUDA(cudaStreamSynchronize(net->stream));
        cudaFree(fs.z1buf); cudaFree(fs.z2buf);
        fs.z1buf = fs.z2buf = nullptr;
        B2_CHECK_CUDA(cudaMalloc(&fs.z1buf, (size_t)2 * B * d1 * sizeof(double)));
        B2_CHECK_CUDA(cudaMalloc(&fs.z2buf, (size_t)2 * B * d2 * sizeof(double)));
        fs.cap_B = B;
    }
    if (!fs.V0t) {
        B2_CHECK_CUDA(cudaMalloc(&fs.V0t, (size_t)d1 * d0 * sizeof(double)));
        B2_CHECK_CUDA(cudaMalloc(&fs.V1b, (size_t)d1 * d2 * sizeof(double)));
        B2_CHECK_CUDA(cudaMalloc(&fs.V2r, (size_t)G * d2 * d3 * sizeof(double)));
        B2_CHECK_CUDA(cudaMalloc(&fs.bar, 2 * sizeof(unsigned long long)));
        B2_CHECK_CUDA(cudaMemsetAsync(fs.bar, 0, 2 * sizeof(unsigned long long), net->stream));
        B2_CHECK_CUDA(cudaFuncSetAttribute(fused_kernel_for(net->inner_act),
                                           cudaFuncAttributeMaxDynamicSharedMemorySize, net->max_smem_optin));
    }
    p.iters = iters;
    p.kind = net->kind; p.inner_act = net->inner_act; p.last_act = net->last_act; p.loss_kind = net->loss;
    p.stable = (net->flags & B2NN_FLAG_STABLE_SOFTMAX) ? 1 : 0;
    p.inner_p = net->inner_param; p.last_p = net->last_param; p.step = step; p.momentum = momentum;
    p.X = Xu8 ? nullptr : X; p.Xu8 = Xu8; p.T = T; p.labels = labels; p.idx = idx_dev;
    p.W0 = net->f64[0].w; p.W1 = net->f64[1].w; p.W2 = net->f64[2].w;
    p.V0 = net->f64[0].v; p.V1 = net->f64[1].v; p.V2 = net->f64[2].v;
    p.V0t = fs.V0t; p.V1b = fs.V1b; p.V2r = fs.V2r; p.z1buf = fs.z1buf; p.z2buf = fs.z2buf; p.bar = fs.bar;
    p.loss_out = loss_dev;
    // cp.async.bulk needs 16-byte aligned rows (global source, shared destination) and sizes
    p.bulk = Xu8 ? (d0 % 16 == 0) : (d0 % 2 == 0);
    if (getenv("B2NN_F64_NO_BULK")) p.bulk = 0;
    // the barrier counter only ever grows: this launch starts from where the previous ones left it (2 barriers of G
    // arrivals per step), so no reset is needed between launches
    p.bar_base = fs.bar_base;
    fs.bar_base += 2ull * (unsigned long long)G * (unsigned long long)iters;
    unsigned long long *&prof_dev = fs.prof;   // per net (hence per device)
    const bool prof = getenv("B2NN_K1_PROF") != nullptr || net->k1_prof;
    if (prof) {
        if (!prof_dev) B2_CHECK_CUDA(cudaMalloc(&prof_dev, 16 * sizeof(unsigned long long)));
        B2_CHECK_CUDA(cudaMemsetAsync(prof_dev, 0, 16 * sizeof(unsigned long long), net->stream));
    }
    p.prof = prof ? prof_dev : nullptr;
    void *args[] = {(void *)&p};
    B2_CHECK_CUDA(cudaLaunchCooperativeKernel(fused_kernel_for(net->inner_act), dim3(G), dim3(512), args,
                                              lo.smem_bytes, net->stream));
    net->launches++;
    if (prof) {
        unsigned long long h[16];
        B2_CHECK_CUDA(cudaMemcpyAsync(h, prof_dev, sizeof h, cudaMemcpyDeviceToHost, net->stream));
        B2_CHECK_CUDA(cudaStreamSynchronize(net->stream));
        // slots 8..14 split phase 3 (a2 load, z3, softmax/delta3, loss; slot 5 = delta2) and phase 5 (W0, W1 columns,
        // W1 rows; slot 7 = W2)
        const char *names[16] = {"stage", "phase1", "barrier1", "phase2", "barrier2", "p3:delta2", "phase4", "p5:W2",
                                 "p3:a2", "p3:z3", "p3:softmax", "p3:loss", "p5:W0", "p5:W1c", "p5:W1r", "stage:wait"};
        for (int i = 0; i < 16; i++) net->k1_prof_cycles[i] = (double)h[i] / iters;
        net->k1_prof_iters = iters;
        if (getenv("B2NN_K1_PROF") != nullptr) {
            fprintf(stderr, "[b2nn K1 prof] cycles per step (CTA 0):");
            for (int i = 0; i < 16; i++) fprintf(stderr, " %s %.0f", names[i], (double)h[i] / iters);
            fprintf(stderr, "\n");
        }
    }
    return B2NN_OK;
}

void f64_fused_free(b2nn_net *net) {
    auto &fs = net->fused;
    cudaFree(fs.V0t); cudaFree(fs.V1b); cudaFree(fs.V2r); cudaFree(fs.z1buf); cudaFree(fs.z2buf); cudaFree(fs.bar); cudaFree(fs.prof);
    fs = {};
}

}  // namespace b2nn
