// wide_path.cu -- host orchestration of the tensor-core path (BF16 or TF32 operands): per-step sequence of
// tcgen05 GEMMs (gemm_tcgen05.cuh) and the HBM-bound kernels around them (wide_kernels.cuh).
// Data layout in HBM (per layer l with W_l: d_l x d_{l+1}):
//   w, v, g  FP32 [d_l][pad8(d_{l+1})]   master weights / velocity / summed gradient (boundary layout, padded)
//   wn       BF16 [d_l][pad8(d_{l+1})]   the GEMM operand copy of w (MN-major B forward, K-major B delta)
//   a[l]     BF16 [B][pad8(d_l)]         activations (K-major A forward, MN-major A gradient)
//   d[l]     BF16 [B][pad8(d_l)]         deltas      (K-major A delta,   MN-major B gradient)
// No transposed copies exist: UMMA reads either major-ness from the same row-major buffer.
// B2NN_PREC_TF32 nets: a[l], d[l] are FP32 arrays of the same shapes, wn does not exist and the GEMMs read the
// FP32 master w directly (tcgen05 kind::tf32 ignores the low mantissa bits TMA did not round away).
#include "wide_path.hpp"

#include "comm.hpp"
#include "dp_kernels.cuh"
#include "gemm_tcgen05.cuh"
#include "wide_kernels.cuh"

namespace b2nn {

typedef CUresult (*PFN_tmapEncodeTiled)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                        const cuuint64_t *, const cuuint32_t *, const cuuint32_t *,
                                        CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                        CUtensorMapFloatOOBfill);

static PFN_tmapEncodeTiled get_encode_fn() {
    static PFN_tmapEncodeTiled fn = nullptr;
    if (fn) return fn;
    void *p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
        return nullptr;
    fn = reinterpret_cast<PFN_tmapEncodeTiled>(p);
    return fn;
}

// One map per operand buffer [rows][ld] with `inner` valid elements per row: a square box whose rows are 128 bytes
// (64 x 64 BF16, 32 x 32 FP32-as-TF32), 128-byte swizzle (matches make_desc_kmajor_sw128 /
// make_desc_mnmajor_sw128; TF32 buffers read MN-major get the 32-byte-atom swizzle of
// make_desc_mnmajor_sw128_base32b instead), out-of-bounds elements read as zero.  The TFLOAT32 map type makes TMA round the FP32
// words to TF32 on the way into shared memory (the tensor core would otherwise truncate them).
int make_tmap(CUtensorMap *m, const void *base, uint64_t inner, uint64_t rows, uint64_t ld, bool tf32, bool mn_major) {
    PFN_tmapEncodeTiled fn = get_encode_fn();
    B2_REQUIRE(fn != nullptr, B2NN_ERR_CUDA, "cuTensorMapEncodeTiled entry point not available");
    const cuuint32_t edge = tf32 ? (cuuint32_t)GemmCfg<true>::CH : (cuuint32_t)GemmCfg<false>::CH;
    cuuint64_t gdim[2] = {inner, rows};
    cuuint64_t gstride[1] = {ld * (tf32 ? 4u : 2u)};
    cuuint32_t box[2] = {edge, edge};
    cuuint32_t estr[2] = {1, 1};
    static const bool plain_f32 = getenv("B2NN_TF32_TMAP") && getenv("B2NN_TF32_TMAP")[0] == 'f';  // debug
    CUresult r = fn(m, tf32 ? (plain_f32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_TFLOAT32)
                            : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2,
                    const_cast<void *>(base), gdim, gstride, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE,
                    // kind::tf32 reads MN-major operands only from the 32-byte-atom flavour of the 128-byte swizzle
                    (tf32 && mn_major) ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B,
                    CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    B2_REQUIRE(r == CUDA_SUCCESS, B2NN_ERR_CUDA, "cuTensorMapEncodeTiled failed: %d (inner=%llu rows=%llu ld=%llu)",
               (int)r, (unsigned long long)inner, (unsigned long long)rows, (unsigned long long)ld);
    return B2NN_OK;
}

// measurement aid (B2NN_TRACE=1): remember a timing event on `st`
void trace_mark(b2nn_net *net, cudaStream_t st, const char *label, int l) {
    if (!net->trace || net->trace_marks.size() > 4000) return;
    cudaEvent_t e;
    if (cudaEventCreate(&e) != cudaSuccess) return;
    cudaEventRecord(e, st);
    char buf[64];
    snprintf(buf, sizeof buf, "%s%d", label, l);
    net->trace_marks.emplace_back(e, buf);
}
void trace_dump(b2nn_net *net, int first, int count) {
    if (!net->trace || net->trace_marks.empty()) return;
    cudaStreamSynchronize(net->stream);
    if (net->aux_stream) cudaStreamSynchronize(net->aux_stream);
    // marks labelled "step<k>" delimit steps: print those of steps [first, first+count)
    int step = -1;
    cudaEvent_t base = nullptr;
    for (auto &m : net->trace_marks) {
        if (m.second.rfind("step", 0) == 0) { step++; if (step == first) base = m.first; }
        if (step >= first && step < first + count && base) {
            float ms = 0.f;
            cudaEventElapsedTime(&ms, base, m.first);
            fprintf(stderr, "[b2nn trace r%d] %8.3f %s\n", net->rank, ms, m.second.c_str());
        }
    }
    for (auto &m : net->trace_marks) cudaEventDestroy(m.first);
    net->trace_marks.clear();
}

// grouped mode: record the stop event of the open contraction window
static int prof_close(b2nn_net *net) {
    if (net->prof_pending) {
        B2_CHECK_CUDA(cudaEventRecord(net->prof_pending, net->prof_pending_stream));
        net->prof_pending = nullptr;
    }
    return B2NN_OK;
}

// fold the recorded event pairs into the accumulators (synchronises the stream)
int profile_flush(b2nn_net *net) {
    B2_TRY(prof_close(net));
    if (net->prof_used) {
        B2_CHECK_CUDA(cudaStreamSynchronize(net->stream));
        if (net->copy_stream) B2_CHECK_CUDA(cudaStreamSynchronize(net->copy_stream));
        for (size_t i = 0; i + 1 < net->prof_used; i += 2) {
            float ms = 0.f;
            B2_CHECK_CUDA(cudaEventElapsedTime(&ms, net->prof_events[i], net->prof_events[i + 1]));
            const int cls = net->prof_cls[i / 2] & 3;
            net->prof_ms_cls[cls] += ms;
            net->prof_n_cls[cls]++;
            if (cls == 0) net->prof_ms_acc += ms;   // launches of class 0 are counted where they are made
        }
        net->prof_used = 0;
    }
    return B2NN_OK;
}

// event pair of class `cls` around whatever is launched on `st` between prof_begin and prof_end (profiling on only).
// A timing event between two kernels costs the stream several microseconds of idle time, part of which lands INSIDE the
// next window (profiles/r02_notes.md: 8 windows per C4 step summed to more than the whole overlapped step).  In grouped
// mode consecutive contraction launches on one stream therefore share ONE window: its stop event stays pending until
// something else is about to be launched (another class, another stream, the end of the forward / backward pass).
struct ProfPair { cudaEvent_t e1 = nullptr; int cls = 0; };
static int prof_begin(b2nn_net *net, int cls, cudaStream_t st, ProfPair *pp) {
    pp->e1 = nullptr;
    pp->cls = cls;
    if (!net->profiling) return B2NN_OK;
    if (net->prof_pending) {
        if (cls == 0 && st == net->prof_pending_stream) {   // the open window goes on
            pp->e1 = net->prof_pending;
            net->prof_pending = nullptr;
            return B2NN_OK;
        }
        B2_TRY(prof_close(net));
    }
    if (net->prof_used + 2 > net->prof_events.size()) B2_TRY(profile_flush(net));
    net->prof_cls[net->prof_used / 2] = (unsigned char)cls;
    cudaEvent_t e0 = net->prof_events[net->prof_used++];
    pp->e1 = net->prof_events[net->prof_used++];
    B2_CHECK_CUDA(cudaEventRecord(e0, st));
    return B2NN_OK;
}
static int prof_end(b2nn_net *net, cudaStream_t st, const ProfPair &pp) {
    if (!pp.e1) return B2NN_OK;
    if (net->prof_grouped && pp.cls == 0) { net->prof_pending = pp.e1; net->prof_pending_stream = st; return B2NN_OK; }
    B2_CHECK_CUDA(cudaEventRecord(pp.e1, st));
    return B2NN_OK;
}

static inline bool is_tf32(const b2nn_net *net) { return net->precision == B2NN_PREC_TF32; }
// B2NN_FLAG_BF16_WIRE: per-GPU gradients stored and exchanged as BF16 (BF16 nets, fused exchange only)
static inline bool bf16_wire(const b2nn_net *net) {
    return (net->flags & B2NN_FLAG_BF16_WIRE) != 0 && net->precision == B2NN_PREC_BF16;
}

typedef void (*GemmKernel)(const CUtensorMap, const CUtensorMap, const GemmArgs);
template <int EPI, bool A_MN, bool B_MN, bool TF32>
static GemmKernel gemm_kernel() {
    if constexpr (TF32) return gemm_tf32_pair_kernel<EPI, A_MN, B_MN>;
    else return gemm_bf16_pair_kernel<EPI, A_MN, B_MN>;
}

template <int EPI, bool A_MN, bool B_MN, bool TF32>
static int launch_gemm_t(b2nn_net *net, const CUtensorMap &ta, const CUtensorMap &tb, const GemmArgs &g) {
    const GemmKernel kernel = gemm_kernel<EPI, A_MN, B_MN, TF32>();
    // function attributes are per DEVICE: one process may hold nets on several GPUs (b2nn_create's `device`)
    static bool attr_done[kMaxDevices] = {};
    B2_REQUIRE(net->device >= 0 && net->device < kMaxDevices, B2NN_ERR_INVALID_ARG, "device ordinal %d", net->device);
    if (!attr_done[net->device]) {
        B2_CHECK_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, GEMM_SMEM_BYTES));
        attr_done[net->device] = true;
    }
    ProfPair pp;
    B2_TRY(prof_begin(net, 0, net->stream, &pp));
    const int tiles = ((g.M + 2 * GEMM_BM - 1) / (2 * GEMM_BM)) * ((g.N + GEMM_BN - 1) / GEMM_BN);
    const int max_pairs = net->num_sms / 2;
    const int pairs = tiles < max_pairs ? tiles : max_pairs;
    GemmArgs ga = g;
    ga.tile_counter = net->dynamic_tiles ? net->tile_counter : nullptr;
    static const int gemm_dbg = getenv("B2NN_GEMM_DBG") ? atoi(getenv("B2NN_GEMM_DBG")) : 0;  // bring-up aid
    ga.dbg = gemm_dbg;
    if (net->prof_stamping && net->prof_stamp_used < net->prof_stamp_cap) {
        ga.stamp = net->prof_stamps + 4 * (size_t)net->prof_stamp_used++;
        net->prof_stamp_flops += 2.0 * (double)g.M * (double)g.N * (double)g.K;
    }
    static const bool want_stalls = getenv("B2NN_GEMM_STALLS") != nullptr;  // debug: where the role warps wait
    static unsigned int *stalls_by_dev[kMaxDevices] = {};
    unsigned int *&stalls_dev = stalls_by_dev[net->device];
    if (want_stalls) {
        if (!stalls_dev) B2_CHECK_CUDA(cudaMalloc(&stalls_dev, 8 * 148 * 2 * sizeof(unsigned int)));
        B2_CHECK_CUDA(cudaMemsetAsync(stalls_dev, 0, 8 * 148 * 2 * sizeof(unsigned int), net->stream));
        ga.stalls = stalls_dev;
    }
    if (net->pdl) {
        // the kernel's prologue (barrier set-up, TMEM allocation, pair rendezvous) may overlap the previous kernel's drain
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(2 * pairs); cfg.blockDim = dim3(GEMM_THREADS);
        cfg.dynamicSmemBytes = GEMM_SMEM_BYTES; cfg.stream = net->stream;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = at; cfg.numAttrs = 1;
        B2_CHECK_CUDA(cudaLaunchKernelEx(&cfg, kernel, ta, tb, ga));
    } else {
        kernel<<<2 * pairs, GEMM_THREADS, GEMM_SMEM_BYTES, net->stream>>>(ta, tb, ga);
    }
    if (want_stalls) {
        static int printed = 0;
        std::vector<unsigned int> h(8 * 2 * pairs);
        B2_CHECK_CUDA(cudaMemcpyAsync(h.data(), stalls_dev, h.size() * sizeof(unsigned int), cudaMemcpyDeviceToHost, net->stream));
        B2_CHECK_CUDA(cudaStreamSynchronize(net->stream));
        double s[8] = {0};
        for (int c = 0; c < 2 * pairs; c++)
            for (int k = 0; k < 8; k++) s[k] += h[8 * c + k];
        const double P = pairs, C = 2.0 * pairs;
        if (printed++ < 64)
            fprintf(stderr, "[stalls] epi %d M %d N %d K %d: per leader: total %.0f clk, items %.2f, wait full %.0f, wait tempty %.0f | "
                            "producer wait empty %.0f | epilogue warp 2: wait tfull %.0f, busy %.0f (per item %.0f)\n",
                    EPI, g.M, g.N, g.K, s[3] / P, s[4] / P, s[1] / P, s[2] / P, s[0] / C, s[5] / C, s[6] / C,
                    s[6] / C / (s[4] / P));
    }
    B2_TRY(prof_end(net, net->stream, pp));
    if (net->profiling) { net->prof_flops += 2.0 * (double)g.M * (double)g.N * (double)g.K; net->prof_launch_acc++; }
    net->launches++;
    B2_CHECK_CUDA(cudaGetLastError());
    return B2NN_OK;
}

template <int EPI, bool A_MN, bool B_MN>
static int launch_gemm(b2nn_net *net, const CUtensorMap &ta, const CUtensorMap &tb, const GemmArgs &g) {
    if constexpr (EPI != EPI_DW_STORE_BF16) {
        if (is_tf32(net)) return launch_gemm_t<EPI, A_MN, B_MN, true>(net, ta, tb, g);
    }
    return launch_gemm_t<EPI, A_MN, B_MN, false>(net, ta, tb, g);
}

static int launch_update(b2nn_net *net, WideLayer &y, float scale, float momentum, int cast_only, cudaStream_t st) {
    const long long n4 = (long long)y.in * y.ld_n / 4;
    long long blocks = (n4 + 255) / 256;
    const long long cap = (long long)net->num_sms * 8;
    if (blocks > cap) blocks = cap;
    ProfPair pp;
    B2_TRY(prof_begin(net, 1, st, &pp));
    wide::sgd_update<<<(int)blocks, 256, 0, st>>>((const float4 *)y.g, (float4 *)y.w, (float4 *)y.v, n4, scale,
                                                  momentum, cast_only, (uint2 *)y.wn);
    B2_TRY(prof_end(net, st, pp));
    net->launches++;
    B2_CHECK_CUDA(cudaGetLastError());
    return B2NN_OK;
}

int wide_upload_weights(b2nn_net *net, const double *w_flat, bool velocity) {
    const int L = net->n_dims - 1;
    if (net->aux_stream) B2_CHECK_CUDA(cudaStreamSynchronize(net->aux_stream));
    for (int l = 0; l < L; l++) {
        WideLayer &y = net->wide[l];
        std::vector<float> tmp((size_t)y.in * y.ld_n, 0.f);
        const double *src = w_flat + net->w_off[l];
        for (int i = 0; i < y.in; i++)
            for (int j = 0; j < y.out; j++) tmp[(size_t)i * y.ld_n + j] = (float)src[(size_t)i * y.out + j];
        B2_CHECK_CUDA(cudaMemcpyAsync(velocity ? y.v : y.w, tmp.data(), tmp.size() * sizeof(float),
                                      cudaMemcpyHostToDevice, net->stream));
        B2_CHECK_CUDA(cudaStreamSynchronize(net->stream));
        if (!velocity && y.wn) B2_TRY(launch_update(net, y, 0.f, 0.f, 1, net->stream));  // rebuild the BF16 copy
    }
    B2_CHECK_CUDA(cudaStreamSynchronize(net->stream));
    return B2NN_OK;
}

int wide_download_weights(b2nn_net *net, double *w_flat, bool velocity) {
    const int L = net->n_dims - 1;
    B2_CHECK_CUDA(cudaStreamSynchronize(net->stream));
    if (net->aux_stream) B2_CHECK_CUDA(cudaStreamSynchronize(net->aux_stream));
    for (int l = 0; l < L; l++) {
        WideLayer &y = net->wide[l];
        std::vector<float> tmp((size_t)y.in * y.ld_n);
        if (net->peer_arena.empty()) {
            B2_CHECK_CUDA(cudaMemcpy(tmp.data(), velocity ? y.v : y.w, tmp.size() * sizeof(float), cudaMemcpyDeviceToHost));
        } else {
            // optimizer state is sharded: slice p of the master weights / velocity lives on rank p
            const long long n4 = (long long)y.in * y.ld_n / 4;
            const size_t off = (size_t)((char *)(velocity ? y.v : y.w) - net->arena);
            for (int p = 0; p < net->world; p++) {
                const long long lo = (long long)p * n4 / net->world, hi = (long long)(p + 1) * n4 / net->world;
                if (hi > lo)
                    B2_CHECK_CUDA(cudaMemcpy(tmp.data() + 4 * lo, (const float *)(net->peer_arena[p] + off) + 4 * lo,
                                             (size_t)(hi - lo) * 16, cudaMemcpyDefault));
            }
        }
        double *dst = w_flat + net->w_off[l];
        for (int i = 0; i < y.in; i++)
            for (int j = 0; j < y.out; j++) dst[(size_t)i * y.out + j] = (double)tmp[(size_t)i * y.ld_n + j];
    }
    return B2NN_OK;
}

// The L1/shared-memory split of an SM can only change while the SM is empty, and the GEMM CTAs need ~198 KB of
// shared memory: a resident block of a kernel that asked for the default (L1-heavy) split keeps GEMM CTAs off
// that SM until it exits.  Every kernel that runs beside or between the GEMMs therefore asks for the
// max-shared carve-out too, so the side-stream update / input prologue really overlap and no
// reconfiguration drain happens at kernel boundaries.
static int prefer_max_shared_carveout(int device) {
    static bool done_dev[kMaxDevices] = {};   // per device, like every function attribute
    B2_REQUIRE(device >= 0 && device < kMaxDevices, B2NN_ERR_INVALID_ARG, "device ordinal %d", device);
    if (done_dev[device]) return B2NN_OK;
    const int c = cudaSharedmemCarveoutMaxShared;
    B2_CHECK_CUDA(cudaFuncSetAttribute(wide::sgd_update, cudaFuncAttributePreferredSharedMemoryCarveout, c));
    B2_CHECK_CUDA(cudaFuncSetAttribute(dp::reduce_update_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, c));
    B2_CHECK_CUDA(cudaFuncSetAttribute(wide::input_prologue<double>, cudaFuncAttributePreferredSharedMemoryCarveout, c));
    B2_CHECK_CUDA(cudaFuncSetAttribute(wide::input_prologue<float>, cudaFuncAttributePreferredSharedMemoryCarveout, c));
    B2_CHECK_CUDA(cudaFuncSetAttribute(wide::input_prologue<__nv_bfloat16>,
                                       cudaFuncAttributePreferredSharedMemoryCarveout, c));
    B2_CHECK_CUDA(cudaFuncSetAttribute(wide::input_prologue<uint8_t>, cudaFuncAttributePreferredSharedMemoryCarveout, c));
    B2_CHECK_CUDA(cudaFuncSetAttribute(wide::input_prologue_u8x16, cudaFuncAttributePreferredSharedMemoryCarveout, c));
    B2_CHECK_CUDA(cudaFuncSetAttribute(wide::output_delta<__nv_bfloat16>, cudaFuncAttributePreferredSharedMemoryCarveout, c));
    B2_CHECK_CUDA(cudaFuncSetAttribute(wide::output_delta<float>, cudaFuncAttributePreferredSharedMemoryCarveout, c));
    B2_CHECK_CUDA(cudaFuncSetAttribute((wide::input_prologue<double, float>), cudaFuncAttributePreferredSharedMemoryCarveout, c));
    B2_CHECK_CUDA(cudaFuncSetAttribute((wide::input_prologue<float, float>), cudaFuncAttributePreferredSharedMemoryCarveout, c));
    B2_CHECK_CUDA(cudaFuncSetAttribute((wide::input_prologue<__nv_bfloat16, float>),
                                       cudaFuncAttributePreferredSharedMemoryCarveout, c));
    B2_CHECK_CUDA(cudaFuncSetAttribute((wide::input_prologue<uint8_t, float>), cudaFuncAttributePreferredSharedMemoryCarveout, c));
    B2_CHECK_CUDA(cudaFuncSetAttribute(wide::reduce_rows, cudaFuncAttributePreferredSharedMemoryCarveout, c));
    B2_CHECK_CUDA(cudaFuncSetAttribute(wide::gather_labels, cudaFuncAttributePreferredSharedMemoryCarveout, c));
    B2_CHECK_CUDA(cudaFuncSetAttribute(wide::gather_targets, cudaFuncAttributePreferredSharedMemoryCarveout, c));
    B2_CHECK_CUDA(cudaFuncSetAttribute(wide::eval_rows, cudaFuncAttributePreferredSharedMemoryCarveout, c));
    B2_CHECK_CUDA(cudaFuncSetAttribute(wide::propagate_rows, cudaFuncAttributePreferredSharedMemoryCarveout, c));
    done_dev[device] = true;
    return B2NN_OK;
}

int wide_init(b2nn_net *net, const std::vector<double> &w_host) {
    const int L = net->n_dims - 1;
    B2_TRY(prefer_max_shared_carveout(net->device));
    net->wide.resize(L);
    // One arena holds the cross-GPU flags and every layer's g / wn / w / v at the same offsets on every rank,
    // so that a single CUDA IPC handle lets the peers of a data-parallel group address all of it (comm.cu).
    auto up = [](size_t x) { return (x + 1023) & ~(size_t)1023; };
    size_t bytes = DP_FLAG_BYTES;
    for (int l = 0; l < L; l++) {
        WideLayer &y = net->wide[l];
        y.in = net->dims[l]; y.out = net->dims[l + 1];
        y.ld_n = pad8(y.out);
        const size_t nf = (size_t)y.in * y.ld_n;
        bytes += 3 * up(nf * sizeof(float)) + (is_tf32(net) ? 0 : up(nf * sizeof(__nv_bfloat16))) +
                 (bf16_wire(net) ? up(nf * sizeof(__nv_bfloat16)) : 0);
    }
    B2_CHECK_CUDA(cudaMalloc(&net->arena, bytes));
    B2_CHECK_CUDA(cudaMemset(net->arena, 0, bytes));
    net->arena_bytes = bytes;
    size_t off = DP_FLAG_BYTES;
    for (int l = 0; l < L; l++) {
        WideLayer &y = net->wide[l];
        const size_t nf = (size_t)y.in * y.ld_n;
        y.g = (float *)(net->arena + off); off += up(nf * sizeof(float));
        y.w = (float *)(net->arena + off); off += up(nf * sizeof(float));
        y.v = (float *)(net->arena + off); off += up(nf * sizeof(float));
        if (bf16_wire(net)) { y.gb = (__nv_bfloat16 *)(net->arena + off); off += up(nf * sizeof(__nv_bfloat16)); }
        if (is_tf32(net)) {
            y.wn = nullptr;  // the FP32 master array is the operand
            B2_TRY(make_tmap(&y.tm_w, y.w, y.out, y.in, y.ld_n, true, false));
            B2_TRY(make_tmap(&y.tm_w_mn, y.w, y.out, y.in, y.ld_n, true, true));
        } else {
            y.wn = (__nv_bfloat16 *)(net->arena + off); off += up(nf * sizeof(__nv_bfloat16));
            B2_TRY(make_tmap(&y.tm_w, y.wn, y.out, y.in, y.ld_n, false));
            y.tm_w_mn = y.tm_w;
        }
    }
    // side stream: the HBM-bound update (and, data-parallel, the all-reduce before it) of layer l runs there
    // while the tensor cores go on with layer l-1's contractions
    B2_CHECK_CUDA(cudaStreamCreateWithFlags(&net->aux_stream, cudaStreamNonBlocking));
    net->layer_events.resize(L);
    net->upd_done.resize(L);
    net->upd_pending.assign(L, 0);
    for (int l = 0; l < L; l++) {
        B2_CHECK_CUDA(cudaEventCreateWithFlags(&net->layer_events[l], cudaEventDisableTiming));
        B2_CHECK_CUDA(cudaEventCreateWithFlags(&net->upd_done[l], cudaEventDisableTiming));
    }
    B2_CHECK_CUDA(cudaEventCreateWithFlags(&net->comm_done, cudaEventDisableTiming));
    B2_CHECK_CUDA(cudaMalloc(&net->tile_counter, sizeof(unsigned int)));
    B2_CHECK_CUDA(cudaMemset(net->tile_counter, 0, sizeof(unsigned int)));
    B2_CHECK_CUDA(cudaStreamCreateWithFlags(&net->copy_stream, cudaStreamNonBlocking));
    for (int q = 0; q < 2; q++) {
        B2_CHECK_CUDA(cudaEventCreateWithFlags(&net->slot_ready[q], cudaEventDisableTiming));
        B2_CHECK_CUDA(cudaEventCreateWithFlags(&net->slot_free[q], cudaEventDisableTiming));
    }
    return wide_upload_weights(net, w_host.data(), false);
}

// the arena moved (comm_vmm.cu): every layer pointer is an offset from its base; the operand maps hold addresses
int wide_rebase_arena(b2nn_net *net, char *new_base) {
    char *old = net->arena;
    for (auto &y : net->wide) {
        y.g = (float *)(new_base + ((char *)y.g - old));
        y.w = (float *)(new_base + ((char *)y.w - old));
        y.v = (float *)(new_base + ((char *)y.v - old));
        if (y.gb) y.gb = (__nv_bfloat16 *)(new_base + ((char *)y.gb - old));
        if (y.wn) {
            y.wn = (__nv_bfloat16 *)(new_base + ((char *)y.wn - old));
            B2_TRY(make_tmap(&y.tm_w, y.wn, y.out, y.in, y.ld_n, false));
            y.tm_w_mn = y.tm_w;
        } else {
            B2_TRY(make_tmap(&y.tm_w, y.w, y.out, y.in, y.ld_n, true, false));
            B2_TRY(make_tmap(&y.tm_w_mn, y.w, y.out, y.in, y.ld_n, true, true));
        }
    }
    net->arena = new_base;
    return B2NN_OK;
}

static void free_batch(b2nn_net *net) {
    WideBatch &b = net->wb;
    if (!b.a.empty()) b.a[0] = nullptr;  // alias of a slot
    b.labels = nullptr;
    for (int q = 0; q < 2; q++) {
        cudaFree(b.a0_slot[q]); cudaFree(b.labels_slot[q]);
        b.a0_slot[q] = nullptr; b.labels_slot[q] = nullptr;
    }
    for (auto p : b.a) cudaFree(p);
    for (auto p : b.d) cudaFree(p);
    b.a.clear(); b.d.clear();
    cudaFree(b.logits); cudaFree(b.row_loss); cudaFree(b.tdense); cudaFree(b.xstage);
    cudaFree(b.loss_sum); cudaFree(b.block_partial); cudaFree(b.block_counter);
    b.block_partial = nullptr; b.block_counter = nullptr;
    b.logits = nullptr; b.row_loss = nullptr; b.tdense = nullptr; b.xstage = nullptr;
    b.loss_sum = nullptr;
    b.B = 0;
}

void wide_free(b2nn_net *net) {
    free_batch(net);
    for (auto e : net->layer_events) cudaEventDestroy(e);
    net->layer_events.clear();
    for (auto e : net->upd_done) cudaEventDestroy(e);
    net->upd_done.clear();
    if (net->comm_done) { cudaEventDestroy(net->comm_done); net->comm_done = nullptr; }
    if (net->aux_stream) { cudaStreamDestroy(net->aux_stream); net->aux_stream = nullptr; }
    if (net->copy_stream) { cudaStreamDestroy(net->copy_stream); net->copy_stream = nullptr; }
    cudaFree(net->tile_counter); net->tile_counter = nullptr;
    for (int q = 0; q < 2; q++) {
        if (net->slot_ready[q]) { cudaEventDestroy(net->slot_ready[q]); net->slot_ready[q] = nullptr; }
        if (net->slot_free[q]) { cudaEventDestroy(net->slot_free[q]); net->slot_free[q] = nullptr; }
    }
    if (net->arena && !net->vmm_arena) cudaFree(net->arena);
    net->arena = nullptr;
    net->wide.clear();
}

// point a[0] / tm_a[0] / labels at input slot q
void wide_use_slot(b2nn_net *net, int q) {
    WideBatch &b = net->wb;
    b.a[0] = b.a0_slot[q];
    b.tm_a[0] = b.tm_a0_slot[q];
    b.tm_a_mn[0] = b.tm_a0_slot_mn[q];
    b.labels = b.labels_slot[q];
}

int wide_ensure_batch(b2nn_net *net, int B) {
    WideBatch &b = net->wb;
    if (b.B == B) return B2NN_OK;
    B2_CHECK_CUDA(cudaStreamSynchronize(net->stream));
    free_batch(net);
    const int nd = net->n_dims, L = nd - 1;
    const bool tf32 = is_tf32(net);
    const size_t es = tf32 ? 4 : 2;  // operand element size
    b.a.assign(nd, nullptr); b.d.assign(nd, nullptr);
    b.tm_a.resize(nd); b.tm_d.resize(nd); b.tm_a_mn.resize(nd); b.tm_d_mn.resize(nd);
    for (int l = 0; l < nd; l++) {
        const int dl = net->dims[l], ld = pad8(dl);
        if (l == 0) {
            for (int q = 0; q < 2; q++) {
                B2_CHECK_CUDA(cudaMalloc(&b.a0_slot[q], (size_t)B * ld * es));
                B2_CHECK_CUDA(cudaMemset(b.a0_slot[q], 0, (size_t)B * ld * es));
                B2_TRY(make_tmap(&b.tm_a0_slot[q], b.a0_slot[q], dl, B, ld, tf32, false));
                B2_TRY(make_tmap(&b.tm_a0_slot_mn[q], b.a0_slot[q], dl, B, ld, tf32, true));
                B2_CHECK_CUDA(cudaMalloc(&b.labels_slot[q], (size_t)B * sizeof(int32_t)));
            }
        } else if (l < L) {
            B2_CHECK_CUDA(cudaMalloc(&b.a[l], (size_t)B * ld * es));
            B2_CHECK_CUDA(cudaMemset(b.a[l], 0, (size_t)B * ld * es));
            B2_TRY(make_tmap(&b.tm_a[l], b.a[l], dl, B, ld, tf32, false));
            B2_TRY(make_tmap(&b.tm_a_mn[l], b.a[l], dl, B, ld, tf32, true));
        }
        if (l >= 1) {
            B2_CHECK_CUDA(cudaMalloc(&b.d[l], (size_t)B * ld * es));
            B2_CHECK_CUDA(cudaMemset(b.d[l], 0, (size_t)B * ld * es));
            B2_TRY(make_tmap(&b.tm_d[l], b.d[l], dl, B, ld, tf32, false));
            B2_TRY(make_tmap(&b.tm_d_mn[l], b.d[l], dl, B, ld, tf32, true));
        }
    }
    const int ldL = pad8(net->dims[L]);
    B2_CHECK_CUDA(cudaMalloc(&b.logits, (size_t)B * ldL * sizeof(float)));
    B2_CHECK_CUDA(cudaMalloc(&b.row_loss, (size_t)B * sizeof(float)));
    B2_CHECK_CUDA(cudaMalloc(&b.tdense, (size_t)B * ldL * sizeof(float)));
    B2_CHECK_CUDA(cudaMalloc(&b.xstage, (size_t)B * net->dims[0] * sizeof(double)));
    B2_CHECK_CUDA(cudaMalloc(&b.loss_sum, 2 * sizeof(double)));
    B2_CHECK_CUDA(cudaMalloc(&b.block_partial, (size_t)((B + 7) / 8) * sizeof(double)));
    B2_CHECK_CUDA(cudaMalloc(&b.block_counter, sizeof(unsigned int)));
    B2_CHECK_CUDA(cudaMemset(b.block_counter, 0, sizeof(unsigned int)));
    b.B = B;
    wide_use_slot(net, 0);
    return B2NN_OK;
}

// K8: a0 = f(X[idx]) as BF16.  X device-visible in dtype x_dtype.
int wide_prologue(b2nn_net *net, const void *X, int x_dtype, const int32_t *idx_dev, int B, cudaStream_t st,
                  int blocks_per_sm) {
    WideBatch &b = net->wb;
    const int d0 = net->dims[0];
    const long long total = (long long)B * ((d0 + 7) / 8);
    long long blocks = (total + 255) / 256;
    if (blocks > (long long)net->num_sms * blocks_per_sm) blocks = (long long)net->num_sms * blocks_per_sm;
    int grid = (int)blocks;
    int thr = 256;
    if (blocks_per_sm == 1) {
        // rows in mapped pinned host memory: one 64-thread block per SM keeps PCIe saturated (~50 GB/s) while
        // bounding the number of multi-microsecond reads in flight; with more, the concurrently running GEMMs
        // slowed down 2.8x (measured, profiles/r01_notes.md)
        thr = 64;
        grid = net->num_sms;
    }
    if (const char *e = getenv("B2NN_PROLOGUE_THREADS")) { thr = atoi(e); grid = net->num_sms * blocks_per_sm; }
    const float p = (float)net->inner_param;
    const int act = net->inner_act, ld = pad8(d0);
    ProfPair pp;
    B2_TRY(prof_begin(net, 3, st, &pp));
    if (is_tf32(net)) {  // FP32 operand rows
        float *a0 = (float *)b.a[0];
        switch (x_dtype) {
        case B2NN_DT_F64: wide::input_prologue<double, float><<<grid, thr, 0, st>>>((const double *)X, idx_dev, B, d0, act, p, a0, ld); break;
        case B2NN_DT_F32: wide::input_prologue<float, float><<<grid, thr, 0, st>>>((const float *)X, idx_dev, B, d0, act, p, a0, ld); break;
        case B2NN_DT_BF16:
            wide::input_prologue<__nv_bfloat16, float><<<grid, thr, 0, st>>>((const __nv_bfloat16 *)X, idx_dev, B, d0, act, p, a0, ld);
            break;
        case B2NN_DT_U8: wide::input_prologue<uint8_t, float><<<grid, thr, 0, st>>>((const uint8_t *)X, idx_dev, B, d0, act, p, a0, ld); break;
        default: B2_REQUIRE(false, B2NN_ERR_INVALID_ARG, "unknown x dtype %d", x_dtype);
        }
    } else {
    __nv_bfloat16 *a0 = (__nv_bfloat16 *)b.a[0];
    switch (x_dtype) {
    case B2NN_DT_F64:
        wide::input_prologue<double><<<grid, thr, 0, st>>>((const double *)X, idx_dev, B, d0, act, p, a0, ld);
        break;
    case B2NN_DT_F32:
        wide::input_prologue<float><<<grid, thr, 0, st>>>((const float *)X, idx_dev, B, d0, act, p, a0, ld);
        break;
    case B2NN_DT_BF16:
        wide::input_prologue<__nv_bfloat16><<<grid, thr, 0, st>>>((const __nv_bfloat16 *)X, idx_dev, B, d0, act, p, a0, ld);
        break;
    case B2NN_DT_U8:
        if (d0 % 16 == 0) {
            wide::input_prologue_u8x16<<<grid, thr, 0, st>>>((const uint8_t *)X, idx_dev, B, d0, act, p, a0, ld);
            break;
        }
        wide::input_prologue<uint8_t><<<grid, thr, 0, st>>>((const uint8_t *)X, idx_dev, B, d0, act, p, a0, ld);
        break;
    default: B2_REQUIRE(false, B2NN_ERR_INVALID_ARG, "unknown x dtype %d", x_dtype);
    }
    }
    B2_TRY(prof_end(net, st, pp));
    net->launches++;
    B2_CHECK_CUDA(cudaGetLastError());
    return B2NN_OK;
}

// forward GEMMs (replaces S:169-193): Z = A_l . W_l, A K-major, W MN-major; the last one keeps FP32 logits
int wide_forward(b2nn_net *net, int B) {
    WideBatch &b = net->wb;
    const int L = net->n_dims - 1;
    for (int l = 0; l < L; l++) {
        WideLayer &y = net->wide[l];
        if (net->upd_pending[l]) {  // layer l's update of the previous step may still be on the side stream
            B2_CHECK_CUDA(cudaStreamWaitEvent(net->stream, net->upd_done[l], 0));
            net->upd_pending[l] = 0;
        }
        GemmArgs g{};
        g.M = B; g.N = y.out; g.K = y.in;
        g.act = net->inner_act; g.act_param = (float)net->inner_param;
        if (l + 1 < L) {
            g.out0 = b.a[l + 1]; g.ld0 = pad8(y.out);
            B2_TRY((launch_gemm<EPI_ACT, false, true>(net, b.tm_a[l], y.tm_w_mn, g)));
        } else {
            g.out0 = b.logits; g.ld0 = pad8(y.out);
            B2_TRY((launch_gemm<EPI_LOGITS, false, true>(net, b.tm_a[l], y.tm_w_mn, g)));
        }
        trace_mark(net, net->stream, "fwd", l);
    }
    return prof_close(net);   // whatever follows on this stream is not a contraction
}

// K5/K6: output delta + per-row loss
int wide_output(b2nn_net *net, int B, bool dense_targets, double *loss_slot_dev) {
    WideBatch &b = net->wb;
    const int L = net->n_dims - 1, dL = net->dims[L], ldL = pad8(dL);
    ProfPair pp;
    B2_TRY(prof_begin(net, 2, net->stream, &pp));
    if (is_tf32(net))
        wide::output_delta<float><<<(B + 7) / 8, 256, 0, net->stream>>>(
            b.logits, ldL, dense_targets ? nullptr : b.labels, dense_targets ? b.tdense : nullptr, ldL, B, dL, net->kind,
            net->last_act, (float)net->last_param, (float *)b.d[L], ldL, b.row_loss, loss_slot_dev, b.block_partial,
            b.block_counter);
    else
        wide::output_delta<__nv_bfloat16><<<(B + 7) / 8, 256, 0, net->stream>>>(
            b.logits, ldL, dense_targets ? nullptr : b.labels, dense_targets ? b.tdense : nullptr, ldL, B, dL, net->kind,
            net->last_act, (float)net->last_param, (__nv_bfloat16 *)b.d[L], ldL, b.row_loss, loss_slot_dev,
            b.block_partial, b.block_counter);
    B2_TRY(prof_end(net, net->stream, pp));
    net->launches++;
    B2_CHECK_CUDA(cudaGetLastError());
    return B2NN_OK;
}

// backward: deltas (S:260-278: D_l = D_l+1 . W_l^T, both operands K-major), gradient GEMMs with the batch sum
// as the K dimension (S:279-283 + S:305-322: G_l = A_l^T . D_l+1, both operands MN-major) and the momentum
// update (S:325-342) -- a streaming kernel after each gradient GEMM (or fused into its epilogue with
// B2NN_FUSED_UPDATE=1); all-reduced per layer on a side stream first when data-parallel.
static int launch_delta(b2nn_net *net, int l, int B) {
    WideBatch &b = net->wb;
    WideLayer &y = net->wide[l];
    GemmArgs g{};
    g.M = B; g.N = y.in; g.K = y.out;
    g.out0 = b.d[l]; g.ld0 = pad8(y.in);
    g.aux = b.a[l]; g.ldaux = pad8(y.in);
    g.act = net->inner_act; g.act_param = (float)net->inner_param;
    B2_TRY((launch_gemm<EPI_DX, false, false>(net, b.tm_d[l + 1], y.tm_w, g)));
    trace_mark(net, net->stream, "dX", l);
    return B2NN_OK;
}

// gradient GEMM of layer l, then (all-reduce and) momentum update on stream `us`; when `us` is the side
// stream, upd_pending[l] tells the next forward GEMM of layer l to wait for upd_done[l]
static int launch_gradient_and_update(b2nn_net *net, int l, int B, float scale, float momentum, cudaStream_t us) {
    WideBatch &b = net->wb;
    WideLayer &y = net->wide[l];
    GemmArgs g{};
    g.M = y.in; g.N = y.out; g.K = B;
    g.ldw = y.ld_n;
    if (net->world == 1 && net->fused_update) {
        // single GPU: the momentum update rides in the gradient GEMM's epilogue (no gradient array, no update kernel)
        g.w = y.w; g.v = y.v;
        g.out0 = y.wn; g.ld0 = y.ld_n;   // BF16 operand copy (null for TF32 nets: w itself is the operand)
        g.scale = scale; g.momentum = momentum;
        return launch_gemm<EPI_DW_UPDATE, true, true>(net, b.tm_a_mn[l], b.tm_d_mn[l + 1], g);
    }
    // BF16 wire (opt-in): the gradient leaves the epilogue as BF16 and the fused exchange reads 2 B per parameter
    const bool wire16 = y.gb != nullptr && net->world > 1 && !net->peer_arena.empty();
    if (wire16) {
        g.out0 = y.gb; g.ld0 = y.ld_n;
        B2_TRY((launch_gemm<EPI_DW_STORE_BF16, true, true>(net, b.tm_a_mn[l], b.tm_d_mn[l + 1], g)));
    } else {
        g.w = y.g;
        B2_TRY((launch_gemm<EPI_DW_STORE, true, true>(net, b.tm_a_mn[l], b.tm_d_mn[l + 1], g)));
    }
    trace_mark(net, net->stream, "dW", l);
    if (us != net->stream) {
        B2_CHECK_CUDA(cudaEventRecord(net->layer_events[l], net->stream));
        B2_CHECK_CUDA(cudaStreamWaitEvent(us, net->layer_events[l], 0));
    }
    trace_mark(net, us, "  side:begin", l);
    if (net->world > 1 && !net->peer_arena.empty()) {
        // fused reduce-scatter + update + all-gather over NVLink peer memory (dp_kernels.cuh)
        dp::Args a{};
        for (int p = 0; p < net->world; p++) a.arena[p] = net->peer_arena[p];
        a.mc = net->use_multimem ? net->mc_base : nullptr;
        a.world = net->world; a.rank = net->rank; a.layer = l;
        a.off_g = (unsigned long long)((char *)y.g - net->arena); a.off_w = (unsigned long long)((char *)y.w - net->arena);
        a.off_v = (unsigned long long)((char *)y.v - net->arena);
        a.off_wn = y.wn ? (unsigned long long)((char *)y.wn - net->arena) : 0ull;
        a.f32_operand = y.wn == nullptr;   // TF32 nets: the FP32 master array is the operand every rank reads
        a.g_bf16 = wire16;
        a.off_gb = wire16 ? (unsigned long long)((char *)y.gb - net->arena) : 0ull;
        a.timeout_ns = net->dp_timeout_ns;
        a.n4 = (long long)y.in * y.ld_n / 4;
        a.scale = scale; a.momentum = momentum;
        a.seq = ++net->dp_seq;
        long long blocks = (a.n4 / net->world + dp::DP_THREADS - 1) / dp::DP_THREADS;
        if (blocks > net->dp_ctas) blocks = net->dp_ctas;
        if (blocks < 1) blocks = 1;
        ProfPair pp;
        B2_TRY(prof_begin(net, 1, us, &pp));
        dp::reduce_update_kernel<<<(int)blocks, dp::DP_THREADS, 0, us>>>(a);
        B2_TRY(prof_end(net, us, pp));
        net->launches++;
        B2_CHECK_CUDA(cudaGetLastError());
    } else {
        if (net->world > 1) B2_TRY(comm_allreduce_sum_f32(net, y.g, (size_t)y.in * y.ld_n, us));
        trace_mark(net, us, "  side:allreduced", l);
        B2_TRY(launch_update(net, y, scale, momentum, 0, us));
    }
    trace_mark(net, us, "  side:updated", l);
    if (us != net->stream) {
        B2_CHECK_CUDA(cudaEventRecord(net->upd_done[l], us));
        net->upd_pending[l] = 1;
    }
    return B2NN_OK;
}

// A bounded cross-GPU wait of the exchange kernel (dp_kernels.cuh) or of K1's grid barrier gave up: the word it left
// behind becomes B2NN_ERR_CUDA here.  Called where the host synchronises anyway (b2nn_sync, calls that return data).
int check_device_error_words(b2nn_net *net) {
    if (net->world > 1 && !net->peer_arena.empty() && net->arena) {
        unsigned int e = 0;
        B2_CHECK_CUDA(cudaMemcpyAsync(&e, net->arena + offsetof(dp::Flags, error), sizeof e, cudaMemcpyDeviceToHost,
                                      net->stream));
        B2_CHECK_CUDA(cudaStreamSynchronize(net->stream));
        if (e != 0) {
            const unsigned int kind = e >> 24, layer = (e >> 8) & 0xffffu, peer = e & 0xffu;
            set_error("data-parallel exchange timed out on rank %d: %s (layer %u, peer rank %u) -- a peer process died or "
                      "stalled; the weights of this net are no longer consistent",
                      net->rank, kind == 1 ? "a peer's gradient never became ready" : kind == 2 ? "a peer never finished "
                      "writing its slice" : "the launch's first CTA never released the others", layer, peer);
            return B2NN_ERR_CUDA;
        }
    }
    if (net->fused.bar) {
        unsigned long long e = 0;
        B2_CHECK_CUDA(cudaMemcpyAsync(&e, net->fused.bar + 1, sizeof e, cudaMemcpyDeviceToHost, net->stream));
        B2_CHECK_CUDA(cudaStreamSynchronize(net->stream));
        B2_REQUIRE(e == 0, B2NN_ERR_CUDA, "the fused FP64 training kernel gave up waiting at a grid barrier (%llu polls)", e);
    }
    return B2NN_OK;
}

// make the main stream wait for every update still running on the side stream
int wide_join(b2nn_net *net) {
    for (size_t l = 0; l < net->upd_pending.size(); l++)
        if (net->upd_pending[l]) {
            B2_CHECK_CUDA(cudaStreamWaitEvent(net->stream, net->upd_done[l], 0));
            net->upd_pending[l] = 0;
        }
    return B2NN_OK;
}

// backward.  Deltas first, top-down (S:260-278: D_l = D_l+1 . W_l^T, both operands K-major; they need the
// OLD weights), then the gradient GEMMs bottom-up with the batch sum as the K dimension (S:279-283 +
// S:305-322: G_l = A_l^T . D_l+1, both operands MN-major), each followed on the side stream by its
// all-reduce (data-parallel) and momentum update (S:325-342).  Bottom-up order puts the big early layers'
// HBM / NVLink work under the remaining gradient GEMMs and leaves only the LAST (smallest) layer's exposed;
// the next step's forward GEMM of layer l waits just for layer l's update (wide_forward).
static int wide_backward_launches(b2nn_net *net, int B, int global_B, double step, double momentum) {
    const int L = net->n_dims - 1;
    const float scale = (float)(step / (double)global_B);  // step*G/batch.size(), S:333
    const float mu = (float)momentum;
    // with b2nn_profile_enable on, everything runs in stream order so that the event pairs time each contraction alone
    const bool side = (net->world > 1 || net->overlap_update) && !net->profiling;
    if (net->backward_reorder && !(net->fused_update && net->world == 1)) {
        for (int l = L - 1; l >= 1; l--) B2_TRY(launch_delta(net, l, B));
        for (int l = 0; l < L; l++)
            B2_TRY(launch_gradient_and_update(net, l, B, scale, mu, side ? net->aux_stream : net->stream));
        return B2NN_OK;
    }
    // layer-by-layer order (the reference's own, S:262-284)
    for (int l = L - 1; l >= 0; l--) {
        if (l >= 1) B2_TRY(launch_delta(net, l, B));
        // layer 0 has nothing left to overlap with on one GPU: update it in stream order
        cudaStream_t us = (net->world > 1 || (l > 0 && net->overlap_update)) ? net->aux_stream : net->stream;
        B2_TRY(launch_gradient_and_update(net, l, B, scale, mu, us));
    }
    return wide_join(net);
}
int wide_backward(b2nn_net *net, int B, int global_B, double step, double momentum) {
    B2_TRY(wide_backward_launches(net, B, global_B, step, momentum));
    return prof_close(net);
}

int wide_eval_rows(b2nn_net *net, int rows, const int32_t *labels_dev, int32_t *pred_dev, bool want_loss) {
    WideBatch &b = net->wb;
    const int L = net->n_dims - 1, dL = net->dims[L];
    wide::eval_rows<<<(rows + 7) / 8, 256, 0, net->stream>>>(b.logits, pad8(dL), labels_dev, rows, dL, net->kind,
                                                             net->last_act, (float)net->last_param, pred_dev,
                                                             want_loss ? b.row_loss : nullptr);
    net->launches++;
    B2_CHECK_CUDA(cudaGetLastError());
    return B2NN_OK;
}

// propagate's output pass (src/NeuralNet.java:16): logits of the last forward -> P_dev[rows][dL] doubles
int wide_propagate_rows(b2nn_net *net, int rows, double *P_dev) {
    const int L = net->n_dims - 1, dL = net->dims[L];
    wide::propagate_rows<<<(rows + 7) / 8, 256, 0, net->stream>>>(net->wb.logits, pad8(dL), rows, dL, net->kind,
                                                                  net->last_act, net->last_param, P_dev);
    net->launches++;
    B2_CHECK_CUDA(cudaGetLastError());
    return B2NN_OK;
}

int wide_reduce(b2nn_net *net, int rows, const int32_t *pred_dev, const int32_t *labels_dev, double *loss_sum_dev,
                long long *hits_dev) {
    wide::reduce_rows<<<1, 256, 0, net->stream>>>(loss_sum_dev ? net->wb.row_loss : nullptr, pred_dev, labels_dev,
                                                  rows, loss_sum_dev, hits_dev);
    net->launches++;
    B2_CHECK_CUDA(cudaGetLastError());
    return B2NN_OK;
}

int wide_gather_labels(b2nn_net *net, const int32_t *labels, const int32_t *idx_dev, int B, cudaStream_t st) {
    wide::gather_labels<<<(B + 255) / 256, 256, 0, st>>>(labels, idx_dev, B, net->wb.labels);
    net->launches++;
    B2_CHECK_CUDA(cudaGetLastError());
    return B2NN_OK;
}

int wide_gather_targets(b2nn_net *net, const double *T, const int32_t *idx_dev, int B) {
    const int dL = net->dims[net->n_dims - 1];
    wide::gather_targets<<<B, 128, 0, net->stream>>>(T, idx_dev, B, dL, net->wb.tdense, pad8(dL));
    net->launches++;
    B2_CHECK_CUDA(cudaGetLastError());
    return B2NN_OK;
}

// the work list of one contraction launch, exactly as the kernel's role warps decode it (b2nn_debug_tile_list)
int wide_tile_list(int M, int N, int pairs, int32_t *m0, int32_t *n0, int32_t *bn, int cap, int32_t *total) {
    B2_REQUIRE(M > 0 && N > 0 && pairs > 0 && total != nullptr, B2NN_ERR_INVALID_ARG, "M, N, pairs must be positive");
    const TileSched sched(M, N, pairs);
    *total = sched.total;
    for (int item = 0; item < sched.total && item < cap; item++) {
        int a, b, c;
        sched.decode(item, a, b, c);
        if (m0) m0[item] = a;
        if (n0) n0[item] = b;
        if (bn) bn[item] = c;
    }
    return B2NN_OK;
}

}  // namespace b2nn
