// common.hpp -- error plumbing, java.util.Random (for the reference-exact weight init) and the
// handle structs shared by the FP64 and BF16 paths of libb2nn.so.
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <string>
#include <utility>
#include <vector>

#include "../../include/b2nn.h"

namespace b2nn {

// ---- thread-local last error ----------------------------------------------------------------
void set_error(const char *fmt, ...);
const char *get_error();

#define B2_CHECK_CUDA(expr)                                                                          \
    do {                                                                                             \
        cudaError_t _e = (expr);                                                                     \
        if (_e != cudaSuccess) {                                                                     \
            b2nn::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e));   \
            return (_e == cudaErrorMemoryAllocation) ? B2NN_ERR_OOM : B2NN_ERR_CUDA;                 \
        }                                                                                            \
    } while (0)

#define B2_REQUIRE(cond, code, ...)            \
    do {                                       \
        if (!(cond)) {                         \
            b2nn::set_error(__VA_ARGS__);      \
            return (code);                     \
        }                                      \
    } while (0)

#define B2_TRY(expr)                 \
    do {                             \
        int _s = (expr);             \
        if (_s != B2NN_OK) return _s;\
    } while (0)

// ---- java.util.Random: 48-bit LCG fixed by the JDK javadoc.  Product copy used only for the
// weight init a2 (src/SoftmaxCrossEntropyNeuralNet.java:111,149); the oracle has its own. ------
struct JavaRandom {
    uint64_t seed;
    explicit JavaRandom(int64_t s) : seed(((uint64_t)s ^ 0x5DEECE66DULL) & ((1ULL << 48) - 1)) {}
    inline int32_t next(int bits) {
        seed = (seed * 0x5DEECE66DULL + 0xBULL) & ((1ULL << 48) - 1);
        return (int32_t)(uint32_t)(seed >> (48 - bits));
    }
    inline double next_double() {
        int64_t hi = next(26);
        int64_t lo = next(27);
        return (double)((hi << 27) + lo) * 0x1.0p-53;
    }
};

// leading dimension of a row-major BF16 / FP32 array with x columns: rows of wide arrays start on 128-byte
// boundaries (64 BF16) so that every 128-byte TMA box row is ONE L2 line -- with a 1000-wide layer padded only to 8
// the 2000-byte pitch made each box row straddle two lines (measured: the MMA warp of the K=1000 delta GEMM waited
// for operands 48 % of its time); small arrays keep the 16-byte granularity TMA requires
inline int pad8(int x) { return x >= 64 ? (x + 63) & ~63 : (x + 7) & ~7; }
// the tensor-core ("wide") path serves both reduced-precision modes; TF32 nets keep FP32 operands
inline bool is_wide_precision(int precision) { return precision == B2NN_PREC_BF16 || precision == B2NN_PREC_TF32; }
constexpr size_t DP_FLAG_BYTES = 8192;  // head of the arena: cross-GPU flags of the fused reduce/update kernel
constexpr int DP_MAX_WORLD = 8;
constexpr int DP_MAX_LAYERS = 32;   // weight matrices the flag block has handshake rows for (dp_kernels.cuh)
constexpr int kMaxDevices = 64;     // per-device bookkeeping of function attributes (one process may drive several GPUs)


// ---- per-layer state ----------------------------------------------------------------------
struct F64Layer {
    double *w = nullptr;  // [in][out] row-major (the boundary layout, src/NeuralNet.java:30-39)
    double *v = nullptr;  // previousUpdate, same layout
};

struct WideLayer {
    int in = 0, out = 0;
    int ld_n = 0;               // pad8(out): leading dim of the [in][out] arrays
    float *w = nullptr;         // FP32 master [in][ld_n]
    float *v = nullptr;         // FP32 velocity [in][ld_n]
    float *g = nullptr;         // FP32 summed gradient [in][ld_n]
    __nv_bfloat16 *gb = nullptr;  // BF16 [in][ld_n]: the summed gradient as it travels with B2NN_FLAG_BF16_WIRE (else null)
    __nv_bfloat16 *wn = nullptr;  // BF16 [in][ld_n]: GEMM operand copy (MN-major B forward, K-major B delta); null for TF32 nets
    CUtensorMap tm_w;             // box map over wn (BF16 nets) or over w itself (TF32 nets), read K-major (delta GEMM)
    CUtensorMap tm_w_mn;          // the map the forward GEMM reads it MN-major with (TF32: 32-byte-atom swizzle; BF16: == tm_w)
};

// activations / deltas for one batch size (row-major only; the GEMMs read either major-ness)
struct WideBatch {
    int B = 0;
    // operand-typed arrays: BF16 for B2NN_PREC_BF16 nets, FP32 (read as TF32) for B2NN_PREC_TF32 nets
    std::vector<void *> a;             // a[l]  [B][pad8(d_l)]   l = 0..L-1 (activations incl. f(x))
    std::vector<void *> d;             // d[l]  [B][pad8(d_l)]   l = 1..L   (deltas)
    float *logits = nullptr;           // [B][pad8(d_L)] FP32
    float *row_loss = nullptr;         // [B]
    int32_t *labels = nullptr;         // [B]
    float *tdense = nullptr;           // [B][pad8(d_L)] FP32 dense targets (optional path)
    void *xstage = nullptr;            // staged raw input rows [B][d0] (any dtype, 8 B / element max)
    std::vector<CUtensorMap> tm_a, tm_d;  // box maps over a[l], d[l] for K-major reads (forward A, delta A)
    std::vector<CUtensorMap> tm_a_mn, tm_d_mn;  // ... for MN-major reads (gradient GEMM; differ from the above for TF32 only)
    double *loss_sum = nullptr;        // device scalar(s)
    double *block_partial = nullptr;   // per-block loss partials of the output kernel
    unsigned int *block_counter = nullptr;
    // input double buffer: while step i computes from slot i&1, the prologue of step i+1 fills the other
    // slot on the copy stream.  a[0] / tm_a[0] / labels above alias the slot in use.
    void *a0_slot[2] = {nullptr, nullptr};
    int32_t *labels_slot[2] = {nullptr, nullptr};
    CUtensorMap tm_a0_slot[2], tm_a0_slot_mn[2];
};

}  // namespace b2nn

struct b2nn_dataset {
    b2nn_net *net = nullptr;  // identity check only (check_pair): never dereferenced at destroy time, the net may be gone
    int device = 0;           // device the rows were staged on
    int64_t n = 0;
    int d0 = 0, dL = 0;
    int location = 0;
    int x_dtype = 0;          // dtype rows are stored in (F64 for F64 nets; as given otherwise)
    void *x = nullptr;        // device rows [n][d0], or pinned host rows (mapped) for HOST_PINNED
    void *x_dev_alias = nullptr;  // device-visible pointer to x
    uint8_t *xu8 = nullptr;   // FP64 nets: device rows as bytes when every x is exactly byte/255.0 (M:97)
    double *t = nullptr;      // device dense targets [n][dL] (F64) or null
    int32_t *labels = nullptr;  // device labels [n] or null (one-hot targets implied)
    int32_t *labels_eval = nullptr;  // labels used for hit counting: == labels, or derived from t (M:183-186)
};

struct b2nn_net {
    int device = 0;
    int n_dims = 0;
    std::vector<int> dims;
    int kind = 0, inner_act = 0, last_act = 0, loss = 0, precision = 0, flags = 0;
    double inner_param = 0, last_param = 0;
    int64_t n_weights = 0;
    int64_t time = 0;  // gradient steps taken (`time`, S:23,343; the reference feeds it to the noise term only, Q12)
    std::vector<int64_t> w_off;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    cudaStream_t aux_stream = nullptr;  // updates / collectives
    cudaStream_t copy_stream = nullptr;  // input prologue of the next step
    cudaEvent_t slot_ready[2] = {nullptr, nullptr}, slot_free[2] = {nullptr, nullptr};
    int num_sms = 0;
    int64_t launches = 0;

    // FP64 path
    std::vector<b2nn::F64Layer> f64;
    struct F64Work {
        int cap_rows = 0;
        std::vector<double *> z, a, d;  // per layer [rows][d_l]
        double *t = nullptr;            // [rows][dL]
        double *row_loss = nullptr;     // [rows]
        int32_t *pred = nullptr;        // [rows]
        int32_t *idx = nullptr;
        int iota_rows = 0;              // idx[0..iota_rows) currently holds 0, 1, 2, .. (f64_iota)
        double *xin = nullptr;          // staged input rows
    } fw;
    double *f64_grad = nullptr;  // [P] scratch for weight_gradient
    // K1 persistent fused kernel (f64_fused.cu)
    struct Fused {
        double *V0t = nullptr, *V1b = nullptr, *V2r = nullptr, *z1buf = nullptr, *z2buf = nullptr;
        unsigned long long *bar = nullptr;
        unsigned long long bar_base = 0;     // arrivals counted in bar[0] by the launches so far
        unsigned long long *prof = nullptr;  // [16] phase cycle sums of CTA 0 (b2nn_k1_profile / B2NN_K1_PROF)
        int cap_B = 0;
    } fused;
    bool k1_prof = false;            // collect K1's phase counters (b2nn_k1_profile)
    double k1_prof_cycles[16] = {};  // ... of the last K1 launch, cycles per step
    int k1_prof_iters = 0;
    bool fused_f64_disabled = false;  // B2NN_F64_FUSED=0 forces the general multi-kernel path
    bool coop_launch = false;
    int max_smem_optin = 0;

    // BF16 path
    unsigned int *tile_counter = nullptr;  // dynamic tile scheduler of the GEMM kernel (all launches are stream-ordered)
    bool dynamic_tiles_forced = false;
    bool dynamic_tiles = false;  // B2NN_DYNAMIC_TILES=1, or data-parallel (NCCL kernels share the SMs)
    bool backward_reorder = true;  // all deltas first, then gradients bottom-up (B2NN_BACKWARD_REORDER=0: layer by layer)
    bool pdl = true;             // (B2NN_PDL=0 switches it off) consecutive contraction launches overlap prologue and drain (programmatic dependent launch)
    bool fused_update = true;   // single GPU: momentum update inside the gradient GEMM's epilogue (B2NN_FUSED_UPDATE=0: separate kernel)
    bool overlap_update = true;  // layer l's update on the side stream under layer l-1's GEMMs (B2NN_OVERLAP_UPDATE=0: off)  // momentum update inside the gradient GEMM's epilogue (B2NN_FUSED_UPDATE=1)
    std::vector<b2nn::WideLayer> wide;
    b2nn::WideBatch wb;

    // data-parallel
    char *arena = nullptr;           // flags + all layers' g / w / v / wn (one allocation, one IPC handle)
    size_t arena_bytes = 0;
    std::vector<char *> peer_arena;  // [world] peers' arenas mapped with CUDA IPC (own entry = arena); empty: NCCL path
    int dp_ctas = 0;                  // CTAs of the fused reduce/update kernel (B2NN_DP_CTAS)
    // NVSwitch multicast path (comm_vmm.cu): arena re-allocated with cuMemCreate, bound to a multicast object
    bool vmm_arena = false;
    CUmemGenericAllocationHandle vmm_handle = 0, mc_handle = 0;
    size_t vmm_bytes = 0, vmm_gran = 0;
    int vmm_fd = -1, mc_fd = -1;
    bool use_multimem = false;       // exchange kernel uses multimem.ld_reduce / multimem.st through mc_base
    char *mc_base = nullptr;         // multicast mapping of all ranks' arenas (multimem.* addresses)
    std::vector<CUmemGenericAllocationHandle> peer_handles;
    unsigned int dp_seq = 0;         // sequence number of the cross-GPU flag handshakes
    unsigned long long dp_timeout_ns = 20ull * 1000000000ull;  // bound of every cross-GPU wait (B2NN_DP_TIMEOUT_MS)
    void *nccl_comm = nullptr;
    int rank = 0, world = 1;
    std::vector<cudaEvent_t> layer_events;
    std::vector<cudaEvent_t> upd_done;   // per layer: side-stream update finished
    std::vector<char> upd_pending;       // per layer: upd_done[l] not yet waited for by the main stream
    cudaEvent_t comm_done = nullptr;

    // B2NN_TRACE=1: timeline marks (timing events) on the main / side streams, dumped by b2nn_train_indexed
    bool trace = false;
    std::vector<std::pair<cudaEvent_t, std::string>> trace_marks;

    // profiling (b2nn_profile_enable): event pairs around the launches of the tensor-core path, by kernel class
    // (0 contraction, 1 momentum update / exchange, 2 output delta, 3 input prologue + label gather)
    bool profiling = false;
    bool prof_grouped = false;      // b2nn_profile_enable(net, 2): ONE pair around each run of consecutive contraction launches
    // b2nn_profile_enable(net, 3): no events, nothing serialised -- every contraction launch stamps %globaltimer (first CTA
    // in, last CTA out) into its slot; b2nn_profile_read_stamps folds the slots into busy time
    bool prof_stamping = false;
    unsigned long long *prof_stamps = nullptr;   // [prof_stamp_cap][2] on the device
    int prof_stamp_cap = 0, prof_stamp_used = 0;
    double prof_stamp_flops = 0;
    double prof_stamp_mhz = 0;    // SM clock of the stamped launches of the last b2nn_profile_read_stamps (cycles / ns of CTA 0)
    cudaEvent_t prof_pending = nullptr;       // stop event of the open contraction window (grouped mode), not yet recorded
    cudaStream_t prof_pending_stream = nullptr;
    std::vector<cudaEvent_t> prof_events;  // start/stop pairs
    std::vector<unsigned char> prof_cls;   // class of pair i (events 2i, 2i+1)
    size_t prof_used = 0;
    double prof_flops = 0;
    double prof_ms_acc = 0;         // class 0
    int64_t prof_launch_acc = 0;    // class 0
    double prof_ms_cls[4] = {0, 0, 0, 0};
    int64_t prof_n_cls[4] = {0, 0, 0, 0};

    // host staging
    void *pinned = nullptr;
    size_t pinned_bytes = 0;
    void *dscratch = nullptr;
    size_t dscratch_bytes = 0;
};
