// checkpoint.cu -- C-ABI of libb2nn.so, part 4: a flat checkpoint file of a net.
//
// The reference has no persistence at all: there is no weight accessor on NeuralNet
// (src/NeuralNet.java:7-67), so a trained net is lost at process exit and the 500 000-iteration run behind
// logs/trainLog.csv:5 cannot be resumed.  The file holds exactly the state gradientStep mutates
// (src/SoftmaxCrossEntropyNeuralNet.java:327-344): the weights, `previousUpdate` (the velocity) and `time`,
// in the boundary layout (flat concatenation over layers of W[l][i][j], src/NeuralNet.java:30-39).
//
// Layout, little-endian, no padding:
//   char[8]  magic "B2NNCKPT"
//   u32      format version (1)
//   u32      n_dims
//   i32      dims[n_dims]
//   i32      net_kind, inner_act, last_act, loss, precision, flags
//   f64      inner_param, last_param
//   i64      time            gradient steps taken
//   i64      n_weights       = sum dims[l] * dims[l+1]
//   f64      W[n_weights]
//   f64      V[n_weights]
//   u64      FNV-1a 64 of every byte before it
// Everything here is host code over b2nn_get_* / b2nn_set_*; only save and load need the GPU.
#include <fcntl.h>
#include <unistd.h>

#include <atomic>
#include <cerrno>
#include <cstring>
#include <memory>

#include "common.hpp"

namespace {
using namespace b2nn;

constexpr char kMagic[8] = {'B', '2', 'N', 'N', 'C', 'K', 'P', 'T'};
constexpr uint32_t kVersion = 1;
constexpr uint32_t kMaxDims = 4096;

struct Fnv {
    uint64_t h = 0xcbf29ce484222325ULL;
    void add(const void *p, size_t n) {
        const unsigned char *b = static_cast<const unsigned char *>(p);
        for (size_t i = 0; i < n; i++) { h ^= b[i]; h *= 0x100000001b3ULL; }
    }
};

struct FileCloser {
    void operator()(FILE *f) const { if (f) fclose(f); }
};
typedef std::unique_ptr<FILE, FileCloser> File;

struct Header {
    uint32_t version = 0, n_dims = 0;
    std::vector<int32_t> dims;
    int32_t kind = 0, inner_act = 0, last_act = 0, loss = 0, precision = 0, flags = 0;
    double inner_param = 0, last_param = 0;
    int64_t time = 0, n_weights = 0;
};

bool put(FILE *f, Fnv &h, const void *p, size_t n) {
    h.add(p, n);
    return fwrite(p, 1, n, f) == n;
}
bool get(FILE *f, Fnv &h, void *p, size_t n) {
    if (fread(p, 1, n, f) != n) return false;
    h.add(p, n);
    return true;
}

// reads and validates everything up to the weight arrays
int read_header(FILE *f, Fnv &h, Header &hd, const char *path) {
    char magic[8];
    B2_REQUIRE(get(f, h, magic, 8) && memcmp(magic, kMagic, 8) == 0, B2NN_ERR_INVALID_ARG,
               "%s is not a b2nn checkpoint (bad magic)", path);
    B2_REQUIRE(get(f, h, &hd.version, 4) && hd.version == kVersion, B2NN_ERR_UNSUPPORTED,
               "%s: checkpoint format version %u, this library reads %u", path, hd.version, kVersion);
    B2_REQUIRE(get(f, h, &hd.n_dims, 4) && hd.n_dims >= 2 && hd.n_dims <= kMaxDims, B2NN_ERR_INVALID_ARG,
               "%s: bad layer count %u", path, hd.n_dims);
    hd.dims.resize(hd.n_dims);
    B2_REQUIRE(get(f, h, hd.dims.data(), 4 * (size_t)hd.n_dims), B2NN_ERR_INVALID_ARG, "%s: truncated header", path);
    int32_t six[6];
    double two[2];
    int64_t tw[2];
    B2_REQUIRE(get(f, h, six, sizeof six) && get(f, h, two, sizeof two) && get(f, h, tw, sizeof tw),
               B2NN_ERR_INVALID_ARG, "%s: truncated header", path);
    hd.kind = six[0]; hd.inner_act = six[1]; hd.last_act = six[2]; hd.loss = six[3]; hd.precision = six[4];
    hd.flags = six[5];
    hd.inner_param = two[0]; hd.last_param = two[1];
    hd.time = tw[0]; hd.n_weights = tw[1];
    int64_t tot = 0;
    for (uint32_t l = 0; l < hd.n_dims; l++) {
        B2_REQUIRE(hd.dims[l] > 0, B2NN_ERR_INVALID_ARG, "%s: layerDims[%u] = %d", path, l, hd.dims[l]);
        if (l + 1 < hd.n_dims) {
            // untrusted file: dims are positive int32, so a product fits int64; bound the running sum (2^40 weights = 8 TiB)
            tot += (int64_t)hd.dims[l] * hd.dims[l + 1];
            B2_REQUIRE(tot <= ((int64_t)1 << 40), B2NN_ERR_INVALID_ARG, "%s: implausible weight count", path);
        }
    }
    B2_REQUIRE(tot == hd.n_weights && hd.time >= 0, B2NN_ERR_INVALID_ARG,
               "%s: weight count %lld does not match the layer dimensions (%lld)", path, (long long)hd.n_weights,
               (long long)tot);
    return B2NN_OK;
}

}  // namespace

extern "C" {

int b2nn_time(const b2nn_net *net, int64_t *time) {
    B2_REQUIRE(net != nullptr && time != nullptr, B2NN_ERR_INVALID_ARG, "null argument");
    *time = net->time;
    return B2NN_OK;
}

int b2nn_checkpoint_save(b2nn_net *net, const char *path) {
    B2_REQUIRE(net != nullptr && path != nullptr, B2NN_ERR_INVALID_ARG, "null argument");
    const size_t n = (size_t)net->n_weights;
    std::vector<double> w(n), v(n);
    B2_TRY(b2nn_get_weights(net, w.data()));   // synchronises the net's streams
    B2_TRY(b2nn_get_velocity(net, v.data()));
    // Write beside the target under a name unique to this process and call (two savers of one path -- e.g. several
    // data-parallel ranks, though only ONE rank should save: they hold identical state -- never share a temp file),
    // fsync the file, rename it over the target and fsync the directory: a crash or power loss leaves either the old
    // checkpoint or the complete new one under `path`.
    static std::atomic<unsigned> counter{0};
    const std::string tmp = std::string(path) + ".tmp." + std::to_string((long long)getpid()) + "." +
                            std::to_string(counter.fetch_add(1));
    File f(fopen(tmp.c_str(), "wb"));
    B2_REQUIRE(f != nullptr, B2NN_ERR_INVALID_ARG, "cannot open %s for writing: %s", tmp.c_str(), strerror(errno));
    Fnv h;
    const uint32_t nd = (uint32_t)net->n_dims;
    const int32_t six[6] = {net->kind, net->inner_act, net->last_act, net->loss, net->precision, net->flags};
    const double two[2] = {net->inner_param, net->last_param};
    const int64_t tw[2] = {net->time, net->n_weights};
    std::vector<int32_t> dims(net->dims.begin(), net->dims.end());
    bool ok = put(f.get(), h, kMagic, 8) && put(f.get(), h, &kVersion, 4) && put(f.get(), h, &nd, 4) &&
              put(f.get(), h, dims.data(), 4 * (size_t)nd) && put(f.get(), h, six, sizeof six) &&
              put(f.get(), h, two, sizeof two) && put(f.get(), h, tw, sizeof tw) &&
              put(f.get(), h, w.data(), n * sizeof(double)) && put(f.get(), h, v.data(), n * sizeof(double));
    const uint64_t sum = h.h;
    ok = ok && fwrite(&sum, 1, 8, f.get()) == 8 && fflush(f.get()) == 0 && fsync(fileno(f.get())) == 0;
    FILE *raw = f.release();
    ok = (fclose(raw) == 0) && ok;
    if (!ok) {
        remove(tmp.c_str());
        set_error("short write to %s: %s", tmp.c_str(), strerror(errno));
        return B2NN_ERR_INVALID_ARG;
    }
    if (rename(tmp.c_str(), path) != 0) {
        set_error("cannot rename %s to %s: %s", tmp.c_str(), path, strerror(errno));
        remove(tmp.c_str());
        return B2NN_ERR_INVALID_ARG;
    }
    {   // make the rename itself durable
        const std::string p(path);
        const size_t slash = p.find_last_of('/');
        const std::string dir = slash == std::string::npos ? "." : (slash == 0 ? "/" : p.substr(0, slash));
        const int dfd = open(dir.c_str(), O_RDONLY | O_DIRECTORY);
        if (dfd >= 0) { fsync(dfd); close(dfd); }
    }
    return B2NN_OK;
}

int b2nn_checkpoint_info(const char *path, int32_t *n_dims, int32_t *dims, int32_t dims_cap, int32_t *net_kind,
                         int32_t *precision, int64_t *time, int64_t *n_weights) {
    B2_REQUIRE(path != nullptr, B2NN_ERR_INVALID_ARG, "path is null");
    File f(fopen(path, "rb"));
    B2_REQUIRE(f != nullptr, B2NN_ERR_INVALID_ARG, "cannot open %s: %s", path, strerror(errno));
    Fnv h;
    Header hd;
    B2_TRY(read_header(f.get(), h, hd, path));
    if (n_dims) *n_dims = (int32_t)hd.n_dims;
    if (dims) {
        B2_REQUIRE(dims_cap >= (int32_t)hd.n_dims, B2NN_ERR_INVALID_ARG, "dims buffer holds %d entries, the file has %u",
                   dims_cap, hd.n_dims);
        memcpy(dims, hd.dims.data(), 4 * (size_t)hd.n_dims);
    }
    if (net_kind) *net_kind = hd.kind;
    if (precision) *precision = hd.precision;
    if (time) *time = hd.time;
    if (n_weights) *n_weights = hd.n_weights;
    return B2NN_OK;
}

int b2nn_checkpoint_load(b2nn_net *net, const char *path) {
    B2_REQUIRE(net != nullptr && path != nullptr, B2NN_ERR_INVALID_ARG, "null argument");
    File f(fopen(path, "rb"));
    B2_REQUIRE(f != nullptr, B2NN_ERR_INVALID_ARG, "cannot open %s: %s", path, strerror(errno));
    Fnv h;
    Header hd;
    B2_TRY(read_header(f.get(), h, hd, path));
    B2_REQUIRE((int)hd.n_dims == net->n_dims, B2NN_ERR_DIM_MISMATCH, "%s holds %u layers, the net has %d", path, hd.n_dims,
               net->n_dims);
    for (int l = 0; l < net->n_dims; l++)
        B2_REQUIRE(hd.dims[l] == net->dims[l], B2NN_ERR_DIM_MISMATCH, "%s: layerDims[%d] = %d, the net has %d", path, l,
                   hd.dims[l], net->dims[l]);
    // weights trained under one set of functions are not a checkpoint of a net with another
    B2_REQUIRE(hd.kind == net->kind && hd.inner_act == net->inner_act && hd.inner_param == net->inner_param &&
                   (hd.kind != B2NN_NET_GENERAL ||
                    (hd.last_act == net->last_act && hd.last_param == net->last_param && hd.loss == net->loss)),
               B2NN_ERR_DIM_MISMATCH, "%s was written by a net of another class or with other activation / loss functions",
               path);
    const size_t n = (size_t)hd.n_weights;
    std::vector<double> w(n), v(n);
    B2_REQUIRE(get(f.get(), h, w.data(), n * sizeof(double)) && get(f.get(), h, v.data(), n * sizeof(double)),
               B2NN_ERR_INVALID_ARG, "%s: truncated weight arrays", path);
    uint64_t sum = 0;
    B2_REQUIRE(fread(&sum, 1, 8, f.get()) == 8 && sum == h.h, B2NN_ERR_INVALID_ARG,
               "%s: checksum mismatch (file damaged or truncated)", path);
    // nothing has touched the net so far: a bad file leaves it as it was
    B2_TRY(b2nn_set_weights(net, w.data()));
    B2_TRY(b2nn_set_velocity(net, v.data()));
    net->time = hd.time;
    return B2NN_OK;
}

}  // extern "C"
