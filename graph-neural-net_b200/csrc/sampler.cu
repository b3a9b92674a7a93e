// sampler.cu -- host-side mini-batch sampler with the reference's exact draw semantics
// (src/NeuralNetTrainer.java:143-168): r = random.nextInt(sampler.size()); take the r-th entry of the
// ArrayList; remove it (later entries shift down); refill from the master order when empty.
// The r-th entry of the shrinking list is the r-th row not yet drawn this epoch, so a Fenwick tree
// over "still in the list" flags reproduces it in O(log n) per draw instead of the O(n) shift.
#include "common.hpp"

struct b2nn_sampler {
    int32_t n = 0, size = 0, top = 1;
    std::vector<int32_t> tree;  // 1-based Fenwick tree of alive counts
    uint64_t seed = 0;          // java.util.Random state

    int32_t next(int bits) {
        seed = (seed * 0x5DEECE66DULL + 0xBULL) & ((1ULL << 48) - 1);
        return (int32_t)(uint32_t)(seed >> (48 - bits));
    }
    int32_t next_int(int32_t bound) {  // java.util.Random.nextInt(int)
        int32_t r = next(31);
        const int32_t m = bound - 1;
        if ((bound & m) == 0) return (int32_t)(((int64_t)bound * (int64_t)r) >> 31);
        for (int32_t u = r;; u = next(31)) {
            r = u % bound;
            if ((int32_t)((uint32_t)u - (uint32_t)r + (uint32_t)m) >= 0) return r;
        }
    }
    void refill() {  // refillSampler, T:164-168
        tree.assign((size_t)n + 1, 0);
        for (int32_t i = 1; i <= n; i++) {
            tree[i] += 1;
            const int32_t j = i + (i & -i);
            if (j <= n) tree[j] += tree[i];
        }
        size = n;
    }
    int32_t take(int32_t r) {  // ArrayList.get(r) + remove(r), T:152-153
        int32_t pos = 0, k = r + 1;
        for (int32_t step = top; step > 0; step >>= 1) {
            const int32_t nx = pos + step;
            if (nx <= n && tree[nx] < k) { pos = nx; k -= tree[nx]; }
        }
        const int32_t row = pos;  // 0-based row id (pos+1 is the 1-based slot)
        for (int32_t i = pos + 1; i <= n; i += i & -i) tree[i] -= 1;
        size--;
        return row;
    }
};

extern "C" {

int b2nn_sampler_create(int32_t n_rows, int64_t seed, b2nn_sampler **out) {
    B2_REQUIRE(out != nullptr, B2NN_ERR_INVALID_ARG, "out is null");
    *out = nullptr;
    B2_REQUIRE(n_rows > 0, B2NN_ERR_INVALID_ARG, "n_rows must be positive");
    b2nn_sampler *s = new b2nn_sampler();
    s->n = n_rows;
    while ((s->top << 1) <= n_rows) s->top <<= 1;
    s->seed = ((uint64_t)seed ^ 0x5DEECE66DULL) & ((1ULL << 48) - 1);
    s->refill();  // T:40
    *out = s;
    return B2NN_OK;
}

int b2nn_sampler_draw(b2nn_sampler *s, int32_t iters, int32_t B, int32_t *idx_out) {
    B2_REQUIRE(s != nullptr && idx_out != nullptr, B2NN_ERR_INVALID_ARG, "null argument");
    B2_REQUIRE(iters > 0 && B > 0 && B < s->n, B2NN_ERR_INVALID_ARG,
               "need iters > 0 and 0 < batchSize < rows (T:145)");
    for (int64_t q = 0; q < (int64_t)iters * B; q++) {
        if (s->size == 0) s->refill();               // T:148-150
        const int32_t r = s->next_int(s->size);      // T:151
        idx_out[q] = s->take(r);
    }
    return B2NN_OK;
}

int b2nn_sampler_destroy(b2nn_sampler *s) {
    delete s;
    return B2NN_OK;
}

}  // extern "C"
