/*
 * b2nn_oracle_trainer.c -- CPU ORACLE (TEST INFRASTRUCTURE, NOT PRODUCT CODE).
 * Trainer / MNIST-harness half of the oracle: sampler, train loop, validate,
 * accuracy evaluation, IDX readers and the trainLog.csv row.
 * T = src/NeuralNetTrainer.java, M = src/MNISTTrainer.java (/root/reference/).
 */
#include "b2nn_oracle.h"

#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* ---- sampler: T:28-43 (ctor), T:143-158 (sample), T:164-168 (refill) ---- */
struct orc_sampler {
    int32_t n_rows;
    int32_t *list;   /* dataSampler: ArrayList of row ids */
    int32_t size;
    orc_jrandom rng; /* new Random(1), T:42 */
};

static void sampler_refill(orc_sampler *s) {      /* T:164-168, master order = IDX file order (Q9) */
    for (int32_t i = 0; i < s->n_rows; i++) s->list[i] = i;
    s->size = s->n_rows;
}

orc_sampler *orc_sampler_create(int32_t n_rows, int64_t seed) {
    orc_sampler *s = (orc_sampler *)calloc(1, sizeof(orc_sampler));
    s->n_rows = n_rows;
    s->list = (int32_t *)malloc(sizeof(int32_t) * (size_t)n_rows);
    sampler_refill(s);                            /* T:40 */
    orc_jrandom_init(&s->rng, seed);              /* T:42 */
    return s;
}

void orc_sampler_destroy(orc_sampler *s) {
    if (!s) return;
    free(s->list); free(s);
}

void orc_sampler_sample(orc_sampler *s, int32_t B, int32_t *idx_out) {
    for (int32_t i = 0; i < B; i++) {
        if (s->size == 0) sampler_refill(s);                          /* T:148-150 */
        int32_t r = orc_jrandom_next_int_bound(&s->rng, s->size);     /* T:151 */
        idx_out[i] = s->list[r];                                      /* T:152 */
        memmove(s->list + r, s->list + r + 1,
                sizeof(int32_t) * (size_t)(s->size - r - 1));         /* ArrayList.remove(r), T:153 */
        s->size--;
    }
}

/* ---- train: T:60-92 (unobserved, unmonitored branch T:88-90) ------------ */
void orc_train(orc_net *n, const double *x, const double *t, int32_t n_rows,
               const int32_t *idx_or_null, int32_t iters, int32_t B,
               double step, double momentum, int32_t noise, int32_t faithful,
               double *per_step_loss_or_null) {
    int d0 = orc_net_input_dim(n), dl = orc_net_output_dim(n);
    orc_sampler *s = idx_or_null ? NULL : orc_sampler_create(n_rows, 1);
    int32_t *idx = (int32_t *)malloc(sizeof(int32_t) * (size_t)B);
    double *bx = (double *)malloc(sizeof(double) * (size_t)B * d0);
    double *bt = (double *)malloc(sizeof(double) * (size_t)B * dl);
    for (int32_t it = 0; it < iters; it++) {
        if (idx_or_null) memcpy(idx, idx_or_null + (int64_t)it * B, sizeof(int32_t) * (size_t)B);
        else orc_sampler_sample(s, B, idx);                           /* T:89 sample(batchSize) */
        for (int b = 0; b < B; b++) {
            memcpy(bx + (int64_t)b * d0, x + (int64_t)idx[b] * d0, sizeof(double) * d0);
            memcpy(bt + (int64_t)b * dl, t + (int64_t)idx[b] * dl, sizeof(double) * dl);
        }
        if (per_step_loss_or_null) {
            double acc = 0;
            for (int b = 0; b < B; b++) acc += orc_net_loss(n, bx + (int64_t)b * d0, bt + (int64_t)b * dl);
            per_step_loss_or_null[it] = acc / B;
        }
        orc_net_gradient_step(n, bx, bt, B, step, momentum, noise, faithful); /* T:89 */
    }
    free(idx); free(bx); free(bt);
    if (s) orc_sampler_destroy(s);
}

/* ---- validate(int): T:102-113 ------------------------------------------- */
double orc_validate(orc_net *n, const double *x, const double *t, int32_t count) {
    int d0 = orc_net_input_dim(n), dl = orc_net_output_dim(n);
    double loss = 0;
    (void)d0; (void)dl;
    double *rows = (double *)malloc(sizeof(double) * (size_t)count);
    orc_net_loss_rows(n, x, t, count, rows);                                   /* T:110 (threads: rows are independent) */
    for (int32_t i = 0; i < count; i++) loss += rows[i];                       /* summed in key order like T:108-111 */
    free(rows);
    return loss / count;                                                       /* T:112 */
}

/* ---- accuracy: M:159-172 / M:180-197 ------------------------------------ */
int64_t orc_evaluate(orc_net *n, const double *x, const int32_t *labels,
                     int64_t n_rows, int32_t *pred_or_null) {
    int d0 = orc_net_input_dim(n), dl = orc_net_output_dim(n);
    (void)d0;
    double *all = (double *)malloc(sizeof(double) * (size_t)dl * (size_t)n_rows);
    orc_net_propagate_rows(n, x, n_rows, all);                                 /* M:163 (threads: rows are independent) */
    int64_t hits = 0;
    for (int64_t r = 0; r < n_rows; r++) {
        const double *out = all + r * dl;
        int actual = 0;
        for (int i = 0; i < dl; i++)
            if (out[i] >= out[actual]) actual = i;                             /* M:166-168 (Q10) */
        if (pred_or_null) pred_or_null[r] = actual;
        if (actual == labels[r]) hits++;                                       /* M:169 */
    }
    free(all);
    return hits;
}

/* ---- IDX readers: M:37-59 (headers), M:76-100 (readInt/readImage) ------- */
static int read_be_int(FILE *f, int32_t *v) {                                  /* M:76-78 */
    unsigned char b[4];
    if (fread(b, 1, 4, f) != 4) return -1;
    *v = (int32_t)(((uint32_t)b[0] << 24) + ((uint32_t)b[1] << 16) + ((uint32_t)b[2] << 8) + (uint32_t)b[3]);
    return 0;
}

int64_t orc_read_idx_images(const char *path, double *out, int64_t max_rows, int32_t *row_len) {
    FILE *f = fopen(path, "rb");
    if (!f) return -1;
    int32_t magic, count, h, w;
    if (read_be_int(f, &magic) || magic != 2051 || read_be_int(f, &count) ||
        read_be_int(f, &h) || read_be_int(f, &w)) { fclose(f); return -2; }    /* M:39,41,43 */
    int32_t bytes = h * w;
    if (row_len) *row_len = bytes;
    int64_t nr = count < max_rows ? count : max_rows;
    unsigned char *raw = (unsigned char *)malloc((size_t)bytes);
    for (int64_t r = 0; r < nr; r++) {
        if (fread(raw, 1, (size_t)bytes, f) != (size_t)bytes) { free(raw); fclose(f); return -3; }
        for (int i = 0; i < bytes; i++) out[r * bytes + i] = (raw[i] & 0xff) / 255.0;  /* M:97 */
    }
    free(raw); fclose(f);
    return nr;
}

int64_t orc_read_idx_labels(const char *path, int32_t *out, int64_t max_rows) {
    FILE *f = fopen(path, "rb");
    if (!f) return -1;
    int32_t magic, count;
    if (read_be_int(f, &magic) || magic != 2049 || read_be_int(f, &count)) { fclose(f); return -2; } /* M:38,40 */
    int64_t nr = count < max_rows ? count : max_rows;
    for (int64_t r = 0; r < nr; r++) {
        int c = fgetc(f);                                                      /* M:114 */
        if (c == EOF) { fclose(f); return -3; }
        out[r] = c;
    }
    fclose(f);
    return nr;
}

/* ---- logTest row: M:211-219 ---------------------------------------------- */
int orc_format_log_row(char *buf, int buflen, const int32_t *dims, int32_t n_dims,
                       int32_t iterations, double step, int32_t batch, double momentum,
                       int32_t noise, double train_acc, double test_acc) {
    int o = snprintf(buf, (size_t)buflen, "%d", dims[0]);                      /* M:212 */
    for (int i = 1; i < n_dims && o < buflen; i++)
        o += snprintf(buf + o, (size_t)(buflen - o), "-%d", dims[i]);          /* M:213-215 */
    o += snprintf(buf + o, (size_t)(buflen - o), ",%d,%.5f,%d,%.3f,", iterations, step, batch, momentum); /* M:216 */
    o += snprintf(buf + o, (size_t)(buflen - o), "%s,", noise ? "true" : "false"); /* M:217 */
    o += snprintf(buf + o, (size_t)(buflen - o), "%.5f,%.5f\n", train_acc, test_acc); /* M:218 */
    return o;
}
