/*
 * b2nn_oracle.c -- CPU ORACLE (TEST INFRASTRUCTURE, NOT PRODUCT CODE).
 * See b2nn_oracle.h for scope, parity status and who may call this.
 * Build: oracle/Makefile (gcc -O2 -ffp-contract=off; no fast-math).
 * Abbreviations in citations: S = src/SoftmaxCrossEntropyNeuralNet.java,
 * G = src/GeneralNeuralNet.java, T = src/NeuralNetTrainer.java,
 * M = src/MNISTTrainer.java (all under /root/reference/).
 */
#include "b2nn_oracle.h"

#include <math.h>
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* ======================================================================== */
/* java.util.Random -- algorithm fixed by the JDK javadoc (not in the       */
/* reference tree; the reference uses it at S:111,149, T:42,152, S:335).    */
/* ======================================================================== */
#define JR_MULT 0x5DEECE66DULL
#define JR_ADD 0xBULL
#define JR_MASK ((1ULL << 48) - 1)

void orc_jrandom_init(orc_jrandom *r, int64_t seed) {
    r->seed = ((uint64_t)seed ^ JR_MULT) & JR_MASK; /* initialScramble */
    r->have_next_gaussian = 0;
    r->next_gaussian = 0.0;
}

static int32_t jr_next(orc_jrandom *r, int bits) {
    r->seed = (r->seed * JR_MULT + JR_ADD) & JR_MASK;
    return (int32_t)(uint32_t)(r->seed >> (48 - bits)); /* (int)(seed >>> (48-bits)) */
}

int32_t orc_jrandom_next_int(orc_jrandom *r) { return jr_next(r, 32); }

int32_t orc_jrandom_next_int_bound(orc_jrandom *r, int32_t bound) {
    int32_t rr = jr_next(r, 31);
    int32_t m = bound - 1;
    if ((bound & m) == 0) { /* power of two */
        rr = (int32_t)(((int64_t)bound * (int64_t)rr) >> 31);
    } else {
        int32_t u = rr;
        for (;;) {
            rr = u % bound;
            /* Java: u - r + m < 0 with int wrap-around */
            uint32_t w = (uint32_t)u - (uint32_t)rr + (uint32_t)m;
            if ((int32_t)w >= 0) break;
            u = jr_next(r, 31);
        }
    }
    return rr;
}

double orc_jrandom_next_double(orc_jrandom *r) {
    int64_t hi = (int64_t)jr_next(r, 26);
    int64_t lo = (int64_t)jr_next(r, 27);
    return (double)((hi << 27) + lo) * 0x1.0p-53;
}

double orc_jrandom_next_gaussian(orc_jrandom *r) {
    if (r->have_next_gaussian) {
        r->have_next_gaussian = 0;
        return r->next_gaussian;
    }
    double v1, v2, s;
    do {
        v1 = 2 * orc_jrandom_next_double(r) - 1;
        v2 = 2 * orc_jrandom_next_double(r) - 1;
        s = v1 * v1 + v2 * v2;
    } while (s >= 1 || s == 0);
    double multiplier = sqrt(-2 * log(s) / s); /* StrictMath in the JDK */
    r->next_gaussian = v2 * multiplier;
    r->have_next_gaussian = 1;
    return v1 * multiplier;
}

uint64_t orc_jrandom_state(const orc_jrandom *r) { return r->seed; }

/* ======================================================================== */
/* scalar functions (L0 interfaces; in-tree instance M:234-235; sigmoid and */
/* half squared error from doc/backprop.pdf)                                */
/* ======================================================================== */
double orc_act(int act, double p, double z) {
    switch (act) {
    case ORC_ACT_LEAKY_RELU: return (z > 0) ? z : p * z;          /* M:234 */
    case ORC_ACT_SIGMOID: return 1.0 / (1.0 + exp(-z));
    case ORC_ACT_RELU: return (z > 0) ? z : 0.0;
    case ORC_ACT_TANH: return tanh(z);
    default: return z;
    }
}

double orc_act_prime(int act, double p, double z) {
    switch (act) {
    case ORC_ACT_LEAKY_RELU: return (z <= 0.0) ? p : 1.0;         /* M:235 */
    case ORC_ACT_SIGMOID: { double s = 1.0 / (1.0 + exp(-z)); return s * (1.0 - s); }
    case ORC_ACT_RELU: return (z <= 0.0) ? 0.0 : 1.0;
    case ORC_ACT_TANH: { double t = tanh(z); return 1.0 - t * t; }
    default: return 1.0;
    }
}

double orc_loss(int loss, double c, double e) {
    (void)loss;
    double d = c - e;
    return 0.5 * (d * d);
}

double orc_loss_prime(int loss, double c, double e) {
    (void)loss;
    return c - e;
}

/* ======================================================================== */
/* the net                                                                  */
/* ======================================================================== */
struct orc_net {
    int32_t n_dims;
    int32_t *dims;
    int32_t kind, inner_act, last_act, loss;
    double inner_param, last_param;
    int64_t n_weights;
    int64_t *w_off;   /* offset of layer l's [in][out] matrix in the flat arrays */
    double *w;        /* weights      (S:15)  */
    double *prev;     /* previousUpdate (S:24) */
    double **neurons; /* pre-activation sums per layer (S:14); [0] aliases the input (TRUST, S:167) */
    int32_t time;     /* S:23 */
    orc_jrandom rng;  /* S:18: init stream, continued by the noise branch */
};

orc_net *orc_net_create(const int32_t *dims, int32_t n_dims, int32_t kind,
                        int32_t inner_act, double inner_param,
                        int32_t last_act, double last_param, int32_t loss,
                        int64_t seed) {
    if (!dims || n_dims < 2) return NULL;
    orc_net *n = (orc_net *)calloc(1, sizeof(orc_net));
    n->n_dims = n_dims;
    n->dims = (int32_t *)malloc(sizeof(int32_t) * n_dims);
    memcpy(n->dims, dims, sizeof(int32_t) * n_dims);
    n->kind = kind; n->inner_act = inner_act; n->inner_param = inner_param;
    n->last_act = last_act; n->last_param = last_param; n->loss = loss;
    n->w_off = (int64_t *)malloc(sizeof(int64_t) * n_dims);
    int64_t tot = 0;
    for (int l = 0; l + 1 < n_dims; l++) { n->w_off[l] = tot; tot += (int64_t)dims[l] * dims[l + 1]; }
    n->w_off[n_dims - 1] = tot;
    n->n_weights = tot;
    n->w = (double *)malloc(sizeof(double) * tot);
    n->prev = (double *)calloc(tot, sizeof(double));         /* S:152 zeros */
    n->neurons = (double **)calloc(n_dims, sizeof(double *));
    for (int l = 1; l < n_dims; l++) n->neurons[l] = (double *)calloc(dims[l], sizeof(double));
    orc_jrandom_init(&n->rng, seed);                          /* S:111 */
    /* appendLayer S:139-156: layer-major, i outer, j inner */
    for (int l = 0; l + 1 < n_dims; l++) {
        double *W = n->w + n->w_off[l];
        for (int i = 0; i < dims[l]; i++)
            for (int j = 0; j < dims[l + 1]; j++)
                W[(int64_t)i * dims[l + 1] + j] = orc_jrandom_next_double(&n->rng) - 0.5; /* S:149 */
    }
    n->time = 0;
    return n;
}

void orc_net_destroy(orc_net *n) {
    if (!n) return;
    for (int l = 1; l < n->n_dims; l++) free(n->neurons[l]);
    free(n->neurons); free(n->w); free(n->prev); free(n->w_off); free(n->dims); free(n);
}

int64_t orc_net_num_weights(const orc_net *n) { return n->n_weights; }
int32_t orc_net_input_dim(const orc_net *n) { return n->dims[0]; }                  /* S:383 */
int32_t orc_net_output_dim(const orc_net *n) { return n->dims[n->n_dims - 1]; }     /* S:392 */
void orc_net_get_weights(const orc_net *n, double *out) { memcpy(out, n->w, sizeof(double) * n->n_weights); }
void orc_net_set_weights(orc_net *n, const double *in) { memcpy(n->w, in, sizeof(double) * n->n_weights); }
void orc_net_get_velocity(const orc_net *n, double *out) { memcpy(out, n->prev, sizeof(double) * n->n_weights); }
void orc_net_set_velocity(orc_net *n, const double *in) { memcpy(n->prev, in, sizeof(double) * n->n_weights); }
uint64_t orc_net_rng_state(const orc_net *n) { return n->rng.seed; }

/* softmaxActivate S:357-376: no max subtraction (Q4); ascending i */
static void softmax_activate(const double *in, int l, double *out) {
    double sum = 0;
    for (int i = 0; i < l; i++) {
        double term = exp(in[i]);
        out[i] = term;
        sum += term;
    }
    for (int i = 0; i < l; i++) out[i] /= sum;
}

/* forward sweep shared by both nets: S:169-193 == G:188-212 */
static void forward_sums(orc_net *n, const double *x) {
    n->neurons[0] = (double *)x; /* TRUST alias, never written (S:167-168) */
    for (int l = 1; l < n->n_dims; l++) {
        const double *prev = n->neurons[l - 1];
        double *cur = n->neurons[l];
        const double *W = n->w + n->w_off[l - 1];
        int np = n->dims[l - 1], nc = n->dims[l];
        for (int i = 0; i < nc; i++) cur[i] = 0.0;                          /* S:180-182 */
        double *pa = (double *)malloc(sizeof(double) * np);                 /* S:183 */
        for (int i = 0; i < np; i++) pa[i] = orc_act(n->inner_act, n->inner_param, prev[i]); /* S:184-186, incl. l-1==0 (Q2) */
        for (int i = 0; i < np; i++)
            for (int j = 0; j < nc; j++)
                cur[j] += pa[i] * W[(int64_t)i * nc + j];                   /* S:187-192 */
        free(pa);
    }
}

void orc_net_propagate(orc_net *n, const double *x, double *y) {
    forward_sums(n, x);
    int L = n->n_dims - 1;
    if (n->kind == ORC_NET_SOFTMAX_CE) {
        softmax_activate(n->neurons[L], n->dims[L], y);                     /* S:197 */
    } else {
        for (int i = 0; i < n->dims[L]; i++)
            y[i] = orc_act(n->last_act, n->last_param, n->neurons[L][i]);   /* G:214-218 */
    }
}

double orc_net_loss(orc_net *n, const double *x, const double *t) {
    int dl = n->dims[n->n_dims - 1];
    double *p = (double *)malloc(sizeof(double) * dl);
    orc_net_propagate(n, x, p);
    double loss = 0;
    if (n->kind == ORC_NET_SOFTMAX_CE) {
        for (int i = 0; i < dl; i++) loss -= t[i] * log(p[i]);              /* S:213-217 (all i, Q11) */
    } else {
        for (int i = 0; i < dl; i++) loss += orc_loss(n->loss, p[i], t[i]); /* G:236-239 */
    }
    free(p);
    return loss;
}

void orc_net_weight_gradient(orc_net *n, const double *x, const double *t, double *grad) {
    int L = n->n_dims - 1;
    int dl = n->dims[L];
    double *out = (double *)malloc(sizeof(double) * dl);
    orc_net_propagate(n, x, out);                                           /* S:237 / G:256 */
    double **dnet = (double **)calloc(n->n_dims, sizeof(double *));
    dnet[L] = (double *)malloc(sizeof(double) * dl);
    /* OUTPUT LAYER CASE */
    if (n->kind == ORC_NET_SOFTMAX_CE) {
        for (int j = 0; j < dl; j++) dnet[L][j] = out[j] - t[j];            /* S:249-251 */
    } else {
        for (int j = 0; j < dl; j++) {
            double ja = orc_act(n->last_act, n->last_param, n->neurons[L][j]);          /* G:269 */
            dnet[L][j] = orc_loss_prime(n->loss, ja, t[j]) *
                         orc_act_prime(n->last_act, n->last_param, n->neurons[L][j]);   /* G:270 */
        }
    }
    {
        const double *in_ = n->neurons[L - 1];
        double *g = grad + n->w_off[L - 1];
        int ni = n->dims[L - 1];
        for (int i = 0; i < ni; i++) {
            double ia = orc_act(n->inner_act, n->inner_param, in_[i]);      /* S:254 */
            for (int j = 0; j < dl; j++) g[(int64_t)i * dl + j] = dnet[L][j] * ia; /* S:255-257 */
        }
    }
    /* INNER LAYER CASE S:260-284 */
    for (int jl = L - 1; jl >= 1; jl--) {
        int il = jl - 1;
        int nj = n->dims[jl], ni = n->dims[il], nk = n->dims[jl + 1];
        dnet[jl] = (double *)malloc(sizeof(double) * nj);
        const double *jn = n->neurons[jl], *in_ = n->neurons[il];
        const double *Wjk = n->w + n->w_off[jl];
        const double *dk = dnet[jl + 1];
        for (int j = 0; j < nj; j++) {
            double sum = 0;
            for (int k = 0; k < nk; k++) sum += Wjk[(int64_t)j * nk + k] * dk[k];       /* S:274-276 */
            dnet[jl][j] = sum * orc_act_prime(n->inner_act, n->inner_param, jn[j]);     /* S:277 */
        }
        double *g = grad + n->w_off[il];
        for (int i = 0; i < ni; i++)
            for (int j = 0; j < nj; j++)
                g[(int64_t)i * nj + j] = dnet[jl][j] * orc_act(n->inner_act, n->inner_param, in_[i]); /* S:281 */
    }
    for (int l = 1; l <= L; l++) free(dnet[l]);
    free(dnet); free(out);
}

/* ---- worker threads (test infrastructure convenience, NOT in the reference, which is serial) --------------------
 * With orc_set_threads(T > 1) the per-sample gradients of a batch (S:305-308) are computed by T threads, each with
 * its own `neurons` scratch, and then summed IN BATCH ORDER (S:309-317) element by element: every floating-point
 * operation and its order per element are those of the serial loop, so the result is bit-identical to it (checked in
 * tests/test_oracle.py).  The timed CPU baseline always runs with 1 thread. */
static int g_threads = 1;
void orc_set_threads(int32_t t) { g_threads = t < 1 ? 1 : t; }
int32_t orc_get_threads(void) { return g_threads; }

/* a shallow copy of the net with private pre-activation scratch (weights are only read by the per-sample pass) */
static void local_net_init(orc_net *dst, const orc_net *src) {
    *dst = *src;
    dst->neurons = (double **)calloc(src->n_dims, sizeof(double *));
    for (int l = 1; l < src->n_dims; l++) dst->neurons[l] = (double *)calloc(src->dims[l], sizeof(double));
}
static void local_net_free(orc_net *n) {
    for (int l = 1; l < n->n_dims; l++) free(n->neurons[l]);
    free(n->neurons);
}

/* static-chunk parallel for over [0, n) on up to nt POSIX threads (the image's gcc has no libgomp) */
typedef void (*par_fn)(int64_t lo, int64_t hi, void *ctx);
typedef struct { par_fn fn; void *ctx; int64_t lo, hi; } par_task;
static void *par_trampoline(void *p) {
    par_task *t = (par_task *)p;
    t->fn(t->lo, t->hi, t->ctx);
    return NULL;
}
static void par_for(int64_t n, int nt, par_fn fn, void *ctx) {
    if (nt > n) nt = (int)n;
    if (nt <= 1) { fn(0, n, ctx); return; }
    pthread_t *th = (pthread_t *)malloc(sizeof(pthread_t) * (size_t)nt);
    par_task *tk = (par_task *)malloc(sizeof(par_task) * (size_t)nt);
    for (int i = 0; i < nt; i++) {
        tk[i].fn = fn; tk[i].ctx = ctx;
        tk[i].lo = n * i / nt; tk[i].hi = n * (i + 1) / nt;
        if (i > 0 && pthread_create(&th[i], NULL, par_trampoline, &tk[i]) != 0) { tk[i].fn(tk[i].lo, tk[i].hi, ctx); tk[i].fn = NULL; }
    }
    par_trampoline(&tk[0]);
    for (int i = 1; i < nt; i++)
        if (tk[i].fn) pthread_join(th[i], NULL);
    free(th); free(tk);
}

typedef struct {
    orc_net *n; const double *x, *t; double *out; int32_t B; double step, momentum;
} par_ctx;

static void par_sample_gradients(int64_t lo, int64_t hi, void *p) {
    par_ctx *c = (par_ctx *)p;
    int d0 = c->n->dims[0], dl = c->n->dims[c->n->n_dims - 1];
    orc_net local;
    local_net_init(&local, c->n);
    for (int64_t b = lo; b < hi; b++)
        orc_net_weight_gradient(&local, c->x + b * d0, c->t + b * dl, c->out + b * c->n->n_weights);
    local_net_free(&local);
}
static void par_sum_and_update(int64_t lo, int64_t hi, void *p) {
    par_ctx *c = (par_ctx *)p;
    orc_net *n = c->n;
    const int64_t P = n->n_weights;
    const int batch_size = c->B;                                            /* batch.size(), S:325 */
    for (int64_t q = lo; q < hi; q++) {
        double a = c->out[q];                                               /* S:319: the first sample's map */
        for (int b = 1; b < c->B; b++) a += c->out[(int64_t)b * P + q];     /* S:309-317, ascending batch order */
        double adj = c->step * a / batch_size + c->momentum * n->prev[q];   /* S:333 */
        n->w[q] -= adj;                                                     /* S:338 */
        n->prev[q] = adj;                                                   /* S:339, S:344 */
    }
}
static void gradient_step_threads(orc_net *n, const double *x, const double *t, int32_t B, double step,
                                  double momentum) {
    int64_t P = n->n_weights;
    par_ctx c = {n, x, t, (double *)malloc(sizeof(double) * (size_t)P * (size_t)B), B, step, momentum};
    par_for(B, g_threads, par_sample_gradients, &c);
    par_for(P, g_threads, par_sum_and_update, &c);
    n->time++;                                                              /* S:343 */
    free(c.out);
}

static void par_loss_rows(int64_t lo, int64_t hi, void *p) {
    par_ctx *c = (par_ctx *)p;
    int d0 = c->n->dims[0], dl = c->n->dims[c->n->n_dims - 1];
    orc_net local;
    local_net_init(&local, c->n);
    for (int64_t i = lo; i < hi; i++) c->out[i] = orc_net_loss(&local, c->x + i * d0, c->t + i * dl);
    local_net_free(&local);
}
/* per-row loss of rows [0, count) into out[] (rows are independent; callers sum in row order) */
void orc_net_loss_rows(orc_net *n, const double *x, const double *t, int64_t count, double *out) {
    par_ctx c = {n, x, t, out, 0, 0.0, 0.0};
    par_for(count, g_threads, par_loss_rows, &c);
}

static void par_propagate_rows(int64_t lo, int64_t hi, void *p) {
    par_ctx *c = (par_ctx *)p;
    int d0 = c->n->dims[0], dl = c->n->dims[c->n->n_dims - 1];
    orc_net local;
    local_net_init(&local, c->n);
    for (int64_t i = lo; i < hi; i++) orc_net_propagate(&local, c->x + i * d0, c->out + i * dl);
    local_net_free(&local);
}
/* propagate rows [0, count) into y [count][dL] */
void orc_net_propagate_rows(orc_net *n, const double *x, int64_t count, double *y) {
    par_ctx c = {n, x, NULL, y, 0, 0.0, 0.0};
    par_for(count, g_threads, par_propagate_rows, &c);
}

void orc_net_gradient_step(orc_net *n, const double *x, const double *t, int32_t B,
                           double step, double momentum, int32_t noise, int32_t faithful) {
    if (g_threads > 1 && !noise && !faithful && B > 1) {
        gradient_step_threads(n, x, t, B, step, momentum);
        return;
    }
    int d0 = n->dims[0], dl = n->dims[n->n_dims - 1];
    int64_t P = n->n_weights;
    double *agg = NULL;
    double *single = faithful ? NULL : (double *)malloc(sizeof(double) * P);
    for (int b = 0; b < B; b++) {                                           /* S:305 (row order replaces keySet order) */
        if (faithful) single = (double *)calloc(P, sizeof(double));        /* fresh zeroed per-sample maps, S:240,264 */
        orc_net_weight_gradient(n, x + (int64_t)b * d0, t + (int64_t)b * dl, single);
        if (agg) {
            for (int64_t q = 0; q < P; q++) agg[q] += single[q];            /* S:309-317 */
            if (faithful) free(single);
        } else {
            if (faithful) { agg = single; }                                 /* S:319 */
            else { agg = (double *)malloc(sizeof(double) * P); memcpy(agg, single, sizeof(double) * P); }
        }
    }
    if (!faithful) free(single);
    int batch_size = B;                                                     /* batch.size(), S:325 */
    for (int64_t q = 0; q < P; q++) {                                       /* S:327-342, layer/j/k order == flat order */
        double adj = step * agg[q] / batch_size + momentum * n->prev[q];    /* S:333 */
        if (noise)
            adj += sqrt(orc_jrandom_next_gaussian(&n->rng) * step / pow(1 + n->time, momentum)); /* S:335 (Q8) */
        n->w[q] -= adj;                                                     /* S:338 */
        agg[q] = adj;                                                       /* S:339 */
    }
    n->time++;                                                              /* S:343 */
    memcpy(n->prev, agg, sizeof(double) * P);                               /* S:344 */
    free(agg);
}
