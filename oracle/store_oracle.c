/*
 * store_oracle.c -- TEST INFRASTRUCTURE ONLY.
 *
 * A CPU restatement, in plain C, of the reference `local-store` backend
 * (/root/reference/backend/go/stores/store.go) and of the Go standard-library
 * routines it leans on (container/heap, slices.Compare / cmp.Compare,
 * slices.BinarySearchFunc).  It exists so that tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs have something to check
 * and time the CUDA path against.  Nothing under localai_tfp_b200/ may import,
 * link or call it.
 *
 * Parity status: the reference is Go and no Go toolchain exists in the build
 * image, so this file cannot be validated against reference *outputs*.  It is
 * pinned against every numeric expectation the reference's own tests hold for
 * this path (tests/integration/stores_test.go, core/http/app_test.go:904-974);
 * see tests/test_oracle_kat.py.
 *
 * Build: gcc -O2 -ffp-contract=off -fno-fast-math (no FMA contraction, which is
 * what Go does on amd64 with GOAMD64=v1) plus -fno-tree-vectorize so the inner
 * loops stay scalar like Go's code generator (see oracle/Makefile).
 *
 * Each function cites the store.go lines it follows.
 */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define ORC_ERRLEN 256

typedef struct {
    /* store.go:17-30 -- sorted keys, parallel values, sticky flag, keyLen */
    float **keys;       /* each row separately allocated, as Go's [][]float32 */
    int64_t *values;    /* opaque value handles (the tests map them to bytes) */
    int64_t n;
    int keys_are_normalized;
    int key_len;
    char err[ORC_ERRLEN];
} orc_store;

/* ---- Go stdlib restatements ------------------------------------------- */

/* cmp.Compare[float32]: NaN sorts lowest, NaN == NaN, -0 == +0. */
static int go_cmp_f32(float x, float y) {
    int xn = (x != x), yn = (y != y);
    if (xn) return yn ? 0 : -1;
    if (yn) return +1;
    if (x < y) return -1;
    if (x > y) return +1;
    return 0;
}

/* slices.Compare for equal-length slices (store.go:48-52 compareSlices). */
static int compare_slices(const float *a, const float *b, int d) {
    for (int i = 0; i < d; i++) {
        int c = go_cmp_f32(a[i], b[i]);
        if (c != 0) return c;
    }
    return 0;
}

/* slices.BinarySearchFunc: leftmost position; found iff equal there
 * (store.go:60-64 findInSortedSlice). */
static int64_t find_in_sorted(float **keys, int64_t n, const float *t, int d, int *found) {
    int64_t lo = 0, hi = n;
    while (lo < hi) {
        int64_t h = (int64_t)(((uint64_t)lo + (uint64_t)hi) >> 1);
        if (compare_slices(keys[h], t, d) < 0) lo = h + 1; else hi = h;
    }
    *found = (lo < n && compare_slices(keys[lo], t, d) == 0);
    return lo;
}

typedef struct { float *key; int64_t value; } orc_pair;
static int g_sort_dim; /* qsort has no context argument; the oracle is single-threaded */
static int pair_cmp(const void *a, const void *b) {
    return compare_slices(((const orc_pair *)a)->key, ((const orc_pair *)b)->key, g_sort_dim);
}
static int keyptr_cmp(const void *a, const void *b) {
    return compare_slices(*(float *const *)a, *(float *const *)b, g_sort_dim);
}

/* ---- lifecycle --------------------------------------------------------- */

orc_store *orc_new(void) {
    /* store.go:39-46 NewStore */
    orc_store *s = (orc_store *)calloc(1, sizeof(orc_store));
    s->keys_are_normalized = 1;
    s->key_len = -1;
    return s;
}

void orc_free(orc_store *s) {
    if (!s) return;
    for (int64_t i = 0; i < s->n; i++) free(s->keys[i]);
    free(s->keys);
    free(s->values);
    free(s);
}

const char *orc_last_error(orc_store *s) { return s->err; }
int64_t orc_len(orc_store *s) { return s->n; }
int orc_key_len(orc_store *s) { return s->key_len; }
int orc_keys_are_normalized(orc_store *s) { return s->keys_are_normalized; }

/* Copy row i (sorted order) out; lets tests look at the store's order. */
void orc_row(orc_store *s, int64_t i, float *out_key, int64_t *out_value) {
    memcpy(out_key, s->keys[i], sizeof(float) * (size_t)s->key_len);
    *out_value = s->values[i];
}

/* ---- arithmetic -------------------------------------------------------- */

/* store.go:313-324 isNormalized: fp64 sum of exact squares, 0.99 <= sqrt <= 1.01 */
int orc_is_normalized(const float *k, int d) {
    double sum = 0.0;
    for (int i = 0; i < d; i++) {
        double v = (double)k[i];
        sum += v * v;
    }
    double s = sqrt(sum);
    return s >= 0.99 && s <= 1.01;
}

/* store.go:327-339 normalizedCosineSimilarity: fp32 accumulator, index order. */
float orc_dot_f32(const float *k1, const float *k2, int d) {
    float dot = 0.0f; /* -ffp-contract=off keeps product and sum separately rounded */
    for (int i = 0; i < d; i++) {
        float prod = k1[i] * k2[i];
        dot = dot + prod;
    }
    return dot;
}

/* store.go:435-439: mag1 = sqrt(sum float64(float32(v*v))) */
double orc_query_mag(const float *q, int d) {
    double m = 0.0;
    for (int i = 0; i < d; i++) {
        float p = q[i] * q[i];
        m += (double)p;
    }
    return sqrt(m);
}

/* store.go:414-428 cosineSimilarity: fp32-rounded products, fp64 accumulation. */
float orc_cosine(const float *k1, const float *k2, int d, double mag1) {
    double dot = 0.0, mag2 = 0.0;
    for (int i = 0; i < d; i++) {
        float p = k1[i] * k2[i];
        float s = k2[i] * k2[i];
        dot += (double)p;
        mag2 += (double)s;
    }
    return (float)(dot / (mag1 * sqrt(mag2)));
}

/* ---- container/heap over (similarity, row) ------------------------------ */

typedef struct { float sim; int64_t row; } orc_item;
typedef struct { orc_item *a; int64_t n; } orc_heap;

/* store.go:351-354 Less: strict < on Similarity (false for any NaN operand). */
static int h_less(orc_heap *h, int64_t i, int64_t j) { return h->a[i].sim < h->a[j].sim; }
static void h_swap(orc_heap *h, int64_t i, int64_t j) { orc_item t = h->a[i]; h->a[i] = h->a[j]; h->a[j] = t; }
static void h_up(orc_heap *h, int64_t j) {
    for (;;) {
        int64_t i = (j - 1) / 2; /* parent; Go integer division truncates toward zero */
        if (i == j || !h_less(h, j, i)) break;
        h_swap(h, i, j);
        j = i;
    }
}
static void h_down(orc_heap *h, int64_t i0, int64_t n) {
    int64_t i = i0;
    for (;;) {
        int64_t j1 = 2 * i + 1;
        if (j1 >= n || j1 < 0) break;
        int64_t j = j1, j2 = j1 + 1;
        if (j2 < n && h_less(h, j2, j1)) j = j2;
        if (!h_less(h, j, i)) break;
        h_swap(h, i, j);
        i = j;
    }
}
static void h_push(orc_heap *h, orc_item x) { h->a[h->n++] = x; h_up(h, h->n - 1); }
static orc_item h_pop(orc_heap *h) {
    int64_t n = h->n - 1;
    h_swap(h, 0, n);
    h_down(h, 0, n);
    h->n = n;
    return h->a[n];
}

/* ---- StoresSet (store.go:106-198) --------------------------------------- */

int orc_set(orc_store *s, const float *keys, int64_t n, int dim, const int64_t *values, int64_t n_values) {
    if (n == 0) { snprintf(s->err, ORC_ERRLEN, "no keys to add"); return 1; }
    if (n != n_values) {
        snprintf(s->err, ORC_ERRLEN, "len(keys) = %lld, len(values) = %lld", (long long)n, (long long)n_values);
        return 1;
    }
    if (s->key_len == -1) s->key_len = dim;
    else if (dim != s->key_len) {
        snprintf(s->err, ORC_ERRLEN, "Try to add key with length %d when existing length is %d", dim, s->key_len);
        return 1;
    }
    orc_pair *kvs = (orc_pair *)malloc(sizeof(orc_pair) * (size_t)n);
    for (int64_t i = 0; i < n; i++) {
        const float *k = keys + (size_t)i * (size_t)dim;
        if (s->keys_are_normalized && !orc_is_normalized(k, dim)) s->keys_are_normalized = 0;
        kvs[i].key = (float *)malloc(sizeof(float) * (size_t)(dim > 0 ? dim : 1));
        memcpy(kvs[i].key, k, sizeof(float) * (size_t)dim);
        kvs[i].value = values[i];
    }
    g_sort_dim = dim;
    qsort(kvs, (size_t)n, sizeof(orc_pair), pair_cmp); /* slices.SortFunc: not stable either */

    int64_t l = n + s->n;
    float **mk = (float **)malloc(sizeof(float *) * (size_t)(l > 0 ? l : 1));
    int64_t *mv = (int64_t *)malloc(sizeof(int64_t) * (size_t)(l > 0 ? l : 1));
    int64_t m = 0, i = 0, j = 0;
    for (;;) {
        if (i + j >= l) break;
        if (i >= n) { mk[m] = s->keys[j]; mv[m] = s->values[j]; m++; j++; continue; }
        if (j >= s->n) { mk[m] = kvs[i].key; mv[m] = kvs[i].value; m++; i++; continue; }
        int c = compare_slices(kvs[i].key, s->keys[j], dim);
        if (c < 0) { mk[m] = kvs[i].key; mv[m] = kvs[i].value; m++; i++; }
        else if (c > 0) { mk[m] = s->keys[j]; mv[m] = s->values[j]; m++; j++; }
        else { /* equal: the new pair replaces the stored one (store.go:183-187) */
            mk[m] = kvs[i].key; mv[m] = kvs[i].value; m++;
            free(s->keys[j]);
            i++; j++;
        }
    }
    /* i+j reaches l only when both inputs are exhausted (i <= n, j <= s->n), so
     * nothing is dropped; m < l whenever a stored key was replaced, which is why
     * the reference's debug-only assert at store.go:191 would fire on overwrite. */
    free(kvs);
    free(s->keys); free(s->values);
    s->keys = mk; s->values = mv; s->n = m;
    return 0;
}

/* ---- StoresDelete (store.go:200-264) ------------------------------------ */

int orc_delete(orc_store *s, const float *keys, int64_t n, int dim) {
    if (n == 0) { snprintf(s->err, ORC_ERRLEN, "no keys to delete"); return 1; }
    if (s->key_len == -1) s->key_len = dim;
    else if (dim != s->key_len) {
        snprintf(s->err, ORC_ERRLEN, "Trying to delete key with length %d when existing length is %d", dim, s->key_len);
        return 1;
    }
    const float **ks = (const float **)malloc(sizeof(float *) * (size_t)n);
    for (int64_t i = 0; i < n; i++) ks[i] = keys + (size_t)i * (size_t)dim;
    g_sort_dim = dim;
    qsort(ks, (size_t)n, sizeof(float *), keyptr_cmp);

    float **mk = (float **)malloc(sizeof(float *) * (size_t)(s->n > 0 ? s->n : 1));
    int64_t *mv = (int64_t *)malloc(sizeof(int64_t) * (size_t)(s->n > 0 ? s->n : 1));
    int64_t m = 0;
    float **tail_k = s->keys; int64_t *tail_v = s->values; int64_t tail_n = s->n;
    for (int64_t i = 0; i < n; i++) {
        int found;
        int64_t j = find_in_sorted(tail_k, tail_n, ks[i], dim, &found);
        if (found) {
            memcpy(mk + m, tail_k, sizeof(float *) * (size_t)j);
            memcpy(mv + m, tail_v, sizeof(int64_t) * (size_t)j);
            m += j;
            free(tail_k[j]);
            tail_k += j + 1; tail_v += j + 1; tail_n -= j + 1;
        }
    }
    memcpy(mk + m, tail_k, sizeof(float *) * (size_t)tail_n);
    memcpy(mv + m, tail_v, sizeof(int64_t) * (size_t)tail_n);
    m += tail_n;
    free(ks);
    free(s->keys); free(s->values);
    s->keys = mk; s->values = mv; s->n = m;
    return 0;
}

/* ---- StoresGet (store.go:266-311) ---------------------------------------
 * out_keys: up to n rows of dim floats; out_values: up to n handles.
 * The reference indexes opts.Keys[0] with no emptiness check (a panic on an
 * empty request, store.go:276); the restatement reports that as return code 2. */
int orc_get(orc_store *s, const float *keys, int64_t n, int dim,
            float *out_keys, int64_t *out_values, int64_t *out_n) {
    *out_n = 0;
    if (n == 0) { snprintf(s->err, ORC_ERRLEN, "panic: index out of range [0] with length 0"); return 2; }
    const float **ks = (const float **)malloc(sizeof(float *) * (size_t)n);
    for (int64_t i = 0; i < n; i++) ks[i] = keys + (size_t)i * (size_t)dim;
    g_sort_dim = dim;
    qsort(ks, (size_t)n, sizeof(float *), keyptr_cmp);
    if (s->key_len == -1) s->key_len = dim;
    else if (dim != s->key_len) {
        snprintf(s->err, ORC_ERRLEN, "Try to get a key with length %d when existing length is %d", dim, s->key_len);
        free(ks);
        return 1;
    }
    float **tail_k = s->keys; int64_t *tail_v = s->values; int64_t tail_n = s->n;
    int64_t m = 0;
    for (int64_t i = 0; i < n; i++) {
        int found;
        int64_t j = find_in_sorted(tail_k, tail_n, ks[i], dim, &found);
        if (found) {
            memcpy(out_keys + (size_t)m * (size_t)dim, ks[i], sizeof(float) * (size_t)dim); /* the REQUEST's key, store.go:289-291 */
            out_values[m] = tail_v[j];
            m++;
            tail_k += j + 1; tail_v += j + 1; tail_n -= j + 1;
        }
    }
    free(ks);
    *out_n = m;
    return 0;
}

/* ---- StoresFind (store.go:477-511, 373-412, 430-475) ---------------------
 * out_rows receives the sorted-order row index of each hit (so that a test can
 * fetch key and value with orc_row); out_path: 1 = normalised, 0 = fallback.
 * `faithful` != 0 allocates one heap item per key like store.go:380-384 (used
 * only by the timed CPU baseline; results are identical). */
int orc_find(orc_store *s, const float *q, int dim, int topk, int faithful,
             int64_t *out_rows, float *out_sims, int64_t *out_n, int *out_path) {
    *out_n = 0;
    if (dim != s->key_len) {
        snprintf(s->err, ORC_ERRLEN, "Try to find key with length %d when existing length is %d", dim, s->key_len);
        return 1;
    }
    if (topk < 1) { snprintf(s->err, ORC_ERRLEN, "opts.TopK = %d, must be >= 1", topk); return 1; }
    int normalized = s->keys_are_normalized && orc_is_normalized(q, dim);
    if (out_path) *out_path = normalized;

    orc_heap h;
    int64_t cap = (int64_t)topk + 1;
    if (cap > s->n + 1) cap = s->n + 1;
    h.a = (orc_item *)malloc(sizeof(orc_item) * (size_t)cap);
    h.n = 0;
    double mag1 = normalized ? 0.0 : orc_query_mag(q, dim);
    for (int64_t i = 0; i < s->n; i++) {
        float sim = normalized ? orc_dot_f32(q, s->keys[i], dim) : orc_cosine(q, s->keys[i], dim, mag1);
        orc_item it; it.sim = sim; it.row = i;
        if (faithful) { /* &PriorityItem{...}: one small allocation per key per query */
            orc_item *p = (orc_item *)malloc(sizeof(orc_item) + 40);
            *p = it; it = *p; free(p);
        }
        h_push(&h, it);
        if (h.n > (int64_t)topk) (void)h_pop(&h);
    }
    int64_t m = h.n;
    for (int64_t i = m - 1; i >= 0; i--) { /* drain: worst first, filled back to front */
        orc_item it = h_pop(&h);
        out_sims[i] = it.sim;
        out_rows[i] = it.row;
    }
    free(h.a);
    *out_n = m;
    return 0;
}

/* ---- contiguous scan for timing and bulk parity -------------------------
 * Scores every row of a contiguous [n x dim] block in the reference arithmetic
 * and keeps the top-k through the same heap.  Scan order is the given row
 * order (the caller passes rows in the order it wants ties resolved). */
int orc_scan_block(const float *keys, int64_t n, int dim, const float *q, int topk, int normalized,
                   int64_t *out_rows, float *out_sims, int64_t *out_n) {
    orc_heap h;
    int64_t cap = (int64_t)topk + 1;
    if (cap > n + 1) cap = n + 1;
    h.a = (orc_item *)malloc(sizeof(orc_item) * (size_t)cap);
    h.n = 0;
    double mag1 = normalized ? 0.0 : orc_query_mag(q, dim);
    for (int64_t i = 0; i < n; i++) {
        const float *k = keys + (size_t)i * (size_t)dim;
        orc_item it;
        it.sim = normalized ? orc_dot_f32(q, k, dim) : orc_cosine(q, k, dim, mag1);
        it.row = i;
        h_push(&h, it);
        if (h.n > (int64_t)topk) (void)h_pop(&h);
    }
    int64_t m = h.n;
    for (int64_t i = m - 1; i >= 0; i--) { orc_item it = h_pop(&h); out_sims[i] = it.sim; out_rows[i] = it.row; }
    free(h.a);
    *out_n = m;
    return 0;
}

/* All scores of a block (no selection): lets tests compare every score the
 * GPU produced, not only the winners. */
void orc_scores_block(const float *keys, int64_t n, int dim, const float *q, int normalized, float *out) {
    double mag1 = normalized ? 0.0 : orc_query_mag(q, dim);
    for (int64_t i = 0; i < n; i++) {
        const float *k = keys + (size_t)i * (size_t)dim;
        out[i] = normalized ? orc_dot_f32(q, k, dim) : orc_cosine(q, k, dim, mag1);
    }
}
