/*
 * oracle.c -- CPU restatement of the doc-agents semantic-search path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under doc_agents_b200/ may include, link,
 * load or call this file.  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs use it, and only as the checker or as
 * the timed CPU baseline -- never as the product.
 *
 * PARITY UNPINNED.  The arithmetic of the reference path does not live in the
 * reference tree: it is pgvector's `<=>` operator inside Postgres
 * (docker image `ankane/pgvector`, no tag: /root/reference/docker-compose.yml:3),
 * and neither Postgres/pgvector nor a Go toolchain exists in this image.  The
 * reference holds no golden vector, known-answer test or fixture for this path
 * (its only TopK tests use a mock: /root/reference/cmd/query/main_test.go:72-79).
 * What follows restates (a) the SQL semantics visible in the reference and
 * (b) pgvector's published cosine_distance algorithm as recalled from its
 * public source (vector.c, `VectorCosineSimilarity`): three float accumulators
 * over the elements, then  dot / sqrt(na*nb)  in double, clamp to [-1,1],
 * distance = 1 - similarity.
 *
 * What each function follows in the reference:
 *   orc_normalize            internal/embeddings/openai.go:146-158
 *   orc_cosine_distance_*    pgvector `<=>` used at internal/store/postgres.go:233,240,241
 *   orc_topk                 internal/store/postgres.go:218-285 (filter :239, floor :223/:240,
 *                            order :241, limit :242, float32 narrowing :259/:263)
 *   orc_upsert_*             internal/store/postgres.go:176-201 (ON CONFLICT (chunk_id) DO UPDATE)
 *   orc_avg_similarity       cmd/query/main.go:161-170
 *
 * The synthetic-corpus generator (orc_gen_*) is this repository's own test
 * data specification; it is integer-exact so that the
 * CUDA generator reproduces it bit for bit (see DESIGN.md "Synthetic corpus").
 *
 * Build: see oracle/Makefile.  With -DORC_FAST the scan entry points are
 * emitted with an `_fast` suffix; that translation unit is compiled with
 * pgvector-style vectorising flags and OpenMP and is what bench.py times.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#ifdef ORC_FAST
#define ORC_NAME(x) x##_fast
#else
#define ORC_NAME(x) x
#endif

/* ------------------------------------------------------------------ */
/* Synthetic corpus specification (integer-exact)                      */
/* ------------------------------------------------------------------ */
#ifndef ORC_FAST

#define ORC_GOLD 0x9E3779B97F4A7C15ULL
#define ORC_PAIR 0xD1B54A32D192ED03ULL

static inline uint64_t orc_mix64(uint64_t z)
{
    z ^= z >> 30; z *= 0xBF58476D1CE4E5B9ULL;
    z ^= z >> 27; z *= 0x94D049BB133111EBULL;
    z ^= z >> 31;
    return z;
}

/* One 64-bit hash serves two adjacent columns (lo word = even col). */
static inline uint64_t orc_stream_key(uint64_t seed, uint64_t index)
{
    return orc_mix64(seed ^ ((index + 1) * ORC_GOLD));
}

/* Irwin-Hall(4) on the four bytes of a 32-bit word, centred: [-510, 510]. */
static inline int32_t orc_ih4(uint32_t w)
{
    return (int32_t)((w & 255u) + ((w >> 8) & 255u) + ((w >> 16) & 255u) + (w >> 24)) - 510;
}

static inline int32_t orc_elem(uint64_t key, uint32_t col)
{
    uint64_t h = orc_mix64(key + (uint64_t)((col >> 1) + 1) * ORC_PAIR);
    uint32_t w = (col & 1u) ? (uint32_t)(h >> 32) : (uint32_t)h;
    return orc_ih4(w);
}

/* u[col] = cmul * centroid[cluster][col] + nmul * noise[stream][col], all int32.
 * x = float(u) / float(sqrt(double(sum u^2)))  -- the normalisation of
 * openai.go:146-158 with the float64 sum of squares computed exactly. */
#define ORC_SUPER_TAG 0x5C0DE5C0DE5C0DE5ULL   /* seed_centroid ^ this = the stream of group directions */

/* Optional second level (smul != 0 and super_group != 0): clusters are grouped super_group at a time and a vector
 * also carries smul x the direction of its cluster's group, so the clusters of one group overlap. */
static void orc_gen_vec(uint64_t seed_noise, uint64_t noise_index, uint64_t seed_centroid,
                        uint64_t cluster, int32_t cmul, int32_t nmul, int32_t smul, uint32_t super_group,
                        int dim, float *out)
{
    uint64_t kn = orc_stream_key(seed_noise, noise_index);
    uint64_t kc = orc_stream_key(seed_centroid, cluster);
    int two_level = smul != 0 && super_group != 0;
    uint64_t ks = two_level ? orc_stream_key(seed_centroid ^ ORC_SUPER_TAG, cluster / super_group) : 0;
    int64_t ss = 0;
    for (int c = 0; c < dim; c++) {
        int32_t u = cmul * orc_elem(kc, (uint32_t)c) + nmul * orc_elem(kn, (uint32_t)c);
        if (two_level) u += smul * orc_elem(ks, (uint32_t)c);
        ss += (int64_t)u * (int64_t)u;
        out[c] = (float)u;
    }
    if (ss == 0) return;                       /* openai.go:151-153 */
    float mag = (float)sqrt((double)ss);       /* openai.go:154 */
    for (int c = 0; c < dim; c++) out[c] /= mag;   /* openai.go:155-157 */
}

typedef struct {
    uint64_t seed_data, seed_centroid, seed_query, seed_ids;
    uint32_t chunks_per_doc, n_clusters;
    int32_t cmul, nmul;
    int32_t smul; uint32_t super_group;
} orc_synth;

uint32_t orc_row_doc(const orc_synth *p, uint64_t row) { return (uint32_t)(row / p->chunks_per_doc); }
uint32_t orc_doc_cluster(const orc_synth *p, uint32_t doc) { return doc % p->n_clusters; }

void orc_gen_rows(const orc_synth *p, uint64_t row_lo, uint64_t n, int dim, float *out)
{
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < (int64_t)n; i++) {
        uint64_t row = row_lo + (uint64_t)i;
        uint32_t cl = orc_doc_cluster(p, orc_row_doc(p, row));
        orc_gen_vec(p->seed_data, row, p->seed_centroid, cl, p->cmul, p->nmul, p->smul, p->super_group, dim,
                    out + (size_t)i * dim);
    }
}

/* The same generator for an arbitrary list of global rows (the members of probed IVFFlat lists, say). */
void orc_gen_rows_at(const orc_synth *p, const uint64_t *rows, uint64_t n, int dim, float *out)
{
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < (int64_t)n; i++) {
        uint32_t cl = orc_doc_cluster(p, orc_row_doc(p, rows[i]));
        orc_gen_vec(p->seed_data, rows[i], p->seed_centroid, cl, p->cmul, p->nmul, p->smul, p->super_group, dim,
                    out + (size_t)i * dim);
    }
}

void orc_gen_query(const orc_synth *p, uint64_t qi, uint32_t cluster, int dim, float *out)
{
    orc_gen_vec(p->seed_query, qi, p->seed_centroid, cluster, p->cmul, p->nmul, p->smul, p->super_group, dim, out);
}

/* 16-byte ids: two hashes, RFC 4122 version-4 / variant bits forced so they
 * parse as the uuid4 values the reference validates (cmd/query/main.go:22). */
static void orc_id16(uint64_t seed, uint64_t tag, uint64_t index, uint8_t out[16])
{
    uint64_t a = orc_mix64(seed ^ orc_mix64(tag + 2 * index));
    uint64_t b = orc_mix64(seed ^ orc_mix64(tag + 2 * index + 1));
    for (int i = 0; i < 8; i++) { out[i] = (uint8_t)(a >> (8 * i)); out[8 + i] = (uint8_t)(b >> (8 * i)); }
    out[6] = (uint8_t)((out[6] & 0x0F) | 0x40);
    out[8] = (uint8_t)((out[8] & 0x3F) | 0x80);
}
void orc_chunk_id(const orc_synth *p, uint64_t row, uint8_t out[16]) { orc_id16(p->seed_ids, 0xC4A9000000000000ULL, row, out); }
void orc_doc_id(const orc_synth *p, uint64_t doc, uint8_t out[16])   { orc_id16(p->seed_ids, 0xD0C0000000000000ULL, doc, out); }

/* internal/embeddings/openai.go:146-158 */
void orc_normalize(float *v, int dim)
{
    double ss = 0.0;
    for (int i = 0; i < dim; i++) ss += (double)v[i] * (double)v[i];
    if (ss == 0.0) return;
    float mag = (float)sqrt(ss);
    for (int i = 0; i < dim; i++) v[i] /= mag;
}

/* cmd/query/main.go:161-170: float32 running sum in result order, / n. */
float orc_avg_similarity(const float *scores, int n)
{
    if (n == 0) return 0.0f;
    float total = 0.0f;
    for (int i = 0; i < n; i++) total += scores[i];
    return total / (float)n;
}
#endif /* !ORC_FAST */

/* ------------------------------------------------------------------ */
/* pgvector `<=>` restated                                             */
/* ------------------------------------------------------------------ */

/* float accumulators, element order, double finish, clamp, 1 - sim. */
static inline double ORC_NAME(orc_cosdist_f32)(const float *a, const float *b, int dim)
{
    float dot = 0.0f, na = 0.0f, nb = 0.0f;
    for (int i = 0; i < dim; i++) {
        dot += a[i] * b[i];
        na += a[i] * a[i];
        nb += b[i] * b[i];
    }
    double sim = (double)dot / sqrt((double)na * (double)nb);
    if (sim > 1.0) sim = 1.0; else if (sim < -1.0) sim = -1.0;   /* NaN falls through */
    return 1.0 - sim;
}

#ifndef ORC_FAST
double orc_cosine_distance_f32(const float *a, const float *b, int dim) { return orc_cosdist_f32(a, b, dim); }

/* Error-bound truth: same formula with double accumulators. */
double orc_cosine_distance_f64(const float *a, const float *b, int dim)
{
    double dot = 0.0, na = 0.0, nb = 0.0;
    for (int i = 0; i < dim; i++) {
        dot += (double)a[i] * (double)b[i];
        na += (double)a[i] * (double)a[i];
        nb += (double)b[i] * (double)b[i];
    }
    double sim = dot / sqrt(na * nb);
    if (sim > 1.0) sim = 1.0; else if (sim < -1.0) sim = -1.0;
    return 1.0 - sim;
}
#endif

/* ------------------------------------------------------------------ */
/* Store.TopK restated (internal/store/postgres.go:218-285)            */
/* ------------------------------------------------------------------ */
typedef struct { double dist; uint32_t row; } orc_cand;

static inline int orc_cand_before(const orc_cand *x, const orc_cand *y)
{
    /* ORDER BY distance ascending (postgres.go:241); Postgres leaves ties
     * unordered, this repository fixes them as lower row first. */
    if (x->dist < y->dist) return 1;
    if (x->dist > y->dist) return 0;
    return x->row < y->row;
}

/* Bounded insertion into a sorted best-first list of at most k. */
static inline void orc_push(orc_cand *best, int *n, int k, orc_cand c)
{
    if (*n == k && !orc_cand_before(&c, &best[k - 1])) return;
    int i = (*n < k) ? (*n)++ : k - 1;
    while (i > 0 && orc_cand_before(&c, &best[i - 1])) { best[i] = best[i - 1]; i--; }
    best[i] = c;
}

/*
 * mat        n x dim row-major float32 (bit-identical to what SaveEmbeddings stored,
 *            since vectorToString is lossless: postgres.go:321-330)
 * row_doc    document index of each row (the JOIN chunks ... document_id, postgres.go:237)
 * doc_allowed  byte per document index, 1 = document_id = ANY($2) (postgres.go:239);
 *            NULL means every document is requested.
 * use_f64    0: pgvector float accumulators; 1: double accumulators (truth)
 * min_sim    the similarity floor, a float8 parameter (0.7 at postgres.go:223)
 * Outputs best-first: rows, score = (float)(1 - dist) (postgres.go:233,259), dist.
 * Returns the number of results (<= k; 0 is the reference's nil slice, postgres.go:250).
 * Rows whose distance is NaN (zero-norm row or query) are skipped: see DESIGN.md
 * "Deviations" (Postgres would keep them: NaN >= 0.7 is true there).
 */
int ORC_NAME(orc_topk)(const float *mat, uint64_t n, int dim, const uint32_t *row_doc,
                       const uint8_t *doc_allowed, const float *q, int k, double min_sim,
                       int use_f64, int n_threads, uint32_t *out_rows, float *out_scores,
                       double *out_dist)
{
    if (k <= 0) return 0;
#ifdef _OPENMP
    int T = n_threads > 0 ? n_threads : omp_get_max_threads();
#else
    int T = 1; (void)n_threads;
#endif
    orc_cand *lists = (orc_cand *)malloc(sizeof(orc_cand) * (size_t)k * (size_t)T);
    int *counts = (int *)calloc((size_t)T, sizeof(int));
#pragma omp parallel num_threads(T)
    {
#ifdef _OPENMP
        int t = omp_get_thread_num();
#else
        int t = 0;
#endif
        orc_cand *best = lists + (size_t)t * k;
        int cnt = 0;
#pragma omp for schedule(static)
        for (int64_t r = 0; r < (int64_t)n; r++) {
            if (doc_allowed && !doc_allowed[row_doc[r]]) continue;
            const float *a = mat + (size_t)r * dim;
            double d;
#ifndef ORC_FAST
            if (use_f64) d = orc_cosine_distance_f64(a, q, dim); else
#endif
            d = ORC_NAME(orc_cosdist_f32)(a, q, dim);
            if (d != d) continue;
            if (!((1.0 - d) >= min_sim)) continue;          /* postgres.go:240 */
            orc_cand c = { d, (uint32_t)r };
            orc_push(best, &cnt, k, c);
        }
        counts[t] = cnt;
    }
    (void)use_f64;
    orc_cand *fin = (orc_cand *)malloc(sizeof(orc_cand) * (size_t)k);
    int nf = 0;
    for (int t = 0; t < T; t++)
        for (int i = 0; i < counts[t]; i++) orc_push(fin, &nf, k, lists[(size_t)t * k + i]);
    for (int i = 0; i < nf; i++) {
        out_rows[i] = fin[i].row;
        out_scores[i] = (float)(1.0 - fin[i].dist);          /* postgres.go:233 then :259 */
        if (out_dist) out_dist[i] = fin[i].dist;
    }
    free(fin); free(lists); free(counts);
    return nf;
}

#ifndef ORC_FAST
/* All similarities of one query (float32-narrowed), for dense checks. */
void orc_all_scores(const float *mat, uint64_t n, int dim, const float *q, int use_f64, float *out)
{
#pragma omp parallel for schedule(static)
    for (int64_t r = 0; r < (int64_t)n; r++) {
        const float *a = mat + (size_t)r * dim;
        double d = use_f64 ? orc_cosine_distance_f64(a, q, dim) : orc_cosine_distance_f32(a, q, dim);
        out[r] = (float)(1.0 - d);
    }
}

/* ------------------------------------------------------------------ */
/* SaveEmbeddings restated (internal/store/postgres.go:176-201)        */
/* A tiny table keyed by chunk_id: insert, or overwrite on conflict.   */
/* ------------------------------------------------------------------ */
typedef struct {
    int dim; uint64_t n, cap;
    uint8_t *ids;      /* n x 16 */
    float *vecs;       /* n x dim */
} orc_table;

orc_table *orc_table_new(int dim)
{
    orc_table *t = (orc_table *)calloc(1, sizeof(orc_table));
    t->dim = dim;
    return t;
}
void orc_table_free(orc_table *t) { if (t) { free(t->ids); free(t->vecs); free(t); } }
uint64_t orc_table_rows(const orc_table *t) { return t->n; }
const float *orc_table_vecs(const orc_table *t) { return t->vecs; }
const uint8_t *orc_table_ids(const orc_table *t) { return t->ids; }

/* Returns 0, or -1 on dimension mismatch (pgvector: "expected N dimensions"). */
int orc_table_upsert(orc_table *t, const uint8_t *chunk_ids, const float *vecs, uint64_t n, int dim)
{
    if (n == 0) return 0;                                   /* postgres.go:177-179 */
    if (dim != t->dim) return -1;
    for (uint64_t i = 0; i < n; i++) {
        uint64_t r = t->n;
        for (uint64_t j = 0; j < t->n; j++)
            if (memcmp(t->ids + 16 * j, chunk_ids + 16 * i, 16) == 0) { r = j; break; }
        if (r == t->n) {
            if (t->n == t->cap) {
                t->cap = t->cap ? 2 * t->cap : 64;
                t->ids = (uint8_t *)realloc(t->ids, 16 * t->cap);
                t->vecs = (float *)realloc(t->vecs, sizeof(float) * (size_t)dim * t->cap);
            }
            memcpy(t->ids + 16 * r, chunk_ids + 16 * i, 16);
            t->n++;
        }
        memcpy(t->vecs + (size_t)r * dim, vecs + (size_t)i * dim, sizeof(float) * (size_t)dim);  /* :197 */
    }
    return 0;
}

int orc_num_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* Threads used by the parallel loops that take no explicit count (the corpus generator).  Harness plumbing:
 * torchrun presets OMP_NUM_THREADS=1 for its workers. */
void orc_set_num_threads(int n)
{
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}
#endif /* !ORC_FAST */
