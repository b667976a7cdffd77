/*
 * oracle.c -- CPU restatement of the retrieval + level-scheduling hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product path:
 * only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
 * reference legs may build, load or call it, and only as the checker.
 *
 * PARITY STATUS
 *   retrieval (distance + telemetry penalty + top-k):  **parity unpinned**.
 *     The reference never issues a ranked query: control_plane.py:51-55 is an
 *     unranked `SELECT name, input_schema_vector FROM service_schemas;` and is
 *     dead code; it has no tests and no golden vectors.  The semantics restated
 *     here are those of the pgvector server extension (third-party C, not under
 *     /root/reference, version unpinned by README.md:23 / :61): `<=>` cosine
 *     distance and `<->` L2 distance, exact scan, ascending order.  SURVEY.md
 *     section 8(c) fixes what pgvector leaves open: fp64 accumulation in index
 *     order, ties -> lowest service id, NaN (zero vector under `<=>`) sorts last.
 *   DAG levels: pinned against NetworkX 3.6.1 topological_generations
 *     (networkx/algorithms/dag.py:166-240), which control_plane.py:104 reaches
 *     through nx.topological_sort (dag.py:308-309).  tests/test_oracle_dag.py
 *     checks this file against the live NetworkX and against tests/golden/.
 *
 * Build: gcc -O2 -ffp-contract=off -shared -fPIC (no libgomp in this image: threading is done by the Python caller) (see oracle/Makefile).
 * -ffp-contract=off matters: the L2 path must round (a-b), (a-b)^2 and the sum
 * separately, exactly like the CUDA rescoring kernel (__dsub_rn/__dmul_rn/__dadd_rn).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define ORACLE_COSINE 0
#define ORACLE_L2 1

/* sum_i a_i*a_i, fp64, index order.  fp32*fp32 is exact in fp64. */
static double norm2_f64(const float *a, int dim) {
    double s = 0.0;
    for (int i = 0; i < dim; ++i) {
        double x = (double)a[i];
        s = s + x * x;
    }
    return s;
}

/*
 * One (query,row) distance with pgvector semantics, fp64 throughout.
 *   cosine (vector.c cosine_distance, `<=>`): 1 - clamp(dot / sqrt(na*nb), -1, 1);
 *     the clamp is two comparisons, so NaN (na*nb == 0) survives it.
 *   l2 (vector.c l2_distance, `<->`): sqrt(sum (a_i-b_i)^2).
 */
static double distance_f64(const float *q, const float *e, int dim, int metric, double nq2) {
    if (metric == ORACLE_COSINE) {
        double dot = 0.0, ne2 = 0.0;
        for (int i = 0; i < dim; ++i) {
            double a = (double)q[i], b = (double)e[i];
            dot = dot + a * b;
            ne2 = ne2 + b * b;
        }
        double sim = dot / sqrt(nq2 * ne2);
        if (sim > 1.0) sim = 1.0;
        else if (sim < -1.0) sim = -1.0;
        return 1.0 - sim;
    } else {
        double s = 0.0;
        for (int i = 0; i < dim; ++i) {
            double d = (double)q[i] - (double)e[i];
            double d2 = d * d;
            s = s + d2;
        }
        return sqrt(s);
    }
}

/*
 * The same two operators with the arithmetic pgvector itself uses (third-party, absent from /root/reference, version unpinned;
 * restated from its published src/vector.c as SURVEY.md section 8(c) records it): `float` accumulators walked in index
 * order, the division / square root in double.
 *   cosine_distance: float dot, norma, normb;  similarity = (double)dot / sqrt((double)norma * (double)normb), clamped to
 *                    [-1, 1];  distance = 1.0 - similarity
 *   l2_distance:     float distance += diff * diff;  sqrt((double)distance)
 * pgvector is built with -ftree-vectorize -fassociative-math, so its own low-order bits depend on the build; this is ONE of
 * the summation orders it can take.  It is not the oracle (the oracle is the fp64 form above, which every such order
 * approximates): tests use it to check north_star's tolerance -- scores within 1e-5 relative of "an exact fp32 brute-force
 * reference of pgvector's semantics", ids equal except between rows whose fp32 scores are that close.
 */
void oracle_pgvector_f32_scores(const float *rows, int64_t n_rows, int dim, int64_t ld, const float *query,
                                int metric, const float *penalty, double *out) {
    for (int64_t r = 0; r < n_rows; ++r) {
        const float *e = rows + r * ld;
        double d;
        if (metric == ORACLE_COSINE) {
            float dot = 0.0f, na = 0.0f, nb = 0.0f;
            for (int i = 0; i < dim; ++i) {
                dot += query[i] * e[i];
                na += query[i] * query[i];
                nb += e[i] * e[i];
            }
            double sim = (double)dot / sqrt((double)na * (double)nb);
            if (sim > 1.0) sim = 1.0;
            else if (sim < -1.0) sim = -1.0;
            d = 1.0 - sim;
        } else {
            float acc = 0.0f;
            for (int i = 0; i < dim; ++i) {
                float diff = query[i] - e[i];
                acc += diff * diff;
            }
            d = sqrt((double)acc);
        }
        out[r] = d + (penalty ? (double)penalty[r] : 0.0);
    }
}

/* scores[n] = distance + (double)penalty (penalty may be NULL => 0). */
void oracle_scores(const float *rows, int64_t n_rows, int dim, int64_t ld, const float *query,
                   int metric, const float *penalty, double *out_scores) {
    double nq2 = norm2_f64(query, dim);
    for (int64_t r = 0; r < n_rows; ++r) {
        double d = distance_f64(query, rows + r * ld, dim, metric, nq2);
        out_scores[r] = penalty ? d + (double)penalty[r] : d;
    }
}

/* (score,id) ascending; NaN ranks after every number; ties -> lowest id. */
static int key_less(double sa, int64_t ia, double sb, int64_t ib) {
    int na = isnan(sa), nb = isnan(sb);
    if (na != nb) return nb;          /* a is a number, b is NaN -> a first */
    if (!na && sa != sb) return sa < sb;
    return ia < ib;
}

/* Exact top-k of one score vector by bounded max-heap; out sorted ascending.
 * Rows beyond n_rows (k > n_rows) are reported as id -1, score +inf. */
static void topk_one(const double *scores, int64_t n_rows, int k, int64_t id_base, int32_t *out_ids,
                     double *out_scores) {
    double *hs = (double *)malloc(sizeof(double) * (size_t)k);
    int64_t *hi = (int64_t *)malloc(sizeof(int64_t) * (size_t)k);
    int m = 0;
    for (int64_t r = 0; r < n_rows; ++r) {
        double s = scores[r];
        if (m < k) {
            int c = m++;
            hs[c] = s; hi[c] = r;
            while (c > 0) {                      /* sift up: max-heap on key */
                int p = (c - 1) / 2;
                if (key_less(hs[p], hi[p], hs[c], hi[c])) {
                    double ts = hs[p]; hs[p] = hs[c]; hs[c] = ts;
                    int64_t ti = hi[p]; hi[p] = hi[c]; hi[c] = ti;
                    c = p;
                } else break;
            }
        } else if (key_less(s, r, hs[0], hi[0])) {
            hs[0] = s; hi[0] = r;
            int c = 0;
            for (;;) {
                int l = 2 * c + 1, rr = l + 1, big = c;
                if (l < m && key_less(hs[big], hi[big], hs[l], hi[l])) big = l;
                if (rr < m && key_less(hs[big], hi[big], hs[rr], hi[rr])) big = rr;
                if (big == c) break;
                double ts = hs[big]; hs[big] = hs[c]; hs[c] = ts;
                int64_t ti = hi[big]; hi[big] = hi[c]; hi[c] = ti;
                c = big;
            }
        }
    }
    /* heap -> ascending: insertion sort (k is small) */
    for (int i = 1; i < m; ++i) {
        double s = hs[i]; int64_t id = hi[i]; int j = i - 1;
        while (j >= 0 && key_less(s, id, hs[j], hi[j])) { hs[j + 1] = hs[j]; hi[j + 1] = hi[j]; --j; }
        hs[j + 1] = s; hi[j + 1] = id;
    }
    for (int i = 0; i < k; ++i) {
        if (i < m) { out_ids[i] = (int32_t)(hi[i] + id_base); out_scores[i] = hs[i]; }
        else       { out_ids[i] = -1;                          out_scores[i] = INFINITY; }
    }
    free(hs); free(hi);
}

/*
 * Exact scan + top-k for a batch of queries (the ORDER BY <op> LIMIT k that the
 * planner's candidate step stands for; call site control_plane.py:58-66).
 * out_scores32 = out_scores64 rounded once to fp32 (may be NULL).
 */
void oracle_topk(const float *rows, int64_t n_rows, int dim, int64_t ld, const float *queries,
                 int n_queries, int k, int metric, const float *penalty, int64_t id_base,
                 int32_t *out_ids, double *out_scores64, float *out_scores32) {
    double *scores = (double *)malloc(sizeof(double) * (size_t)(n_rows > 0 ? n_rows : 1));
    for (int q = 0; q < n_queries; ++q) {
        oracle_scores(rows, n_rows, dim, ld, queries + (int64_t)q * dim, metric, penalty, scores);
        topk_one(scores, n_rows, k, id_base, out_ids + (int64_t)q * k, out_scores64 + (int64_t)q * k);
        if (out_scores32)
            for (int i = 0; i < k; ++i)
                out_scores32[(int64_t)q * k + i] = (float)out_scores64[(int64_t)q * k + i];
    }
    free(scores);
}

/* Merge S sorted per-shard lists [S][Q][k] into the global top-k with the same key. */
void oracle_merge(const double *scores, const int32_t *ids, int n_shards, int n_queries, int k,
                  int32_t *out_ids, double *out_scores) {
    for (int q = 0; q < n_queries; ++q) {
        int *pos = (int *)calloc((size_t)n_shards, sizeof(int));
        for (int i = 0; i < k; ++i) {
            int best = -1;
            for (int s = 0; s < n_shards; ++s) {
                if (pos[s] >= k) continue;
                int64_t o = ((int64_t)s * n_queries + q) * k + pos[s];
                if (ids[o] < 0) { pos[s] = k; continue; }
                if (best < 0) { best = s; continue; }
                int64_t ob = ((int64_t)best * n_queries + q) * k + pos[best];
                if (key_less(scores[o], ids[o], scores[ob], ids[ob])) best = s;
            }
            if (best < 0) { out_ids[(int64_t)q * k + i] = -1; out_scores[(int64_t)q * k + i] = INFINITY; }
            else {
                int64_t ob = ((int64_t)best * n_queries + q) * k + pos[best]++;
                out_ids[(int64_t)q * k + i] = ids[ob]; out_scores[(int64_t)q * k + i] = scores[ob];
            }
        }
        free(pos);
    }
}

/*
 * Kahn levels for a batch of DAGs, following networkx dag.py:218-240 step by step
 * on the graph control_plane.py:94-100 builds (nx.DiGraph: duplicate edges collapse,
 * adjacency keeps first-insertion order).
 *   n_nodes[b] <= 64; edges are (src,dst) byte pairs, DAG b owns pairs
 *   edge_off[b] .. edge_off[b+1].
 *   out_level[b*64+v]  generation index of node v, 0xFF if v >= n_nodes or v is
 *                      never emitted (in or behind a cycle);
 *   out_order[b*64+i]  i-th node nx.topological_sort yields (0xFF past the end);
 *   out_n_levels[b]    number of generations yielded before the generator ends;
 *   out_status[b]      0 ok, 1 NetworkXUnfeasible (nodes remain).
 */
void oracle_dag_levels(const uint8_t *n_nodes, const uint32_t *edge_off, const uint8_t *edges,
                       int64_t n_dags, uint8_t *out_level, uint8_t *out_order,
                       uint8_t *out_n_levels, uint8_t *out_status) {
    for (int64_t b = 0; b < n_dags; ++b) {
        int n = n_nodes[b];
        uint8_t adj[64][64]; int nadj[64]; int indeg[64]; uint64_t seen[64];
        for (int v = 0; v < n; ++v) { nadj[v] = 0; indeg[v] = 0; seen[v] = 0; }
        for (uint32_t e = edge_off[b]; e < edge_off[b + 1]; ++e) {
            int u = edges[2 * (size_t)e], v = edges[2 * (size_t)e + 1];
            if (u >= n || v >= n) continue;                 /* encoder never emits these */
            if (seen[u] >> v & 1) continue;                 /* DiGraph: same edge again */
            seen[u] |= 1ull << v;
            adj[u][nadj[u]++] = (uint8_t)v;
            indeg[v]++;
        }
        uint8_t *lvl = out_level + b * 64, *ord = out_order ? out_order + b * 64 : 0;
        memset(lvl, 0xFF, 64);
        if (ord) memset(ord, 0xFF, 64);
        uint8_t cur[64], nxt[64]; int ncur = 0, nnxt, done = 0, level = 0;
        for (int v = 0; v < n; ++v) if (indeg[v] == 0) cur[ncur++] = (uint8_t)v;
        while (ncur) {
            nnxt = 0;
            for (int i = 0; i < ncur; ++i) {
                int u = cur[i];
                lvl[u] = (uint8_t)level;
                if (ord) ord[done] = (uint8_t)u;
                done++;
                for (int j = 0; j < nadj[u]; ++j) {
                    int c = adj[u][j];
                    if (--indeg[c] == 0) nxt[nnxt++] = (uint8_t)c;
                }
            }
            memcpy(cur, nxt, (size_t)nnxt); ncur = nnxt; level++;
        }
        out_n_levels[b] = (uint8_t)level;
        out_status[b] = (uint8_t)(done != n);
    }
}
