/*
 * Plain-C restatement of the hot path's arithmetic. TEST INFRASTRUCTURE ONLY (see
 * oracle/__init__.py): tests/, __graft_entry__.smoke() and bench.py's CPU legs may load it,
 * nothing under super_rag_b200/ does.
 *
 * Reference anchors (paths relative to /root/reference):
 *   cosine:  graphiti_core/search/search_utils.py:56-67  dot/(|a||b|), 0 if a norm is 0
 *            (the production arithmetic sits in the external SeekDB server reached at
 *            super_rag/vectorstore/seekdb_connector.py:103-107 — PARITY UNPINNED)
 *   rrf:     graphiti_core/search/search_utils.py:1763-1778 (PINNED via tests/golden)
 *   bm25:    no in-tree arithmetic — published Okapi/Lucene BM25, contract in
 *            oracle/bm25.py (PARITY UNPINNED)
 * Build: make -C oracle   ->  oracle/_build/liboracle.so
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

/* scores[b*n + r] = cosine (metric 0) or dot (metric 1), fp32 inputs, sequential fp32 dot,
 * norms accumulated in double. */
void oracle_scores_f32(const float* corpus, int64_t n, const float* queries, int b, int dim, int metric,
                       float* scores) {
    double* cn = (double*)malloc(sizeof(double) * (size_t)(n > 0 ? n : 1));
    for (int64_t r = 0; r < n; ++r) {
        double s = 0;
        for (int d = 0; d < dim; ++d) s += (double)corpus[r * dim + d] * (double)corpus[r * dim + d];
        cn[r] = sqrt(s);
    }
    for (int q = 0; q < b; ++q) {
        double qn = 0;
        for (int d = 0; d < dim; ++d) qn += (double)queries[(int64_t)q * dim + d] * (double)queries[(int64_t)q * dim + d];
        qn = sqrt(qn);
        for (int64_t r = 0; r < n; ++r) {
            float dot = 0.f;
            for (int d = 0; d < dim; ++d) dot += queries[(int64_t)q * dim + d] * corpus[r * dim + d];
            float out = dot;
            if (metric == 0) {
                if (qn == 0 || cn[r] == 0) out = 0.f;
                else out = (float)((double)dot / (qn * cn[r]));
            }
            scores[(int64_t)q * n + r] = out;
        }
    }
    free(cn);
}

/* top-k of one score row under (score desc, id asc); ids are 0..n-1; valid may be NULL.
 * O(n*k) insertion — small cases only. Returns the number of results. */
int oracle_topk_row(const float* scores, const uint8_t* valid, int64_t n, int k, float* out_s, int64_t* out_i) {
    int m = 0;
    for (int64_t r = 0; r < n; ++r) {
        if (valid && !valid[r]) continue;
        float s = scores[r];
        if (s != s) continue;
        int pos = m;
        while (pos > 0 && (out_s[pos - 1] < s)) --pos; /* equal scores keep the earlier (smaller) id first */
        if (pos >= k) continue;
        int last = m < k ? m : k - 1;
        for (int j = last; j > pos; --j) { out_s[j] = out_s[j - 1]; out_i[j] = out_i[j - 1]; }
        out_s[pos] = s; out_i[pos] = r;
        if (m < k) ++m;
    }
    return m;
}

/* BM25 impacts, one IEEE fp32 operation per step (compile without -ffast-math / FMA contraction). */
void oracle_bm25_impacts(const int64_t* term_ptr, int vocab, const int32_t* post_doc, const int32_t* post_tf,
                         const float* idf, const int32_t* dl, float avgdl, float* imp) {
    const volatile float k1 = 1.2f, b = 0.75f;
    const volatile float one_minus_b = 1.0f - b;
    const volatile float k1p1 = k1 + 1.0f;
    for (int t = 0; t < vocab; ++t) {
        for (int64_t p = term_ptr[t]; p < term_ptr[t + 1]; ++p) {
            volatile float tf = (float)post_tf[p];
            volatile float r = (float)dl[post_doc[p]] / avgdl;
            volatile float tt = b * r;
            volatile float u = one_minus_b + tt;
            volatile float nrm = k1 * u;
            volatile float a = tf * k1p1;
            volatile float num = idf[t] * a;
            volatile float den = tf + nrm;
            imp[p] = num / den;
        }
    }
}

/* acc[d] += imp for every posting of the given (unique, ascending) terms; hit[d] = 1 */
void oracle_bm25_accumulate(const int64_t* term_ptr, const int32_t* post_doc, const float* post_imp,
                            const int32_t* terms, int n_terms, float* acc, uint8_t* hit) {
    for (int j = 0; j < n_terms; ++j) {
        int t = terms[j];
        if (t < 0) continue;
        for (int64_t p = term_ptr[t]; p < term_ptr[t + 1]; ++p) {
            volatile float s = acc[post_doc[p]] + post_imp[p];
            acc[post_doc[p]] = s;
            hit[post_doc[p]] = 1;
        }
    }
}

/* rrf over n_lists padded id lists (id < 0 = padding). out arrays sized n_lists*list_len.
 * Scores as in search_utils.py:1763-1778 (double, list order); order (score desc, id asc).
 * Returns the number of fused ids with score >= min_score. */
int oracle_rrf(const int64_t* ids, int n_lists, int list_len, int rank_const, double min_score,
               int64_t* out_ids, double* out_scores) {
    int n = n_lists * list_len, m = 0;
    for (int e = 0; e < n; ++e) {
        int64_t id = ids[e];
        if (id < 0) continue;
        int seen = 0;
        for (int o = 0; o < e; ++o) if (ids[o] == id) { seen = 1; break; }
        if (seen) continue;
        double acc = 0.0;
        for (int o = e; o < n; ++o) if (ids[o] == id) acc += 1.0 / (double)((o % list_len) + rank_const);
        if (!(acc >= min_score)) continue;
        int pos = m;
        while (pos > 0 && (out_scores[pos - 1] < acc || (out_scores[pos - 1] == acc && out_ids[pos - 1] > id))) --pos;
        for (int j = m; j > pos; --j) { out_scores[j] = out_scores[j - 1]; out_ids[j] = out_ids[j - 1]; }
        out_scores[pos] = acc; out_ids[pos] = id;
        ++m;
    }
    return m;
}
