// K3: RRF / weighted fusion + dedup of candidate lists, and K4: cross-shard candidate merge.
// One CTA per query; lists are tiny (<= 2048 entries) so everything lives in shared memory and
// ordering is rank-by-counting under (score desc, id asc) — deterministic, no atomics.
//
// RRF follows graphiti_core/search/search_utils.py:1763-1778 of the reference exactly:
//   float64, score[id] += 1/(i + rank_const) with i 0-based, lists visited in order, duplicates
//   inside a list add again, keep score >= min_score. The `merge` node strategies "rrf" /
//   "weighted" (super_rag/nodeflow/runners/merge.py:40-65 only knows "union") call this.
#include "common.cuh"

namespace {

constexpr int FUSE_THREADS = 256;
constexpr int FUSE_MAX_N = 2048;

struct FuseWeights {
    double w[SRAG_MAX_LISTS];
};

__global__ void __launch_bounds__(FUSE_THREADS)
fuse_kernel(const int64_t* __restrict__ ids, const float* __restrict__ scores, int n_lists, int n_queries,
            int list_len, int mode, int rank_const, double min_score, FuseWeights wts, int k_out,
            int64_t* __restrict__ out_ids, double* __restrict__ out_scores) {
    __shared__ int64_t s_id[FUSE_MAX_N];
    __shared__ double s_val[FUSE_MAX_N];    // contribution, later the fused score
    __shared__ uint8_t s_keep[FUSE_MAX_N];  // canonical (first occurrence) and >= min_score
    const int q = blockIdx.x;
    const int n = n_lists * list_len;
    for (int e = threadIdx.x; e < n; e += blockDim.x) {
        const int l = e / list_len, i = e - l * list_len;
        const size_t g = ((size_t)l * n_queries + q) * list_len + i;
        const int64_t id = ids[g];
        s_id[e] = id;
        double c = 0.0;
        if (id >= 0) {
            if (mode == SRAG_FUSE_RRF) {
                c = 1.0 / (double)(i + rank_const);
            } else {
                c = (double)__fmul_rn((float)wts.w[l], scores[g]);
            }
        }
        s_val[e] = c;
    }
    for (int j = threadIdx.x; j < k_out; j += blockDim.x) {
        out_ids[(size_t)q * k_out + j] = -1;
        out_scores[(size_t)q * k_out + j] = -INFINITY;
    }
    __syncthreads();
    // fused score of each first occurrence, accumulated in (list, position) order
    double my_score[FUSE_MAX_N / FUSE_THREADS];
    bool my_keep[FUSE_MAX_N / FUSE_THREADS];
#pragma unroll
    for (int r = 0; r < FUSE_MAX_N / FUSE_THREADS; ++r) {
        const int e = threadIdx.x + r * FUSE_THREADS;
        my_keep[r] = false;
        my_score[r] = 0.0;
        if (e < n && s_id[e] >= 0) {
            const int64_t id = s_id[e];
            bool first = true;
            for (int o = 0; o < e; ++o)
                if (s_id[o] == id) { first = false; break; }
            if (first) {
                if (mode == SRAG_FUSE_RRF) {
                    double acc = 0.0;
                    for (int o = e; o < n; ++o)
                        if (s_id[o] == id) acc = __dadd_rn(acc, s_val[o]);
                    my_score[r] = acc;
                } else {
                    float acc = 0.0f;
                    for (int o = e; o < n; ++o)
                        if (s_id[o] == id) acc = __fadd_rn(acc, (float)s_val[o]);
                    my_score[r] = (double)acc;
                }
                my_keep[r] = my_score[r] >= min_score;
            }
        }
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < FUSE_MAX_N / FUSE_THREADS; ++r) {
        const int e = threadIdx.x + r * FUSE_THREADS;
        if (e < n) {
            s_keep[e] = my_keep[r] ? 1 : 0;
            s_val[e] = my_score[r];
        }
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < FUSE_MAX_N / FUSE_THREADS; ++r) {
        const int e = threadIdx.x + r * FUSE_THREADS;
        if (e < n && my_keep[r]) {
            const double s = my_score[r];
            const int64_t id = s_id[e];
            int rank = 0;
            for (int o = 0; o < n; ++o) {
                if (!s_keep[o]) continue;
                const double so = s_val[o];
                rank += (so > s || (so == s && s_id[o] < id)) ? 1 : 0;
            }
            if (rank < k_out) {
                out_ids[(size_t)q * k_out + rank] = id;
                out_scores[(size_t)q * k_out + rank] = s;
            }
        }
    }
}

__global__ void __launch_bounds__(FUSE_THREADS)
merge_candidates_kernel(const float* __restrict__ scores, const int64_t* __restrict__ ids, int n_shards,
                        int n_queries, int kk, int k, float* __restrict__ out_scores,
                        int64_t* __restrict__ out_ids) {
    __shared__ int64_t s_id[FUSE_MAX_N];
    __shared__ float s_sc[FUSE_MAX_N];
    const int q = blockIdx.x;
    const int n = n_shards * kk;
    for (int e = threadIdx.x; e < n; e += blockDim.x) {
        const int g = e / kk, i = e - g * kk;
        const size_t src = ((size_t)g * n_queries + q) * kk + i;
        s_id[e] = ids[src];
        s_sc[e] = scores[src];
    }
    for (int j = threadIdx.x; j < k; j += blockDim.x) {
        out_ids[(size_t)q * k + j] = -1;
        out_scores[(size_t)q * k + j] = -INFINITY;
    }
    __syncthreads();
    for (int e = threadIdx.x; e < n; e += blockDim.x) {
        const int64_t id = s_id[e];
        if (id < 0) continue;
        const float s = s_sc[e];
        int rank = 0;
        for (int o = 0; o < n; ++o) {
            const int64_t io = s_id[o];
            if (io < 0) continue;
            const float so = s_sc[o];
            rank += (so > s || (so == s && io < id)) ? 1 : 0;
        }
        if (rank < k) {
            out_ids[(size_t)q * k + rank] = id;
            out_scores[(size_t)q * k + rank] = s;
        }
    }
}

}  // namespace

extern "C" int srag_fuse(const int64_t* ids, const float* scores, int n_lists, int n_queries, int list_len,
                         int mode, int rank_const, double min_score, const double* weights_host, int k_out,
                         int64_t* out_ids, double* out_scores, void* stream) {
    SRAG_CHECK_ARG(n_queries >= 0 && k_out >= 1, "bad sizes n_queries=%d k_out=%d", n_queries, k_out);
    SRAG_CHECK_ARG(n_lists >= 1 && n_lists <= SRAG_MAX_LISTS && list_len >= 1 && n_lists * list_len <= FUSE_MAX_N,
                   "bad list shape n_lists=%d list_len=%d", n_lists, list_len);
    SRAG_CHECK_ARG(mode == SRAG_FUSE_RRF || mode == SRAG_FUSE_WEIGHTED, "bad mode %d", mode);
    SRAG_CHECK_ARG(mode != SRAG_FUSE_WEIGHTED || (scores && weights_host), "weighted fusion needs scores and weights");
    SRAG_CHECK_ARG(mode != SRAG_FUSE_RRF || rank_const >= 1, "rank_const must be >= 1");
    if (n_queries == 0) return SRAG_OK;
    SRAG_CHECK_ARG(ids && out_ids && out_scores, "null pointer argument");
    FuseWeights w;
    for (int l = 0; l < SRAG_MAX_LISTS; ++l) w.w[l] = (weights_host && l < n_lists) ? weights_host[l] : 1.0;
    fuse_kernel<<<n_queries, FUSE_THREADS, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
        ids, scores, n_lists, n_queries, list_len, mode, rank_const, min_score, w, k_out, out_ids, out_scores);
    SRAG_CUDA(cudaGetLastError());
    return SRAG_OK;
}

extern "C" int srag_merge_candidates(const float* scores, const int64_t* ids, int n_shards, int n_queries, int kk,
                                     int k, float* out_scores, int64_t* out_ids, void* stream) {
    SRAG_CHECK_ARG(n_queries >= 0 && k >= 1, "bad sizes n_queries=%d k=%d", n_queries, k);
    SRAG_CHECK_ARG(n_shards >= 1 && kk >= 1 && n_shards * kk <= FUSE_MAX_N, "bad candidate shape n_shards=%d kk=%d",
                   n_shards, kk);
    if (n_queries == 0) return SRAG_OK;
    SRAG_CHECK_ARG(scores && ids && out_scores && out_ids, "null pointer argument");
    merge_candidates_kernel<<<n_queries, FUSE_THREADS, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
        scores, ids, n_shards, n_queries, kk, k, out_scores, out_ids);
    SRAG_CUDA(cudaGetLastError());
    return SRAG_OK;
}
