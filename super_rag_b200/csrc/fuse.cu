// K3: RRF / weighted fusion + dedup of candidate lists, and K4: cross-shard candidate merge.
// One CTA per query; lists are tiny (<= 2048 entries), so everything lives in shared memory. Both kernels are
// SEGMENTED SORTS (north_star): a bitonic sort by (id, position) brings the occurrences of a chunk together in
// their original (list, position) order, the head of each segment sums its segment sequentially — the summation
// order of the reference — and a second bitonic sort by (score desc, id asc) orders the survivors.
// O(n log^2 n) per query; deterministic, no atomics. (Round 1 ranked by counting, O(n^2).)
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

// Bitonic network over P (power of two) slots; `less(i, j)` says whether slot i must come before slot j in the final
// order, `swap(i, j)` exchanges them. All threads of the CTA take part.
template <typename Before, typename Swap>
__device__ __forceinline__ void block_bitonic(int P, Before before, Swap swap) {
    for (int size = 2; size <= P; size <<= 1) {
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            for (int t = threadIdx.x; t < (P >> 1); t += blockDim.x) {
                const int i = 2 * t - (t & (stride - 1));
                const int j = i + stride;
                const bool up = (i & size) == 0;  // this pair sorts towards the final order
                if (before(j, i) == up) swap(i, j);
            }
            __syncthreads();
        }
    }
}

__device__ __forceinline__ int pow2_at_least(int n) {
    int p = 32;
    while (p < n) p <<= 1;
    return p;
}

__global__ void __launch_bounds__(FUSE_THREADS)
fuse_kernel(const int64_t* __restrict__ ids, const float* __restrict__ scores, int n_lists, int n_queries,
            int list_len, int mode, int rank_const, double min_score, FuseWeights wts, int k_out,
            int64_t* __restrict__ out_ids, double* __restrict__ out_scores) {
    __shared__ int64_t s_id[FUSE_MAX_N];
    __shared__ double s_val[FUSE_MAX_N];     // contribution of the entry, later the fused score of a segment head
    __shared__ uint16_t s_pos[FUSE_MAX_N];   // (list, position) order of the entry = its index in the input
    const int q = blockIdx.x;
    const int n = n_lists * list_len;
    const int P = pow2_at_least(n);
    constexpr int64_t ID_NONE = 0x7FFFFFFFFFFFFFFFll;  // padding sorts last
    for (int e = threadIdx.x; e < P; e += blockDim.x) {
        int64_t id = ID_NONE;
        double c = 0.0;
        if (e < n) {
            const int l = e / list_len, i = e - l * list_len;
            const size_t g = ((size_t)l * n_queries + q) * list_len + i;
            const int64_t v = ids[g];
            if (v >= 0) {
                id = v;
                c = mode == SRAG_FUSE_RRF ? 1.0 / (double)(i + rank_const) : (double)__fmul_rn((float)wts.w[l], scores[g]);
            }
        }
        s_id[e] = id;
        s_val[e] = c;
        s_pos[e] = (uint16_t)e;
    }
    for (int j = threadIdx.x; j < k_out; j += blockDim.x) {
        out_ids[(size_t)q * k_out + j] = -1;
        out_scores[(size_t)q * k_out + j] = -INFINITY;
    }
    __syncthreads();
    auto swap_all = [&](int i, int j) {
        const int64_t a = s_id[i]; s_id[i] = s_id[j]; s_id[j] = a;
        const double b = s_val[i]; s_val[i] = s_val[j]; s_val[j] = b;
        const uint16_t c = s_pos[i]; s_pos[i] = s_pos[j]; s_pos[j] = c;
    };
    // 1. segments: (id asc, position asc)
    block_bitonic(P, [&](int i, int j) { return s_id[i] < s_id[j] || (s_id[i] == s_id[j] && s_pos[i] < s_pos[j]); }, swap_all);
    // 2. the head of each segment adds its members in (list, position) order — exactly the reference's `+=` sequence;
    //    everything else (and heads below min_score) becomes padding
    double head_score[FUSE_MAX_N / FUSE_THREADS];
    bool head[FUSE_MAX_N / FUSE_THREADS];
#pragma unroll
    for (int r = 0; r < FUSE_MAX_N / FUSE_THREADS; ++r) {
        const int e = threadIdx.x + r * FUSE_THREADS;
        head[r] = false;
        head_score[r] = 0.0;
        if (e < P && s_id[e] != ID_NONE && (e == 0 || s_id[e - 1] != s_id[e])) {
            const int64_t id = s_id[e];
            if (mode == SRAG_FUSE_RRF) {
                double acc = 0.0;
                for (int o = e; o < P && s_id[o] == id; ++o) acc = __dadd_rn(acc, s_val[o]);
                head_score[r] = acc;
            } else {
                float acc = 0.0f;
                for (int o = e; o < P && s_id[o] == id; ++o) acc = __fadd_rn(acc, (float)s_val[o]);
                head_score[r] = (double)acc;
            }
            head[r] = head_score[r] >= min_score;
        }
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < FUSE_MAX_N / FUSE_THREADS; ++r) {
        const int e = threadIdx.x + r * FUSE_THREADS;
        if (e < P) {
            if (head[r]) {
                s_val[e] = head_score[r];
            } else {
                s_id[e] = ID_NONE;
                s_val[e] = -INFINITY;
            }
        }
    }
    __syncthreads();
    // 3. final order: (score desc, id asc); padding (id = ID_NONE) last whatever its score
    block_bitonic(P, [&](int i, int j) {
        const bool pi = s_id[i] == ID_NONE, pj = s_id[j] == ID_NONE;
        if (pi != pj) return pj;
        return s_val[i] > s_val[j] || (s_val[i] == s_val[j] && s_id[i] < s_id[j]);
    }, swap_all);
    for (int j = threadIdx.x; j < k_out && j < P; j += blockDim.x) {
        if (s_id[j] != ID_NONE) {
            out_ids[(size_t)q * k_out + j] = s_id[j];
            out_scores[(size_t)q * k_out + j] = s_val[j];
        }
    }
}

// K4 on (score, int64 id) pairs: any id width (the fallback exchange format).
__global__ void __launch_bounds__(FUSE_THREADS)
merge_candidates_kernel(const float* __restrict__ scores, const int64_t* __restrict__ ids, int n_shards,
                        int n_queries, int kk, int k, float* __restrict__ out_scores,
                        int64_t* __restrict__ out_ids) {
    __shared__ int64_t s_id[FUSE_MAX_N];
    __shared__ float s_sc[FUSE_MAX_N];
    const int q = blockIdx.x;
    const int n = n_shards * kk;
    const int P = pow2_at_least(n);
    for (int e = threadIdx.x; e < P; e += blockDim.x) {
        int64_t id = -1;
        float sc = -INFINITY;
        if (e < n) {
            const int g = e / kk, i = e - g * kk;
            const size_t src = ((size_t)g * n_queries + q) * kk + i;
            id = ids[src];
            sc = scores[src];
        }
        s_id[e] = id;
        s_sc[e] = sc;
    }
    __syncthreads();
    block_bitonic(P, [&](int i, int j) {
        const bool pi = s_id[i] < 0, pj = s_id[j] < 0;
        if (pi != pj) return pj;
        return s_sc[i] > s_sc[j] || (s_sc[i] == s_sc[j] && s_id[i] < s_id[j]);
    }, [&](int i, int j) {
        const int64_t a = s_id[i]; s_id[i] = s_id[j]; s_id[j] = a;
        const float b = s_sc[i]; s_sc[i] = s_sc[j]; s_sc[j] = b;
    });
    for (int j = threadIdx.x; j < k; j += blockDim.x) {
        const bool ok = j < P && s_id[j] >= 0;
        out_ids[(size_t)q * k + j] = ok ? s_id[j] : -1;
        out_scores[(size_t)q * k + j] = ok ? s_sc[j] : -INFINITY;
    }
}

// K4 on exchange keys (the format the finalize kernel writes into the all-gather send buffer): one 64-bit sortable
// key per candidate = orderable(final score) << 32 | (0xFFFFFFFF - global row); key descending <=> (score desc, id
// asc); 0 = empty. g * kk keys per query, one CTA per query, bitonic sort, decode.
__global__ void __launch_bounds__(FUSE_THREADS)
merge_keys_kernel(const uint64_t* __restrict__ keys, int n_shards, int n_queries, int kk, int k,
                  float* __restrict__ out_scores, int64_t* __restrict__ out_ids) {
    __shared__ uint64_t s_key[FUSE_MAX_N];
    const int q = blockIdx.x;
    const int n = n_shards * kk;
    const int P = pow2_at_least(n);
    for (int e = threadIdx.x; e < P; e += blockDim.x) {
        uint64_t key = SRAG_KEY_EMPTY;
        if (e < n) {
            const int g = e / kk, i = e - g * kk;
            key = keys[((size_t)g * n_queries + q) * kk + i];
        }
        s_key[e] = key;
    }
    __syncthreads();
    block_bitonic(P, [&](int i, int j) { return s_key[i] > s_key[j]; },
                  [&](int i, int j) { const uint64_t a = s_key[i]; s_key[i] = s_key[j]; s_key[j] = a; });
    for (int j = threadIdx.x; j < k; j += blockDim.x) {
        const uint64_t key = j < P ? s_key[j] : SRAG_KEY_EMPTY;
        const bool ok = key != SRAG_KEY_EMPTY;
        out_ids[(size_t)q * k + j] = ok ? (int64_t)srag_key_idx(key) : -1;
        out_scores[(size_t)q * k + j] = ok ? __uint_as_float(srag_key_score_bits(key)) : -INFINITY;
    }
}

__global__ void pack_keys_kernel(const float* __restrict__ scores, const int64_t* __restrict__ ids, int64_t n,
                                 uint64_t* __restrict__ keys, int* __restrict__ err) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t id = ids[i];
        uint64_t key = SRAG_KEY_EMPTY;
        if (id >= 0) {
            if (id >= (int64_t)0xFFFFFFFFll) {
                if (err) *err = 1;
            } else {
                key = srag_pack_key(__float_as_uint(scores[i]), (uint32_t)id);
            }
        }
        keys[i] = key;
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

extern "C" int srag_merge_keys(const uint64_t* keys, int n_shards, int n_queries, int kk, int k, float* out_scores,
                               int64_t* out_ids, void* stream) {
    SRAG_CHECK_ARG(n_queries >= 0 && k >= 1, "bad sizes n_queries=%d k=%d", n_queries, k);
    SRAG_CHECK_ARG(n_shards >= 1 && kk >= 1 && n_shards * kk <= FUSE_MAX_N, "bad candidate shape n_shards=%d kk=%d",
                   n_shards, kk);
    if (n_queries == 0) return SRAG_OK;
    SRAG_CHECK_ARG(keys && out_scores && out_ids, "null pointer argument");
    merge_keys_kernel<<<n_queries, FUSE_THREADS, 0, reinterpret_cast<cudaStream_t>(stream)>>>(keys, n_shards, n_queries, kk,
                                                                                            k, out_scores, out_ids);
    SRAG_CUDA(cudaGetLastError());
    return SRAG_OK;
}

extern "C" int srag_pack_keys(const float* scores, const int64_t* ids, int64_t n, uint64_t* keys, int* err_flag,
                              void* stream) {
    SRAG_CHECK_ARG(n >= 0, "bad size n=%lld", (long long)n);
    if (n == 0) return SRAG_OK;
    SRAG_CHECK_ARG(scores && ids && keys, "null pointer argument");
    int64_t blocks = (n + 255) / 256;
    if (blocks > 148 * 8) blocks = 148 * 8;
    pack_keys_kernel<<<(int)blocks, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(scores, ids, n, keys, err_flag);
    SRAG_CUDA(cudaGetLastError());
    return SRAG_OK;
}
