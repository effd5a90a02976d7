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

// ---------------------------------------------------------------------------------------
// K4x: candidate exchange + merge in ONE kernel over peer memory (NVLink / NVSwitch P2P stores), no collective call.
// Every rank owns a symmetric buffer (same layout on every rank, mapped into every peer):
//     keys  uint64 [2 parities][world][cap_keys]      flags uint32 [world][cap_q]
// Step `epoch` (1, 2, ... the same on every rank): phase 1 — each CTA stores its queries' k exchange keys into slot
// [epoch & 1][my rank] of EVERY rank's buffer (its own included), fences system-wide, then sets flag [my rank][q] = epoch
// on every rank; phase 2 — for each of its queries the CTA waits until all `world` flags of the query have reached
// `epoch`, reads the world * k keys (bypassing L1) and merges them like srag_merge_keys. Pushes never wait for anything,
// so no set of resident CTAs can deadlock another rank. Keys are double-buffered by parity: a rank cannot start step
// e + 2 (which reuses the parity of e) before every rank has pushed e + 1, i.e. finished merging e.
// ---------------------------------------------------------------------------------------
namespace {

constexpr int P2P_MAX_WORLD = 8;
struct P2pPeers {
    uint64_t* keys[P2P_MAX_WORLD];
    uint32_t* flags[P2P_MAX_WORLD];
};

__global__ void __launch_bounds__(FUSE_THREADS)
exchange_merge_p2p_kernel(const uint64_t* __restrict__ local_keys, P2pPeers peers, int world, int rank, int n_queries, int k,
                          uint32_t epoch, size_t cap_keys, int cap_q, float* __restrict__ out_scores,
                          int64_t* __restrict__ out_ids) {
    __shared__ uint64_t s_key[FUSE_MAX_N];
    const size_t par_off = (size_t)(epoch & 1u) * world * cap_keys;
    // ---- phase 1: push this CTA's queries to every rank ----
    for (int q = blockIdx.x; q < n_queries; q += gridDim.x) {
        for (int e = threadIdx.x; e < world * k; e += blockDim.x) {
            const int dst = e / k, j = e - dst * k;
            peers.keys[dst][par_off + (size_t)rank * cap_keys + (size_t)q * k + j] = local_keys[(size_t)q * k + j];
        }
        __threadfence_system();   // every thread: its stores are visible system-wide before the flags below
        __syncthreads();
        if (threadIdx.x < world)
            *reinterpret_cast<volatile uint32_t*>(peers.flags[threadIdx.x] + (size_t)rank * cap_q + q) = epoch;
    }
    // ---- phase 2: wait for every rank's keys of a query, merge ----
    const volatile uint32_t* my_flags = peers.flags[rank];
    const uint64_t* my_keys = peers.keys[rank] + par_off;
    const int n = world * k;
    const int P = pow2_at_least(n);
    for (int q = blockIdx.x; q < n_queries; q += gridDim.x) {
        if (threadIdx.x < world) {
            uint32_t spins = 0;
            // flags only ever rise (a peer one step ahead writes epoch + 1); bounded: a lost peer traps instead of hanging
            while ((int32_t)(my_flags[(size_t)threadIdx.x * cap_q + q] - epoch) < 0) {
                __nanosleep(64);
                if (++spins > (1u << 24)) __trap();
            }
            __threadfence_system();
        }
        __syncthreads();
        for (int e = threadIdx.x; e < P; e += blockDim.x) {
            uint64_t key = SRAG_KEY_EMPTY;
            if (e < n) {
                const int g = e / k, i = e - g * k;
                key = __ldcv(my_keys + (size_t)g * cap_keys + (size_t)q * k + i);   // written by a peer: not through L1
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
        __syncthreads();
    }
}

}  // namespace

extern "C" int srag_exchange_merge_p2p(const uint64_t* local_keys, const uint64_t* peer_keys_host,
                                       const uint64_t* peer_flags_host, int world, int rank, int n_queries, int k,
                                       uint32_t epoch, size_t cap_keys, int cap_q, float* out_scores, int64_t* out_ids,
                                       void* stream) {
    SRAG_CHECK_ARG(world >= 1 && world <= P2P_MAX_WORLD && rank >= 0 && rank < world, "bad world=%d rank=%d", world, rank);
    SRAG_CHECK_ARG(n_queries >= 0 && k >= 1 && world * k <= FUSE_MAX_N, "bad sizes n_queries=%d k=%d", n_queries, k);
    SRAG_CHECK_ARG(epoch >= 1 && n_queries <= cap_q && (size_t)n_queries * k <= cap_keys,
                   "epoch must be >= 1 and the batch must fit the symmetric buffer (cap_q=%d cap_keys=%zu)", cap_q, cap_keys);
    if (n_queries == 0) return SRAG_OK;
    SRAG_CHECK_ARG(local_keys && peer_keys_host && peer_flags_host && out_scores && out_ids, "null pointer argument");
    P2pPeers peers;
    for (int r = 0; r < P2P_MAX_WORLD; ++r) {
        peers.keys[r] = r < world ? reinterpret_cast<uint64_t*>(peer_keys_host[r]) : nullptr;
        peers.flags[r] = r < world ? reinterpret_cast<uint32_t*>(peer_flags_host[r]) : nullptr;
        SRAG_CHECK_ARG(r >= world || (peers.keys[r] && peers.flags[r]), "null peer pointer for rank %d", r);
    }
    int sms = 0;
    const int rc = srag_current_device_props(&sms, nullptr, nullptr);
    if (rc) return rc;
    const int resident = sms * 4;   // every CTA pushes before it waits, so the grid size is a throughput choice only
    const int grid = n_queries < resident ? n_queries : resident;
    exchange_merge_p2p_kernel<<<grid, FUSE_THREADS, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
        local_keys, peers, world, rank, n_queries, k, epoch, cap_keys, cap_q, out_scores, out_ids);
    SRAG_CUDA(cudaGetLastError());
    return SRAG_OK;
}
