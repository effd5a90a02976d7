// K2 index build from tokens: tokens + document boundaries -> CSR-by-term postings
// (term_ptr, post_doc, post_tf, dl) and the sub-block offset table block_ptr, without a global sort.
// HBM-bound integer work; every step is one pass over its input.
//
//   build_doc_terms_kernel   one CTA per document (a second launch with 224 KB of shared memory takes the documents
//                            of more than 8192 tokens): its tokens are sorted in shared memory (bitonic) and run-length
//                            encoded IN PLACE into (term, tf) pairs at the start of the document's token range; every
//                            pair is counted into cnt[term][sub-block of the document].
//   build_row_sum / build_scan_terms / build_row_scan
//                            cnt -> term_ptr (exclusive scan of the row sums) and, in place, block_ptr[term][j] =
//                            term_ptr[term] + postings of the term in sub-blocks < j. That table is exactly what the
//                            scoring kernels index (bm25.cu), so no bisection pass is needed afterwards.
//   build_scatter_kernel     one warp per document: posting (doc, tf) goes to block_ptr[t][sub] + (a running count per
//                            slice). Slices fill in arbitrary order ...
//   build_sort_slices_kernel ... so every (term, sub-block) slice is then sorted by document (one warp per slice), which
//                            makes the result deterministic: postings ascending by document inside a term, as
//                            oracle/bm25.py builds them.
// The reference has no index build; contract: oracle/bm25.py build_postings.
#include "common.cuh"

namespace {

constexpr int BUILD_SUB_DOCS = 2048;   // = BM25_SUB_DOCS (bm25.cu); srag_bm25_block_docs() is checked by the caller
constexpr int BUILD_MAX_DL = 8192;     // tokens of one document in the main pass (32 KB of shared memory, 7 CTAs per SM)
constexpr int BUILD_LONG_DL = 57344;   // ... in the pass for long documents (224 KB, one CTA per SM)
constexpr int BUILD_THREADS = 128;
constexpr uint32_t PAD = 0xFFFFFFFFu;

__global__ void __launch_bounds__(BUILD_THREADS)
build_doc_terms_kernel(int32_t* __restrict__ tokens, const int64_t* __restrict__ doc_ptr, int64_t n_docs, int vocab,
                       int n_sub, int32_t* __restrict__ tf_out, int32_t* __restrict__ doc_uniq, int32_t* __restrict__ dl_out,
                       unsigned long long* __restrict__ cnt, int* __restrict__ err,
                       int min_len /* documents of min_len < tokens <= max_len are this launch's */, int max_len) {
    extern __shared__ uint32_t buf[];
    __shared__ int warp_heads[BUILD_THREADS / 32];
    __shared__ int running;
    __shared__ uint32_t sel[BUILD_THREADS / 32];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;

    // the whole CTA on one document: sort, run-length encode in place, count the pairs
    auto process_doc = [&](int64_t d) {
        const int64_t a = doc_ptr[d];
        const int len = (int)(doc_ptr[d + 1] - a);
        if (threadIdx.x == 0) dl_out[d] = len;
        int P = 2;
        while (P < len) P <<= 1;
        for (int i = threadIdx.x; i < P; i += BUILD_THREADS) buf[i] = i < len ? (uint32_t)tokens[a + i] : PAD;
        if (threadIdx.x == 0) running = 0;
        __syncthreads();
        for (int size = 2; size <= P; size <<= 1) {
            for (int stride = size >> 1; stride > 0; stride >>= 1) {
                for (int t = threadIdx.x; t < (P >> 1); t += BUILD_THREADS) {
                    const int i = 2 * t - (t & (stride - 1));
                    const int j = i + stride;
                    const bool asc = (i & size) == 0;
                    const uint32_t x = buf[i], y = buf[j];
                    if ((x > y) == asc) { buf[i] = y; buf[j] = x; }
                }
                __syncthreads();
            }
        }
        // a head is the first token of a run of equal terms
        const int sub = (int)(d / BUILD_SUB_DOCS);
        for (int base = 0; base < len; base += BUILD_THREADS) {
            const int i = base + threadIdx.x;
            const bool head = i < len && (i == 0 || buf[i] != buf[i - 1]);
            const uint32_t bal = __ballot_sync(0xFFFFFFFFu, head);
            if (lane == 0) warp_heads[w] = __popc(bal);
            __syncthreads();
            int before = running + __popc(bal & ((1u << lane) - 1u));
            for (int v = 0; v < w; ++v) before += warp_heads[v];
            if (head) {
                const uint32_t t = buf[i];
                int j = i + 1;
                while (j < len && buf[j] == t) ++j;
                if (t >= (uint32_t)vocab) {
                    atomicExch(err, 2);
                } else {
                    tokens[a + before] = (int32_t)t;   // all of the document's tokens are in shared memory by now
                    tf_out[a + before] = j - i;
                    atomicAdd(&cnt[(size_t)t * (n_sub + 1) + sub], 1ull);
                }
            }
            __syncthreads();
            if (threadIdx.x == 0) {
                int tot = 0;
                for (int v = 0; v < BUILD_THREADS / 32; ++v) tot += warp_heads[v];
                running += tot;
            }
            __syncthreads();
        }
        if (threadIdx.x == 0) doc_uniq[d] = running;
        __syncthreads();
    };

    // A CTA takes BUILD_THREADS consecutive documents at a time: every thread looks at one document's length, the ones
    // that belong to this launch are then processed one after the other.
    for (int64_t d0 = (int64_t)blockIdx.x * BUILD_THREADS; d0 < n_docs; d0 += (int64_t)gridDim.x * BUILD_THREADS) {
        const int64_t dd = d0 + threadIdx.x;
        bool mine = false;
        if (dd < n_docs) {
            const int64_t l = doc_ptr[dd + 1] - doc_ptr[dd];
            if (l > BUILD_LONG_DL || l < 0) {
                if (min_len == 0) { atomicExch(err, 1); doc_uniq[dd] = 0; dl_out[dd] = 0; }
            } else {
                mine = (l > min_len || (min_len == 0 && l == 0)) && l <= max_len;
            }
        }
        const uint32_t bal = __ballot_sync(0xFFFFFFFFu, mine);
        if (lane == 0) sel[w] = bal;
        __syncthreads();
        for (int sw = 0; sw < BUILD_THREADS / 32; ++sw) {
            uint32_t todo = sel[sw];
            while (todo) {
                process_doc(d0 + 32 * sw + (__ffs(todo) - 1));
                todo &= todo - 1;
            }
        }
        __syncthreads();  // sel is rewritten by the next round
    }
}

// df[t] = sum_j cnt[t][j], one warp per term
__global__ void build_row_sum_kernel(const unsigned long long* __restrict__ cnt, int vocab, int n_sub,
                                     int64_t* __restrict__ df) {
    const int lane = threadIdx.x & 31;
    const int64_t warps = (int64_t)gridDim.x * (blockDim.x >> 5);
    for (int64_t t = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); t < vocab; t += warps) {
        const unsigned long long* row = cnt + (size_t)t * (n_sub + 1);
        unsigned long long s = 0;
        for (int j = lane; j < n_sub; j += 32) s += row[j];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xFFFFFFFFu, s, o);
        if (lane == 0) df[t] = (int64_t)s;
    }
}

// term_ptr = exclusive scan of df (one CTA; every thread owns a contiguous segment)
__global__ void __launch_bounds__(1024) build_scan_terms_kernel(const int64_t* __restrict__ df, int vocab,
                                                                int64_t* __restrict__ term_ptr) {
    __shared__ int64_t part[1024];
    const int per = (vocab + 1023) / 1024;
    const int lo = threadIdx.x * per, hi = min(vocab, lo + per);
    int64_t s = 0;
    for (int i = lo; i < hi; ++i) s += df[i];
    part[threadIdx.x] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        int64_t run = 0;
        for (int i = 0; i < 1024; ++i) {
            const int64_t v = part[i];
            part[i] = run;
            run += v;
        }
        term_ptr[vocab] = run;
    }
    __syncthreads();
    int64_t run = part[threadIdx.x];
    for (int i = lo; i < hi; ++i) {
        term_ptr[i] = run;
        run += df[i];
    }
}

// cnt[t][j] -> block_ptr[t][j] = term_ptr[t] + sum_{j' < j} cnt[t][j'] in place; [t][n_sub] = term_ptr[t + 1]
__global__ void build_row_scan_kernel(unsigned long long* __restrict__ cnt, int vocab, int n_sub,
                                      const int64_t* __restrict__ term_ptr) {
    const int lane = threadIdx.x & 31;
    const int64_t warps = (int64_t)gridDim.x * (blockDim.x >> 5);
    for (int64_t t = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); t < vocab; t += warps) {
        unsigned long long* row = cnt + (size_t)t * (n_sub + 1);
        unsigned long long carry = (unsigned long long)term_ptr[t];
        for (int base = 0; base <= n_sub; base += 32) {
            const int j = base + lane;
            const unsigned long long v = j < n_sub ? row[j] : 0ull;
            unsigned long long inc = v;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const unsigned long long u = __shfl_up_sync(0xFFFFFFFFu, inc, o);
                if (lane >= o) inc += u;
            }
            if (j <= n_sub) row[j] = carry + inc - v;
            carry += __shfl_sync(0xFFFFFFFFu, inc, 31);
        }
    }
}

// one warp per document: its (term, tf) pairs go to their slices
__global__ void build_scatter_kernel(const int32_t* __restrict__ uniq_term, const int32_t* __restrict__ uniq_tf,
                                     const int64_t* __restrict__ doc_ptr, const int32_t* __restrict__ doc_uniq,
                                     int64_t n_docs, int n_sub, const int64_t* __restrict__ block_ptr,
                                     uint32_t* __restrict__ fill, int32_t* __restrict__ post_doc,
                                     int32_t* __restrict__ post_tf) {
    const int lane = threadIdx.x & 31;
    const int64_t warps = (int64_t)gridDim.x * (blockDim.x >> 5);
    for (int64_t d = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); d < n_docs; d += warps) {
        const int64_t a = doc_ptr[d];
        const int n = doc_uniq[d];
        const int sub = (int)(d / BUILD_SUB_DOCS);
        for (int i = lane; i < n; i += 32) {
            const int32_t t = uniq_term[a + i];
            const int64_t pos = block_ptr[(size_t)t * (n_sub + 1) + sub] + atomicAdd(&fill[(size_t)t * n_sub + sub], 1u);
            post_doc[pos] = (int32_t)d;
            post_tf[pos] = uniq_tf[a + i];
        }
    }
}

// one warp per (term, sub-block) slice: ascending documents. Slices of up to 32 postings are ranked by counting in
// registers, longer ones (<= BUILD_SUB_DOCS) by a bitonic sort of (doc << 32 | tf) keys in the warp's shared memory.
constexpr int SORT_WARPS = 8;
__global__ void __launch_bounds__(32 * SORT_WARPS)
build_sort_slices_kernel(const int64_t* __restrict__ block_ptr, int vocab, int n_sub, int32_t* __restrict__ post_doc,
                         int32_t* __restrict__ post_tf) {
    extern __shared__ unsigned long long sort_buf[];  // [SORT_WARPS][BUILD_SUB_DOCS]
    constexpr uint32_t FULL = 0xFFFFFFFFu;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    unsigned long long* keys = sort_buf + (size_t)w * BUILD_SUB_DOCS;
    const int64_t n_slices = (int64_t)vocab * n_sub;
    const int64_t n_warps = (int64_t)gridDim.x * SORT_WARPS;
    for (int64_t s0 = ((int64_t)blockIdx.x * SORT_WARPS + w) * 32; s0 < n_slices; s0 += n_warps * 32) {
        int64_t lo_l = 0, hi_l = 0;
        if (s0 + lane < n_slices) {
            const int64_t t = (s0 + lane) / n_sub, j = (s0 + lane) - t * n_sub;
            const int64_t* bp = block_ptr + t * (n_sub + 1) + j;
            lo_l = __ldg(bp);
            hi_l = __ldg(bp + 1);
        }
        uint32_t todo = __ballot_sync(FULL, hi_l - lo_l > 1);
        while (todo) {
            const int src = __ffs(todo) - 1;
            todo &= todo - 1;
            const int64_t lo = __shfl_sync(FULL, lo_l, src), hi = __shfl_sync(FULL, hi_l, src);
            const int n = (int)(hi - lo);
            if (n <= 32) {
                const bool valid = lane < n;
                const int32_t d = valid ? post_doc[lo + lane] : 0x7FFFFFFF;
                const int32_t tf = valid ? post_tf[lo + lane] : 0;
                int rank = 0;
                for (int j = 0; j < n; ++j) rank += __shfl_sync(FULL, d, j) < d;  // documents of a slice are distinct
                __syncwarp();
                if (valid) {
                    post_doc[lo + rank] = d;
                    post_tf[lo + rank] = tf;
                }
            } else {
                int P = 64;
                while (P < n) P <<= 1;
                for (int i = lane; i < P; i += 32)
                    keys[i] = i < n ? ((unsigned long long)(uint32_t)post_doc[lo + i] << 32) | (uint32_t)post_tf[lo + i]
                                    : ~0ull;
                __syncwarp();
                for (int size = 2; size <= P; size <<= 1) {
                    for (int stride = size >> 1; stride > 0; stride >>= 1) {
                        for (int t = lane; t < (P >> 1); t += 32) {
                            const int i = 2 * t - (t & (stride - 1));
                            const int j = i + stride;
                            const bool asc = (i & size) == 0;
                            const unsigned long long x = keys[i], y = keys[j];
                            if ((x > y) == asc) { keys[i] = y; keys[j] = x; }
                        }
                        __syncwarp();
                    }
                }
                for (int i = lane; i < n; i += 32) {
                    post_doc[lo + i] = (int32_t)(keys[i] >> 32);
                    post_tf[lo + i] = (int32_t)(uint32_t)keys[i];
                }
            }
            __syncwarp();
        }
    }
}

int grid_for(int64_t items, int per_cta, int cap) {
    int64_t g = (items + per_cta - 1) / per_cta;
    if (g < 1) g = 1;
    return (int)(g < cap ? g : cap);
}

}  // namespace

extern "C" int srag_bm25_build_plan(int32_t* tokens, const int64_t* doc_ptr, int64_t n_docs, int vocab, int32_t* uniq_tf,
                                    int32_t* doc_uniq, int32_t* dl, int64_t* df, int64_t* term_ptr, int64_t* block_ptr,
                                    int* err_flag, void* stream) {
    SRAG_CHECK_ARG(n_docs >= 0 && n_docs < 0x7FFFFFFF && vocab >= 1, "bad sizes n_docs=%lld vocab=%d", (long long)n_docs,
                   vocab);
    SRAG_CHECK_ARG(doc_ptr && doc_uniq && dl && df && term_ptr && block_ptr && err_flag, "null pointer argument");
    SRAG_CHECK_ARG(n_docs == 0 || (tokens && uniq_tf), "null token arrays");
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    int n_sub = (int)((n_docs + BUILD_SUB_DOCS - 1) / BUILD_SUB_DOCS);
    if (n_sub < 1) n_sub = 1;
    {
        int rc = srag_func_smem_optin(reinterpret_cast<const void*>(&build_doc_terms_kernel),
                                      BUILD_LONG_DL * (int)sizeof(uint32_t));
        if (!rc) rc = srag_func_smem_optin(reinterpret_cast<const void*>(&build_sort_slices_kernel),
                                           SORT_WARPS * BUILD_SUB_DOCS * (int)sizeof(unsigned long long));
        if (rc) return rc;
    }
    unsigned long long* cnt = reinterpret_cast<unsigned long long*>(block_ptr);
    SRAG_CUDA(cudaMemsetAsync(cnt, 0, (size_t)vocab * (n_sub + 1) * sizeof(unsigned long long), st));
    SRAG_CUDA(cudaMemsetAsync(err_flag, 0, sizeof(int), st));
    if (n_docs > 0) {
        build_doc_terms_kernel<<<grid_for(n_docs, BUILD_THREADS, 148 * 16), BUILD_THREADS, BUILD_MAX_DL * sizeof(uint32_t), st>>>(
            tokens, doc_ptr, n_docs, vocab, n_sub, uniq_tf, doc_uniq, dl, cnt, err_flag, 0, BUILD_MAX_DL);
        SRAG_CUDA(cudaGetLastError());
        // documents of more than 8192 tokens (rare): same kernel, one CTA per SM with 224 KB of shared memory
        build_doc_terms_kernel<<<grid_for(n_docs, BUILD_THREADS, 148), BUILD_THREADS, BUILD_LONG_DL * sizeof(uint32_t), st>>>(
            tokens, doc_ptr, n_docs, vocab, n_sub, uniq_tf, doc_uniq, dl, cnt, err_flag, BUILD_MAX_DL, BUILD_LONG_DL);
        SRAG_CUDA(cudaGetLastError());
    }
    build_row_sum_kernel<<<grid_for(vocab, 8, 148 * 8), 256, 0, st>>>(cnt, vocab, n_sub, df);
    SRAG_CUDA(cudaGetLastError());
    build_scan_terms_kernel<<<1, 1024, 0, st>>>(df, vocab, term_ptr);
    SRAG_CUDA(cudaGetLastError());
    build_row_scan_kernel<<<grid_for(vocab, 8, 148 * 8), 256, 0, st>>>(cnt, vocab, n_sub, term_ptr);
    SRAG_CUDA(cudaGetLastError());
    return SRAG_OK;
}

extern "C" int srag_bm25_build_fill(const int32_t* uniq_term, const int32_t* uniq_tf, const int64_t* doc_ptr,
                                    const int32_t* doc_uniq, int64_t n_docs, int vocab, const int64_t* block_ptr,
                                    uint32_t* fill, int32_t* post_doc, int32_t* post_tf, void* stream) {
    SRAG_CHECK_ARG(n_docs >= 0 && n_docs < 0x7FFFFFFF && vocab >= 1, "bad sizes n_docs=%lld vocab=%d", (long long)n_docs,
                   vocab);
    if (n_docs == 0) return SRAG_OK;
    SRAG_CHECK_ARG(uniq_term && uniq_tf && doc_ptr && doc_uniq && block_ptr && fill && post_doc && post_tf,
                   "null pointer argument");
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    const int n_sub = (int)((n_docs + BUILD_SUB_DOCS - 1) / BUILD_SUB_DOCS);
    SRAG_CUDA(cudaMemsetAsync(fill, 0, (size_t)vocab * n_sub * sizeof(uint32_t), st));
    build_scatter_kernel<<<grid_for(n_docs, 8, 148 * 16), 256, 0, st>>>(uniq_term, uniq_tf, doc_ptr, doc_uniq, n_docs, n_sub,
                                                                       block_ptr, fill, post_doc, post_tf);
    SRAG_CUDA(cudaGetLastError());
    const int64_t slices = (int64_t)vocab * n_sub;
    build_sort_slices_kernel<<<grid_for(slices, 32 * SORT_WARPS, 148 * 8), 32 * SORT_WARPS,
                               SORT_WARPS * BUILD_SUB_DOCS * sizeof(unsigned long long), st>>>(block_ptr, vocab, n_sub,
                                                                                                post_doc, post_tf);
    SRAG_CUDA(cudaGetLastError());
    return SRAG_OK;
}
