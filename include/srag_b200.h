/*
 * srag_b200.h — C ABI of the B200 hybrid-retrieval kernels (libsrag_b200.so).
 *
 * The reference (promoteAI/super-rag, 100% Python) has no FFI; each entry point below
 * names the reference interface whose arithmetic it replaces. Paths are relative to the
 * reference root. INTEGRATION.md shows the ctypes binding a maintainer would add.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer owned by the caller (a torch.Tensor's data_ptr())
 *     unless the name ends in _host; the library never allocates or frees caller-visible
 *     memory; scratch space is a caller-provided workspace sized by *_workspace_bytes();
 *   - `stream` is a cudaStream_t passed as void*; launches are asynchronous with respect
 *     to the host, the caller synchronises;
 *   - return value: 0 = ok, negative = error (SRAG_E_*); srag_last_error() gives the text
 *     for the calling thread. No exceptions cross the boundary;
 *   - ordering everywhere: score descending, id ascending on ties (north_star);
 *   - empty result slots are (score = -inf, id = -1).
 */
#ifndef SRAG_B200_H
#define SRAG_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SRAG_ABI_VERSION 2

#define SRAG_OK              0
#define SRAG_E_INVALID      -1   /* bad argument (null pointer, size, alignment, k) */
#define SRAG_E_CUDA         -2   /* a CUDA runtime / driver call failed */
#define SRAG_E_WORKSPACE    -3   /* workspace too small */
#define SRAG_E_UNSUPPORTED  -4   /* device is not sm_100 or driver lacks a needed entry point */

#define SRAG_METRIC_COSINE 0     /* seekdb_connector.py:38  distance="cosine" */
#define SRAG_METRIC_IP     1

#define SRAG_FUSE_RRF      0     /* graphiti_core/search/search_utils.py:1763-1778 */
#define SRAG_FUSE_WEIGHTED 1

#define SRAG_MAX_TOPK      128   /* k <= 128 for vector search */
#define SRAG_MAX_BM25_K    256
#define SRAG_MAX_LISTS     8     /* fused lists */

int         srag_abi_version(void);
const char* srag_last_error(void);
/* sm count, compute capability of the current device */
int         srag_device_info(int* sm_count, int* cc_major, int* cc_minor);

/* ---- K0: ingest -------------------------------------------------------------------
 * Replaces the server side of `collection.add(ids, embeddings=...)`
 * (super_rag/vectorstore/seekdb_connector.py:68-85): cast fp32 rows to bf16 and compute the
 * per-row score multiplier (1/|row| of the bf16-ROUNDED row for cosine, 0 for a zero row —
 * graphiti_core/search/search_utils.py:63-64; 1.0 for ip).
 *   src_f32  [n, dim] fp32      dst_bf16 [n, dim] bf16      scale [n] fp32            */
int srag_rows_ingest_f32(const float* src_f32, int64_t n, int dim, void* dst_bf16,
                         float* scale, int metric, void* stream);
/* same, for rows that are already bf16 */
int srag_rows_scale_bf16(const void* rows_bf16, int64_t n, int dim, float* scale,
                         int metric, void* stream);

/* ---- K1: exact top-k vector search --------------------------------------------------
 * Replaces `collection.query(query_embeddings, n_results)` called at
 * super_rag/vectorstore/seekdb_connector.py:103-107 (one query) with a batched exact scan:
 *   score[b, r] = (q_b . c_r) * q_scale[b] * row_scale[r]
 * computed as a tcgen05 bf16 GEMM with fp32 accumulation in TMEM; scores never reach HBM.
 * A row whose row_scale is NaN is a tombstone (connector.delete, seekdb_connector.py:90-96)
 * and is never returned.
 *   corpus_bf16 [n_rows, dim] bf16 row-major, 16-byte aligned, dim % 8 == 0
 *   queries_bf16 [n_queries, dim] bf16, same constraints
 *   out_scores [n_queries, k] fp32, out_ids [n_queries, k] int64 = id_base + row
 *   1 <= k <= SRAG_MAX_TOPK                                                           */
int srag_topk_workspace_bytes(int64_t n_rows, int n_queries, int k, size_t* bytes);
/* Same plus room for a zero-padded copy of the queries (whole 128 / 256-row query tiles). Give the
 * search a workspace of this size and small or ragged batches avoid mostly-out-of-bounds TMA boxes
 * (1.6 x faster at B = 1); with the smaller size above the search still works, unpadded.          */
int srag_topk_workspace_bytes_ex(int64_t n_rows, int dim, int n_queries, int k, size_t* bytes);
int srag_topk_search(const void* corpus_bf16, const float* row_scale, int64_t n_rows, int dim,
                     const void* queries_bf16, const float* q_scale, int n_queries, int k,
                     int64_t id_base, float* out_scores, int64_t* out_ids,
                     void* workspace, size_t workspace_bytes, void* stream);

/* Same scan, additionally (or only: out_scores = out_ids = NULL) emitting the cross-shard EXCHANGE KEYS of SURVEY §8(e):
 * out_keys uint64 [n_queries, k], key = orderable(final fp32 score) << 32 | (0xFFFFFFFF - (id_base + row)), 0 = empty
 * slot, so that key descending <=> (score descending, id ascending). The finalize kernel writes them straight into
 * the buffer the caller hands to the all-gather; srag_merge_keys consumes the gathered buffer. Needs
 * id_base + n_rows < 2^32 - 1 (SRAG_E_INVALID otherwise: use out_scores / out_ids + srag_merge_candidates).       */
int srag_topk_search_keys(const void* corpus_bf16, const float* row_scale, int64_t n_rows, int dim,
                          const void* queries_bf16, const float* q_scale, int n_queries, int k,
                          int64_t id_base, uint64_t* out_keys, float* out_scores, int64_t* out_ids,
                          void* workspace, size_t workspace_bytes, void* stream);

/* The same scan for callers that may not allocate device memory (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc). It has NO COUNTERPART IN THE REFERENCE.
 * Identical to the plain search entry point above except that the scratch memory (per-thread heaps, shared
 * thresholds, pacing counters, the zero-padded query copy) is a file-scope `static __device__` array inside the library of
 * SRAG_TOPK_STATIC_SCRATCH_BYTES bytes — every GPU that loads the library holds that much until the process exits — so the
 * call needs no workspace and makes no cudaMalloc / cudaFree / allocating library call (the array's device address is
 * looked up once per device with cudaGetSymbolAddress). A search whose scratch need (what srag_topk_workspace_bytes
 * reports) exceeds the array is rejected with SRAG_E_WORKSPACE and a message that names both sizes: 8 MiB covers
 * n_queries <= 1024 at k <= 10 over any corpus size (5.9 MiB), or k <= 16 at n_queries <= 512; larger k needs the caller-
 * workspace entry point. ALL calls on a device share the one array and it keeps the last call's contents: every call
 * clears the part it uses on `stream` before the scan, and two calls that overlap on the GPU (different streams, or
 * different host threads) RACE — serialise them.                                                                        */
#define SRAG_TOPK_STATIC_SCRATCH_BYTES (8u << 20)
int srag_topk_search_static(const void* corpus_bf16, const float* row_scale, int64_t n_rows, int dim,
                            const void* queries_bf16, const float* q_scale, int n_queries, int k,
                            int64_t id_base, float* out_scores, int64_t* out_ids, void* stream);

/* Same scan with the metadata pre-filter the reference BUILDS but its connector drops
 * (ContextManager._create_combined_filter, super_rag/context/context.py:74-111;
 * seekdb_connector.py:99-100): row r is eligible for query b iff
 *   (row_tags  == NULL || (row_tags[r] & q_tag_mask[b]) != 0)            -- `indexer` $in / $exists:false
 *   (row_group == NULL || q_group[b] < 0 || row_group[r] == q_group[b])  -- `chat_id` equality
 * Ineligible rows are never returned; the result is the exact top-k of the eligible rows.
 *   row_tags [n_rows] uint32, row_group [n_rows] int32, q_tag_mask [n_queries] uint32,
 *   q_group [n_queries] int32 (device pointers; each row_* comes with its q_* or both are NULL) */
int srag_topk_search_filtered(const void* corpus_bf16, const float* row_scale, int64_t n_rows, int dim,
                              const void* queries_bf16, const float* q_scale, int n_queries, int k,
                              int64_t id_base, const uint32_t* row_tags, const int32_t* row_group,
                              const uint32_t* q_tag_mask, const int32_t* q_group, float* out_scores,
                              int64_t* out_ids, void* workspace, size_t workspace_bytes, void* stream);

/* Device-side timing of K1 for roofline reporting: when enabled, srag_topk_search records CUDA
 * events around its main kernel (memset + fused GEMM/top-k) and its finalize kernel on the
 * launching stream (up to 256 calls). srag_prof_read waits for them and returns the sums.      */
int srag_prof_enable(int on);
int srag_prof_read(double* main_ms_sum, double* finalize_ms_sum, int* n_calls);

/* ---- K2: BM25 -----------------------------------------------------------------------
 * The `fulltext_search` node the reference declares but does not implement
 * (super_rag/nodeflow/runners/merge.py:18-20, super_rag/schema/view_models.py:1043-1047).
 * Index layout (CSR by term). While it is BUILT: term_ptr int64 [vocab+1], post_doc int32 [nnz] ascending inside a
 * term, post_tf int32 [nnz], block_ptr int64 [vocab, n_sub+1]. As it is SEARCHED (srag_bm25_pack_plan / _fill):
 *   slice_ptr uint32 [vocab, n_sub+1]  start of every (term, 2048-document sub-block) slice in GROUPS of 8 postings: every
 *                                      slice starts on a group and is padded to whole groups with zero-impact postings;
 *   post_off  uint16 [8 * groups]      (doc & 16383) * 4 = byte offset of the document's fp32 accumulator inside the tile
 *                                      of its 16384-document block (the slice names the block); padding: a trash offset;
 *   post_imp  fp32   [8 * groups]      impact (> 0; exactly 0 marks padding)
 * = 6 bytes per posting in HBM. Both arrays are padded by 8 entries. Numeric contract: oracle/bm25.py docstring.
 *
 * srag_bm25_impacts: imp = idf[term] * tf*(k1+1) / (tf + k1*(1-b+b*dl/avgdl)), one IEEE
 * fp32 op per step; the term of posting p is found by bisection in term_ptr.
 * post_tf int32 [nnz]; idf fp32 [vocab]; dl int32 [n_docs].                                */
int srag_bm25_impacts(const int64_t* term_ptr, int vocab, const int32_t* post_doc,
                      const int32_t* post_tf, int64_t nnz, const float* idf, const int32_t* dl,
                      float avgdl, float* post_imp, void* stream);
/* block_ptr int64 [vocab, n_blocks+1]: absolute offset of the first posting of term t whose
 * doc >= j * block_docs (j = n_blocks gives term_ptr[t+1]).                               */
int srag_bm25_block_index(const int64_t* term_ptr, const int32_t* post_doc, int vocab,
                          int64_t n_docs, int block_docs, int64_t* block_ptr, void* stream);
int srag_bm25_block_docs(void);   /* doc-range width one CTA accumulates in shared memory */
/* Index-build steps after srag_bm25_block_index: canonical postings -> the searched layout.
 *   srag_bm25_pack_plan: slice_ptr from block_ptr (three small kernels: per-term group counts, a scan over the
 *     vocabulary, a scan along every term). scratch int64 [2 * vocab + 1] (device); afterwards scratch[2 * vocab] is the
 *     total number of groups (the caller reads it back to size post_off / post_imp = 8 * groups + 8 entries).
 *   srag_bm25_pack_fill: writes post_off / post_imp. With permute != 0 the postings INSIDE every slice are laid out so that
 *     the 32 accumulators one warp of the scoring kernel touches per shared-memory access fall on distinct banks where
 *     the slice allows it (every posting of a slice is a distinct document that receives one add per query, so scores
 *     are unchanged bit for bit); permute = 0 keeps ascending documents.                                            */
int srag_bm25_pack_plan(const int64_t* block_ptr, int vocab, int64_t n_docs, int64_t* scratch, uint32_t* slice_ptr,
                        void* stream);
int srag_bm25_pack_fill(const int64_t* block_ptr, const uint32_t* slice_ptr, int vocab, int64_t n_docs,
                        const int32_t* doc_in, const float* imp_in, uint16_t* off_out, float* imp_out, int permute,
                        void* stream);
/* The inverse: doc_out int32 [8 * groups] = document of every stored posting (-1 for padding), term_out (may be NULL) its
 * term. Not on the search path: tests, canonical postings, shard conversion.                                      */
int srag_bm25_unpack_docs(const uint32_t* slice_ptr, int vocab, int64_t n_docs, const uint16_t* post_off,
                          const float* post_imp, int32_t* doc_out, int32_t* term_out, void* stream);
int srag_bm25_workspace_bytes(int64_t n_docs, int n_queries, int k, size_t* bytes);
/* Index build from tokens without a global sort (the reference has no index build, contract:
 * oracle/bm25.py build_postings). Two calls, because the number of postings is only known after the first:
 *   srag_bm25_build_plan: tokens int32 [n_tokens] with term ids in [0, vocab), documents d = tokens[doc_ptr[d] ..
 *     doc_ptr[d+1]) of at most 57344 tokens. Every document's tokens are sorted and run-length encoded IN PLACE:
 *     afterwards tokens[doc_ptr[d] + i], uniq_tf[doc_ptr[d] + i] (i < doc_uniq[d]) are its distinct terms and their
 *     frequencies. Also written: dl int32 [n_docs], df int64 [vocab], term_ptr int64 [vocab+1] (nnz = term_ptr[vocab])
 *     and block_ptr int64 [vocab, n_sub+1] (n_sub = ceil(n_docs / srag_bm25_block_docs()), as srag_bm25_block_index
 *     would produce it). *err_flag (device int): 0 ok, 1 a document is too long, 2 a term id is out of range.
 *   srag_bm25_build_fill: post_doc / post_tf int32 [nnz] (ascending documents inside a term); fill uint32
 *     [vocab * n_sub] is scratch.                                                                              */
int srag_bm25_build_plan(int32_t* tokens, const int64_t* doc_ptr, int64_t n_docs, int vocab, int32_t* uniq_tf,
                         int32_t* doc_uniq, int32_t* dl, int64_t* df, int64_t* term_ptr, int64_t* block_ptr,
                         int* err_flag, void* stream);
int srag_bm25_build_fill(const int32_t* uniq_term, const int32_t* uniq_tf, const int64_t* doc_ptr,
                         const int32_t* doc_uniq, int64_t n_docs, int vocab, const int64_t* block_ptr,
                         uint32_t* fill, int32_t* post_doc, int32_t* post_tf, void* stream);
/* q_terms int32 [n_queries, max_terms]: unique term ids in ASCENDING order, -1 padded.
 * out_* [n_queries, k]; only documents matching >= 1 term are candidates.               */
int srag_bm25_search(const uint32_t* slice_ptr, const uint16_t* post_off, const float* post_imp,
                     int vocab, int64_t n_docs, const int32_t* q_terms, int n_queries,
                     int max_terms, int k, int64_t id_base, float* out_scores, int64_t* out_ids,
                     void* workspace, size_t workspace_bytes, void* stream);
/* Same with the metadata pre-filter of srag_topk_search_filtered applied to documents: only admitted
 * documents are scored into the result (doc_tags / doc_group are indexed by the document's position in
 * THIS index, 0 .. n_docs-1; q_tag_mask / q_group per query).                                    */
int srag_bm25_search_filtered(const uint32_t* slice_ptr, const uint16_t* post_off, const float* post_imp,
                              int vocab, int64_t n_docs, const int32_t* q_terms, int n_queries,
                              int max_terms, int k, int64_t id_base, const uint32_t* doc_tags,
                              const int32_t* doc_group, const uint32_t* q_tag_mask, const int32_t* q_group,
                              float* out_scores, int64_t* out_ids, void* workspace,
                              size_t workspace_bytes, void* stream);

/* ---- K3: fusion ---------------------------------------------------------------------
 * RRF exactly as graphiti_core/search/search_utils.py:1763-1778 (float64,
 * score[id] += 1/(i + rank_const), lists accumulated in order, keep score >= min_score) or
 * weighted fp32 fusion, for the `merge` node (super_rag/nodeflow/runners/merge.py:40-65)
 * strategies "rrf" / "weighted". Final order (score desc, id asc), cut to k_out.
 *   ids    int64 [n_lists, n_queries, list_len]  (id < 0 = padding)
 *   scores fp32  [n_lists, n_queries, list_len]  (may be NULL for RRF)
 *   weights_host  double[n_lists] on the HOST (NULL for RRF)
 *   out_ids int64 [n_queries, k_out], out_scores double [n_queries, k_out]
 *   n_lists <= SRAG_MAX_LISTS, n_lists * list_len <= 2048                               */
int srag_fuse(const int64_t* ids, const float* scores, int n_lists, int n_queries, int list_len,
              int mode, int rank_const, double min_score, const double* weights_host,
              int k_out, int64_t* out_ids, double* out_scores, void* stream);

/* ---- K5: maximal-marginal-relevance rerank ---------------------------------------------
 * Replaces `maximal_marginal_relevance(query_vector, candidates, mmr_lambda, min_score)`
 * (graphiti_core/search/search_utils.py:1884-1922) for candidates that are rows of the device
 * corpus (ids as returned by srag_topk_search / srag_fuse):
 *   c_i = row_i * row_scale[row_i]   (row_scale from srag_rows_ingest_f32 with the cosine metric)
 *   mmr_i = lambda * (q . c_i) * q_scale + (lambda - 1) * max(0, max_{j != i} c_i . c_j)
 * out_ids / out_scores [n_queries, m]: candidates with mmr_i >= min_score by score descending,
 * ties in candidate order (the reference's stable sort); the rest of each row is -1 / -inf.
 * Ids that are negative, out of range or tombstoned are skipped. q_scale = 1 keeps the
 * reference's un-normalised query. 1 <= m <= SRAG_MAX_MMR.                                   */
#define SRAG_MAX_MMR 256
int srag_mmr_rerank(const void* corpus_bf16, const float* row_scale, int64_t n_rows, int dim,
                    const void* queries_bf16, const float* q_scale, int n_queries,
                    const int64_t* cand_ids, int m, int64_t id_base, float mmr_lambda, float min_score,
                    int64_t* out_ids, float* out_scores, void* stream);

/* ---- K4: cross-shard candidate merge -------------------------------------------------
 * Final step after the all-gather of per-shard candidates.
 *   scores fp32 [n_shards, n_queries, kk], ids int64 same shape (id < 0 = padding)
 *   out_* [n_queries, k];  n_shards * kk <= 2048                                        */
/* The same merge on exchange keys (srag_topk_search_keys / srag_pack_keys): keys uint64 [n_shards, n_queries, kk] as the
 * all-gather leaves them; 8 bytes per candidate on the wire instead of 12.                                   */
int srag_merge_keys(const uint64_t* keys, int n_shards, int n_queries, int kk, int k, float* out_scores,
                    int64_t* out_ids, void* stream);
/* The exchange AND the merge in one kernel over peer memory (NVLink / NVSwitch P2P stores; no collective call, SURVEY
 * §5 "peer-to-peer ... is a legitimate alternative to measure against NCCL"). Every rank owns a buffer of the SAME layout
 * that every peer can address (torch symmetric memory / cudaIpc):
 *     keys uint64 [2][world][cap_keys] followed anywhere by flags uint32 [world][cap_q], zero-initialised once.
 * peer_keys_host / peer_flags_host: HOST arrays of `world` device addresses (as uint64) — this process's mapping of every
 * rank's key / flag region, index = rank. Step `epoch` = 1, 2, 3, ... (the same value on every rank, one step at a
 * time per buffer): each CTA first stores its queries' keys (local_keys uint64 [n_queries, k], as written by
 * srag_topk_search_keys) into slot [epoch & 1][rank] of every rank and then raises flag [rank][q] = epoch there; then
 * it waits for the `world` flags of each of its queries and merges world * k keys (<= 2048) into out_* [n_queries, k].   */
int srag_exchange_merge_p2p(const uint64_t* local_keys, const uint64_t* peer_keys_host, const uint64_t* peer_flags_host,
                            int world, int rank, int n_queries, int k, uint32_t epoch, size_t cap_keys, int cap_q,
                            float* out_scores, int64_t* out_ids, void* stream);
/* (score, id) -> exchange key, for candidate lists that were not produced by srag_topk_search_keys (BM25 shards).
 * ids must be < 2^32 - 1; *err_flag (device int, may be NULL) is set to 1 otherwise. id < 0 -> empty key.        */
int srag_pack_keys(const float* scores, const int64_t* ids, int64_t n, uint64_t* keys, int* err_flag, void* stream);
int srag_merge_candidates(const float* scores, const int64_t* ids, int n_shards, int n_queries,
                          int kk, int k, float* out_scores, int64_t* out_ids, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* SRAG_B200_H */
