// K2: BM25 scoring over CSR postings. HBM-bound integer/float streaming work — no tensor cores.
//
// Postings are CSR by term: post_off uint16 [nnz] = the document's BYTE offset inside its 16384-document block's fp32
// accumulator tile ((doc & 16383) * 4; the slice a posting belongs to names the block), post_imp fp32 [nnz] = its impact:
// 6 bytes per posting in HBM. block_ptr gives the first posting of every (term, BM25_SUB_DOCS = 2048 documents) slice.
// Query terms are applied in ASCENDING term id; inside a term every posting hits a distinct document, so plain
// (non-atomic) read-add-write is race-free and the per-document sum order is fixed -> bit-exact with oracle/bm25.py.
// Selection is two-pass: a SAMPLE pass over the sub-blocks of every 32nd / 16th block reports a few true document scores
// per sampled sub-block, their k-th best tau bounds the query's global k-th score from below, and the MAIN pass over
// all documents only reports the accumulators that cross tau.
//
//   bm25_sub_kernel    both passes, nearly all of the bytes: one WARP per (query, 2048-document sub-block), 8 KB of private
//                      accumulators, no CTA barrier, no producer; documents crossing tau are appended to the query's
//                      candidate region. Overflowing queries are flagged and redone by bm25_items_kernel.
//   bm25_items_kernel  exact top-k of every (query, 16384-document block): indexes too small to sample, flagged queries
//                      (and the whole of the first design, SRAG_BM25_SUBKERNEL=0). Persistent CTAs (2 per SM); the block's
//                      fp32 accumulators live in 64 KB of shared memory; two producer warps stream the posting slices
//                      through a packed shared-memory ring with 1-D bulk copies, eight consumer warps accumulate and select.
//   bm25_bank_order_kernel  index build: permutes every slice so that the accumulator accesses of a warp fall on
//                      distinct shared-memory banks (scores unchanged: the order inside a slice is free).
//   srag_launch_finalize(_ex/_live) (topk_gemm.cu) merges the lists / candidate region of a query.
//
// Algorithmic bytes per query = 8 * sum_t df(t)  (int32 doc + fp32 impact per posting); the index itself
// stores 6 (16-bit offset + fp32 impact), so HBM traffic is below the algorithmic figure.
// The reference has no BM25 arithmetic; contract: oracle/bm25.py docstring.
#include <stdlib.h>

#include "common.cuh"

namespace {

// Geometry (compile time): consumer warps per CTA decide the rest.
//   8 warps: 16384-document tile (64 KB), 36 KB posting ring, 2 CTAs per SM   <- default (13.1 ms at C3)
//   4 warps:  8192-document tile (32 KB), 24 KB posting ring, 3 CTAs per SM   (18.2 ms: per-item costs dominate)
#ifndef SRAG_BM25_WARPS
#define SRAG_BM25_WARPS 8
#endif
constexpr int BM25_BLOCK_DOCS = SRAG_BM25_WARPS == 4 ? 8192 : 16384;  // fp32 accumulators per CTA
// Index granularity: block_ptr holds one offset per (term, BM25_SUB_DOCS documents). A tile of bm25_items_kernel is
// BM25_SUBS consecutive sub-blocks; bm25_sub_kernel gives one sub-block to one warp.
#ifndef SRAG_BM25_SUB_DOCS
#define SRAG_BM25_SUB_DOCS 2048
#endif
constexpr int BM25_SUB_DOCS = SRAG_BM25_SUB_DOCS;   // (a build-time knob for A/B runs: the index layout follows it)
// Slices of the searched layout start on, and are padded to, whole GROUPS of BM25_GROUP postings (16 bytes of 16-bit
// offsets, 32 bytes of impacts: the unit of a bulk copy); slice_ptr counts groups.
constexpr int BM25_GROUP = 8;
constexpr int BM25_GROUP_LOG = 3;
constexpr int BM25_SUBS = BM25_BLOCK_DOCS / BM25_SUB_DOCS;
constexpr int BM25_THREADS = 32 * SRAG_BM25_WARPS;  // consumer threads
constexpr int RADIX_BINS = 2048;
constexpr int SLOTS_PER_THREAD = BM25_BLOCK_DOCS / BM25_THREADS;  // 64 contiguous slots (ordered tie path only)
constexpr int BM25_CTA_THREADS = BM25_THREADS + 64;  // consumer warps + 2 producer warps (alternate items)
constexpr int BM25_CHUNK = 4 * BM25_THREADS;         // postings per chunk: one aligned group of 4 per consumer thread
constexpr int BM25_CTAS_PER_SM = SRAG_BM25_WARPS == 4 ? 3 : 2;
// posting ring capacity (x 8 bytes: doc ids + impacts)
constexpr int BM25_RING = SRAG_BM25_WARPS == 4 ? 3072 : 4608;
constexpr int BM25_SLOTS = 32;                       // chunk descriptors in flight
constexpr int BM25_CAND_MAX = 1024;

// stream flags of a ring slot
constexpr uint32_t SLOT_NEW_ITEM = 1u;  // first chunk of a (query, block) item
constexpr uint32_t SLOT_NEW_TERM = 2u;  // first chunk of a term that needs a barrier before it (see the producer)
constexpr uint32_t SLOT_END = 4u;       // no data: the CTA's item list is exhausted

struct SlotMeta {  // written by the producer before it issues the slot's copies
    int32_t ql, blk;         // query (inside the chunk) and document block
    int32_t lo_rel, hi_rel;  // valid postings of the chunk: [lo_rel, hi_rel)
    uint32_t flags;
    float tau;               // sample k-th best score of the query (0 = none)
    uint32_t pos;            // ring position of the chunk's first posting (multiple of 8)
    uint32_t n;              // postings copied (multiple of 8)
};

__device__ __forceinline__ void consumer_sync() { ptx::named_bar_sync(1, BM25_THREADS); }

struct ProducerState {  // ring bookkeeping, handed between the two producer warps with the turn
    uint32_t slot, pos, inflight, free_post, tail_slot, tail_par;
    uint32_t turn;                    // sequence number (in the CTA's item list) of the item whose chunks go next
    uint32_t pad;
    uint32_t slot_post[BM25_SLOTS];   // postings of the chunk occupying each descriptor slot
};

struct Bm25Smem {
    float acc[BM25_BLOCK_DOCS];
    uint32_t hist[RADIX_BINS];
    alignas(128) uint16_t ring_doc[BM25_RING];   // byte offsets into acc[] (block-relative)
    alignas(128) float ring_imp[BM25_RING];
    SlotMeta meta[BM25_SLOTS];
    uint64_t full[BM25_SLOTS], empty[BM25_SLOTS];
    ProducerState prod;
    uint32_t warp_tmp[BM25_THREADS / 32];
    uint32_t sel_bin;
    uint32_t sel_above;
    uint32_t out_count;
    uint32_t nonzero;
    uint32_t cand_count;              // documents whose accumulator crossed tau while terms were applied
    uint16_t cand[BM25_CAND_MAX];     // their offsets inside the block
};

__device__ __forceinline__ uint32_t block_excl_scan(uint32_t v, uint32_t* warp_tmp, uint32_t* total) {
    // exclusive prefix sum over the CTA in thread order
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    uint32_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, inc, o);
        if (lane >= o) inc += t;
    }
    consumer_sync();  // protect warp_tmp reuse
    if (lane == 31) warp_tmp[w] = inc;
    consumer_sync();
    uint32_t base = 0, tot = 0;
#pragma unroll
    for (int i = 0; i < BM25_THREADS / 32; ++i) {
        const uint32_t x = warp_tmp[i];
        if (i < w) base += x;
        tot += x;
    }
    *total = tot;
    return base + inc - v;
}

// One radix pass: histogram digit (bits >> shift) & mask of the accumulator entries whose bits
// above this digit equal `prefix` (prefix_mask selects them), then locate the bin in which the
// `want`-th largest such entry falls. Returns through smem: sel_bin, sel_above (# in higher bins).
__device__ void radix_pass(Bm25Smem* sm, uint32_t prefix, uint32_t prefix_mask, int shift, uint32_t mask,
                           uint32_t want) {
    for (int i = threadIdx.x; i < RADIX_BINS; i += BM25_THREADS) sm->hist[i] = 0;
    consumer_sync();
    const float4* a4 = reinterpret_cast<const float4*>(sm->acc);
    for (int i = threadIdx.x; i < BM25_BLOCK_DOCS / 4; i += BM25_THREADS) {
        const float4 v = a4[i];
        const uint32_t b[4] = {__float_as_uint(v.x), __float_as_uint(v.y), __float_as_uint(v.z), __float_as_uint(v.w)};
#pragma unroll
        for (int j = 0; j < 4; ++j)
            if (b[j] != 0 && (b[j] & prefix_mask) == prefix) atomicAdd(&sm->hist[(b[j] >> shift) & mask], 1u);
    }
    consumer_sync();
    // suffix scan from the top bin: thread t owns bins [2047 - 8t - 7, 2047 - 8t]
    constexpr int PER = RADIX_BINS / BM25_THREADS;
    uint32_t mine = 0;
    const int top = RADIX_BINS - 1 - threadIdx.x * PER;
#pragma unroll
    for (int i = 0; i < PER; ++i) mine += sm->hist[top - i];
    uint32_t total;
    const uint32_t above = block_excl_scan(mine, sm->warp_tmp, &total);
    if (above < want && above + mine >= want) {
        uint32_t run = above;
        for (int i = 0; i < PER; ++i) {
            const uint32_t h = sm->hist[top - i];
            if (run + h >= want) {
                sm->sel_bin = (uint32_t)(top - i);
                sm->sel_above = run;
                break;
            }
            run += h;
        }
    }
    consumer_sync();
}

// Exact top-k of the accumulator tile -> k sortable keys at `out` (unordered; finalize sorts).
// 11/11/10-bit radix select finds the k-th largest value; a tie that straddles position k is
// resolved towards the lowest doc ids (ordered path, rare).
__device__ void select_exact(Bm25Smem* sm, int k, int32_t doc_base, uint32_t total_nz, uint64_t* out) {
    const float4* a4 = reinterpret_cast<const float4*>(sm->acc);
    uint32_t thr_bits = 0;           // k-th largest value; everything above it is taken
    uint32_t need_eq = 0xFFFFFFFFu;  // how many entries == thr_bits to take (lowest doc ids first)
    bool ordered = false;
    if (total_nz > (uint32_t)k) {
        radix_pass(sm, 0u, 0u, 21, 0x7FFu, (uint32_t)k);
        const uint32_t b1 = sm->sel_bin, above1 = sm->sel_above;
        consumer_sync();
        radix_pass(sm, b1 << 21, 0xFFE00000u, 10, 0x7FFu, (uint32_t)k - above1);
        const uint32_t b2 = sm->sel_bin, above2 = above1 + sm->sel_above;
        consumer_sync();
        radix_pass(sm, (b1 << 21) | (b2 << 10), 0xFFFFFC00u, 0, 0x3FFu, (uint32_t)k - above2);
        const uint32_t b3 = sm->sel_bin, above3 = above2 + sm->sel_above;
        consumer_sync();
        thr_bits = (b1 << 21) | (b2 << 10) | b3;
        need_eq = (uint32_t)k - above3;
        // last-level bin == exact value, so its population is the number of entries == thr_bits
        ordered = sm->hist[b3] > need_eq;
    }
    if (!ordered) {
        // take every entry >= thr_bits (exactly k of them, or all matches when fewer than k): bank-conflict-free
        const uint32_t lo = thr_bits ? thr_bits : 1u;
        for (int i = threadIdx.x; i < BM25_BLOCK_DOCS / 4; i += BM25_THREADS) {
            const float4 v = a4[i];
            const uint32_t b[4] = {__float_as_uint(v.x), __float_as_uint(v.y), __float_as_uint(v.z), __float_as_uint(v.w)};
#pragma unroll
            for (int j = 0; j < 4; ++j)
                if (b[j] >= lo) {
                    const uint32_t slot = atomicAdd(&sm->out_count, 1u);
                    if (slot < (uint32_t)k) out[slot] = srag_pack_key(b[j], (uint32_t)(doc_base + 4 * i + j));
                }
        }
        return;
    }
    // ordered path: thread t owns 64 consecutive documents, so a prefix sum in thread order is doc order
    const int s0 = threadIdx.x * SLOTS_PER_THREAD;
    uint32_t eq = 0;
    for (int i = 0; i < SLOTS_PER_THREAD; ++i) eq += (__float_as_uint(sm->acc[s0 + i]) == thr_bits);
    uint32_t eq_total;
    uint32_t eq_rank = block_excl_scan(eq, sm->warp_tmp, &eq_total);
    for (int i = 0; i < SLOTS_PER_THREAD; ++i) {
        const uint32_t b = __float_as_uint(sm->acc[s0 + i]);
        if (b == 0) continue;
        bool take = b > thr_bits;
        if (b == thr_bits) { take = eq_rank < need_eq; ++eq_rank; }
        if (take) {
            const uint32_t slot = atomicAdd(&sm->out_count, 1u);
            if (slot < (uint32_t)k) out[slot] = srag_pack_key(b, (uint32_t)(doc_base + s0 + i));
        }
    }
}

// Fast exact top-k for a tile whose matches are dense (phase 0, and phase-1 fallbacks): BM25 scores of one
// block sit in a few binades, so a radix pass over the top bits funnels every match into a handful of
// histogram bins (serialised shared-memory atomics). Here the histogram is taken over the DISTANCE to the
// tile maximum in units of 2^13 ulps: the k best are spread over many bins near 0, everything further than
// two binades below the maximum is only counted. One scan for the maximum, one for the histogram, one to
// emit; the bin holding the k-th value is resolved exactly by ranking its few members (ties -> lowest doc id).
// Returns false, with nothing emitted, when the k-th value is not within reach (caller falls back to the
// radix select). Precondition: more than k matches in the tile, out_count == 0.
constexpr int NEAR_SHIFT = 13;
__device__ bool select_near_max(Bm25Smem* sm, int k, int32_t doc_base, uint64_t* out) {
    const float4* a4 = reinterpret_cast<const float4*>(sm->acc);
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    uint32_t mx = 0;
    for (int i = threadIdx.x; i < BM25_BLOCK_DOCS / 4; i += BM25_THREADS) {
        const float4 v = a4[i];
        mx = max(max(mx, __float_as_uint(v.x)), max(__float_as_uint(v.y), max(__float_as_uint(v.z), __float_as_uint(v.w))));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = max(mx, __shfl_xor_sync(0xFFFFFFFFu, mx, o));
    for (int i = threadIdx.x; i < RADIX_BINS; i += BM25_THREADS) sm->hist[i] = 0;
    consumer_sync();  // also protects warp_tmp reuse
    if (lane == 0) sm->warp_tmp[w] = mx;
    consumer_sync();
#pragma unroll
    for (int i = 0; i < BM25_THREADS / 32; ++i) mx = max(mx, sm->warp_tmp[i]);
    for (int i = threadIdx.x; i < BM25_BLOCK_DOCS / 4; i += BM25_THREADS) {
        const float4 v = a4[i];
        const uint32_t b[4] = {__float_as_uint(v.x), __float_as_uint(v.y), __float_as_uint(v.z), __float_as_uint(v.w)};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const uint32_t bin = (mx - b[j]) >> NEAR_SHIFT;
            if (b[j] != 0 && bin < (uint32_t)RADIX_BINS) atomicAdd(&sm->hist[bin], 1u);
        }
    }
    consumer_sync();
    // prefix scan from bin 0 (nearest the maximum): thread t owns bins [PER t, PER t + PER)
    constexpr int PER = RADIX_BINS / BM25_THREADS;
    uint32_t mine = 0;
#pragma unroll
    for (int i = 0; i < PER; ++i) mine += sm->hist[threadIdx.x * PER + i];
    uint32_t total;
    const uint32_t above = block_excl_scan(mine, sm->warp_tmp, &total);
    if (total < (uint32_t)k) return false;  // uniform: the k-th value lies beyond the histogram's reach
    if (above < (uint32_t)k && above + mine >= (uint32_t)k) {
        uint32_t run = above;
        for (int i = 0; i < PER; ++i) {
            const uint32_t h = sm->hist[threadIdx.x * PER + i];
            if (run + h >= (uint32_t)k) {
                sm->sel_bin = (uint32_t)(threadIdx.x * PER + i);
                sm->sel_above = run;
                break;
            }
            run += h;
        }
        sm->cand_count = 0;
    }
    consumer_sync();
    const uint32_t sel_bin = sm->sel_bin, need = (uint32_t)k - sm->sel_above;
    const uint32_t pop = sm->hist[sel_bin];
    if (pop > (uint32_t)BM25_CAND_MAX) return false;  // uniform
    for (int i = threadIdx.x; i < BM25_BLOCK_DOCS / 4; i += BM25_THREADS) {
        const float4 v = a4[i];
        const uint32_t b[4] = {__float_as_uint(v.x), __float_as_uint(v.y), __float_as_uint(v.z), __float_as_uint(v.w)};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            if (b[j] == 0) continue;
            const uint32_t bin = (mx - b[j]) >> NEAR_SHIFT;
            if (bin < sel_bin) {
                const uint32_t slot = atomicAdd(&sm->out_count, 1u);
                out[slot] = srag_pack_key(b[j], (uint32_t)(doc_base + 4 * i + j));
            } else if (bin == sel_bin) {
                sm->cand[atomicAdd(&sm->cand_count, 1u)] = (uint16_t)(4 * i + j);
            }
        }
    }
    consumer_sync();
    // the bin of the k-th value: take its `need` best members, equal scores by ascending doc id
    for (uint32_t e = threadIdx.x; e < pop; e += BM25_THREADS) {
        const uint32_t de = sm->cand[e], be = __float_as_uint(sm->acc[de]);
        uint32_t rank = 0;
        for (uint32_t f = 0; f < pop; ++f) {
            const uint32_t df = sm->cand[f], bf = __float_as_uint(sm->acc[df]);
            rank += (bf > be) || (bf == be && df < de);
        }
        if (rank < need) {
            const uint32_t slot = atomicAdd(&sm->out_count, 1u);
            out[slot] = srag_pack_key(be, (uint32_t)(doc_base + de));
        }
    }
    return true;
}

// One work item = (query, 16384-document block). Persistent CTAs walk the items of one phase:
//   phase 0  every `stride`-th block of each query, exact block top-k (select_exact);
//   phase 1  the remaining blocks, filtered by tau[q] = k-th best score over the phase-0 blocks of the
//            query (a lower bound of its global k-th score): documents crossing tau are recorded while
//            the terms are applied, so usually there is nothing to scan and nothing to emit; a block
//            with more than k entries >= tau falls back to select_exact.
//
// Warp 8 is the PRODUCER: it walks the CTA's items, looks up the posting slice of every query term inside
// the block (block_ptr) and streams the slices, in ascending-term order and cut into chunks of at most
// 1024 postings, through a ring of shared-memory slots with 1-D bulk copies (cp.async.bulk + mbarrier
// complete_tx). It runs as far ahead as the ring allows, across item boundaries, so HBM requests stay
// in flight while the consumers zero, accumulate and select.
// Warps 0..7 are the CONSUMERS: they follow the slot stream; the slot's metadata says where an item or a
// term starts. A barrier separates terms (two terms may hit the same document; inside a term every
// posting is a distinct document), which keeps the per-document summation order = ascending term id.
struct Bm25Filter {  // metadata pre-filter, same rule as K1 (srag_topk_search_filtered); null pointers = no constraint
    const uint32_t* doc_tags;    // [n_docs]
    const int32_t* doc_group;    // [n_docs]
    const uint32_t* q_tag_mask;  // [n_queries]
    const int32_t* q_group;      // [n_queries]
};

template <bool FILT>  // FILT: the unfiltered instantiation carries none of the filter code
__global__ void __launch_bounds__(BM25_CTA_THREADS, BM25_CTAS_PER_SM)
bm25_items_kernel(const uint32_t* __restrict__ slice_ptr /* [vocab][n_sub + 1], groups of 8 postings */,
                  const uint16_t* __restrict__ post_doc,
                  const float* __restrict__ post_imp, int n_blocks, int n_sub, const int32_t* __restrict__ q_terms,
                  int max_terms, int k, int q_offset, int nq, int phase, int stride,
                  const float* __restrict__ tau /* [nq][k] sorted sample scores, phase 1 */,
                  uint64_t* __restrict__ partial /* phase 1: [nq][n_blocks][k]; phase 0: [nq][ceil(n_blocks/stride)][k] */,
                  int debug /* SRAG_BM25_DEBUG */, Bm25Filter flt, int64_t n_docs,
                  const uint32_t* __restrict__ q_flag /* non-null: only the flagged queries (fallback of bm25_sub_kernel) */,
                  const uint32_t* __restrict__ n_flagged) {
    extern __shared__ uint8_t bm25_smem_raw[];
    Bm25Smem* sm = reinterpret_cast<Bm25Smem*>(bm25_smem_raw);
    if (q_flag && *n_flagged == 0) return;  // nothing overflowed: the usual case
    const int nb0 = (n_blocks + stride - 1) / stride;
    const int nbx = phase == 0 ? nb0 : n_blocks - nb0;
    const long long n_items = (long long)nq * nbx;
    const int lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        for (int s = 0; s < BM25_SLOTS; ++s) {
            ptx::mbar_init(&sm->full[s], 1);
            ptx::mbar_init(&sm->empty[s], BM25_THREADS / 32);
        }
        ptx::fence_mbar_init();
        sm->prod.slot = 0; sm->prod.pos = 0; sm->prod.inflight = 0; sm->prod.free_post = BM25_RING;
        sm->prod.tail_slot = 0; sm->prod.tail_par = 0; sm->prod.turn = 0;
    }
    __syncthreads();

    if (threadIdx.x >= BM25_THREADS) {
        // ------------------------------ producer warp ------------------------------------------
        // Items of this CTA: blockIdx.x, + gridDim.x, ... as (query, j-th block of the phase); stepping by
        // gridDim.x is an add with carry, so no 64-bit division per item.
        struct ItemPos { int ql, j; };
        const int step_q = (int)(gridDim.x / (unsigned)nbx), step_j = (int)(gridDim.x % (unsigned)nbx);
        auto advance = [&](ItemPos p) {
            p.j += step_j;
            p.ql += step_q;
            if (p.j >= nbx) { p.j -= nbx; ++p.ql; }
            return p;
        };
        auto block_of = [&](int j_) { return phase == 0 ? j_ * stride : j_ + j_ / (stride - 1) + 1; };
        // lane l looks after term slot j0 + l of the query
        auto term_at = [&](ItemPos p, int j0, float& tq) -> int32_t {
            tq = 0.f;
            if (p.ql >= nq) return -1;
            if (q_flag && q_flag[p.ql] == 0) return -1;  // an item without terms costs nothing downstream
            if (phase == 1) tq = __ldg(tau + (size_t)p.ql * k + (k - 1));
            return j0 + lane < max_terms ? __ldg(q_terms + (size_t)(q_offset + p.ql) * max_terms + j0 + lane) : -1;
        };
        auto ptrs_at = [&](ItemPos p, int32_t t, int64_t& lo, int64_t& hi) {
            lo = 0;
            hi = 0;
            if (p.ql < nq && t >= 0) {
                const int s0 = block_of(p.j) * BM25_SUBS;  // the tile = sub-blocks s0 .. s0 + BM25_SUBS - 1
                const uint32_t* sp = slice_ptr + (size_t)t * (n_sub + 1);
                lo = (int64_t)__ldg(sp + s0) << BM25_GROUP_LOG;
                hi = (int64_t)__ldg(sp + (s0 + BM25_SUBS < n_sub ? s0 + BM25_SUBS : n_sub)) << BM25_GROUP_LOG;
            }
        };
        // The two producer warps take ALTERNATE items of the CTA's list: planning an item (slice lookup, scans,
        // chunk table) is most of a producer's work and runs in parallel on both; the issue of its chunks into
        // the shared rings is serialised by a turn counter so that the stream stays in item order. Ring state
        // (head / tail of the descriptor ring, posting ring position and free space) lives in shared memory and
        // is handed over with the turn.
        const uint32_t pw = (threadIdx.x - BM25_THREADS) >> 5;  // 0 or 1
        ProducerState& ps = sm->prod;
        uint32_t slot = 0, tail_slot = 0, tail_par = 0, inflight = 0, pos = 0, free_post = BM25_RING;
        static_assert(BM25_SLOTS == 32, "descriptor slots are addressed with 5-bit arithmetic");
        static_assert(BM25_BLOCK_DOCS * 4 <= 65536, "byte offsets inside a block's tile must fit 16 bits");
        auto acquire_turn = [&](uint32_t seq) {
            uint32_t spins = 0;
            while (*reinterpret_cast<volatile uint32_t*>(&ps.turn) != seq) {
                __nanosleep(32);
                if (++spins > (1u << 26)) __trap();
            }
            __threadfence_block();  // the other warp's state stores are visible
            slot = ps.slot; pos = ps.pos; inflight = ps.inflight; free_post = ps.free_post;
            tail_slot = ps.tail_slot; tail_par = ps.tail_par;
        };
        auto release_turn = [&](uint32_t seq) {
            __syncwarp();
            if (lane == 0) {
                ps.slot = slot; ps.pos = pos; ps.inflight = inflight; ps.free_post = free_post;
                ps.tail_slot = tail_slot; ps.tail_par = tail_par;
                __threadfence_block();
                *reinterpret_cast<volatile uint32_t*>(&ps.turn) = seq + 1;
            }
        };
        auto reserve = [&](uint32_t n_slots, uint32_t n_post) {  // wait for descriptors and ring space
            while (inflight + n_slots > (uint32_t)BM25_SLOTS || free_post < n_post) {
                ptx::mbar_wait_lean(&sm->empty[tail_slot], tail_par);
                free_post += ps.slot_post[tail_slot];
                tail_slot = (tail_slot + 1) & (BM25_SLOTS - 1);
                tail_par ^= tail_slot == 0;
                --inflight;
            }
        };
        constexpr uint32_t FULL = 0xFFFFFFFFu;
        auto warp_incl_scan = [&](uint32_t v) {
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t t = __shfl_up_sync(FULL, v, o);
                if (lane >= o) v += t;
            }
            return v;
        };
        auto advance2 = [&](ItemPos p) { return advance(advance(p)); };
        // two-deep lookahead (in this warp's own item sequence) so that neither dependent load is waited for
        ItemPos p0{(int)(blockIdx.x / (unsigned)nbx), (int)(blockIdx.x % (unsigned)nbx)};
        if (pw) p0 = advance(p0);
        ItemPos p1 = advance2(p0), p2 = advance2(p1), p3 = advance2(p2);
        float tau0, tau1, tau2;
        int64_t lo, hi, nlo, nhi;
        int32_t t0 = term_at(p0, 0, tau0);
        int32_t t1 = term_at(p1, 0, tau1);
        int32_t t2 = term_at(p2, 0, tau2);
        ptrs_at(p0, t0, lo, hi);
        ptrs_at(p1, t1, nlo, nhi);
        uint32_t seq = pw;  // position of the current item in the CTA's list
        for (; p0.ql < nq; p0 = p1, p1 = p2, p2 = p3, p3 = advance2(p3), seq += 2) {
            const int ql = p0.ql, blk = block_of(p0.j);
            bool first_of_item = true, have_turn = false;
            uint32_t last_hr = 0;  // valid postings of the previous chunk of this item
            for (int j0 = 0; j0 < max_terms; j0 += 32) {
                if (j0 > 0) {  // queries with more than 32 terms: fetched in line
                    float unused;
                    const int32_t t = term_at(p0, j0, unused);
                    ptrs_at(p0, t, lo, hi);
                }
                const int nt = max_terms - j0 < 32 ? max_terms - j0 : 32;
                // Lane u plans term u: the slice rounded out to groups of 8 postings (16 bytes of 16-bit offsets: the unit
                // of a bulk copy; the arrays are padded) and cut into chunks. A term has at most BM25_BLOCK_DOCS postings
                // inside a block.
                const int64_t a = lo & ~(int64_t)7;
                const uint32_t hi_off = (uint32_t)(hi - a), lo_off = (uint32_t)(lo - a);
                const uint32_t len = lo == hi ? 0u : (hi_off + 7u) & ~7u;
                const uint32_t cnt = (len + BM25_CHUNK - 1) / BM25_CHUNK;
                const uint32_t incl = warp_incl_scan(cnt);
                const uint32_t total = __shfl_sync(FULL, incl, 31);
                // Lane i then issues chunk base + i of the item's stream (term-major order)
                for (uint32_t base = 0; base < total; base += 32) {
                    const uint32_t i = base + lane;
                    const bool valid = i < total;
                    uint32_t u = 0;
                    for (int v = 0; v < nt; ++v) u += __shfl_sync(FULL, incl, v) <= i;
                    if (!valid) u = 0;
                    const uint32_t j = i - __shfl_sync(FULL, incl - cnt, u);  // chunk inside its term
                    const uint32_t off = j * BM25_CHUNK;
                    const int64_t a_u = __shfl_sync(FULL, a, u);
                    const uint32_t len_u = __shfl_sync(FULL, len, u), hi_u = __shfl_sync(FULL, hi_off, u);
                    const uint32_t lo_u = __shfl_sync(FULL, lo_off, u);
                    const uint32_t n = !valid ? 0u : (len_u - off < (uint32_t)BM25_CHUNK ? len_u - off : (uint32_t)BM25_CHUNK);
                    const uint32_t cum = warp_incl_scan(n);
                    const uint32_t in_round = total - base < 32u ? total - base : 32u;
                    // A term boundary needs a CTA barrier only if either side of it involves more than the first
                    // consumer warp (threads 0..31 own postings 0..127 of a chunk): a run of short slices is
                    // applied by that one warp in program order.
                    const uint32_t hr = !valid ? 0u : (hi_u - off < (uint32_t)BM25_CHUNK ? hi_u - off : (uint32_t)BM25_CHUNK);
                    uint32_t prev_hr = __shfl_up_sync(FULL, hr, 1);
                    if (lane == 0) prev_hr = last_hr;
                    last_hr = __shfl_sync(FULL, hr, in_round - 1);
                    const bool term_barrier = j == 0 && (hr > 128u || prev_hr > 128u);
                    // the chunk's descriptor, complete but for its ring position
                    SlotMeta md;
                    md.ql = ql;
                    md.blk = blk;
                    md.lo_rel = j == 0 ? (int32_t)lo_u : 0;
                    md.hi_rel = (int32_t)hr;
                    md.flags = (first_of_item && i == 0) ? SLOT_NEW_ITEM : (term_barrier ? SLOT_NEW_TERM : 0u);
                    md.tau = tau0;
                    md.n = n;
                    const uint16_t* src_d = post_doc + a_u + off;
                    const float* src_w = post_imp + a_u + off;
                    if (!have_turn) {  // everything above ran while the other warp was issuing its item
                        acquire_turn(seq);
                        have_turn = true;
                    }
                    uint32_t done = 0, cum_base = 0;
                    while (done < in_round) {
                        // as many chunks as fit half the ring at once (a chunk is at most a ninth of it)
                        const uint32_t fits = __ballot_sync(FULL, valid && lane >= done && cum - cum_base <= (uint32_t)BM25_RING / 2);
                        const uint32_t take = __popc(fits);
                        const uint32_t need = __shfl_sync(FULL, cum, done + take - 1) - cum_base;
                        reserve(take, need);
                        if (lane >= done && lane < done + take) {
                            const uint32_t my_slot = (slot + (lane - done)) & (BM25_SLOTS - 1);
                            uint32_t my_pos = pos + (cum - n - cum_base);
                            if (my_pos >= (uint32_t)BM25_RING) my_pos -= BM25_RING;
                            uint64_t* bar = &sm->full[my_slot];
                            md.pos = my_pos;
                            sm->meta[my_slot] = md;
                            ps.slot_post[my_slot] = n;
                            ptx::mbar_arrive_expect_tx(bar, n * 6u);
                            const uint32_t n1 = my_pos + n <= (uint32_t)BM25_RING ? n : BM25_RING - my_pos;  // up to the ring's end
                            ptx::bulk_load_1d(&sm->ring_doc[my_pos], src_d, n1 * 2u, bar);
                            ptx::bulk_load_1d(&sm->ring_imp[my_pos], src_w, n1 * 4u, bar);
                            if (n1 < n) {  // wrapped remainder
                                ptx::bulk_load_1d(&sm->ring_doc[0], src_d + n1, (n - n1) * 2u, bar);
                                ptx::bulk_load_1d(&sm->ring_imp[0], src_w + n1, (n - n1) * 4u, bar);
                            }
                        }
                        __syncwarp();
                        slot = (slot + take) & (BM25_SLOTS - 1);
                        pos += need;
                        if (pos >= (uint32_t)BM25_RING) pos -= BM25_RING;
                        free_post -= need;
                        inflight += take;
                        done += take;
                        cum_base += need;
                    }
                }
                if (total > 0) first_of_item = false;
            }
            if (!have_turn) acquire_turn(seq);  // an item without postings still passes the turn on
            release_turn(seq);
            // rotate the lookahead
            lo = nlo;
            hi = nhi;
            tau0 = tau1;
            tau1 = tau2;
            ptrs_at(p2, t2, nlo, nhi);
            t2 = term_at(p3, 0, tau2);
        }
        // the warp whose turn follows the CTA's last item closes the stream
        const uint32_t n_cta = (uint32_t)((n_items - (long long)blockIdx.x + gridDim.x - 1) / gridDim.x);
        if (seq == n_cta) {
            acquire_turn(seq);
            reserve(1, 0);
            if (lane == 0) {
                sm->meta[slot].flags = SLOT_END;
                sm->meta[slot].n = 0;
                ptx::mbar_arrive(&sm->full[slot]);
            }
        }
        return;
    }

    // ---------------------------------- consumer warps -----------------------------------------
    float4* a4 = reinterpret_cast<float4*>(sm->acc);
    uint32_t slot = 0, par = 0;
    bool have_item = false;
    int cur_ql = 0, cur_blk = 0;
    float cur_tau = 0.f;
    uint32_t cross_bits = 0xFFFFFFFFu;  // tau as orderable bits when candidates are tracked, else never crossed
    uint32_t f_mask = 0xFFFFFFFFu;      // this item's query filter (FILT only)
    int32_t f_group = -1;
    auto eligible = [&](int32_t doc) -> bool {  // doc = document index inside this shard
        if (flt.doc_tags && (__ldg(flt.doc_tags + doc) & f_mask) == 0) return false;
        if (f_group >= 0 && __ldg(flt.doc_group + doc) != f_group) return false;
        return true;
    };

    // top-k of the finished tile -> partial[cur_ql][cur_blk][*] (all consumers; adds are complete)
    auto select_item = [&]() {
        const int32_t doc_base = cur_blk * BM25_BLOCK_DOCS;
        uint64_t* out = phase == 0 ? partial + ((size_t)cur_ql * nb0 + cur_blk / stride) * k
                                   : partial + ((size_t)cur_ql * n_blocks + cur_blk) * k;
        if (cross_bits != 0xFFFFFFFFu) {
            const uint32_t nc = sm->cand_count;
            if (nc <= (uint32_t)k && nc <= (uint32_t)BM25_CAND_MAX) {
                if (threadIdx.x < nc) {
                    const uint32_t r = sm->cand[threadIdx.x];
                    out[threadIdx.x] = srag_pack_key(__float_as_uint(sm->acc[r]), (uint32_t)(doc_base + r));
                }
                return;
            }
        }
        if constexpr (FILT) {
            // scan paths select over the whole tile: drop the matches the query's filter does not admit first
            for (int i = threadIdx.x; i < BM25_BLOCK_DOCS / 4; i += BM25_THREADS) {
                float4 v = a4[i];
                if (v.x == 0.f && v.y == 0.f && v.z == 0.f && v.w == 0.f) continue;
                const int32_t d0 = doc_base + 4 * i;
                if (v.x != 0.f && !(d0 < n_docs && eligible(d0))) v.x = 0.f;
                if (v.y != 0.f && !(d0 + 1 < n_docs && eligible(d0 + 1))) v.y = 0.f;
                if (v.z != 0.f && !(d0 + 2 < n_docs && eligible(d0 + 2))) v.z = 0.f;
                if (v.w != 0.f && !(d0 + 3 < n_docs && eligible(d0 + 3))) v.w = 0.f;
                a4[i] = v;
            }
            consumer_sync();
        }
        // impacts are > 0, so "matched" <=> accumulator bits != 0, and positive floats order as uints
        uint32_t lo_bits = 1u;
        if (cur_tau > 0.f) lo_bits = __float_as_uint(cur_tau);  // ties with the sample's k-th best are kept
        uint32_t cnt = 0;
        for (int i = threadIdx.x; i < BM25_BLOCK_DOCS / 4; i += BM25_THREADS) {
            const float4 v = a4[i];
            cnt += (__float_as_uint(v.x) >= lo_bits) + (__float_as_uint(v.y) >= lo_bits) +
                   (__float_as_uint(v.z) >= lo_bits) + (__float_as_uint(v.w) >= lo_bits);
        }
        uint32_t total;
        (void)block_excl_scan(cnt, sm->warp_tmp, &total);
        if (total == 0) return;  // `partial` is pre-zeroed (= empty keys)
        if (total <= (uint32_t)k) {
            for (int i = threadIdx.x; i < BM25_BLOCK_DOCS / 4; i += BM25_THREADS) {
                const float4 v = a4[i];
                const uint32_t b[4] = {__float_as_uint(v.x), __float_as_uint(v.y), __float_as_uint(v.z), __float_as_uint(v.w)};
#pragma unroll
                for (int e = 0; e < 4; ++e)
                    if (b[e] >= lo_bits) {
                        const uint32_t s_ = atomicAdd(&sm->out_count, 1u);
                        out[s_] = srag_pack_key(b[e], (uint32_t)(doc_base + 4 * i + e));
                    }
            }
            return;
        }
        // more than k survivors (always so in phase 0 for a dense block): exact top-k of ALL matches
        if (lo_bits > 1u) {
            uint32_t nz = 0;
            for (int i = threadIdx.x; i < BM25_BLOCK_DOCS / 4; i += BM25_THREADS) {
                const float4 v = a4[i];
                nz += (v.x != 0.f) + (v.y != 0.f) + (v.z != 0.f) + (v.w != 0.f);
            }
            (void)block_excl_scan(nz, sm->warp_tmp, &total);
        }
        if ((debug & 64) || !select_near_max(sm, k, doc_base, out)) select_exact(sm, k, doc_base, total, out);
    };

    for (;;) {
        ptx::mbar_wait(&sm->full[slot], par);
        const SlotMeta m = sm->meta[slot];
        if (m.flags & (SLOT_NEW_ITEM | SLOT_END)) {
            consumer_sync();  // every add of the previous item has landed
            // A tracked item in which nothing crossed tau (the usual case in phase 1) has nothing to select and
            // nothing to read back from the tile: go straight to zeroing. cand_count is final after the barrier
            // above, and the only write that can race with this read is thread 0 storing the same 0 below.
            const bool nothing = cross_bits != 0xFFFFFFFFu && sm->cand_count == 0;
            if (have_item && !nothing && !(debug & 2)) {
                select_item();
                consumer_sync();  // selection has read the tile
            }
            if (m.flags & SLOT_END) break;
            if (!(debug & 4))
                for (int i = threadIdx.x; i < BM25_BLOCK_DOCS / 4; i += BM25_THREADS) a4[i] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (threadIdx.x == 0) { sm->out_count = 0; sm->cand_count = 0; }
            have_item = true;
            cur_ql = m.ql;
            cur_blk = m.blk;
            cur_tau = m.tau;
            cross_bits = (cur_tau > 0.f && !(debug & 16)) ? __float_as_uint(cur_tau) : 0xFFFFFFFFu;
            if constexpr (FILT) {
                f_mask = flt.doc_tags ? __ldg(flt.q_tag_mask + q_offset + cur_ql) : 0xFFFFFFFFu;
                f_group = flt.doc_group ? __ldg(flt.q_group + q_offset + cur_ql) : -1;
            }
            consumer_sync();
        } else if ((m.flags & SLOT_NEW_TERM) && !(debug & 8)) {
            consumer_sync();
        }
        const int rel = 4 * (int)threadIdx.x;
        uint2 dd = make_uint2(0u, 0u);   // four 16-bit byte offsets into acc[]
        float4 ww = make_float4(0.f, 0.f, 0.f, 0.f);
        const bool active = rel < m.hi_rel && rel + 4 > m.lo_rel;
        if (active) {
            uint32_t idx = m.pos + (uint32_t)rel;
            if (idx >= (uint32_t)BM25_RING) idx -= BM25_RING;
            dd = *reinterpret_cast<const uint2*>(&sm->ring_doc[idx]);
            ww = *reinterpret_cast<const float4*>(&sm->ring_imp[idx]);
        }
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(&sm->empty[slot]);  // the slot's data is in registers: hand it back
        if (active && !(debug & 1)) {
            const int32_t doc_base = cur_blk * BM25_BLOCK_DOCS;
            // (slices are padded to whole groups of 8 with zero-impact postings whose offsets only mean something to
            // bm25_sub_kernel: skipped here — a real impact is > 0)
            const bool m0 = rel >= m.lo_rel && rel < m.hi_rel && ww.x != 0.f, m1 = rel + 1 >= m.lo_rel && rel + 1 < m.hi_rel && ww.y != 0.f;
            const bool m2 = rel + 2 >= m.lo_rel && rel + 2 < m.hi_rel && ww.z != 0.f, m3 = rel + 3 >= m.lo_rel && rel + 3 < m.hi_rel && ww.w != 0.f;
            const int r0 = m0 ? (int)((dd.x & 0xFFFFu) >> 2) : 0, r1 = m1 ? (int)(dd.x >> 18) : 0;
            const int r2 = m2 ? (int)((dd.y & 0xFFFFu) >> 2) : 0, r3 = m3 ? (int)(dd.y >> 18) : 0;
            // the four reads are issued together, then the adds, then the writes (one latency, not four)
            const float a0 = sm->acc[r0], a1 = sm->acc[r1], a2 = sm->acc[r2], a3 = sm->acc[r3];
            const float n0 = __fadd_rn(a0, ww.x), n1 = __fadd_rn(a1, ww.y), n2 = __fadd_rn(a2, ww.z), n3 = __fadd_rn(a3, ww.w);
            if (m0) sm->acc[r0] = n0;
            if (m1) sm->acc[r1] = n1;
            if (m2) sm->acc[r2] = n2;
            if (m3) sm->acc[r3] = n3;
            // Impacts are > 0, so an accumulator only grows and crosses tau at most once: the add that takes
            // it from below tau to >= tau records the document.
            const bool c0 = m0 && __float_as_uint(n0) >= cross_bits && __float_as_uint(a0) < cross_bits;
            const bool c1 = m1 && __float_as_uint(n1) >= cross_bits && __float_as_uint(a1) < cross_bits;
            const bool c2 = m2 && __float_as_uint(n2) >= cross_bits && __float_as_uint(a2) < cross_bits;
            const bool c3 = m3 && __float_as_uint(n3) >= cross_bits && __float_as_uint(a3) < cross_bits;
            if (c0 | c1 | c2 | c3) {
                const int rr[4] = {r0, r1, r2, r3};
                const bool cc[4] = {c0, c1, c2, c3};
#pragma unroll
                for (int e = 0; e < 4; ++e)
                    if (cc[e] && (!FILT || eligible(doc_base + rr[e]))) {
                        const uint32_t s_ = atomicAdd(&sm->cand_count, 1u);
                        if (s_ < (uint32_t)BM25_CAND_MAX) sm->cand[s_] = (uint16_t)rr[e];
                    }
            }
        }
        if (++slot == BM25_SLOTS) { slot = 0; par ^= 1; }
    }
}

// ---------------------------------------------------------------------------------------------------------
// bm25_sub_kernel: scoring at WARP granularity (sample pass and main pass). A work item is (query, sub-block of
// BM25_SUB_DOCS documents); the sub-block's fp32 accumulators are 8 KB of shared memory private to ONE warp, so every add
// to a document is made by the same warp: terms are applied in ascending order with nothing but __syncwarp() between
// terms — no CTA barrier, no producer warps. 20 independent warps per SM walk the items round-robin; neighbouring warps
// take neighbouring sub-blocks, i.e. neighbouring pieces of the same posting lists.
//   * lane u holds the bounds of query term u inside the sub-block (slice_ptr), fetched two items ahead (term ids three
//     ahead), so no dependent load is waited for;
//   * a round = up to 128 postings of one term: 8 bytes of offsets + 16 bytes of impacts per lane. Slices start on a
//     group of 8 postings and are padded to whole groups (padding adds +0.0 to a trash slot), so a round has no masks;
//     the rounds of all terms form one stream in ascending-term order. The stream is described 32 rounds at a time —
//     lane i works out where round i starts and how long it is by bisection in the running totals, ONCE per chunk — and
//     the round loop only broadcasts two words per round;
//   * the postings of a round travel global -> shared by cp.async (one group per round) SUB_DEPTH - 1 rounds ahead of the
//     adds, and cp.async.wait_group gives a true software pipeline (see SubWarpTile for why registers cannot);
//   * main pass: a document whose accumulator crosses tau is remembered (<= SUB_CAND per item); at the end of the item
//     its final score goes into the query's candidate region — the query's slice of `partial`, filled from the front
//     through one atomic counter per query. tau is the k-th best of the scores a 1/stride sample reported, so a query
//     collects about stride * k candidates over all its items;
//   * more crossings in one sub-block than SUB_CAND, or a full region, flags the QUERY; flagged queries are redone
//     by bm25_items_kernel (exact top-k of every block) after their region is cleared. Massive ties end up there.
// Per-document sums are formed exactly as in bm25_items_kernel: bit-exact with the oracle.
// History at C3 (ms per batch): 10.85 round 1 (8 B / posting, masked register-staged rounds, 8.35e9 warp instructions) ->
// 10.97 with 6-byte padded slices and chunked round descriptors (5.84e9 instructions, but every round still waited a full
// L2 round trip: one scoreboard for all LDGs) -> 9.87 with the cp.async pipeline; stop-word queries 71.9 -> 35.1 ms.
#ifndef SRAG_BM25_SUB_DEPTH
#define SRAG_BM25_SUB_DEPTH 4
#endif
constexpr int SUB_DEPTH = SRAG_BM25_SUB_DEPTH;  // stages of the round pipeline: SUB_DEPTH - 1 rounds in flight while one is applied
#ifndef SRAG_BM25_SUB_WARPS
#define SRAG_BM25_SUB_WARPS (SRAG_BM25_SUB_DEPTH <= 2 ? 23 : SRAG_BM25_SUB_DEPTH == 3 ? 21 : 20)
#endif
constexpr int SUB_WARPS = SRAG_BM25_SUB_WARPS;   // warps per CTA (one CTA per SM): as many as the shared memory takes
#ifndef SRAG_BM25_PF
#define SRAG_BM25_PF 0
#endif
// L2 prefetch of a later item's slices: 0 off, 1 the next item, 2 the item after next. Off since the cp.async pipeline
// covers the latency by itself: C3 9.87 ms without against 10.28 ms with it, the stop-word queries 35.1 ms against 52.1 ms
// (round 1's register-staged loads needed it: 12.97 ms without, 10.97 ms with).
constexpr int SUB_PF = SRAG_BM25_PF;
constexpr int SUB_CAND = 64;
#ifndef SRAG_BM25_STAGE_TMA
#define SRAG_BM25_STAGE_TMA 0
#endif
// How a round's postings travel global -> shared: 0 (default) = cp.async (LDGSTS) per lane with wait_group; 1 = two 1-D
// bulk copies (cp.async.bulk, TMA) issued by one lane and tracked by an mbarrier per stage. The kernel's busiest unit is
// the SM's LSU data pipe (ncu, C3: 70 % of its wavefront slots — per 128-posting round ~11.7 accumulator reads + writes at
// the bank-conflict floor max_b count(b), ~4.2 stage reads, ~6 LDGSTS writes, ~5.2 zeroing); bulk copies bypass it for
// the writes (Mem Busy 76 % -> 55 %), but 256 / 512-byte copies are too small for the TMA path and the one-lane issue
// block + try_wait loop add instructions: C3 12.3 ms against 9.86 ms with LDGSTS. Kept selectable for A/B runs.
constexpr bool SUB_STAGE_TMA = SRAG_BM25_STAGE_TMA != 0;
#ifndef SRAG_BM25_ZERO_TMA
#define SRAG_BM25_ZERO_TMA 1
#endif
// How a warp's 2048 accumulators are cleared between items: 0 = 16 STS.128 per lane (64 LSU wavefronts per item: 17 % of
// all the kernel's shared-memory wavefronts on C3, the pipe it is bound by); 1 = ONE 8 KB bulk copy (cp.async.bulk, the
// TMA / async proxy — no LSU wavefronts) from a zero page that stays in L2, issued when the item's candidates have been
// read out and awaited behind the first fetches of the next item, so its latency hides under theirs.
constexpr bool SUB_ZERO_TMA = SRAG_BM25_ZERO_TMA != 0;
// 128 windows of 8 KB, one per warp (mod 128): 3000 warps copying the SAME 64 lines would queue on a few L2 slices
constexpr int SUB_ZERO_WINDOWS = 128;
__device__ __align__(128) float g_bm25_zero_page[SUB_ZERO_WINDOWS * BM25_SUB_DOCS];   // never written

// One warp's accumulators with a trash slot on either side, and the staging buffers of its round pipeline.
// Slices are padded to whole groups of 8 postings; a padding posting carries impact +0.0 and an offset that lands on a
// trash slot of whichever warp processes it:
//   sub-block 0 of a block: offset 8192 = the float right AFTER the warp's 2048 accumulators (acc_b = acc)
//   sub-block m > 0:        offset m * 8192 - 4 = the float right BEFORE them (acc_b = acc - m * 8192 bytes)
// so padding needs no mask anywhere in the round loop (sub_trash_offset below is what the index build writes).
// Staging: a round = up to 128 postings = 8 bytes of offsets + 16 bytes of impacts per lane, copied global -> shared by
// cp.async (LDGSTS) SUB_DEPTH - 1 rounds before they are applied. The copies of a round are one cp.async GROUP, and
// cp.async.wait_group waits for "all but the N newest groups": a true software pipeline. (Register-staged loads cannot
// be pipelined here: ptxas puts the LDGs of every stage on ONE scoreboard, so the first use of round r's data also waits
// for the loads of rounds r + 1, r + 2 issued after it — measured: depth 2, 3 and 4 all ran at the speed of depth 1.)
struct SubWarpTile {
    float before[4];
    float acc[BM25_SUB_DOCS];
    float after[4];
    alignas(16) float4 stage_imp[SUB_DEPTH][32];
    alignas(16) uint2 stage_off[SUB_DEPTH][32];
    alignas(8) uint64_t stage_bar[SUB_DEPTH];   // TMA staging: one mbarrier per stage (count 1 + transaction bytes)
    alignas(8) uint64_t zero_bar;               // bulk-copy zeroing of acc
};
__host__ __device__ constexpr uint32_t sub_trash_offset(int sub_in_block) {
    return sub_in_block == 0 ? (uint32_t)BM25_SUB_DOCS * 4u : (uint32_t)sub_in_block * (BM25_SUB_DOCS * 4u) - 4u;
}
static_assert(sub_trash_offset(BM25_SUBS - 1) < 65536u, "trash offsets must fit 16 bits");

struct SubSmem {
    SubWarpTile tile[SUB_WARPS];
    uint16_t cand[SUB_WARPS][SUB_CAND];
    uint32_t cand_count[SUB_WARPS];
};

__device__ __forceinline__ void cp_async_8(void* smem_dst, const void* gsrc) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(ptx::smem_u32(smem_dst)), "l"(gsrc) : "memory");
}
// (.cg = L1 bypass. ncu counts 6.7 shared-memory wavefronts per LDGSTS.E.BYPASS.128 against 2.6 for the LDS.128 that reads
// the same bytes back — 27 % of the kernel's shared-memory wavefronts; the .ca form through L1 has the ideal count but
// thrashes the 29 KB of L1 left beside 222 KB of shared memory: C3 9.50 -> 10.01 ms, stop-word queries 36.8 -> 40.7 ms.)
__device__ __forceinline__ void cp_async_16(void* smem_dst, const void* gsrc) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(ptx::smem_u32(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// round descriptor word: bits 0..7 = postings of the round (4 .. 128, a multiple of 4: lane l holds postings 4 l .. 4 l + 3
// when 4 l < count), bit 17 = first round of a term (a __syncwarp() goes before it)
constexpr uint32_t STAGE_NEW_TERM = 1u << 17;

// A document crossed tau: remember its tile offset (rare: ~32 k crossings per query over all its sub-blocks). Out of
// line, so that the round loop stays small.
__device__ __noinline__ void sub_record(uint32_t* cand_count, uint16_t* cand, uint32_t off_bytes) {
    const uint32_t c = atomicAdd(cand_count, 1u);
    if (c < (uint32_t)SUB_CAND) cand[c] = (uint16_t)off_bytes;
}
__device__ __noinline__ void sub_record_filtered(uint32_t* cand_count, uint16_t* cand, uint32_t off_bytes, int32_t blk_doc0,
                                                 const uint32_t* doc_tags, const int32_t* doc_group, uint32_t f_mask,
                                                 int32_t f_group) {
    const int32_t doc = blk_doc0 + (int32_t)(off_bytes >> 2);
    if (doc_tags && (__ldg(doc_tags + doc) & f_mask) == 0) return;
    if (f_group >= 0 && __ldg(doc_group + doc) != f_group) return;
    const uint32_t c = atomicAdd(cand_count, 1u);
    if (c < (uint32_t)SUB_CAND) cand[c] = (uint16_t)off_bytes;
}

// SAMPLE = the pass that yields tau (nothing tracked while terms are applied, a scan at the end reports true scores);
// otherwise the MAIN pass (documents crossing tau are recorded by the add that takes them across).
template <bool FILT, bool SAMPLE>
__global__ void __launch_bounds__(32 * SUB_WARPS, 1)
bm25_sub_kernel(const uint32_t* __restrict__ slice_ptr /* [vocab][n_sub + 1], in groups of 8 postings */,
                const uint16_t* __restrict__ post_off, const float* __restrict__ post_imp, int n_blocks, int n_sub,
                const int32_t* __restrict__ q_terms, int max_terms, int k, int q_offset, int nq, int stride,
                const float* __restrict__ tau /* [nq][k], main pass */,
                uint64_t* __restrict__ partial /* [nq][n_blocks * k]; sample pass: [nq][nb0 * BM25_SUBS * 32] */,
                uint32_t* __restrict__ q_count /* [nq], zeroed */, uint32_t* __restrict__ q_flag /* [nq], zeroed */,
                uint32_t* __restrict__ n_flagged, Bm25Filter flt,
                int sample_group /* sample pass: 1 or 4 lanes share one report (32 or 8 reports per sub-block) */) {
    extern __shared__ uint8_t bm25_smem_raw[];
    SubSmem* sm = reinterpret_cast<SubSmem*>(bm25_smem_raw);
    constexpr uint32_t FULL = 0xFFFFFFFFu;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    SubWarpTile* tile = &sm->tile[w];
    float* acc = tile->acc;
    uint16_t* cand = sm->cand[w];
    uint32_t* cand_count = &sm->cand_count[w];
    const int nb0 = (n_blocks + stride - 1) / stride;
    const int per_q = (SAMPLE ? nb0 : n_blocks) * BM25_SUBS;   // sub-blocks of one query in this pass
    const long long step = (long long)gridDim.x * SUB_WARPS;
    const uint32_t cap = (uint32_t)n_blocks * (uint32_t)k;     // key slots of one query in `partial`
    // the next item's slices are pulled into L2 by lpt lanes per term (1, 2, 4 ... 32): the longest slice sets the trip count
    int lpt_log = 0;
    while (lpt_log < 5 && (max_terms << (lpt_log + 1)) <= 32) ++lpt_log;
    const int pf_term = lane >> lpt_log, pf_j = lane & ((1 << lpt_log) - 1);
    if (lane == 0) {  // the trash slots only ever receive +0.0
        sm->tile[w].before[3] = 0.f;
        sm->tile[w].after[0] = 0.f;
        if constexpr (SUB_STAGE_TMA) {
#pragma unroll
            for (int i = 0; i < SUB_DEPTH; ++i) ptx::mbar_init(&sm->tile[w].stage_bar[i], 1);
            ptx::fence_mbar_init();
        }
        if constexpr (SUB_ZERO_TMA) {
            ptx::mbar_init(&sm->tile[w].zero_bar, 1);
            ptx::fence_mbar_init();
        }
    }
    __syncwarp();
    uint32_t stage_phase = 0;   // TMA staging: bit i = parity the next wait on stage i expects
    // bulk-copy zeroing: acc_clean = the accumulators are all zero (or a copy that makes them so is in flight: zero_wait)
    bool acc_clean = false, zero_wait = false;
    uint32_t zero_phase = 0;
    const float* zero_src = g_bm25_zero_page + (size_t)((blockIdx.x * SUB_WARPS + w) % SUB_ZERO_WINDOWS) * BM25_SUB_DOCS;
    auto zero_issue = [&]() {   // every lane is done with acc (reads and writes): order them before the async-proxy write
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
        if (lane == 0) {
            ptx::mbar_arrive_expect_tx(&tile->zero_bar, BM25_SUB_DOCS * 4u);
            ptx::bulk_load_1d(acc, zero_src, BM25_SUB_DOCS * 4u, &tile->zero_bar);
        }
        acc_clean = true;
        zero_wait = true;
    };
    auto zero_await = [&]() {
        if (zero_wait) {
            ptx::mbar_wait_lean(&tile->zero_bar, zero_phase);
            zero_phase ^= 1u;
            zero_wait = false;
        }
    };

    // Items of this warp: first, first + step, ... as (query, j-th sub-block of the pass); stepping is an add with carry.
    struct Item { int q, j; };
    const int step_q = (int)(step / per_q), step_j = (int)(step % per_q);
    auto advance = [&](Item x) {
        x.j += step_j;
        x.q += step_q;
        if (x.j >= per_q) { x.j -= per_q; ++x.q; }
        return x;
    };
    auto sub_of = [&](Item x) -> int {  // sub-block index, -1 past the end of the queries or of the documents
        if (x.q >= nq) return -1;
        const int sub = SAMPLE ? (x.j / BM25_SUBS) * stride * BM25_SUBS + x.j % BM25_SUBS : x.j;
        return sub < n_sub ? sub : -1;
    };
    auto term_of = [&](Item x, int sub, int j0) -> int32_t {
        if (sub < 0 || j0 + lane >= max_terms) return -1;
        return __ldg(q_terms + (size_t)(q_offset + x.q) * max_terms + j0 + lane);
    };
    // first and one-past-last group of 8 postings of the slice; kept RAW (the length is only formed where it is used, an
    // item later: a subtraction right behind the loads would wait for them)
    auto bounds_of = [&](int sub, int32_t t, uint32_t& g_lo, uint32_t& g_hi) {
        g_lo = 0;
        g_hi = 0;
        if (sub >= 0 && t >= 0) {
            const uint32_t* sp = slice_ptr + (size_t)t * (n_sub + 1) + sub;
            g_lo = __ldg(sp);
            g_hi = __ldg(sp + 1);
        }
    };
    auto flag_query = [&](int q) {
        if (lane == 0 && atomicExch(&q_flag[q], 1u) == 0u) atomicAdd(n_flagged, 1u);
    };

    // Lookahead: slice bounds one and two items ahead (the second so that the item after next can be pulled into L2 a
    // whole item early), term ids three ahead, so that no dependent load is ever waited for.
    const long long first = (long long)blockIdx.x * SUB_WARPS + w;
    Item x0{(int)(first / per_q), (int)(first % per_q)};
    Item x1 = advance(x0), x2 = advance(x1), x3 = advance(x2);
    int s0 = sub_of(x0), s1 = sub_of(x1), s2 = sub_of(x2), s3 = sub_of(x3);
    int32_t t3 = term_of(x3, s3, 0);
    uint32_t g0, g1, g2;   // slice starts, in groups of 8 postings
    uint32_t h0, h1, h2;   // slice ends (exclusive), in groups
    bounds_of(s0, term_of(x0, s0, 0), g0, h0);
    bounds_of(s1, term_of(x1, s1, 0), g1, h1);
    bounds_of(s2, term_of(x2, s2, 0), g2, h2);
    float tau0 = s0 >= 0 && !SAMPLE ? __ldg(tau + (size_t)x0.q * k + (k - 1)) : 0.f;
    float tau1 = s1 >= 0 && !SAMPLE ? __ldg(tau + (size_t)x1.q * k + (k - 1)) : 0.f;

    for (; x0.q < nq;) {
        if (s0 >= 0) {
            const int q = x0.q;
            // post_off holds byte offsets inside the 16384-document block's tile; this warp's 2048 accumulators are the
            // piece of that tile that starts at byte sub_byte0, so acc_b + offset is the accumulator's address
            const uint32_t sub_byte0 = (uint32_t)(s0 & (BM25_SUBS - 1)) * (BM25_SUB_DOCS * 4u);
            char* acc_b = reinterpret_cast<char*>(acc) - sub_byte0;
            const int32_t blk_doc0 = (s0 / BM25_SUBS) * BM25_BLOCK_DOCS;   // document of byte offset 0
            // tau = 0 (the sample matched fewer than k documents): every matched document is a candidate
            const uint32_t cross_bits = SAMPLE ? 0xFFFFFFFFu : tau0 > 0.f ? __float_as_uint(tau0) : 1u;
            uint32_t f_mask = 0xFFFFFFFFu;
            int32_t f_group = -1;
            if constexpr (FILT) {
                f_mask = flt.doc_tags ? __ldg(flt.q_tag_mask + q_offset + q) : 0xFFFFFFFFu;
                f_group = flt.doc_group ? __ldg(flt.q_group + q_offset + q) : -1;
            }
            if constexpr (SUB_PF > 0) {
                // Pull the NEXT item's slices into L2 now (one 128-byte line per request): its loads then see L2 latency
                // instead of HBM latency, and nothing has to be held in registers meanwhile. Lane group pf_term looks
                // after term pf_term, its lanes take every (1 << lpt_log)-th line.
                const uint32_t g_p = __shfl_sync(FULL, SUB_PF == 2 ? g2 : g1, pf_term);
                const uint32_t n_p = (__shfl_sync(FULL, SUB_PF == 2 ? h2 : h1, pf_term) - g_p) << BM25_GROUP_LOG;
                if (n_p > 0) {
                    const char* po = reinterpret_cast<const char*>(post_off) + (size_t)g_p * (BM25_GROUP * 2u);
                    const char* pi = reinterpret_cast<const char*>(post_imp) + (size_t)g_p * (BM25_GROUP * 4u);
                    const int bytes_o = (int)n_p * 2 + (int)(reinterpret_cast<uintptr_t>(po) & 127);
                    const int bytes_i = (int)n_p * 4 + (int)(reinterpret_cast<uintptr_t>(pi) & 127);
                    po -= reinterpret_cast<uintptr_t>(po) & 127;
                    pi -= reinterpret_cast<uintptr_t>(pi) & 127;
                    const int stride_b = 128 << lpt_log;
                    for (int o = pf_j * 128; o < bytes_i; o += stride_b) {
                        asm volatile("prefetch.global.L2 [%0];" ::"l"(pi + o));
                        if (o < bytes_o) asm volatile("prefetch.global.L2 [%0];" ::"l"(po + o));
                    }
                }
            }
            bool zeroed = false;
            auto record = [&](uint32_t off_bytes) {
                if constexpr (FILT) sub_record_filtered(cand_count, cand, off_bytes, blk_doc0, flt.doc_tags, flt.doc_group, f_mask, f_group);
                else sub_record(cand_count, cand, off_bytes);
            };
            for (int j0 = 0; j0 < max_terms; j0 += 32) {
                if (j0 > 0) bounds_of(s0, term_of(x0, s0, j0), g0, h0);  // queries with more than 32 terms: fetched in line
                // lane u = term j0 + u: its rounds of up to 128 postings and the running total over the terms (ascending
                // term id = lane order)
                const uint32_t n_l = (h0 - g0) << BM25_GROUP_LOG;   // postings of the (padded) slice, a multiple of 8
                const uint32_t cnt = (n_l + 127u) >> 7;
                uint32_t incl = cnt;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const uint32_t v = __shfl_up_sync(FULL, incl, o);
                    if (lane >= o) incl += v;
                }
                const uint32_t rounds = __shfl_sync(FULL, incl, 31);
                if (rounds == 0) continue;
                if (!zeroed) {
                    if (!SUB_ZERO_TMA || !acc_clean) {
                        float4* a4 = reinterpret_cast<float4*>(acc);
#pragma unroll
                        for (int i = 0; i < BM25_SUB_DOCS / 128; ++i) a4[lane + 32 * i] = make_float4(0.f, 0.f, 0.f, 0.f);
                    }
                    acc_clean = false;
                    if (lane == 0) *cand_count = 0;
                    zeroed = true;
                }
                __syncwarp();  // zeroing (and the previous group of terms) done before the first add
                // The round stream is described 32 rounds at a time, one descriptor per lane: lane i of a chunk finds the
                // term of round base + i by bisection in the running totals (once per CHUNK, not per round) and works out
                // where the round's postings start and how many there are; the round loop then only broadcasts two words.
                for (uint32_t base = 0; base < rounds; base += 32) {
                    const uint32_t g = base + (uint32_t)lane;
                    int u = 0;
#pragma unroll
                    for (int stepb = 16; stepb > 0; stepb >>= 1) {
                        const uint32_t v = __shfl_sync(FULL, incl, u + stepb - 1);
                        if (v <= g) u += stepb;
                    }
                    u = u > 31 ? 31 : u;                                   // lanes past the last round (g >= rounds)
                    const uint32_t incl_u = __shfl_sync(FULL, incl, u), cnt_u = __shfl_sync(FULL, cnt, u);
                    const uint32_t g0_u = __shfl_sync(FULL, g0, u), n_u = __shfl_sync(FULL, n_l, u);
                    const uint32_t r_in = g - (incl_u - cnt_u);            // round inside its term
                    const uint32_t d_grp = g0_u + (r_in << (7 - BM25_GROUP_LOG));   // first group of 8 postings of the round
                    const uint32_t left = n_u - (r_in << 7);
                    const uint32_t d_meta = (left < 128u ? left : 128u) | (r_in == 0 ? STAGE_NEW_TERM : 0u);
                    const int n_chunk = (int)(rounds - base < 32u ? rounds - base : 32u);

                    // fetch: start the copies of round r (inside the chunk, warp-uniform) into stage r % SUB_DEPTH
                    auto fetch = [&](int stage, int r, uint32_t& meta_out) {
                        const uint32_t grp = __shfl_sync(FULL, d_grp, r);
                        const uint32_t meta = __shfl_sync(FULL, d_meta, r);
                        meta_out = meta;
                        if constexpr (SUB_STAGE_TMA) {
                            // every lane has read this stage's previous contents (apply() of SUB_DEPTH - 1 rounds ago)
                            __syncwarp();
                            if (lane == 0) {
                                const uint32_t count = meta & 0xFFu;   // a multiple of 8: 16 / 32-byte granules
                                const size_t p = (size_t)grp * BM25_GROUP;
                                uint64_t* bar = &tile->stage_bar[stage];
                                ptx::mbar_arrive_expect_tx(bar, count * 6u);
                                ptx::bulk_load_1d(&tile->stage_off[stage][0], post_off + p, count * 2u, bar);
                                ptx::bulk_load_1d(&tile->stage_imp[stage][0], post_imp + p, count * 4u, bar);
                            }
                        } else if (4u * (uint32_t)lane < (meta & 0xFFu)) {
                            const size_t p = (size_t)grp * BM25_GROUP + (uint32_t)lane * 4u;
                            cp_async_8(&tile->stage_off[stage][lane], post_off + p);
                            cp_async_16(&tile->stage_imp[stage][lane], post_imp + p);
                        }
                    };
                    auto apply = [&](int stage, uint32_t meta) {
                        if constexpr (SUB_STAGE_TMA) {
                            ptx::mbar_wait_lean(&tile->stage_bar[stage], (stage_phase >> stage) & 1u);
                            stage_phase ^= 1u << stage;
                        }
                        if (meta & STAGE_NEW_TERM) __syncwarp();  // the previous term's writes are visible to every lane
                        if (4u * (uint32_t)lane < (meta & 0xFFu)) {
                            // four postings per lane, no masks (padding postings add +0.0 to a trash slot)
                            const uint2 o = tile->stage_off[stage][lane];
                            const float4 wv = tile->stage_imp[stage][lane];
                            const uint32_t o0 = o.x & 0xFFFFu, o1 = o.x >> 16, o2 = o.y & 0xFFFFu, o3 = o.y >> 16;
                            float* p0 = reinterpret_cast<float*>(acc_b + o0);
                            float* p1 = reinterpret_cast<float*>(acc_b + o1);
                            float* p2 = reinterpret_cast<float*>(acc_b + o2);
                            float* p3 = reinterpret_cast<float*>(acc_b + o3);
                            // the four reads are issued together, then the adds, then the writes
                            const float a0 = *p0, a1 = *p1, a2 = *p2, a3 = *p3;
                            const float v0 = __fadd_rn(a0, wv.x), v1 = __fadd_rn(a1, wv.y);
                            const float v2 = __fadd_rn(a2, wv.z), v3 = __fadd_rn(a3, wv.w);
                            *p0 = v0;
                            *p1 = v1;
                            *p2 = v2;
                            *p3 = v3;
                            if constexpr (!SAMPLE) {
                                // Impacts are > 0: an accumulator only grows and crosses tau at most once. Cheap test first.
                                if (__float_as_uint(fmaxf(fmaxf(v0, v1), fmaxf(v2, v3))) >= cross_bits) {
                                    if (__float_as_uint(v0) >= cross_bits && __float_as_uint(a0) < cross_bits) record(o0);
                                    if (__float_as_uint(v1) >= cross_bits && __float_as_uint(a1) < cross_bits) record(o1);
                                    if (__float_as_uint(v2) >= cross_bits && __float_as_uint(a2) < cross_bits) record(o2);
                                    if (__float_as_uint(v3) >= cross_bits && __float_as_uint(a3) < cross_bits) record(o3);
                                }
                            }
                        }
                    };
                    // SUB_DEPTH stages in rotation. Every step commits exactly one cp.async group (an empty one past the
                    // end of the chunk), so "all but the SUB_DEPTH - 1 newest groups are complete" always means "the round
                    // about to be applied has landed". A lane only ever reads what it copied itself: no barrier needed.
                    uint32_t metas[SUB_DEPTH];
#pragma unroll
                    for (int i = 0; i < SUB_DEPTH; ++i) metas[i] = 0u;
#pragma unroll
                    for (int i = 0; i < SUB_DEPTH - 1; ++i) {
                        if (i < n_chunk) fetch(i, i, metas[i]);
                        if constexpr (!SUB_STAGE_TMA) cp_async_commit();
                    }
                    if constexpr (SUB_ZERO_TMA) zero_await();   // the zeros have landed before the first add
                    for (int r = 0; r < n_chunk; r += SUB_DEPTH) {
#pragma unroll
                        for (int i = 0; i < SUB_DEPTH; ++i) {
                            if (r + i < n_chunk) {
                                const int nxt = (i + SUB_DEPTH - 1) % SUB_DEPTH;
                                if (r + i + SUB_DEPTH - 1 < n_chunk) fetch(nxt, r + i + SUB_DEPTH - 1, metas[nxt]);
                                if constexpr (!SUB_STAGE_TMA) {
                                    cp_async_commit();
                                    cp_async_wait<SUB_DEPTH - 1>();
                                }
                                apply(i, metas[i]);
                            }
                        }
                    }
                    if constexpr (!SUB_STAGE_TMA) cp_async_wait<0>();   // (only empty groups can be left)
                }
            }
            if constexpr (SAMPLE) {
                if (zeroed) {
                    // lane l scans accumulators l, l + 32, ... (conflict-free) and reports its best (ties: lowest document)
                    __syncwarp();
                    const int32_t doc_base = s0 * BM25_SUB_DOCS;
                    uint32_t best = 0, best_rel = 0;
#pragma unroll 8
                    for (int i = 0; i < BM25_SUB_DOCS / 32; ++i) {
                        const uint32_t rel = (uint32_t)(lane + 32 * i);
                        bool ok = true;
                        const uint32_t v = __float_as_uint(acc[rel]);
                        if constexpr (FILT) {
                            if (v > best) {
                                const int32_t doc = doc_base + (int32_t)rel;
                                if (flt.doc_tags && (__ldg(flt.doc_tags + doc) & f_mask) == 0) ok = false;
                                if (f_group >= 0 && __ldg(flt.doc_group + doc) != f_group) ok = false;
                            }
                        }
                        if (v > best && ok) { best = v; best_rel = rel; }
                    }
                    const int ssub = (s0 / (stride * BM25_SUBS)) * BM25_SUBS + s0 % BM25_SUBS;  // index among the sampled ones
                    uint64_t key = best ? srag_pack_key(best, (uint32_t)doc_base + best_rel) : SRAG_KEY_EMPTY;
                    if (sample_group == 4) {  // the best of four lanes: fewer keys for the tau merge to sort
                        uint64_t o = __shfl_xor_sync(FULL, key, 1);
                        key = o > key ? o : key;
                        o = __shfl_xor_sync(FULL, key, 2);
                        key = o > key ? o : key;
                    }
                    const int per_sub = 32 / sample_group;
                    if ((lane & (sample_group - 1)) == 0)
                        partial[((size_t)q * per_q + ssub) * per_sub + lane / sample_group] = key;
                    __syncwarp();
                    if constexpr (SUB_ZERO_TMA) zero_issue();
                }
            } else if (zeroed) {
                __syncwarp();
                const uint32_t nc = *cand_count;
                if (nc > (uint32_t)SUB_CAND) {
                    flag_query(q);
                } else if (nc > 0) {
                    uint32_t base = 0;
                    if (lane == 0) base = atomicAdd(&q_count[q], nc);
                    base = __shfl_sync(FULL, base, 0);
                    if (base + nc > cap) {
                        flag_query(q);
                    } else {
                        uint64_t* out = partial + (size_t)q * cap + base;  // the region fills from the front
                        for (uint32_t i = lane; i < nc; i += 32) {
                            const uint32_t ob = cand[i];
                            out[i] = srag_pack_key(__float_as_uint(*reinterpret_cast<const float*>(acc_b + ob)),
                                                   (uint32_t)blk_doc0 + (ob >> 2));
                        }
                    }
                }
                __syncwarp();
                if constexpr (SUB_ZERO_TMA) zero_issue();
            }
        }
        // rotate the lookahead
        x0 = x1;
        x1 = x2;
        x2 = x3;
        x3 = advance(x3);
        s0 = s1;
        s1 = s2;
        s2 = s3;
        s3 = sub_of(x3);
        g0 = g1;
        h0 = h1;
        g1 = g2;
        h1 = h2;
        tau0 = tau1;
        bounds_of(s2, t3, g2, h2);
        tau1 = s1 >= 0 && !SAMPLE ? __ldg(tau + (size_t)x1.q * k + (k - 1)) : 0.f;
        t3 = term_of(x3, s3, 0);
    }
    if constexpr (SUB_ZERO_TMA) zero_await();   // no copy into this CTA's shared memory may outlive it
}

// Flagged queries are redone by bm25_items_kernel into per-block lists: clear everything but the sample's top-k.
__global__ void bm25_clear_flagged_kernel(uint64_t* __restrict__ partial, int k, uint32_t cap, const uint32_t* __restrict__ q_flag,
                                          const uint32_t* __restrict__ n_flagged) {
    if (*n_flagged == 0 || q_flag[blockIdx.x] == 0) return;
    uint64_t* p = partial + (size_t)blockIdx.x * cap;
    for (uint32_t i = (uint32_t)k + threadIdx.x; i < cap; i += blockDim.x) p[i] = SRAG_KEY_EMPTY;
}

// imp = idf[t] * (tf * (k1 + 1)) / (tf + k1 * ((1 - b) + b * (dl / avgdl))), one rounding per op
__global__ void bm25_impacts_kernel(const int64_t* __restrict__ term_ptr, int vocab,
                                    const int32_t* __restrict__ post_doc, const int32_t* __restrict__ post_tf,
                                    int64_t nnz, const float* __restrict__ idf, const int32_t* __restrict__ dl,
                                    float avgdl, float* __restrict__ post_imp) {
    const float k1 = 1.2f, b = 0.75f;
    const float one_minus_b = __fadd_rn(1.0f, -b);
    const float k1p1 = __fadd_rn(k1, 1.0f);
    for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < nnz; p += (int64_t)gridDim.x * blockDim.x) {
        int lo = 0, hi = vocab;  // largest t with term_ptr[t] <= p
        while (hi - lo > 1) {
            const int mid = (lo + hi) >> 1;
            if (__ldg(term_ptr + mid) <= p) lo = mid; else hi = mid;
        }
        const float tf = (float)post_tf[p];
        const float r = __fdiv_rn((float)dl[post_doc[p]], avgdl);
        const float t = __fmul_rn(b, r);
        const float u = __fadd_rn(one_minus_b, t);
        const float nrm = __fmul_rn(k1, u);
        const float a = __fmul_rn(tf, k1p1);
        const float num = __fmul_rn(idf[lo], a);
        const float den = __fadd_rn(tf, nrm);
        post_imp[p] = __fdiv_rn(num, den);
    }
}

__global__ void bm25_block_index_kernel(const int64_t* __restrict__ term_ptr, const int32_t* __restrict__ post_doc,
                                        int vocab, int n_blocks, int block_docs, int64_t* __restrict__ block_ptr) {
    const int64_t total = (int64_t)vocab * (n_blocks + 1);
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int t = (int)(i / (n_blocks + 1));
        const int j = (int)(i - (int64_t)t * (n_blocks + 1));
        int64_t lo = term_ptr[t], hi = term_ptr[t + 1];
        const int64_t target = (int64_t)j * block_docs;  // first posting with doc >= target
        while (lo < hi) {
            const int64_t mid = (lo + hi) >> 1;
            if ((int64_t)__ldg(post_doc + mid) < target) lo = mid + 1; else hi = mid;
        }
        block_ptr[i] = lo;
    }
}

// ---------------------------------------------------------------------------------------------------------
// Packing (index build): canonical CSR postings (int32 documents ascending inside a term, fp32 impacts, block_ptr int64
// [vocab][n_sub + 1]) -> the SEARCHED layout:
//   * every (term, sub-block) slice starts on a group of 8 postings and is padded to whole groups with zero-impact
//     postings that land on a trash slot (sub_trash_offset), so the scoring kernel loads 8 / 16 aligned bytes per lane
//     and needs no mask anywhere; slice_ptr uint32 [vocab][n_sub + 1] holds the slice starts in GROUPS (so 2^32 groups =
//     1.7e10 postings per shard);
//   * post_off uint16 = byte offset of the document's accumulator inside its 16384-document block's tile;
//   * BANK ORDER: inside a slice every posting is a distinct document and receives exactly one add per query, so the
//     ORDER of a slice's postings is free: sums, and therefore scores, do not depend on it. The scoring kernels
//     read-add-write acc[doc] with lane l holding postings 4l .. 4l+3 of a group of 128, i.e. one shared-memory access
//     covers the 32 postings at positions 128g + 4l + e (fixed g, e). With postings in ascending document order their
//     banks (doc & 31) are as good as random: ~3.4 wavefronts per access. The fill kernel lays each slice out "round-robin
//     over banks" (all first-of-their-bank postings in bank order, then all second ones, ...; a window of 32 consecutive
//     entries of full rounds hits every bank once) and writes that sequence to the slice's positions in ACCESS order
//     (g, e, l). The floor is max_b count(b) wavefronts per slice; simulation: 3.2 -> 2.1 wavefronts per access for slices
//     of 100 postings, 3.4 -> 1.6 for 800. One warp per slice, deterministic (stable inside a bank).
// ---------------------------------------------------------------------------------------------------------
constexpr int BANK_ORDER_WARPS = 8;

// groups[t] = sum over the term's slices of ceil(n / 8)
__global__ void pack_row_sum_kernel(const int64_t* __restrict__ block_ptr, int vocab, int n_sub, int64_t* __restrict__ groups) {
    const int lane = threadIdx.x & 31;
    const int64_t warps = (int64_t)gridDim.x * (blockDim.x >> 5);
    for (int64_t t = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); t < vocab; t += warps) {
        const int64_t* row = block_ptr + (size_t)t * (n_sub + 1);
        int64_t s = 0;
        for (int j = lane; j < n_sub; j += 32) s += (__ldg(row + j + 1) - __ldg(row + j) + (BM25_GROUP - 1)) >> BM25_GROUP_LOG;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xFFFFFFFFu, s, o);
        if (lane == 0) groups[t] = s;
    }
}

// term_group[t] = exclusive scan of groups (one CTA; every thread owns a contiguous segment); [vocab] = total
__global__ void __launch_bounds__(1024) pack_scan_terms_kernel(const int64_t* __restrict__ groups, int vocab,
                                                               int64_t* __restrict__ term_group) {
    __shared__ int64_t part[1024];
    const int per = (vocab + 1023) / 1024;
    const int lo = threadIdx.x * per, hi = min(vocab, lo + per);
    int64_t s = 0;
    for (int i = lo; i < hi; ++i) s += groups[i];
    part[threadIdx.x] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        int64_t run = 0;
        for (int i = 0; i < 1024; ++i) {
            const int64_t v = part[i];
            part[i] = run;
            run += v;
        }
        term_group[vocab] = run;
    }
    __syncthreads();
    int64_t run = part[threadIdx.x];
    for (int i = lo; i < hi; ++i) {
        const int64_t v = groups[i];   // (groups and term_group may not alias)
        term_group[i] = run;
        run += v;
    }
}

// slice_ptr[t][j] = term_group[t] + sum_{j' < j} ceil(n_{t j'} / 8); [t][n_sub] = term_group[t + 1]
__global__ void pack_row_scan_kernel(const int64_t* __restrict__ block_ptr, int vocab, int n_sub,
                                     const int64_t* __restrict__ term_group, uint32_t* __restrict__ slice_ptr) {
    const int lane = threadIdx.x & 31;
    const int64_t warps = (int64_t)gridDim.x * (blockDim.x >> 5);
    for (int64_t t = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); t < vocab; t += warps) {
        const int64_t* row = block_ptr + (size_t)t * (n_sub + 1);
        uint32_t* out = slice_ptr + (size_t)t * (n_sub + 1);
        int64_t carry = term_group[t];
        for (int base = 0; base <= n_sub; base += 32) {
            const int j = base + lane;
            const int64_t v = j < n_sub ? (__ldg(row + j + 1) - __ldg(row + j) + (BM25_GROUP - 1)) >> BM25_GROUP_LOG : 0;
            int64_t inc = v;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int64_t u = __shfl_up_sync(0xFFFFFFFFu, inc, o);
                if (lane >= o) inc += u;
            }
            if (j <= n_sub) out[j] = (uint32_t)(carry + inc - v);
            carry += __shfl_sync(0xFFFFFFFFu, inc, 31);
        }
    }
}

__global__ void __launch_bounds__(32 * BANK_ORDER_WARPS)
bm25_pack_fill_kernel(const int64_t* __restrict__ block_ptr, const uint32_t* __restrict__ slice_ptr, int vocab, int n_sub,
                      const int32_t* __restrict__ doc_in, const float* __restrict__ imp_in,
                      uint16_t* __restrict__ off_out, float* __restrict__ imp_out, int permute) {
    __shared__ uint32_t cnt_s[BANK_ORDER_WARPS][32];
    constexpr uint32_t FULL = 0xFFFFFFFFu;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    uint32_t* cnt = cnt_s[w];
    const int64_t n_slices = (int64_t)vocab * n_sub;
    const int64_t n_warps = (int64_t)gridDim.x * BANK_ORDER_WARPS;
    for (int64_t s0 = ((int64_t)blockIdx.x * BANK_ORDER_WARPS + w) * 32; s0 < n_slices; s0 += n_warps * 32) {
        // the bounds of 32 consecutive slices, one per lane
        int64_t lo_l = 0, hi_l = 0, dst_l = 0;
        int sub_l = 0;
        if (s0 + lane < n_slices) {
            const int64_t t = (s0 + lane) / n_sub, j = (s0 + lane) - t * n_sub;
            const int64_t* bp = block_ptr + t * (n_sub + 1) + j;
            lo_l = __ldg(bp);
            hi_l = __ldg(bp + 1);
            dst_l = (int64_t)__ldg(slice_ptr + t * (n_sub + 1) + j) << BM25_GROUP_LOG;
            sub_l = (int)(j & (BM25_SUBS - 1));
        }
        uint32_t todo = __ballot_sync(FULL, hi_l > lo_l);
        while (todo) {
            const int src = __ffs(todo) - 1;
            todo &= todo - 1;
            const int64_t lo = __shfl_sync(FULL, lo_l, src), hi = __shfl_sync(FULL, hi_l, src);
            const int64_t a = __shfl_sync(FULL, dst_l, src);   // first position of the padded slice
            const int sub_in_block = __shfl_sync(FULL, sub_l, src);
            const int n = (int)(hi - lo);
            const int n_pad = (n + BM25_GROUP - 1) & ~(BM25_GROUP - 1);
            // padding: zero impact, offset = the trash slot of whichever warp scores this sub-block
            if (lane < n_pad - n) {
                off_out[a + n + lane] = (uint16_t)sub_trash_offset(sub_in_block);
                imp_out[a + n + lane] = 0.f;
            }
            // pass 1: postings per bank (lane b ends up holding the count of bank b)
            cnt[lane] = 0;
            __syncwarp();
            for (int i = lane; i < n; i += 32) atomicAdd(&cnt[__ldg(doc_in + lo + i) & 31], 1u);
            __syncwarp();
            const uint32_t c = cnt[lane];
            __syncwarp();
            cnt[lane] = 0;  // pass 2: postings of each bank placed so far
            __syncwarp();
            for (int base = 0; base < n; base += 32) {
                const int i = base + lane;
                const bool valid = i < n;
                const int32_t d = valid ? __ldg(doc_in + lo + i) : 0;
                const float im = valid ? __ldg(imp_in + lo + i) : 0.f;
                const uint32_t b = (uint32_t)d & 31u;
                const uint32_t peers = __match_any_sync(FULL, valid ? b : 32u + (uint32_t)lane);
                // r = how many postings of this bank come before this one (stable: slice order)
                const uint32_t r = cnt[b] + __popc(peers & ((1u << lane) - 1u));
                __syncwarp();
                if (valid && (peers >> lane) == 1u) cnt[b] = r + 1;  // the highest lane of the group
                __syncwarp();
                // index in the round-robin sequence: every bank contributes min(count, r) entries to rounds 0..r-1,
                // and the banks below b that still have an r-th posting come first in round r
                uint32_t sidx = 0;
#pragma unroll
                for (int bb = 0; bb < 32; ++bb) {
                    const uint32_t cb = __shfl_sync(FULL, c, bb);
                    sidx += (cb < r ? cb : r) + (((uint32_t)bb < b && cb > r) ? 1u : 0u);
                }
                if (valid) {
                    int pos = i;   // permute == 0: ascending documents, only packed
                    if (permute) {
                        // the sidx-th position of the slice in access order (group g of 128, element e, lane l)
                        const int g = (int)sidx >> 7;
                        int kk = (int)sidx - 128 * g;
                        const int tg = n - 128 * g < 128 ? n - 128 * g : 128;
                        pos = -1;
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            const int ne = (tg - e + 3) >> 2;   // lanes that hold an element e in this group
                            if (pos < 0) {
                                if (kk < ne) pos = 128 * g + 4 * kk + e;
                                else kk -= ne;
                            }
                        }
                    }
                    off_out[a + pos] = (uint16_t)(((uint32_t)d & (uint32_t)(BM25_BLOCK_DOCS - 1)) << 2);
                    imp_out[a + pos] = im;
                }
            }
            __syncwarp();
        }
    }
}

// The inverse of the packing: document id of every stored posting (-1 for padding) from (slice, byte offset). One warp
// per 32 slices; used by tests, `canonical_postings` and shard conversion — never on the search path.
__global__ void __launch_bounds__(32 * BANK_ORDER_WARPS)
bm25_unpack_docs_kernel(const uint32_t* __restrict__ slice_ptr, int vocab, int n_sub, const uint16_t* __restrict__ post_off,
                        const float* __restrict__ post_imp, int32_t* __restrict__ doc_out, int32_t* __restrict__ term_out) {
    constexpr uint32_t FULL = 0xFFFFFFFFu;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int64_t n_slices = (int64_t)vocab * n_sub;
    const int64_t n_warps = (int64_t)gridDim.x * BANK_ORDER_WARPS;
    for (int64_t s0 = ((int64_t)blockIdx.x * BANK_ORDER_WARPS + w) * 32; s0 < n_slices; s0 += n_warps * 32) {
        int64_t lo_l = 0, hi_l = 0;
        int32_t base_l = 0, term_l = 0;
        if (s0 + lane < n_slices) {
            const int64_t t = (s0 + lane) / n_sub, j = (s0 + lane) - t * n_sub;
            const uint32_t* sp = slice_ptr + t * (n_sub + 1) + j;
            lo_l = (int64_t)__ldg(sp) << BM25_GROUP_LOG;
            hi_l = (int64_t)__ldg(sp + 1) << BM25_GROUP_LOG;
            base_l = (int32_t)(j / BM25_SUBS) * BM25_BLOCK_DOCS;
            term_l = (int32_t)t;
        }
        uint32_t todo = __ballot_sync(FULL, hi_l > lo_l);
        while (todo) {
            const int src = __ffs(todo) - 1;
            todo &= todo - 1;
            const int64_t lo = __shfl_sync(FULL, lo_l, src), hi = __shfl_sync(FULL, hi_l, src);
            const int32_t base = __shfl_sync(FULL, base_l, src), term = __shfl_sync(FULL, term_l, src);
            for (int64_t p = lo + lane; p < hi; p += 32) {
                doc_out[p] = post_imp[p] != 0.f ? base + (int32_t)(post_off[p] >> 2) : -1;
                if (term_out) term_out[p] = term;
            }
        }
    }
}

int bm25_blocks(int64_t n_docs) { return (int)((n_docs + BM25_BLOCK_DOCS - 1) / BM25_BLOCK_DOCS); }  // tiles
int bm25_subs(int64_t n_docs) { return (int)((n_docs + BM25_SUB_DOCS - 1) / BM25_SUB_DOCS); }       // index granularity
constexpr int BM25_Q_CHUNK = 1024;  // queries whose per-block lists are resident at once (0.5 GB at 10 M docs, k = 100)

}  // namespace

extern "C" int srag_bm25_block_docs(void) { return BM25_SUB_DOCS; }

extern "C" int srag_bm25_impacts(const int64_t* term_ptr, int vocab, const int32_t* post_doc, const int32_t* post_tf,
                                 int64_t nnz, const float* idf, const int32_t* dl, float avgdl, float* post_imp,
                                 void* stream) {
    SRAG_CHECK_ARG(nnz >= 0 && vocab >= 1, "bad sizes nnz=%lld vocab=%d", (long long)nnz, vocab);
    if (nnz == 0) return SRAG_OK;  // a collection whose texts all tokenise to nothing: avgdl = 0 is fine, no postings
    SRAG_CHECK_ARG(avgdl > 0.f, "bad avgdl=%f with nnz=%lld", (double)avgdl, (long long)nnz);
    SRAG_CHECK_ARG(term_ptr && post_doc && post_tf && idf && dl && post_imp, "null pointer argument");
    int64_t blocks = (nnz + 255) / 256;
    if (blocks > 148 * 32) blocks = 148 * 32;
    bm25_impacts_kernel<<<(int)blocks, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
        term_ptr, vocab, post_doc, post_tf, nnz, idf, dl, avgdl, post_imp);
    SRAG_CUDA(cudaGetLastError());
    return SRAG_OK;
}

extern "C" int srag_bm25_block_index(const int64_t* term_ptr, const int32_t* post_doc, int vocab, int64_t n_docs,
                                     int block_docs, int64_t* block_ptr, void* stream) {
    SRAG_CHECK_ARG(vocab >= 1 && n_docs >= 0 && n_docs < 0x7FFFFFFF && block_docs == BM25_SUB_DOCS,
                   "bad sizes vocab=%d n_docs=%lld block_docs=%d", vocab, (long long)n_docs, block_docs);
    SRAG_CHECK_ARG(term_ptr && block_ptr, "null pointer argument");
    const int nb = bm25_subs(n_docs) > 0 ? bm25_subs(n_docs) : 1;
    int64_t total = (int64_t)vocab * (nb + 1);
    int64_t blocks = (total + 255) / 256;
    if (blocks > 148 * 32) blocks = 148 * 32;
    bm25_block_index_kernel<<<(int)blocks, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
        term_ptr, post_doc, vocab, nb, block_docs, block_ptr);
    SRAG_CUDA(cudaGetLastError());
    return SRAG_OK;
}

static int pack_grid(int64_t n_slices) {
    const int64_t rounds = (n_slices + 32 * BANK_ORDER_WARPS - 1) / (32 * BANK_ORDER_WARPS);
    return (int)(rounds < 148 * 64 ? (rounds > 0 ? rounds : 1) : 148 * 64);
}

extern "C" int srag_bm25_pack_plan(const int64_t* block_ptr, int vocab, int64_t n_docs, int64_t* scratch, uint32_t* slice_ptr,
                                   void* stream) {
    SRAG_CHECK_ARG(vocab >= 1 && n_docs >= 0 && n_docs < 0x7FFFFFFF, "bad sizes vocab=%d n_docs=%lld", vocab,
                   (long long)n_docs);
    SRAG_CHECK_ARG(block_ptr && scratch && slice_ptr, "null pointer argument");
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    const int nb = bm25_subs(n_docs) > 0 ? bm25_subs(n_docs) : 1;
    int64_t* groups = scratch;                 // [vocab]
    int64_t* term_group = scratch + vocab;     // [vocab + 1]; term_group[vocab] = total groups
    int64_t blocks = ((int64_t)vocab + 7) / 8;
    if (blocks > 148 * 8) blocks = 148 * 8;
    pack_row_sum_kernel<<<(int)blocks, 256, 0, st>>>(block_ptr, vocab, nb, groups);
    SRAG_CUDA(cudaGetLastError());
    pack_scan_terms_kernel<<<1, 1024, 0, st>>>(groups, vocab, term_group);
    SRAG_CUDA(cudaGetLastError());
    pack_row_scan_kernel<<<(int)blocks, 256, 0, st>>>(block_ptr, vocab, nb, term_group, slice_ptr);
    SRAG_CUDA(cudaGetLastError());
    return SRAG_OK;
}

extern "C" int srag_bm25_pack_fill(const int64_t* block_ptr, const uint32_t* slice_ptr, int vocab, int64_t n_docs,
                                   const int32_t* doc_in, const float* imp_in, uint16_t* off_out, float* imp_out,
                                   int permute, void* stream) {
    SRAG_CHECK_ARG(vocab >= 1 && n_docs >= 0 && n_docs < 0x7FFFFFFF, "bad sizes vocab=%d n_docs=%lld", vocab,
                   (long long)n_docs);
    if (n_docs == 0) return SRAG_OK;
    SRAG_CHECK_ARG(block_ptr && slice_ptr && doc_in && imp_in && off_out && imp_out, "null pointer argument");
    SRAG_CHECK_ARG(imp_in != imp_out, "packing is not in place");
    const int nb = bm25_subs(n_docs);
    bm25_pack_fill_kernel<<<pack_grid((int64_t)vocab * nb), 32 * BANK_ORDER_WARPS, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
        block_ptr, slice_ptr, vocab, nb, doc_in, imp_in, off_out, imp_out, permute);
    SRAG_CUDA(cudaGetLastError());
    return SRAG_OK;
}

extern "C" int srag_bm25_unpack_docs(const uint32_t* slice_ptr, int vocab, int64_t n_docs, const uint16_t* post_off,
                                     const float* post_imp, int32_t* doc_out, int32_t* term_out, void* stream) {
    SRAG_CHECK_ARG(vocab >= 1 && n_docs >= 0 && n_docs < 0x7FFFFFFF, "bad sizes vocab=%d n_docs=%lld", vocab,
                   (long long)n_docs);
    if (n_docs == 0) return SRAG_OK;
    SRAG_CHECK_ARG(slice_ptr && post_off && post_imp && doc_out, "null pointer argument");
    const int nb = bm25_subs(n_docs);
    bm25_unpack_docs_kernel<<<pack_grid((int64_t)vocab * nb), 32 * BANK_ORDER_WARPS, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
        slice_ptr, vocab, nb, post_off, post_imp, doc_out, term_out);
    SRAG_CUDA(cudaGetLastError());
    return SRAG_OK;
}

// workspace = per-block lists of one query chunk + the sample top-k (scores, ids) that yields tau
static size_t bm25_partial_bytes(int qc, int nb, int k) { return (((size_t)qc * nb * k * sizeof(uint64_t)) + 255) & ~(size_t)255; }
// sample every stride-th block first when there are enough blocks for the sample to be worth a launch
static int bm25_stride(int nb) {
    static int forced = -1;
    if (forced < 0) {
        const char* e = getenv("SRAG_BM25_STRIDE");
        forced = e ? atoi(e) : 0;
    }
    if (forced >= 2) return nb >= 2 * forced ? forced : 1;
    // every 32nd block when there are plenty (C3: 12.12 ms at 16, 11.64 at 32, 12.06 at 64: the sample costs an exact
    // selection per block, a lower tau costs candidates), every 16th for mid-sized indexes, no sample below 32 blocks
    return nb >= 64 ? 32 : nb >= 32 ? 16 : 1;
}
constexpr int SUB_SAMPLE_KEYS = BM25_SUBS * 32;  // sample reports per sampled block on the sub-kernel path (32 per sub-block)
static size_t bm25_sample_bytes(int qc, int nb, int k) {
    const int stride = bm25_stride(nb);
    return stride > 1 ? bm25_partial_bytes(qc, (nb + stride - 1) / stride, k > SUB_SAMPLE_KEYS ? k : SUB_SAMPLE_KEYS) : 0;
}
static size_t bm25_tau_bytes(int qc, int k) { return (((size_t)qc * k * (sizeof(float) + sizeof(int64_t))) + 255) & ~(size_t)255; }
// bm25_sub_kernel: candidate counter and overflow flag per query + the number of flagged queries
static size_t bm25_ctr_bytes(int qc) { return (((size_t)(2 * qc + 1) * sizeof(uint32_t)) + 255) & ~(size_t)255; }

extern "C" int srag_bm25_workspace_bytes(int64_t n_docs, int n_queries, int k, size_t* bytes) {
    SRAG_CHECK_ARG(bytes && n_queries >= 1 && k >= 1 && k <= SRAG_MAX_BM25_K && n_docs >= 0, "bad sizes");
    const int nb = bm25_blocks(n_docs) > 0 ? bm25_blocks(n_docs) : 1;
    const int qc = n_queries < BM25_Q_CHUNK ? n_queries : BM25_Q_CHUNK;
    *bytes = bm25_partial_bytes(qc, nb, k) + bm25_sample_bytes(qc, nb, k) + bm25_tau_bytes(qc, k) + bm25_ctr_bytes(qc) + 256;
    return SRAG_OK;
}

static int bm25_search_impl(const uint32_t* block_ptr, const uint16_t* post_doc, const float* post_imp, int vocab,
                            int64_t n_docs, const int32_t* q_terms, int n_queries, int max_terms, int k, int64_t id_base,
                            Bm25Filter flt, float* out_scores, int64_t* out_ids, void* workspace, size_t workspace_bytes,
                            void* stream) {
    SRAG_CHECK_ARG((flt.doc_tags == nullptr) == (flt.q_tag_mask == nullptr) &&
                       (flt.doc_group == nullptr) == (flt.q_group == nullptr),
                   "doc_tags / q_tag_mask and doc_group / q_group come in pairs");
    const bool filtered = flt.doc_tags != nullptr || flt.doc_group != nullptr;
    SRAG_CHECK_ARG(n_queries >= 1 && max_terms >= 1 && k >= 1 && k <= SRAG_MAX_BM25_K, "bad sizes n_queries=%d k=%d",
                   n_queries, k);
    SRAG_CHECK_ARG(n_docs >= 0 && n_docs < 0x7FFFFFFF && vocab >= 1, "bad index sizes");
    SRAG_CHECK_ARG(block_ptr && q_terms && out_scores && out_ids && workspace, "null pointer argument");
    SRAG_CHECK_ARG(n_docs == 0 || (post_doc && post_imp), "null postings");
    SRAG_CHECK_ARG(((uintptr_t)post_doc & 15) == 0 && ((uintptr_t)post_imp & 15) == 0 && ((uintptr_t)workspace & 15) == 0,
                   "post_off / post_imp / workspace must be 16-byte aligned");
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    const int nb = bm25_blocks(n_docs) > 0 ? bm25_blocks(n_docs) : 1;
    const int qc = n_queries < BM25_Q_CHUNK ? n_queries : BM25_Q_CHUNK;
    const int n_sub = bm25_subs(n_docs) > 0 ? bm25_subs(n_docs) : 1;
    const size_t need = bm25_partial_bytes(qc, nb, k) + bm25_sample_bytes(qc, nb, k) + bm25_tau_bytes(qc, k) + bm25_ctr_bytes(qc);
    if (workspace_bytes < need) {
        srag_set_error("workspace too small: %zu < %zu", workspace_bytes, need);
        return SRAG_E_WORKSPACE;
    }
    int sm_count = 0;
    {
        int rc = srag_func_smem_optin(reinterpret_cast<const void*>(&bm25_items_kernel<false>), (int)sizeof(Bm25Smem));
        if (!rc) rc = srag_func_smem_optin(reinterpret_cast<const void*>(&bm25_items_kernel<true>), (int)sizeof(Bm25Smem));
        if (!rc) rc = srag_func_smem_optin(reinterpret_cast<const void*>(&bm25_sub_kernel<false, false>), (int)sizeof(SubSmem));
        if (!rc) rc = srag_func_smem_optin(reinterpret_cast<const void*>(&bm25_sub_kernel<false, true>), (int)sizeof(SubSmem));
        if (!rc) rc = srag_func_smem_optin(reinterpret_cast<const void*>(&bm25_sub_kernel<true, false>), (int)sizeof(SubSmem));
        if (!rc) rc = srag_func_smem_optin(reinterpret_cast<const void*>(&bm25_sub_kernel<true, true>), (int)sizeof(SubSmem));
        if (!rc) rc = srag_current_device_props(&sm_count, nullptr, nullptr);
        if (rc) return rc;
        // SRAG_BM25_SMS: scan on fewer SMs (see SRAG_TOPK_SMS)
        if (const char* e = getenv("SRAG_BM25_SMS")) {
            const int lim = atoi(e);
            if (lim >= 1 && lim < sm_count) sm_count = lim;
        }
    }
    uint8_t* ws = reinterpret_cast<uint8_t*>(workspace);
    uint64_t* partial = reinterpret_cast<uint64_t*>(ws);
    uint64_t* sample = reinterpret_cast<uint64_t*>(ws + bm25_partial_bytes(qc, nb, k));
    // ids first: the region starts 256-byte aligned, so the 8-byte id stores of the finalize kernel stay aligned for any
    // qc * k (scores first put the ids on a 4-byte boundary whenever qc * k was odd, e.g. B = 1, k = 5)
    int64_t* tau_ids = reinterpret_cast<int64_t*>(ws + bm25_partial_bytes(qc, nb, k) + bm25_sample_bytes(qc, nb, k));
    float* tau_scores = reinterpret_cast<float*>(tau_ids + (size_t)qc * k);
    uint32_t* q_count = reinterpret_cast<uint32_t*>(ws + bm25_partial_bytes(qc, nb, k) + bm25_sample_bytes(qc, nb, k) +
                                                    bm25_tau_bytes(qc, k));
    uint32_t* q_flag = q_count + qc;
    uint32_t* n_flagged = q_flag + qc;
    static int use_sub = -1;  // SRAG_BM25_SUBKERNEL=0: phase 1 on bm25_items_kernel alone (A/B measurements)
    if (use_sub < 0) {
        const char* e = getenv("SRAG_BM25_SUBKERNEL");
        use_sub = e ? atoi(e) != 0 : 1;
    }
    const int stride = bm25_stride(nb);
    const int nb0 = (nb + stride - 1) / stride;
    const char* dbg_env = getenv("SRAG_BM25_DEBUG");
    const int dbg = dbg_env ? atoi(dbg_env) : 0;
    const int resident = sm_count * BM25_CTAS_PER_SM;
    for (int q0 = 0; q0 < n_queries; q0 += qc) {
        const int nq = (n_queries - q0) < qc ? (n_queries - q0) : qc;
        const bool sub_path = use_sub && stride > 1 && n_docs > 0;
        if (sub_path) {
            // Warp-granular path. SAMPLE pass over every stride-th block -> 32 true scores per sampled sub-block -> tau =
            // the k-th best of them; MAIN pass over ALL sub-blocks (the sampled ones again: 1/stride more postings, but no
            // exact per-block selection anywhere) appends the documents that cross tau to the query's region, which is
            // filled from the front and read up to its fill level, so it needs no zeroing. Queries whose candidates
            // overflow are flagged, their region cleared, and redone by the exact per-block kernel over all blocks.
            auto kern = filtered ? bm25_items_kernel<true> : bm25_items_kernel<false>;
            auto skern_sample = filtered ? bm25_sub_kernel<true, true> : bm25_sub_kernel<false, true>;
            auto skern = filtered ? bm25_sub_kernel<true, false> : bm25_sub_kernel<false, false>;
            // 8 reports per sampled sub-block when that still leaves plenty more than k of them, else 32
            const int sgroup = (long long)nb0 * BM25_SUBS * 8 >= 8ll * k ? 4 : 1;
            const size_t sample_keys = (size_t)nb0 * (SUB_SAMPLE_KEYS / sgroup);
            SRAG_CUDA(cudaMemsetAsync(sample, 0, (size_t)nq * sample_keys * sizeof(uint64_t), st));
            SRAG_CUDA(cudaMemsetAsync(q_count, 0, bm25_ctr_bytes(qc), st));
            const long long w0 = ((long long)nq * nb0 * BM25_SUBS + SUB_WARPS - 1) / SUB_WARPS;
            skern_sample<<<(int)(w0 < sm_count ? w0 : sm_count), 32 * SUB_WARPS, sizeof(SubSmem), st>>>(
                block_ptr, post_doc, post_imp, nb, n_sub, q_terms, max_terms, k, q0, nq, stride, nullptr, sample, q_count,
                q_flag, n_flagged, flt, sgroup);
            SRAG_CUDA(cudaGetLastError());
            int rc = srag_launch_finalize_ex(sample, sample_keys, nq, (int)sample_keys, k, nullptr, 0, tau_scores, tau_ids,
                                             nullptr, 0, st);
            if (rc) return rc;
            const long long w1 = ((long long)nq * nb * BM25_SUBS + SUB_WARPS - 1) / SUB_WARPS;
            skern<<<(int)(w1 < sm_count ? w1 : sm_count), 32 * SUB_WARPS, sizeof(SubSmem), st>>>(
                block_ptr, post_doc, post_imp, nb, n_sub, q_terms, max_terms, k, q0, nq, stride, tau_scores, partial, q_count,
                q_flag, n_flagged, flt, 1);
            SRAG_CUDA(cudaGetLastError());
            bm25_clear_flagged_kernel<<<nq, 256, 0, st>>>(partial, 0, (uint32_t)nb * (uint32_t)k, q_flag, n_flagged);
            SRAG_CUDA(cudaGetLastError());
            const long long itemsf = (long long)nq * nb;
            kern<<<(int)(itemsf < resident ? itemsf : resident), BM25_CTA_THREADS, sizeof(Bm25Smem), st>>>(
                block_ptr, post_doc, post_imp, nb, n_sub, q_terms, max_terms, k, q0, nq, 0, 1, nullptr, partial, dbg, flt,
                n_docs, q_flag, n_flagged);
            SRAG_CUDA(cudaGetLastError());
            rc = srag_launch_finalize_live(partial, (size_t)nb * k, nq, nb * k, k, nullptr, id_base,
                                           out_scores + (size_t)q0 * k, out_ids + (size_t)q0 * k, nullptr, 0, q_count, q_flag,
                                           0, st);
            if (rc) return rc;
            continue;
        }
        // producer / consumer kernel only (indexes too small to sample, SRAG_BM25_SUBKERNEL=0): the memset covers the
        // sample lists too
        SRAG_CUDA(cudaMemsetAsync(partial, 0, bm25_partial_bytes(qc, nb, k) + (stride > 1 ? (size_t)nq * nb0 * k * sizeof(uint64_t) : 0), st));
        if (n_docs > 0) {
            const long long items0 = (long long)nq * nb0;
            const int grid0 = (int)(items0 < resident ? items0 : resident);
            auto kern = filtered ? bm25_items_kernel<true> : bm25_items_kernel<false>;
            kern<<<grid0, BM25_CTA_THREADS, sizeof(Bm25Smem), st>>>(block_ptr, post_doc, post_imp, nb, n_sub, q_terms, max_terms,
                                                                    k, q0, nq, 0, stride, nullptr,
                                                                    stride > 1 ? sample : partial, dbg, flt, n_docs, nullptr,
                                                                    nullptr);
            SRAG_CUDA(cudaGetLastError());
            if (stride > 1) {
                // top-k of the sample: scores give tau; the keys seed the final merge in block 0's slot of
                // `partial` (block 0 always belongs to the sample, so phase 1 never writes there)
                int rc = srag_launch_finalize_ex(sample, (size_t)nb0 * k, nq, nb0 * k, k, nullptr, 0, tau_scores, tau_ids,
                                                 partial, (size_t)nb * k, st);
                if (rc) return rc;
                const long long items1 = (long long)nq * (nb - nb0);
                const int grid1 = (int)(items1 < resident ? items1 : resident);
                kern<<<grid1, BM25_CTA_THREADS, sizeof(Bm25Smem), st>>>(block_ptr, post_doc, post_imp, nb, n_sub, q_terms,
                                                                        max_terms, k, q0, nq, 1, stride, tau_scores, partial,
                                                                        dbg, flt, n_docs, nullptr, nullptr);
                SRAG_CUDA(cudaGetLastError());
            }
        }
        int rc = srag_launch_finalize(partial, nq, nb * k, k, nullptr, id_base, out_scores + (size_t)q0 * k,
                                      out_ids + (size_t)q0 * k, st);
        if (rc) return rc;
    }
    return SRAG_OK;
}

extern "C" int srag_bm25_search(const uint32_t* block_ptr, const uint16_t* post_doc, const float* post_imp, int vocab,
                                int64_t n_docs, const int32_t* q_terms, int n_queries, int max_terms, int k,
                                int64_t id_base, float* out_scores, int64_t* out_ids, void* workspace,
                                size_t workspace_bytes, void* stream) {
    return bm25_search_impl(block_ptr, post_doc, post_imp, vocab, n_docs, q_terms, n_queries, max_terms, k, id_base,
                            Bm25Filter{nullptr, nullptr, nullptr, nullptr}, out_scores, out_ids, workspace,
                            workspace_bytes, stream);
}

extern "C" int srag_bm25_search_filtered(const uint32_t* block_ptr, const uint16_t* post_doc, const float* post_imp,
                                         int vocab, int64_t n_docs, const int32_t* q_terms, int n_queries, int max_terms,
                                         int k, int64_t id_base, const uint32_t* doc_tags, const int32_t* doc_group,
                                         const uint32_t* q_tag_mask, const int32_t* q_group, float* out_scores,
                                         int64_t* out_ids, void* workspace, size_t workspace_bytes, void* stream) {
    return bm25_search_impl(block_ptr, post_doc, post_imp, vocab, n_docs, q_terms, n_queries, max_terms, k, id_base,
                            Bm25Filter{doc_tags, doc_group, q_tag_mask, q_group}, out_scores, out_ids, workspace,
                            workspace_bytes, stream);
}
