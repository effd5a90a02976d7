// K1: exact top-k vector search = tcgen05 bf16 GEMM (fp32 accumulators in TMEM) fused with
// per-row scaling (L2 normalisation) and a per-query threshold filter + candidate compaction.
// Scores never reach HBM; the only HBM traffic is the corpus (once), the queries and a few
// candidate lists.
//
// Replaces `collection.query(query_embeddings, n_results)` at
// super_rag/vectorstore/seekdb_connector.py:103-107 of the reference (see include/srag_b200.h).
//
// Work decomposition
//   unit      = one CTA (cta_group 1) or one CTA pair (cta_group 2, tcgen05.mma M = 256)
//   q-tile    = 128 * cta_group consecutive queries; unit u owns q-tile (u % T_q) for the whole
//               kernel, so one TMEM lane == one query
//   slab      = u / T_q; the unit scans corpus tiles  slab, slab + S, slab + 2S, ...
//               ids ascending, so "score > threshold" is the exact tie rule
//   warps     0: TMA producer   1: MMA issuer + TMEM owner   2..9: epilogue. Epilogue warp w may
//               read TMEM lane quadrant w % 4; the two warps of a quadrant split the tile's columns,
//               so a query is filtered by two threads with independent thresholds / candidate lists
//   pipelines smem full/empty ring (TMA <-> MMA), TMEM full/empty x2 (MMA <-> epilogue)
//
// Two operand plans (template RQ):
//   RQ = true  (dim <= 1024) "resident queries": the unit's q-tile never moves again after
//              set-up — k-blocks 0..7 sit in TMEM columns 256..511 (tcgen05.mma with A from TMEM),
//              k-blocks 8..15 in 128 KB of shared memory. Only corpus tiles (128 rows) stream
//              through the ring, which halves the bytes every SM has to ingest per flop: the
//              streaming plan needs 64 B/clk/SM at full tensor rate — the whole L2->SM port.
//   RQ = false streaming: query and corpus k-blocks both go through the ring (256-row tiles);
//              any dim.
#include <cuda.h>

#include <mutex>

#include "common.cuh"

namespace {

constexpr int BLOCK_M = 128;  // queries per CTA  (TMEM lanes)
constexpr int BLOCK_K = 64;   // bf16 per 128-byte swizzle row
constexpr int UMMA_K = 16;
constexpr int MAX_EPI_WARPS = 8;  // epilogue warps per CTA: 4 or 8 (template EW)
constexpr int TMEM_COLS = 512;
constexpr int PROG_SLOTS = 8;   // q-tile units per slab that pace each other (one 32-byte sector)
constexpr int MAX_STAGES = 12;
constexpr int RQ_TMEM_KB = 8;   // query k-blocks resident in TMEM (8 * 32 columns at column 256)
constexpr int RQ_SMEM_KB = 8;   // query k-blocks resident in shared memory
constexpr int RQ_MAX_DIM = (RQ_TMEM_KB + RQ_SMEM_KB) * BLOCK_K;
constexpr int RQ_A_TMEM_COL = 256;

template <int G, bool RQ>
struct Cfg {
    static constexpr int BLOCK_N = RQ ? 128 : 256;  // corpus rows per tile = fp32 columns per accumulator
    static constexpr int B_ROWS = BLOCK_N / G;      // corpus rows THIS CTA stages per k-block
    static constexpr int A_BYTES = BLOCK_M * BLOCK_K * 2;
    static constexpr int B_BYTES = B_ROWS * BLOCK_K * 2;
    static constexpr int STAGE_BYTES = RQ ? B_BYTES : A_BYTES + B_BYTES;
    static constexpr int QRES_BYTES = RQ ? RQ_SMEM_KB * A_BYTES : 0;
    static constexpr int STAGES = RQ ? (G == 2 ? 11 : 5) : (G == 1 ? 4 : 6);
    static constexpr int RING_BYTES = STAGES * STAGE_BYTES;
    static_assert(STAGES <= MAX_STAGES, "barrier arrays too small");
};

struct __align__(16) SmemTail {
    uint64_t full[MAX_STAGES];
    uint64_t empty[MAX_STAGES];
    uint64_t tmem_full[2];
    uint64_t tmem_empty[2];
    uint64_t q_full;   // resident query k-blocks landed in shared memory (TMA)
    uint64_t q_tmem;   // resident query k-blocks written to TMEM by the epilogue warps
    uint32_t tmem_base;
    uint32_t pad;
    alignas(16) float scale[1024];  // per-warp copies of the row multipliers of its columns (EW * BLOCK_N / (EW / 4))
};

template <int G, bool RQ>
constexpr size_t smem_bytes() {
    return 1024 /*alignment slack*/ + Cfg<G, RQ>::QRES_BYTES + Cfg<G, RQ>::RING_BYTES + sizeof(SmemTail);
}

struct TopkParams {
    const float* row_scale;  // [n_rows]
    const __nv_bfloat16* queries;  // [n_queries][dim] (RQ: read directly for the TMEM-resident part)
    int64_t n_rows;
    int dim;
    int n_queries;
    int k;
    int num_q_tiles;  // T_q
    int num_slabs;    // S
    uint64_t* partial;  // [n_queries][S * EW/4][heap_stride(k)] per-thread 4-ary min-heaps of sortable keys, pre-zeroed (= empty)
    uint32_t* gthr;   // [S * EW/4][n_queries] j-th best score of every list (orderable bits), pre-zeroed
    int share_j;      // every list publishes its j-th best score, 1..8; 0 = threshold sharing off
    int share_need;   // the shared bound is the share_need-th largest published value, share_need = ceil(k / j) <= 16
    int share_period; // the bound is re-read every tile for the first share_early tiles, then every share_period-th
    int share_early;
    int prescan;      // 1: tile 0 is read twice from TMEM — maxima first, pushes once the shared bound exists
    int* progress;    // [S][2 ranks][PROG_SLOTS] tiles fully issued per q-tile unit, pre-zeroed; null = no pacing
    int window;       // tiles the leader may run ahead of the slowest follower
    int warm;         // warm-up tiles per list (0 = off): see "Warm-up" in the epilogue
    int debug;        // SRAG_TOPK_DEBUG perf-dissection bits (results are WRONG when set): 2 no MMA, 4 no epilogue work
    // metadata pre-filter (all null = off): a row is eligible for query q iff
    //   (row_tags == null  || (row_tags[row] & q_tag_mask[q]) != 0) &&
    //   (row_group == null || q_group[q] < 0 || row_group[row] == q_group[q])
    const uint32_t* row_tags;
    const int32_t* row_group;
    const uint32_t* q_tag_mask;
    const int32_t* q_group;
};

// ---------------------------------------------------------------------------------------
// Per-thread running top-k: a 4-ary MIN-heap of k sortable keys that lives directly in this
// thread's slice of `partial` (pre-zeroed = EMPTY keys, which is a valid heap). A candidate that
// beats the threshold replaces the root (the current k-th best) and sifts down: thread-private
// loads/stores, no warp cooperation, and the threshold is always the exact k-th best seen so far.
// The lists are larger than L1 at k = 100 (100 KB per CTA), so every level of the sift is an L2
// round trip: four children per node halve the depth of a binary heap, and the list is laid out
// so that a node's four children are ONE aligned 32-byte sector: logical node j sits in slot
// j + 3 (slots 0..2 stay EMPTY), children 4j+1..4j+4 -> slots 4j+4..4j+7. A list is
// heap_stride(k) slots so the last (partial) child group stays inside it.
// Precondition: key > root. Returns the new threshold.
// ---------------------------------------------------------------------------------------
__host__ __device__ constexpr int heap_stride(int k) { return (k + 6 + 3) & ~3; }

__device__ __noinline__ float heap_replace_root(uint64_t* heap, int k, uint32_t score_bits, uint32_t row) {
    const uint64_t key = srag_pack_key(score_bits, row);
    uint64_t root = key;
    int i = 0;  // logical node
    for (;;) {
        const int c0 = 4 * i + 1;
        if (c0 >= k) break;
        const ulonglong2 a = *reinterpret_cast<const ulonglong2*>(heap + c0 + 3);
        const ulonglong2 b = *reinterpret_cast<const ulonglong2*>(heap + c0 + 5);
        uint64_t kc = a.x;
        int c = c0;
        if (c0 + 1 < k && a.y < kc) { kc = a.y; c = c0 + 1; }
        if (c0 + 2 < k && b.x < kc) { kc = b.x; c = c0 + 2; }
        if (c0 + 3 < k && b.y < kc) { kc = b.y; c = c0 + 3; }
        if (kc >= key) break;
        heap[i + 3] = kc;
        if (i == 0) root = kc;
        i = c;
    }
    heap[i + 3] = key;
    return root == SRAG_KEY_EMPTY ? -INFINITY : __uint_as_float(srag_key_score_bits(root));
}


// While a list holds fewer than k keys it is only APPENDED to (slot count + 3: one store, no loads; the finalize kernel
// reads every slot whatever its order). When the k-th key arrives the slots are turned into the heap above in place,
// bottom-up, once. With threshold sharing most lists never fill (~j ln(n/j) pushes against k slots), so their pushes
// never pay a sift. Returns the threshold (the root's score).
__device__ __noinline__ float heap_build(uint64_t* heap, int k) {
    for (int start = (k - 2) >> 2; start >= 0; --start) {
        const uint64_t key = heap[start + 3];
        int i = start;
        for (;;) {
            const int c0 = 4 * i + 1;
            if (c0 >= k) break;
            const ulonglong2 a = *reinterpret_cast<const ulonglong2*>(heap + c0 + 3);
            const ulonglong2 b = *reinterpret_cast<const ulonglong2*>(heap + c0 + 5);
            uint64_t kc = a.x;
            int c = c0;
            if (c0 + 1 < k && a.y < kc) { kc = a.y; c = c0 + 1; }
            if (c0 + 2 < k && b.x < kc) { kc = b.x; c = c0 + 2; }
            if (c0 + 3 < k && b.y < kc) { kc = b.y; c = c0 + 3; }
            if (kc >= key) break;
            heap[i + 3] = kc;
            i = c;
        }
        heap[i + 3] = key;
    }
    const uint64_t root = heap[3];
    return root == SRAG_KEY_EMPTY ? -INFINITY : __uint_as_float(srag_key_score_bits(root));
}

// sv[e] for a run-time e without touching memory: a 5-level select tree (31 FSEL) over registers
__device__ __forceinline__ float select32(const float (&sv)[32], int e) {
    float a[16], b[8], c[4];
#pragma unroll
    for (int i = 0; i < 16; ++i) a[i] = (e & 1) ? sv[2 * i + 1] : sv[2 * i];
#pragma unroll
    for (int i = 0; i < 8; ++i) b[i] = (e & 2) ? a[2 * i + 1] : a[2 * i];
#pragma unroll
    for (int i = 0; i < 4; ++i) c[i] = (e & 4) ? b[2 * i + 1] : b[2 * i];
    const float d0 = (e & 8) ? c[1] : c[0], d1 = (e & 8) ? c[3] : c[2];
    return (e & 16) ? d1 : d0;
}

// ---------------------------------------------------------------------------------------
// main kernel
// ---------------------------------------------------------------------------------------
// FILT: metadata pre-filter compiled into the push path. WARM: the warm-up pass (large k) compiled in — a template
// parameter because merely carrying the run-time `warm` logic cost the HBM-bound plans 15-25 % (B = 128, k = 10: 3.7 ->
// 4.8 ms on one box), like the run-time filter flag did in round 1.
template <int G, bool RQ, int EW, bool FILT = false, bool WARM = false>
__global__ void __launch_bounds__(64 + 32 * EW, 1)
topk_gemm_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_c,
                 const TopkParams p) {
    using C = Cfg<G, RQ>;
    constexpr int BLOCK_N = C::BLOCK_N;
    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const uint32_t cta_rank = (G == 2) ? ptx::cluster_ctarank() : 0u;
    const int unit = (G == 2) ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
    const int q_tile = unit % p.num_q_tiles;
    const int slab = unit / p.num_q_tiles;
    if (slab >= p.num_slabs) return;  // spare units (both CTAs of a pair leave together)

    const int num_kb = (p.dim + BLOCK_K - 1) / BLOCK_K;
    const int num_c_tiles = (int)((p.n_rows + BLOCK_N - 1) / BLOCK_N);
    const int q_row = (q_tile * G + (int)cta_rank) * BLOCK_M;  // first query of this CTA
    // The unit's tile sequence: its first `warm` tiles once for the warm-up (epilogue: thresholds only), then ALL its
    // tiles slab, slab + S, ... Producer, MMA issuer and epilogue walk the same sequence.
    const int n_my = num_c_tiles > slab ? (num_c_tiles - slab + p.num_slabs - 1) / p.num_slabs : 0;
    const int warm = WARM ? (p.warm < n_my ? p.warm : n_my) : 0;
    const int n_iter = n_my + warm;
    auto tile_of = [&](int it_) { return slab + (it_ < warm ? it_ : it_ - warm) * p.num_slabs; };

    extern __shared__ uint8_t smem_raw[];
    uint8_t* qres = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* ring = qres + C::QRES_BYTES;
    SmemTail* tail = reinterpret_cast<SmemTail*>(ring + C::RING_BYTES);

    // ---- one-time setup ----------------------------------------------------------------
    if (warp == 0 && lane == 0) {
        ptx::prefetch_tmap(&tmap_q);
        ptx::prefetch_tmap(&tmap_c);
        for (int s = 0; s < C::STAGES; ++s) {
            ptx::mbar_init(&tail->full[s], 1);
            ptx::mbar_init(&tail->empty[s], 1);
        }
        for (int b = 0; b < 2; ++b) {
            ptx::mbar_init(&tail->tmem_full[b], 1);
            ptx::mbar_init(&tail->tmem_empty[b], EW * G);
        }
        ptx::mbar_init(&tail->q_full, 1);
        ptx::mbar_init(&tail->q_tmem, EW * G);
        ptx::fence_mbar_init();
    }
    if (warp == 1) {
        ptx::tmem_alloc<G>(&tail->tmem_base, TMEM_COLS);
        ptx::tmem_relinquish<G>();
    }
    ptx::tc_fence_before();
    if constexpr (G == 2) ptx::cluster_sync_all(); else __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *(volatile uint32_t*)&tail->tmem_base;
    const int num_qres = RQ ? max(0, num_kb - RQ_TMEM_KB) : 0;  // query k-blocks resident in shared memory

    // Producer and MMA warps run their loops warp-uniformly and predicate only the single-thread
    // instructions with elect.sync: addresses and descriptors then live in uniform registers and the
    // issue loop is a handful of SASS instructions per tcgen05.mma / TMA (one thread looping alone
    // paid an ELECT + R2UR per operand, ~130 instructions per k-block: issue-bound).
    if (warp == 0) {
        // ===== TMA producer ===============================================================
        if (RQ && num_qres > 0) {
            if (ptx::elect_one()) {
                if constexpr (G == 1) {
                    ptx::mbar_arrive_expect_tx(&tail->q_full, (uint32_t)num_qres * C::A_BYTES);
                    for (int j = 0; j < num_qres; ++j)
                        ptx::tma_load_2d(qres + j * C::A_BYTES, &tmap_q, &tail->q_full, (RQ_TMEM_KB + j) * BLOCK_K,
                                         q_row);
                } else {
                    if (cta_rank == 0) ptx::mbar_arrive_expect_tx(&tail->q_full, 2u * (uint32_t)num_qres * C::A_BYTES);
                    for (int j = 0; j < num_qres; ++j)
                        ptx::tma_load_2d_2sm(qres + j * C::A_BYTES, &tmap_q, &tail->q_full,
                                             (RQ_TMEM_KB + j) * BLOCK_K, q_row, 0x1000000000000000ull);
                }
            }
            __syncwarp();
        }
        uint32_t stage = 0, phase = 0;
        // L2 pacing: the T_q units of a slab stream the SAME corpus tiles. Unit 0
        // (leader) pulls a tile from HBM; the others (followers) start it one tile later, when it
        // sits in L2, and the leader never runs more than `window` tiles ahead of the slowest
        // follower. A hint only: waits are bounded, correctness never depends on them.
        volatile int* prog = p.progress ? p.progress + ((size_t)slab * 2 + cta_rank) * PROG_SLOTS : nullptr;
        const bool paced = prog != nullptr && p.num_q_tiles > 1 && p.num_q_tiles <= PROG_SLOTS;
        for (int it = 0; it < n_iter; ++it) {
            const int c = tile_of(it);
            const int c_row = c * BLOCK_N + (int)cta_rank * C::B_ROWS;
            if (paced) {
                uint32_t spins = 0;
                if (q_tile == 0) {
                    for (;;) {
                        int lo = 0x7FFFFFFF;
                        for (int j = 1; j < p.num_q_tiles; ++j) lo = min(lo, prog[j]);
                        if (lo >= it - p.window || ++spins > 4096u) break;
                        __nanosleep(64);
                    }
                } else {
                    while (prog[0] < it + 1 && ++spins <= 4096u) __nanosleep(64);
                }
            }
            for (int kb = 0; kb < num_kb; ++kb) {
                ptx::mbar_wait(&tail->empty[stage], phase ^ 1);
                if (ptx::elect_one()) {
                    uint8_t* s_dst = ring + stage * C::STAGE_BYTES;
                    uint8_t* b_dst = RQ ? s_dst : s_dst + C::A_BYTES;
                    if constexpr (G == 1) {
                        ptx::mbar_arrive_expect_tx(&tail->full[stage], C::STAGE_BYTES);
                        if constexpr (!RQ)
                            ptx::tma_load_2d_hint(s_dst, &tmap_q, &tail->full[stage], kb * BLOCK_K, q_row,
                                                  SRAG_L2_EVICT_LAST);
                        ptx::tma_load_2d(b_dst, &tmap_c, &tail->full[stage], kb * BLOCK_K, c_row);
                    } else {
                        // bytes of BOTH CTAs are counted on the leader's barrier
                        if (cta_rank == 0) ptx::mbar_arrive_expect_tx(&tail->full[stage], 2 * C::STAGE_BYTES);
                        if constexpr (!RQ)
                            ptx::tma_load_2d_2sm(s_dst, &tmap_q, &tail->full[stage], kb * BLOCK_K, q_row,
                                                 SRAG_L2_EVICT_LAST);
                        ptx::tma_load_2d_2sm(b_dst, &tmap_c, &tail->full[stage], kb * BLOCK_K, c_row,
                                             0x1000000000000000ull /* evict normal */);
                    }
                }
                __syncwarp();
                if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
            }
            if (paced && lane == 0) prog[q_tile] = it + 1;
        }
    } else if (warp == 1) {
        // ===== MMA issuer (leader CTA) =====================================================
        if (cta_rank == 0) {
            constexpr uint32_t idesc = ptx::umma_idesc_bf16(BLOCK_M * G, BLOCK_N);
            constexpr uint32_t KSTEP = (UMMA_K * 2) >> 4;  // descriptor start-address units per UMMA_K
            if constexpr (RQ) {
                ptx::mbar_wait(&tail->q_tmem, 0);
                if (num_qres > 0) ptx::mbar_wait(&tail->q_full, 0);
                ptx::tc_fence_after();
            }
            const uint64_t desc0 = ptx::umma_desc_k128(0);
            const uint32_t ring_lo = (ptx::smem_u32(ring) & 0x3FFFFu) >> 4;
            const uint32_t qres_lo = (ptx::smem_u32(qres) & 0x3FFFFu) >> 4;
            uint32_t stage = 0, phase = 0;
            for (int it = 0; it < n_iter; ++it) {
                const int buf = it & 1;
                const uint32_t use = (uint32_t)(it >> 1);
                ptx::mbar_wait(&tail->tmem_empty[buf], (use & 1) ^ 1);
                ptx::tc_fence_after();
                const uint32_t d_tmem = tmem_base + (uint32_t)buf * BLOCK_N;
                for (int kb = 0; kb < num_kb; ++kb) {
                    ptx::mbar_wait(&tail->full[stage], phase);
                    ptx::tc_fence_after();
                    const uint32_t s_lo = ring_lo + stage * (uint32_t)(C::STAGE_BYTES >> 4);
                    if (ptx::elect_one()) {
                        if (!(p.debug & 2)) {
                            if constexpr (RQ) {
                                const uint64_t bd = desc0 | (uint64_t)s_lo;
                                if (kb < RQ_TMEM_KB) {
                                    const uint32_t a_tmem = tmem_base + RQ_A_TMEM_COL + (uint32_t)kb * (BLOCK_K / 2);
#pragma unroll
                                    for (int k = 0; k < BLOCK_K / UMMA_K; ++k)
                                        ptx::umma_bf16_ts<G>(d_tmem, a_tmem + k * (UMMA_K / 2), bd + k * KSTEP, idesc,
                                                             (uint32_t)((kb | k) != 0));
                                } else {
                                    const uint64_t ad =
                                        desc0 | (uint64_t)(qres_lo + (uint32_t)(kb - RQ_TMEM_KB) * (C::A_BYTES >> 4));
#pragma unroll
                                    for (int k = 0; k < BLOCK_K / UMMA_K; ++k)
                                        ptx::umma_bf16<G>(d_tmem, ad + k * KSTEP, bd + k * KSTEP, idesc, 1u);
                                }
                            } else {
                                const uint64_t ad = desc0 | (uint64_t)s_lo;
                                const uint64_t bd = ad + (uint64_t)(C::A_BYTES >> 4);
#pragma unroll
                                for (int k = 0; k < BLOCK_K / UMMA_K; ++k)
                                    ptx::umma_bf16<G>(d_tmem, ad + k * KSTEP, bd + k * KSTEP, idesc,
                                                      (uint32_t)((kb | k) != 0));
                            }
                        }
                        ptx::umma_commit<G>(&tail->empty[stage]);  // smem slot reusable once these MMAs retire
                        if (kb == num_kb - 1) ptx::umma_commit<G>(&tail->tmem_full[buf]);  // accumulator ready
                    }
                    __syncwarp();
                    if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else {
        // ===== epilogue: lane == query; EW/4 warps (column slices) per TMEM lane quadrant ======
        constexpr int H = EW / 4;             // warps (threads) that share one query
        constexpr int CPW = BLOCK_N / H;      // tile columns per warp
        constexpr int CHUNKS = CPW / 32;      // 32-column TMEM reads per warp and tile
        constexpr int SPL = CPW / 32;         // consecutive row multipliers each lane fetches
        constexpr int LPC = 32 / SPL;         // lanes whose multipliers cover one 32-column chunk
        const int ew = warp - 2;        // 0..EW-1
        const int quad = warp & 3;      // TMEM lane quadrant this warp may access
        const int half = ew >> 2;       // which column slice of the tile
        const int row_in_cta = quad * 32 + lane;
        const int q_global = q_row + row_in_cta;
        const uint32_t lane_addr = tmem_base + ((uint32_t)(quad * 32) << 16);

        if constexpr (RQ) {
            // park query k-blocks 0..7 in TMEM: each thread copies ITS row, 64 bf16 (32 columns) at a time
            const uint4* qsrc = reinterpret_cast<const uint4*>(p.queries + (size_t)q_global * p.dim);
            constexpr int PER = RQ_TMEM_KB / H;
            for (int cch = half * PER; cch < half * PER + PER; ++cch) {
                if (cch * BLOCK_K >= p.dim) break;  // warp-uniform
                uint32_t r[32];
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    uint4 v = make_uint4(0u, 0u, 0u, 0u);
                    if (q_global < p.n_queries && cch * BLOCK_K + j * 8 < p.dim) v = __ldg(qsrc + cch * 8 + j);
                    r[4 * j + 0] = v.x; r[4 * j + 1] = v.y; r[4 * j + 2] = v.z; r[4 * j + 3] = v.w;
                }
                ptx::tmem_st_32x32(lane_addr + RQ_A_TMEM_COL + (uint32_t)cch * 32u, r);
            }
            ptx::tmem_st_wait();
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) {
                if constexpr (G == 2) ptx::mbar_arrive_cluster(&tail->q_tmem, 0);
                else ptx::mbar_arrive(&tail->q_tmem);
            }
        }

        // this thread's heap: partial[q][slab * H + half][heap_stride(k)]; threads of padding queries never insert
        uint64_t* heap = p.partial + ((size_t)min(q_global, p.n_queries - 1) * (p.num_slabs * H) + slab * H + half) * heap_stride(p.k);
        float* sc = tail->scale + ew * CPW;
        const float nanf_ = __uint_as_float(0x7FC00000u);
        const int col0 = half * CPW;
        // Threshold sharing. The query is scanned by L = S * H lists (one per unit and column slice) over disjoint rows.
        // Each list publishes its j-th best score. T = the n-th largest published value, n = ceil(k / j): at least n
        // lists hold >= j rows scoring >= T, i.e. >= k rows overall score >= T, so a row of the global top-k scores >= T
        // and is never dropped by "> thr" with thr = max(own k-th best, T - 2 ulp). (Round 1 used the MIN over all lists
        // with j = ceil(k / L): the same argument with n = L, but held down by the unluckiest list and undefined until
        // every list had published. With j = 1, n = k the bound after one tile is already ~ the k-th best of everything
        // scanned so far by anyone.) Cuts heap pushes per thread from ~k ln(n/k) to a handful.
        const bool valid_q = q_global < p.n_queries;
        const int qx = min(q_global, p.n_queries - 1);
        const int num_lists = p.num_slabs * H;
        const int my_list = slab * H + half;
        float topj[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) topj[i] = -INFINITY;
        float thr_own = -INFINITY, thr_shared = -INFINITY;
        uint32_t published = 0;
        int heap_cnt = 0;   // WARM kernels: keys appended so far; == k: the list is a heap
        float thr = valid_q ? -INFINITY : INFINITY;
        // metadata pre-filter: checked only for the few scores that beat the running threshold
        constexpr bool filtered = FILT;  // a template parameter: the unfiltered kernel carries none of this
        const uint32_t f_mask = p.row_tags ? __ldg(p.q_tag_mask + qx) : 0xFFFFFFFFu;
        const int32_t f_group = p.row_group ? __ldg(p.q_group + qx) : -1;
        auto eligible = [&](uint32_t row) -> bool {
            if (p.row_tags && (__ldg(p.row_tags + row) & f_mask) == 0) return false;
            if (f_group >= 0 && __ldg(p.row_group + row) != f_group) return false;
            return true;
        };

        auto load_scales = [&](int c, float (&r)[SPL]) {
            const int64_t row0 = (int64_t)c * BLOCK_N + col0 + lane * SPL;
            if (row0 + SPL <= p.n_rows) {
                if constexpr (SPL == 2) {
                    const float2 a = __ldg(reinterpret_cast<const float2*>(p.row_scale + row0));
                    r[0] = a.x; r[1] = a.y;
                } else {
#pragma unroll
                    for (int i = 0; i < SPL; i += 4) {
                        const float4 a = __ldg(reinterpret_cast<const float4*>(p.row_scale + row0 + i));
                        r[i] = a.x; r[i + 1] = a.y; r[i + 2] = a.z; r[i + 3] = a.w;
                    }
                }
            } else {
#pragma unroll
                for (int i = 0; i < SPL; ++i) r[i] = (row0 + i < p.n_rows) ? __ldg(p.row_scale + row0 + i) : nanf_;
            }
        };

        // Warm-up (k >= 32 with threshold sharing). With an empty heap every early row is a push, and a push at k = 100 is a
        // sift of dependent L2 round trips: ~330 of a thread's ~370 pushes fell into its first four tiles, before the
        // first shared threshold existed, and stalled the MMA pipe through tmem_empty (18.6 ms per batch at k = 100 against
        // 15.7 ms at k = 10). So the first `warm` tiles are scanned TWICE: the first time only the j best scores are kept in
        // registers (topj) and published; then every list waits (bounded) for the others' j-th best, and the real scan
        // starts over from the first tile with T = min over lists already in place: ~j pushes for those tiles instead of
        // ~k ln(rows / k). T is valid for the same reason as ever: it counts j DISTINCT rows per list (topj is cleared
        // before the rescan, `published` only ever rises), so >= k rows overall score >= T.
        bool warm_mode = WARM && warm > 0;
        float nxt[SPL];
        if (n_iter > 0) load_scales(tile_of(0), nxt);

        // share_need-th largest published value of this query's lists (orderable bits; <= 0x00800001 = not there yet)
        // and how many lists have REPORTED at all (a value, or the mark 1 = "scanned, nothing to publish yet")
        auto shared_bound = [&](int& n_pub) -> uint32_t {
            uint32_t best[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) best[i] = 0u;
            const volatile uint32_t* gp = p.gthr + qx;
            int cnt = 0;
#pragma unroll 1
            for (int l0 = 0; l0 < num_lists; l0 += 8) {   // 8 slots per L2 round trip (the loads are issued together)
                uint32_t vv[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) vv[u] = l0 + u < num_lists ? gp[(size_t)(l0 + u) * p.n_queries] : 0u;
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    uint32_t v = vv[u];
                    cnt += v != 0u;   // reported: a real value, or the "nothing yet" mark 1
#pragma unroll
                    for (int i = 0; i < 16; ++i) {  // keep the 16 largest sorted (descending)
                        const uint32_t hi = max(best[i], v);
                        v = min(best[i], v);
                        best[i] = hi;
                    }
                }
            }
            uint32_t r = best[0];
#pragma unroll
            for (int i = 1; i < 16; ++i) r = (i < p.share_need) ? best[i] : r;
            n_pub = cnt;
            return r;
        };
        auto publish = [&]() {
            float tj = topj[0];
#pragma unroll
            for (int i = 1; i < 8; ++i) tj = (i < p.share_j) ? topj[i] : tj;
            if (tj > -INFINITY) {
                const uint32_t o = srag_f32_orderable(__float_as_uint(tj));
                if (o > published) {  // single writer per slot: a plain (volatile) store
                    *reinterpret_cast<volatile uint32_t*>(p.gthr + (size_t)my_list * p.n_queries + qx) = o;
                    published = o;
                }
            } else if (published == 0u) {
                // fewer than j eligible rows so far (deleted rows, a selective metadata filter): say so, or everyone who
                // waits for all lists after the prescan would wait out the whole bound on every search
                *reinterpret_cast<volatile uint32_t*>(p.gthr + (size_t)my_list * p.n_queries + qx) = 1u;
                published = 1u;
            }
        };
        auto now_ns = [&]() -> uint64_t {
            uint64_t t;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
            return t;
        };

        // (Tried and dropped: the push path out of line, once per kernel — spill the chunk's 32 scores to local memory when
        // one passes and call a routine that loops over them, state in local memory. The kernels shrank from 5.1 K / 7.9 K to
        // 2.8 K instructions, but a push became ~2000 cycles of dependent local-memory accesses: 1.25 M rows, B = 1024:
        // k = 10 1.87 -> 2.54 ms, k = 100 3.04 -> 6.8 ms. The cold start is bound by the LATENCY of the push path in the
        // epilogue warps, not by instruction fetch.)
        // Cold start. With empty heaps and no bound every early score is a push, and because the 32 lanes of a warp
        // (32 queries) push at different columns, a warp pays a sift's dependent L2 round trips at nearly EVERY column of
        // its first tiles: one tile per unit took 0.41 ms against 0.03 ms in steady state — a fixed ~0.25 ms per search,
        // 12 % of a 1.25 M-row shard's scan (the N = 8 strong-scaling point). So tile 0 is read from TMEM TWICE (the
        // accumulator is simply not released in between; the MMAs of tile 1 run into the other buffer meanwhile): first
        // only the j best scores are kept in registers and published; every list then waits (bounded — a hint, never a
        // correctness condition) for the others and takes T as above; then the normal pass over tile 0 pushes only what
        // beats T (~k rows per QUERY, not per thread). topj is cleared in between, `published` only ever rises, so the
        // j rows a published value stands for are distinct rows of this list.
        if (p.prescan && p.share_j > 0 && n_iter > 0) {
            __syncwarp();
            if constexpr (SPL == 2) {
                *reinterpret_cast<float2*>(sc + lane * 2) = make_float2(nxt[0], nxt[1]);
            } else {
#pragma unroll
                for (int i = 0; i < SPL; i += 4)
                    *reinterpret_cast<float4*>(sc + lane * SPL + i) = make_float4(nxt[i], nxt[i + 1], nxt[i + 2], nxt[i + 3]);
            }
            __syncwarp();
            ptx::mbar_wait(&tail->tmem_full[0], 0);
            ptx::tc_fence_after();
            const uint32_t t_addr0 = lane_addr + (uint32_t)col0;
            const uint32_t row_base0 = (uint32_t)tile_of(0) * BLOCK_N + (uint32_t)col0;
            float tj = valid_q ? -INFINITY : INFINITY;
#pragma unroll 1
            for (int ch = 0; ch < CHUNKS; ++ch) {
                uint32_t v[32];
                ptx::tmem_ld_32x32(t_addr0 + (uint32_t)ch * 32u, v);
                ptx::tmem_ld_wait_dep(v);
                float x[32];
                float mx = -INFINITY;
#pragma unroll
                for (int e = 0; e < 32; ++e) {
                    x[e] = __uint_as_float(v[e]) * sc[ch * 32 + e];   // the product the main pass computes (NaN never passes)
                    mx = fmaxf(mx, x[e]);
                }
                if (mx > tj) {
                    uint32_t mask = 0u;
#pragma unroll
                    for (int e = 0; e < 32; ++e) mask |= (x[e] > tj ? 1u : 0u) << e;
#pragma unroll 1
                    while (mask) {
                        const int e = __ffs(mask) - 1;
                        mask &= mask - 1u;
                        float y = select32(x, e);
                        if (y > tj && (!filtered || eligible(row_base0 + (uint32_t)(ch * 32 + e)))) {
#pragma unroll
                            for (int i = 0; i < 8; ++i) {
                                const float hi = fmaxf(topj[i], y);
                                y = fminf(topj[i], y);
                                topj[i] = hi;
                            }
                            tj = topj[0];
#pragma unroll
                            for (int i = 1; i < 8; ++i) tj = (i < p.share_j) ? topj[i] : tj;
                        }
                    }
                }
            }
            if (valid_q) {
                publish();
                const int lists_with_rows = (num_c_tiles < p.num_slabs ? num_c_tiles : p.num_slabs) * H;
                const uint64_t t_end = now_ns() + 40000u;   // <= 40 us: the lists' units start together, tile 0 takes ~10
                uint32_t b;
                for (;;) {
                    int n_pub;
                    b = shared_bound(n_pub);
                    if (n_pub >= lists_with_rows || now_ns() > t_end) break;
                    __nanosleep(64);
                }
                if (b > 0x00800001u) {
                    thr_shared = __uint_as_float(srag_f32_from_orderable(b - 2u));
                    thr = fmaxf(thr_own, thr_shared);
                }
            }
#pragma unroll
            for (int i = 0; i < 8; ++i) topj[i] = -INFINITY;
            __syncwarp();
        }
        for (int it = 0; it < n_iter; ++it) {
            const int c = tile_of(it);
            const int buf = it & 1;
            const uint32_t use = (uint32_t)(it >> 1);
            __syncwarp();
            if constexpr (SPL == 2) {
                *reinterpret_cast<float2*>(sc + lane * 2) = make_float2(nxt[0], nxt[1]);
            } else {
#pragma unroll
                for (int i = 0; i < SPL; i += 4)
                    *reinterpret_cast<float4*>(sc + lane * SPL + i) = make_float4(nxt[i], nxt[i + 1], nxt[i + 2], nxt[i + 3]);
            }
            // per-chunk bounds of the row multipliers (NaN = tombstone / padding is ignored by fmaxf/fminf;
            // a chunk of only such rows gets NaN bounds and is skipped by the filter below)
            float s_hi_l = nxt[0], s_lo_l = nxt[0];  // after the butterfly: bounds of this lane's chunk
#pragma unroll
            for (int i = 1; i < SPL; ++i) { s_hi_l = fmaxf(s_hi_l, nxt[i]); s_lo_l = fminf(s_lo_l, nxt[i]); }
#pragma unroll
            for (int o = 1; o < LPC; o <<= 1) {
                s_hi_l = fmaxf(s_hi_l, __shfl_xor_sync(0xFFFFFFFFu, s_hi_l, o));
                s_lo_l = fminf(s_lo_l, __shfl_xor_sync(0xFFFFFFFFu, s_lo_l, o));
            }
            __syncwarp();
            if (it + 1 < n_iter) load_scales(tile_of(it + 1), nxt);

            if (WARM && warm_mode && it == warm) {  // the warm-up is over: the real scan starts with an empty heap and empty topj
                warm_mode = false;
#pragma unroll
                for (int i = 0; i < 8; ++i) topj[i] = -INFINITY;
                if (valid_q) thr = -INFINITY;
            }
            // refresh the shared bound: every tile while it still moves fast, then every 4th
            if (p.share_j > 0 && valid_q && !warm_mode && (it - warm < p.share_early || (it - warm) % p.share_period == 0)) {
                // right after a warm-up pass: wait (bounded — a hint, never a correctness condition) until every list of
                // this query has published
                const bool after_warm = warm > 0 && it == warm;
                const uint64_t t_end = after_warm ? now_ns() + 100000u : 0u;
                const int lists_with_rows = (num_c_tiles < p.num_slabs ? num_c_tiles : p.num_slabs) * H;
                uint32_t b;
                for (;;) {
                    int n_pub;
                    b = shared_bound(n_pub);
                    if (!after_warm || n_pub >= lists_with_rows || now_ns() > t_end) break;
                    __nanosleep(128);
                }
                if (b > 0x00800001u) {  // share_need lists have published something above -inf
                    thr_shared = fmaxf(thr_shared, __uint_as_float(srag_f32_from_orderable(b - 2u)));
                    thr = fmaxf(thr_own, thr_shared);
                }
            }
            ptx::mbar_wait(&tail->tmem_full[buf], use & 1);
            ptx::tc_fence_after();
            const uint32_t t_addr = lane_addr + (uint32_t)buf * BLOCK_N + (uint32_t)col0;
            const uint32_t row_base = (uint32_t)c * BLOCK_N + (uint32_t)col0;
            const uint32_t sc_addr = ptx::smem_u32(sc);
            // hand the accumulator back to the MMA issuer (leader CTA's barrier) as soon as this warp's
            // last TMEM read has landed in registers
            auto release = [&]() {
                ptx::tc_fence_before();
                __syncwarp();
                if (lane == 0) {
                    if constexpr (G == 2) ptx::mbar_arrive_cluster(&tail->tmem_empty[buf], 0);
                    else ptx::mbar_arrive(&tail->tmem_empty[buf]);
                }
            };
            // Filter per 32-column chunk: m = max of the RAW accumulators (FMNMX3 tree only). Every scaled
            // score of the chunk is <= max(m * smax, m * smin) (IEEE multiply is monotone), so the exact
            // per-element scaling + append scan only runs when that bound beats this thread's threshold.
            auto process = [&](uint32_t (&v)[32], int ch, float s_hi, float s_lo) {
                float t0 = __uint_as_float(v[0]), t1 = __uint_as_float(v[1]);
                float t2 = __uint_as_float(v[2]), t3 = __uint_as_float(v[3]);
#pragma unroll
                for (int j = 4; j < 32; j += 4) {
                    t0 = fmaxf(t0, __uint_as_float(v[j]));
                    t1 = fmaxf(t1, __uint_as_float(v[j + 1]));
                    t2 = fmaxf(t2, __uint_as_float(v[j + 2]));
                    t3 = fmaxf(t3, __uint_as_float(v[j + 3]));
                }
                const float m = fmaxf(fmaxf(t0, t1), fmaxf(t2, t3));
                const float bound = fmaxf(m * s_hi, m * s_lo);
                if (!__any_sync(0xFFFFFFFFu, bound > thr)) return;
                // exact scores of the chunk (FMUL2) and the maxima of its four 8-column groups
                float sv[32], g[4];
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    float m0, m1, m2, m3, m4, m5, m6, m7;
                    ptx::lds_v4(sc_addr + (uint32_t)(ch * 32 + q * 8) * 4u, m0, m1, m2, m3);
                    ptx::lds_v4(sc_addr + (uint32_t)(ch * 32 + q * 8 + 4) * 4u, m4, m5, m6, m7);
                    ptx::mul2(sv[q * 8 + 0], sv[q * 8 + 1], v[q * 8 + 0], v[q * 8 + 1], m0, m1);
                    ptx::mul2(sv[q * 8 + 2], sv[q * 8 + 3], v[q * 8 + 2], v[q * 8 + 3], m2, m3);
                    ptx::mul2(sv[q * 8 + 4], sv[q * 8 + 5], v[q * 8 + 4], v[q * 8 + 5], m4, m5);
                    ptx::mul2(sv[q * 8 + 6], sv[q * 8 + 7], v[q * 8 + 6], v[q * 8 + 7], m6, m7);
                    // fmaxf ignores NaN (tombstone / padding row); an all-NaN group never passes "> thr"
                    g[q] = fmaxf(fmaxf(fmaxf(sv[q * 8 + 0], sv[q * 8 + 1]), fmaxf(sv[q * 8 + 2], sv[q * 8 + 3])),
                                 fmaxf(fmaxf(sv[q * 8 + 4], sv[q * 8 + 5]), fmaxf(sv[q * 8 + 6], sv[q * 8 + 7])));
                }
                const float mx = fmaxf(fmaxf(g[0], g[1]), fmaxf(g[2], g[3]));
                if (!__any_sync(0xFFFFFFFFu, mx > thr)) return;
                if (mx > thr) {
                    // Rare path, ONE copy of the push code per chunk reader instead of one per column: a bit mask of the
                    // columns that beat the threshold, then a loop over its set bits in ascending row order (the score is
                    // re-read from the registers with a select tree and re-checked: the threshold only rises). It used to be
                    // inlined at each of the 64 unrolled compare sites — 2 K (plain) to 4.7 K (append + filter)
                    // instructions that run a few times per tile and are never in the instruction cache (L0 6 KB, L1.5
                    // 32 KB): ncu on a cold-start-dominated run (k = 100, 1.25 M rows) put 30 % of all stall samples on
                    // `no_inst` at those sites. A loop also lets every lane work on ITS next passing column in the same
                    // iteration, where the unrolled sites serialised a warp over the union of its lanes' columns.
                    const uint32_t r0 = row_base + (uint32_t)ch * 32u;
                    uint32_t mask = 0u;
#pragma unroll
                    for (int e = 0; e < 32; ++e) mask |= (sv[e] > thr ? 1u : 0u) << e;
#pragma unroll 1
                    while (mask) {
                        const int e = __ffs(mask) - 1;
                        mask &= mask - 1u;
                        float v = select32(sv, e);
                        if (v > thr && (!filtered || eligible(r0 + (uint32_t)e))) {
                            if (!WARM || !warm_mode) {
                                if constexpr (WARM) {
                                    if (heap_cnt < p.k) {   // not full yet: append (the threshold stays where it is)
                                        heap[heap_cnt + 3] = srag_pack_key(__float_as_uint(v), r0 + (uint32_t)e);
                                        if (++heap_cnt == p.k) thr_own = heap_build(heap, p.k);
                                    } else {
                                        thr_own = heap_replace_root(heap, p.k, __float_as_uint(v), r0 + (uint32_t)e);
                                    }
                                } else {
                                    thr_own = heap_replace_root(heap, p.k, __float_as_uint(v), r0 + (uint32_t)e);
                                }
                                thr = fmaxf(thr_own, thr_shared);
                            }
#pragma unroll
                            for (int i = 0; i < 8; ++i) {  // keep the 8 best scores sorted (descending)
                                const float hi = fmaxf(topj[i], v);
                                v = fminf(topj[i], v);
                                topj[i] = hi;
                            }
                            if (WARM && warm_mode) {  // warm-up: only the j best matter; the filter is the j-th best so far
                                float tj = topj[0];
#pragma unroll
                                for (int i = 1; i < 8; ++i) tj = (i < p.share_j) ? topj[i] : tj;
                                thr = tj;
                            }
                        }
                    }
                }
            };
            uint32_t va[32], vb[32];
            if (p.debug & 4) {
                release();
            } else {
                ptx::tmem_ld_32x32(t_addr, va);
#pragma unroll 1
                for (int ch = 0; ch < CHUNKS; ch += 2) {
                    const float h0 = __shfl_sync(0xFFFFFFFFu, s_hi_l, ch * LPC);
                    const float l0 = __shfl_sync(0xFFFFFFFFu, s_lo_l, ch * LPC);
                    const float h1 = __shfl_sync(0xFFFFFFFFu, s_hi_l, (ch + 1) * LPC);
                    const float l1 = __shfl_sync(0xFFFFFFFFu, s_lo_l, (ch + 1) * LPC);
                    ptx::tmem_ld_wait_dep(va);
                    ptx::tmem_ld_32x32(t_addr + (ch + 1) * 32, vb);
                    process(va, ch, h0, l0);
                    ptx::tmem_ld_wait_dep(vb);
                    if (ch + 2 < CHUNKS) ptx::tmem_ld_32x32(t_addr + (ch + 2) * 32, va);
                    else release();
                    process(vb, ch + 1, h1, l1);
                }
            }
            if (p.share_j > 0 && valid_q) publish();
        }
        // nothing to flush: every thread's heap already is its slice of `partial`
    }

    // ---- teardown ------------------------------------------------------------------------
    ptx::tc_fence_before();
    if constexpr (G == 2) ptx::cluster_sync_all(); else __syncthreads();
    if (warp == 1) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc<G>(tmem_base, TMEM_COLS);
    }
}


// ---------------------------------------------------------------------------------------
// finalize: merge the S per-slab sorted lists of one query (bitonic sort of sortable keys in
// shared memory, carried top-k across chunks), apply the query multiplier, emit (score, id).
// Shared with K2 (bm25.cu) through srag_launch_finalize(); K4 has its own int64-id merge.
// ---------------------------------------------------------------------------------------
constexpr int FIN_THREADS = 256;
constexpr int FIN_MAX_P = 8192;  // keys sorted per pass (64 KB of shared memory)

__device__ void block_bitonic_desc(uint64_t* a, int P) {
    for (int size = 2; size <= P; size <<= 1) {
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            for (int t = threadIdx.x; t < (P >> 1); t += blockDim.x) {
                const int i = 2 * t - (t & (stride - 1));
                const int j = i + stride;
                const bool desc = (i & size) == 0;
                const uint64_t x = a[i], y = a[j];
                if ((x < y) == desc) { a[i] = y; a[j] = x; }
            }
            __syncthreads();
        }
    }
}

__global__ void __launch_bounds__(FIN_THREADS)
topk_finalize_kernel(const uint64_t* __restrict__ partial, size_t src_stride, int total_max /* lists * k */, int k, int kcap,
                     int P, const float* __restrict__ q_scale, int64_t id_base, float* __restrict__ out_scores,
                     int64_t* __restrict__ out_ids, uint64_t* __restrict__ out_keys, size_t out_keys_stride,
                     const uint32_t* __restrict__ live_count, const uint32_t* __restrict__ live_all, int live_base,
                     uint64_t* __restrict__ out_gkeys /* [n_queries][k] exchange keys (final score, global id), or null */) {
    extern __shared__ uint64_t fin_keys[];
    __shared__ int fin_count;
    const int q = blockIdx.x;
    const uint64_t* src = partial + (size_t)q * src_stride;
    const int chunk = P - kcap;
    bool sorted = false;
    // K2's candidate regions are filled from the front: only the first live_base + live_count[q] keys of a query exist
    // (nothing beyond them was ever written), unless the query was redone into per-block lists (live_all[q]).
    int total = total_max;
    if (live_count && !live_all[q]) {
        const long long t = (long long)live_base + live_count[q];
        total = t < total_max ? (int)t : total_max;
    }
    {
        // Sparse lists (K2 with a sample threshold leaves most block lists empty): compact the non-empty keys
        // in ONE pass; when they all fit, sort one small power of two instead of total/chunk passes.
        for (int i = threadIdx.x; i < P; i += blockDim.x) fin_keys[i] = SRAG_KEY_EMPTY;
        if (threadIdx.x == 0) fin_count = 0;
        __syncthreads();
        auto keep = [&](uint64_t key) {
            if (key != SRAG_KEY_EMPTY) {
                const int slot = atomicAdd(&fin_count, 1);  // order is irrelevant: sorted next
                if (slot < chunk) fin_keys[slot] = key;
            }
        };
        if ((reinterpret_cast<uintptr_t>(src) & 15) == 0) {
            const ulonglong2* src2 = reinterpret_cast<const ulonglong2*>(src);
            const int pairs = total >> 1;
#pragma unroll 8
            for (int i = threadIdx.x; i < pairs; i += FIN_THREADS) {
                const ulonglong2 v = __ldg(src2 + i);
                keep(v.x);
                keep(v.y);
            }
            if ((total & 1) && threadIdx.x == 0) keep(src[total - 1]);
        } else {
#pragma unroll 8
            for (int i = threadIdx.x; i < total; i += FIN_THREADS) keep(__ldg(src + i));
        }
        __syncthreads();
        const int live = fin_count;
        __syncthreads();
        if (live <= chunk) {
            int P2 = 2 * kcap;
            while (P2 < live + kcap) P2 <<= 1;
            block_bitonic_desc(fin_keys, P2);
            sorted = true;
        }
    }
    if (!sorted) {
        for (int i = threadIdx.x; i < kcap; i += blockDim.x) fin_keys[i] = SRAG_KEY_EMPTY;
        for (int off = 0; off < total; off += chunk) {
            for (int i = threadIdx.x; i < chunk; i += blockDim.x)
                fin_keys[kcap + i] = (off + i < total) ? src[off + i] : SRAG_KEY_EMPTY;
            __syncthreads();
            block_bitonic_desc(fin_keys, P);
            for (int i = k + threadIdx.x; i < kcap; i += blockDim.x) fin_keys[i] = SRAG_KEY_EMPTY;
            __syncthreads();
        }
    }
    if (out_keys)  // the merged list as sortable keys (K2 seeds its second phase with the sample's top-k)
        for (int j = threadIdx.x; j < k; j += blockDim.x) out_keys[(size_t)q * out_keys_stride + j] = fin_keys[j];
    const float qs = q_scale ? q_scale[q] : 1.0f;
    for (int j = threadIdx.x; j < k; j += blockDim.x) {
        const uint64_t key = fin_keys[j];
        float s = -INFINITY;
        int64_t id = -1;
        if (key != SRAG_KEY_EMPTY) {
            s = __uint_as_float(srag_key_score_bits(key)) * qs;
            id = id_base + (int64_t)srag_key_idx(key);
        }
        if (out_scores) out_scores[(size_t)q * k + j] = s;
        if (out_ids) out_ids[(size_t)q * k + j] = id;
        // the cross-shard exchange format: ONE sortable 64-bit key per candidate, carrying the final
        // score bits and the GLOBAL row, written straight into the all-gather send buffer
        if (out_gkeys)
            out_gkeys[(size_t)q * k + j] = id >= 0 ? srag_pack_key(__float_as_uint(s), (uint32_t)id) : SRAG_KEY_EMPTY;
    }
}

// ---------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

PFN_encodeTiled get_encode_fn() {
    static PFN_encodeTiled fn = nullptr;
    if (fn) return fn;
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) != cudaSuccess ||
        qres != cudaDriverEntryPointSuccess)
        return nullptr;
    fn = reinterpret_cast<PFN_encodeTiled>(ptr);
    return fn;
}

// rows x dim bf16 row-major -> box of (64 columns = 128 bytes) x box_rows, 128-byte swizzle
int make_tmap(CUtensorMap* tm, const void* base, int64_t rows, int dim, int box_rows, bool promote) {
    PFN_encodeTiled enc = get_encode_fn();
    if (!enc) {
        srag_set_error("cuTensorMapEncodeTiled entry point not available");
        return SRAG_E_UNSUPPORTED;
    }
    cuuint64_t gdim[2] = {(cuuint64_t)dim, (cuuint64_t)rows};
    cuuint64_t gstride[1] = {(cuuint64_t)dim * 2};
    cuuint32_t box[2] = {(cuuint32_t)BLOCK_K, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), gdim, gstride, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                     promote ? CU_TENSOR_MAP_L2_PROMOTION_L2_256B : CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        srag_set_error("cuTensorMapEncodeTiled failed with CUresult %d (rows=%lld dim=%d box_rows=%d)", (int)r,
                       (long long)rows, dim, box_rows);
        return SRAG_E_CUDA;
    }
    return SRAG_OK;
}

struct Plan {
    int cta_group;  // 1 or 2
    int grid;       // CTAs launched
    int num_q_tiles;
    int num_slabs;
    size_t partial_bytes;   // padded to 256
    size_t gthr_bytes;      // shared per-query thresholds, zeroed together with `partial`
    size_t progress_bytes;  // pacing counters, zeroed together with `partial`
    int window;             // 0 = pacing off
};

int device_sm_count(int* sms) {
    int major = 0;
    const int rc = srag_current_device_props(sms, &major, nullptr);
    if (rc) return rc;
    if (major != 10) {
        srag_set_error("device compute capability %d.x is not sm_100", major);
        return SRAG_E_UNSUPPORTED;
    }
    // SRAG_TOPK_SMS: scan on fewer SMs (an even number), leaving the rest to a kernel on another stream (§6b: K1 and K2
    // of the hybrid pipeline side by side). Read per call; sizes the plan AND the workspace, so set it before both.
    if (const char* e = getenv("SRAG_TOPK_SMS")) {
        const int lim = atoi(e) & ~1;
        if (lim >= 2 && lim < *sms) *sms = lim;
    }
    return SRAG_OK;
}

int env_int(const char* name, int dflt) {
    const char* e = getenv(name);
    if (e && ((e[0] >= '0' && e[0] <= '9') || e[0] == '-')) return atoi(e);
    return dflt;
}

int forced_cta_group() {
    const int g = env_int("SRAG_TOPK_CTA_GROUP", 0);
    return (g == 1 || g == 2) ? g : 0;
}

// SRAG_TOPK_WINDOW: tiles the leader unit of a slab may run ahead of its followers (0 = no pacing)
int pacing_window() { return env_int("SRAG_TOPK_WINDOW", 3); }

// SRAG_TOPK_RQ=1 selects the resident-query operand plan (dim <= 1024); default is the streaming plan
bool use_resident_queries(int dim) { return dim <= RQ_MAX_DIM && env_int("SRAG_TOPK_RQ", 0) != 0; }

// SRAG_TOPK_EPI_WARPS: 4 or 8 epilogue warps per CTA
int epilogue_warps() { return env_int("SRAG_TOPK_EPI_WARPS", 4) == 8 ? 8 : 4; }

int make_plan(int sms, int n_queries, int k, Plan* pl) {
    int g = forced_cta_group();
    if (g == 0) g = (n_queries > BLOCK_M) ? 2 : 1;
    const int units = sms / g;
    const int tq = (n_queries + BLOCK_M * g - 1) / (BLOCK_M * g);
    if (tq > units) {
        srag_set_error("n_queries=%d needs %d query tiles but the device has %d units; split the batch", n_queries,
                       tq, units);
        return SRAG_E_INVALID;
    }
    pl->cta_group = g;
    pl->grid = units * g;
    pl->num_q_tiles = tq;
    pl->num_slabs = units / tq;
    pl->partial_bytes = (((size_t)n_queries * pl->num_slabs * 2 * heap_stride(k) * sizeof(uint64_t)) + 255) & ~(size_t)255;
    pl->gthr_bytes = (((size_t)n_queries * pl->num_slabs * 2 * sizeof(uint32_t)) + 255) & ~(size_t)255;
    pl->progress_bytes = (size_t)pl->num_slabs * 2 * PROG_SLOTS * sizeof(int);
    pl->window = pacing_window();
    return SRAG_OK;
}

// Optional device-side timing of the two K1 kernels (bench.py's roofline leg): events are
// recorded on the launching stream, never synchronised here.
constexpr int PROF_RING = 256;
struct ProfState {
    bool on = false;
    int n = 0;
    int dev = -1;  // events belong to the device that was current at srag_prof_enable(1); other devices are not timed
    cudaEvent_t ev[PROF_RING][3];
    bool created = false;
} g_prof;
std::mutex g_prof_mu;  // the ABI may be entered from several host threads

// Claims the next event triple when profiling is on for the current device; -1 otherwise.
int prof_claim() {
    std::lock_guard<std::mutex> lk(g_prof_mu);
    if (!g_prof.on || g_prof.n >= PROF_RING) return -1;
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev != g_prof.dev) return -1;
    return g_prof.n++;
}

template <int G, bool RQ, int EW, bool FILT = false, bool WARM = false>
int launch_main(const Plan& pl, const CUtensorMap& tq, const CUtensorMap& tc, const TopkParams& prm,
                cudaStream_t st) {
    {
        const int rc = srag_func_smem_optin(reinterpret_cast<const void*>(&topk_gemm_kernel<G, RQ, EW, FILT, WARM>),
                                            (int)smem_bytes<G, RQ>());
        if (rc) return rc;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(pl.grid);
    cfg.blockDim = dim3(64 + 32 * EW);
    cfg.dynamicSmemBytes = smem_bytes<G, RQ>();
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = G;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    SRAG_CUDA(cudaLaunchKernelEx(&cfg, topk_gemm_kernel<G, RQ, EW, FILT, WARM>, tq, tc, prm));
    return SRAG_OK;
}

}  // namespace

extern "C" int srag_topk_workspace_bytes(int64_t n_rows, int n_queries, int k, size_t* bytes) {
    SRAG_CHECK_ARG(bytes != nullptr, "bytes is null");
    SRAG_CHECK_ARG(n_queries >= 1 && k >= 1 && k <= SRAG_MAX_TOPK && n_rows >= 0, "bad sizes n_queries=%d k=%d",
                   n_queries, k);
    int sms = 0;
    int rc = device_sm_count(&sms);
    if (rc) return rc;
    Plan pl;
    rc = make_plan(sms, n_queries, k, &pl);
    if (rc) return rc;
    *bytes = pl.partial_bytes + pl.gthr_bytes + pl.progress_bytes + 256;
    return SRAG_OK;
}

// Queries padded with zero rows to whole q-tiles: a TMA box that is mostly out of bounds (B = 1..32 of a
// 128-row tile) loads far slower than one over real zero rows (4.8 ms vs 2.9 ms per 10 M x 1024 scan at B = 1).
static size_t padded_query_rows(const Plan& pl, int n_queries) {
    const size_t tile = (size_t)BLOCK_M * pl.cta_group;
    return ((size_t)n_queries + tile - 1) / tile * tile;
}

extern "C" int srag_topk_workspace_bytes_ex(int64_t n_rows, int dim, int n_queries, int k, size_t* bytes) {
    int rc = srag_topk_workspace_bytes(n_rows, n_queries, k, bytes);
    if (rc) return rc;
    SRAG_CHECK_ARG(dim >= 8 && dim % 8 == 0, "dim=%d must be a positive multiple of 8", dim);
    int sms = 0;
    rc = device_sm_count(&sms);
    if (rc) return rc;
    Plan pl;
    rc = make_plan(sms, n_queries, k, &pl);
    if (rc) return rc;
    const size_t rows = padded_query_rows(pl, n_queries);
    if (rows != (size_t)n_queries) *bytes += rows * dim * sizeof(__nv_bfloat16) + 256;
    return SRAG_OK;
}

static int topk_search_impl(const void* corpus_bf16, const float* row_scale, int64_t n_rows, int dim,
                            const void* queries_bf16, const float* q_scale, int n_queries, int k, int64_t id_base,
                            const uint32_t* row_tags, const int32_t* row_group, const uint32_t* q_tag_mask,
                            const int32_t* q_group, float* out_scores, int64_t* out_ids, uint64_t* out_gkeys,
                            void* workspace, size_t workspace_bytes, void* stream) {
    SRAG_CHECK_ARG(queries_bf16 && workspace && ((out_scores && out_ids) || out_gkeys), "null pointer argument");
    SRAG_CHECK_ARG(!out_gkeys || (id_base >= 0 && id_base + n_rows <= (int64_t)0xFFFFFFFEll),
                   "exchange keys hold 32-bit global rows: id_base + n_rows = %lld does not fit", (long long)(id_base + n_rows));
    SRAG_CHECK_ARG((row_tags == nullptr) == (q_tag_mask == nullptr) && (row_group == nullptr) == (q_group == nullptr),
                   "row_tags / q_tag_mask and row_group / q_group come in pairs");
    SRAG_CHECK_ARG(n_rows == 0 || (corpus_bf16 && row_scale), "null corpus / row_scale");
    SRAG_CHECK_ARG(n_queries >= 1 && k >= 1 && k <= SRAG_MAX_TOPK, "bad sizes n_queries=%d k=%d", n_queries, k);
    SRAG_CHECK_ARG(dim >= 8 && dim % 8 == 0, "dim=%d must be a positive multiple of 8", dim);
    SRAG_CHECK_ARG(n_rows >= 0 && n_rows < (int64_t)0xFFFFFF00ll, "n_rows out of range");
    SRAG_CHECK_ARG(((uintptr_t)corpus_bf16 & 15) == 0 && ((uintptr_t)queries_bf16 & 15) == 0 &&
                       ((uintptr_t)row_scale & 15) == 0,
                   "corpus / queries / row_scale must be 16-byte aligned");
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    int sms = 0;
    int rc = device_sm_count(&sms);
    if (rc) return rc;
    Plan pl;
    rc = make_plan(sms, n_queries, k, &pl);
    if (rc) return rc;
    const size_t need = pl.partial_bytes + pl.gthr_bytes + pl.progress_bytes;
    if (workspace_bytes < need) {
        srag_set_error("workspace too small: %zu < %zu", workspace_bytes, need);
        return SRAG_E_WORKSPACE;
    }
    uint8_t* ws = reinterpret_cast<uint8_t*>(workspace);
    SRAG_CHECK_ARG(((uintptr_t)ws & 15) == 0, "workspace must be 16-byte aligned");
    const bool rq = use_resident_queries(dim);
    const int ew = epilogue_warps();
    TopkParams prm;
    prm.row_scale = row_scale;
    prm.queries = reinterpret_cast<const __nv_bfloat16*>(queries_bf16);
    prm.n_rows = n_rows;
    prm.dim = dim;
    prm.n_queries = n_queries;
    prm.k = k;
    prm.num_q_tiles = pl.num_q_tiles;
    prm.num_slabs = pl.num_slabs;
    prm.partial = reinterpret_cast<uint64_t*>(ws);
    prm.progress = pl.window > 0 ? reinterpret_cast<int*>(ws + pl.partial_bytes + pl.gthr_bytes) : nullptr;
    prm.gthr = reinterpret_cast<uint32_t*>(ws + pl.partial_bytes);
    {
        // sharing: every list tracks its j best scores (j <= 8 registers) and the bound is the ceil(k / j)-th largest
        // published value (<= 16 registers, <= lists): the smallest such j
        const int lists = pl.num_slabs * (ew / 4);
        int j = (k + lists - 1) / lists;
        if (j < (k + 15) / 16) j = (k + 15) / 16;
        // Polling many lists costs ~33 instructions per list and refresh. With one q-tile (148 lists) the bound is
        // therefore re-read on the first two tiles and then every 64th only — the prescan's bound does most of the work:
        // B = 256, k = 10, 10 M rows: 4.70 -> 4.02 ms (0.80 of the roofline), B = 128: 3.38 -> 3.08, B = 32: 3.27 -> 3.12;
        // a handful of queries (B = 8: 2.96 -> 3.08 ms, HBM-bound at 0.98 already) gain nothing and stay without.
        const int share_env = env_int("SRAG_TOPK_SHARE", 1);   // 0 off, 1 default rule, 2 wherever possible (A/B runs)
        const bool worth = lists <= 80 || (lists <= 160 && (k >= 32 || n_queries >= 24)) || share_env == 2;
        prm.share_j = (share_env != 0 && lists >= 2 && worth && j <= 8) ? j : 0;
        prm.share_period = env_int("SRAG_TOPK_SHARE_PERIOD", lists <= 80 ? 4 : k >= 32 ? 16 : 64);   // (large k: the bound moves longer)
        if (prm.share_period < 1) prm.share_period = 1;
        prm.share_early = env_int("SRAG_TOPK_SHARE_EARLY", lists <= 80 ? 8 : 2);
        prm.share_need = prm.share_j > 0 ? (k + prm.share_j - 1) / prm.share_j : 0;
        prm.prescan = prm.share_j > 0 && env_int("SRAG_TOPK_PRESCAN", 1) != 0;
    }
    prm.window = pl.window;
    // large k (a push is a deep sift): the instantiation with append-until-full lists, plus `warm` warm-up tiles (see the
    // epilogue; SRAG_TOPK_WARM=0 leaves the cold start to the tile-0 prescan alone)
    const bool warm_inst = prm.share_j > 0 && k >= env_int("SRAG_TOPK_WARM_MIN_K", 32) && !rq && ew == 4;
    prm.warm = warm_inst ? env_int("SRAG_TOPK_WARM", 0) : 0;
    if (prm.warm < 0) prm.warm = 0;
    prm.debug = env_int("SRAG_TOPK_DEBUG", 0);
    prm.row_tags = n_rows > 0 ? row_tags : nullptr;
    prm.row_group = n_rows > 0 ? row_group : nullptr;
    prm.q_tag_mask = q_tag_mask;
    prm.q_group = q_group;
    SRAG_CUDA(cudaMemsetAsync(prm.partial, 0, pl.partial_bytes + pl.gthr_bytes + pl.progress_bytes, st));
    const int prof_slot = prof_claim();
    const bool prof = prof_slot >= 0;
    if (prof) SRAG_CUDA(cudaEventRecord(g_prof.ev[prof_slot][0], st));
    if (n_rows > 0) {
        CUtensorMap tq, tc;
        // zero-padded copy of the queries when the batch is not whole q-tiles and the caller sized the
        // workspace with srag_topk_workspace_bytes_ex (otherwise the out-of-bounds box path is used)
        const size_t q_rows = padded_query_rows(pl, n_queries);
        const size_t q_off = (need + 255) & ~(size_t)255;
        const size_t row_bytes = (size_t)dim * sizeof(__nv_bfloat16);
        if (q_rows != (size_t)n_queries && workspace_bytes >= q_off + q_rows * row_bytes) {
            uint8_t* qpad = ws + q_off;
            SRAG_CUDA(cudaMemcpyAsync(qpad, queries_bf16, n_queries * row_bytes, cudaMemcpyDeviceToDevice, st));
            SRAG_CUDA(cudaMemsetAsync(qpad + n_queries * row_bytes, 0, (q_rows - n_queries) * row_bytes, st));
            prm.queries = reinterpret_cast<const __nv_bfloat16*>(qpad);
            rc = make_tmap(&tq, qpad, (int64_t)q_rows, dim, BLOCK_M, false);
        } else {
            rc = make_tmap(&tq, queries_bf16, n_queries, dim, BLOCK_M, false);
        }
        if (rc) return rc;
        const int block_n = rq ? 128 : 256;
        rc = make_tmap(&tc, corpus_bf16, n_rows, dim, block_n / pl.cta_group, true);
        if (rc) return rc;
        const int sel = (pl.cta_group == 2 ? 4 : 0) | (rq ? 2 : 0) | (ew == 8 ? 1 : 0);
        if (prm.row_tags || prm.row_group) {
            // the filtered kernel exists for the default plans only (streaming queries, 4 epilogue warps)
            if (rq || ew != 4) {
                srag_set_error("the metadata filter is not available with SRAG_TOPK_RQ / SRAG_TOPK_EPI_WARPS overrides");
                return SRAG_E_INVALID;
            }
            if (warm_inst)
                rc = pl.cta_group == 2 ? launch_main<2, false, 4, true, true>(pl, tq, tc, prm, st)
                                       : launch_main<1, false, 4, true, true>(pl, tq, tc, prm, st);
            else
                rc = pl.cta_group == 2 ? launch_main<2, false, 4, true>(pl, tq, tc, prm, st)
                                       : launch_main<1, false, 4, true>(pl, tq, tc, prm, st);
        } else
        switch (sel) {
            case 0: rc = warm_inst ? launch_main<1, false, 4, false, true>(pl, tq, tc, prm, st)
                                      : launch_main<1, false, 4>(pl, tq, tc, prm, st); break;
            case 1: rc = launch_main<1, false, 8>(pl, tq, tc, prm, st); break;
            case 2: rc = launch_main<1, true, 4>(pl, tq, tc, prm, st); break;
            case 3: rc = launch_main<1, true, 8>(pl, tq, tc, prm, st); break;
            case 4: rc = warm_inst ? launch_main<2, false, 4, false, true>(pl, tq, tc, prm, st)
                                      : launch_main<2, false, 4>(pl, tq, tc, prm, st); break;
            case 5: rc = launch_main<2, false, 8>(pl, tq, tc, prm, st); break;
            case 6: rc = launch_main<2, true, 4>(pl, tq, tc, prm, st); break;
            default: rc = launch_main<2, true, 8>(pl, tq, tc, prm, st); break;
        }
        if (rc) return rc;
    }
    if (prof) SRAG_CUDA(cudaEventRecord(g_prof.ev[prof_slot][1], st));
    {
        const int total = pl.num_slabs * (ew / 4) * heap_stride(k);
        rc = srag_launch_finalize_gkeys(prm.partial, (size_t)total, n_queries, total, k, q_scale, id_base, out_scores, out_ids,
                                        nullptr, 0, nullptr, nullptr, 0, out_gkeys, st);
    }
    if (rc) return rc;
    if (prof) SRAG_CUDA(cudaEventRecord(g_prof.ev[prof_slot][2], st));
    return SRAG_OK;
}

extern "C" int srag_topk_search(const void* corpus_bf16, const float* row_scale, int64_t n_rows, int dim,
                                const void* queries_bf16, const float* q_scale, int n_queries, int k,
                                int64_t id_base, float* out_scores, int64_t* out_ids, void* workspace,
                                size_t workspace_bytes, void* stream) {
    return topk_search_impl(corpus_bf16, row_scale, n_rows, dim, queries_bf16, q_scale, n_queries, k, id_base, nullptr,
                            nullptr, nullptr, nullptr, out_scores, out_ids, nullptr, workspace, workspace_bytes, stream);
}

extern "C" int srag_topk_search_keys(const void* corpus_bf16, const float* row_scale, int64_t n_rows, int dim,
                                     const void* queries_bf16, const float* q_scale, int n_queries, int k,
                                     int64_t id_base, uint64_t* out_keys, float* out_scores, int64_t* out_ids,
                                     void* workspace, size_t workspace_bytes, void* stream) {
    SRAG_CHECK_ARG(out_keys != nullptr, "out_keys is null");
    SRAG_CHECK_ARG((out_scores == nullptr) == (out_ids == nullptr), "out_scores / out_ids come together (or both NULL)");
    return topk_search_impl(corpus_bf16, row_scale, n_rows, dim, queries_bf16, q_scale, n_queries, k, id_base, nullptr,
                            nullptr, nullptr, nullptr, out_scores, out_ids, out_keys, workspace, workspace_bytes, stream);
}

// ---------------------------------------------------------------------------------------
// srag_topk_search_static: the same scan with its scratch in a file-scope device array, for callers that may not
// allocate (include/srag_b200.h). In host code `g_topk_static_scratch` is only a host-side stand-in for the symbol: its
// DEVICE address comes from cudaGetSymbolAddress, looked up once per device with that device current.
// ---------------------------------------------------------------------------------------
static __device__ __align__(256) unsigned char g_topk_static_scratch[SRAG_TOPK_STATIC_SCRATCH_BYTES];

static int topk_static_scratch_ptr(void** out) {
    constexpr int MAX_DEVICES = 64;
    static std::mutex mu;
    static void* ptrs[MAX_DEVICES] = {};
    int dev = 0;
    SRAG_CUDA(cudaGetDevice(&dev));
    SRAG_CHECK_ARG(dev >= 0 && dev < MAX_DEVICES, "device ordinal %d out of range", dev);
    std::lock_guard<std::mutex> lk(mu);
    if (ptrs[dev] == nullptr) {
        void* p = nullptr;
        SRAG_CUDA(cudaGetSymbolAddress(&p, g_topk_static_scratch));
        SRAG_CHECK_ARG(p != nullptr && ((uintptr_t)p & 255) == 0, "static scratch has no usable device address");
        ptrs[dev] = p;
    }
    *out = ptrs[dev];
    return SRAG_OK;
}

extern "C" int srag_topk_search_static(const void* corpus_bf16, const float* row_scale, int64_t n_rows, int dim,
                                       const void* queries_bf16, const float* q_scale, int n_queries, int k,
                                       int64_t id_base, float* out_scores, int64_t* out_ids, void* stream) {
    SRAG_CHECK_ARG(out_scores && out_ids, "null pointer argument");
    size_t need = 0;
    int rc = srag_topk_workspace_bytes(n_rows, n_queries, k, &need);   // validates n_queries / k / n_rows as well
    if (rc) return rc;
    if (need > (size_t)SRAG_TOPK_STATIC_SCRATCH_BYTES) {
        srag_set_error("srag_topk_search_static: n_queries=%d k=%d needs %zu bytes of scratch, the static array holds %zu "
                       "(SRAG_TOPK_STATIC_SCRATCH_BYTES); use srag_topk_search with a caller workspace",
                       n_queries, k, need, (size_t)SRAG_TOPK_STATIC_SCRATCH_BYTES);
        return SRAG_E_WORKSPACE;
    }
    void* scratch = nullptr;
    rc = topk_static_scratch_ptr(&scratch);
    if (rc) return rc;
    // the impl clears (cudaMemsetAsync on `stream`) the heaps, thresholds and counters it uses before the scan, so the last
    // call's contents never leak into this one; the padded query copy is used when it still fits behind them
    return topk_search_impl(corpus_bf16, row_scale, n_rows, dim, queries_bf16, q_scale, n_queries, k, id_base, nullptr,
                            nullptr, nullptr, nullptr, out_scores, out_ids, nullptr, scratch,
                            (size_t)SRAG_TOPK_STATIC_SCRATCH_BYTES, stream);
}

extern "C" int srag_topk_search_filtered(const void* corpus_bf16, const float* row_scale, int64_t n_rows, int dim,
                                         const void* queries_bf16, const float* q_scale, int n_queries, int k,
                                         int64_t id_base, const uint32_t* row_tags, const int32_t* row_group,
                                         const uint32_t* q_tag_mask, const int32_t* q_group, float* out_scores,
                                         int64_t* out_ids, void* workspace, size_t workspace_bytes, void* stream) {
    return topk_search_impl(corpus_bf16, row_scale, n_rows, dim, queries_bf16, q_scale, n_queries, k, id_base, row_tags,
                            row_group, q_tag_mask, q_group, out_scores, out_ids, nullptr, workspace, workspace_bytes, stream);
}

extern "C" int srag_prof_enable(int on) {
    std::lock_guard<std::mutex> lk(g_prof_mu);
    int dev = 0;
    SRAG_CUDA(cudaGetDevice(&dev));
    if (on && g_prof.created && g_prof.dev != dev) {  // re-home the events on the device now current
        for (int i = 0; i < PROF_RING; ++i)
            for (int j = 0; j < 3; ++j) cudaEventDestroy(g_prof.ev[i][j]);
        g_prof.created = false;
    }
    if (on && !g_prof.created) {
        for (int i = 0; i < PROF_RING; ++i)
            for (int j = 0; j < 3; ++j) SRAG_CUDA(cudaEventCreate(&g_prof.ev[i][j]));
        g_prof.created = true;
        g_prof.dev = dev;
    }
    g_prof.on = on != 0;
    g_prof.n = 0;
    return SRAG_OK;
}

extern "C" int srag_prof_read(double* main_ms_sum, double* finalize_ms_sum, int* n_calls) {
    std::lock_guard<std::mutex> lk(g_prof_mu);
    double a = 0, b = 0;
    for (int i = 0; i < g_prof.n; ++i) {
        float x = 0, y = 0;
        SRAG_CUDA(cudaEventSynchronize(g_prof.ev[i][2]));
        SRAG_CUDA(cudaEventElapsedTime(&x, g_prof.ev[i][0], g_prof.ev[i][1]));
        SRAG_CUDA(cudaEventElapsedTime(&y, g_prof.ev[i][1], g_prof.ev[i][2]));
        a += x;
        b += y;
    }
    if (main_ms_sum) *main_ms_sum = a;
    if (finalize_ms_sum) *finalize_ms_sum = b;
    if (n_calls) *n_calls = g_prof.n;
    return SRAG_OK;
}

// Shared with K2 (bm25.cu): per-query lists of sortable keys -> sorted top-k (score, id).
int srag_launch_finalize_live(const uint64_t* partial, size_t src_stride, int n_queries, int total, int k,
                              const float* q_scale, int64_t id_base, float* out_scores, int64_t* out_ids,
                              uint64_t* out_keys, size_t out_keys_stride, const uint32_t* live_count,
                              const uint32_t* live_all, int live_base, cudaStream_t st) {
    return srag_launch_finalize_gkeys(partial, src_stride, n_queries, total, k, q_scale, id_base, out_scores, out_ids,
                                      out_keys, out_keys_stride, live_count, live_all, live_base, nullptr, st);
}

int srag_launch_finalize_gkeys(const uint64_t* partial, size_t src_stride, int n_queries, int total, int k,
                               const float* q_scale, int64_t id_base, float* out_scores, int64_t* out_ids,
                               uint64_t* out_keys, size_t out_keys_stride, const uint32_t* live_count,
                               const uint32_t* live_all, int live_base, uint64_t* out_gkeys, cudaStream_t st) {
    const int kcap = (k <= 128) ? 128 : 256;
    int P = 2 * kcap;
    while (P < total + kcap && P < FIN_MAX_P) P <<= 1;
    {
        const int rc = srag_func_smem_optin(reinterpret_cast<const void*>(&topk_finalize_kernel),
                                            FIN_MAX_P * (int)sizeof(uint64_t));
        if (rc) return rc;
    }
    topk_finalize_kernel<<<n_queries, FIN_THREADS, (size_t)P * sizeof(uint64_t), st>>>(
        partial, src_stride, total, k, kcap, P, q_scale, id_base, out_scores, out_ids, out_keys, out_keys_stride, live_count,
        live_all, live_base, out_gkeys);
    SRAG_CUDA(cudaGetLastError());
    return SRAG_OK;
}

int srag_launch_finalize_ex(const uint64_t* partial, size_t src_stride, int n_queries, int total, int k,
                            const float* q_scale, int64_t id_base, float* out_scores, int64_t* out_ids,
                            uint64_t* out_keys, size_t out_keys_stride, cudaStream_t st) {
    return srag_launch_finalize_live(partial, src_stride, n_queries, total, k, q_scale, id_base, out_scores, out_ids, out_keys,
                                     out_keys_stride, nullptr, nullptr, 0, st);
}

int srag_launch_finalize(const uint64_t* partial, int n_queries, int total, int k, const float* q_scale,
                         int64_t id_base, float* out_scores, int64_t* out_ids, cudaStream_t st) {
    return srag_launch_finalize_ex(partial, (size_t)total, n_queries, total, k, q_scale, id_base, out_scores, out_ids,
                                   nullptr, 0, st);
}
