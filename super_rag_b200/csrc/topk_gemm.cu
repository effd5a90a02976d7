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
//               kernel, so one epilogue thread == one query (TMEM lane == query)
//   slab      = u / T_q; the unit scans corpus tiles  slab, slab + S, slab + 2S, ...
//               (256 rows each), ids ascending, so "score > threshold" is the exact tie rule
//   warps     0: TMA producer   1: MMA issuer + TMEM owner   2..5: epilogue (one TMEM lane
//               quadrant each)
//   pipelines smem full/empty ring (TMA <-> MMA), TMEM full/empty x2 (MMA <-> epilogue)
#include <cuda.h>

#include "common.cuh"

namespace {

constexpr int BLOCK_M = 128;  // queries per CTA  (TMEM lanes)
constexpr int BLOCK_N = 256;  // corpus rows per tile (TMEM fp32 columns per accumulator)
constexpr int BLOCK_K = 64;   // bf16 per 128-byte swizzle row
constexpr int UMMA_K = 16;
constexpr int CAND_CAP = 256;  // candidate entries per (CTA, query)
constexpr int NUM_THREADS = 192;
constexpr int NUM_EPI_WARPS = 4;
constexpr int TMEM_COLS = 512;  // two 256-column accumulators

template <int G>
struct Cfg {
    static constexpr int B_ROWS = BLOCK_N / G;  // corpus rows THIS CTA stages per k-block
    static constexpr int A_BYTES = BLOCK_M * BLOCK_K * 2;
    static constexpr int B_BYTES = B_ROWS * BLOCK_K * 2;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int STAGES = (G == 1) ? 4 : 6;  // 192 KB of operand staging either way
    static constexpr int RING_BYTES = STAGES * STAGE_BYTES;
};

struct __align__(16) SmemTail {
    uint64_t full[8];
    uint64_t empty[8];
    uint64_t tmem_full[2];
    uint64_t tmem_empty[2];
    uint32_t tmem_base;
    uint32_t pad;
    float thr_xchg[NUM_EPI_WARPS];
    alignas(16) float scale[NUM_EPI_WARPS][BLOCK_N];     // per-warp copy of the tile's row multipliers
    alignas(16) uint64_t keys[NUM_EPI_WARPS][CAND_CAP];  // per-warp compaction scratch
};

template <int G>
constexpr size_t smem_bytes() {
    return 1024 /*alignment slack*/ + Cfg<G>::RING_BYTES + sizeof(SmemTail);
}

struct TopkParams {
    const float* row_scale;  // [n_rows]
    int64_t n_rows;
    int dim;
    int n_queries;
    int k;
    int num_q_tiles;  // T_q
    int num_slabs;    // S
    int trigger;      // compaction trigger (entries)
    uint2* cand;      // [grid * 128][CAND_CAP]  (score bits, local row)
    uint64_t* partial;  // [n_queries][S][k] sortable keys, pre-zeroed (= empty)
};

// ---------------------------------------------------------------------------------------
// Warp-cooperative compaction of lane `src`'s candidate buffer: keep the k best under
// (score desc, row asc), sorted. Rank-by-counting: keys are unique (rows are unique), so
// ranks are a permutation. Cost ~ n * ceil(n/32) comparisons per lane.
// If out_keys != nullptr the sorted keys go there (final flush), else the sorted raw pairs
// go back to the front of the buffer. Returns the new threshold for lane src (all lanes
// return the same value; only src uses it) and the new count through *new_cnt.
// ---------------------------------------------------------------------------------------
__device__ __noinline__ float warp_compact(uint2* buf, int n, int k, float old_thr, uint64_t* scratch,
                                           float* thr_xchg, uint64_t* out_keys, int* new_cnt) {
    const int lane = ptx::lane_id();
    uint64_t key[CAND_CAP / 32];
    uint2 raw[CAND_CAP / 32];
#pragma unroll
    for (int i = 0; i < CAND_CAP / 32; ++i) {
        int e = lane + 32 * i;
        key[i] = SRAG_KEY_EMPTY;
        if (e < n) {
            raw[i] = __ldcv(buf + e);  // written by another lane of this warp: bypass L1
            key[i] = srag_pack_key(raw[i].x, raw[i].y);
            scratch[e] = key[i];
        }
    }
    __syncwarp();
    int rank[CAND_CAP / 32];
#pragma unroll
    for (int i = 0; i < CAND_CAP / 32; ++i) rank[i] = 0;
    const int groups = (n + 31) >> 5;
    for (int j = 0; j < n; ++j) {
        uint64_t kj = scratch[j];
#pragma unroll
        for (int i = 0; i < CAND_CAP / 32; ++i)
            if (i < groups) rank[i] += (kj > key[i]) ? 1 : 0;
    }
    __syncwarp();
#pragma unroll
    for (int i = 0; i < CAND_CAP / 32; ++i) {
        int e = lane + 32 * i;
        if (e < n && rank[i] < k) {
            if (out_keys != nullptr)
                out_keys[rank[i]] = key[i];
            else
                buf[rank[i]] = raw[i];
            if (rank[i] == k - 1) *thr_xchg = __uint_as_float(raw[i].x);
        }
    }
    __syncwarp();
    float thr = old_thr;
    if (n >= k) thr = *(volatile float*)thr_xchg;
    *new_cnt = n < k ? n : k;
    __syncwarp();
    return thr;
}


// ---------------------------------------------------------------------------------------
// main kernel
// ---------------------------------------------------------------------------------------
template <int G>
__global__ void __launch_bounds__(NUM_THREADS, 1)
topk_gemm_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_c,
                 const TopkParams p) {
    using C = Cfg<G>;
    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const uint32_t cta_rank = (G == 2) ? ptx::cluster_ctarank() : 0u;
    const int unit = (G == 2) ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
    const int q_tile = unit % p.num_q_tiles;
    const int slab = unit / p.num_q_tiles;
    if (slab >= p.num_slabs) return;  // spare units (both CTAs of a pair leave together)

    const int num_kb = (p.dim + BLOCK_K - 1) / BLOCK_K;
    const int num_c_tiles = (int)((p.n_rows + BLOCK_N - 1) / BLOCK_N);

    extern __shared__ uint8_t smem_raw[];
    uint8_t* ring = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
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
            ptx::mbar_init(&tail->tmem_empty[b], NUM_EPI_WARPS * G);
        }
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

    if (warp == 0) {
        // ===== TMA producer (one lane) ====================================================
        if (lane == 0) {
            const int q_row = (q_tile * G + (int)cta_rank) * BLOCK_M;
            uint32_t stage = 0, phase = 0;
            for (int c = slab; c < num_c_tiles; c += p.num_slabs) {
                const int c_row = c * BLOCK_N + (int)cta_rank * C::B_ROWS;
                for (int kb = 0; kb < num_kb; ++kb) {
                    ptx::mbar_wait(&tail->empty[stage], phase ^ 1);
                    uint8_t* a_dst = ring + stage * C::STAGE_BYTES;
                    uint8_t* b_dst = a_dst + C::A_BYTES;
                    if constexpr (G == 1) {
                        ptx::mbar_arrive_expect_tx(&tail->full[stage], C::STAGE_BYTES);
                        ptx::tma_load_2d_hint(a_dst, &tmap_q, &tail->full[stage], kb * BLOCK_K, q_row,
                                              SRAG_L2_EVICT_LAST);
                        ptx::tma_load_2d(b_dst, &tmap_c, &tail->full[stage], kb * BLOCK_K, c_row);
                    } else {
                        // bytes of BOTH CTAs are counted on the leader's barrier
                        if (cta_rank == 0) ptx::mbar_arrive_expect_tx(&tail->full[stage], 2 * C::STAGE_BYTES);
                        ptx::tma_load_2d_2sm(a_dst, &tmap_q, &tail->full[stage], kb * BLOCK_K, q_row,
                                             SRAG_L2_EVICT_LAST);
                        ptx::tma_load_2d_2sm(b_dst, &tmap_c, &tail->full[stage], kb * BLOCK_K, c_row,
                                             0x1000000000000000ull /* evict normal */);
                    }
                    if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ===== MMA issuer (one lane of the leader CTA) ====================================
        if (cta_rank == 0 && lane == 0) {
            constexpr uint32_t idesc = ptx::umma_idesc_bf16(BLOCK_M * G, BLOCK_N);
            uint32_t stage = 0, phase = 0;
            int it = 0;
            for (int c = slab; c < num_c_tiles; c += p.num_slabs, ++it) {
                const int buf = it & 1;
                const uint32_t use = (uint32_t)(it >> 1);
                ptx::mbar_wait(&tail->tmem_empty[buf], (use & 1) ^ 1);
                ptx::tc_fence_after();
                const uint32_t d_tmem = tmem_base + (uint32_t)buf * BLOCK_N;
                for (int kb = 0; kb < num_kb; ++kb) {
                    ptx::mbar_wait(&tail->full[stage], phase);
                    ptx::tc_fence_after();
                    const uint32_t a_addr = ptx::smem_u32(ring + stage * C::STAGE_BYTES);
                    const uint32_t b_addr = a_addr + C::A_BYTES;
#pragma unroll
                    for (int k = 0; k < BLOCK_K / UMMA_K; ++k) {
                        ptx::umma_bf16<G>(d_tmem, ptx::umma_desc_k128(a_addr + k * UMMA_K * 2),
                                          ptx::umma_desc_k128(b_addr + k * UMMA_K * 2), idesc,
                                          (uint32_t)((kb | k) != 0));
                    }
                    ptx::umma_commit<G>(&tail->empty[stage]);  // smem slot reusable once these MMAs retire
                    if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
                }
                ptx::umma_commit<G>(&tail->tmem_full[buf]);  // accumulator ready for the epilogue
            }
        }
        __syncwarp();
    } else {
        // ===== epilogue: thread == query =================================================
        const int ew = warp - 2;        // 0..3: private scratch index
        const int quad = warp & 3;      // TMEM lane quadrant this warp may read
        const int row_in_cta = quad * 32 + lane;
        const int q_global = (q_tile * G + (int)cta_rank) * BLOCK_M + row_in_cta;
        uint2* my_buf = p.cand + ((size_t)blockIdx.x * BLOCK_M + row_in_cta) * CAND_CAP;
        float* sc = tail->scale[ew];
        uint64_t* scratch = tail->keys[ew];
        float* thr_x = &tail->thr_xchg[ew];
        const float nanf_ = __uint_as_float(0x7FC00000u);

        float thr = -INFINITY;
        int cnt = 0;

        auto load_scales = [&](int c, float (&r)[8]) {
            const int64_t row0 = (int64_t)c * BLOCK_N + lane * 8;
            if (row0 + 8 <= p.n_rows) {
                const float4* src = reinterpret_cast<const float4*>(p.row_scale + row0);
                float4 a = __ldg(src), b = __ldg(src + 1);
                r[0] = a.x; r[1] = a.y; r[2] = a.z; r[3] = a.w;
                r[4] = b.x; r[5] = b.y; r[6] = b.z; r[7] = b.w;
            } else {
#pragma unroll
                for (int i = 0; i < 8; ++i) r[i] = (row0 + i < p.n_rows) ? __ldg(p.row_scale + row0 + i) : nanf_;
            }
        };

        float nxt[8];
        if (slab < num_c_tiles) load_scales(slab, nxt);
        int it = 0;
        for (int c = slab; c < num_c_tiles; c += p.num_slabs, ++it) {
            const int buf = it & 1;
            const uint32_t use = (uint32_t)(it >> 1);
            __syncwarp();
            *reinterpret_cast<float4*>(sc + lane * 8) = make_float4(nxt[0], nxt[1], nxt[2], nxt[3]);
            *reinterpret_cast<float4*>(sc + lane * 8 + 4) = make_float4(nxt[4], nxt[5], nxt[6], nxt[7]);
            __syncwarp();
            if (c + p.num_slabs < num_c_tiles) load_scales(c + p.num_slabs, nxt);

            ptx::mbar_wait(&tail->tmem_full[buf], use & 1);
            ptx::tc_fence_after();
            const uint32_t t_addr = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)buf * BLOCK_N;
            const uint32_t row_base = (uint32_t)c * BLOCK_N;
            // Software-pipelined TMEM reads: chunk ch+1 is in flight while chunk ch is filtered.
            // Fast path per chunk: 32 x (FMUL, FMNMX) and one compare; the append scan only runs
            // when some score of the chunk beats this thread's threshold.
            const uint32_t sc_addr = ptx::smem_u32(sc);
            uint32_t va[32], vb[32];
            ptx::tmem_ld_32x32(t_addr, va);
            auto process = [&](uint32_t (&v)[32], int ch) {
                float sv[32];
                float mx = -INFINITY;
#pragma unroll
                for (int j4 = 0; j4 < 8; ++j4) {
                    float m0, m1, m2, m3;
                    ptx::lds_v4(sc_addr + (uint32_t)(ch * 32 + j4 * 4) * 4u, m0, m1, m2, m3);
                    sv[j4 * 4 + 0] = __uint_as_float(v[j4 * 4 + 0]) * m0;
                    sv[j4 * 4 + 1] = __uint_as_float(v[j4 * 4 + 1]) * m1;
                    sv[j4 * 4 + 2] = __uint_as_float(v[j4 * 4 + 2]) * m2;
                    sv[j4 * 4 + 3] = __uint_as_float(v[j4 * 4 + 3]) * m3;
                }
#pragma unroll
                for (int j = 0; j < 32; ++j) mx = fmaxf(mx, sv[j]);  // NaN (tombstone / padding row) is ignored
                if (__any_sync(0xFFFFFFFFu, mx > thr)) {
                    if (mx > thr) {
                        const uint32_t r0 = row_base + (uint32_t)ch * 32u;
#pragma unroll
                        for (int j = 0; j < 32; ++j) {
                            if (sv[j] > thr) {
                                my_buf[cnt] = make_uint2(__float_as_uint(sv[j]), r0 + j);
                                ++cnt;
                            }
                        }
                    }
                    uint32_t need = __ballot_sync(0xFFFFFFFFu, cnt > p.trigger);
                    while (need) {
                        const int src = __ffs(need) - 1;
                        need &= need - 1;
                        uint2* b = reinterpret_cast<uint2*>(__shfl_sync(0xFFFFFFFFu, (unsigned long long)my_buf, src));
                        const int n = __shfl_sync(0xFFFFFFFFu, cnt, src);
                        const float t0 = __shfl_sync(0xFFFFFFFFu, thr, src);
                        int ncnt;
                        const float t1 = warp_compact(b, n, p.k, t0, scratch, thr_x, nullptr, &ncnt);
                        if (lane == src) { thr = t1; cnt = ncnt; }
                    }
                }
            };
#pragma unroll 1
            for (int ch = 0; ch < BLOCK_N / 32; ch += 2) {
                ptx::tmem_ld_wait_dep(va);
                ptx::tmem_ld_32x32(t_addr + (ch + 1) * 32, vb);
                process(va, ch);
                ptx::tmem_ld_wait_dep(vb);
                if (ch + 2 < BLOCK_N / 32) ptx::tmem_ld_32x32(t_addr + (ch + 2) * 32, va);
                process(vb, ch + 1);
            }
            // hand the accumulator back to the MMA issuer (leader CTA's barrier)
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) {
                if constexpr (G == 2) ptx::mbar_arrive_cluster(&tail->tmem_empty[buf], 0);
                else ptx::mbar_arrive(&tail->tmem_empty[buf]);
            }
        }
        // final flush: sorted top-k keys of every lane's query -> partial[q][slab][k]
        for (int src = 0; src < 32; ++src) {
            uint2* b = reinterpret_cast<uint2*>(__shfl_sync(0xFFFFFFFFu, (unsigned long long)my_buf, src));
            const int n = __shfl_sync(0xFFFFFFFFu, cnt, src);
            const int qg = __shfl_sync(0xFFFFFFFFu, q_global, src);
            if (qg >= p.n_queries) continue;  // warp-uniform
            uint64_t* out = p.partial + ((size_t)qg * p.num_slabs + slab) * p.k;
            int ncnt;
            (void)warp_compact(b, n, p.k, 0.f, scratch, thr_x, out, &ncnt);
        }
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
topk_finalize_kernel(const uint64_t* __restrict__ partial, int total /* lists * k */, int k, int kcap, int P,
                     const float* __restrict__ q_scale, int64_t id_base, float* __restrict__ out_scores,
                     int64_t* __restrict__ out_ids) {
    extern __shared__ uint64_t fin_keys[];
    const int q = blockIdx.x;
    const uint64_t* src = partial + (size_t)q * total;
    const int chunk = P - kcap;
    for (int i = threadIdx.x; i < kcap; i += blockDim.x) fin_keys[i] = SRAG_KEY_EMPTY;
    for (int off = 0; off < total; off += chunk) {
        for (int i = threadIdx.x; i < chunk; i += blockDim.x)
            fin_keys[kcap + i] = (off + i < total) ? src[off + i] : SRAG_KEY_EMPTY;
        __syncthreads();
        block_bitonic_desc(fin_keys, P);
        for (int i = k + threadIdx.x; i < kcap; i += blockDim.x) fin_keys[i] = SRAG_KEY_EMPTY;
        __syncthreads();
    }
    const float qs = q_scale ? q_scale[q] : 1.0f;
    for (int j = threadIdx.x; j < k; j += blockDim.x) {
        const uint64_t key = fin_keys[j];
        float s = -INFINITY;
        int64_t id = -1;
        if (key != SRAG_KEY_EMPTY) {
            s = __uint_as_float(srag_key_score_bits(key)) * qs;
            id = id_base + (int64_t)srag_key_idx(key);
        }
        out_scores[(size_t)q * k + j] = s;
        out_ids[(size_t)q * k + j] = id;
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
    int trigger;
    size_t cand_bytes;
    size_t partial_bytes;
};

int device_sm_count(int* sms) {
    static int cached_dev = -1, cached_sms = 0, cached_major = 0;
    int dev = 0;
    SRAG_CUDA(cudaGetDevice(&dev));
    if (dev != cached_dev) {
        cudaDeviceProp prop;
        SRAG_CUDA(cudaGetDeviceProperties(&prop, dev));
        cached_dev = dev;
        cached_sms = prop.multiProcessorCount;
        cached_major = prop.major;
    }
    if (cached_major != 10) {
        srag_set_error("device compute capability %d.x is not sm_100", cached_major);
        return SRAG_E_UNSUPPORTED;
    }
    *sms = cached_sms;
    return SRAG_OK;
}

int forced_cta_group() {
    const char* e = getenv("SRAG_TOPK_CTA_GROUP");
    if (e && (e[0] == '1' || e[0] == '2')) return e[0] - '0';
    return 0;
}

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
    int trig = 2 * k + 32;
    if (trig < 64) trig = 64;
    if (trig > CAND_CAP - 32) trig = CAND_CAP - 32;
    pl->trigger = trig;
    pl->cand_bytes = (size_t)pl->grid * BLOCK_M * CAND_CAP * sizeof(uint2);
    pl->partial_bytes = (size_t)n_queries * pl->num_slabs * k * sizeof(uint64_t);
    return SRAG_OK;
}

// Optional device-side timing of the two K1 kernels (bench.py's roofline leg): events are
// recorded on the launching stream, never synchronised here.
constexpr int PROF_RING = 256;
struct ProfState {
    bool on = false;
    int n = 0;
    cudaEvent_t ev[PROF_RING][3];
    bool created = false;
} g_prof;

template <int G>
int launch_main(const Plan& pl, const CUtensorMap& tq, const CUtensorMap& tc, const TopkParams& prm,
                cudaStream_t st) {
    static bool attr_done = false;
    if (!attr_done) {
        SRAG_CUDA(cudaFuncSetAttribute(topk_gemm_kernel<G>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)smem_bytes<G>()));
        attr_done = true;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(pl.grid);
    cfg.blockDim = dim3(NUM_THREADS);
    cfg.dynamicSmemBytes = smem_bytes<G>();
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = G;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    SRAG_CUDA(cudaLaunchKernelEx(&cfg, topk_gemm_kernel<G>, tq, tc, prm));
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
    *bytes = pl.cand_bytes + ((pl.partial_bytes + 255) & ~(size_t)255) + 256;
    return SRAG_OK;
}

extern "C" int srag_topk_search(const void* corpus_bf16, const float* row_scale, int64_t n_rows, int dim,
                                const void* queries_bf16, const float* q_scale, int n_queries, int k,
                                int64_t id_base, float* out_scores, int64_t* out_ids, void* workspace,
                                size_t workspace_bytes, void* stream) {
    SRAG_CHECK_ARG(queries_bf16 && out_scores && out_ids && workspace, "null pointer argument");
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
    const size_t need = pl.cand_bytes + ((pl.partial_bytes + 255) & ~(size_t)255);
    if (workspace_bytes < need) {
        srag_set_error("workspace too small: %zu < %zu", workspace_bytes, need);
        return SRAG_E_WORKSPACE;
    }
    uint8_t* ws = reinterpret_cast<uint8_t*>(workspace);
    SRAG_CHECK_ARG(((uintptr_t)ws & 15) == 0, "workspace must be 16-byte aligned");
    TopkParams prm;
    prm.row_scale = row_scale;
    prm.n_rows = n_rows;
    prm.dim = dim;
    prm.n_queries = n_queries;
    prm.k = k;
    prm.num_q_tiles = pl.num_q_tiles;
    prm.num_slabs = pl.num_slabs;
    prm.trigger = pl.trigger;
    prm.cand = reinterpret_cast<uint2*>(ws);
    prm.partial = reinterpret_cast<uint64_t*>(ws + pl.cand_bytes);
    SRAG_CUDA(cudaMemsetAsync(prm.partial, 0, pl.partial_bytes, st));
    const bool prof = g_prof.on && g_prof.n < PROF_RING;
    if (prof) SRAG_CUDA(cudaEventRecord(g_prof.ev[g_prof.n][0], st));
    if (n_rows > 0) {
        CUtensorMap tq, tc;
        rc = make_tmap(&tq, queries_bf16, n_queries, dim, BLOCK_M, false);
        if (rc) return rc;
        rc = make_tmap(&tc, corpus_bf16, n_rows, dim, BLOCK_N / pl.cta_group, true);
        if (rc) return rc;
        rc = (pl.cta_group == 2) ? launch_main<2>(pl, tq, tc, prm, st) : launch_main<1>(pl, tq, tc, prm, st);
        if (rc) return rc;
    }
    if (prof) SRAG_CUDA(cudaEventRecord(g_prof.ev[g_prof.n][1], st));
    rc = srag_launch_finalize(prm.partial, n_queries, pl.num_slabs * k, k, q_scale, id_base, out_scores, out_ids, st);
    if (rc) return rc;
    if (prof) {
        SRAG_CUDA(cudaEventRecord(g_prof.ev[g_prof.n][2], st));
        ++g_prof.n;
    }
    return SRAG_OK;
}

extern "C" int srag_prof_enable(int on) {
    if (on && !g_prof.created) {
        for (int i = 0; i < PROF_RING; ++i)
            for (int j = 0; j < 3; ++j) SRAG_CUDA(cudaEventCreate(&g_prof.ev[i][j]));
        g_prof.created = true;
    }
    g_prof.on = on != 0;
    g_prof.n = 0;
    return SRAG_OK;
}

extern "C" int srag_prof_read(double* main_ms_sum, double* finalize_ms_sum, int* n_calls) {
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
int srag_launch_finalize(const uint64_t* partial, int n_queries, int total, int k, const float* q_scale,
                         int64_t id_base, float* out_scores, int64_t* out_ids, cudaStream_t st) {
    const int kcap = (k <= 128) ? 128 : 256;
    int P = 2 * kcap;
    while (P < total + kcap && P < FIN_MAX_P) P <<= 1;
    static bool fin_attr = false;
    if (!fin_attr) {
        SRAG_CUDA(cudaFuncSetAttribute(topk_finalize_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       FIN_MAX_P * (int)sizeof(uint64_t)));
        fin_attr = true;
    }
    topk_finalize_kernel<<<n_queries, FIN_THREADS, (size_t)P * sizeof(uint64_t), st>>>(
        partial, total, k, kcap, P, q_scale, id_base, out_scores, out_ids);
    SRAG_CUDA(cudaGetLastError());
    return SRAG_OK;
}
