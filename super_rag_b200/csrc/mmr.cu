// K5: maximal-marginal-relevance rerank of a fused candidate list.
//
// Reference: maximal_marginal_relevance, graphiti_core/search/search_utils.py:1884-1922 — a ONE-SHOT score,
// not the greedy textbook loop:
//     c_i   = candidate_i / |candidate_i|                (zero vector stays zero)
//     S_ij  = c_i . c_j (i != j), S_ii = 0
//     mmr_i = lambda * (q . c_i) + (lambda - 1) * max_j S_ij        (the row max includes the zero diagonal)
//     order = stable sort by mmr descending, keep mmr_i >= min_score
//
// Here the candidates are rows of the device corpus named by id (the output of K1 / K3), so the gather of the
// embeddings, the m x m Gram matrix, the row maxima, the scores and the sort all happen in one CTA per
// query and nothing but m ids + m scores leaves the SM. The Gram matrix is a register-tiled fp32 product
// over shared-memory tiles (m <= 256, D ~ 1024: 67 MFLOP per query at most — CUDA cores, no tensor cores;
// the tile never exists in memory, only its row maxima do). 1/|c_i| is the row multiplier K0 stored at
// ingest, folded into the tile as it is staged.
#include <stdlib.h>

#include "common.cuh"

namespace {

constexpr int MMR_THREADS = 256;  // a 16 x 16 grid of thread tiles
constexpr int MMR_DK = 16;        // embedding dimensions staged per step
constexpr int MMR_MAX = 256;      // candidates per query

__device__ __forceinline__ void bf16x8_to_f32(const uint4& raw, float scale, float (&out)[8]) {
    const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        out[2 * i] = __uint_as_float(w[i] << 16) * scale;
        out[2 * i + 1] = __uint_as_float(w[i] & 0xFFFF0000u) * scale;
    }
}

template <int TM>  // every thread owns a TM x TM tile of a (16 TM) x (16 TM) block of the Gram matrix
__global__ void __launch_bounds__(MMR_THREADS)
mmr_kernel(const __nv_bfloat16* __restrict__ corpus, const float* __restrict__ row_scale, int64_t n_rows, int dim,
           const __nv_bfloat16* __restrict__ queries, const float* __restrict__ q_scale,
           const int64_t* __restrict__ cand_ids, int m, int64_t id_base, float lambda, float min_score,
           int64_t* __restrict__ out_ids, float* __restrict__ out_scores) {
    constexpr int BLK = 16 * TM;
    __shared__ float As[MMR_DK][BLK];
    __shared__ float Bs[MMR_DK][BLK];
    __shared__ int64_t s_row[MMR_MAX];   // corpus row of candidate i, -1 = absent / tombstone
    __shared__ float s_scale[MMR_MAX];   // 1 / |c_i| (0 for a zero vector)
    __shared__ float s_rel[MMR_MAX];     // q . c_i
    __shared__ int s_rowmax[MMR_MAX];    // bits of max(0, max_j S_ij): non-negative floats order as ints
    __shared__ uint64_t s_key[MMR_MAX];
    const int q = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t* ids = cand_ids + (size_t)q * m;

    for (int i = tid; i < MMR_MAX; i += MMR_THREADS) {
        int64_t row = -1;
        float sc = 0.f;
        if (i < m) {
            const int64_t id = ids[i];
            if (id >= id_base && id - id_base < n_rows) {
                row = id - id_base;
                sc = __ldg(row_scale + row);
                if (!(sc == sc)) row = -1;  // NaN multiplier = deleted row
            }
        }
        s_row[i] = row;
        s_scale[i] = row >= 0 ? sc : 0.f;
        s_rel[i] = 0.f;
        s_rowmax[i] = 0;
    }
    __syncthreads();

    // relevance q . c_i: one warp per candidate, 8 dimensions per lane and step
    const __nv_bfloat16* qv = queries + (size_t)q * dim;
    for (int i = warp; i < m; i += MMR_THREADS / 32) {
        const int64_t row = s_row[i];
        if (row < 0) continue;
        const __nv_bfloat16* cv = corpus + (size_t)row * dim;
        float acc = 0.f;
        for (int d = 8 * lane; d < dim; d += 256) {
            float a[8], b[8];
            bf16x8_to_f32(__ldg(reinterpret_cast<const uint4*>(qv + d)), 1.0f, a);
            bf16x8_to_f32(__ldg(reinterpret_cast<const uint4*>(cv + d)), 1.0f, b);
#pragma unroll
            for (int e = 0; e < 8; ++e) acc = fmaf(a[e], b[e], acc);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xFFFFFFFFu, acc, o);
        if (lane == 0) s_rel[i] = acc * __ldg(q_scale + q) * s_scale[i];
    }

    // Gram matrix block by block; only the row maxima survive
    const int tx = tid & 15, ty = tid >> 4;
    const int nblk = (m + BLK - 1) / BLK;
    auto stage = [&](float (&T)[MMR_DK][BLK], int blk, int d0) {
        for (int e = tid; e < BLK * 2; e += MMR_THREADS) {  // 8 dimensions (16 bytes) of one row per thread
            const int r = e >> 1, h = e & 1;
            const int i = blk * BLK + r;
            const int d = d0 + 8 * h;
            float v[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
            if (i < m && d < dim) {
                const int64_t row = s_row[i];
                if (row >= 0) bf16x8_to_f32(__ldg(reinterpret_cast<const uint4*>(corpus + (size_t)row * dim + d)), s_scale[i], v);
            }
#pragma unroll
            for (int t = 0; t < 8; ++t) T[8 * h + t][r] = v[t];
        }
    };
    for (int rb = 0; rb < nblk; ++rb) {
        for (int cb = 0; cb < nblk; ++cb) {
            float acc[TM][TM];
#pragma unroll
            for (int a = 0; a < TM; ++a)
#pragma unroll
                for (int b = 0; b < TM; ++b) acc[a][b] = 0.f;
            for (int d0 = 0; d0 < dim; d0 += MMR_DK) {
                __syncthreads();  // previous step's tiles are consumed
                stage(As, rb, d0);
                stage(Bs, cb, d0);
                __syncthreads();
#pragma unroll
                for (int kk = 0; kk < MMR_DK; ++kk) {
                    float av[TM], bv[TM];
#pragma unroll
                    for (int a = 0; a < TM; ++a) av[a] = As[kk][ty + 16 * a];
#pragma unroll
                    for (int b = 0; b < TM; ++b) bv[b] = Bs[kk][tx + 16 * b];
#pragma unroll
                    for (int a = 0; a < TM; ++a)
#pragma unroll
                        for (int b = 0; b < TM; ++b) acc[a][b] = fmaf(av[a], bv[b], acc[a][b]);
                }
            }
#pragma unroll
            for (int a = 0; a < TM; ++a) {
                const int i = rb * BLK + ty + 16 * a;
                if (i >= m || s_row[i] < 0) continue;
                float mx = 0.f;
#pragma unroll
                for (int b = 0; b < TM; ++b) {
                    const int j = cb * BLK + tx + 16 * b;
                    if (j < m && j != i && s_row[j] >= 0) mx = fmaxf(mx, acc[a][b]);
                }
                if (mx > 0.f) atomicMax(&s_rowmax[i], __float_as_int(mx));
            }
        }
    }
    __syncthreads();

    // scores -> sortable keys (score desc, candidate position asc = the reference's stable sort)
    for (int i = tid; i < MMR_MAX; i += MMR_THREADS) {
        uint64_t key = SRAG_KEY_EMPTY;
        if (i < m && s_row[i] >= 0) {
            const float s = lambda * s_rel[i] + (lambda - 1.0f) * __int_as_float(s_rowmax[i]);
            if (s >= min_score) key = srag_pack_key(__float_as_uint(s), (uint32_t)i);
        }
        s_key[i] = key;
    }
    __syncthreads();
    for (int size = 2; size <= MMR_MAX; size <<= 1)
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            for (int t = tid; t < MMR_MAX / 2; t += MMR_THREADS) {
                const int i = 2 * t - (t & (stride - 1)), j = i + stride;
                const bool desc = (i & size) == 0;
                const uint64_t x = s_key[i], y = s_key[j];
                if ((x < y) == desc) { s_key[i] = y; s_key[j] = x; }
            }
            __syncthreads();
        }
    for (int j = tid; j < m; j += MMR_THREADS) {
        const uint64_t key = s_key[j];
        int64_t id = -1;
        float s = -INFINITY;
        if (key != SRAG_KEY_EMPTY) {
            id = ids[srag_key_idx(key)];
            s = __uint_as_float(srag_key_score_bits(key));
        }
        out_ids[(size_t)q * m + j] = id;
        out_scores[(size_t)q * m + j] = s;
    }
}


// ---------------------------------------------------------------------------------------------------------
// Tensor-core variant (round 2; m > 32). The Gram matrix is the one contraction of the path that ran on CUDA cores
// (3.7 ms per 1024 queries at m = 256). Here one CTA per query computes it with tcgen05.mma, accumulators in TMEM:
//   * the candidates' raw bf16 rows are gathered by id, 64 dimensions (128 bytes) at a time, straight into the K-major
//     128-byte-swizzled shared-memory layout the MMA reads (what TMA would write: 16-byte chunk c of row r at
//     r * 128 + ((c ^ (r & 7)) << 4)); two stages, so the gather of k-block kb + 1 overlaps the MMAs of kb;
//   * A = candidate rows [128 h, 128 h + 128) (M = 128, h = 0 / 1), B = all candidate rows (N = 128 or 256): D_h[i][j] =
//     c_i . c_j for the unscaled rows, fp32, in TMEM columns [h N, h N + N);
//   * epilogue: a thread owns one row (TMEM lane), reads it 32 columns at a time and keeps max_j D_ij s_j over the valid
//     j != i; times s_i (>= 0) that is max_j S_ij. Products of bf16 values are exact in fp32, so the only rounding is the
//     accumulation — closer to the float64 reference than the scaled-fp32 tiles of the CUDA-core kernel.
// Relevance, scoring and the sort are the code above.
// ---------------------------------------------------------------------------------------------------------
constexpr int MMRT_THREADS = 256;
constexpr int MMRT_KB = 64;                       // dimensions per k-block: one 128-byte swizzle row
constexpr int MMRT_STAGES = 4;                    // k-blocks in flight (cp.async groups)

__device__ __forceinline__ void mmr_cp_async_16(uint32_t smem_dst, const void* gsrc, uint32_t src_bytes) {
    // src_bytes = 16 copies, 0 zero-fills the 16 bytes (the source address must still be valid)
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_dst), "l"(gsrc), "r"(src_bytes) : "memory");
}

__global__ void __launch_bounds__(MMRT_THREADS)
mmr_tc_kernel(const __nv_bfloat16* __restrict__ corpus, const float* __restrict__ row_scale, int64_t n_rows, int dim,
              const __nv_bfloat16* __restrict__ queries, const float* __restrict__ q_scale,
              const int64_t* __restrict__ cand_ids, int m, int64_t id_base, float lambda, float min_score,
              int64_t* __restrict__ out_ids, float* __restrict__ out_scores) {
    extern __shared__ uint8_t mmr_smem_raw[];
    uint8_t* stage_base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(mmr_smem_raw) + 1023) & ~(uintptr_t)1023);
    __shared__ int64_t s_row[MMR_MAX];
    __shared__ float s_scale[MMR_MAX];
    __shared__ float s_rel[MMR_MAX];
    __shared__ float s_rowmax[MMR_MAX];
    __shared__ uint64_t s_key[MMR_MAX];
    __shared__ uint64_t mma_done[MMRT_STAGES];
    __shared__ uint32_t tmem_slot;
    const int q = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t* ids = cand_ids + (size_t)q * m;
    const int halves = m > 128 ? 2 : 1;
    const int N = m > 128 ? 256 : 128;
    const int rows_pad = N;                               // rows staged per k-block
    const int stage_bytes = rows_pad * 128;
    const int num_kb = (dim + MMRT_KB - 1) / MMRT_KB;
    const uint32_t tmem_cols = (uint32_t)(halves * N);   // 128 or 512: a power of two >= 32
    uint8_t* q_s = stage_base + MMRT_STAGES * stage_bytes;   // the query, raw bf16, zero-padded to num_kb * 64 dimensions

    for (int i = tid; i < MMR_MAX; i += MMRT_THREADS) {
        int64_t row = -1;
        float sc = 0.f;
        if (i < m) {
            const int64_t id = ids[i];
            if (id >= id_base && id - id_base < n_rows) {
                row = id - id_base;
                sc = __ldg(row_scale + row);
                if (!(sc == sc)) row = -1;  // NaN multiplier = deleted row
            }
        }
        s_row[i] = row;
        s_scale[i] = row >= 0 ? sc : 0.f;
        s_rel[i] = 0.f;
        s_rowmax[i] = 0.f;
    }
    for (int c = tid; c < num_kb * 8; c += MMRT_THREADS) {
        uint4 v = make_uint4(0u, 0u, 0u, 0u);
        if (c * 8 < dim) v = __ldg(reinterpret_cast<const uint4*>(queries + (size_t)q * dim + c * 8));
        *reinterpret_cast<uint4*>(q_s + c * 16) = v;
    }
    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < MMRT_STAGES; ++s) ptx::mbar_init(&mma_done[s], 1);
        ptx::fence_mbar_init();
    }
    if (warp == 0) {
        ptx::tmem_alloc<1>(&tmem_slot, tmem_cols);
        ptx::tmem_relinquish<1>();
    }
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = tmem_slot;

    // ---- Gram matrix on the tensor cores, relevance from the same staged tiles ----
    // cp.async group g = k-block g (an empty group past the end), MMRT_STAGES - 1 k-blocks ahead of the MMAs.
    const uint32_t stage_u32 = ptx::smem_u32(stage_base);
    auto issue = [&](int kb) {
        const uint32_t dst = stage_u32 + (uint32_t)((kb % MMRT_STAGES) * stage_bytes);
        for (int e = tid; e < rows_pad * 8; e += MMRT_THREADS) {
            const int r = e >> 3, c = e & 7;
            const int d = kb * MMRT_KB + c * 8;
            const int64_t row = r < m ? s_row[r] : -1;
            const bool ok = row >= 0 && d < dim;
            mmr_cp_async_16(dst + (uint32_t)(r * 128 + ((c ^ (r & 7)) << 4)), ok ? corpus + (size_t)row * dim + d : corpus,
                            ok ? 16u : 0u);
        }
    };
    for (int kb = 0; kb < MMRT_STAGES - 1; ++kb) {
        if (kb < num_kb) issue(kb);
        asm volatile("cp.async.commit_group;" ::: "memory");
    }
    const uint32_t idesc = ptx::umma_idesc_bf16(128, (uint32_t)N);
    constexpr uint32_t KSTEP = (16 * 2) >> 4;                   // descriptor units per UMMA_K = 16 bf16
    float rel_acc = 0.f;                                        // thread tid: q . (raw row tid)
    for (int kb = 0; kb < num_kb; ++kb) {
        const int st = kb % MMRT_STAGES;
        uint8_t* tile = stage_base + st * stage_bytes;
        asm volatile("cp.async.wait_group %0;" ::"n"(MMRT_STAGES - 2) : "memory");   // this thread's part of k-block kb landed
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy writes -> visible to the tensor core
        __syncthreads();                                        // ... everyone's part; and relevance(kb - 1) is over
        if (warp == 0) {
            ptx::tc_fence_after();
            if (ptx::elect_one()) {
                const uint64_t b_desc = ptx::umma_desc_k128(ptx::smem_u32(tile));
                for (int h = 0; h < halves; ++h) {
                    const uint64_t a_desc = ptx::umma_desc_k128(ptx::smem_u32(tile + h * 128 * 128));
#pragma unroll
                    for (int k = 0; k < MMRT_KB / 16; ++k)
                        ptx::umma_bf16<1>(tmem_base + (uint32_t)(h * N), a_desc + k * KSTEP, b_desc + k * KSTEP, idesc,
                                          (uint32_t)((kb | k) != 0));
                }
                ptx::umma_commit<1>(&mma_done[st]);
            }
            __syncwarp();
        }
        // refill the stage k-block kb - 1 used (its MMAs must have read it; its relevance pass ended before the barrier)
        const int nxt = kb + MMRT_STAGES - 1;
        if (nxt < num_kb) {
            if (kb >= 1) ptx::mbar_wait(&mma_done[(kb - 1) % MMRT_STAGES], (uint32_t)(((kb - 1) / MMRT_STAGES) & 1));
            issue(nxt);
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
        // relevance: thread = candidate row; the swizzle makes the 16-byte reads of 8 neighbouring rows conflict-free
        if (tid < rows_pad) {
            const uint8_t* rowp = tile + tid * 128;
            const uint8_t* qp = q_s + kb * 128;
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                float a[8], b[8];
                bf16x8_to_f32(*reinterpret_cast<const uint4*>(rowp + ((c ^ (tid & 7)) << 4)), 1.0f, a);
                bf16x8_to_f32(*reinterpret_cast<const uint4*>(qp + c * 16), 1.0f, b);
#pragma unroll
                for (int e = 0; e < 8; ++e) rel_acc = fmaf(a[e], b[e], rel_acc);
            }
        }
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    if (tid < m && s_row[tid] >= 0) s_rel[tid] = rel_acc * __ldg(q_scale + q) * s_scale[tid];
    // all MMAs retired: the last commit covers every earlier one
    {
        const int last = num_kb - 1;
        ptx::mbar_wait(&mma_done[last % MMRT_STAGES], (uint32_t)((last / MMRT_STAGES) & 1));
        ptx::tc_fence_after();
    }
    // epilogue: thread = row (TMEM lane) of half `warp >> 2`, lane quadrant `warp & 3`
    {
        const int h = warp >> 2, quad = warp & 3;
        if (h < halves) {
            const int i = h * 128 + quad * 32 + lane;
            const uint32_t t_addr = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(h * N);
            float mx = 0.f;
            for (int ch = 0; ch < N / 32; ++ch) {
                uint32_t v[32];
                ptx::tmem_ld_32x32(t_addr + (uint32_t)(ch * 32), v);
                ptx::tmem_ld_wait_dep(v);
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    const int col = ch * 32 + j;
                    if (col < m && col != i && s_row[col] >= 0) mx = fmaxf(mx, __uint_as_float(v[j]) * s_scale[col]);
                }
            }
            if (i < m && s_row[i] >= 0) s_rowmax[i] = fmaxf(0.f, mx * s_scale[i]);
        }
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc<1>(tmem_base, tmem_cols);
    }

    // scores -> sortable keys (score desc, candidate position asc = the reference's stable sort)
    for (int i = tid; i < MMR_MAX; i += MMRT_THREADS) {
        uint64_t key = SRAG_KEY_EMPTY;
        if (i < m && s_row[i] >= 0) {
            const float s = lambda * s_rel[i] + (lambda - 1.0f) * s_rowmax[i];
            if (s >= min_score) key = srag_pack_key(__float_as_uint(s), (uint32_t)i);
        }
        s_key[i] = key;
    }
    __syncthreads();
    for (int size = 2; size <= MMR_MAX; size <<= 1)
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            for (int t = tid; t < MMR_MAX / 2; t += MMRT_THREADS) {
                const int i = 2 * t - (t & (stride - 1)), j = i + stride;
                const bool desc = (i & size) == 0;
                const uint64_t x = s_key[i], y = s_key[j];
                if ((x < y) == desc) { s_key[i] = y; s_key[j] = x; }
            }
            __syncthreads();
        }
    for (int j = tid; j < m; j += MMRT_THREADS) {
        const uint64_t key = s_key[j];
        int64_t id = -1;
        float s = -INFINITY;
        if (key != SRAG_KEY_EMPTY) {
            id = ids[srag_key_idx(key)];
            s = __uint_as_float(srag_key_score_bits(key));
        }
        out_ids[(size_t)q * m + j] = id;
        out_scores[(size_t)q * m + j] = s;
    }
}

}  // namespace

extern "C" int srag_mmr_rerank(const void* corpus_bf16, const float* row_scale, int64_t n_rows, int dim,
                               const void* queries_bf16, const float* q_scale, int n_queries, const int64_t* cand_ids,
                               int m, int64_t id_base, float mmr_lambda, float min_score, int64_t* out_ids,
                               float* out_scores, void* stream) {
    SRAG_CHECK_ARG(n_queries >= 1 && m >= 1 && m <= MMR_MAX, "bad sizes n_queries=%d m=%d (m <= %d)", n_queries, m, MMR_MAX);
    SRAG_CHECK_ARG(dim >= 8 && dim % 8 == 0, "dim=%d must be a positive multiple of 8", dim);
    SRAG_CHECK_ARG(queries_bf16 && q_scale && cand_ids && out_ids && out_scores, "null pointer argument");
    SRAG_CHECK_ARG(n_rows == 0 || (corpus_bf16 && row_scale), "null corpus / row_scale");
    SRAG_CHECK_ARG(((uintptr_t)corpus_bf16 & 15) == 0 && ((uintptr_t)queries_bf16 & 15) == 0,
                   "corpus / queries must be 16-byte aligned");
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    const __nv_bfloat16* c = reinterpret_cast<const __nv_bfloat16*>(corpus_bf16);
    const __nv_bfloat16* qv = reinterpret_cast<const __nv_bfloat16*>(queries_bf16);
    static const bool use_tc = [] { const char* e = getenv("SRAG_MMR_TC"); return !(e && e[0] == '0'); }();
    const int smem = MMRT_STAGES * (m > 128 ? 256 : 128) * 128 + ((dim + MMRT_KB - 1) / MMRT_KB) * 128 + 1024;
    // tensor-core Gram matrix (SRAG_MMR_TC=0, or a dim whose staged query does not fit: the CUDA-core kernel)
    if (m > 32 && use_tc && smem <= 220 * 1024) {
        const int rc = srag_func_smem_optin(reinterpret_cast<const void*>(&mmr_tc_kernel), smem);
        if (rc) return rc;
        mmr_tc_kernel<<<n_queries, MMRT_THREADS, smem, st>>>(c, row_scale, n_rows, dim, qv, q_scale, cand_ids, m, id_base,
                                                             mmr_lambda, min_score, out_ids, out_scores);
        SRAG_CUDA(cudaGetLastError());
        return SRAG_OK;
    }
    if (m <= 32)
        mmr_kernel<2><<<n_queries, MMR_THREADS, 0, st>>>(c, row_scale, n_rows, dim, qv, q_scale, cand_ids, m, id_base,
                                                         mmr_lambda, min_score, out_ids, out_scores);
    else if (m <= 64)
        mmr_kernel<4><<<n_queries, MMR_THREADS, 0, st>>>(c, row_scale, n_rows, dim, qv, q_scale, cand_ids, m, id_base,
                                                         mmr_lambda, min_score, out_ids, out_scores);
    else
        mmr_kernel<8><<<n_queries, MMR_THREADS, 0, st>>>(c, row_scale, n_rows, dim, qv, q_scale, cand_ids, m, id_base,
                                                         mmr_lambda, min_score, out_ids, out_scores);
    SRAG_CUDA(cudaGetLastError());
    return SRAG_OK;
}
