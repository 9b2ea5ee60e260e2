// tcgen05 / TMEM / TMA bf16 GEMM for the decoder linears (sm_100a):
//     out[m][n] = epi( sum_k A[m][k] * W[n][k] + bias[n] ),   A = activations [M][K], W = nn.Linear weight [N][K]
//
// Decode-shaped problem: M (rows = sentences*beam) is small (60..1000) while N*K is the whole weight
// matrix, so the kernel is WEIGHT-STREAMING and HBM-bound for M below the ridge (~250 rows).  Hence
// "swap-AB": the weight tile is the UMMA A operand (128 weight rows -> 128 TMEM lanes) and the
// activation rows are the UMMA N dimension (MT = 32..256 columns), so small M does not waste the
// 128-row MMA atom and every weight byte is read from HBM exactly once.
//
//   warp 0   TMA producer  : cp.async.bulk.tensor 2D (SWIZZLE_128B) W[128x64] + A[MTx64] per stage
//   warp 1   MMA issuer    : one thread, 4 x tcgen05.mma (K=16) per stage, fp32 accumulators in TMEM
//   warp 2   TMEM allocator
//   warps 4-11 epilogue    : tcgen05.ld 32x32b -> +bias (-> GELU) -> coalesced global stores; two warps per TMEM lane
//                            quarter, each taking half of the tile's columns
// Persistent CTAs (grid = #SMs) walk the tile list; TMEM accumulators are double-buffered so the
// epilogue of tile i overlaps the MMAs/loads of tile i+1.
#include <cuda.h>
#include "common.cuh"
#include "launch.h"

#define TC_BN 128          // weight rows per tile (UMMA M)
#define TC_BK 64           // K per stage: 64 bf16 = one 128-byte swizzle row
#define TC_THREADS 384      // warp 0 TMA, 1 MMA, 2 TMEM alloc, 3 idle, 4-11 epilogue (two warps per TMEM lane quarter)
#define TC_SPIN_LIMIT (1u << 28)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    uint32_t done = 0, spins = 0;
    const uint32_t addr = smem_u32(bar);
    while (true) {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done) : "r"(addr), "r"(parity) : "memory");
        if (done) break;
        if (++spins > TC_SPIN_LIMIT) __trap();     // a protocol bug must fault, never hang the GPU
    }
}
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* map, int c0, int c1, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
        : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// D[tmem] (+)= A[smem desc] * B[smem desc]; kind::f16 covers bf16 inputs with fp32 accumulation.
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accum) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accum) : "memory");
}
// ---- CTA-pair (cta_group::2) variants: the two CTAs of a cluster on one TPC run ONE 256-row MMA; each stages its own
//      128 UMMA-A rows and HALF of the UMMA-B rows, so a CTA pulls 32 KB instead of 48 KB through L2 per k-block ----
__device__ __forceinline__ uint32_t cluster_addr_of(uint32_t smem_addr, uint32_t cta_rank) {   // same offset, CTA `cta_rank`
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_addr), "r"(cta_rank));
    return r;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_bar_addr) {
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_bar_addr) : "memory");
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// TMA load into THIS CTA's shared memory whose byte count is credited to the barrier at `cluster_bar_addr` (the leader's)
__device__ __forceinline__ void tma_load_2d_pair(void* smem_dst, const CUtensorMap* map, int c0, int c1, uint32_t cluster_bar_addr) {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(cluster_bar_addr), "r"(c0), "r"(c1)
        : "memory");
}
// arrives on the barrier at this shared-memory offset in BOTH CTAs of the pair once the MMAs issued so far retire
__device__ __forceinline__ void umma_commit_pair(uint64_t* bar) {
    asm volatile(
        "{\n\t.reg .b16 m;\n\tmov.b16 m, 3;\n\t"
        "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], m;\n\t}"
        ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void umma_bf16_pair(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accum) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accum) : "memory");
}
// K-major, SWIZZLE_128B shared-memory matrix descriptor (sm_100 format: version=1 at bit 46, layout type 2 at bit 61).
// Rows are 128 bytes; 8-row groups are 1024 bytes apart (SBO); LBO unused for swizzled K-major.
__device__ __forceinline__ uint64_t make_sw128_desc(uint32_t smem_addr) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr & 0x3FFFFu) >> 4);
    d |= (uint64_t)(1024u >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}
__device__ __forceinline__ float tc_ex2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t* v) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
          "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
          "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
          "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// registers -> TMEM, same 32 lanes x 32 columns shape as tmem_ld32 (the epilogue parks its biased logits in the accumulator)
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const float* x) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
        "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
        "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"
        ::"r"(taddr), "f"(x[0]), "f"(x[1]), "f"(x[2]), "f"(x[3]), "f"(x[4]), "f"(x[5]), "f"(x[6]), "f"(x[7]),
          "f"(x[8]), "f"(x[9]), "f"(x[10]), "f"(x[11]), "f"(x[12]), "f"(x[13]), "f"(x[14]), "f"(x[15]),
          "f"(x[16]), "f"(x[17]), "f"(x[18]), "f"(x[19]), "f"(x[20]), "f"(x[21]), "f"(x[22]), "f"(x[23]),
          "f"(x[24]), "f"(x[25]), "f"(x[26]), "f"(x[27]), "f"(x[28]), "f"(x[29]), "f"(x[30]), "f"(x[31])
        : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

template <int MT> struct TcCfg {
    static constexpr int kStageBytes = TC_BN * TC_BK * 2 + MT * TC_BK * 2;
    // small-MT tiles (latency-bound decode GEMMs) keep ~100 KB in flight so that two CTAs (the tail of one GEMM and
    // the PDL-launched head of the next) can share an SM; large tiles take the whole SM.
    static constexpr int kBudget = (MT <= 64 ? 100 : 200) * 1024;
    static constexpr int kStagesRaw = kBudget / kStageBytes;
    static constexpr int kStages = kStagesRaw > 8 ? 8 : kStagesRaw;
    // two accumulator buffers of MT columns, rounded up to the power of two tcgen05.alloc wants (MT = 192 -> 512)
    static constexpr int kTmemCols = (2 * MT) <= 32 ? 32 : (2 * MT) <= 64 ? 64 : (2 * MT) <= 128 ? 128 : (2 * MT) <= 256 ? 256 : 512;
    static constexpr int kSmemBytes = kStages * kStageBytes + 1024 /*align*/ + 256 /*barriers*/;
};

// Tile t -> (split-K slice, tile of the MT-row operand, tile of the 128-row operand).  Normal launches walk the MT-row
// (activation) tiles fastest so that concurrently running CTAs share one weight tile; the vocabulary launch (operands
// swapped, see tc_epilogue_vocab) walks the 128-row (activation) tiles fastest for the same reason.
template <bool SWAPPED>
__device__ __forceinline__ void tc_tile_of(int t, int nsplit, int m_tiles, int n_tiles, int& ks, int& mt, int& nt) {
    ks = t % nsplit;
    const int u = t / nsplit;
    if (SWAPPED) { nt = u % n_tiles; mt = u / n_tiles; }
    else { mt = u % m_tiles; nt = u / m_tiles; }
}

// Rows [rows_act, rows) of the activation operand belong to groups whose search has ended (ItbCtl::rows_active): a tile
// whose activation rows all lie there is skipped -- by EVERY role of the CTA alike (they walk the same filtered tile list).
template <bool SWAPPED>
__device__ __forceinline__ bool tc_tile_dead(int t, int nsplit, int m_tiles, int n_tiles, int mt_rows, int rows_act) {
    int ks, mt, nt;
    tc_tile_of<SWAPPED>(t, nsplit, m_tiles, n_tiles, ks, mt, nt);
    return SWAPPED ? (nt * TC_BN >= rows_act) : (mt * mt_rows >= rows_act);
}

// Output of the fused vocabulary projection (EPI >= ITB_EPI_VOCAB_K4): per (decoder row, chunk of 128 vocabulary
// entries) the chunk's max logit, sum exp(logit - max) and top-KT (logit, index) -- what the beam kernel needs to finish
// log_softmax + top-k (reference src/seq_gen.py:194,203) without an (M,V) logits tensor ever reaching HBM.
struct TcVocabPart {
    float* cand_val; int* cand_idx;      // [rows][n_chunk][KT]
    float* chunk_max; float* chunk_sum;  // [rows][n_chunk]
    float* logits; int ldl;              // optional full logits [rows][ldl] (parity traces); null in production
    int n_chunk;
};

// Vocabulary launch: the operands are SWAPPED by the host -- the 128-row UMMA-A operand is the activation tile (TMEM lane
// = decoder row) and the MT = 256 row UMMA-B operand is the weight tile (TMEM column = vocabulary entry) -- so each
// epilogue thread owns ONE decoder row and walks its 128 logits of the tile in registers: + bias, running max / sum-exp
// (EX2 with pre-scaled argument), sorted top-KT insertion (strict >, ascending index => lowest index wins ties).
// Warps 4-7 take columns [0,128), warps 8-11 columns [128,256) of the tile; chunk index = 2 * vocab tile + half.
#define TC_VOC_EPI_WARPS 16 // epilogue warps of a vocabulary launch: 4 per TMEM lane quarter, 64 of the tile's 256 columns each
#define TC_VOC_THREADS (128 + 32 * TC_VOC_EPI_WARPS)
#define TC_CAND 12          // candidates per epilogue thread the KT = 4 path can take before it falls back to the full walk
#define TC_CAND_SLOTS 16    // slots per thread: the write pointer is clamped once per 4 entries; [slots][512] floats = 32 KB per CTA
#define TC_SLOT_STRIDE (32u * TC_VOC_EPI_WARPS * 4u)      // bytes between consecutive slots of one thread

// logits of 32 consecutive vocabulary entries from v0: accumulator + bias; entries past V read as -inf
__device__ __forceinline__ void tc_vocab_bias(const uint32_t* v, float* x, const float* __restrict__ bias, int v0, int V) {
    if (v0 + 32 <= V) {                                             // whole chunk inside the vocabulary: no per-entry checks
        const float4* bp = reinterpret_cast<const float4*>(bias + v0);                   // same address in every lane
#pragma unroll
        for (int j4 = 0; j4 < 8; ++j4) {
            const float4 b4 = __ldg(bp + j4);
            x[4 * j4] = __uint_as_float(v[4 * j4]) + b4.x;
            x[4 * j4 + 1] = __uint_as_float(v[4 * j4 + 1]) + b4.y;
            x[4 * j4 + 2] = __uint_as_float(v[4 * j4 + 2]) + b4.z;
            x[4 * j4 + 3] = __uint_as_float(v[4 * j4 + 3]) + b4.w;
        }
    } else {
#pragma unroll
        for (int j = 0; j < 32; ++j) {
            const bool ok = v0 + j < V;
            x[j] = ok ? __uint_as_float(v[j]) + __ldg(bias + (ok ? v0 + j : 0)) : -INFINITY;
        }
    }
}

// sorted insertion, strict >: among equal values the one inserted first (the lower index) stays ahead
template <int KT>
__device__ __forceinline__ void tc_topk_insert(float (&tv)[KT], int (&ti)[KT], float xv, int idx) {
    if (xv > tv[KT - 1]) {
        tv[KT - 1] = xv; ti[KT - 1] = idx;
#pragma unroll
        for (int u = KT - 1; u > 0; --u) {
            if (tv[u] > tv[u - 1]) {
                const float fv = tv[u]; tv[u] = tv[u - 1]; tv[u - 1] = fv;
                const int fi = ti[u]; ti[u] = ti[u - 1]; ti[u - 1] = fi;
            }
        }
    }
}

__device__ __forceinline__ void sts_f32(uint32_t addr, float v) { asm volatile("st.shared.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory"); }
__device__ __forceinline__ float lds_f32(uint32_t addr) {
    float v;
    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr) : "memory");
    return v;
}
__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// Vocabulary launch: the operands are SWAPPED by the host -- the 128-row UMMA-A operand is the activation tile (TMEM lane
// = decoder row) and the MT = 256 row UMMA-B operand is the weight tile (TMEM column = vocabulary entry) -- so each
// epilogue thread owns ONE decoder row and 64 of the tile's logits: + bias, running max / sum-exp (EX2 with pre-scaled
// argument), top-KT (value desc, index asc).  The epilogue is the launch's bottleneck (instruction issue), hence 16
// epilogue warps: warp 4 + 4 cp + q drains columns [64 cp, 64 cp + 64) of TMEM lane quarter q.  The two warps of one
// 128-entry chunk (cp = 2 half, 2 half + 1) merge through shared memory; the even one writes the chunk's partials
// (chunk index = 2 * vocab tile + half).
// PAIR: the 2-CTA launch (gemm_tc2_vocab_kernel): tiles are walked per CTA pair, this CTA owns row tile 2 * (pair's row
// tile) + rank, and the "accumulator drained" arrival goes to the LEADER CTA's barrier (tempty_remote = its
// shared::cluster address).
template <int MT, bool PAIR = false>
__device__ __forceinline__ void tc_epilogue_vocab4(uint32_t tmem_base, uint64_t* tfull, uint64_t* tempty, const float* __restrict__ bias,
                                                  const TcVocabPart vp, int V, int rows, int total, int m_tiles, int n_tiles,
                                                  int warp, int lane, float* cbuf, uint32_t tempty_remote = 0u, int rows_act = 0x7fffffff) {
    static_assert(MT == 256, "the vocabulary epilogue is written for 256-entry tiles");
    constexpr int KT = 4;
    static_assert(2 + 2 * KT <= TC_CAND_SLOTS, "the exchange of the partials uses the candidate slots");
    constexpr float kLog2e = 1.4426950408889634f;
    const int q = warp & 3, cp = (warp - 4) >> 2, half = cp >> 1, sub = cp & 1;
    // this thread's column of the candidate buffer: slot k at cslot + k * TC_SLOT_STRIDE; the partner thread (same row,
    // other 64 entries of the chunk) is 128 threads away
    const uint32_t cslot = smem_u32(cbuf + (threadIdx.x - 128));
    const uint32_t pslot = cslot + 128u * 4u;                       // sub == 0 only: where the sub == 1 partner writes
    const int bar_id = 1 + q * 2 + half;                            // named barrier of the two partner warps (64 threads)
    uint32_t as = 0, aphase = 0;
    const int t0 = PAIR ? (int)(blockIdx.x >> 1) : (int)blockIdx.x, tstep = PAIR ? (int)(gridDim.x >> 1) : (int)gridDim.x;
    for (int t = t0; t < total; t += tstep) {
        int ks, mt, nt;
        tc_tile_of<true>(t, 1, m_tiles, n_tiles, ks, mt, nt);
        if (!PAIR && nt * TC_BN >= rows_act) continue;              // rows of finished groups only (see tc_tile_dead)
        if (PAIR) nt = 2 * nt + (int)(blockIdx.x & 1);
        mbar_wait(&tfull[as], aphase);
        tc_fence_after();
        const int r = nt * TC_BN + q * 32 + lane;                   // decoder row of this thread
        const int vbase = mt * MT + cp * 64;                        // first vocabulary entry of this warp's columns
        const bool live = nt * TC_BN + q * 32 < rows;               // warp-uniform (and equal in both partner warps)
        float tv[KT]; int ti[KT];
#pragma unroll
        for (int j = 0; j < KT; ++j) { tv[j] = -INFINITY; ti[j] = -1; }
        float m = -INFINITY, s = 0.f;
        if (live) {
            const uint32_t tcol = tmem_base + ((uint32_t)(q * 32) << 16) + as * MT + cp * 64;
            bool scan_all;
            {
                // ---- KT = 4: two passes, so that the data-dependent insertion runs over ~6 candidates per row
                //      instead of being dragged through every entry by whichever lane of the warp inserts ----
                // pass 1: log-sum-exp statistics and the maxima of the 8 groups of 8 entries
                float gm[8];
#pragma unroll
                for (int ci = 0; ci < 2; ++ci) {
                    const int v0 = vbase + 32 * ci;
#pragma unroll
                    for (int g = 0; g < 4; ++g) gm[4 * ci + g] = -INFINITY;
                    if (v0 < V) {
                        uint32_t v[32];
                        tmem_ld32(tcol + 32 * ci, v);
                        float x[32];
                        tc_vocab_bias(v, x, bias, v0, V);
                        tmem_st32(tcol + 32 * ci, x);                // park the biased logits for pass 2 (and the full walk)
                        if (vp.logits != nullptr && r < rows) {
                            float* lp = vp.logits + (size_t)r * vp.ldl + v0;
#pragma unroll
                            for (int j = 0; j < 32; ++j) if (v0 + j < V) lp[j] = x[j];
                        }
#pragma unroll
                        for (int g = 0; g < 4; ++g) {
                            float gg = x[8 * g];
#pragma unroll
                            for (int j = 1; j < 8; ++j) gg = fmaxf(gg, x[8 * g + j]);
                            gm[4 * ci + g] = gg;
                        }
                        const float bm = fmaxf(fmaxf(gm[4 * ci], gm[4 * ci + 1]), fmaxf(gm[4 * ci + 2], gm[4 * ci + 3]));
                        if (bm > m) {
                            s = (m == -INFINITY) ? 0.f : s * tc_ex2((m - bm) * kLog2e);
                            m = bm;
                        }
                        const float nm = -m * kLog2e;
#pragma unroll
                        for (int j = 0; j < 32; ++j) s += tc_ex2(fmaf(x[j], kLog2e, nm));   // padding: ex2(-inf) = 0
                    }
                }
                // T = 4th largest group maximum: four distinct entries are >= T, so every top-4 entry is >= T
                float t0v = -INFINITY, t1v = -INFINITY, t2v = -INFINITY, t3v = -INFINITY;
#pragma unroll
                for (int g = 0; g < 8; ++g) {
                    float cgm = gm[g], hi;
                    hi = fmaxf(t0v, cgm); cgm = fminf(t0v, cgm); t0v = hi;
                    hi = fmaxf(t1v, cgm); cgm = fminf(t1v, cgm); t1v = hi;
                    hi = fmaxf(t2v, cgm); cgm = fminf(t2v, cgm); t2v = hi;
                    t3v = fmaxf(t3v, cgm);
                }
                t3v = fmaxf(t3v, -3.0e38f);                         // fewer than 4 groups: every real entry, never the -inf padding
                // pass 2: entries >= T go to this thread's column of the candidate buffer, their positions into one bit
                // mask per 32 entries
                uint32_t mk[2] = {0u, 0u};
                uint32_t wp = cslot;                                 // write pointer, clamped to slot TC_CAND every 4 entries
                tmem_st_wait();
#pragma unroll
                for (int ci = 0; ci < 2; ++ci) {
                    const int v0 = vbase + 32 * ci;
                    if (v0 < V) {
                        uint32_t v[32];
                        tmem_ld32(tcol + 32 * ci, v);                // biased logits (-inf past V), parked by pass 1
#pragma unroll
                        for (int j = 0; j < 32; ++j) {
                            const float xj = __uint_as_float(v[j]);
                            if (xj >= t3v) {
                                sts_f32(wp, xj);
                                wp += TC_SLOT_STRIDE;
                                mk[ci] |= 1u << j;
                            }
                            if ((j & 3) == 3) wp = min(wp, cslot + TC_CAND * TC_SLOT_STRIDE);
                        }
                    }
                }
                const int ncand = __popc(mk[0]) + __popc(mk[1]);
                // more than TC_CAND candidates in some lane (many equal logits; practically never): the whole warp walks
                // everything -- warp-uniform, because the walk below re-reads TMEM with .sync.aligned loads
                scan_all = __any_sync(0xffffffffu, ncand > TC_CAND);
                if (!scan_all) {
                    // pass 3: sorted insertion of the candidates in ascending index order
                    uint32_t cur = mk[0];
                    int ci = 0;
                    for (int k = 0; k < ncand; ++k) {
                        if (cur == 0u) { ci = 1; cur = mk[1]; }
                        const int j = __ffs(cur) - 1;
                        cur &= cur - 1u;
                        tc_topk_insert<KT>(tv, ti, lds_f32(cslot + TC_SLOT_STRIDE * (uint32_t)k), vbase + 32 * ci + j);
                    }
                }
            }
            // the full walk (m, s and the logits are done above; the accumulator holds the biased logits)
#pragma unroll 1
            for (int c0 = 0; c0 < (scan_all ? 64 : 0); c0 += 32) {
                const int v0 = vbase + c0;
                if (v0 >= V) break;                                 // entries past V: the chunk stays (or ends) empty
                uint32_t v[32];
                tmem_ld32(tcol + c0, v);
#pragma unroll
                for (int j = 0; j < 32; ++j) tc_topk_insert<KT>(tv, ti, __uint_as_float(v[j]), v0 + j);
            }
        }
        // the accumulator is drained: hand it back before the (cheap) merge
        tc_fence_before();
        if (PAIR) mbar_arrive_cluster(tempty_remote + 8u * as);
        else mbar_arrive(&tempty[as]);
        if (++as == 2) { as = 0; aphase ^= 1; }

        if (live) {
            // ---- merge the two 64-entry halves of the chunk: the upper one hands {max, sum, top-KT} over in its own
            //      candidate slots; all its indices are higher, so strict-> insertion keeps the lowest index first ----
            if (sub == 1) {
                sts_f32(cslot, m);
                sts_f32(cslot + TC_SLOT_STRIDE, s);
#pragma unroll
                for (int j = 0; j < KT; ++j) {
                    sts_f32(cslot + (2 + 2 * j) * TC_SLOT_STRIDE, tv[j]);
                    sts_f32(cslot + (3 + 2 * j) * TC_SLOT_STRIDE, __int_as_float(ti[j]));
                }
            }
            named_bar_sync(bar_id, 64);
            if (sub == 0) {
                const float m1 = lds_f32(pslot), s1 = lds_f32(pslot + TC_SLOT_STRIDE);
#pragma unroll
                for (int j = 0; j < KT; ++j)
                    tc_topk_insert<KT>(tv, ti, lds_f32(pslot + (2 + 2 * j) * TC_SLOT_STRIDE),
                                       __float_as_int(lds_f32(pslot + (3 + 2 * j) * TC_SLOT_STRIDE)));
                const float mm = fmaxf(m, m1);
                if (mm > -INFINITY) s = s * tc_ex2((m - mm) * kLog2e) + s1 * tc_ex2((m1 - mm) * kLog2e);   // ex2(-inf) = 0
                m = mm;
            }
            named_bar_sync(bar_id, 64);                              // the slots may be overwritten by the next tile's candidates
            if (sub == 0 && r < rows) {
                const size_t ci = (size_t)r * vp.n_chunk + (size_t)(mt * 2 + half);
                vp.chunk_max[ci] = m;
                vp.chunk_sum[ci] = s;
#pragma unroll
                for (int j = 0; j < KT; j += 4) {
                    *reinterpret_cast<float4*>(vp.cand_val + ci * KT + j) = make_float4(tv[j], tv[j + 1], tv[j + 2], tv[j + 3]);
                    *reinterpret_cast<int4*>(vp.cand_idx + ci * KT + j) = make_int4(ti[j], ti[j + 1], ti[j + 2], ti[j + 3]);
                }
            }
        }
    }
}

// ---- beam widths 5..16 (KT = 8, 16): 8 epilogue warps, each thread walks its 128 logits with a sorted insertion ----
// PAIR: the 2-CTA launch (gemm_tc2_vocab_kernel): tiles are walked per CTA pair, this CTA owns row tile 2 * (pair's row
// tile) + rank, and the "accumulator drained" arrival goes to the LEADER CTA's barrier (tempty_remote = its
// shared::cluster address).
template <int MT, int KT, bool PAIR = false>
__device__ __forceinline__ void tc_epilogue_vocab(uint32_t tmem_base, uint64_t* tfull, uint64_t* tempty, const float* __restrict__ bias,
                                                  const TcVocabPart vp, int V, int rows, int total, int m_tiles, int n_tiles,
                                                  int warp, int lane, uint32_t tempty_remote = 0u, int rows_act = 0x7fffffff) {
    static_assert(MT == 256, "the vocabulary epilogue is written for 256-entry tiles");
    constexpr float kLog2e = 1.4426950408889634f;
    const int q = warp & 3, half = (warp - 4) >> 2;
    uint32_t as = 0, aphase = 0;
    const int t0 = PAIR ? (int)(blockIdx.x >> 1) : (int)blockIdx.x, tstep = PAIR ? (int)(gridDim.x >> 1) : (int)gridDim.x;
    for (int t = t0; t < total; t += tstep) {
        int ks, mt, nt;
        tc_tile_of<true>(t, 1, m_tiles, n_tiles, ks, mt, nt);
        if (!PAIR && nt * TC_BN >= rows_act) continue;              // rows of finished groups only (see tc_tile_dead)
        if (PAIR) nt = 2 * nt + (int)(blockIdx.x & 1);
        mbar_wait(&tfull[as], aphase);
        tc_fence_after();
        const int r = nt * TC_BN + q * 32 + lane;                   // decoder row of this thread
        const int vbase = mt * MT + half * 128;                     // first vocabulary entry of this warp's columns
        if (nt * TC_BN + q * 32 < rows) {                           // warp-uniform: some row of this warp is real
            float tv[KT]; int ti[KT];
#pragma unroll
            for (int j = 0; j < KT; ++j) { tv[j] = -INFINITY; ti[j] = -1; }
            float m = -INFINITY, s = 0.f;
#pragma unroll 1
            for (int c0 = 0; c0 < 128; c0 += 32) {
                const int v0 = vbase + c0;
                if (v0 >= V) break;                                 // entries past V: the chunk stays (or ends) empty
                uint32_t v[32];
                tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + as * MT + half * 128 + c0, v);
                float x[32];
                float bm = -INFINITY;
                if (v0 + 32 <= V) {                                 // whole chunk inside the vocabulary: no per-entry checks
                    const float4* bp = reinterpret_cast<const float4*>(bias + v0);       // same address in every lane
#pragma unroll
                    for (int j4 = 0; j4 < 8; ++j4) {
                        const float4 b4 = __ldg(bp + j4);
                        x[4 * j4] = __uint_as_float(v[4 * j4]) + b4.x;
                        x[4 * j4 + 1] = __uint_as_float(v[4 * j4 + 1]) + b4.y;
                        x[4 * j4 + 2] = __uint_as_float(v[4 * j4 + 2]) + b4.z;
                        x[4 * j4 + 3] = __uint_as_float(v[4 * j4 + 3]) + b4.w;
                    }
#pragma unroll
                    for (int j = 0; j < 32; ++j) bm = fmaxf(bm, x[j]);
                } else {
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        const bool ok = v0 + j < V;
                        x[j] = ok ? __uint_as_float(v[j]) + __ldg(bias + (ok ? v0 + j : 0)) : -INFINITY;
                        bm = fmaxf(bm, x[j]);
                    }
                }
                if (vp.logits != nullptr && r < rows) {
                    float* lp = vp.logits + (size_t)r * vp.ldl + v0;
#pragma unroll
                    for (int j = 0; j < 32; ++j) if (v0 + j < V) lp[j] = x[j];
                }
                if (bm > m) {
                    s = (m == -INFINITY) ? 0.f : s * tc_ex2((m - bm) * kLog2e);
                    m = bm;
                }
                const float nm = -m * kLog2e;
#pragma unroll
                for (int j = 0; j < 32; ++j) s += tc_ex2(fmaf(x[j], kLog2e, nm));       // padding: ex2(-inf) = 0
                if (bm > tv[KT - 1]) {
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        if (x[j] > tv[KT - 1]) {
                            tv[KT - 1] = x[j]; ti[KT - 1] = v0 + j;
#pragma unroll
                            for (int u = KT - 1; u > 0; --u) {
                                if (tv[u] > tv[u - 1]) {
                                    const float fv = tv[u]; tv[u] = tv[u - 1]; tv[u - 1] = fv;
                                    const int fi = ti[u]; ti[u] = ti[u - 1]; ti[u - 1] = fi;
                                }
                            }
                        }
                    }
                }
            }
            if (r < rows) {
                const size_t ci = (size_t)r * vp.n_chunk + (size_t)(mt * 2 + half);
                vp.chunk_max[ci] = m;
                vp.chunk_sum[ci] = s;
#pragma unroll
                for (int j = 0; j < KT; j += 4) {
                    *reinterpret_cast<float4*>(vp.cand_val + ci * KT + j) = make_float4(tv[j], tv[j + 1], tv[j + 2], tv[j + 3]);
                    *reinterpret_cast<int4*>(vp.cand_idx + ci * KT + j) = make_int4(ti[j], ti[j + 1], ti[j + 2], ti[j + 3]);
                }
            }
        }
        tc_fence_before();
        if (PAIR) mbar_arrive_cluster(tempty_remote + 8u * as);
        else mbar_arrive(&tempty[as]);
        if (++as == 2) { as = 0; aphase ^= 1; }
    }
}

// Epilogue of warps 4-11: TMEM lane = weight row n (this lane), TMEM column = activation row m.  Per 32-column chunk one
// tcgen05.ld, then 32 stores out[m][n] that are coalesced across the warp (consecutive n); the row pointer advances by
// ldo per column, full chunks take a path without per-element bounds checks.
// PAIR (gemm_tc2_kernel): tiles are walked per CTA pair (256 weight rows x 256 activation rows); this CTA owns the weight
// rows of 128-row tile 2 * (pair's tile) + rank, and hands the accumulator back on the LEADER's barrier (tempty_remote).
template <int MT, int EPI, typename OT, bool PAIR = false>
__device__ __forceinline__ void tc_epilogue(uint32_t tmem_base, uint64_t* tfull, uint64_t* tempty, const float* __restrict__ bias,
                                            OT* __restrict__ out, int M, int N, int ldo, int nsplit, size_t part_stride,
                                            int total, int m_tiles, int warp, int lane, int rows_act, uint32_t tempty_remote = 0u) {
    const int q = warp & 3;                                 // TMEM lane quarter this warp may access
    const int half = (warp - 4) >> 2;                       // which half of the tile's columns this warp drains
    constexpr int CH = (MT / 2 < 32) ? 32 : MT / 2;         // MT = 32: the second warp of the quarter has no columns
    uint32_t as = 0, aphase = 0;
    const int t0 = PAIR ? (int)(blockIdx.x >> 1) : (int)blockIdx.x, tstep = PAIR ? (int)(gridDim.x >> 1) : (int)gridDim.x;
    for (int t = t0; t < total; t += tstep) {
        const int ks = PAIR ? 0 : t % nsplit, mt = PAIR ? t % m_tiles : (t / nsplit) % m_tiles;
        const int nt = PAIR ? 2 * (t / m_tiles) + (int)(blockIdx.x & 1) : (t / nsplit) / m_tiles;
        if (mt * MT >= rows_act) continue;                  // rows of finished groups only (see tc_tile_dead)
        mbar_wait(&tfull[as], aphase);
        tc_fence_after();
        const int n = nt * TC_BN + q * 32 + lane;
        const float bv = (n < N && bias != nullptr && ks == 0) ? bias[n] : 0.f;
        OT* outp = out + (size_t)ks * part_stride;          // split-K partial (f32 epilogue only)
#pragma unroll 1
        for (int c0 = half * CH; c0 < (half + 1) * CH && c0 < MT; c0 += 32) {
            const int mbase = mt * MT + c0;
            if (mbase >= M) break;                              // rows past M (zero-filled by TMA): nothing to store
            uint32_t v[32];
            tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + as * MT + c0, v);
            if (n < N) {
                OT* p = outp + (size_t)mbase * ldo + n;
                if (mbase + 32 <= M) {
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        float x = __uint_as_float(v[j]) + bv;
                        if (EPI == ITB_EPI_BIAS_GELU_ACT) x = gelu_erf_fast(x);
                        p[(size_t)j * ldo] = from_f32<OT>(x);
                    }
                } else {
                    const int rows = M - mbase;
#pragma unroll
                    for (int j = 0; j < 32; ++j) {          // fully unrolled: v[] must stay in registers
                        if (j < rows) {
                            float x = __uint_as_float(v[j]) + bv;
                            if (EPI == ITB_EPI_BIAS_GELU_ACT) x = gelu_erf_fast(x);
                            p[(size_t)j * ldo] = from_f32<OT>(x);
                        }
                    }
                }
            }
        }
        tc_fence_before();
        if (PAIR) mbar_arrive_cluster(tempty_remote + 8u * as);
        else mbar_arrive(&tempty[as]);
        if (++as == 2) { as = 0; aphase ^= 1; }
    }
}

template <int MT, int EPI>
__global__ void __launch_bounds__(EPI == ITB_EPI_VOCAB_K4 ? TC_VOC_THREADS : TC_THREADS, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmW, const __grid_constant__ CUtensorMap tmA,
               const ItbCtl* __restrict__ ctl, const float* __restrict__ bias, void* __restrict__ out, int M, int N,
               int K, int ldo, int act_f16, int nsplit, size_t part_stride, const TcVocabPart vp) {
    using Cfg = TcCfg<MT>;
    constexpr bool SWAPPED = (EPI >= ITB_EPI_VOCAB_K4);   // vocabulary launch: tmW = activations, tmA = weights
    extern __shared__ uint8_t smem_raw[];
    pdl_trigger();                                     // the next kernel of the chain may start its own prologue

    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* smem_w = smem;                                           // [stages][128][64] bf16, 1024-aligned
    uint8_t* smem_a = smem + Cfg::kStages * (TC_BN * TC_BK * 2);      // [stages][MT][64] bf16, 1024-aligned
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + Cfg::kStages * Cfg::kStageBytes);
    uint64_t* full = bars;
    uint64_t* empty = bars + Cfg::kStages;
    uint64_t* tfull = bars + 2 * Cfg::kStages;
    uint64_t* tempty = tfull + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);
    float* cbuf = reinterpret_cast<float*>(smem + Cfg::kStages * Cfg::kStageBytes + 256);   // ITB_EPI_VOCAB_K4 launches only

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n_tiles = (N + TC_BN - 1) / TC_BN, m_tiles = (M + MT - 1) / MT;
    // split-K: tile t = ((nt * m_tiles) + mt) * nsplit + ks computes k-blocks [ks*KB, (ks+1)*KB) into partial output ks
    const int total = n_tiles * m_tiles * nsplit, KB = K / TC_BK / nsplit;

    if (warp == 0 && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmW)) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmA)) : "memory");
    }
    if (warp == 1 && lane == 0) {
        for (int i = 0; i < Cfg::kStages; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
        for (int i = 0; i < 2; ++i) { mbar_init(&tfull[i], 1); mbar_init(&tempty[i], EPI == ITB_EPI_VOCAB_K4 ? 32 * TC_VOC_EPI_WARPS : 256); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 2) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"((uint32_t)Cfg::kTmemCols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    // Live rows of the activation operand.  Read BEFORE the programmatic-dependency wait, once per CTA, and shared by all
    // roles: the value may lag the current step (the beam kernel of the previous step may still be in flight), which is
    // safe -- rows_active never grows during a (super-)batch, so a stale value only keeps a few finished rows alive -- and
    // it lets the producer pick its first tile, whose weights it streams ahead of the wait, from the live ones.
    if (threadIdx.x == 96) tmem_slot[1] = (uint32_t)itb_rows_active(ctl);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const int rows_act = (int)tmem_slot[1];
    int tfirst = blockIdx.x;                           // this CTA's first live tile
    while (tfirst < total && tc_tile_dead<SWAPPED>(tfirst, nsplit, m_tiles, n_tiles, MT, rows_act)) tfirst += gridDim.x;

    // Weights never depend on earlier kernels: the producer starts streaming the first weight tiles BEFORE the
    // programmatic-dependency wait, so the HBM latency of the weight stream hides behind the predecessor's tail.
    int pre = 0;
    if (warp == 0 && lane == 0 && tfirst < total) {
        int ks, mt, nt;
        tc_tile_of<SWAPPED>(tfirst, nsplit, m_tiles, n_tiles, ks, mt, nt);
        pre = KB < Cfg::kStages ? KB : Cfg::kStages;
        for (int kb = 0; kb < pre; ++kb) {
            mbar_expect_tx(&full[kb], Cfg::kStageBytes);
            if (SWAPPED) tma_load_2d(smem_a + kb * (MT * TC_BK * 2), &tmA, (ks * KB + kb) * TC_BK, mt * MT, &full[kb]);
            else tma_load_2d(smem_w + kb * (TC_BN * TC_BK * 2), &tmW, (ks * KB + kb) * TC_BK, nt * TC_BN, &full[kb]);
        }
    }
    pdl_wait();                                        // activations / ctl are valid from here on
    const bool done = (ctl != nullptr) && (ctl->done != 0);

    if (warp == 0) {
        if (lane == 0) {
            if (tfirst < total) {
                int ks, mt, nt;
                tc_tile_of<SWAPPED>(tfirst, nsplit, m_tiles, n_tiles, ks, mt, nt);
                for (int kb = 0; kb < pre; ++kb) {
                    if (SWAPPED) tma_load_2d(smem_w + kb * (TC_BN * TC_BK * 2), &tmW, (ks * KB + kb) * TC_BK, nt * TC_BN, &full[kb]);
                    else tma_load_2d(smem_a + kb * (MT * TC_BK * 2), &tmA, (ks * KB + kb) * TC_BK, mt * MT, &full[kb]);
                }
                if (done)                              // search already terminated: drain the copies, then leave
                    for (int kb = 0; kb < pre; ++kb) mbar_wait(&full[kb], 0);
            }
            if (!done) {
                uint32_t stage = (pre == Cfg::kStages) ? 0 : pre, phase = (pre == Cfg::kStages) ? 1 : 0;
                int kb0 = pre;
                for (int t = tfirst; t < total; t += gridDim.x) {
                    if (tc_tile_dead<SWAPPED>(t, nsplit, m_tiles, n_tiles, MT, rows_act)) continue;
                    int ks, mt, nt;
                    tc_tile_of<SWAPPED>(t, nsplit, m_tiles, n_tiles, ks, mt, nt);
                    for (int kb = kb0; kb < KB; ++kb) {
                        mbar_wait(&empty[stage], phase ^ 1);
                        mbar_expect_tx(&full[stage], Cfg::kStageBytes);
                        tma_load_2d(smem_w + stage * (TC_BN * TC_BK * 2), &tmW, (ks * KB + kb) * TC_BK, nt * TC_BN, &full[stage]);
                        tma_load_2d(smem_a + stage * (MT * TC_BK * 2), &tmA, (ks * KB + kb) * TC_BK, mt * MT, &full[stage]);
                        if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
                    }
                    kb0 = 0;
                }
            }
        }
    } else if (done) {
        // nothing to compute
    } else if (warp == 1) {
        if (lane == 0) {
            // instruction descriptor: D=f32 (bit 4); UMMA-A (weights) @7 and UMMA-B (activations) @10: bf16 = 1, f16 = 0 -- the two
            // formats must be EQUAL (a mixed bf16 x f16 descriptor raises an illegal-instruction fault on B200); both K-major; N=MT (>>3 @17); M=128 (>>4 @24)
            const uint32_t idesc = (1u << 4) | ((act_f16 ? 0u : 1u) << 7) | ((act_f16 ? 0u : 1u) << 10) | ((uint32_t)(MT >> 3) << 17) | ((uint32_t)(TC_BN >> 4) << 24);
            uint32_t stage = 0, phase = 0, as = 0, aphase = 0;
            for (int t = blockIdx.x; t < total; t += gridDim.x) {
                if (tc_tile_dead<SWAPPED>(t, nsplit, m_tiles, n_tiles, MT, rows_act)) continue;
                mbar_wait(&tempty[as], aphase ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + as * MT;
                for (int kb = 0; kb < KB; ++kb) {
                    mbar_wait(&full[stage], phase);
                    tc_fence_after();
                    const uint64_t dw = make_sw128_desc(smem_u32(smem_w + stage * (TC_BN * TC_BK * 2)));
                    const uint64_t da = make_sw128_desc(smem_u32(smem_a + stage * (MT * TC_BK * 2)));
#pragma unroll
                    for (int k = 0; k < TC_BK / 16; ++k)       // +32 bytes (>>4 = 2) per K=16 slice inside the swizzle row
                        umma_bf16(d_tmem, dw + 2 * k, da + 2 * k, idesc, (kb | k) != 0 ? 1u : 0u);
                    umma_commit(&empty[stage]);                 // frees the smem stage when the MMAs retire
                    if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
                }
                umma_commit(&tfull[as]);                        // accumulator ready for the epilogue
                if (++as == 2) { as = 0; aphase ^= 1; }
            }
        }
    } else if (warp >= 4) {
        if constexpr (EPI == ITB_EPI_VOCAB_K4) tc_epilogue_vocab4<MT>(tmem_base, tfull, tempty, bias, vp, M, N, total, m_tiles, n_tiles, warp, lane, cbuf, 0u, rows_act);
        else if constexpr (EPI == ITB_EPI_VOCAB_K8) tc_epilogue_vocab<MT, 8>(tmem_base, tfull, tempty, bias, vp, M, N, total, m_tiles, n_tiles, warp, lane, 0u, rows_act);
        else if constexpr (EPI == ITB_EPI_VOCAB_K16) tc_epilogue_vocab<MT, 16>(tmem_base, tfull, tempty, bias, vp, M, N, total, m_tiles, n_tiles, warp, lane, 0u, rows_act);
        else if constexpr (EPI == ITB_EPI_BIAS_F32) tc_epilogue<MT, EPI, float>(tmem_base, tfull, tempty, bias, reinterpret_cast<float*>(out), M, N, ldo, nsplit, part_stride, total, m_tiles, warp, lane, rows_act);
        else if (act_f16) tc_epilogue<MT, EPI, __half>(tmem_base, tfull, tempty, bias, reinterpret_cast<__half*>(out), M, N, ldo, 1, 0, total, m_tiles, warp, lane, rows_act);
        else tc_epilogue<MT, EPI, __nv_bfloat16>(tmem_base, tfull, tempty, bias, reinterpret_cast<__nv_bfloat16*>(out), M, N, ldo, 1, 0, total, m_tiles, warp, lane, rows_act);
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)Cfg::kTmemCols) : "memory");
    }
}

// ---------------------------------- fused vocabulary projection on CTA pairs ----------------------------------
// Same math and output as gemm_tc_kernel<256, ITB_EPI_VOCAB_*>, but ONE tile (256 decoder rows x 256 vocabulary entries)
// per CTA PAIR: tcgen05.mma.cta_group::2 with M = 256.  Each CTA stages its own 128 decoder rows and 128 of the tile's
// 256 vocabulary rows per k-block (32 KB instead of 48 KB per 128x256x64 MACs), which is what the single-CTA launch is
// bound by (L2 -> SM operand feed).  The leader CTA (rank 0) owns the "stage full" and "accumulator drained" barriers
// and issues the MMAs; "stage free" and "accumulator ready" are multicast to both CTAs by tcgen05.commit.
#define TC2_HALF_BYTES (128 * TC_BK * 2)          // one 128-row operand tile of a stage (16 KB)
#define TC2_STAGES 6
#define TC2_SMEM (TC2_STAGES * 2 * TC2_HALF_BYTES + 1024 + 256 + TC_CAND_SLOTS * 32 * TC_VOC_EPI_WARPS * 4)

template <int KT>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(KT == 4 ? TC_VOC_THREADS : TC_THREADS, 1)
gemm_tc2_vocab_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmWv,
                      const ItbCtl* __restrict__ ctl, const float* __restrict__ bias, int V, int rows, int K, int act_f16,
                      const TcVocabPart vp) {
    extern __shared__ uint8_t smem_raw[];
    pdl_trigger();
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* smem_x = smem;                                           // [stages][128 decoder rows][64]
    uint8_t* smem_w = smem + TC2_STAGES * TC2_HALF_BYTES;             // [stages][128 vocabulary rows][64]
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + TC2_STAGES * 2 * TC2_HALF_BYTES);
    uint64_t* full = bars;                     // leader's is used: expects both CTAs' bytes
    uint64_t* empty = bars + TC2_STAGES;       // per CTA, multicast arrival
    uint64_t* tfull = bars + 2 * TC2_STAGES;   // per CTA, multicast arrival
    uint64_t* tempty = tfull + 2;              // leader's is used: 256 epilogue threads of each CTA
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);
    float* cbuf = reinterpret_cast<float*>(smem + TC2_STAGES * 2 * TC2_HALF_BYTES + 256);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint32_t rank;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
    const int pair = blockIdx.x >> 1, npair = gridDim.x >> 1;
    const int rp_tiles = (rows + 255) / 256, v_tiles = (V + 255) / 256, total = rp_tiles * v_tiles, KB = K / TC_BK;

    if (warp == 0 && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmX)) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmWv)) : "memory");
    }
    if (warp == 1 && lane == 0) {
        for (int i = 0; i < TC2_STAGES; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
        for (int i = 0; i < 2; ++i) { mbar_init(&tfull[i], 1); mbar_init(&tempty[i], KT == 4 ? 2 * 32 * TC_VOC_EPI_WARPS : 512); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    cluster_sync_all();                        // both CTAs' barriers exist before any remote arrival / TMEM pairing
    if (warp == 2) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(512u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t full_l = cluster_addr_of(smem_u32(full), 0), tempty_l = cluster_addr_of(smem_u32(tempty), 0);

    // the weight halves of the first stages are requested before the programmatic-dependency wait (weights are static)
    int pre = 0;
    if (warp == 0 && lane == 0 && pair < total) {
        const int vt = pair / rp_tiles;
        pre = KB < TC2_STAGES ? KB : TC2_STAGES;
        for (int kb = 0; kb < pre; ++kb) {
            if (rank == 0) mbar_expect_tx(&full[kb], 4 * TC2_HALF_BYTES);
            tma_load_2d_pair(smem_w + kb * TC2_HALF_BYTES, &tmWv, kb * TC_BK, vt * 256 + (int)rank * 128, full_l + 8u * kb);
        }
    }
    pdl_wait();
    const bool done = (ctl != nullptr) && (ctl->done != 0);

    if (warp == 0) {
        if (lane == 0) {
            if (pair < total) {
                const int rp = pair % rp_tiles;
                for (int kb = 0; kb < pre; ++kb)
                    tma_load_2d_pair(smem_x + kb * TC2_HALF_BYTES, &tmX, kb * TC_BK, (2 * rp + (int)rank) * 128, full_l + 8u * kb);
                if (done && rank == 0)                 // search already terminated: drain both CTAs' copies, then leave
                    for (int kb = 0; kb < pre; ++kb) mbar_wait(&full[kb], 0);
            }
            if (!done) {
                uint32_t stage = (pre == TC2_STAGES) ? 0 : pre, phase = (pre == TC2_STAGES) ? 1 : 0;
                int kb0 = pre;
                for (int t = pair; t < total; t += npair) {
                    const int rp = t % rp_tiles, vt = t / rp_tiles;
                    for (int kb = kb0; kb < KB; ++kb) {
                        mbar_wait(&empty[stage], phase ^ 1);
                        if (rank == 0) mbar_expect_tx(&full[stage], 4 * TC2_HALF_BYTES);
                        tma_load_2d_pair(smem_w + stage * TC2_HALF_BYTES, &tmWv, kb * TC_BK, vt * 256 + (int)rank * 128, full_l + 8u * stage);
                        tma_load_2d_pair(smem_x + stage * TC2_HALF_BYTES, &tmX, kb * TC_BK, (2 * rp + (int)rank) * 128, full_l + 8u * stage);
                        if (++stage == TC2_STAGES) { stage = 0; phase ^= 1; }
                    }
                    kb0 = 0;
                }
            }
        }
    } else if (done) {
        // nothing to compute
    } else if (warp == 1) {
        if (lane == 0 && rank == 0) {
            // D = f32; A, B formats equal (bf16 = 1 / f16 = 0); both K-major; N = 256 (>>3 @17); M = 256 over the pair (>>4 @24)
            const uint32_t idesc = (1u << 4) | ((act_f16 ? 0u : 1u) << 7) | ((act_f16 ? 0u : 1u) << 10) | ((uint32_t)(256 >> 3) << 17) | ((uint32_t)(256 >> 4) << 24);
            uint32_t stage = 0, phase = 0, as = 0, aphase = 0;
            for (int t = pair; t < total; t += npair) {
                mbar_wait(&tempty[as], aphase ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + as * 256;
                for (int kb = 0; kb < KB; ++kb) {
                    mbar_wait(&full[stage], phase);
                    tc_fence_after();
                    const uint64_t dx = make_sw128_desc(smem_u32(smem_x + stage * TC2_HALF_BYTES));
                    const uint64_t dw = make_sw128_desc(smem_u32(smem_w + stage * TC2_HALF_BYTES));
#pragma unroll
                    for (int k = 0; k < TC_BK / 16; ++k)
                        umma_bf16_pair(d_tmem, dx + 2 * k, dw + 2 * k, idesc, (kb | k) != 0 ? 1u : 0u);
                    umma_commit_pair(&empty[stage]);
                    if (++stage == TC2_STAGES) { stage = 0; phase ^= 1; }
                }
                umma_commit_pair(&tfull[as]);
                if (++as == 2) { as = 0; aphase ^= 1; }
            }
        }
    } else if (warp >= 4) {
        if constexpr (KT == 4) tc_epilogue_vocab4<256, true>(tmem_base, tfull, tempty, bias, vp, V, rows, total, v_tiles, rp_tiles, warp, lane, cbuf, tempty_l);
        else tc_epilogue_vocab<256, KT, true>(tmem_base, tfull, tempty, bias, vp, V, rows, total, v_tiles, rp_tiles, warp, lane, tempty_l);
    }

    tc_fence_before();
    __syncthreads();
    cluster_sync_all();                        // the peer's shared memory / barriers stay alive until both CTAs are through
    if (warp == 2) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
    }
}

// ---------------------------------- layer GEMMs on CTA pairs (large M; opt-in, see itb_tc_pair_applies) ----------------------------------
// out[M][N] = epi(A . W^T + b) like gemm_tc_kernel, but ONE tile of 256 weight rows x 256 activation rows per CTA PAIR
// (tcgen05.mma.cta_group::2, M = 256, N = 256).  A single-CTA 128 x 256 tile pulls (128 + 256) x 128 B per k-block through L2
// for 128 x 256 x 64 MACs (85 flop/B), the pair tile (128 + 128) x 128 B per CTA for the same MACs (128 flop/B).  Each CTA stages its own 128 weight rows (UMMA A: TMEM lane =
// weight row, as in gemm_tc_kernel, so the epilogue is the same) and 128 of the tile's 256 activation rows (its half of
// UMMA B).  Barrier protocol as in gemm_tc2_vocab_kernel: the leader (rank 0) owns "stage full" and "accumulator drained"
// and issues the MMAs; "stage free" and "accumulator ready" are multicast to both CTAs by tcgen05.commit.  No split-K.
#define TC2L_STAGES 6
#define TC2L_SMEM (TC2L_STAGES * 2 * TC2_HALF_BYTES + 1024 + 256)

__device__ __forceinline__ uint32_t ld_shared_cluster_u32(uint32_t cluster_addr) {
    uint32_t v;
    asm volatile("ld.shared::cluster.u32 %0, [%1];" : "=r"(v) : "r"(cluster_addr) : "memory");
    return v;
}

template <int EPI>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(TC_THREADS, 1)
gemm_tc2_kernel(const __grid_constant__ CUtensorMap tmW, const __grid_constant__ CUtensorMap tmA,
                const ItbCtl* __restrict__ ctl, const float* __restrict__ bias, void* __restrict__ out, int M, int N, int K,
                int ldo, int act_f16) {
    extern __shared__ uint8_t smem_raw[];
    pdl_trigger();
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* smem_w = smem;                                           // [stages][128 weight rows][64]      (UMMA A, this CTA's rows)
    uint8_t* smem_a = smem + TC2L_STAGES * TC2_HALF_BYTES;            // [stages][128 activation rows][64]  (this CTA's half of UMMA B)
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + TC2L_STAGES * 2 * TC2_HALF_BYTES);
    uint64_t* full = bars;                     // leader's is used: expects both CTAs' bytes
    uint64_t* empty = bars + TC2L_STAGES;      // per CTA, multicast arrival
    uint64_t* tfull = bars + 2 * TC2L_STAGES;  // per CTA, multicast arrival
    uint64_t* tempty = tfull + 2;              // leader's is used: 256 epilogue threads of each CTA
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);    // [0] TMEM base, [1] rows_active (the leader's reading, for both)

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint32_t rank;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
    const int pair = blockIdx.x >> 1, npair = gridDim.x >> 1;
    const int m_tiles = (M + 255) / 256, n_tiles = (N + 255) / 256, total = m_tiles * n_tiles, KB = K / TC_BK;

    if (warp == 0 && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmW)) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmA)) : "memory");
    }
    if (warp == 1 && lane == 0) {
        for (int i = 0; i < TC2L_STAGES; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
        for (int i = 0; i < 2; ++i) { mbar_init(&tfull[i], 1); mbar_init(&tempty[i], 512); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    // Both CTAs must agree on which tiles are dead: the leader reads rows_active (before the dependency wait: a stale value
    // is a LARGER one, see gemm_tc_kernel) into its shared memory ahead of the cluster barrier; the peer reads it from there.
    if (threadIdx.x == 96 && rank == 0) tmem_slot[1] = (uint32_t)itb_rows_active(ctl);
    __syncthreads();
    cluster_sync_all();                        // both CTAs' barriers exist before any remote arrival / TMEM pairing
    if (warp == 2) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(512u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const int rows_act = (int)ld_shared_cluster_u32(cluster_addr_of(smem_u32(&tmem_slot[1]), 0));
    const uint32_t full_l = cluster_addr_of(smem_u32(full), 0), tempty_l = cluster_addr_of(smem_u32(tempty), 0);
    int tfirst = pair;                         // the pair's first live tile (tile t: activation tile t % m_tiles, weight tile t / m_tiles)
    while (tfirst < total && (tfirst % m_tiles) * 256 >= rows_act) tfirst += npair;

    // the weight rows of the first stages are requested before the programmatic-dependency wait (weights are static)
    int pre = 0;
    if (warp == 0 && lane == 0 && tfirst < total) {
        const int nt = tfirst / m_tiles;
        pre = KB < TC2L_STAGES ? KB : TC2L_STAGES;
        for (int kb = 0; kb < pre; ++kb) {
            if (rank == 0) mbar_expect_tx(&full[kb], 4 * TC2_HALF_BYTES);
            tma_load_2d_pair(smem_w + kb * TC2_HALF_BYTES, &tmW, kb * TC_BK, nt * 256 + (int)rank * 128, full_l + 8u * kb);
        }
    }
    pdl_wait();
    const bool done = (ctl != nullptr) && (ctl->done != 0);

    if (warp == 0) {
        if (lane == 0) {
            if (tfirst < total) {
                const int mt = tfirst % m_tiles;
                for (int kb = 0; kb < pre; ++kb)
                    tma_load_2d_pair(smem_a + kb * TC2_HALF_BYTES, &tmA, kb * TC_BK, mt * 256 + (int)rank * 128, full_l + 8u * kb);
                if (done && rank == 0)                 // search already terminated: drain both CTAs' copies, then leave
                    for (int kb = 0; kb < pre; ++kb) mbar_wait(&full[kb], 0);
            }
            if (!done) {
                uint32_t stage = (pre == TC2L_STAGES) ? 0 : pre, phase = (pre == TC2L_STAGES) ? 1 : 0;
                int kb0 = pre;
                for (int t = tfirst; t < total; t += npair) {
                    const int mt = t % m_tiles, nt = t / m_tiles;
                    if (mt * 256 >= rows_act) continue;
                    for (int kb = kb0; kb < KB; ++kb) {
                        mbar_wait(&empty[stage], phase ^ 1);
                        if (rank == 0) mbar_expect_tx(&full[stage], 4 * TC2_HALF_BYTES);
                        tma_load_2d_pair(smem_w + stage * TC2_HALF_BYTES, &tmW, kb * TC_BK, nt * 256 + (int)rank * 128, full_l + 8u * stage);
                        tma_load_2d_pair(smem_a + stage * TC2_HALF_BYTES, &tmA, kb * TC_BK, mt * 256 + (int)rank * 128, full_l + 8u * stage);
                        if (++stage == TC2L_STAGES) { stage = 0; phase ^= 1; }
                    }
                    kb0 = 0;
                }
            }
        }
    } else if (done) {
        // nothing to compute
    } else if (warp == 1) {
        if (lane == 0 && rank == 0) {
            // D = f32; A, B formats equal (bf16 = 1 / f16 = 0); both K-major; N = 256 (>>3 @17); M = 256 over the pair (>>4 @24)
            const uint32_t idesc = (1u << 4) | ((act_f16 ? 0u : 1u) << 7) | ((act_f16 ? 0u : 1u) << 10) | ((uint32_t)(256 >> 3) << 17) | ((uint32_t)(256 >> 4) << 24);
            uint32_t stage = 0, phase = 0, as = 0, aphase = 0;
            for (int t = tfirst; t < total; t += npair) {
                if ((t % m_tiles) * 256 >= rows_act) continue;
                mbar_wait(&tempty[as], aphase ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + as * 256;
                for (int kb = 0; kb < KB; ++kb) {
                    mbar_wait(&full[stage], phase);
                    tc_fence_after();
                    const uint64_t dw = make_sw128_desc(smem_u32(smem_w + stage * TC2_HALF_BYTES));
                    const uint64_t da = make_sw128_desc(smem_u32(smem_a + stage * TC2_HALF_BYTES));
#pragma unroll
                    for (int k = 0; k < TC_BK / 16; ++k)
                        umma_bf16_pair(d_tmem, dw + 2 * k, da + 2 * k, idesc, (kb | k) != 0 ? 1u : 0u);
                    umma_commit_pair(&empty[stage]);
                    if (++stage == TC2L_STAGES) { stage = 0; phase ^= 1; }
                }
                umma_commit_pair(&tfull[as]);
                if (++as == 2) { as = 0; aphase ^= 1; }
            }
        }
    } else if (warp >= 4) {
        if constexpr (EPI == ITB_EPI_BIAS_F32)
            tc_epilogue<256, EPI, float, true>(tmem_base, tfull, tempty, bias, reinterpret_cast<float*>(out), M, N, ldo, 1, (size_t)0, total, m_tiles, warp, lane, rows_act, tempty_l);
        else if (act_f16)
            tc_epilogue<256, EPI, __half, true>(tmem_base, tfull, tempty, bias, reinterpret_cast<__half*>(out), M, N, ldo, 1, (size_t)0, total, m_tiles, warp, lane, rows_act, tempty_l);
        else
            tc_epilogue<256, EPI, __nv_bfloat16, true>(tmem_base, tfull, tempty, bias, reinterpret_cast<__nv_bfloat16*>(out), M, N, ldo, 1, (size_t)0, total, m_tiles, warp, lane, rows_act, tempty_l);
    }

    tc_fence_before();
    __syncthreads();
    cluster_sync_all();                        // the peer's shared memory / barriers stay alive until both CTAs are through
    if (warp == 2) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
    }
}

// ------------------------------------------------ host side ------------------------------------------------
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled get_encode() {
    static PFN_encodeTiled fn = nullptr;
    if (fn == nullptr) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<PFN_encodeTiled>(p);
    }
    return fn;
}

// 2-D bf16 row-major [rows][K] tensor, box = [box_rows][64], 128-byte swizzle, OOB rows read as zero.
int itb_make_tmap_16b(CUtensorMap* map, const void* base, int rows, int K, int box_rows, int is_f16) {
    PFN_encodeTiled enc = get_encode();
    if (enc == nullptr) { itb_set_error("cuTensorMapEncodeTiled entry point not available (no CUDA driver?)"); return -3; }
    cuuint64_t gdim[2] = {(cuuint64_t)K, (cuuint64_t)rows};
    cuuint64_t gstr[1] = {(cuuint64_t)K * 2};
    cuuint32_t box[2] = {(cuuint32_t)TC_BK, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(map, is_f16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), gdim, gstr, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { itb_set_error("cuTensorMapEncodeTiled failed with CUresult %d (rows=%d K=%d box=%d)", (int)r, rows, K, box_rows); return -3; }
    return 0;
}



static int num_sms() { return itb_num_sms(); }

// Tile plan.  The decode GEMMs are weight-streaming and latency-bound at small M, so the activation rows per tile
// (UMMA N = mt) and the split-K factor are chosen to spread the weight matrix over as many SMs as possible:
//   cost = waves * (bytes staged per tile + bytes of the fp32 output tile written and re-read by the consumer + fixed).
// Every split-K partial is a full fp32 output written by the GEMM and summed in add_ln, so at large M (where an
// unsplit launch already fills most SMs) the output term makes split-K lose.  Split-K (partials summed by the
// consumer in a fixed order -> deterministic) only when the caller allows it.
void itb_tc_plan(int M, int N, int K, int max_split, int* mt_out, int* split_out) {
    const int n_tiles = (N + TC_BN - 1) / TC_BN, KB = K / TC_BK;
    long best_cost = -1;
    *mt_out = 32; *split_out = 1;
    static const int kTileRows[] = {32, 64, 128, 192, 256};                    // 192: 22 row tiles x 6 = 132 CTAs for N = 768 at M = 4096
    for (int ti = 0; ti < 5; ++ti) {
        const int mt = kTileRows[ti];
        for (int sp = 1; sp <= max_split; ++sp) {
            if (KB % sp != 0) continue;
            const long tiles = (long)n_tiles * ((M + mt - 1) / mt) * sp;
            const long waves = (tiles + num_sms() - 1) / num_sms();
            const int rows = mt < M ? mt : M;                                // real activation rows in a tile
            const long staged = (long)(KB / sp) * (TC_BN + mt) * (TC_BK * 2);
            const long outb = (long)TC_BN * rows * 8;                          // fp32 write + re-read
            const long cost = waves * (staged + outb + 6 * 160 * 128);
            if (best_cost < 0 || cost < best_cost) { *mt_out = mt; *split_out = sp; best_cost = cost; }
        }
        if (mt >= M) break;
    }
}

template <int MT, int EPI>
static int launch_tc(const CUtensorMap& tmW, const CUtensorMap& tmA, const ItbCtl* ctl, const float* bias, void* out,
                     int M, int N, int K, int ldo, int act_f16, int nsplit, cudaStream_t s, const TcVocabPart vp = TcVocabPart()) {
    constexpr int kSmem = TcCfg<MT>::kSmemBytes + (EPI == ITB_EPI_VOCAB_K4 ? TC_CAND_SLOTS * 32 * TC_VOC_EPI_WARPS * 4 : 0);   // + candidate buffer
    constexpr int kThreads = EPI == ITB_EPI_VOCAB_K4 ? TC_VOC_THREADS : TC_THREADS;
    static bool attr_set[ITB_MAX_DEVICES] = {};
    {
        cudaError_t e = itb_smem_attr_once(attr_set, gemm_tc_kernel<MT, EPI>, kSmem);
        if (e != cudaSuccess) { itb_set_error("gemm_tc: cudaFuncSetAttribute: %s", cudaGetErrorString(e)); return -1; }
    }
    const int total = ((N + TC_BN - 1) / TC_BN) * ((M + MT - 1) / MT) * nsplit;
    const int grid = total < num_sms() ? total : num_sms();
    cudaError_t e = itb_launch(gemm_tc_kernel<MT, EPI>, dim3(grid), dim3(kThreads), (size_t)kSmem, s, tmW, tmA,
                               ctl, bias, out, M, N, K, ldo, act_f16, nsplit, (size_t)M * (size_t)ldo, vp);
    if (e != cudaSuccess) { itb_set_error("gemm_tc: launch failed: %s", cudaGetErrorString(e)); return -1; }
    return itb_check_launch("gemm_tc");
}

template <int MT>
static int launch_tc_epi(int epi, const CUtensorMap& tmW, const CUtensorMap& tmA, const ItbCtl* ctl, const float* bias,
                         void* out, int M, int N, int K, int ldo, int act_f16, int nsplit, cudaStream_t s) {
    if (epi == ITB_EPI_BIAS_F32) return launch_tc<MT, ITB_EPI_BIAS_F32>(tmW, tmA, ctl, bias, out, M, N, K, ldo, act_f16, nsplit, s);
    if (epi == ITB_EPI_BIAS_GELU_ACT) return launch_tc<MT, ITB_EPI_BIAS_GELU_ACT>(tmW, tmA, ctl, bias, out, M, N, K, ldo, act_f16, 1, s);
    return launch_tc<MT, ITB_EPI_BIAS_ACT>(tmW, tmA, ctl, bias, out, M, N, K, ldo, act_f16, 1, s);
}

// tmW: [N][K] box 128 rows; tmA: [M][K] box `mt` rows; (mt, nsplit) from itb_tc_plan.  With nsplit > 1 (f32 epilogue
// only) `out` receives nsplit partial outputs [nsplit][M][ldo]; the bias is added to partial 0.
int itb_gemm_tc(int epi, const CUtensorMap* tmW, const CUtensorMap* tmA, const void* ctl, const float* bias, void* out,
                int M, int N, int K, int ldo, int act_f16, int mt, int nsplit, cudaStream_t s) {
    ITB_REQUIRE(K % TC_BK == 0 && K >= TC_BK, "linear(tcgen05): K must be a multiple of 64");
    ITB_REQUIRE(nsplit >= 1 && (K / TC_BK) % nsplit == 0 && (nsplit == 1 || epi == ITB_EPI_BIAS_F32), "linear(tcgen05): bad split-K factor");
    const ItbCtl* c = reinterpret_cast<const ItbCtl*>(ctl);
    switch (mt) {
        case 32: return launch_tc_epi<32>(epi, *tmW, *tmA, c, bias, out, M, N, K, ldo, act_f16, nsplit, s);
        case 64: return launch_tc_epi<64>(epi, *tmW, *tmA, c, bias, out, M, N, K, ldo, act_f16, nsplit, s);
        case 128: return launch_tc_epi<128>(epi, *tmW, *tmA, c, bias, out, M, N, K, ldo, act_f16, nsplit, s);
        case 192: return launch_tc_epi<192>(epi, *tmW, *tmA, c, bias, out, M, N, K, ldo, act_f16, nsplit, s);
        case 256: return launch_tc_epi<256>(epi, *tmW, *tmA, c, bias, out, M, N, K, ldo, act_f16, nsplit, s);
        default: itb_set_error("linear(tcgen05): unsupported tile height %d", mt); return -2;
    }
}

// CTA-pair launch of a layer GEMM (gemm_tc2_kernel): tmW and tmA128 both with 128-row boxes.
template <int EPI>
static int launch_tc2(const CUtensorMap& tmW, const CUtensorMap& tmA128, const ItbCtl* ctl, const float* bias, void* out, int M, int N,
                      int K, int ldo, int act_f16, cudaStream_t s) {
    static bool attr_set[ITB_MAX_DEVICES] = {};
    {
        cudaError_t e = itb_smem_attr_once(attr_set, gemm_tc2_kernel<EPI>, TC2L_SMEM);
        if (e != cudaSuccess) { itb_set_error("gemm_tc2: cudaFuncSetAttribute: %s", cudaGetErrorString(e)); return -1; }
    }
    const int total = ((N + 255) / 256) * ((M + 255) / 256);
    const int cap = num_sms() & ~1;
    const int grid = 2 * total < cap ? 2 * total : cap;
    cudaError_t e = itb_launch(gemm_tc2_kernel<EPI>, dim3(grid), dim3(TC_THREADS), (size_t)TC2L_SMEM, s, tmW, tmA128, ctl, bias, out, M, N, K,
                               ldo, act_f16);
    if (e != cudaSuccess) { itb_set_error("gemm_tc2: launch failed: %s", cudaGetErrorString(e)); return -1; }
    return itb_check_launch("gemm_tc2");
}

int itb_gemm_tc_pair(int epi, const CUtensorMap* tmW, const CUtensorMap* tmA128, const void* ctl, const float* bias, void* out,
                     int M, int N, int K, int ldo, int act_f16, cudaStream_t s) {
    ITB_REQUIRE(K % TC_BK == 0 && K >= TC_BK, "linear(tcgen05 pair): K must be a multiple of 64");
    const ItbCtl* c = reinterpret_cast<const ItbCtl*>(ctl);
    if (epi == ITB_EPI_BIAS_F32) return launch_tc2<ITB_EPI_BIAS_F32>(*tmW, *tmA128, c, bias, out, M, N, K, ldo, act_f16, s);
    if (epi == ITB_EPI_BIAS_GELU_ACT) return launch_tc2<ITB_EPI_BIAS_GELU_ACT>(*tmW, *tmA128, c, bias, out, M, N, K, ldo, act_f16, s);
    return launch_tc2<ITB_EPI_BIAS_ACT>(*tmW, *tmA128, c, bias, out, M, N, K, ldo, act_f16, s);
}

// Whether a layer GEMM of this shape runs on CTA pairs.  Opt-in: ITB_GEMM_PAIR=1 every shape (tests), =2 wide outputs at
// large M (N >= 1536, M >= 1024: QKV and FFN1 of a super-batch; N = 768 makes only 3 x ceil(M / 256) pair tiles, too few for
// 74 pairs).  Measured at 4096 rows (profiles/r2_gemm_pair_ab.txt): QKV 15.6 -> 15.9 us, FFN1 + GELU 25.2 -> 25.6 us -- the
// single-CTA launches are not bound by the operand feed the pairing halves (tensor pipe active 39-44 % with the issue
// slots 17 % busy: what is left is fill / drain of 2-3 tiles per CTA), so the default stays on the single-CTA plan.
bool itb_tc_pair_applies(int M, int N, int K) {
    static int mode = -1;
    if (mode < 0) {
        const char* e = getenv("ITB_GEMM_PAIR");
        mode = (e == nullptr) ? 0 : atoi(e);
    }
    if (mode == 0 || K % TC_BK != 0) return false;
    if (mode == 1) return true;
    return N >= 1536 && M >= 1024;
}

// Fused vocabulary projection (operands swapped: tmX = activations [rows][K] box 128, tmWv = weights [V][K] box 256,
// tmWv128 = the same weights with a 128-row box for the CTA-pair launch, or null).
// Writes per (row, 128-entry chunk) max / sum-exp / top-KT into vp; n_chunk must be itb_vocab_chunks(V).
template <int KT>
static int launch_tc2_vocab(const CUtensorMap& tmX, const CUtensorMap& tmWv128, const ItbCtl* ctl, const float* bias, int rows, int V,
                            int K, int act_f16, cudaStream_t s, const TcVocabPart& vp) {
    static bool attr_set[ITB_MAX_DEVICES] = {};
    {
        cudaError_t e = itb_smem_attr_once(attr_set, gemm_tc2_vocab_kernel<KT>, TC2_SMEM);
        if (e != cudaSuccess) { itb_set_error("gemm_tc2: cudaFuncSetAttribute: %s", cudaGetErrorString(e)); return -1; }
    }
    const int total = ((rows + 255) / 256) * ((V + 255) / 256);
    const int cap = num_sms() & ~1;
    const int grid = 2 * total < cap ? 2 * total : cap;
    cudaError_t e = itb_launch(gemm_tc2_vocab_kernel<KT>, dim3(grid), dim3(KT == 4 ? TC_VOC_THREADS : TC_THREADS), (size_t)TC2_SMEM, s, tmX, tmWv128, ctl, bias, V,
                               rows, K, act_f16, vp);
    if (e != cudaSuccess) { itb_set_error("gemm_tc2: launch failed: %s", cudaGetErrorString(e)); return -1; }
    return itb_check_launch("gemm_tc2");
}

// CTA-pair launch (needs at least two 128-row tiles of decoder rows): opt-in with ITB_VOCAB_PAIR=1 -- measured 1% SLOWER
// than the single-CTA launch at 4096 rows, because the launch is bound by the epilogue's instruction issue, not by
// the operand feed the pairing halves.
static bool use_pair(int rows) {
    static int v = -1;
    if (v < 0) {
        const char* e = getenv("ITB_VOCAB_PAIR");
        v = (e != nullptr && e[0] == '1') ? 1 : 0;
    }
    return v != 0 && rows > 128;
}

int itb_gemm_tc_vocab(int kt, const CUtensorMap* tmX, const CUtensorMap* tmWv, const CUtensorMap* tmWv128, const void* ctl,
                      const float* bias, int rows, int V, int K, int act_f16, float* cand_val, int* cand_idx, float* chunk_max,
                      float* chunk_sum, float* logits, int ldl, int n_chunk, cudaStream_t s) {
    ITB_REQUIRE(K % TC_BK == 0 && K >= TC_BK, "vocab(tcgen05): K must be a multiple of 64");
    ITB_REQUIRE(n_chunk == 2 * ((V + 255) / 256), "vocab(tcgen05): n_chunk must be 2 * ceil(V / 256)");
    TcVocabPart vp;
    vp.cand_val = cand_val; vp.cand_idx = cand_idx; vp.chunk_max = chunk_max; vp.chunk_sum = chunk_sum;
    vp.logits = logits; vp.ldl = ldl; vp.n_chunk = n_chunk;
    const ItbCtl* c = reinterpret_cast<const ItbCtl*>(ctl);
    if (tmWv128 != nullptr && use_pair(rows)) {
        if (kt == 4) return launch_tc2_vocab<4>(*tmX, *tmWv128, c, bias, rows, V, K, act_f16, s, vp);
        if (kt == 8) return launch_tc2_vocab<8>(*tmX, *tmWv128, c, bias, rows, V, K, act_f16, s, vp);
        return launch_tc2_vocab<16>(*tmX, *tmWv128, c, bias, rows, V, K, act_f16, s, vp);
    }
    // kernel view: the 128-row operand ("N") is the activation matrix, the 256-row operand ("M") the weights
    if (kt == 4) return launch_tc<256, ITB_EPI_VOCAB_K4>(*tmX, *tmWv, c, bias, nullptr, V, rows, K, 0, act_f16, 1, s, vp);
    if (kt == 8) return launch_tc<256, ITB_EPI_VOCAB_K8>(*tmX, *tmWv, c, bias, nullptr, V, rows, K, 0, act_f16, 1, s, vp);
    return launch_tc<256, ITB_EPI_VOCAB_K16>(*tmX, *tmWv, c, bias, nullptr, V, rows, K, 0, act_f16, 1, s, vp);
}
