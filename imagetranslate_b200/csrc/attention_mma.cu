// Single-query attention on the tensor cores for the 16-bit decode modes (bf16 / fp16 K,V cache).
// HF BertSelfAttention on the last position only; head_dim 64.
//
// One WARP per (sentence, group of <=4 beam slots, head); no shared memory, no block barrier.  The <=4 queries of the
// group are rows 0..3 of the 16-row A operand of mma.sync.m16n8k16, so that K and V rows shared by the beam slots are
// fetched ONCE and used for all of them:
//   * S = Q.K^T : B fragments come straight from global memory -- the dot product does not care about the order of
//     the 64 head dims, so lane (n = lane/4, c = lane%4) loads the 32 contiguous bytes [16c,16c+16) dims of K row n and
//     uses them as its k-slice fragments; Q is permuted the same way.  Q (fp32 from the QKV GEMM) is split into
//     hi + lo 16-bit parts (two MMAs), so scores are as accurate as fp32-q x 16-bit-k.
//   * softmax in registers on the C fragments (FlashAttention-2 style), online across chunks of 32 positions.
//   * O += P.V : P is split hi + lo as well; V fragments (pairs of POSITIONS per register) are built from four 16-byte
//     row loads with byte permutes; the output dim order is chosen so that each lane ends up with 16 contiguous dims.
// "Virtual positions": the cross-attention memory is simply positions [0, mem_len).  For self-attention the beam
// slots of a sentence address the SAME physical cache rows (through the ancestor table written by the beam kernel) for
// every position before their lineages split, so the list is  [0,c) shared by all slots  +  for each slot its own tail
// [c, step) (owner mask = that slot only; foreign rows get -inf).  A beam reorder never copies K/V, and the K/V
// traffic of a beam-4 step is close to that of ONE hypothesis.  The rows' new K,V are appended (rounded to the cache
// type) before the list is walked.
// HBM-bound by design: 2*d*2 bytes per (virtual position, layer); the MMAs cut the instruction count ~5x against the
// CUDA-core formulation, which was issue-bound at ~17 % of HBM bandwidth (profiles/r1_attn_*.txt).
#include "common.cuh"
#include "launch.h"
#include "mma16.cuh"

#define DH 64
#define AM_WARPS 4
#define AM_THREADS (AM_WARPS * 32)
#define AM_ROWB 272            // bytes per staged row: K 128 + V 128 + 16 pad
// STAGE = virtual positions staged in shared memory per round trip (template parameter of the kernel).  Two builds:
// STAGE 64 / 3 CTAs per SM (fewest round trips; launches of a wave or two, where latency is everything) and
// STAGE 32 / 4 CTAs per SM (a third more warps in flight; many-wave launches, where the SM's issue rate is the limit).
#define AM_NCH (STAGE / 32)     // 32-position chunks per stage
#define AM_SMEM(STAGE) (AM_WARPS * (STAGE) * AM_ROWB)
#define AM_MINB(STAGE) ((STAGE) == 64 ? 3 : 4)

struct AttnMmaArgs {
    const ItbCtl* ctl;
    const int* sent_done;                // [B] or null
    const int* mem_len;                  // cross: [B] or null (= S)
    int early_kv;                        // cross: K,V and mask were complete before the last full stream dependency
    const float* q; int ldq;             // f32 queries (self: qkv buffer, k at +d, v at +2d)
    void* kv;                            // self: [M][T_max][2][d]; cross: [B][S][L][2][d]
    const unsigned char* anc;            // self: [2][M][T_max]
    const unsigned char* mask;           // cross: [B][S] or null
    void* ctx;                           // act [M][d]
    int M, beam, d, T_max, S, L, layer, n_warps;
};

// NSPLIT = 2 or 4 (small launches, where a step is bound by the length of one warp's walk and most SMs idle): NSPLIT
// consecutive warps of a CTA share ONE item; warp w of the team walks stages w, w + NSPLIT, ... of the virtual position
// list and the partial softmax states (max, sum, unnormalised output) are merged through shared memory by the team's
// first warp (split-K for attention).  The team synchronises on its own named barrier: teams of one CTA work on
// different sentences and leave independently.
template <typename T, bool SELF, int STAGE, int NSPLIT = 1>
__global__ void __launch_bounds__(AM_THREADS, AM_MINB(STAGE)) attn_mma_kernel(AttnMmaArgs a) {
    pdl_trigger();
    constexpr bool SPLIT = NSPLIT > 1;
    const int lane = threadIdx.x & 31, nl = lane >> 2, c = lane & 3;       // nl: B-fragment column / C row; c: quad lane
    const int wsub = (int)(threadIdx.x >> 5) % NSPLIT;                     // this warp's share of the stages
    const int team = (int)(threadIdx.x >> 5) / NSPLIT;                     // teams per CTA: AM_WARPS / NSPLIT
    const int sb0 = wsub * STAGE;                                          // first virtual position of this warp's first stage
    const int wg = (int)blockIdx.x * (AM_WARPS / NSPLIT) + team;
    if (wg >= a.n_warps) return;                                           // uniform over the team
    const int d = a.d, H = d / DH, nz = (a.beam + 3) >> 2;
    const int h = wg % H, z = (wg / H) % nz, b = wg / (H * nz);
    const int q0 = z * 4, row0 = b * a.beam + q0, nq = min(4, a.beam - q0);
    const int hoff = h * DH;
    // Per-warp staging buffer: STAGE virtual rows of [K 128 B | V 128 B], padded to 272 B so that the fragment reads
    // below are bank-conflict free.  All rows of a stage are requested with cp.async (16 B each, zero-filled when the
    // virtual position does not exist) BEFORE anything is consumed: one memory round trip per 64 positions, no
    // registers tied up by loads in flight.
    extern __shared__ __align__(16) unsigned char am_smem[];
    unsigned char* ws = am_smem + (size_t)(threadIdx.x >> 5) * (STAGE * AM_ROWB);
    const uint32_t ws_u32 = (uint32_t)__cvta_generic_to_shared(ws);
    // copies: one cp.async instruction moves the K (or the V) halves of 4 rows, 8 lanes x 16 bytes per row.  K and V of a
    // stage travel as two cp.async groups, so that the K half of the buffer can be refilled as soon as the scores are
    // done and the V half as soon as P.V is done (software pipeline over the stages, see the main loop)
    const int piece = lane & 7, rsub = lane >> 3;
    const uint32_t lane_dst = ws_u32 + rsub * AM_ROWB + piece * 16;
    // Cross-attention: the encoder-side K,V and the source mask were written once per batch, many launches ago, and
    // their addresses depend on the warp index only.  The first stage is therefore requested BEFORE the wait on the
    // predecessor grid and before the control words are known (rows past the source length are fetched too and
    // masked below; rows past S are zero-filled) -- it overlaps the tail of the previous kernel and the query loads.
    int mk0[AM_NCH];
    if (SELF || a.early_kv == 0) pdl_wait();
    if (!SELF) {
        const T* cbase = reinterpret_cast<const T*>(a.kv) + ((size_t)b * a.S * a.L + a.layer) * 2 * d + hoff;
        const size_t cstride = (size_t)a.L * 2 * d;
#pragma unroll
        for (int kv = 0; kv < 2; ++kv) {                              // group 0: K halves, group 1: V halves
#pragma unroll
            for (int k = 0; k < STAGE / 4; ++k) {
                const int row = sb0 + 4 * k + rsub;
                if ((row & ~31) < a.S) {
                    const bool ok = row < a.S;
                    const T* src = cbase + (ok ? (size_t)row * cstride + (size_t)kv * d + piece * 8 : (size_t)0);
                    const int nbytes = ok ? 16 : 0;
                    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(lane_dst + 4 * k * AM_ROWB + kv * 128), "l"(src), "r"(nbytes) : "memory");
                }
            }
            asm volatile("cp.async.commit_group;" ::: "memory");
        }
#pragma unroll
        for (int hv = 0; hv < AM_NCH; ++hv) {
            const int v = sb0 + 32 * hv + lane;
            mk0[hv] = (a.mask != nullptr && v < a.S) ? (int)a.mask[(size_t)b * a.S + v] : 1;
        }
        if (a.early_kv != 0) pdl_wait();
    }
    // ---- first round trip: everything whose address depends on the warp index only (control words, the query row,
    //      the row's new K,V) is requested before anything is consumed ----
    int4 c4 = make_int4(1, 0, 0, 0);                     // {step, done, steps_run, cur}
    if (a.ctl != nullptr) c4 = *reinterpret_cast<const int4*>(a.ctl);
    const int sdone = (a.sent_done != nullptr) ? a.sent_done[b] : 0;
    int n = a.S;
    if (!SELF && a.mem_len != nullptr) n = a.mem_len[b];
    float4 qx[4], kx[4], vx[4];
    {
        const float4* qp = reinterpret_cast<const float4*>(a.q + (size_t)(row0 + min(nl, nq - 1)) * a.ldq + hoff + 16 * c);
#pragma unroll
        for (int i = 0; i < 4; ++i) qx[i] = qp[i];
        if (SELF) {
#pragma unroll
            for (int i = 0; i < 4; ++i) { kx[i] = qp[d / 4 + i]; vx[i] = qp[d / 2 + i]; }
        }
    }
    // self: the ancestor bytes of positions [0,32) of the group's rows, for BOTH parities of the ancestor table (which
    // one is current is in the control word that is still in flight) -- short contexts need no further table reads
    int pa[2][4];
    if (SELF) {
#pragma unroll
        for (int par = 0; par < 2; ++par)
#pragma unroll
            for (int qi = 0; qi < 4; ++qi)
                pa[par][qi] = (lane < a.T_max) ? (int)a.anc[((size_t)par * a.M + row0 + min(qi, nq - 1)) * a.T_max + lane] : 0;
    }
    if (c4.y != 0 || sdone != 0) {
        asm volatile("cp.async.wait_all;" ::: "memory");
        return;
    }
    if (SELF) n = c4.x;                                  // self: positions 0..step-1

    // ---- Q fragments.  A-operand quad of k-slice ks = {a0: row nl, k 0-7 | a1: row nl+8, k 0-7 | a2: row nl, k 8-15 |
    //      a3: row nl+8, k 8-15}.  Rows 0..3 carry the hi 16-bit part of the 4 queries and rows 8..11 the lo part, so ONE
    //      mma yields hi.K in c0,c1 and lo.K in c2,c3 (no wasted accumulator rows, no zero registers).  Dims [16c,16c+16)
    //      of the query in k-slice order, scaled by 1/8 (exact). ----
    uint32_t aq[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        split2<T>(qx[i].x * 0.125f, qx[i].y * 0.125f, aq[i][0], aq[i][1]);
        split2<T>(qx[i].z * 0.125f, qx[i].w * 0.125f, aq[i][2], aq[i][3]);
    }
    if (nl >= nq) {
#pragma unroll
        for (int i = 0; i < 4; ++i) { aq[i][0] = 0u; aq[i][1] = 0u; aq[i][2] = 0u; aq[i][3] = 0u; }
    }

    // ---- geometry of the virtual position list ----
    const T* base;                       // element pointer of row index 0, this head
    size_t pos_stride;                   // elements between consecutive row indices
    int cshared = n, tl = 0, nv = n;     // self: [0,cshared) shared, then nq tails of tl positions each
    const unsigned char* anc0 = nullptr;
    int a0b[4] = {0, 0, 0, 0};           // self: ancestor slot of position `lane` for the group's 4 rows (current parity)
    if (SELF) {
        const int pc = n - 1;
        base = reinterpret_cast<const T*>(a.kv) + hoff;
        pos_stride = (size_t)2 * d;
        anc0 = a.anc + ((size_t)c4.w * a.M + row0) * a.T_max;
        // append the new K,V of the group's rows at position pc: lane (nl < nq, c) owns dims [16c,16c+16) of row nl
        if (nl < nq && wsub == 0) {
            T* dst = reinterpret_cast<T*>(a.kv) + ((size_t)(row0 + nl) * a.T_max + pc) * pos_stride + hoff + 16 * c;
            uint32_t w[8];
#pragma unroll
            for (int i = 0; i < 4; ++i) { w[2 * i] = Mma16<T>::pack(kx[i].x, kx[i].y); w[2 * i + 1] = Mma16<T>::pack(kx[i].z, kx[i].w); }
            reinterpret_cast<uint4*>(dst)[0] = make_uint4(w[0], w[1], w[2], w[3]);
            reinterpret_cast<uint4*>(dst)[1] = make_uint4(w[4], w[5], w[6], w[7]);
#pragma unroll
            for (int i = 0; i < 4; ++i) { w[2 * i] = Mma16<T>::pack(vx[i].x, vx[i].y); w[2 * i + 1] = Mma16<T>::pack(vx[i].z, vx[i].w); }
            reinterpret_cast<uint4*>(dst + d)[0] = make_uint4(w[0], w[1], w[2], w[3]);
            reinterpret_cast<uint4*>(dst + d)[1] = make_uint4(w[4], w[5], w[6], w[7]);
        }
        // first position whose physical row differs between the group's slots
#pragma unroll
        for (int qi = 0; qi < 4; ++qi) a0b[qi] = (c4.w != 0) ? pa[1][qi] : pa[0][qi];
        cshared = pc;
        if (nq > 1) {
            for (int blk = 0; blk * 32 < pc; ++blk) {
                const int p = blk * 32 + lane;
                bool same = true;
                if (p < pc) {
                    if (blk == 0) {
                        same = (a0b[1] == a0b[0]) && (a0b[2] == a0b[0]) && (a0b[3] == a0b[0]);   // rows >= nq mirror row nq-1
                    } else {
                        const int a0 = anc0[p];
                        for (int qi = 1; qi < nq; ++qi) same = same && (anc0[(size_t)qi * a.T_max + p] == a0);
                    }
                }
                const unsigned bal = __ballot_sync(0xffffffffu, !same);
                if (bal != 0u) { cshared = blk * 32 + __ffs(bal) - 1; break; }
            }
        }
        tl = pc - cshared + 1;
        nv = cshared + nq * tl;
        if (SPLIT) asm volatile("bar.sync %0, %1;" ::"r"(1 + team), "r"(32 * NSPLIT) : "memory");   // the appended rows are read back
        else __syncwarp();                                                                           // below by other lanes (of the team)
    } else {
        base = reinterpret_cast<const T*>(a.kv) + ((size_t)b * a.S * a.L + a.layer) * 2 * d + hoff;
        pos_stride = (size_t)a.L * 2 * d;
    }
    const unsigned char* mrow = (!SELF && a.mask != nullptr) ? a.mask + (size_t)b * a.S : nullptr;

    float o[8][4];                       // O fragments: tile nt -> dims 16c+nt (o[nt][0]) and 16c+8+nt (o[nt][1]) of row nl
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) { o[nt][0] = 0.f; o[nt][1] = 0.f; o[nt][2] = 0.f; o[nt][3] = 0.f; }
    float m_run = -INFINITY, l_run = 0.f;
    const int qrow = nl & 3;             // rows 4..7 mirror rows 0..3 (their results are discarded)

    // copy loop constants: a row offset travels as ONE 32-bit count of 16-byte units (row index x stride16), so a copy
    // costs a shuffle, one wide multiply-add and the cp.async
    constexpr uint32_t kNoRow = 0xFFFFFFFFu;
    const uint32_t stride16 = (uint32_t)(pos_stride * sizeof(T) / 16);
    const unsigned char* lane_src = reinterpret_cast<const unsigned char*>(base) + piece * 16;
    const size_t v_off = (size_t)d * sizeof(T);                      // V half of a cache row, in bytes

    // virtual positions sb + 32*hv + lane: row offset and meta = owner mask (bits 0-3) | masked-source (bit 4)
    auto build_list = [&](int sb, uint32_t (&rowidx)[AM_NCH], int (&meta)[AM_NCH]) {
#pragma unroll
        for (int hv = 0; hv < AM_NCH; ++hv) {
            rowidx[hv] = kNoRow; meta[hv] = 0;
            const int v = sb + 32 * hv + lane;
            if (SELF) {
                const bool in = v < nv, tail = in && v >= cshared;
                const int vt = v - cshared;
                const int qi = tail ? (vt >= tl) + (vt >= 2 * tl) + (vt >= 3 * tl) : 0;
                const int p = tail ? cshared + vt - qi * tl : v;                 // cache position of this virtual position
                // ancestor slot of (row qi, position p): in registers for p < 32, in the table above that
                const int s0 = __shfl_sync(0xffffffffu, a0b[0], p & 31), s1 = __shfl_sync(0xffffffffu, a0b[1], p & 31);
                const int s2 = __shfl_sync(0xffffffffu, a0b[2], p & 31), s3 = __shfl_sync(0xffffffffu, a0b[3], p & 31);
                int slot = (qi == 0) ? s0 : (qi == 1) ? s1 : (qi == 2) ? s2 : s3;
                if (in && p >= 32) slot = anc0[(size_t)qi * a.T_max + p];
                if (tail && p == n - 1) slot = q0 + qi;                          // the row appended above
                if (in) { rowidx[hv] = (uint32_t)((b * a.beam + slot) * a.T_max + p) * stride16; meta[hv] = tail ? (1 << qi) : 0xF; }
            } else if (v < nv) {
                rowidx[hv] = (uint32_t)v * stride16;
                const int mk = (sb == sb0) ? mk0[hv] : ((mrow != nullptr) ? (int)mrow[v] : 1);
                meta[hv] = 0xF | ((mk == 0) ? 0x10 : 0);
            }
        }
    };
    // one cp.async group: the K (kv = 0) or V (kv = 1) halves of the rows of a stage's first `nch` chunks
    auto issue_half = [&](const uint32_t (&rowidx)[AM_NCH], int nch, int kv) {
#pragma unroll
        for (int k = 0; k < STAGE / 4; ++k) {
            if (k < 8 * nch) {
                const uint32_t o = __shfl_sync(0xffffffffu, rowidx[k >> 3], (4 * k + rsub) & 31);
                const bool ok = o != kNoRow;
                const unsigned char* src = lane_src + (size_t)(ok ? o : 0u) * 16 + (kv != 0 ? v_off : (size_t)0);
                const int nbytes = ok ? 16 : 0;
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(lane_dst + 4 * k * AM_ROWB + kv * 128), "l"(src), "r"(nbytes) : "memory");
            }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };

    // ---- software pipeline over the stages.  Groups are committed in the order K0 V0 K1 V1 ...; K(s+1) is requested
    //      right after the last scores of stage s (the K half is free then), V(s+1) after its last P.V, so a warp keeps
    //      copies in flight while it computes.  The position list of stage s+1 is built at the top of stage s. ----
    uint32_t rowidx[AM_NCH], rowidx_n[AM_NCH];
    int meta[AM_NCH], meta_n[AM_NCH];
    constexpr int STEP = NSPLIT * STAGE;                             // distance between two stages of one warp
    build_list(sb0, rowidx, meta);
    if (SELF) {                                                       // cross: this warp's first stage was requested at the top
        const int nch0 = min(AM_NCH, (nv - sb0 + 31) >> 5);          // <= 0: this warp has no stage (nothing is copied)
        issue_half(rowidx, nch0, 0);
        issue_half(rowidx, nch0, 1);
    }
    for (int sb = sb0; sb < nv; sb += STEP) {
        const int nchunk = min(AM_NCH, (nv - sb + 31) >> 5);         // 32-position chunks in this stage (warp-uniform)
        const bool has_next = sb + STEP < nv;
        const int nchunk_n = min(AM_NCH, (nv - sb - STEP + 31) >> 5);
        if (has_next) build_list(sb + STEP, rowidx_n, meta_n);
        asm volatile("cp.async.wait_group 1;" ::: "memory");        // K of this stage has landed (V may still be in flight)
        __syncwarp();
        bool k_next = false;

#pragma unroll
        for (int cb = 0; cb < AM_NCH; ++cb) {
            if (cb < nchunk) {
                const unsigned char* wc = ws + cb * 32 * AM_ROWB;
                // ---- S = Q.K^T for 4 tiles of 8 positions ----
                float s[4][4];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    s[j][0] = 0.f; s[j][1] = 0.f; s[j][2] = 0.f; s[j][3] = 0.f;
                    const uint4* kp = reinterpret_cast<const uint4*>(wc + (8 * j + nl) * AM_ROWB + 32 * c);
                    const uint4 k0 = kp[0], k1 = kp[1];
                    const uint32_t kw[8] = {k0.x, k0.y, k0.z, k0.w, k1.x, k1.y, k1.z, k1.w};
#pragma unroll
                    for (int ks = 0; ks < 4; ++ks) Mma16<T>::mma(s[j], aq[ks], kw[2 * ks], kw[2 * ks + 1]);
                    s[j][0] += s[j][2];                      // hi.K (rows 0..3) + lo.K (rows 8..11)
                    s[j][1] += s[j][3];
                }
                // ---- mask, online softmax on the C fragments (row qrow, positions 8j + 2c + {0,1} of the chunk) ----
                float mc = -INFINITY;
#pragma unroll
                for (int j = 0; j < 4; ++j)
#pragma unroll
                    for (int e = 0; e < 2; ++e) {
                        const int mt = __shfl_sync(0xffffffffu, meta[cb], 8 * j + 2 * c + e);
                        float x = s[j][e];
                        if (mt & 0x10) x += -10000.0f;       // additive mask (1 - m) * -10000: reference src/bert_seq2seq.py:36-37
                        x = ((mt >> qrow) & 1) ? x : -INFINITY;
                        s[j][e] = x;
                        mc = fmaxf(mc, x);
                    }
                mc = fmaxf(mc, __shfl_xor_sync(0xffffffffu, mc, 1));
                mc = fmaxf(mc, __shfl_xor_sync(0xffffffffu, mc, 2));
                const float m_new = fmaxf(m_run, mc);
                const float m_use = (m_new == -INFINITY) ? 0.f : m_new;  // nothing visible yet for this row: all p = 0
                const float corr = __expf(m_run - m_use);                // m_run = -inf -> 0
                m_run = m_new;
                float psum = 0.f;
                uint32_t ap[2][4];                           // P as A quads: hi in rows 0..3, lo in rows 8..11 (as for Q)
#pragma unroll
                for (int t = 0; t < 2; ++t) {
                    const float p0 = __expf(s[2 * t][0] - m_use), p1 = __expf(s[2 * t][1] - m_use);
                    const float p2 = __expf(s[2 * t + 1][0] - m_use), p3 = __expf(s[2 * t + 1][1] - m_use);
                    psum += (p0 + p1) + (p2 + p3);
                    split2<T>(p0, p1, ap[t][0], ap[t][1]);
                    split2<T>(p2, p3, ap[t][2], ap[t][3]);
                }
                l_run = l_run * corr + psum;
                if (corr != 1.0f) {
#pragma unroll
                    for (int nt = 0; nt < 8; ++nt) { o[nt][0] *= corr; o[nt][1] *= corr; o[nt][2] *= corr; o[nt][3] *= corr; }
                }
                if (cb == nchunk - 1 && has_next) {          // the K half is free: request K of the next stage
                    __syncwarp();
                    issue_half(rowidx_n, nchunk_n, 0);
                    k_next = true;
                }
                if (cb == 0) {                               // V of this stage (everything older than the group just issued)
                    if (k_next) asm volatile("cp.async.wait_group 1;" ::: "memory");
                    else asm volatile("cp.async.wait_group 0;" ::: "memory");
                    __syncwarp();
                }
                // ---- O += P.V : rows 16t + {2c, 2c+1, 2c+8, 2c+9}, dims [8nl, 8nl+8) ----
#pragma unroll
                for (int t = 0; t < 2; ++t) {
                    const unsigned char* vp = wc + (16 * t + 2 * c) * AM_ROWB + 128 + 16 * nl;
                    const uint4 v0 = *reinterpret_cast<const uint4*>(vp), v1 = *reinterpret_cast<const uint4*>(vp + AM_ROWB);
                    const uint4 v2 = *reinterpret_cast<const uint4*>(vp + 8 * AM_ROWB), v3 = *reinterpret_cast<const uint4*>(vp + 9 * AM_ROWB);
                    const uint32_t va[4] = {v0.x, v0.y, v0.z, v0.w}, vb_[4] = {v1.x, v1.y, v1.z, v1.w};
                    const uint32_t vc[4] = {v2.x, v2.y, v2.z, v2.w}, vd[4] = {v3.x, v3.y, v3.z, v3.w};
#pragma unroll
                    for (int nt = 0; nt < 8; ++nt) {
                        const uint32_t sel = (nt & 1) ? 0x7632u : 0x5410u;
                        const uint32_t b0 = __byte_perm(va[nt >> 1], vb_[nt >> 1], sel);
                        const uint32_t b1 = __byte_perm(vc[nt >> 1], vd[nt >> 1], sel);
                        Mma16<T>::mma(o[nt], ap[t], b0, b1);
                    }
                }
            }
        }
        __syncwarp();                    // the V half is free: request V of the next stage
        if (has_next) {
            issue_half(rowidx_n, nchunk_n, 1);
#pragma unroll
            for (int hv = 0; hv < AM_NCH; ++hv) { rowidx[hv] = rowidx_n[hv]; meta[hv] = meta_n[hv]; }
        }
    }
    asm volatile("cp.async.wait_all;" ::: "memory");     // nothing may stay in flight (empty position lists)
    // ---- normalise and store: lane (nl < nq, c) holds dims [16c,16c+16) of row nl ----
    l_run += __shfl_xor_sync(0xffffffffu, l_run, 1);
    l_run += __shfl_xor_sync(0xffffffffu, l_run, 2);
    // this lane's 16 output dims, unnormalised (hi + lo accumulator rows): f[i] / f[4 + i] pair up as in the stores below
    float f[16];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        f[2 * i] = o[2 * i][0] + o[2 * i][2];     f[2 * i + 1] = o[2 * i + 1][0] + o[2 * i + 1][2];
        f[8 + 2 * i] = o[2 * i][1] + o[2 * i][3]; f[8 + 2 * i + 1] = o[2 * i + 1][1] + o[2 * i + 1][3];
    }
    if (SPLIT) {
        // every warp publishes {m, l, f[16]} of its lanes in its own (now idle) staging buffer; warp 0 merges
        __syncwarp();
        float* mine = reinterpret_cast<float*>(ws) + lane * 18;
        mine[0] = m_run; mine[1] = l_run;
#pragma unroll
        for (int i = 0; i < 16; ++i) mine[2 + i] = f[i];
        asm volatile("bar.sync %0, %1;" ::"r"(1 + team), "r"(32 * NSPLIT) : "memory");
        if (wsub != 0) return;
        const unsigned char* team_ws = am_smem + (size_t)(team * NSPLIT) * (STAGE * AM_ROWB);
        float mm = m_run;
#pragma unroll
        for (int w2 = 1; w2 < NSPLIT; ++w2)
            mm = fmaxf(mm, reinterpret_cast<const float*>(team_ws + (size_t)w2 * (STAGE * AM_ROWB))[lane * 18]);
        const float mu = (mm == -INFINITY) ? 0.f : mm;
        const float c0s = __expf(m_run - mu);                             // -inf -> 0: a warp that saw nothing visible
        l_run *= c0s;
#pragma unroll
        for (int i = 0; i < 16; ++i) f[i] *= c0s;
#pragma unroll
        for (int w2 = 1; w2 < NSPLIT; ++w2) {
            const float* oth = reinterpret_cast<const float*>(team_ws + (size_t)w2 * (STAGE * AM_ROWB)) + lane * 18;
            const float cs = __expf(oth[0] - mu);
            l_run += oth[1] * cs;
#pragma unroll
            for (int i = 0; i < 16; ++i) f[i] += oth[2 + i] * cs;
        }
    }
    if (nl < nq) {
        const float inv = 1.0f / l_run;
        uint32_t w[8];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            w[i] = Mma16<T>::pack(f[2 * i] * inv, f[2 * i + 1] * inv);
            w[4 + i] = Mma16<T>::pack(f[8 + 2 * i] * inv, f[8 + 2 * i + 1] * inv);
        }
        uint4* dp = reinterpret_cast<uint4*>(reinterpret_cast<T*>(a.ctx) + (size_t)(row0 + nl) * d + hoff + 16 * c);
        dp[0] = make_uint4(w[0], w[1], w[2], w[3]);
        dp[1] = make_uint4(w[4], w[5], w[6], w[7]);
    }
}

template <typename T, int STAGE>
static int launch_attn_mma_stage(const AttnMmaArgs& a, bool self, cudaStream_t s) {
    const dim3 grid((a.n_warps + AM_WARPS - 1) / AM_WARPS), block(AM_THREADS);
    static bool attr_self[ITB_MAX_DEVICES] = {}, attr_cross[ITB_MAX_DEVICES] = {};
    {
        cudaError_t e1 = itb_smem_attr_once(attr_self, attn_mma_kernel<T, true, STAGE>, AM_SMEM(STAGE));
        cudaError_t e2 = itb_smem_attr_once(attr_cross, attn_mma_kernel<T, false, STAGE>, AM_SMEM(STAGE));
        if (e1 != cudaSuccess || e2 != cudaSuccess) { itb_set_error("attention: cudaFuncSetAttribute failed"); return -1; }
    }
    cudaError_t e = self ? itb_launch(attn_mma_kernel<T, true, STAGE>, grid, block, (size_t)AM_SMEM(STAGE), s, a)
                         : itb_launch(attn_mma_kernel<T, false, STAGE>, grid, block, (size_t)AM_SMEM(STAGE), s, a);
    if (e != cudaSuccess) { itb_set_error("%s: launch failed: %s", self ? "self_attention" : "cross_attention", cudaGetErrorString(e)); return -1; }
    return itb_check_launch(self ? "self_attention" : "cross_attention");
}

// Split launch: NSPLIT warps per item, AM_WARPS / NSPLIT items per CTA
template <typename T, int NSPLIT>
static int launch_attn_mma_split(const AttnMmaArgs& a, bool self, cudaStream_t s) {
    constexpr int kItems = AM_WARPS / NSPLIT;
    const dim3 grid((a.n_warps + kItems - 1) / kItems), block(AM_THREADS);
    static bool attr_self[ITB_MAX_DEVICES] = {}, attr_cross[ITB_MAX_DEVICES] = {};
    {
        cudaError_t e1 = itb_smem_attr_once(attr_self, attn_mma_kernel<T, true, 64, NSPLIT>, AM_SMEM(64));
        cudaError_t e2 = itb_smem_attr_once(attr_cross, attn_mma_kernel<T, false, 64, NSPLIT>, AM_SMEM(64));
        if (e1 != cudaSuccess || e2 != cudaSuccess) { itb_set_error("attention: cudaFuncSetAttribute failed"); return -1; }
    }
    cudaError_t e = self ? itb_launch(attn_mma_kernel<T, true, 64, NSPLIT>, grid, block, (size_t)AM_SMEM(64), s, a)
                         : itb_launch(attn_mma_kernel<T, false, 64, NSPLIT>, grid, block, (size_t)AM_SMEM(64), s, a);
    if (e != cudaSuccess) { itb_set_error("%s: launch failed: %s", self ? "self_attention" : "cross_attention", cudaGetErrorString(e)); return -1; }
    return itb_check_launch(self ? "self_attention" : "cross_attention");
}

template <typename T>
static int launch_attn_mma(AttnMmaArgs a, bool self, cudaStream_t s) {
    const int B = a.M / a.beam, H = a.d / DH;
    a.n_warps = B * H * ((a.beam + 3) / 4);
    // Small launches: the step waits for the longest walk of a single warp while most of the GPU idles, so an item's stages
    // are shared by four warps (up to one wave of one-item CTAs, 3 CTAs per SM) or two (up to one wave of two-item CTAs).
    // ITB_ATTN_SPLIT = item count up to which the 4-way split applies, the 2-way split up to twice that (0: never).
    static int split_max = -1;
    if (split_max < 0) {
        const char* e = getenv("ITB_ATTN_SPLIT");
        split_max = (e != nullptr) ? atoi(e) : 3 * itb_num_sms();
    }
    static int forced = -1;              // ITB_ATTN_STAGE = 32 / 64: that unsplit build for every launch (tests, A/B timing)
    if (forced < 0) {
        const char* e = getenv("ITB_ATTN_STAGE");
        forced = (e != nullptr) ? atoi(e) : 0;
    }
    if (forced == 0 && a.n_warps <= split_max) return launch_attn_mma_split<T, 4>(a, self, s);
    if (forced == 0 && a.n_warps <= 2 * split_max) return launch_attn_mma_split<T, 2>(a, self, s);
    // more than ~2 waves of the 3-CTA build (148 SMs x 12 warps): occupancy beats round trips (measured at 4096 rows:
    // step -28 us; at 244 rows the 64-position build is 25 us per step faster)
    const bool small_stage = forced != 0 ? forced == 32 : a.n_warps > 2 * itb_num_sms() * 12;
    return small_stage ? launch_attn_mma_stage<T, 32>(a, self, s) : launch_attn_mma_stage<T, 64>(a, self, s);
}

bool itb_attn_bulk_applies(bool self, int M, int beam, int d, int T_max, int S);
int itb_attn_bulk(int dtype, bool self, const void* ctl, const int* sent_done, const int* mem_len, const float* q, int ldq,
                  void* kv, const unsigned char* anc, const unsigned char* mask, void* ctx, int M, int beam, int d, int T_max,
                  int S, int L, int layer, cudaStream_t s);

// Called by itb_self_attention / itb_cross_attention (attention.cu) for the 16-bit dtypes.  Large launches go to the
// bulk-copy producer / consumer kernel (attention_bulk.cu), small ones stay on the warp-per-head kernel above.
int itb_attn_mma(int dtype, bool self, const void* ctl, const int* sent_done, const int* mem_len, const float* q, int ldq,
                 void* kv, const unsigned char* anc, const unsigned char* mask, void* ctx, int M, int beam, int d, int T_max,
                 int S, int L, int layer, int early_kv, cudaStream_t s) {
    if (itb_attn_bulk_applies(self, M, beam, d, T_max, S))
        return itb_attn_bulk(dtype, self, ctl, sent_done, mem_len, q, ldq, kv, anc, mask, ctx, M, beam, d, T_max, S, L, layer, s);
    AttnMmaArgs a = {};
    a.early_kv = early_kv;
    a.ctl = reinterpret_cast<const ItbCtl*>(ctl);
    a.sent_done = sent_done; a.mem_len = mem_len; a.q = q; a.ldq = ldq; a.kv = kv; a.anc = anc; a.mask = mask; a.ctx = ctx;
    a.M = M; a.beam = beam; a.d = d; a.T_max = T_max; a.S = S; a.L = L; a.layer = layer;
    if (dtype == ITB_BF16) return launch_attn_mma<__nv_bfloat16>(a, self, s);
    return launch_attn_mma<__half>(a, self, s);
}
