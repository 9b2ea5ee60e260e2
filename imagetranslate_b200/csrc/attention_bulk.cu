// Single-query attention for large launches of the 16-bit decode modes: bulk-copy staged, producer / consumer.
// Same arithmetic as attention_mma.cu (HF BertSelfAttention on the last position; head_dim 64;
// fp32 queries split hi + lo, mma.sync.m16n8k16, online softmax in fp32), different data movement:
//
//   * one PERSISTENT CTA per SM; a work item is (sentence, group of <= 4 beam slots) -- ALL heads at once;
//   * one PRODUCER warp walks the item's virtual position list (cross-attention: the memory positions of the sentence;
//     self-attention: the prefix the beam slots share ONCE + each slot's divergent tail, through the ancestor table) and
//     issues ONE bulk copy per position: `cp.async.bulk.shared.global` of the whole [K | V] cache row of that position
//     (2 * d * 2 B = 3 KB, contiguous in both cache layouts) into a ring of stages of 16 positions, completion counted on an
//     mbarrier (complete_tx).  The producer runs ahead of the consumers across items, so K,V of the next sentences are in
//     flight while the current one is computed -- the copy engine, not the warps' own load instructions, keeps HBM busy;
//   * H CONSUMER warps (one per head) wait for a stage, run Q.K^T / softmax / P.V for their 128-byte slices of the staged
//     rows, and release the stage (mbarrier arrive).  Rows are staged 16 bytes apart from a multiple of 128 (row pitch
//     2*d*2 + 16), which makes the fragment reads of all heads bank-conflict free.
//   * self-attention: the rows' NEW K,V (position step-1) are appended to the cache by the consumers and enter the
//     softmax straight from registers (they are never read back through the copy engine).
// Per stage the producer publishes the positions' owner masks / source-mask bits and the stage header in shared memory
// before it arms the barrier.  Stage slots without a position keep stale (finite) data and are masked by their zero
// owner mask; the ring is zeroed once per kernel so that nothing non-finite can ever sit in a V slot.
// HBM-bound by design: 2*d*2 bytes per (virtual position, layer), every byte requested exactly once per item.
#include <stdlib.h>
#include "common.cuh"
#include "launch.h"
#include "mma16.cuh"

#define DH 64
#define AB_CH 16               // virtual positions per stage
#define AB_MAX_STAGES 4
#define AB_LOOKAHEAD 16       // producer look-ahead registers per row: positions lane + 32k, k < 16 (T_max, S <= 512)
#define AB_SPIN_LIMIT (1u << 28)
#ifndef AB_QSTAGE
#define AB_QSTAGE 0            // 1: query rows staged by the producer's copy engine one item ahead; 0: consumers load them
                               // (measured on B200: no gain -- the extra stage the ring loses costs what the staging saves)
#endif

struct AttnBulkArgs {
    const ItbCtl* ctl;
    const int* sent_done;                // [B] or null
    const int* mem_len;                  // cross: [B] or null (= S)
    const float* q; int ldq;             // f32 queries (self: qkv buffer, k at +d, v at +2d)
    void* kv;                            // self: [M][T_max][2][d]; cross: [B][S][L][2][d]
    const unsigned char* anc;            // self: [2][M][T_max]
    const unsigned char* mask;           // cross: [B][S] or null
    void* ctx;                           // act [M][d]
    int M, beam, d, T_max, S, L, layer;
    int n_items, nst, rowb, anc_cap;     // work items, ring stages, staged row pitch (bytes), bytes per ancestor row in smem
};

__device__ __forceinline__ uint32_t ab_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void ab_mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(ab_smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void ab_mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(ab_smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void ab_mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(ab_smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void ab_mbar_wait(uint64_t* bar, uint32_t parity) {
    uint32_t done = 0, spins = 0;
    const uint32_t addr = ab_smem_u32(bar);
    while (true) {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done) : "r"(addr), "r"(parity) : "memory");
        if (done) break;
        if (++spins > AB_SPIN_LIMIT) __trap();     // a protocol bug must fault, never hang the GPU
    }
}
// 1-D bulk copy global -> shared, completion (bytes) signalled on `bar`
__device__ __forceinline__ void ab_bulk_load(uint32_t smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_dst), "l"(gsrc), "r"(bytes), "r"(ab_smem_u32(bar)) : "memory");
}

template <typename T, bool SELF>
__global__ void __launch_bounds__(416, 1) attn_bulk_kernel(AttnBulkArgs a) {
    extern __shared__ __align__(128) unsigned char ab_smem[];
    pdl_trigger();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int d = a.d, H = d / DH, nz = (a.beam + 3) >> 2;
    const int rowbytes = 2 * d * (int)sizeof(T);             // [K | V] of one position
    const int stage_bytes = AB_CH * a.rowb, ring_bytes = a.nst * stage_bytes;
    // query rows of the item the consumers will work on NEXT (fp32; self: the whole [q | k | v] row of the QKV buffer),
    // fetched by the producer's copy engine one item ahead, so the consumers never wait on a global load of their own
    const int qrow_bytes = AB_QSTAGE ? (SELF ? 3 : 1) * d * (int)sizeof(float) : 0;
    unsigned char* qbuf = ab_smem + ring_bytes;                               // [4][qrow_bytes]
    uint64_t* full = reinterpret_cast<uint64_t*>(qbuf + 4 * qrow_bytes);
    uint64_t* empty = full + AB_MAX_STAGES;
    uint64_t* qfull = empty + AB_MAX_STAGES;
    uint64_t* qempty = qfull + 1;
    int* meta = reinterpret_cast<int*>(qempty + 1);                           // [AB_MAX_STAGES][AB_CH]
    int* hdr = meta + AB_MAX_STAGES * AB_CH;                                  // [AB_MAX_STAGES] positions | last << 8
    unsigned char* aux_s = reinterpret_cast<unsigned char*>(hdr + AB_MAX_STAGES);   // self: ancestor bytes [4][anc_cap]; cross: mask bytes [S]
    if (threadIdx.x == 0) {
        for (int i = 0; i < a.nst; ++i) { ab_mbar_init(&full[i], 1); ab_mbar_init(&empty[i], (uint32_t)H); }
        ab_mbar_init(qfull, 1);
        ab_mbar_init(qempty, (uint32_t)H);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    for (int i = threadIdx.x; i < ring_bytes / 16; i += blockDim.x) reinterpret_cast<uint4*>(ab_smem)[i] = make_uint4(0u, 0u, 0u, 0u);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");      // generic-proxy zeroes before async-proxy copies
    __syncthreads();
    pdl_wait();
    int4 c4 = make_int4(1, 0, 0, 0);                      // {step, done, steps_run, cur}
    if (a.ctl != nullptr) c4 = *reinterpret_cast<const int4*>(a.ctl);
    if (c4.y != 0) return;
    const uint32_t ring_u32 = ab_smem_u32(ab_smem);
    const int pc = c4.x - 1;                              // self: positions 0..step-1, the new token sits at pc

    if (warp == H) {
        // ------------------------------------------------------------------ producer
        // Everything the producer reads from global memory for an item -- the sentence's group flag, its memory length and
        // source-mask bytes (cross) or the ancestor bytes of the group's rows (self) -- is requested ONE ITEM AHEAD into
        // registers (positions lane + 32k), so those round trips overlap the copies of the current item instead of
        // standing between two stages.
        constexpr int KM = AB_LOOKAHEAD;
        unsigned char rb[SELF ? 4 * KM : KM];
        int r_skip = 0, r_n = 0;
        auto prefetch = [&](int item) {
            const int b = item / nz, z = item - b * nz;
            r_skip = (a.sent_done != nullptr) ? a.sent_done[b] : 0;
            if (SELF) {
                const int q0 = z * 4, row0 = b * a.beam + q0, nq = min(4, a.beam - q0);
                const unsigned char* anc0 = a.anc + ((size_t)c4.w * a.M + row0) * a.T_max;
#pragma unroll
                for (int k = 0; k < KM; ++k) {
                    if (32 * k < pc) {                    // warp-uniform: short contexts touch one or two blocks only
                        const int p = lane + 32 * k;
#pragma unroll
                        for (int qi = 0; qi < 4; ++qi)
                            rb[qi * KM + k] = (p < pc) ? anc0[(size_t)min(qi, nq - 1) * a.T_max + p] : (unsigned char)0;
                    }
                }
            } else {
                r_n = (a.mem_len != nullptr) ? a.mem_len[b] : a.S;
#pragma unroll
                for (int k = 0; k < KM; ++k) {
                    if (32 * k < a.S) {
                        const int p = lane + 32 * k;
                        rb[k] = (a.mask != nullptr && p < a.S) ? a.mask[(size_t)b * a.S + p] : (unsigned char)1;
                    }
                }
            }
        };
        uint32_t seq = 0, nit = 0;
        if ((int)blockIdx.x < a.n_items) prefetch(blockIdx.x);
        for (int item = blockIdx.x; item < a.n_items; item += gridDim.x) {
            const int b = item / nz, z = item - b * nz;
            const int q0 = z * 4, nq = min(4, a.beam - q0);
            // this item's bytes: registers -> shared memory (aux_s), then the next item's loads are put in flight
            const int skip = r_skip;
            int nv = r_n, cshared = 0, tl = 0;
            __syncwarp();
            if (SELF) {
#pragma unroll
                for (int k = 0; k < KM; ++k) {
                    if (32 * k < pc) {
#pragma unroll
                        for (int qi = 0; qi < 4; ++qi)
                            if (lane + 32 * k < pc) aux_s[qi * a.anc_cap + lane + 32 * k] = rb[qi * KM + k];
                    }
                }
            } else {
#pragma unroll
                for (int k = 0; k < KM; ++k)
                    if (32 * k < a.S && lane + 32 * k < a.S) aux_s[lane + 32 * k] = rb[k];
            }
            __syncwarp();
            if (item + (int)gridDim.x < a.n_items) prefetch(item + gridDim.x);
            if (skip != 0) continue;
            if (AB_QSTAGE) {   // the item's query rows: one bulk copy per row, as soon as the consumers have taken the previous item's
                const int row0 = b * a.beam + q0;
                ab_mbar_wait(qempty, (nit & 1u) ^ 1u);
                if (lane == 0) ab_mbar_expect_tx(qfull, (uint32_t)(nq * qrow_bytes));
                __syncwarp();
                if (lane < nq)
                    ab_bulk_load(ab_smem_u32(qbuf + lane * qrow_bytes), a.q + (size_t)(row0 + lane) * a.ldq, (uint32_t)qrow_bytes, qfull);
                ++nit;
            }
            const unsigned char* base;                    // byte address of row index 0
            size_t pitch;                                 // bytes between consecutive row indices
            if (SELF) {
                base = reinterpret_cast<const unsigned char*>(a.kv);
                pitch = (size_t)rowbytes;
                cshared = pc;                             // first position whose physical row differs between the slots
                if (nq > 1) {
                    for (int blk = 0; blk * 32 < pc; ++blk) {
                        const int p = blk * 32 + lane;
                        bool same = true;
                        if (p < pc) {
                            const int a0 = aux_s[p];
                            for (int qi = 1; qi < nq; ++qi) same = same && (aux_s[qi * a.anc_cap + p] == a0);
                        }
                        const unsigned bal = __ballot_sync(0xffffffffu, !same);
                        if (bal != 0u) { cshared = blk * 32 + __ffs(bal) - 1; break; }
                    }
                }
                tl = pc - cshared;
                nv = cshared + nq * tl;
            } else {
                base = reinterpret_cast<const unsigned char*>(a.kv) + ((size_t)b * a.S * a.L + a.layer) * rowbytes;
                pitch = (size_t)a.L * rowbytes;
            }
            const int nstage = max(1, (nv + AB_CH - 1) / AB_CH);
            for (int sidx = 0; sidx < nstage; ++sidx, ++seq) {
                const int st = (int)(seq % (uint32_t)a.nst);
                const uint32_t ph = (seq / (uint32_t)a.nst) & 1u;
                ab_mbar_wait(&empty[st], ph ^ 1u);        // the consumers are done with what this slot held
                const int v = sidx * AB_CH + lane;
                const unsigned char* src = nullptr;
                int mt = 0;
                if (lane < AB_CH && v < nv) {
                    if (SELF) {
                        const bool tail = v >= cshared;
                        const int vt = v - cshared;
                        const int qi = tail ? vt / tl : 0;
                        const int p = tail ? cshared + vt - qi * tl : v;
                        const int slot = aux_s[qi * a.anc_cap + p];
                        src = base + ((size_t)(b * a.beam + slot) * a.T_max + p) * pitch;
                        mt = tail ? (1 << qi) : 0xF;
                    } else {
                        src = base + (size_t)v * pitch;
                        mt = 0xF | ((aux_s[v] == 0) ? 0x10 : 0);
                    }
                }
                if (lane < AB_CH) meta[st * AB_CH + lane] = mt;
                const int npos = max(0, min(AB_CH, nv - sidx * AB_CH));
                if (lane == 0) hdr[st] = npos | ((sidx == nstage - 1) ? 0x100 : 0);
                __syncwarp();                             // meta / hdr of all lanes precede the arming arrive (release)
                if (lane == 0) {
                    if (npos > 0) ab_mbar_expect_tx(&full[st], (uint32_t)(npos * rowbytes));
                    else ab_mbar_arrive(&full[st]);
                }
                __syncwarp();
                if (src != nullptr) ab_bulk_load(ring_u32 + (uint32_t)(st * stage_bytes + lane * a.rowb), src, (uint32_t)rowbytes, &full[st]);
            }
        }
        return;
    }
    if (warp > H) return;

    // ---------------------------------------------------------------------- consumers: warp = head
    const int h = warp, nl = lane >> 2, c = lane & 3, qrow = nl & 3;
    const int hoff = h * DH;
    const int koff = h * (DH * (int)sizeof(T)), voff = d * (int)sizeof(T) + koff;     // byte offsets inside a staged row
    uint32_t seq = 0, nit = 0;
    // cross-attention: the queries of the NEXT item are requested before this item's stages are consumed (their round
    // trip hides behind the stage loop); qn_item says which item the registers hold
    float4 qn[4];
    int qn_item = -1;
    for (int item = blockIdx.x; item < a.n_items; item += gridDim.x) {
        const int b = item / nz, z = item - b * nz;
        if (a.sent_done != nullptr && a.sent_done[b] != 0) continue;
        const int q0 = z * 4, row0 = b * a.beam + q0, nq = min(4, a.beam - q0);
        float4 qx[4], kx[4], vx[4];
        if (AB_QSTAGE) {
            ab_mbar_wait(qfull, nit & 1u);
            const float4* qp = reinterpret_cast<const float4*>(qbuf + min(nl, nq - 1) * qrow_bytes) + (hoff + 16 * c) / 4;
#pragma unroll
            for (int i = 0; i < 4; ++i) qx[i] = qp[i];
            if (SELF) {
#pragma unroll
                for (int i = 0; i < 4; ++i) { kx[i] = qp[d / 4 + i]; vx[i] = qp[d / 2 + i]; }
            }
            __syncwarp();                                // every lane holds its values: the buffer may be refilled
            if (lane == 0) ab_mbar_arrive(qempty);
            ++nit;
        } else {
            const float4* qp = reinterpret_cast<const float4*>(a.q + (size_t)(row0 + min(nl, nq - 1)) * a.ldq + hoff + 16 * c);
            if (!SELF && qn_item == item) {
#pragma unroll
                for (int i = 0; i < 4; ++i) qx[i] = qn[i];
            } else {
#pragma unroll
                for (int i = 0; i < 4; ++i) qx[i] = qp[i];
            }
            if (SELF) {
#pragma unroll
                for (int i = 0; i < 4; ++i) { kx[i] = qp[d / 4 + i]; vx[i] = qp[d / 2 + i]; }
            } else {
                const int nx = item + (int)gridDim.x;
                if (nx < a.n_items) {
                    const int bn = nx / nz, zn = nx - bn * nz;
                    const int rown = bn * a.beam + zn * 4 + min(nl, min(4, a.beam - zn * 4) - 1);
                    const float4* qpn = reinterpret_cast<const float4*>(a.q + (size_t)rown * a.ldq + hoff + 16 * c);
#pragma unroll
                    for (int i = 0; i < 4; ++i) qn[i] = qpn[i];
                    qn_item = nx;
                }
            }
        }
        // Accumulators.  Self-attention starts them from the row's own NEW token (position pc): its K,V are in registers
        // (rounded to the cache type, exactly as appended for the steps to come), so it never travels through the copy
        // engine -- and q, k, v are dead before the stage loop starts.  Lane (nl, c) holds dims [16c, 16c+16) of row nl of
        // q, k, v; its O fragments cover the same dims (o[nt][0] = dim 16c + nt, o[nt][1] = dim 16c + 8 + nt).
        float o[8][4];
        float m_run = -INFINITY, l_run = 0.f;
#pragma unroll
        for (int nt = 0; nt < 8; ++nt) { o[nt][0] = 0.f; o[nt][1] = 0.f; o[nt][2] = 0.f; o[nt][3] = 0.f; }
        if (SELF) {
            uint32_t kw_new[8], vw_new[8];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                kw_new[2 * i] = Mma16<T>::pack(kx[i].x, kx[i].y); kw_new[2 * i + 1] = Mma16<T>::pack(kx[i].z, kx[i].w);
                vw_new[2 * i] = Mma16<T>::pack(vx[i].x, vx[i].y); vw_new[2 * i + 1] = Mma16<T>::pack(vx[i].z, vx[i].w);
            }
            if (nl < nq) {
                T* dst = reinterpret_cast<T*>(a.kv) + ((size_t)(row0 + nl) * a.T_max + pc) * (size_t)(2 * d) + hoff + 16 * c;
                reinterpret_cast<uint4*>(dst)[0] = make_uint4(kw_new[0], kw_new[1], kw_new[2], kw_new[3]);
                reinterpret_cast<uint4*>(dst)[1] = make_uint4(kw_new[4], kw_new[5], kw_new[6], kw_new[7]);
                reinterpret_cast<uint4*>(dst + d)[0] = make_uint4(vw_new[0], vw_new[1], vw_new[2], vw_new[3]);
                reinterpret_cast<uint4*>(dst + d)[1] = make_uint4(vw_new[4], vw_new[5], vw_new[6], vw_new[7]);
            }
            float sn = 0.f;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const float2 k01 = Mma16<T>::unpack(kw_new[2 * i]), k23 = Mma16<T>::unpack(kw_new[2 * i + 1]);
                sn = fmaf(qx[i].x * 0.125f, k01.x, sn); sn = fmaf(qx[i].y * 0.125f, k01.y, sn);
                sn = fmaf(qx[i].z * 0.125f, k23.x, sn); sn = fmaf(qx[i].w * 0.125f, k23.y, sn);
            }
            sn += __shfl_xor_sync(0xffffffffu, sn, 1);
            sn += __shfl_xor_sync(0xffffffffu, sn, 2);
            m_run = sn;                                  // p = exp(sn - m_run) = 1
            l_run = (c == 0) ? 1.f : 0.f;                // l_run is a per-lane partial, summed over the quad at the end
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const float2 v01 = Mma16<T>::unpack(vw_new[2 * i]), v23 = Mma16<T>::unpack(vw_new[2 * i + 1]);
                const float vv[4] = {v01.x, v01.y, v23.x, v23.y};
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const int dim = 4 * i + e;           // 0..15 within the lane's slice
                    if (dim < 8) o[dim][0] = vv[e];
                    else o[dim - 8][1] = vv[e];
                }
            }
        }
        // Q fragments: rows 0..3 = hi parts of the group's queries, rows 8..11 = lo parts (see attention_mma.cu)
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
        while (true) {
            const int st = (int)(seq % (uint32_t)a.nst);
            const uint32_t ph = (seq / (uint32_t)a.nst) & 1u;
            ab_mbar_wait(&full[st], ph);
            const int hd = hdr[st];
            if ((hd & 0xFF) > 0) {
                const unsigned char* wc = ab_smem + (size_t)st * stage_bytes;
                const int* mp = meta + st * AB_CH;
                // ---- S = Q.K^T for 2 tiles of 8 positions ----
                float s[2][4];
#pragma unroll
                for (int j = 0; j < 2; ++j) {
                    s[j][0] = 0.f; s[j][1] = 0.f; s[j][2] = 0.f; s[j][3] = 0.f;
                    const uint4* kp = reinterpret_cast<const uint4*>(wc + (8 * j + nl) * a.rowb + koff + 32 * c);
                    const uint4 k0 = kp[0], k1 = kp[1];
                    const uint32_t kw[8] = {k0.x, k0.y, k0.z, k0.w, k1.x, k1.y, k1.z, k1.w};
#pragma unroll
                    for (int ks = 0; ks < 4; ++ks) Mma16<T>::mma(s[j], aq[ks], kw[2 * ks], kw[2 * ks + 1]);
                    s[j][0] += s[j][2];                      // hi.K (rows 0..3) + lo.K (rows 8..11)
                    s[j][1] += s[j][3];
                }
                // ---- mask, online softmax on the C fragments (row qrow, positions 8j + 2c + {0,1} of the stage) ----
                float mc = -INFINITY;
#pragma unroll
                for (int j = 0; j < 2; ++j)
#pragma unroll
                    for (int e = 0; e < 2; ++e) {
                        const int mt = mp[8 * j + 2 * c + e];
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
                const float p0 = __expf(s[0][0] - m_use), p1 = __expf(s[0][1] - m_use);
                const float p2 = __expf(s[1][0] - m_use), p3 = __expf(s[1][1] - m_use);
                uint32_t ap[4];                              // P as an A quad: hi in rows 0..3, lo in rows 8..11 (as for Q)
                split2<T>(p0, p1, ap[0], ap[1]);
                split2<T>(p2, p3, ap[2], ap[3]);
                l_run = l_run * corr + ((p0 + p1) + (p2 + p3));
                if (corr != 1.0f) {
#pragma unroll
                    for (int nt = 0; nt < 8; ++nt) { o[nt][0] *= corr; o[nt][1] *= corr; o[nt][2] *= corr; o[nt][3] *= corr; }
                }
                // ---- O += P.V : rows {2c, 2c+1, 2c+8, 2c+9}, dims [8nl, 8nl+8) of this head ----
                const unsigned char* vp = wc + (2 * c) * a.rowb + voff + 16 * nl;
                const uint4 v0 = *reinterpret_cast<const uint4*>(vp), v1 = *reinterpret_cast<const uint4*>(vp + a.rowb);
                const uint4 v2 = *reinterpret_cast<const uint4*>(vp + 8 * a.rowb), v3 = *reinterpret_cast<const uint4*>(vp + 9 * a.rowb);
                const uint32_t va[4] = {v0.x, v0.y, v0.z, v0.w}, vb_[4] = {v1.x, v1.y, v1.z, v1.w};
                const uint32_t vc[4] = {v2.x, v2.y, v2.z, v2.w}, vd[4] = {v3.x, v3.y, v3.z, v3.w};
#pragma unroll
                for (int nt = 0; nt < 8; ++nt) {
                    const uint32_t sel = (nt & 1) ? 0x7632u : 0x5410u;
                    const uint32_t b0 = __byte_perm(va[nt >> 1], vb_[nt >> 1], sel);
                    const uint32_t b1 = __byte_perm(vc[nt >> 1], vd[nt >> 1], sel);
                    Mma16<T>::mma(o[nt], ap, b0, b1);
                }
            }
            __syncwarp();                                // every lane is done reading the stage
            if (lane == 0) ab_mbar_arrive(&empty[st]);
            ++seq;
            if (hd & 0x100) break;
        }
        // ---- normalise and store: lane (nl < nq, c) holds dims [16c,16c+16) of row nl ----
        l_run += __shfl_xor_sync(0xffffffffu, l_run, 1);
        l_run += __shfl_xor_sync(0xffffffffu, l_run, 2);
        if (nl < nq) {
            const float inv = 1.0f / l_run;
            uint32_t w[8];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                w[i] = Mma16<T>::pack((o[2 * i][0] + o[2 * i][2]) * inv, (o[2 * i + 1][0] + o[2 * i + 1][2]) * inv);
                w[4 + i] = Mma16<T>::pack((o[2 * i][1] + o[2 * i][3]) * inv, (o[2 * i + 1][1] + o[2 * i + 1][3]) * inv);
            }
            uint4* dp = reinterpret_cast<uint4*>(reinterpret_cast<T*>(a.ctx) + (size_t)(row0 + nl) * d + hoff + 16 * c);
            dp[0] = make_uint4(w[0], w[1], w[2], w[3]);
            dp[1] = make_uint4(w[4], w[5], w[6], w[7]);
        }
    }
}

// ITB_ATTN_BULK: unset = cross-attention launches of >= 96 work items (the default), 0 = never, 1 = every 16-bit attention
// launch whose shape allows it (tests), 3 = cross- and self-attention launches of >= 96 work items.
static int bulk_mode() {
    static int v = -2;
    if (v == -2) {
        const char* e = getenv("ITB_ATTN_BULK");
        v = (e == nullptr) ? -1 : (e[0] == '0' ? 0 : (e[0] == '3' ? 3 : 1));
    }
    return v;
}

// The bulk-copy kernel serves a launch when its shape fits (<= 12 heads, row pitch small enough for two stages, context
// within the producer's look-ahead) and it has enough work items to occupy the SMs (an item is a whole sentence: small
// batches keep the warp-per-head kernel).  Measured on B200 at 4096 beam rows (tools/kernel_bench.py, same box, same run):
// cross-attention 40.8 -> 30.5 us (S = 40) and 64.4 -> 58.0 us (S = 100, 83 % of the HBM copy bandwidth); self-attention is
// within noise of the warp-per-head kernel (t = 30: 44.7 vs 46.7 us, t = 80: 75.7 vs 75.6 us), and end to end the
// one-CTA-per-SM self kernel takes away the overlap between the two lanes of forward_many (bench: 422 k tok/s with the
// cross kernel only, 414 k with both, 412 k with neither), so by default only cross-attention runs on it.
bool itb_attn_bulk_applies(bool self, int M, int beam, int d, int T_max, int S) {
    const int mode = bulk_mode();
    if (mode == 0 || beam > 16 || d % DH != 0 || d / DH > 12) return false;       // <= 12 consumer warps + 1 producer (416 threads)
    const int rowb = 2 * d * 2 + 16;
    if (2 * AB_CH * rowb + 12 * d * 4 > 200 * 1024) return false;
    if ((self ? T_max : S) > 32 * AB_LOOKAHEAD) return false;
    if (mode == 1) return true;
    if (self && mode != 3) return false;
    const int n_items = (M / beam) * ((beam + 3) / 4);
    return n_items >= 96;
}

template <typename T>
static int launch_bulk(AttnBulkArgs a, bool self, cudaStream_t s) {
    const int H = a.d / DH;
    a.rowb = 2 * a.d * (int)sizeof(T) + 16;
    const int stage = AB_CH * a.rowb;
    a.anc_cap = self ? ((a.T_max + 15) & ~15) : 0;
    const size_t aux = self ? 4 * (size_t)a.anc_cap : (size_t)((a.S + 15) & ~15);
    const size_t qbuf = AB_QSTAGE ? 4 * (size_t)(self ? 3 : 1) * a.d * sizeof(float) : 0;
    const size_t fixed = qbuf + (2 * AB_MAX_STAGES + 2) * 8 + AB_MAX_STAGES * AB_CH * 4 + AB_MAX_STAGES * 4 + aux + 16;
    a.nst = (int)((227 * 1024 - fixed) / stage);
    if (a.nst > AB_MAX_STAGES) a.nst = AB_MAX_STAGES;
    if (a.nst < 2) { itb_set_error("attention(bulk): shape does not fit two stages"); return -2; }
    a.n_items = (a.M / a.beam) * ((a.beam + 3) / 4);
    const size_t smem = (size_t)a.nst * stage + fixed;
    static bool attr_self[ITB_MAX_DEVICES] = {}, attr_cross[ITB_MAX_DEVICES] = {};
    cudaError_t e = self ? itb_smem_attr_once(attr_self, attn_bulk_kernel<T, true>, 227 * 1024)
                         : itb_smem_attr_once(attr_cross, attn_bulk_kernel<T, false>, 227 * 1024);
    if (e != cudaSuccess) { itb_set_error("attention(bulk): cudaFuncSetAttribute: %s", cudaGetErrorString(e)); return -1; }
    if (smem > 227 * 1024) { itb_set_error("attention(bulk): %zu bytes of shared memory needed", smem); return -2; }
    const int grid = a.n_items < itb_num_sms() ? a.n_items : itb_num_sms();
    const dim3 g(grid), blk((H + 1) * 32);
    e = self ? itb_launch(attn_bulk_kernel<T, true>, g, blk, smem, s, a) : itb_launch(attn_bulk_kernel<T, false>, g, blk, smem, s, a);
    if (e != cudaSuccess) { itb_set_error("%s(bulk): launch failed: %s", self ? "self_attention" : "cross_attention", cudaGetErrorString(e)); return -1; }
    return itb_check_launch(self ? "self_attention(bulk)" : "cross_attention(bulk)");
}

int itb_attn_bulk(int dtype, bool self, const void* ctl, const int* sent_done, const int* mem_len, const float* q, int ldq,
                  void* kv, const unsigned char* anc, const unsigned char* mask, void* ctx, int M, int beam, int d, int T_max,
                  int S, int L, int layer, cudaStream_t s) {
    AttnBulkArgs a = {};
    a.ctl = reinterpret_cast<const ItbCtl*>(ctl);
    a.sent_done = sent_done; a.mem_len = mem_len; a.q = q; a.ldq = ldq; a.kv = kv; a.anc = anc; a.mask = mask; a.ctx = ctx;
    a.M = M; a.beam = beam; a.d = d; a.T_max = T_max; a.S = S; a.L = L; a.layer = layer;
    if (dtype == ITB_BF16) return launch_bulk<__nv_bfloat16>(a, self, s);
    return launch_bulk<__half>(a, self, s);
}
