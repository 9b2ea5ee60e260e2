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

#define DH 64
#define AM_WARPS 4
#define AM_THREADS (AM_WARPS * 32)

struct AttnMmaArgs {
    const ItbCtl* ctl;
    const int* sent_done;                // [B] or null
    const int* mem_len;                  // cross: [B] or null (= S)
    const float* q; int ldq;             // f32 queries (self: qkv buffer, k at +d, v at +2d)
    void* kv;                            // self: [M][T_max][2][d]; cross: [B][S][L][2][d]
    const unsigned char* anc;            // self: [2][M][T_max]
    const unsigned char* mask;           // cross: [B][S] or null
    void* ctx;                           // act [M][d]
    int M, beam, d, T_max, S, L, layer, n_warps;
};

template <typename T> struct Mma16;
template <> struct Mma16<__nv_bfloat16> {
    static __device__ __forceinline__ void mma(float* c, const uint32_t* a, uint32_t b0, uint32_t b1) {
        asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                     : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                     : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
    }
    static __device__ __forceinline__ uint32_t pack(float lo, float hi) {
        __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
        return *reinterpret_cast<uint32_t*>(&h);
    }
    static __device__ __forceinline__ float2 unpack(uint32_t u) { return __bfloat1622float2(*reinterpret_cast<__nv_bfloat162*>(&u)); }
};
template <> struct Mma16<__half> {
    static __device__ __forceinline__ void mma(float* c, const uint32_t* a, uint32_t b0, uint32_t b1) {
        asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                     : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                     : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
    }
    static __device__ __forceinline__ uint32_t pack(float lo, float hi) {
        __half2 h = __floats2half2_rn(lo, hi);
        return *reinterpret_cast<uint32_t*>(&h);
    }
    static __device__ __forceinline__ float2 unpack(uint32_t u) { return __half22float2(*reinterpret_cast<__half2*>(&u)); }
};

// hi/lo split of a pair of fp32 values into two packed 16-bit registers: x = hi + lo up to ~2^-17 relative.
template <typename T>
__device__ __forceinline__ void split2(float x, float y, uint32_t& hi, uint32_t& lo) {
    hi = Mma16<T>::pack(x, y);
    const float2 h = Mma16<T>::unpack(hi);
    lo = Mma16<T>::pack(x - h.x, y - h.y);
}

template <typename T, bool SELF>
__global__ void __launch_bounds__(AM_THREADS, 4) attn_mma_kernel(AttnMmaArgs a) {
    pdl_trigger();
    pdl_wait();
    const int lane = threadIdx.x & 31, nl = lane >> 2, c = lane & 3;       // nl: B-fragment column / C row; c: quad lane
    const int wg = blockIdx.x * AM_WARPS + (threadIdx.x >> 5);
    if (wg >= a.n_warps) return;
    int4 c4 = make_int4(1, 0, 0, 0);                     // {step, done, steps_run, cur}
    if (a.ctl != nullptr) c4 = *reinterpret_cast<const int4*>(a.ctl);
    if (c4.y != 0) return;
    const int d = a.d, H = d / DH, nz = (a.beam + 3) >> 2;
    const int h = wg % H, z = (wg / H) % nz, b = wg / (H * nz);
    if (a.sent_done != nullptr && a.sent_done[b] != 0) return;
    const int q0 = z * 4, row0 = b * a.beam + q0, nq = min(4, a.beam - q0);
    const int n = SELF ? c4.x : ((a.mem_len != nullptr) ? a.mem_len[b] : a.S);   // self: positions 0..step-1
    const int hoff = h * DH;

    // ---- Q fragments: row nl (< nq) of the A operand, dims [16c,16c+16) in k-slice order, scaled by 1/8, hi + lo ----
    uint32_t qhi[8], qlo[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) { qhi[i] = 0u; qlo[i] = 0u; }
    if (nl < nq) {
        const float4* qp = reinterpret_cast<const float4*>(a.q + (size_t)(row0 + nl) * a.ldq + hoff + 16 * c);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const float4 x = qp[i];
            split2<T>(x.x * 0.125f, x.y * 0.125f, qhi[2 * i], qlo[2 * i]);
            split2<T>(x.z * 0.125f, x.w * 0.125f, qhi[2 * i + 1], qlo[2 * i + 1]);
        }
    }

    // ---- geometry of the virtual position list ----
    const T* base;                       // element pointer of row index 0, this head
    size_t pos_stride;                   // elements between consecutive row indices
    int cshared = n, tl = 0, nv = n;     // self: [0,cshared) shared, then nq tails of tl positions each
    const unsigned char* anc0 = nullptr;
    if (SELF) {
        const int pc = n - 1;
        base = reinterpret_cast<const T*>(a.kv) + hoff;
        pos_stride = (size_t)2 * d;
        anc0 = a.anc + ((size_t)c4.w * a.M + row0) * a.T_max;
        // append the new K,V of the group's rows at position pc: lane (nl < nq, c) owns dims [16c,16c+16) of row nl
        if (nl < nq) {
            const float* kp = a.q + (size_t)(row0 + nl) * a.ldq + d + hoff + 16 * c;
            T* dst = reinterpret_cast<T*>(a.kv) + ((size_t)(row0 + nl) * a.T_max + pc) * pos_stride + hoff + 16 * c;
#pragma unroll
            for (int kvsel = 0; kvsel < 2; ++kvsel) {
                const float4* sp = reinterpret_cast<const float4*>(kp + kvsel * d);
                uint32_t w[8];
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const float4 x = sp[i];
                    w[2 * i] = Mma16<T>::pack(x.x, x.y);
                    w[2 * i + 1] = Mma16<T>::pack(x.z, x.w);
                }
                uint4* dp = reinterpret_cast<uint4*>(dst + kvsel * d);
                dp[0] = make_uint4(w[0], w[1], w[2], w[3]);
                dp[1] = make_uint4(w[4], w[5], w[6], w[7]);
            }
        }
        // first position whose physical row differs between the group's slots
        cshared = pc;
        if (nq > 1) {
            for (int blk = 0; blk * 32 < pc; ++blk) {
                const int p = blk * 32 + lane;
                bool same = true;
                if (p < pc) {
                    const int a0 = anc0[p];
                    for (int qi = 1; qi < nq; ++qi) same = same && (anc0[(size_t)qi * a.T_max + p] == a0);
                }
                const unsigned bal = __ballot_sync(0xffffffffu, !same);
                if (bal != 0u) { cshared = blk * 32 + __ffs(bal) - 1; break; }
            }
        }
        tl = pc - cshared + 1;
        nv = cshared + nq * tl;
        __syncwarp();                    // the appended rows are read back below by other lanes of this warp
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

    for (int vb = 0; vb < nv; vb += 32) {
        // ---- virtual position vb + lane: row index and meta = owner mask (bits 0-3) | masked-source flag (bit 4) ----
        int rowidx = -1, meta = 0;
        {
            const int v = vb + lane;
            if (v < nv) {
                if (SELF) {
                    if (v < cshared) { rowidx = (b * a.beam + anc0[v]) * a.T_max + v; meta = 0xF; }
                    else {
                        const int vt = v - cshared;
                        const int qi = (vt >= tl) + (vt >= 2 * tl) + (vt >= 3 * tl);
                        const int p = cshared + vt - qi * tl;
                        const int slot = (p == n - 1) ? (q0 + qi) : anc0[(size_t)qi * a.T_max + p];
                        rowidx = (b * a.beam + slot) * a.T_max + p;
                        meta = 1 << qi;
                    }
                } else {
                    rowidx = v;
                    meta = 0xF | ((mrow != nullptr && mrow[v] == 0) ? 0x10 : 0);
                }
            }
        }
        // ---- S = Q.K^T for 4 tiles of 8 positions ----
        uint4 kr[4][2];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int r = __shfl_sync(0xffffffffu, rowidx, 8 * j + nl);
            kr[j][0] = make_uint4(0u, 0u, 0u, 0u); kr[j][1] = make_uint4(0u, 0u, 0u, 0u);
            if (r >= 0) {
                const uint4* kp = reinterpret_cast<const uint4*>(base + (size_t)r * pos_stride + 16 * c);
                kr[j][0] = kp[0]; kr[j][1] = kp[1];
            }
        }
        float s[4][4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            s[j][0] = 0.f; s[j][1] = 0.f; s[j][2] = 0.f; s[j][3] = 0.f;
            const uint32_t kw[8] = {kr[j][0].x, kr[j][0].y, kr[j][0].z, kr[j][0].w, kr[j][1].x, kr[j][1].y, kr[j][1].z, kr[j][1].w};
#pragma unroll
            for (int ks = 0; ks < 4; ++ks) {
                const uint32_t ah[4] = {qhi[2 * ks], 0u, qhi[2 * ks + 1], 0u};
                const uint32_t al[4] = {qlo[2 * ks], 0u, qlo[2 * ks + 1], 0u};
                Mma16<T>::mma(s[j], ah, kw[2 * ks], kw[2 * ks + 1]);
                Mma16<T>::mma(s[j], al, kw[2 * ks], kw[2 * ks + 1]);
            }
        }
        // ---- V rows of this chunk (issued before the softmax arithmetic) ----
        uint4 vr[2][4];
#pragma unroll
        for (int t = 0; t < 2; ++t)
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int r = __shfl_sync(0xffffffffu, rowidx, 16 * t + 2 * c + (i & 1) + 8 * (i >> 1));
                vr[t][i] = make_uint4(0u, 0u, 0u, 0u);
                if (r >= 0) vr[t][i] = *reinterpret_cast<const uint4*>(base + (size_t)r * pos_stride + d + 8 * nl);
            }
        // ---- mask, online softmax on the C fragments (row qrow, positions vb + 8j + 2c + {0,1}) ----
        float mc = -INFINITY;
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const int mt = __shfl_sync(0xffffffffu, meta, 8 * j + 2 * c + e);
                float x = s[j][e];
                if (mt & 0x10) x += -10000.0f;               // additive mask (1 - m) * -10000: reference src/bert_seq2seq.py:36-37
                x = ((mt >> qrow) & 1) ? x : -INFINITY;
                s[j][e] = x;
                mc = fmaxf(mc, x);
            }
        mc = fmaxf(mc, __shfl_xor_sync(0xffffffffu, mc, 1));
        mc = fmaxf(mc, __shfl_xor_sync(0xffffffffu, mc, 2));
        const float m_new = fmaxf(m_run, mc);
        const float m_use = (m_new == -INFINITY) ? 0.f : m_new;          // nothing visible yet for this row: all p = 0
        const float corr = __expf(m_run - m_use);                        // m_run = -inf -> 0
        m_run = m_new;
        float psum = 0.f;
        uint32_t phi[2][4], plo[2][4];
#pragma unroll
        for (int t = 0; t < 2; ++t) {
            const float p0 = __expf(s[2 * t][0] - m_use), p1 = __expf(s[2 * t][1] - m_use);
            const float p2 = __expf(s[2 * t + 1][0] - m_use), p3 = __expf(s[2 * t + 1][1] - m_use);
            psum += (p0 + p1) + (p2 + p3);
            split2<T>(p0, p1, phi[t][0], plo[t][0]);
            split2<T>(p2, p3, phi[t][2], plo[t][2]);
            phi[t][1] = 0u; phi[t][3] = 0u; plo[t][1] = 0u; plo[t][3] = 0u;
        }
        l_run = l_run * corr + psum;
#pragma unroll
        for (int nt = 0; nt < 8; ++nt) { o[nt][0] *= corr; o[nt][1] *= corr; }
        // ---- O += P.V ----
#pragma unroll
        for (int t = 0; t < 2; ++t) {
            const uint32_t va[4] = {vr[t][0].x, vr[t][0].y, vr[t][0].z, vr[t][0].w};
            const uint32_t vb_[4] = {vr[t][1].x, vr[t][1].y, vr[t][1].z, vr[t][1].w};
            const uint32_t vc[4] = {vr[t][2].x, vr[t][2].y, vr[t][2].z, vr[t][2].w};
            const uint32_t vd[4] = {vr[t][3].x, vr[t][3].y, vr[t][3].z, vr[t][3].w};
#pragma unroll
            for (int nt = 0; nt < 8; ++nt) {
                const uint32_t sel = (nt & 1) ? 0x7632u : 0x5410u;
                const uint32_t b0 = __byte_perm(va[nt >> 1], vb_[nt >> 1], sel);
                const uint32_t b1 = __byte_perm(vc[nt >> 1], vd[nt >> 1], sel);
                Mma16<T>::mma(o[nt], phi[t], b0, b1);
                Mma16<T>::mma(o[nt], plo[t], b0, b1);
            }
        }
    }
    // ---- normalise and store: lane (nl < nq, c) holds dims [16c,16c+16) of row nl ----
    l_run += __shfl_xor_sync(0xffffffffu, l_run, 1);
    l_run += __shfl_xor_sync(0xffffffffu, l_run, 2);
    if (nl < nq) {
        const float inv = 1.0f / l_run;
        uint32_t w[8];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            w[i] = Mma16<T>::pack(o[2 * i][0] * inv, o[2 * i + 1][0] * inv);
            w[4 + i] = Mma16<T>::pack(o[2 * i][1] * inv, o[2 * i + 1][1] * inv);
        }
        uint4* dp = reinterpret_cast<uint4*>(reinterpret_cast<T*>(a.ctx) + (size_t)(row0 + nl) * d + hoff + 16 * c);
        dp[0] = make_uint4(w[0], w[1], w[2], w[3]);
        dp[1] = make_uint4(w[4], w[5], w[6], w[7]);
    }
}

template <typename T>
static int launch_attn_mma(AttnMmaArgs a, bool self, cudaStream_t s) {
    const int B = a.M / a.beam, H = a.d / DH;
    a.n_warps = B * H * ((a.beam + 3) / 4);
    const dim3 grid((a.n_warps + AM_WARPS - 1) / AM_WARPS), block(AM_THREADS);
    cudaError_t e = self ? itb_launch(attn_mma_kernel<T, true>, grid, block, (size_t)0, s, a)
                         : itb_launch(attn_mma_kernel<T, false>, grid, block, (size_t)0, s, a);
    if (e != cudaSuccess) { itb_set_error("%s: launch failed: %s", self ? "self_attention" : "cross_attention", cudaGetErrorString(e)); return -1; }
    return itb_check_launch(self ? "self_attention" : "cross_attention");
}

// Called by itb_self_attention / itb_cross_attention (attention.cu) for the 16-bit dtypes.
int itb_attn_mma(int dtype, bool self, const void* ctl, const int* sent_done, const int* mem_len, const float* q, int ldq,
                 void* kv, const unsigned char* anc, const unsigned char* mask, void* ctx, int M, int beam, int d, int T_max,
                 int S, int L, int layer, cudaStream_t s) {
    AttnMmaArgs a = {};
    a.ctl = reinterpret_cast<const ItbCtl*>(ctl);
    a.sent_done = sent_done; a.mem_len = mem_len; a.q = q; a.ldq = ldq; a.kv = kv; a.anc = anc; a.mask = mask; a.ctx = ctx;
    a.M = M; a.beam = beam; a.d = d; a.T_max = T_max; a.S = S; a.L = L; a.layer = layer;
    if (dtype == ITB_BF16) return launch_attn_mma<__nv_bfloat16>(a, self, s);
    return launch_attn_mma<__half>(a, self, s);
}
