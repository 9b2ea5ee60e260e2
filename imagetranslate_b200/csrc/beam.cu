// Fused search step: reference src/seq_gen.py:134-143,194-227 in ONE launch.
//
// Phase 1 (every CTA = one (row, vocab-chunk)): streams the fp32 logits ONCE (float4, coalesced),
//   keeping per-thread online-softmax (max, sum-exp) and a sorted top-K list; warp-shuffle merge,
//   then a block merge.  No (M,V) log-prob tensor is ever written.
// Phase 2 (the last CTA to finish for a sentence -- atomic ticket, nobody waits): finishes
//   log-sum-exp per row, turns the row top-K logits into candidates
//   (score + logp) / ((size+6)/6)^alpha, handles finished / over-length rows (all-zero rows tie;
//   deterministic order = lowest flat index), picks the sentence top-`beam`, applies the forced-PAD
//   rewrites, and writes the new prefixes, sizes, scores, EOS flags and the ANCESTOR TABLE that
//   replaces the KV-cache gather.  The last sentence advances the device-side step counter and
//   evaluates the termination rule, so the host never synchronises inside the search loop.
// HBM roofline: 4*M*V bytes read per launch.
#include <float.h>
#include "common.cuh"
#include "launch.h"

struct BeamArgs {
    ItbCtl* ctl;
    ItbGrp* grp; const int* sent_grp; int* sent_done; int G;
    const float* logits; int ldl;
    int B, beam, V, T_max, n_chunk, chunk, pad_id, eos_id;
    float alpha;
    const int* max_lens;
    int* tokens; unsigned char* anc; float* scores; float* sizes; unsigned char* fin;
    float* cand_val; int* cand_idx; float* chunk_max; float* chunk_sum; int* sent_counter;
    int fused;       // 1: phase 1 already ran inside the vocabulary GEMM epilogue (gemm_tc.cu); grid = one CTA per sentence
};

__device__ __forceinline__ float fast_ex2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

__device__ __forceinline__ bool better(float v, int i, float bv, int bi) { return v > bv || (v == bv && i < bi); }

// One warp selects the top-K of (val, idx)[0..n) in shared memory by (val desc, idx asc).
// Entries are consumed (idx set to -1).  Results to out_val/out_idx[0..K).
__device__ __forceinline__ void warp_select(float* val, int* idx, int n, int K, float* out_val, int* out_idx, int lane) {
    for (int k = 0; k < K; ++k) {
        float bv = -INFINITY; int bi = INT_MAX; int bp = -1;
        for (int p = lane; p < n; p += 32) {
            int i = idx[p];
            if (i >= 0 && (bp < 0 || better(val[p], i, bv, bi))) { bv = val[p]; bi = i; bp = p; }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            float ov = __shfl_xor_sync(0xffffffffu, bv, o);
            int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            int op = __shfl_xor_sync(0xffffffffu, bp, o);
            if (op >= 0 && (bp < 0 || better(ov, oi, bv, bi))) { bv = ov; bi = oi; bp = op; }
        }
        if (lane == 0) {
            out_val[k] = bv; out_idx[k] = (bp >= 0) ? bi : -1;
            if (bp >= 0) idx[bp] = -1;
        }
        __syncwarp();
    }
}

// PRECISE: expf for the fp32 parity mode; otherwise EX2 with a pre-scaled argument (16-bit modes).
template <int KT, bool PRECISE>
__global__ void __launch_bounds__(256) beam_step_kernel(BeamArgs a) {
    __shared__ float s_val[8 * KT > 16 * 16 ? 8 * KT : 16 * 16];   // phase 1: per-warp lists; phase 2: row candidates
    __shared__ int s_idx[8 * KT > 16 * 16 ? 8 * KT : 16 * 16];
    __shared__ float s_m[8], s_s[8];
    __shared__ float s_top_val[16];
    __shared__ int s_top_idx[16];
    __shared__ float s_rtop_val[8][16];
    __shared__ int s_rtop_idx[8][16];
    __shared__ int s_src[16], s_word[16];
    __shared__ int s_q[8][64];            // phase 2: per-warp queue of qualifying chunk indices
    __shared__ int s_last;

    pdl_trigger();
    pdl_wait();
    ItbCtl* ctl = a.ctl;
    if (ctl->done) return;
    const int step = ctl->step, cur = ctl->cur, nxt = cur ^ 1;
    const bool first = (step == 1);
    const int M = a.B * a.beam, K = a.beam, V = a.V;
    const bool fused = a.fused != 0;
    const int c = blockIdx.x, r = fused ? blockIdx.y * a.beam : blockIdx.y, b = r / a.beam, kslot = r % a.beam;
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;

    const int gid = a.sent_grp[b];
    const bool gdone = a.sent_done[b] != 0;                                  // this sentence's reference batch has ended
    const bool lim = (!first) && (a.max_lens[b] < step + 1);                 // seq_gen.py:138
    const bool row_zero = (a.fin[(size_t)cur * M + r] != 0) || lim;          // :195-197
    const bool skip = fused || gdone || row_zero || (first && kslot != 0);            // step 1 has one row per sentence (:146,221)

    if (!skip) {
        // The warp keeps ONE sorted top-KT list, entry j in lane j (by value desc, index asc), and a warp-uniform
        // threshold thr = its KT-th value: a logit is looked at again only when it reaches thr, which happens
        // ~KT*ln(n/KT) times per warp over n logits instead of per thread.
        float lv = -INFINITY; int li = INT_MAX;
        float thr = -INFINITY;
        float m = -INFINITY, s = 0.f;
        const int lo = c * a.chunk, hi = min(V, lo + a.chunk);
        const float* row = a.logits + (size_t)r * a.ldl;
        // 4 float4 per thread in flight; per batch of 16 logits: one max tree, a (rare) rescale of the running sum, one
        // FFMA + EX2 + FADD per element, and the top-K insertion only when the batch maximum beats the K-th best so far.
        constexpr float kLog2e = 1.4426950408889634f;
        // software pipeline: the 4 float4 of the NEXT batch are requested before the current batch is consumed
        // (8 x 16 B in flight per thread)
        const float4 kNegInf4 = make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY);
        float4 nx[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int i = lo + tid * 4 + u * 256 * 4;
            nx[u] = (i < hi) ? __ldcs(reinterpret_cast<const float4*>(row + i)) : kNegInf4;
        }
        // the loop bound is warp-uniform (the body holds warp collectives); lanes past `hi` carry -inf
        for (int iw = lo + w * 128; iw < hi; iw += 4 * 256 * 4) {
            const int i0 = iw + lane * 4;
            float xs[16];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int i = i0 + u * 256 * 4;
                float4 x4 = nx[u];
                if (i + 4 > hi && i < hi) {                  // ragged end of the chunk (V not a multiple of 4)
                    if (i + 1 >= hi) x4.y = -INFINITY;
                    if (i + 2 >= hi) x4.z = -INFINITY;
                    if (i + 3 >= hi) x4.w = -INFINITY;
                }
                xs[4 * u] = x4.x; xs[4 * u + 1] = x4.y; xs[4 * u + 2] = x4.z; xs[4 * u + 3] = x4.w;
                const int in = i + 4 * 256 * 4;
                nx[u] = (in < hi) ? __ldcs(reinterpret_cast<const float4*>(row + in)) : kNegInf4;
            }
            float bm = xs[0];
#pragma unroll
            for (int e = 1; e < 16; ++e) bm = fmaxf(bm, xs[e]);
            if (bm > m) {                                                               // rare once the maximum settled
                s = (m == -INFINITY) ? 0.f : s * (PRECISE ? expf(m - bm) : fast_ex2((m - bm) * kLog2e));
                m = bm;
            }
            const float mref = (m == -INFINITY) ? 0.f : m;                              // lane entirely past `hi`: all terms are 0
            if (PRECISE) {
#pragma unroll
                for (int e = 0; e < 16; ++e) s += expf(xs[e] - mref);                   // exp(-inf) = 0 for the padding
            } else {
                const float nm = -mref * kLog2e;
#pragma unroll
                for (int e = 0; e < 16; ++e) s += fast_ex2(fmaf(xs[e], kLog2e, nm));
            }
            if (__any_sync(0xffffffffu, bm >= thr && bm > -INFINITY)) {
#pragma unroll
                for (int e = 0; e < 16; ++e) {
                    unsigned q = __ballot_sync(0xffffffffu, xs[e] >= thr && xs[e] > -INFINITY);
                    while (q != 0u) {
                        const int src = __ffs(q) - 1;
                        q &= q - 1;
                        const float x = __shfl_sync(0xffffffffu, xs[e], src);
                        const int xi = i0 + (src - lane) * 4 + (e >> 2) * 256 * 4 + (e & 3);     // index of lane src's element e
                        // sorted insert: entries better than (x, xi) form a prefix of the list
                        const unsigned bet = __ballot_sync(0xffffffffu, better(lv, li, x, xi)) & ((1u << KT) - 1u);
                        const int pos = __popc(bet);
                        const float uv = __shfl_up_sync(0xffffffffu, lv, 1);
                        const int ui = __shfl_up_sync(0xffffffffu, li, 1);
                        if (lane < KT) {
                            if (lane == pos) { lv = x; li = xi; }
                            else if (lane > pos) { lv = uv; li = ui; }
                        }
                        thr = __shfl_sync(0xffffffffu, lv, KT - 1);
                    }
                }
            }
        }
        // warp merge of the online-softmax pairs
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            float om = __shfl_xor_sync(0xffffffffu, m, o), os = __shfl_xor_sync(0xffffffffu, s, o);
            float nm = fmaxf(m, om);
            if (nm == -INFINITY) { s = 0.f; } else { s = s * expf(m - nm) + os * expf(om - nm); }
            m = nm;
        }
        // the warp's list goes to shared memory as it is (lane k holds entry k)
        if (lane < K) { s_val[w * KT + lane] = lv; s_idx[w * KT + lane] = (li == INT_MAX) ? -1 : li; }
        if (lane == 0) {
            for (int k = K; k < KT; ++k) { s_val[w * KT + k] = -INFINITY; s_idx[w * KT + k] = -1; }
            s_m[w] = m; s_s[w] = s;
        }
        __syncthreads();
        if (w == 0) {
            const size_t o = ((size_t)r * a.n_chunk + c) * KT;
            warp_select(s_val, s_idx, 8 * KT, K, a.cand_val + o, a.cand_idx + o, lane);
            if (lane >= K && lane < KT) { a.cand_val[o + lane] = -INFINITY; a.cand_idx[o + lane] = -1; }   // unused slots
            if (lane == 0) {
                float mm = -INFINITY, ss = 0.f;
                for (int j = 0; j < 8; ++j) mm = fmaxf(mm, s_m[j]);
                for (int j = 0; j < 8; ++j) if (s_m[j] != -INFINITY) ss += s_s[j] * expf(s_m[j] - mm);
                a.chunk_max[(size_t)r * a.n_chunk + c] = mm;
                a.chunk_sum[(size_t)r * a.n_chunk + c] = ss;
            }
        }
    }

    // ---- ticket: the last CTA of this sentence runs phase 2 (fused: every CTA is its sentence's only one) ----
    if (!fused) {
        __syncthreads();
        if (tid == 0) {
            __threadfence();
            int t = atomicAdd(&a.sent_counter[b], 1);
            s_last = (t == a.beam * a.n_chunk - 1);
            if (s_last) __threadfence();
        }
        __syncthreads();
        if (!s_last) return;
    }

    // ---------------------------------- phase 2 (one CTA per sentence) ----------------------------------
    ITB_TL(1);
    if (!gdone) {
    const int nrows = first ? 1 : K;
    float* rc_val = s_val;      // [nrows][K] row candidates (sentence-level values)
    int* rc_idx = s_idx;
    // one warp per beam row, all rows of the sentence in flight at once (the critical path of the step ends here)
    for (int kr = w; kr < nrows; kr += 8) {
        const int rr = b * a.beam + kr;
        const float sc = a.scores[(size_t)cur * M + rr];
        const float sz = a.sizes[(size_t)cur * M + rr];
        const bool rz = (a.fin[(size_t)cur * M + rr] != 0) || lim;
        float lp = 1.f;
        if (K > 1) lp = powf((sz + 6.0f) / 6.0f, a.alpha);                                  // seq_gen.py:43
        if (rz) {
            // whole row is 0 => V exact ties at (score + 0)/lp; lowest flat index first
            if (lane < K) {
                float cv = sc + 0.0f;
                if (K > 1) cv = cv / lp;
                rc_val[kr * K + lane] = cv;
                rc_idx[kr * K + lane] = kr * V + lane;
            }
        } else {
            // log-sum-exp over the chunk partials, then a streaming top-K over the n_chunk * KT chunk candidates with the
            // warp-level sorted list (entry j in lane j) and its threshold -- any number of chunks, no staging buffer
            const int nc = a.n_chunk;
            const float* cmp = a.chunk_max + (size_t)rr * nc;
            const float* csp = a.chunk_sum + (size_t)rr * nc;
            float mm = -INFINITY, ss = 0.f, logz = 0.f;
            float lv = -INFINITY; int li = INT_MAX;                  // the row's sorted top-KT list: entry j in lane j
            float thr = -INFINITY;
            // inserts the KT candidates of up to 32 chunks (lane l < k owns chunk s_q[w][l]): one round trip for all of
            // them, then rank by rank with the list's threshold
            auto insert_chunks = [&](int base, int k) {
                const bool mine = lane < k;
                const int ci = mine ? s_q[w][base + lane] : 0;
                float xv[KT]; int xj[KT];
#pragma unroll
                for (int u = 0; u < KT; u += 4) {
                    float4 v4 = make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY);
                    int4 j4 = make_int4(-1, -1, -1, -1);
                    if (mine) {
                        v4 = __ldcg(reinterpret_cast<const float4*>(a.cand_val + ((size_t)rr * nc + ci) * KT + u));
                        j4 = __ldcg(reinterpret_cast<const int4*>(a.cand_idx + ((size_t)rr * nc + ci) * KT + u));
                    }
                    xv[u] = v4.x; xv[u + 1] = v4.y; xv[u + 2] = v4.z; xv[u + 3] = v4.w;
                    xj[u] = j4.x; xj[u + 1] = j4.y; xj[u + 2] = j4.z; xj[u + 3] = j4.w;
                }
#pragma unroll
                for (int u = 0; u < KT; ++u) {
                    unsigned q = __ballot_sync(0xffffffffu, xj[u] >= 0 && xv[u] >= thr && xv[u] > -INFINITY);
                    while (q != 0u) {
                        const int src = __ffs(q) - 1;
                        q &= q - 1;
                        const float cx = __shfl_sync(0xffffffffu, xv[u], src);
                        const int cxi = __shfl_sync(0xffffffffu, xj[u], src);
                        const unsigned bet = __ballot_sync(0xffffffffu, better(lv, li, cx, cxi)) & ((1u << KT) - 1u);
                        const int pos = __popc(bet);
                        const float uv = __shfl_up_sync(0xffffffffu, lv, 1);
                        const int ui = __shfl_up_sync(0xffffffffu, li, 1);
                        if (lane < KT) {
                            if (lane == pos) { lv = cx; li = cxi; }
                            else if (lane > pos) { lv = uv; li = ui; }
                        }
                        thr = __shfl_sync(0xffffffffu, lv, KT - 1);
                    }
                }
            };
            constexpr int CPL = 16;                                  // chunks per lane held in registers
            bool streamed = nc > 32 * CPL;                           // the generic, streaming form below
            if (!streamed) {
                // ---- the usual case (V <= 65536): every chunk maximum / sum of the row is loaded ONCE, all loads in
                //      flight together; the passes below (max, sum, threshold, queue) run from registers -- the looped
                //      form paid a dependent memory round trip per 32 chunks and pass, which was the whole kernel ----
                float cmr[CPL], csr[CPL];
#pragma unroll
                for (int j = 0; j < CPL; ++j) {
                    const int ci = j * 32 + lane;
                    cmr[j] = ci < nc ? __ldcg(cmp + ci) : -INFINITY;
                    csr[j] = ci < nc ? __ldcg(csp + ci) : 0.f;
                }
#pragma unroll
                for (int j = 0; j < CPL; ++j) mm = fmaxf(mm, cmr[j]);
                mm = warp_max(mm);
                ITB_TL(10);
#pragma unroll
                for (int j = 0; j < CPL; ++j)
                    if (cmr[j] != -INFINITY) ss += csr[j] * expf(cmr[j] - mm);        // same order as the looped form
                ss = warp_sum(ss);
                logz = logf(ss);
                ITB_TL(11);
                // T = K-th largest chunk maximum (a real logit each): a lower bound of the row's K-th largest logit, so
                // only chunks whose maximum reaches T can hold a top-K entry -- typically K..2K of them.  K rounds of
                // "largest remaining maximum" (local maximum of the lane's 16, warp maximum, the lowest lane that holds it
                // gives up one slot) name the first K chunks of the queue directly; the chunks that TIE with T follow.
                float cw[CPL];                                       // working copy: extracted slots become -inf
#pragma unroll
                for (int j = 0; j < CPL; ++j) cw[j] = cmr[j];
                float T = -INFINITY;
                int cnt = 0;
                for (int k = 0; k < K; ++k) {
                    float lm = cw[0];
#pragma unroll
                    for (int j = 1; j < CPL; ++j) lm = fmaxf(lm, cw[j]);
                    const float wm = warp_max(lm);
                    if (wm == -INFINITY) break;                      // fewer than K chunks hold anything (warp-uniform)
                    T = wm;
                    const int wl = __ffs(__ballot_sync(0xffffffffu, lm == wm)) - 1;
                    if (lane == wl) {
                        int js = 0;
                        bool found = false;
#pragma unroll
                        for (int j = 0; j < CPL; ++j) {
                            const bool hit = !found && cw[j] == wm;
                            if (hit) { js = j; cw[j] = -INFINITY; }
                            found = found || hit;
                        }
                        s_q[w][cnt] = js * 32 + lane;
                    }
                    ++cnt;
                }
                ITB_TL(12);
                // chunks whose maximum ties with T (equal logits in different chunks: the lower index must win, so they count).
                // The queue holds 64 chunks; a row with more (hundreds of equal maxima) takes the streaming form below.
                // (One copy of the insertion code for the whole kernel path: 17 inlined copies cost more in instruction
                // fetch than the ranking itself.)
#pragma unroll
                for (int j = 0; j < CPL; ++j) {
                    const bool mine = cw[j] == T && T > -INFINITY;
                    const unsigned bal = __ballot_sync(0xffffffffu, mine);
                    if (bal != 0u) {
                        if (cnt + __popc(bal) > 64) streamed = true;
                        else {
                            if (mine) s_q[w][cnt + __popc(bal & ((1u << lane) - 1u))] = j * 32 + lane;
                            cnt += __popc(bal);
                        }
                    }
                }
                __syncwarp();
                ITB_TL(13);
                if (!streamed)
                    for (int base = 0; base < cnt; base += 32) insert_chunks(base, min(32, cnt - base));
                ITB_TL(14);
            }
            if (streamed) {
                mm = -INFINITY;
                for (int ci = lane; ci < nc; ci += 32) mm = fmaxf(mm, __ldcg(cmp + ci));
                mm = warp_max(mm);
                ss = 0.f;
                for (int ci = lane; ci < nc; ci += 32) {
                    const float cm = __ldcg(cmp + ci);
                    if (cm != -INFINITY) ss += __ldcg(csp + ci) * expf(cm - mm);
                }
                ss = warp_sum(ss);
                logz = logf(ss);
                // T = K-th largest chunk maximum (a real logit each): a lower bound of the row's K-th largest logit, so only
                // chunks whose maximum reaches T can hold a top-K entry -- typically K..2K of the n_chunk chunks.  Found with
                // the warp-level sorted list (entry j in lane j) over the maxima alone.
                float tl = -INFINITY, T = -INFINITY;
                for (int ci0 = 0; ci0 < nc; ci0 += 32) {
                    const int ci = ci0 + lane;
                    const float cm = ci < nc ? __ldcg(cmp + ci) : -INFINITY;
                    unsigned q = __ballot_sync(0xffffffffu, cm > T);
                    while (q != 0u) {
                        const int src = __ffs(q) - 1;
                        q &= q - 1;
                        const float cx = __shfl_sync(0xffffffffu, cm, src);
                        if (cx > T) {                                            // warp-uniform; T may have risen meanwhile
                            const int pos = __popc(__ballot_sync(0xffffffffu, lane < K && tl >= cx));
                            const float uv = __shfl_up_sync(0xffffffffu, tl, 1);
                            if (lane < K) {
                                if (lane == pos) tl = cx;
                                else if (lane > pos) tl = uv;
                            }
                            T = __shfl_sync(0xffffffffu, tl, K - 1);
                        }
                    }
                }
                // streaming top-K over the candidates of the qualifying chunks: in every batch of 32 chunks the lane of a
                // qualifying chunk loads that chunk's KT (value, index) pairs (one round trip for the batch), then the warp
                // inserts rank by rank with the list's threshold -- any number of chunks, no staging buffer
                for (int ci0 = 0; ci0 < nc; ci0 += 32) {
                    const int ci = ci0 + lane;
                    const float cm = ci < nc ? __ldcg(cmp + ci) : -INFINITY;
                    const bool mine = cm >= T && cm > -INFINITY;
                    if (__ballot_sync(0xffffffffu, mine) == 0u) continue;
                    float xv[KT]; int xj[KT];
    #pragma unroll
                    for (int u = 0; u < KT; u += 4) {
                        float4 v4 = make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY);
                        int4 j4 = make_int4(-1, -1, -1, -1);
                        if (mine) {
                            v4 = __ldcg(reinterpret_cast<const float4*>(a.cand_val + ((size_t)rr * nc + ci) * KT + u));
                            j4 = __ldcg(reinterpret_cast<const int4*>(a.cand_idx + ((size_t)rr * nc + ci) * KT + u));
                        }
                        xv[u] = v4.x; xv[u + 1] = v4.y; xv[u + 2] = v4.z; xv[u + 3] = v4.w;
                        xj[u] = j4.x; xj[u + 1] = j4.y; xj[u + 2] = j4.z; xj[u + 3] = j4.w;
                    }
    #pragma unroll
                    for (int u = 0; u < KT; ++u) {
                        unsigned q = __ballot_sync(0xffffffffu, xj[u] >= 0 && xv[u] >= thr && xv[u] > -INFINITY);
                        while (q != 0u) {
                            const int src = __ffs(q) - 1;
                            q &= q - 1;
                            const float cx = __shfl_sync(0xffffffffu, xv[u], src);
                            const int cxi = __shfl_sync(0xffffffffu, xj[u], src);
                            const unsigned bet = __ballot_sync(0xffffffffu, better(lv, li, cx, cxi)) & ((1u << KT) - 1u);
                            const int pos = __popc(bet);
                            const float uv = __shfl_up_sync(0xffffffffu, lv, 1);
                            const int ui = __shfl_up_sync(0xffffffffu, li, 1);
                            if (lane < KT) {
                                if (lane == pos) { lv = cx; li = cxi; }
                                else if (lane > pos) { lv = uv; li = ui; }
                            }
                            thr = __shfl_sync(0xffffffffu, lv, KT - 1);
                        }
                    }
                }
            }
            if (lane < K) { s_rtop_val[w][lane] = lv; s_rtop_idx[w][lane] = li; }
            __syncwarp();
            if (lane < K) {
                // log_softmax as ATen computes it: (x - max) - log(sum exp(x - max))   (seq_gen.py:194)
                float lpz = (s_rtop_val[w][lane] - mm) - logz;
                float cv = sc + lpz;                                                        // :199 / :201
                if (K > 1) cv = cv / lp;
                rc_val[kr * K + lane] = cv;
                rc_idx[kr * K + lane] = kr * V + s_rtop_idx[w][lane];
            }
        }
    }
    __syncthreads();
    ITB_TL(2);
    if (w == 0) {
        warp_select(rc_val, rc_idx, nrows * K, K, s_top_val, s_top_idx, lane);              // :203
        if (lane < K) {
            int idx = s_top_idx[lane];
            if (!first) {
                if (lim) idx = a.pad_id;                                                    // :207
                if (a.fin[(size_t)cur * M + b * a.beam + lane] != 0) idx = a.pad_id;        // :213 (new slot, old flag)
            }
            const int src = idx / V, word = idx % V;                                        // :216, :223
            s_src[lane] = src; s_word[lane] = word;
            const int rn = b * a.beam + lane, ro = b * a.beam + src;
            a.scores[(size_t)nxt * M + rn] = s_top_val[lane];                               // :227
            a.sizes[(size_t)nxt * M + rn] = a.sizes[(size_t)cur * M + ro] + ((word != a.pad_id) ? 1.0f : 0.0f);   // :226
            unsigned char f = (a.fin[(size_t)cur * M + ro] != 0 || word == a.eos_id) ? 1 : 0;
            a.fin[(size_t)nxt * M + rn] = f;
            unsigned nf = __ballot_sync(__activemask(), f != 0);
            if (lane == 0) atomicAdd(&a.grp[gid].fin_rows, __popc(nf));
        }
    }
    __syncthreads();
    ITB_TL(3);
    // new prefixes (:217-224) and ancestor table (replaces the KV gather)
    for (int j = 0; j < K; ++j) {
        const int src = s_src[j];
        const int* told = a.tokens + ((size_t)cur * M + b * a.beam + src) * a.T_max;
        int* tnew = a.tokens + ((size_t)nxt * M + b * a.beam + j) * a.T_max;
        const unsigned char* aold = a.anc + ((size_t)cur * M + b * a.beam + src) * a.T_max;
        unsigned char* anew = a.anc + ((size_t)nxt * M + b * a.beam + j) * a.T_max;
        for (int p = tid; p < step; p += blockDim.x) {
            tnew[p] = told[p];
            anew[p] = (p == step - 1) ? (unsigned char)src : aold[p];
        }
        if (tid == 0) tnew[step] = s_word[j];
    }
    }   // !gdone
    __syncthreads();
    ITB_TL(4);
    if (tid == 0) {
        a.sent_counter[b] = 0;
        __threadfence();
        int t = atomicAdd(&ctl->sent_done, 1);
        s_last = (t == a.B - 1);
        if (s_last) __threadfence();
    }
    __syncthreads();
    if (!s_last) return;
    ITB_TL_ANY(5);
    // ---- the last sentence of the step: per-group termination (seq_gen.py:131,134), then advance the step ----
    __shared__ int s_all_done;
    if (tid == 0) s_all_done = 1;
    __syncthreads();
    for (int g = tid; g < a.G; g += blockDim.x) {
        ItbGrp* gp = a.grp + g;
        if (gp->done == 0) {
            const int fin_rows = atomicAdd(&gp->fin_rows, 0);
            gp->steps_run = step;
            gp->fin_rows = 0;
            if (fin_rows == gp->n_rows || step + 1 >= gp->max_len) { gp->done = 1; gp->final_cur = nxt; }
            else atomicAnd(&s_all_done, 0);
        }
    }
    __syncthreads();
    __shared__ int s_hi;                 // 1 + index of the last sentence whose group is still searching
    if (tid == 0) s_hi = 0;
    __syncthreads();
    for (int i = tid; i < a.B; i += blockDim.x) {
        if (a.sent_done[i] == 0 && a.grp[a.sent_grp[i]].done != 0) a.sent_done[i] = 1;
        if (a.sent_done[i] == 0) atomicMax(&s_hi, i + 1);
    }
    __syncthreads();
    if (tid == 0) {
        ctl->rows_active = max(s_hi, 1) * a.beam;
        ctl->steps_run = step;
        ctl->step = step + 1;
        ctl->cur = nxt;
        ctl->sent_done = 0;
        ctl->done = s_all_done;
    }
    ITB_TL_ANY(6);
}

static int kt_for(int beam) { return beam <= 4 ? 4 : (beam <= 8 ? 8 : 16); }

extern "C" {

int itb_topk_capacity(int beam) { return kt_for(beam); }

int itb_beam_step(const itb_config* c, void* stream) {
    ITB_REQUIRE(c->beam >= 1 && c->beam <= 16, "beam_step: beam width must be in [1,16]");
    ITB_REQUIRE(c->V >= c->beam, "beam_step: vocabulary smaller than the beam");
    ITB_REQUIRE(c->n_chunk >= 1 && (c->fused_vocab != 0 || c->n_chunk <= 32), "beam_step: n_chunk must be in [1,32]");
    BeamArgs a;
    a.ctl = reinterpret_cast<ItbCtl*>(c->ctl);
    a.grp = reinterpret_cast<ItbGrp*>(c->grp); a.sent_grp = c->sent_grp; a.sent_done = c->sent_done; a.G = c->G;
    ITB_REQUIRE(a.grp != nullptr && a.sent_grp != nullptr && a.sent_done != nullptr && a.G >= 1, "beam_step: group arrays missing");
    a.logits = c->logits; a.ldl = (c->V + 3) / 4 * 4;
    a.B = c->B; a.beam = c->beam; a.V = c->V; a.T_max = c->T_max; a.n_chunk = c->n_chunk;
    a.chunk = (((c->V + c->n_chunk - 1) / c->n_chunk) + 3) / 4 * 4;
    a.pad_id = c->pad_id; a.eos_id = c->eos_id; a.alpha = c->len_alpha;
    a.max_lens = c->max_lens; a.tokens = c->tokens; a.anc = c->anc; a.scores = c->scores; a.sizes = c->sizes;
    a.fin = c->fin; a.cand_val = c->cand_val; a.cand_idx = c->cand_idx; a.chunk_max = c->chunk_max;
    a.chunk_sum = c->chunk_sum; a.sent_counter = c->sent_counter;
    a.fused = c->fused_vocab != 0 ? 1 : 0;
    dim3 grid(a.fused ? 1 : c->n_chunk, a.fused ? c->B : c->B * c->beam);
    cudaStream_t s = (cudaStream_t)stream;
    cudaError_t e;
    const bool precise = (c->dtype == ITB_F32);
    switch (kt_for(c->beam)) {
        case 4: e = precise ? itb_launch(beam_step_kernel<4, true>, grid, dim3(256), (size_t)0, s, a)
                            : itb_launch(beam_step_kernel<4, false>, grid, dim3(256), (size_t)0, s, a); break;
        case 8: e = precise ? itb_launch(beam_step_kernel<8, true>, grid, dim3(256), (size_t)0, s, a)
                            : itb_launch(beam_step_kernel<8, false>, grid, dim3(256), (size_t)0, s, a); break;
        default: e = precise ? itb_launch(beam_step_kernel<16, true>, grid, dim3(256), (size_t)0, s, a)
                             : itb_launch(beam_step_kernel<16, false>, grid, dim3(256), (size_t)0, s, a); break;
    }
    if (e != cudaSuccess) { itb_set_error("beam_step: launch failed: %s", cudaGetErrorString(e)); return -1; }
    return itb_check_launch("beam_step");
}

}  // extern "C"
