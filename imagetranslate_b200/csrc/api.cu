// C-ABI glue: error reporting, GEMM dispatch, the whole-step driver (include/itb.h).
#include <cuda.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <vector>
#include "common.cuh"
#include "launch.h"

int itb_gemm_simt(int dtype, int epi, const void* ctl, const void* A, const void* W, const float* bias, void* out,
                  int M, int N, int K, int ldo, cudaStream_t s);
int itb_gemm_tc(int epi, const CUtensorMap* tmW, const CUtensorMap* tmA, const void* ctl, const float* bias, void* out,
                int M, int N, int K, int ldo, int act_f16, int mt, int nsplit, cudaStream_t s);
int itb_make_tmap_16b(CUtensorMap* map, const void* base, int rows, int K, int box_rows, int is_f16);
void itb_tc_plan(int M, int N, int K, int max_split, int* mt_out, int* split_out);
bool itb_tc_pair_applies(int M, int N, int K);
int itb_gemm_tc_pair(int epi, const CUtensorMap* tmW, const CUtensorMap* tmA128, const void* ctl, const float* bias, void* out,
                     int M, int N, int K, int ldo, int act_f16, cudaStream_t s);
int itb_gemm_tc_vocab(int kt, const CUtensorMap* tmX, const CUtensorMap* tmWv, const CUtensorMap* tmWv128, const void* ctl,
                      const float* bias, int rows, int V, int K, int act_f16, float* cand_val, int* cand_idx, float* chunk_max,
                      float* chunk_sum, float* logits, int ldl, int n_chunk, cudaStream_t s);

static thread_local char g_err[512] = "";

void itb_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int itb_check_launch(const char* what) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        itb_set_error("%s: launch failed: %s", what, cudaGetErrorString(e));
        return -1;
    }
    return 0;
}

bool itb_pdl_enabled() {
    static int v = -1;
    if (v < 0) {
        const char* e = getenv("ITB_PDL");
        v = (e != nullptr && e[0] == '0') ? 0 : 1;
    }
    return v != 0;
}

static bool use_tc(int dtype, int impl) {
    if (impl == ITB_GEMM_SIMT) return false;
    return dtype != ITB_F32;       // fp32 (parity mode) always runs on the CUDA cores
}

// One linear with its cached TMA descriptors.
struct Linear {
    const void* W = nullptr; const float* bias = nullptr; const void* A = nullptr; void* out = nullptr;
    int M = 0, N = 0, K = 0, ldo = 0, epi = 0;
    int max_split = 1;       // > 1: `out` is [max_split][M][ldo] and the consumer sums `nsplit` partial outputs
    int mt = 32, nsplit = 1;
    bool tc = false, pair = false;   // pair: CTA-pair launch (256 x 256 tiles, no split-K; gemm_tc.cu)
    CUtensorMap tmW, tmA;
    int prepare(int dtype, int impl) {
        tc = use_tc(dtype, impl);
        if (!tc) return 0;
        pair = itb_tc_pair_applies(M, N, K);
        if (pair) { mt = 128; nsplit = 1; }
        else itb_tc_plan(M, N, K, max_split, &mt, &nsplit);
        int r = itb_make_tmap_16b(&tmW, W, N, K, 128, dtype == ITB_F16 ? 1 : 0);
        if (r != 0) return r;
        return itb_make_tmap_16b(&tmA, A, M, K, mt, dtype == ITB_F16 ? 1 : 0);
    }
    int run(int dtype, const void* ctl, cudaStream_t s) const {
        if (tc && pair) return itb_gemm_tc_pair(epi, &tmW, &tmA, ctl, bias, out, M, N, K, ldo, dtype == ITB_F16 ? 1 : 0, s);
        if (tc) return itb_gemm_tc(epi, &tmW, &tmA, ctl, bias, out, M, N, K, ldo, dtype == ITB_F16 ? 1 : 0, mt, nsplit, s);
        return itb_gemm_simt(dtype, epi, ctl, A, W, bias, out, M, N, K, ldo, s);
    }
};

static Linear make_linear(const void* A, const void* W, const float* bias, void* out, int M, int N, int K, int ldo, int epi,
                          int max_split = 1) {
    Linear l;
    l.A = A; l.W = W; l.bias = bias; l.out = out; l.M = M; l.N = N; l.K = K; l.ldo = ldo; l.epi = epi; l.max_split = max_split;
    return l;
}

// One decoder layer stack: its weights and the six linears of every layer with their cached TMA descriptors.
struct LayerSet {
    std::vector<itb_layer> layers;
    std::vector<Linear> qkv, o, cq, co, f1, f2;
    int build(const itb_config* c, const itb_layer* src) {
        const int M = c->B * c->beam, D = c->d, F = c->F;
        layers.assign(src, src + c->L);
        int r = 0;
        for (int l = 0; l < c->L && r == 0; ++l) {
            const itb_layer& w = layers[l];
            qkv.push_back(make_linear(c->x_act, w.w_qkv, w.b_qkv, c->qkv, M, 3 * D, D, 3 * D, ITB_EPI_BIAS_F32));
            o.push_back(make_linear(c->ctx, w.w_o, w.b_o, c->y, M, D, D, D, ITB_EPI_BIAS_F32, ITB_Y_PARTIALS));
            cq.push_back(make_linear(c->x_act, w.w_cq, w.b_cq, c->qkv, M, D, D, 3 * D, ITB_EPI_BIAS_F32));
            co.push_back(make_linear(c->ctx, w.w_co, w.b_co, c->y, M, D, D, D, ITB_EPI_BIAS_F32, ITB_Y_PARTIALS));
            f1.push_back(make_linear(c->x_act, w.w_1, w.b_1, c->ffn_h, M, F, D, F, ITB_EPI_BIAS_GELU_ACT));
            f2.push_back(make_linear(c->ffn_h, w.w_2, w.b_2, c->y, M, D, F, D, ITB_EPI_BIAS_F32, ITB_Y_PARTIALS));
            Linear* ls[6] = {&qkv.back(), &o.back(), &cq.back(), &co.back(), &f1.back(), &f2.back()};
            for (int i = 0; i < 6 && r == 0; ++i) r = ls[i]->prepare(c->dtype, c->gemm_impl);
        }
        return r;
    }
};

struct itb_decoder {
    itb_config cfg;
    LayerSet first, second;              // second: obj_decoder / image pass of the same decoder (cfg.two_stream)
    Linear vocab;
    bool vocab_fused = false;            // 16-bit modes: vocabulary GEMM with the softmax / top-k partials in its epilogue
    CUtensorMap tm_vx, tm_vw, tm_vw2;            // fused launch: activations [M][d] box 128, vocabulary weights [V][d] box 256
    cudaGraphExec_t exec = nullptr;      // `graph_steps` search steps captured once, replayed for the whole batch
    int graph_steps = 0;
    cudaStream_t cap_stream = nullptr;
};

__global__ void stream_fence_kernel() {}

int itb_cross_attention_impl(int dtype, const itb_ctl* ctl, const int* sent_done, const int* mem_len, const float* q, int ldq,
                             const void* cross_kv, int layer, const unsigned char* src_mask, void* ctx, int M, int beam, int d,
                             int H, int S, int L, int early_kv, void* stream);

extern "C" {

int itb_abi_version(void) { return ITB_ABI_VERSION; }
const char* itb_last_error(void) { return g_err; }
int itb_sizeof_config(void) { return (int)sizeof(itb_config); }
int itb_sizeof_layer(void) { return (int)sizeof(itb_layer); }
int itb_vocab_chunks(int V) { return 2 * ((V + 255) / 256); }

int itb_linear(int dtype, int impl, int epi, const void* A, const void* W, const float* bias, void* out, int M, int N,
               int K, int ldo, void* stream) {
    ITB_REQUIRE(dtype == ITB_F32 || dtype == ITB_BF16 || dtype == ITB_F16, "linear: unknown dtype");
    ITB_REQUIRE(!(impl == ITB_GEMM_TCGEN05 && dtype == ITB_F32), "linear: the tcgen05 path needs 16-bit operands");
    Linear l = make_linear(A, W, bias, out, M, N, K, ldo, epi);
    int r = l.prepare(dtype, impl);
    if (r != 0) return r;
    return l.run(dtype, nullptr, (cudaStream_t)stream);
}

int itb_vocab_partials(int dtype, int beam, const void* x, const void* w_vocab, const float* b_vocab, int rows, int V, int K,
                       float* cand_val, int* cand_idx, float* chunk_max, float* chunk_sum, float* logits, void* stream) {
    ITB_REQUIRE(dtype == ITB_BF16 || dtype == ITB_F16, "vocab_partials: the fused vocabulary projection needs 16-bit operands");
    ITB_REQUIRE(beam >= 1 && beam <= 16 && rows >= 1 && V >= 1, "vocab_partials: bad sizes");
    CUtensorMap tx, tw, tw2;
    int r = itb_make_tmap_16b(&tx, x, rows, K, 128, dtype == ITB_F16 ? 1 : 0);
    if (r == 0) r = itb_make_tmap_16b(&tw, w_vocab, V, K, 256, dtype == ITB_F16 ? 1 : 0);
    if (r == 0) r = itb_make_tmap_16b(&tw2, w_vocab, V, K, 128, dtype == ITB_F16 ? 1 : 0);
    if (r != 0) return r;
    return itb_gemm_tc_vocab(itb_topk_capacity(beam), &tx, &tw, &tw2, nullptr, b_vocab, rows, V, K, dtype == ITB_F16 ? 1 : 0, cand_val, cand_idx,
                             chunk_max, chunk_sum, logits, (V + 3) / 4 * 4, itb_vocab_chunks(V), (cudaStream_t)stream);
}

int itb_cross_kv(int dtype, int impl, const void* memory, const void* w_ckv, const float* b_ckv, void* cross_kv,
                 int rows, int d, int L, void* stream) {
    int r = itb_linear(dtype, impl, ITB_EPI_BIAS_ACT, memory, w_ckv, b_ckv, cross_kv, rows, L * 2 * d, d, L * 2 * d, stream);
    if (r != 0) return r;
    // A plain (non-programmatic) launch: everything after it in the stream starts only once cross_kv is complete, which
    // is what lets the decoder's cross-attention fetch K,V ahead of its programmatic-dependency wait.
    stream_fence_kernel<<<1, 32, 0, (cudaStream_t)stream>>>();
    return itb_check_launch("cross_kv fence");
}

int itb_decoder_create(const itb_config* c, itb_decoder** out) {
    ITB_REQUIRE(c != nullptr && out != nullptr, "decoder_create: null argument");
    ITB_REQUIRE(c->dtype == ITB_F32 || c->dtype == ITB_BF16 || c->dtype == ITB_F16, "decoder_create: unknown dtype");
    ITB_REQUIRE(c->d == c->H * 64, "decoder_create: only head_dim 64 is supported (hidden = 64 * heads)");
    ITB_REQUIRE(c->d % 64 == 0 && c->F % 64 == 0, "decoder_create: hidden and FFN sizes must be multiples of 64");
    ITB_REQUIRE(c->beam >= 1 && c->beam <= 16, "decoder_create: beam width must be in [1,16]");
    ITB_REQUIRE(c->B >= 1 && c->L >= 1 && c->V >= c->beam && c->S >= 1 && c->T_max >= 2, "decoder_create: bad sizes");
    if (c->two_stream != 0)
        ITB_REQUIRE(c->layers2_host != nullptr && c->cross_kv2 != nullptr && c->kv_self2 != nullptr && c->stream_gate != nullptr &&
                    c->x_first != nullptr && c->mem_len2 != nullptr && c->emb2_word != nullptr && c->S2 >= 1,
                    "decoder_create: two_stream needs layers2 / emb2 / cross_kv2 / kv_self2 / mem_len2 / stream_gate / x_first");
    if (c->n_prop != 0)
        ITB_REQUIRE(c->n_prop > 0 && c->proposals != nullptr && c->prop_emb != nullptr && c->lex_gate != nullptr &&
                    c->lex_ln_g != nullptr && c->lex_ln_b != nullptr, "decoder_create: proposals need prop_emb / lex_gate / lex_ln");
    itb_decoder* d = new itb_decoder();
    d->cfg = *c;
    const int M = c->B * c->beam, D = c->d;
    const int ldl = (c->V + 3) / 4 * 4;
    int r = d->first.build(c, c->layers_host);
    d->cfg.layers_host = d->first.layers.data();
    if (r == 0 && c->two_stream != 0) {
        r = d->second.build(c, c->layers2_host);
        d->cfg.layers2_host = d->second.layers.data();
    }
    if (r == 0 && c->fused_vocab != 0) {
        if (!use_tc(c->dtype, c->gemm_impl) || c->n_chunk != itb_vocab_chunks(c->V)) {
            itb_set_error("decoder_create: fused_vocab needs a 16-bit tcgen05 decoder and n_chunk == itb_vocab_chunks(V)");
            r = -2;
        } else {
            d->vocab_fused = true;
            r = itb_make_tmap_16b(&d->tm_vx, c->x_act, M, D, 128, c->dtype == ITB_F16 ? 1 : 0);
            if (r == 0) r = itb_make_tmap_16b(&d->tm_vw, c->w_vocab, c->V, D, 256, c->dtype == ITB_F16 ? 1 : 0);
            if (r == 0) r = itb_make_tmap_16b(&d->tm_vw2, c->w_vocab, c->V, D, 128, c->dtype == ITB_F16 ? 1 : 0);
        }
    } else if (r == 0 && c->logits == nullptr) {
        itb_set_error("decoder_create: logits buffer missing");
        r = -2;
    } else if (r == 0) {
        d->vocab = make_linear(c->x_act, c->w_vocab, c->b_vocab, c->logits, M, c->V, D, ldl, ITB_EPI_BIAS_F32);
        r = d->vocab.prepare(c->dtype, c->gemm_impl);
    }
    if (r != 0) { delete d; return r; }
    *out = d;
    return 0;
}

int itb_decoder_destroy(itb_decoder* dec) {
    if (dec != nullptr) {
        if (dec->exec != nullptr) cudaGraphExecDestroy(dec->exec);
        if (dec->cap_stream != nullptr) cudaStreamDestroy(dec->cap_stream);
    }
    delete dec;
    return 0;
}

#define ITB_TRY(x) do { int r_ = (x); if (r_ != 0) return r_; } while (0)

}  // extern "C"

// One pass of a decoder layer stack for the ONE new token of every row: embeddings + LN, then per layer
// QKV -> self-attention (appends K,V) -> O + LN -> CQ -> cross-attention -> CO + LN -> FFN1 (GELU) -> FFN2 + LN.
static int run_stack(itb_decoder* dec, const LayerSet& st, const float* emb_word, const float* emb_pos, const float* emb_type,
                     const float* emb_ln_g, const float* emb_ln_b, void* kv_self, const void* cross_kv,
                     const unsigned char* src_mask, const int* mem_len, int S, cudaStream_t s) {
    const itb_config& c = dec->cfg;
    void* stream = (void*)s;
    const int M = c.B * c.beam, D = c.d;
    const size_t esz = (c.dtype == ITB_F32) ? 4 : 2;
    ITB_TRY(itb_embed_ln(c.dtype, c.ctl, c.tokens, c.tgt_lang, emb_word, emb_pos, emb_type, emb_ln_g, emb_ln_b,
                         c.ln_eps, c.x_f32, c.x_act, M, c.beam, D, c.T_max, stream));
    for (int l = 0; l < c.L; ++l) {
        const itb_layer& w = st.layers[l];
        char* kv_l = reinterpret_cast<char*>(kv_self) + (size_t)l * M * c.T_max * 2 * D * esz;
        ITB_TRY(st.qkv[l].run(c.dtype, c.ctl, s));
        ITB_TRY(itb_self_attention(c.dtype, c.ctl, c.sent_done, c.qkv, kv_l, c.anc, c.ctx, M, c.beam, D, c.H, c.T_max, stream));
        ITB_TRY(st.o[l].run(c.dtype, c.ctl, s));
        ITB_TRY(itb_add_ln(c.dtype, c.ctl, c.y, st.o[l].nsplit, w.ln1_g, w.ln1_b, c.ln_eps, c.x_f32, c.x_act, M, D, stream));
        ITB_TRY(st.cq[l].run(c.dtype, c.ctl, s));
        ITB_TRY(itb_cross_attention_impl(c.dtype, c.ctl, c.sent_done, mem_len, c.qkv, 3 * D, cross_kv, l, src_mask, c.ctx, M, c.beam, D,
                                         c.H, S, c.L, 1, stream));
        ITB_TRY(st.co[l].run(c.dtype, c.ctl, s));
        ITB_TRY(itb_add_ln(c.dtype, c.ctl, c.y, st.co[l].nsplit, w.ln2_g, w.ln2_b, c.ln_eps, c.x_f32, c.x_act, M, D, stream));
        ITB_TRY(st.f1[l].run(c.dtype, c.ctl, s));
        ITB_TRY(st.f2[l].run(c.dtype, c.ctl, s));
        ITB_TRY(itb_add_ln(c.dtype, c.ctl, c.y, st.f2[l].nsplit, w.ln3_g, w.ln3_b, c.ln_eps, c.x_f32, c.x_act, M, D, stream));
    }
    return 0;
}

extern "C" {

int itb_decoder_step(itb_decoder* dec, void* stream) {
    ITB_REQUIRE(dec != nullptr, "decoder_step: null decoder");
    const itb_config& c = dec->cfg;
    cudaStream_t s = (cudaStream_t)stream;
    const int M = c.B * c.beam, D = c.d;
    ITB_TRY(run_stack(dec, dec->first, c.emb_word, c.emb_pos, c.emb_type, c.emb_ln_g, c.emb_ln_b, c.kv_self, c.cross_kv, c.src_mask,
                      c.mem_len, c.S, s));
    if (c.two_stream != 0) {             // seq_gen.py:167-188
        ITB_TRY(itb_save_state(c.ctl, c.x_f32, c.x_first, M, D, stream));
        ITB_TRY(run_stack(dec, dec->second, c.emb2_word, c.emb2_pos, c.emb2_type, c.emb2_ln_g, c.emb2_ln_b, c.kv_self2, c.cross_kv2,
                          c.src_mask2, c.mem_len2, c.S2, s));
        ITB_TRY(itb_gate_mix(c.dtype, c.ctl, c.x_first, c.stream_gate, c.gate_eps, c.x_f32, c.x_act, M, D, stream));
    }
    if (c.n_prop != 0)                   // seq_gen.py:191-192
        ITB_TRY(itb_attend_proposal(c.dtype, c.ctl, c.proposals, c.n_prop, c.pad_id, c.prop_emb, c.lex_gate, c.lex_ln_g, c.lex_ln_b,
                                    c.lex_ln_eps, c.x_f32, c.x_act, M, c.beam, D, stream));
    if (dec->vocab_fused)
        ITB_TRY(itb_gemm_tc_vocab(itb_topk_capacity(c.beam), &dec->tm_vx, &dec->tm_vw, &dec->tm_vw2, c.ctl, c.b_vocab, M, c.V, D, c.dtype == ITB_F16 ? 1 : 0,
                                  c.cand_val, c.cand_idx, c.chunk_max, c.chunk_sum, c.logits, (c.V + 3) / 4 * 4, c.n_chunk, s));
    else
        ITB_TRY(dec->vocab.run(c.dtype, c.ctl, s));
    ITB_TRY(itb_beam_step(&c, stream));
    return 0;
}

// Captures `n_steps` consecutive search steps into one CUDA graph (the step index, ping-pong parity and the
// termination flag live in device memory -- ItbCtl -- so the same graph serves every step of every batch of
// this shape, and steps replayed past termination are no-ops).
int itb_decoder_capture(itb_decoder* dec, int n_steps) {
    ITB_REQUIRE(dec != nullptr && n_steps >= 1 && n_steps <= 64, "decoder_capture: bad arguments");
    if (dec->exec != nullptr && dec->graph_steps == n_steps) return 0;
    if (dec->exec != nullptr) { cudaGraphExecDestroy(dec->exec); dec->exec = nullptr; }
    if (dec->cap_stream == nullptr) {
        cudaError_t e = cudaStreamCreateWithFlags(&dec->cap_stream, cudaStreamNonBlocking);
        if (e != cudaSuccess) { itb_set_error("decoder_capture: cudaStreamCreate: %s", cudaGetErrorString(e)); return -1; }
    }
    cudaError_t e = cudaStreamBeginCapture(dec->cap_stream, cudaStreamCaptureModeRelaxed);
    if (e != cudaSuccess) { itb_set_error("decoder_capture: begin capture: %s", cudaGetErrorString(e)); return -1; }
    int r = 0;
    for (int i = 0; i < n_steps && r == 0; ++i) r = itb_decoder_step(dec, dec->cap_stream);
    cudaGraph_t graph = nullptr;
    e = cudaStreamEndCapture(dec->cap_stream, &graph);
    if (r != 0) { if (graph != nullptr) cudaGraphDestroy(graph); return r; }
    if (e != cudaSuccess || graph == nullptr) { itb_set_error("decoder_capture: end capture: %s", cudaGetErrorString(e)); return -1; }
    e = cudaGraphInstantiate(&dec->exec, graph, 0);
    cudaGraphDestroy(graph);
    if (e != cudaSuccess) { dec->exec = nullptr; itb_set_error("decoder_capture: instantiate: %s", cudaGetErrorString(e)); return -1; }
    dec->graph_steps = n_steps;
    return 0;
}

// Replays the captured graph once (= graph_steps search steps) on `stream`.
int itb_decoder_launch(itb_decoder* dec, void* stream) {
    ITB_REQUIRE(dec != nullptr && dec->exec != nullptr, "decoder_launch: no captured graph (call itb_decoder_capture)");
    cudaError_t e = cudaGraphLaunch(dec->exec, (cudaStream_t)stream);
    if (e != cudaSuccess) { itb_set_error("decoder_launch: %s", cudaGetErrorString(e)); return -1; }
    return 0;
}

}  // extern "C"
