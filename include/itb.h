/* itb.h -- C ABI of the B200-native ImageTranslate beam-search decode path.
 *
 * The reference (rasoolims/ImageTranslate) is pure Python: it has no FFI / plugin
 * registry, its boundary for this path is the Python class
 * `BeamDecoder` (reference src/seq_gen.py:27-48).  This header is therefore the
 * boundary a maintainer binds with ctypes; every entry point
 * names the reference lines whose arithmetic it replaces.
 *
 * Conventions: plain pointers and sizes only (no torch types); every pointer is a
 * DEVICE pointer unless its name ends in `_host`; `stream` is a cudaStream_t passed
 * as void*; functions only ENQUEUE work (no sync, no allocation of device memory)
 * and return 0 on success or a negative code, with text from itb_last_error().
 * Matrices are row-major; "act" means the compute dtype of the decoder
 * (ITB_F32: float, ITB_BF16: __nv_bfloat16, ITB_F16: __half); linear weights are [out][in] exactly
 * as torch.nn.Linear stores them (HF BertSelfAttention.query.weight etc.).
 */
#ifndef ITB_H_
#define ITB_H_
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define ITB_ABI_VERSION 4
#define ITB_STATIC_SCRATCH_FLOATS (2 * 1024 * 1024)   /* floats of the library's file-scope device scratch (8 MiB per GPU):
                                                       * itb_linear_add_ln_static accepts M * d <= this / ITB_Y_PARTIALS */
#define ITB_Y_PARTIALS 4   /* cfg.y holds up to this many split-K partial outputs: [ITB_Y_PARTIALS][M][d] floats */
/* dtype codes: (linear-weight type, activation / KV-cache type).  Accumulation, softmax, LayerNorm statistics,
 * logits and beam scores are always fp32. */
enum { ITB_F32 = 0,        /* float weights,  float activations   (parity mode, CUDA cores)                 */
       ITB_BF16 = 1,       /* bf16 weights,   bf16 activations    (tcgen05 kind::f16)                        */
       ITB_F16 = 2 };      /* fp16 weights,   fp16 activations    (tcgen05 kind::f16; the reference's --fp16 / apex O2) */
enum { ITB_GEMM_AUTO = 0, ITB_GEMM_SIMT = 1, ITB_GEMM_TCGEN05 = 2 };
/* GEMM epilogues */
enum { ITB_EPI_BIAS_F32 = 0,      /* out f32  = acc + bias                                  */
       ITB_EPI_BIAS_GELU_ACT = 1, /* out act  = gelu_erf(acc + bias)  (HF BertIntermediate) */
       ITB_EPI_BIAS_ACT = 2,      /* out act  = acc + bias                                  */
       /* internal to the fused vocabulary projection (csrc/gemm_tc.cu): per-chunk max / sum-exp / top-K, K = 4, 8, 16 */
       ITB_EPI_VOCAB_K4 = 3, ITB_EPI_VOCAB_K8 = 4, ITB_EPI_VOCAB_K16 = 5 };

/* Control block living in device memory (see csrc/common.cuh for field meaning). */
typedef struct { int step, done, steps_run, cur, fin_rows, sent_done, max_len, rows_active; } itb_ctl;   /* rows_active: rows of unfinished
 * groups form a prefix [0, rows_active) when the groups are laid out longest loop bound first; dense kernels skip the rest */
/* One GROUP = one reference batch (one BeamDecoder.forward call of the reference).  Several groups may be decoded
 * in lockstep as one super-batch: rows of all groups share every kernel launch (and every weight byte streamed), but
 * each group keeps the reference's per-batch loop bound and termination rule (seq_gen.py:113-117,131-136). */
typedef struct { int done, steps_run, max_len, fin_rows, n_rows, final_cur, pad0_, pad1_; } itb_grp;

/* One decoder layer of HF-2.9.0 BertLayer(is_decoder=True). */
typedef struct {
  const void* w_qkv;  const float* b_qkv;   /* [3d][d]: attention.self.{query,key,value} stacked      */
  const void* w_o;    const float* b_o;     /* attention.output.dense                                   */
  const float* ln1_g; const float* ln1_b;   /* attention.output.LayerNorm                               */
  const void* w_cq;   const float* b_cq;    /* crossattention.self.query                                */
  const void* w_co;   const float* b_co;    /* crossattention.output.dense                              */
  const float* ln2_g; const float* ln2_b;   /* crossattention.output.LayerNorm                          */
  const void* w_1;    const float* b_1;     /* intermediate.dense [F][d]                                */
  const void* w_2;    const float* b_2;     /* output.dense [d][F]                                      */
  const float* ln3_g; const float* ln3_b;   /* output.LayerNorm                                         */
} itb_layer;

/* Everything one decode call needs.  Filled by the host (imagetranslate_b200/engine.py). */
typedef struct {
  int dtype, gemm_impl;
  int B, beam, d, H, F, L, V, S, T_max;     /* M = B*beam rows; S = memory length; T_max = max tokens/row */
  int pad_id, eos_id, n_chunk, fused_vocab; /* fused_vocab: the tcgen05 vocabulary GEMM emits the per-chunk partials itself
                                             * (16-bit modes); then n_chunk = itb_vocab_chunks(V) and `logits` may be NULL */
  float len_alpha, ln_eps;                  /* length-penalty exponent (seq_gen.py:43); LayerNorm eps     */
  /* decoder embeddings (f32 tables) -- HF BertEmbeddings */
  const float *emb_word, *emb_pos, *emb_type, *emb_ln_g, *emb_ln_b;
  const itb_layer* layers_host;             /* HOST array of L entries (copied at create)                 */
  const void* w_vocab; const float* b_vocab;/* BertOutputLayer.layer (bert_seq2seq.py:6-12) [V][d], [V]   */
  /* per-call inputs */
  const int* tgt_lang;                      /* [B] token-type id of the target language (seq_gen.py:148)  */
  const int* max_lens;                      /* [B] (seq_gen.py:119-121)                                    */
  const int* sent_grp;                      /* [B] group (reference batch) of each sentence                */
  const int* mem_len;                       /* [B] memory positions of the sentence's group (<= S)         */
  itb_grp* grp;                             /* [G] per-group control blocks                                */
  int* sent_done;                           /* [B] 1 once the sentence's group has terminated              */
  int G, reserved1;
  const unsigned char* src_mask;            /* [B][S] 1 = attend, or NULL (image memory, seq_gen.py:153)   */
  const void* cross_kv;                     /* act [B][S][L][2][d]: cross K,V of every layer                */
  /* search state */
  itb_ctl* ctl;
  int* tokens;                              /* [2][M][T_max]   ping-pong prefixes                          */
  unsigned char* anc;                       /* [2][M][T_max]   beam slot that physically holds position p  */
  float* scores;                            /* [2][M]                                                      */
  float* sizes;                             /* [2][M]                                                      */
  unsigned char* fin;                       /* [2][M]          prefix holds EOS                            */
  void* kv_self;                            /* act [L][M][T_max][2][d]                                     */
  /* scratch */
  float* x_f32; void* x_act;                /* [M][d] residual stream and its act-typed copy               */
  float* qkv;                               /* [M][3d]                                                     */
  void* ctx;                                /* act [M][d]                                                  */
  float* y;                                 /* [ITB_Y_PARTIALS][M][d] split-K partials of the d-wide GEMMs */
  void* ffn_h;                              /* act [M][F]                                                  */
  float* logits;                            /* [M][V]                                                      */
  float* cand_val; int* cand_idx;           /* [M][n_chunk][KT]                                            */
  float* chunk_max; float* chunk_sum;       /* [M][n_chunk]                                                */
  int* sent_counter;                        /* [B] zero-initialised                                        */
  /* ---- optional SECOND decoder stream (reference src/seq_gen.py:167-188): the object-feature stream through
   * `obj_decoder` (:167-179), or the image memory through the SAME decoder in the text+image mode (:180-188).  Each step
   * runs a second pass of a decoder layer stack over a second memory with its own KV cache (the ancestor table is
   * shared: same beam); the two last states are mixed by sigmoid(stream_gate + gate_eps) before the vocabulary
   * projection. ---- */
  int two_stream, S2;                       /* S2 = second memory length                                   */
  const itb_layer* layers2_host;            /* HOST array of L entries                                     */
  const float *emb2_word, *emb2_pos, *emb2_type, *emb2_ln_g, *emb2_ln_b;   /* embeddings of the second stack */
  const void* cross_kv2;                    /* act [B][S2][L][2][d]                                        */
  const unsigned char* src_mask2;           /* [B][S2] or NULL                                             */
  const int* mem_len2;                      /* [B]                                                         */
  void* kv_self2;                           /* act [L][M][T_max][2][d]                                     */
  const float* stream_gate;                 /* [d] multistream_attention_gate / multimodal_attention_gate  */
  float* x_first;                           /* [M][d] scratch: last state of the first stream              */
  /* ---- optional lexical proposals (reference src/seq2seq.py:110-144, src/seq_gen.py:191-192) ---- */
  int n_prop, reserved2;                    /* proposals per sentence (0 = off)                            */
  const int* proposals;                     /* [B][n_prop] token ids                                       */
  const float* prop_emb;                    /* f32 [V][d]: model.proposal_embedding (= encoder word table) */
  const float *lex_gate, *lex_ln_g, *lex_ln_b;   /* lexical_gate [d], lexical_layer_norm                   */
  float lex_ln_eps, gate_eps;               /* gate_eps: 1e-7 (src/seq_gen.py:176,186)                     */
} itb_config;

typedef struct itb_decoder itb_decoder;

int         itb_abi_version(void);
const char* itb_last_error(void);
int         itb_topk_capacity(int beam);
int         itb_vocab_chunks(int V);      /* n_chunk of the fused vocabulary projection: 2 * ceil(V / 256) */
int         itb_sizeof_config(void);      /* sizeof(itb_config) / sizeof(itb_layer): binding self-check */
int         itb_sizeof_layer(void);   /* KT: candidate slots per chunk for this beam width */

/* -------- single ops (kernel-level entry points; also used by the parity tests) -------- */

/* out[M][N] = epi(A[M][K] . W[N][K]^T + bias[N]).  Replaces every nn.Linear on the path
 * (HF BertSelfAttention.query/key/value, BertSelfOutput.dense, BertIntermediate.dense,
 * BertOutput.dense; reference src/bert_seq2seq.py:9-12 for the vocabulary projection).
 * dtype = type of A and W; ldo = row stride of `out` in elements. */
int itb_linear(int dtype, int impl, int epi, const void* A, const void* W, const float* bias, void* out,
               int M, int N, int K, int ldo, void* stream);

/* Fused vocabulary projection of the 16-bit modes (tcgen05): logits = x[rows][K] . w_vocab[V][K]^T + b_vocab are produced
 * in TMEM and reduced in the GEMM epilogue to, per (row, chunk c of 128 vocabulary entries; n_chunk = itb_vocab_chunks(V)):
 *   chunk_max[row][c], chunk_sum[row][c] = sum exp(logit - chunk_max), and the chunk's top-KT (logit, index) pairs
 *   cand_val / cand_idx [row][c][KT], KT = itb_topk_capacity(beam) -- everything log_softmax + topk need
 * (reference src/bert_seq2seq.py:6-12 + src/seq_gen.py:194,203) without an (M,V) tensor in HBM.
 * `logits` ([rows][ceil4(V)] fp32) is optional (parity traces); pass NULL in production. */
int itb_vocab_partials(int dtype, int beam, const void* x, const void* w_vocab, const float* b_vocab, int rows, int V, int K,
                       float* cand_val, int* cand_idx, float* chunk_max, float* chunk_sum, float* logits, void* stream);

/* x = LayerNorm(word[tok] + pos[step-1] + type[lang]); HF BertEmbeddings (Appendix A.2), for the ONE
 * new token of each row (reference recomputes the whole prefix: src/seq_gen.py:164-166). */
int itb_embed_ln(int dtype, const itb_ctl* ctl, const int* tokens, const int* tgt_lang, const float* word,
                 const float* pos, const float* type, const float* g, const float* b, float eps, float* x_f32,
                 void* x_act, int M, int beam, int d, int T_max, void* stream);

/* Encoder embeddings for the whole padded source batch: x[b*S+p] = LayerNorm(word[tokens[b][p]] + pos[p] + type[lang[b]]).
 * HF BertEmbeddings as called by the reference encoder (src/bert_seq2seq.py:103-129; token types = source language,
 * src/seq_gen.py:94-96).  The rest of the encoder layer stack runs on itb_linear / itb_cross_attention (queries = the S
 * positions of a sentence, beam := S, L := 1, layer := 0 over a [B][S][1][2][d] K,V buffer) / itb_add_ln. */
int itb_embed_seq_ln(int dtype, const int* tokens, const int* lang, const float* word, const float* pos, const float* type,
                     const float* g, const float* b, float eps, float* x_f32, void* x_act, int B, int S, int d, void* stream);

/* x = LayerNorm(sum_{s<n_partial} y[s] + x) in place on x_f32, and act copy.  HF BertSelfOutput / BertOutput
 * (A.4, A.6).  y is [n_partial][M][d]: the split-K partial outputs of the producing GEMM, summed in a fixed order. */
int itb_add_ln(int dtype, const itb_ctl* ctl, const float* y, int n_partial, const float* g, const float* b, float eps,
               float* x_f32, void* x_act, int M, int d, void* stream);

/* Single-query causal self-attention over the KV cache through the ancestor table; appends the
 * row's new K,V at position step-1.  HF BertSelfAttention (A.3) on the last position only. */
int itb_self_attention(int dtype, const itb_ctl* ctl, const int* sent_done, const float* qkv, void* kv_layer,
                       const unsigned char* anc, void* ctx, int M, int beam, int d, int H, int T_max, void* stream);

/* Single-query cross-attention over the per-sentence memory K,V with the additive -10000 source mask
 * (reference src/bert_seq2seq.py:25-38). q has row stride ldq. */
int itb_cross_attention(int dtype, const itb_ctl* ctl, const int* sent_done, const int* mem_len, const float* q, int ldq,
                        const void* cross_kv, int layer, const unsigned char* src_mask, void* ctx, int M, int beam, int d,
                        int H, int S, int L, void* stream);

/* Two-stream decoding (reference src/seq_gen.py:167-188): x_first := x_f32 (parks the first stream's last state), and
 * x = sigmoid(gate + gate_eps) * x_first + (1 - sigmoid(gate + gate_eps)) * x, written to x_f32 and its act copy. */
int itb_save_state(const itb_ctl* ctl, const float* x_f32, float* x_first, int M, int d, void* stream);
int itb_gate_mix(int dtype, const itb_ctl* ctl, const float* x_first, const float* gate, float gate_eps, float* x_f32,
                 void* x_act, int M, int d, void* stream);

/* Lexical proposals on the last state (reference src/seq2seq.py:110-144 attend_proposal; src/seq_gen.py:191-192):
 * x[r] = LayerNorm(sigmoid(lex_gate + 1e-8) * x[r] + (1 - sigmoid(..)) * sum_p softmax_p(x[r] . E[prop[r/beam][p]]) E[..]),
 * context := 1e-8 for sentences whose proposals are all PAD.  The reference's pad mask is a no-op (it fills a copy); so here. */
int itb_attend_proposal(int dtype, const itb_ctl* ctl, const int* proposals, int n_prop, int pad_id, const float* prop_emb,
                        const float* lex_gate, const float* ln_g, const float* ln_b, float eps, float* x_f32, void* x_act,
                        int M, int beam, int d, void* stream);

/* Physical KV-cache reorder by back-pointer: kv_dst[l][r][p] = kv_src[l][src_row[r]][p], p < n_pos, caches [L][M][T_max][2][d].
 * What a literal KV-cached port of reference src/seq_gen.py:217-219 (gather of the prefixes by beam index) would run every
 * step.  NOT on the decode path (the ancestor table in itb_beam_step / itb_self_attention replaces it); exported so the two
 * variants can be measured side by side. */
int itb_kv_gather(int dtype, const void* kv_src, void* kv_dst, const int* src_row, int M, int T_max, int n_pos, int d, int L,
                  void* stream);

/* Fused log-softmax + finished/over-length row zeroing + score add + length-penalty divide + top-k over
 * beam*V + forced-PAD rewrites + back-pointer / prefix / size / score / ancestor-table update + termination
 * bookkeeping.  Replaces reference src/seq_gen.py:134-143 and :194-227. */
int itb_beam_step(const itb_config* cfg_host, void* stream);

/* Writes slot 0 of every sentence: out_tokens[B][T_max] and out_len[B] (first-EOS cut or max_lens;
 * reference src/seq_gen.py:229-231 and :6-24).  out_len2[B] = number of tokens produced (unpad_output=False). */
int itb_finalize(const itb_ctl* ctl, const itb_grp* grp, const int* sent_grp, const int* tokens, const int* max_lens,
                 int eos, int* out_tokens, int* out_len, int* out_len_full, int B, int beam, int T_max, void* stream);

/* Resets ctl, the group blocks and the search state for a new (super-)batch (seq_gen.py:109-123).
 * first_tokens: [B]; grp_max_len: [G] loop bound of each group (seq_gen.py:113-117), all <= T_max. */
int itb_init_state(const itb_config* cfg_host, const int* first_tokens, const int* grp_max_len, void* stream);

/* -------- whole decoder step -------- */
int itb_decoder_create(const itb_config* cfg_host, itb_decoder** out);
int itb_decoder_destroy(itb_decoder* dec);
/* Enqueues one full search step i (embed -> L decoder layers -> vocab projection -> beam step).
 * Graph-capturable: the step index lives in cfg->ctl. */
int itb_decoder_step(itb_decoder* dec, void* stream);
/* Captures n_steps (1..64) consecutive itb_decoder_step()s into one CUDA graph owned by `dec`; itb_decoder_launch
 * replays it on `stream`.  Steps replayed after the search terminated (seq_gen.py:134-136) are no-ops, so the host
 * may over-enqueue and only has to look at ctl.done now and then (the reference synchronises every step). */
int itb_decoder_capture(itb_decoder* dec, int n_steps);
int itb_decoder_launch(itb_decoder* dec, void* stream);
/* One-off per batch: cross K,V of all layers = memory[B*S][d] . Wckv[L*2d][d]^T + b  (HF BertSelfAttention
 * key/value on encoder_hidden_states; the reference redoes it every step: SURVEY K5). */
int itb_cross_kv(int dtype, int impl, const void* memory, const void* w_ckv, const float* b_ckv, void* cross_kv,
                 int rows, int d, int L, void* stream);


/* -------- allocation-free variant -------- */
/* x = LayerNorm(A[M][K] . W[d][K]^T + bias + x), in place on x_f32 with its act copy: itb_linear (f32 epilogue, the decode
 * path's split-K plan) followed by itb_add_ln, with the split-K partial outputs held in a file-scope device array of the
 * library instead of a caller-allocated buffer.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot allocate device memory (CUDA graph capture, serving loops, harnesses that forbid
 * cudaMalloc).  Allocates nothing (no cudaMalloc / cudaFree / allocating library call).  Limit: M * d <=
 * ITB_STATIC_SCRATCH_FLOATS / ITB_Y_PARTIALS = 524288 (682 rows at d = 768); larger problems are rejected with -2 and a
 * message in itb_last_error().  The array (8 MiB) is held by every GPU that loads the library until the process exits and is
 * shared by every call on that GPU: calls that overlap on the GPU race on it (calls on one stream are ordered).  It keeps
 * the last call's partials; nothing in it needs to start at zero (the GEMM overwrites all that add_ln reads). */
int itb_linear_add_ln_static(int dtype, int impl, const void* A, const void* W, const float* bias, const float* g,
                             const float* b, float eps, float* x_f32, void* x_act, int M, int d, int K, void* stream);
int itb_static_scratch_floats(void);   /* ITB_STATIC_SCRATCH_FLOATS of the loaded library */

#ifdef __cplusplus
}
#endif
#endif
