"""ctypes binding of the C-ABI in include/itb.h (imagetranslate_b200/libitb.so).

There is NO fallback: if the CUDA extension is missing or a call fails, a RuntimeError is
raised (the reference drivers catch RuntimeError per batch: src/translate.py:110-120).
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("ITB_LIB") or os.path.join(_HERE, "libitb.so")      # ITB_LIB: A/B builds of the same ABI

F32, BF16, F16 = 0, 1, 2      # (weights, activations): (f32,f32) / (bf16,bf16) / (fp16,fp16); include/itb.h
GEMM_AUTO, GEMM_SIMT, GEMM_TCGEN05 = 0, 1, 2
EPI_BIAS_F32, EPI_BIAS_GELU_ACT, EPI_BIAS_ACT = 0, 1, 2
ABI_VERSION = 4
Y_PARTIALS = 4

P = C.c_void_p


class Ctl(C.Structure):
    _fields_ = [(n, C.c_int) for n in ("step", "done", "steps_run", "cur", "fin_rows", "sent_done", "max_len", "rows_active")]


class Grp(C.Structure):
    _fields_ = [(n, C.c_int) for n in ("done", "steps_run", "max_len", "fin_rows", "n_rows", "final_cur", "pad0_", "pad1_")]


class Layer(C.Structure):
    _fields_ = [(n, P) for n in ("w_qkv", "b_qkv", "w_o", "b_o", "ln1_g", "ln1_b", "w_cq", "b_cq", "w_co", "b_co",
                                 "ln2_g", "ln2_b", "w_1", "b_1", "w_2", "b_2", "ln3_g", "ln3_b")]


class Config(C.Structure):
    _fields_ = (
        [(n, C.c_int) for n in ("dtype", "gemm_impl", "B", "beam", "d", "H", "F", "L", "V", "S", "T_max", "pad_id",
                                "eos_id", "n_chunk", "fused_vocab")]
        + [("len_alpha", C.c_float), ("ln_eps", C.c_float)]
        + [(n, P) for n in ("emb_word", "emb_pos", "emb_type", "emb_ln_g", "emb_ln_b", "layers_host", "w_vocab",
                            "b_vocab", "tgt_lang", "max_lens", "sent_grp", "mem_len", "grp", "sent_done")]
        + [("G", C.c_int), ("reserved1", C.c_int)]
        + [(n, P) for n in ("src_mask", "cross_kv", "ctl", "tokens", "anc", "scores",
                            "sizes", "fin", "kv_self", "x_f32", "x_act", "qkv", "ctx", "y", "ffn_h", "logits",
                            "cand_val", "cand_idx", "chunk_max", "chunk_sum", "sent_counter")]
        + [("two_stream", C.c_int), ("S2", C.c_int)]
        + [(n, P) for n in ("layers2_host", "emb2_word", "emb2_pos", "emb2_type", "emb2_ln_g", "emb2_ln_b", "cross_kv2",
                            "src_mask2", "mem_len2", "kv_self2", "stream_gate", "x_first")]
        + [("n_prop", C.c_int), ("reserved2", C.c_int)]
        + [(n, P) for n in ("proposals", "prop_emb", "lex_gate", "lex_ln_g", "lex_ln_b")]
        + [("lex_ln_eps", C.c_float), ("gate_eps", C.c_float)])


_SIGS = {
    "itb_abi_version": (C.c_int, []),
    "itb_last_error": (C.c_char_p, []),
    "itb_topk_capacity": (C.c_int, [C.c_int]),
    "itb_vocab_chunks": (C.c_int, [C.c_int]),
    "itb_sizeof_config": (C.c_int, []),
    "itb_sizeof_layer": (C.c_int, []),
    "itb_linear": (C.c_int, [C.c_int, C.c_int, C.c_int, P, P, P, P, C.c_int, C.c_int, C.c_int, C.c_int, P]),
    "itb_vocab_partials": (C.c_int, [C.c_int, C.c_int, P, P, P, C.c_int, C.c_int, C.c_int, P, P, P, P, P, P]),
    "itb_embed_ln": (C.c_int, [C.c_int, P, P, P, P, P, P, P, P, C.c_float, P, P, C.c_int, C.c_int, C.c_int, C.c_int, P]),
    "itb_embed_seq_ln": (C.c_int, [C.c_int, P, P, P, P, P, P, P, C.c_float, P, P, C.c_int, C.c_int, C.c_int, P]),
    "itb_add_ln": (C.c_int, [C.c_int, P, P, C.c_int, P, P, C.c_float, P, P, C.c_int, C.c_int, P]),
    "itb_self_attention": (C.c_int, [C.c_int, P, P, P, P, P, P, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, P]),
    "itb_cross_attention": (C.c_int, [C.c_int, P, P, P, P, C.c_int, P, C.c_int, P, P, C.c_int, C.c_int, C.c_int, C.c_int,
                                      C.c_int, C.c_int, P]),
    "itb_save_state": (C.c_int, [P, P, P, C.c_int, C.c_int, P]),
    "itb_gate_mix": (C.c_int, [C.c_int, P, P, P, C.c_float, P, P, C.c_int, C.c_int, P]),
    "itb_attend_proposal": (C.c_int, [C.c_int, P, P, C.c_int, C.c_int, P, P, P, P, C.c_float, P, P, C.c_int, C.c_int, C.c_int, P]),
    "itb_kv_gather": (C.c_int, [C.c_int, P, P, P, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, P]),
    "itb_beam_step": (C.c_int, [C.POINTER(Config), P]),
    "itb_finalize": (C.c_int, [P, P, P, P, P, C.c_int, P, P, P, C.c_int, C.c_int, C.c_int, P]),
    "itb_init_state": (C.c_int, [C.POINTER(Config), P, P, P]),
    "itb_decoder_create": (C.c_int, [C.POINTER(Config), C.POINTER(P)]),
    "itb_decoder_destroy": (C.c_int, [P]),
    "itb_decoder_step": (C.c_int, [P, P]),
    "itb_decoder_capture": (C.c_int, [P, C.c_int]),
    "itb_decoder_launch": (C.c_int, [P, P]),
    "itb_cross_kv": (C.c_int, [C.c_int, C.c_int, P, P, P, P, C.c_int, C.c_int, C.c_int, P]),
    "itb_linear_add_ln_static": (C.c_int, [C.c_int, C.c_int, P, P, P, P, P, C.c_float, P, P, C.c_int, C.c_int, C.c_int, P]),
    "itb_static_scratch_floats": (C.c_int, []),
}
EXPORTS = tuple(sorted(_SIGS))
_lib = None


def lib():
    """Loads libitb.so once; raises RuntimeError when it is missing (no silent fallback)."""
    global _lib
    if _lib is None:
        if not os.path.isfile(LIB_PATH):
            raise RuntimeError("imagetranslate_b200: CUDA extension %s is missing; build it with "
                               "`python -c 'import __graft_entry__ as g; g.build()'` or "
                               "imagetranslate_b200/csrc/build.sh -- there is no CPU fallback" % LIB_PATH)
        l = C.CDLL(LIB_PATH)
        for name, (res, args) in _SIGS.items():
            fn = getattr(l, name)
            fn.restype, fn.argtypes = res, args
        if l.itb_abi_version() != ABI_VERSION:
            raise RuntimeError("imagetranslate_b200: libitb.so ABI %d != expected %d" % (l.itb_abi_version(), ABI_VERSION))
        if l.itb_sizeof_config() != C.sizeof(Config) or l.itb_sizeof_layer() != C.sizeof(Layer):
            raise RuntimeError("imagetranslate_b200: ctypes struct layout does not match include/itb.h")
        _lib = l
    return _lib


def check(code, what=""):
    if code != 0:
        msg = lib().itb_last_error().decode("utf-8", "replace")
        raise RuntimeError("imagetranslate_b200 %s failed (%d): %s" % (what, code, msg))


def ptr(t):
    return None if t is None else C.c_void_p(t.data_ptr())
