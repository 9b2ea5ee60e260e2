"""ORACLE -- TEST INFRASTRUCTURE ONLY.  Generates tests/golden/*.npz by running the REAL
reference BeamDecoder (/root/reference/src/seq_gen.py via oracle/reference_loader.py) on the
seeded cases of oracle/cases.py.  Run here (build container):  python -m oracle.make_golden
The fixtures travel to the GPU box; /root/reference does not."""
import json
import os

import numpy as np
import torch

from . import cases, reference_loader

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


SUB_STRIDE = 251          # vocabulary subsample of the per-step log-probs kept in the recorded fixtures
TOPN = 8


def pack_trace(trace, M, V):
    """Per-step tensors of a recorded run -> small arrays: log-probs at every SUB_STRIDE-th vocabulary entry and the
    TOPN largest (value, index) of every row, the new prefixes and the new beam scores.  Step 1 has one row per
    sentence (src/seq_gen.py:146): rows beyond it are left zero."""
    T = len(trace)
    sub = np.arange(0, V, SUB_STRIDE)
    lp_sub = np.zeros((T, M, len(sub)), dtype=np.float32)
    top_val = np.zeros((T, M, TOPN), dtype=np.float32)
    top_idx = np.zeros((T, M, TOPN), dtype=np.int64)
    n_rows = np.zeros((T,), dtype=np.int64)
    scores = np.zeros((T, M), dtype=np.float32)
    L = trace[-1]["prefixes"].shape[-1]
    prefixes = np.zeros((T, M, L), dtype=np.int64)
    for k, t in enumerate(trace):
        lp = t["log_probs"].float()
        n = lp.shape[0]
        n_rows[k] = n
        lp_sub[k, :n] = lp[:, sub].numpy()
        v, ix = lp.topk(TOPN, dim=-1)
        top_val[k, :n], top_idx[k, :n] = v.numpy(), ix.numpy()
        pre = t["prefixes"].reshape(M, -1)
        prefixes[k, :, :pre.shape[1]] = pre.numpy()
        scores[k] = t["scores"].reshape(-1).numpy()
    return dict(step_logp_sub=lp_sub, step_top_val=top_val, step_top_idx=top_idx, step_rows=n_rows,
                step_prefixes=prefixes, step_scores=scores, sub_index=sub)


def run_reference(name, record=False):
    ns = reference_loader.load()
    model, kw, c = cases.build(name)
    ref = ns.mass_seq2seq.MassSeq2Seq(model.text_processor, lang_dec=c["lang_dec"], tie_embed=c["tie"],
                                      enc_layer=c["enc"], dec_layer=c["dec"], use_proposals="proposals" in c)
    if c["kind"] == "image_obj":
        cases.attach_obj_stream(ref, ns.bert_seq2seq.BertDecoderModel, ref.decoder.config, c["seed"] + 500)
        ref.encode = model.encode                 # the same fixed (grid, object) memories on both sides
    ref = ref.eval()
    missing = ref.load_state_dict(model.state_dict())
    assert not missing.missing_keys and not missing.unexpected_keys
    kw = dict(kw)
    beam = kw.pop("beam_width")
    bd = ns.seq_gen.BeamDecoder(ref, beam_width=beam, **cases.GEN)
    with torch.no_grad():
        outs = bd(**kw)
        full = bd(unpad_output=False, **kw)
    extra = {}
    if record:
        ns.seq_gen_rec.TRACE.clear()
        with torch.no_grad():
            again = ns.seq_gen_rec.BeamDecoder(ref, beam_width=beam, **cases.GEN)(**kw)
        assert [a.tolist() for a in again] == [o.tolist() for o in outs]
        extra = pack_trace(ns.seq_gen_rec.TRACE, len(outs) * beam, c["V"])
        ns.seq_gen_rec.TRACE.clear()
    return model, outs, full, extra


def main():
    os.makedirs(OUT, exist_ok=True)
    import sys
    only = sys.argv[1:]
    for name in cases.CASES:
        if only and name not in only:
            continue
        model, outs, full, extra = run_reference(name, record=cases.CASES[name].get("record", False))
        L = max(len(o) for o in outs)
        toks = np.zeros((len(outs), L), dtype=np.int64)
        lens = np.array([len(o) for o in outs], dtype=np.int64)
        for r, o in enumerate(outs):
            toks[r, :len(o)] = o.numpy()
        np.savez_compressed(os.path.join(OUT, name + ".npz"), tokens=toks, lengths=lens,
                            full=torch.stack(full).numpy(), **extra,
                            meta=json.dumps(dict(case=name, recipe=cases.CASES[name], gen=cases.GEN,
                                                 weights_sha256=cases.weights_digest(model),
                                                 torch=torch.__version__,
                                                 source="/root/reference/src/seq_gen.py BeamDecoder (line 216 '//' patch)")))
        print(name, "lengths", lens.tolist(), "steps", full[0].numel() - 1)


if __name__ == "__main__":
    main()
