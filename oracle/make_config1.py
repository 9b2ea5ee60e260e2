"""ORACLE -- TEST INFRASTRUCTURE ONLY.  BASELINE.json configs[0] ("Seq2Seq 6+6 layer d=768 random-init, tokenizer trained on
src/sample fa-en, beam-4 translate of sample sentences on CPU") run through the REAL reference's own driver helpers in the
build container (SURVEY.md §8(d) "Config 1", Appendix C.8):

  1. tokenizer trained by the reference's TextProcessor.train_tokenizer (src/textprocessor.py:34-38) on src/sample/fa.txt +
     en.txt, vocab 1000, languages {<en>: 0, <fa>: 1} as README.md:131-147 prescribes for the sample;
  2. Seq2Seq(lang_dec=False, tie_embed=False, enc 6, dec 6, d 768) with the seeded weights of oracle/cases.py's recipe, saved
     by the reference's Seq2Seq.save (src/seq2seq.py:183-196);
  3. src/translate.py: build_model (Seq2Seq.load + BeamDecoder, :85-98), build_data_loader (:59-82) on the first N lines of
     src/sample/fa.txt, translate_batch (:34-56), --beam 4 --batch 4000 --capacity 600, fa -> en, fp32, CPU.

Writes tests/golden/config1/: tokenizer.json + langs (the tokenizer directory), input.txt (the N source lines), and
config1.npz (batched source ids, pad_idx, output token ids per sentence, output strings, weights digest).  The checkpoint
itself (167 M parameters) is not committed: tests rebuild it from the seed and check the digest.

Run here (build container):  python -m oracle.make_config1 [N=24]
"""
import json
import os
import shutil
import sys
import tempfile
import types

import numpy as np
import torch

from . import cases, reference_loader, restated as R

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "config1")
SEED, VOCAB = 0, 1000


def seeded_model(text_processor):
    torch.manual_seed(SEED)
    return R.Seq2SeqModel(text_processor, lang_dec=False, tie_embed=False, enc_layer=6, dec_layer=6).eval()


def main():
    n_lines = int(sys.argv[1]) if len(sys.argv) > 1 else 24
    ns = reference_loader.load()
    os.makedirs(OUT, exist_ok=True)
    work = tempfile.mkdtemp()
    sample = os.path.join(reference_loader.REF_SRC, "sample")
    # 1. tokenizer (README.md:131-147: raw text of both languages, language ids as special tokens)
    raw = os.path.join(work, "all.txt")
    with open(raw, "w") as wf:
        for name in ("fa.txt", "en.txt"):
            wf.write(open(os.path.join(sample, name)).read())
    tp = ns.TextProcessor()
    tp.train_tokenizer(paths=[raw], vocab_size=VOCAB, to_save_dir=OUT, languages={"<en>": 0, "<fa>": 1})
    tp = ns.TextProcessor(OUT)
    # 2. checkpoint written by the reference
    model = seeded_model(tp)
    ref = ns.seq2seq.Seq2Seq(tp, lang_dec=False, tie_embed=False, enc_layer=6, dec_layer=6)
    res = ref.load_state_dict(model.state_dict())
    assert not res.missing_keys and not res.unexpected_keys
    model_dir = os.path.join(work, "model")
    ref.save(model_dir)
    # 3. the reference's translate.py flow
    lines = [l for l in open(os.path.join(sample, "fa.txt")).read().split("\n") if l.strip()][:n_lines]
    inp = os.path.join(OUT, "input.txt")
    open(inp, "w").write("\n".join(lines) + "\n")
    options = types.SimpleNamespace(input_path=inp, src_lang="fa", target_lang="en", output_path=None, batch=4000,
                                    tokenizer_path=OUT, cache_size=300, model_path=model_dir, verbose=False, beam_width=4,
                                    max_len_a=1.3, max_len_b=5, len_penalty_ratio=0.8, total_capacity=600, fp16=False)
    generator, text_processor = ns.translate.build_model(options)
    loader = ns.translate.build_data_loader(options, text_processor)
    captured = []
    inner = generator.forward

    def recording_forward(*a, **kw):
        outs = inner(*a, **kw)
        captured.append([o.clone() for o in outs])
        return outs

    generator.forward = recording_forward
    strings, batches = [], []
    with torch.no_grad():
        for batch in loader:
            mt_output, _ = ns.translate.translate_batch(batch, generator, text_processor, False)
            strings += mt_output
            batches.append({k: v.squeeze(0).numpy() for k, v in batch.items() if k in ("src_texts", "pad_idx", "dst_texts")})
    outs = [o for b in captured for o in b]
    width = max(len(o) for o in outs)
    tok = np.zeros((len(outs), width), dtype=np.int64)
    for r, o in enumerate(outs):
        tok[r, :len(o)] = o.numpy()
    arrays = {}
    for i, b in enumerate(batches):
        for k, v in b.items():
            arrays["batch%d_%s" % (i, k)] = v
    np.savez_compressed(os.path.join(OUT, "config1.npz"), out_tokens=tok, out_lengths=np.array([len(o) for o in outs]),
                        n_batches=len(batches), **arrays,
                        meta=json.dumps(dict(strings=strings, weights_sha256=cases.weights_digest(model), seed=SEED, vocab=VOCAB,
                                             n_lines=n_lines, torch=torch.__version__,
                                             source="reference src/translate.py build_model / build_data_loader / translate_batch")))
    shutil.rmtree(work)
    print("config 1:", len(outs), "sentences in", len(batches), "batches; source widths", [b["src_texts"].shape for b in batches])
    print("first outputs:", strings[:2])


if __name__ == "__main__":
    main()
