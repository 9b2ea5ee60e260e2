"""ORACLE -- TEST INFRASTRUCTURE ONLY.  Imports the REAL reference modules from
/root/reference/src (present only in the build container, never on the GPU box).

Fixes applied, all documented in SURVEY.md §8(c) / Appendix D:
  1. `transformers` -> oracle/shim/transformers (2.9.0 restatement) while importing.
  2. src/seq_gen.py:216 `indices / V` -> `indices // V` (true division on torch>=1.5 breaks gather).
  3. `apex` stub (src/translate.py:6), `faster_rcnn_feats` stub (src/image_model.py:9).
Nothing from the reference is copied into the repo; files are exec'd where they lie.
"""
import importlib
import os
import sys
import types

REF_SRC = "/root/reference/src"
SHIM = os.path.join(os.path.dirname(os.path.abspath(__file__)), "shim")
_cache = {}


def available():
    return os.path.isfile(os.path.join(REF_SRC, "seq_gen.py"))


def _patch_text_processor(ns):
    """tokenizers 0.22 cannot `save(directory)` (src/textprocessor.py:43) nor re-read a merges.txt holding a newline merge:
    the oracle's TextProcessor serialises tokenizer.json instead (SURVEY §8c blocker 4, Appendix D.5).  Tokenisation is
    outside the hot path; both sides of a parity run consume the same ids."""
    import os as _os
    import pickle
    import tokenizers
    base = ns.textprocessor.TextProcessor

    class TextProcessor(base):
        def __init__(self, tok_model_path=None):
            self.languages = {}
            if tok_model_path is not None:
                self.tokenizer = tokenizers.Tokenizer.from_file(_os.path.join(tok_model_path, "tokenizer.json"))
                with open(_os.path.join(tok_model_path, "langs"), "rb") as fp:
                    self.languages = pickle.load(fp)
            self.init_properties(self.languages)

        def save(self, directory):
            _os.makedirs(directory, exist_ok=True)
            self.tokenizer.save(_os.path.join(directory, "tokenizer.json"))
            with open(_os.path.join(directory, "langs"), "wb") as fp:
                pickle.dump(self.languages, fp)

    ns.textprocessor.TextProcessor = TextProcessor
    ns.seq2seq.TextProcessor = TextProcessor
    ns.TextProcessor = TextProcessor


def load():
    """Returns a namespace with the reference modules: seq_gen, seq2seq, mass_seq2seq, bert_seq2seq."""
    if "ns" in _cache:
        return _cache["ns"]
    if not available():
        raise RuntimeError("reference sources not present (expected only in the build container)")
    saved = {k: v for k, v in sys.modules.items() if k == "transformers" or k.startswith("transformers.")}
    for k in saved:
        del sys.modules[k]
    sys.path.insert(0, REF_SRC)
    sys.path.insert(0, SHIM)
    try:
        apex = types.ModuleType("apex")
        apex.amp = types.ModuleType("apex.amp")
        apex.amp.initialize = lambda m, **kw: m
        sys.modules["apex"] = apex
        sys.modules["apex.amp"] = apex.amp
        import torch
        frc = types.ModuleType("faster_rcnn_feats")
        frc.fasterrcnn_resnet50_fpn = lambda *a, **k: torch.nn.Identity()       # callable returning a dummy module (D.3)
        sys.modules["faster_rcnn_feats"] = frc

        src = open(os.path.join(REF_SRC, "seq_gen.py")).read()
        needle = "beam_indices = indices / output.size(-1)"
        assert src.count(needle) == 1
        src = src.replace(needle, "beam_indices = indices // output.size(-1)")
        seq_gen = types.ModuleType("seq_gen")
        seq_gen.__file__ = os.path.join(REF_SRC, "seq_gen.py")
        exec(compile(src, seq_gen.__file__, "exec"), seq_gen.__dict__)
        sys.modules["seq_gen"] = seq_gen

        # Recorder: the same file with ONE extra statement after `top_beam_scores = top_scores`
        # (src/seq_gen.py:227) that appends clones of the step's tensors to the module-level list TRACE.  `output` is the
        # (M,V) log-softmax AFTER the finished / over-length rows were zeroed (:194-197).
        hook = "            top_beam_scores = top_scores\n"
        assert src.count(hook) == 1
        rec_src = src.replace(hook, hook + "            TRACE.append(dict(i=i, log_probs=output.clone(), prefixes=top_beam_outputs.clone(), "
                              "scores=top_scores.clone(), raw_idx=indices.clone(), "
                              "sizes=None if cur_size is None else cur_size.clone()))\n")
        seq_gen_rec = types.ModuleType("seq_gen_rec")
        seq_gen_rec.__file__ = seq_gen.__file__
        seq_gen_rec.TRACE = []
        exec(compile(rec_src, seq_gen.__file__, "exec"), seq_gen_rec.__dict__)

        ns = types.SimpleNamespace(seq_gen=seq_gen, seq_gen_rec=seq_gen_rec)
        for name in ("bert_seq2seq", "seq2seq", "mass_seq2seq", "dataset", "textprocessor"):
            setattr(ns, name, importlib.import_module(name))
        _patch_text_processor(ns)
        # image_model.init_net asks torchvision for pretrained weights (src/image_model.py:87-97): no network here, so the
        # torchvision constructors are wrapped to build the same architectures with random init
        import torchvision.models as tvm
        for fn in ("resnet18", "resnet34", "resnet50", "resnet101", "resnet152", "resnext101_32x8d"):
            orig = getattr(tvm, fn)
            if not getattr(orig, "_itb_wrapped", False):
                def wrapped(pretrained=False, _orig=orig, **kw):
                    return _orig(weights=None, **kw)
                wrapped._itb_wrapped = True
                setattr(tvm, fn, wrapped)
        for name in ("image_model", "translate", "caption"):
            setattr(ns, name, importlib.import_module(name))
    finally:
        sys.path.remove(SHIM)
        sys.path.remove(REF_SRC)
        for k in [k for k in sys.modules if k == "transformers" or k.startswith("transformers.")]:
            del sys.modules[k]
        sys.modules.update(saved)
    _cache["ns"] = ns
    return ns
