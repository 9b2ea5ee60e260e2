"""C-ABI boundary checks that need no GPU: the library loads, exports every symbol include/itb.h declares,
the ctypes structs match, and the product never touches the oracle."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    hdr = open(os.path.join(ROOT, "include", "itb.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    return sorted(set(re.findall(r"\b(itb_[a-z0-9_]+)\s*\(", hdr)))


def test_library_exports_every_declared_symbol():
    from imagetranslate_b200 import _lib
    assert os.path.isfile(_lib.LIB_PATH), "build with `python -c 'import __graft_entry__ as g; g.build()'`"
    raw = ctypes.CDLL(_lib.LIB_PATH)
    names = _declared()
    assert len(names) >= 15
    for n in names:
        assert hasattr(raw, n), "libitb.so does not export %s" % n
    assert sorted(_lib.EXPORTS) == names, "ctypes binding and header disagree"


def test_struct_layouts_match_header():
    from imagetranslate_b200 import _lib
    l = _lib.lib()
    assert l.itb_sizeof_config() == ctypes.sizeof(_lib.Config)
    assert l.itb_sizeof_layer() == ctypes.sizeof(_lib.Layer)
    assert l.itb_abi_version() == _lib.ABI_VERSION
    assert [l.itb_topk_capacity(b) for b in (1, 4, 5, 8, 12)] == [4, 4, 8, 8, 16]


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "imagetranslate_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dp, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", txt, flags=re.M), f
                if f.endswith(".py"):   # kernels may CITE reference lines in comments; host code must never read them
                    assert "/root/reference" not in txt, f


def test_cpu_model_is_refused_loudly():
    import torch
    from imagetranslate_b200 import BeamDecoder
    from oracle import restated as R
    m = R.Seq2SeqModel(R.FakeTextProcessor(50), lang_dec=False, enc_layer=1, dec_layer=1, embed_dim=64,
                       intermediate_dim=64).eval()
    bd = BeamDecoder(m, beam_width=2)
    with pytest.raises(RuntimeError):
        bd(image_embed=torch.zeros(1, 3, 64), first_tokens=torch.tensor([6]), tgt_langs=torch.tensor([1]), pad_idx=0,
           max_len=4)
