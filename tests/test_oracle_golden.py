"""The travel-capable oracle (oracle/restated.py) must reproduce the fixtures that the REAL reference
BeamDecoder produced in the build container (tests/golden/*.npz, made by oracle/make_golden.py)."""
import json
import os

import numpy as np
import pytest
import torch

from oracle import cases, restated as R

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load_golden(name):
    z = np.load(os.path.join(GOLD, name + ".npz"))
    return z["tokens"], z["lengths"], z["full"], json.loads(str(z["meta"]))


@pytest.mark.parametrize("name", sorted(cases.CASES))
def test_restated_oracle_matches_reference_fixture(name):
    tokens, lengths, full, meta = load_golden(name)
    model, kw, c = cases.build(name)
    assert cases.weights_digest(model) == meta["weights_sha256"], "seeded weights differ from the fixture's"
    outs = R.beam_search(model, **cases.GEN, **kw)
    assert [len(o) for o in outs] == lengths.tolist()
    for r, o in enumerate(outs):
        assert o.tolist() == tokens[r, :lengths[r]].tolist()
    outs_full = R.beam_search(model, unpad_output=False, **cases.GEN, **kw)
    assert torch.stack(outs_full).numpy().tolist() == full.tolist()


def test_fixture_cases_cover_the_edge_cases():
    """ragged lengths, early EOS next to running rows, beam 1, wide beam, image memory."""
    _, l_eos, full_eos, _ = load_golden("text_eos")
    assert l_eos.min() <= 2 and l_eos.max() >= 10
    assert cases.CASES["text_b1_langdec"]["beam"] == 1 and cases.CASES["text_b8_tie"]["beam"] == 8
    assert cases.CASES["image_b3"]["kind"] == "image"
