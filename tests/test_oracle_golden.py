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
    outs = cases.run_oracle(model, kw)
    assert [len(o) for o in outs] == lengths.tolist()
    for r, o in enumerate(outs):
        assert o.tolist() == tokens[r, :lengths[r]].tolist()
    outs_full = cases.run_oracle(model, kw, unpad_output=False)
    assert torch.stack(outs_full).numpy().tolist() == full.tolist()


def test_fixture_cases_cover_the_edge_cases():
    """ragged lengths, early EOS next to running rows, beam 1, wide beam, image memory."""
    _, l_eos, full_eos, _ = load_golden("text_eos")
    assert l_eos.min() <= 2 and l_eos.max() >= 10
    assert cases.CASES["text_b1_langdec"]["beam"] == 1 and cases.CASES["text_b8_tie"]["beam"] == 8
    assert cases.CASES["image_b3"]["kind"] == "image"


def test_max_lens_follow_the_float32_rule_of_the_reference():
    """src/seq_gen.py:113-121 maps max_len_func over 0-dim LongTensors: a * s is float32.  The real reference's output
    lengths (no EOS: length = max_lens[r]) for src_sizes 179 / 90 / 170 at a = 1.3, b = 5."""
    _, lengths, full, _ = load_golden("text_maxlens_f32")
    assert lengths.tolist() == [237, 121, 225]            # double arithmetic would give 237 / 122 / 226
    assert full.shape[1] == 239


def test_restated_per_step_log_probs_match_the_recorded_reference_run():
    """real size (6+6 layers, d 768, V 60000): per-step log-softmax rows, beam scores and prefixes recorded inside the
    REAL reference's loop (oracle/reference_loader.py recorder) against the restatement's own trace."""
    z = np.load(os.path.join(GOLD, "real_6x6.npz"))
    model, kw, c = cases.build("real_6x6")
    trace = []
    R.beam_search(model, trace=trace, **cases.GEN, **kw)
    assert len(trace) == z["step_rows"].shape[0] == c["max_len"] - 1
    sub = torch.from_numpy(z["sub_index"])
    for k, t in enumerate(trace):
        n = int(z["step_rows"][k])
        lp = t["log_probs"]
        assert lp.shape[0] == n
        # the recorder stores rows AFTER finished / over-length rows were zeroed; this case has neither
        assert float((lp[:, sub] - torch.from_numpy(z["step_logp_sub"][k, :n])).abs().max()) < 2e-5
        tv, ti = lp.topk(8, -1)
        assert float((tv - torch.from_numpy(z["step_top_val"][k, :n])).abs().max()) < 2e-5
        assert ti.tolist() == z["step_top_idx"][k, :n].tolist()
        pre = t["prefixes"].reshape(-1, t["prefixes"].shape[-1])
        assert pre.tolist() == z["step_prefixes"][k, :, :pre.shape[1]].tolist()
        assert float((t["scores"].reshape(-1) - torch.from_numpy(z["step_scores"][k])).abs().max()) < 2e-5
