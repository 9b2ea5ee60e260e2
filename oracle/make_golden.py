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


def run_reference(name):
    ns = reference_loader.load()
    model, kw, c = cases.build(name)
    ref = ns.mass_seq2seq.MassSeq2Seq(model.text_processor, lang_dec=c["lang_dec"], tie_embed=c["tie"],
                                      enc_layer=c["enc"], dec_layer=c["dec"]).eval()
    missing = ref.load_state_dict(model.state_dict())
    assert not missing.missing_keys and not missing.unexpected_keys
    kw = dict(kw)
    beam = kw.pop("beam_width")
    bd = ns.seq_gen.BeamDecoder(ref, beam_width=beam, **cases.GEN)
    with torch.no_grad():
        outs = bd(**kw)
        full = bd(unpad_output=False, **kw)
    return model, outs, full


def main():
    os.makedirs(OUT, exist_ok=True)
    for name in cases.CASES:
        model, outs, full = run_reference(name)
        L = max(len(o) for o in outs)
        toks = np.zeros((len(outs), L), dtype=np.int64)
        lens = np.array([len(o) for o in outs], dtype=np.int64)
        for r, o in enumerate(outs):
            toks[r, :len(o)] = o.numpy()
        np.savez_compressed(os.path.join(OUT, name + ".npz"), tokens=toks, lengths=lens,
                            full=torch.stack(full).numpy(),
                            meta=json.dumps(dict(case=name, recipe=cases.CASES[name], gen=cases.GEN,
                                                 weights_sha256=cases.weights_digest(model),
                                                 torch=torch.__version__,
                                                 source="/root/reference/src/seq_gen.py BeamDecoder (line 216 '//' patch)")))
        print(name, "lengths", lens.tolist(), "steps", full[0].numel() - 1)


if __name__ == "__main__":
    main()
