"""The translate.py / caption.py-compatible drivers end to end (-m gpu).

  * BASELINE.json configs[0]: tokenizer trained on src/sample fa-en, Seq2Seq 6+6 d=768, beam 4, --batch 4000 --capacity 600:
    tests/golden/config1 holds what the REAL reference's build_model / build_data_loader / translate_batch produced in the
    build container (oracle/make_config1.py); here the same tokenizer directory, the same 24 source lines and a checkpoint
    in the reference's on-disk format go through THIS package's helpers of the same names.
  * caption driver: image directory -> ImageDataset -> ImageCaptioning.encode(images=) (ResNet grid + fused fc/location
    GEMM) -> beam search, against the oracle decoding the same memory.
"""
import json
import os
import types

import numpy as np
import pytest
import torch

from imagetranslate_b200 import caption as cap_driver
from imagetranslate_b200 import translate as tr_driver
from imagetranslate_b200.models import ImageCaptioning, Seq2Seq
from imagetranslate_b200.textprocessor import TextProcessor
from oracle import cases, restated as R

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "config1")


def _options(**kw):
    base = dict(input_path=None, src_lang="fa", target_lang="en", output_path=None, batch=4000, tokenizer_path=GOLD, cache_size=300,
                model_path=None, verbose=False, beam_width=4, max_len_a=1.3, max_len_b=5, len_penalty_ratio=0.8, total_capacity=600,
                fp16=False, rows=4096, lanes=3, window=256)
    base.update(kw)
    return types.SimpleNamespace(**base)


@pytest.fixture(scope="module")
def config1_checkpoint(tmp_path_factory):
    """The seeded 6+6 model of the fixture, written in the reference's checkpoint format (mt_config + mt_model.state_dict)."""
    z = np.load(os.path.join(GOLD, "config1.npz"))
    meta = json.loads(str(z["meta"]))
    tp = TextProcessor(GOLD)
    torch.manual_seed(meta["seed"])
    oracle_model = R.Seq2SeqModel(tp, lang_dec=False, tie_embed=False, enc_layer=6, dec_layer=6).eval()
    assert cases.weights_digest(oracle_model) == meta["weights_sha256"]
    product = Seq2Seq(tp, lang_dec=False, tie_embed=False, enc_layer=6, dec_layer=6)
    res = product.load_state_dict(oracle_model.state_dict())
    assert not res.missing_keys and not res.unexpected_keys
    out = str(tmp_path_factory.mktemp("config1_model"))
    product.save(out)
    return out, z, meta


def test_config1_translate_flow_matches_the_reference_run(config1_checkpoint, tmp_path):
    model_dir, z, meta = config1_checkpoint
    opts = _options(input_path=os.path.join(GOLD, "input.txt"), model_path=model_dir, output_path=str(tmp_path / "out.txt"))
    generator, text_processor = tr_driver.build_model(opts)
    loader = tr_driver.build_data_loader(opts, text_processor)
    assert len(loader) == int(z["n_batches"])
    strings, ids = [], []
    with torch.no_grad():
        for i, batch in enumerate(loader):
            assert torch.equal(batch["src_texts"].squeeze(0), torch.from_numpy(z["batch%d_src_texts" % i]))      # same tokenisation
            assert torch.equal(batch["pad_idx"].squeeze(0), torch.from_numpy(z["batch%d_pad_idx" % i]))
            out, _ = tr_driver.translate_batch(batch, generator, text_processor)
            strings += out
            ids += generator(**tr_driver.batch_kwargs(batch, text_processor))
    want = [z["out_tokens"][r, :z["out_lengths"][r]].tolist() for r in range(len(ids))]
    assert [o.tolist() for o in ids] == want                          # fp32: token-identical to the reference run
    assert strings == meta["strings"]
    # the whole driver (windowed forward_many, writer): same file as the reference's loop would write
    n = tr_driver.run(opts)
    assert n == len(want)
    # (decoded strings may hold newlines of their own -- the sample tokenizer learns a newline merge -- so compare the bytes)
    assert open(opts.output_path).read() == "\n".join(meta["strings"]) + "\n"
    # --fp16: the production kernels on the same checkpoint
    opts16 = _options(input_path=opts.input_path, model_path=model_dir, output_path=str(tmp_path / "out16.txt"), fp16=True)
    tr_driver.run(opts16)
    gen16, _ = tr_driver.build_model(opts16)
    with torch.no_grad():
        ids16 = [o for batch in loader for o in gen16(**tr_driver.batch_kwargs(batch, text_processor))]
    same = sum(int(a.tolist() == b) for a, b in zip(ids16, want))
    print("config 1 --fp16: %d/%d sentences token-identical to the reference's fp32 run" % (same, len(want)))
    assert len(ids16) == len(want) and same >= 0.9 * len(want)
    assert open(opts16.output_path).read() == "\n".join(text_processor.tokenizer.decode(o[1:].numpy()) for o in ids16) + "\n"


def _tiny_tokenizer(tmp_path):
    import tokenizers
    text = tmp_path / "train.txt"
    words = ["a", "dog", "cat", "runs", "on", "the", "grass", "red", "ball", "two", "people", "street"]
    rng = np.random.RandomState(1)
    text.write_text("\n".join(" ".join(rng.choice(words, size=rng.randint(3, 9))) for _ in range(300)))
    tok = tokenizers.SentencePieceBPETokenizer()
    tok.train(files=[str(text)], vocab_size=90, min_frequency=2, special_tokens=["<pad>", "<s>", "<unk>", "<mask>", "</s>", "<en>", "<fa>"])
    tp = TextProcessor()
    tp.tokenizer = tok
    tp.init_properties({"<en>": 0, "<fa>": 1})
    out = str(tmp_path / "tok")
    tp.save(out)
    return out


def test_caption_driver_end_to_end(tmp_path):
    from PIL import Image
    tok_dir = _tiny_tokenizer(tmp_path)
    tp = TextProcessor(tok_dir)
    torch.manual_seed(5)
    model = ImageCaptioning(tp, resnet_depth=1, lang_dec=False, tie_embed=False, enc_layer=1, dec_layer=2, use_obj=False).eval()
    with torch.no_grad():
        model.image_model.fc.weight.mul_(0.05)                       # O(1) memory entries from a random-init trunk
        model.output_layer[0].layer.bias[4] += 1.5                   # captions that end
    model_dir = str(tmp_path / "model")
    model.save(model_dir)
    img_dir = tmp_path / "images"
    img_dir.mkdir()
    rng = np.random.RandomState(2)
    for i in range(7):
        Image.fromarray(rng.randint(0, 255, (240 + 10 * i, 320, 3), dtype=np.uint8)).save(str(img_dir / ("p%d.jpg" % i)))
    opts = types.SimpleNamespace(input_path=str(img_dir), target_lang="en", output_path=str(tmp_path / "cap.txt"), batch=3,
                                 tokenizer_path=tok_dir, model_path=model_dir, beam_width=4, max_len_a=1.3, max_len_b=5,
                                 len_penalty_ratio=0.8, fp16=False, obj=False, rows=4096, lanes=2, window=64)
    n = cap_driver.run(opts)
    assert n == 7
    written = open(opts.output_path).read()
    assert written.count("\t") == 7
    # batch by batch through caption_batch (the reference's call pattern) and through the oracle on the same memory
    generator, text_processor = cap_driver.build_model(opts)
    loader = cap_driver.build_data_loader(opts, text_processor)
    gpu_model = generator.seq2seq_model
    oracle = R.Seq2SeqModel(tp, lang_dec=False, tie_embed=False, enc_layer=1, dec_layer=2).eval()
    res = oracle.load_state_dict({k: v.cpu() for k, v in gpu_model.state_dict().items()}, strict=False)
    assert not res.missing_keys
    k, expect = 0, ""
    with torch.no_grad():
        for batch in loader:
            out, paths = cap_driver.caption_batch(batch, generator, text_processor)
            kw = cap_driver.batch_kwargs(batch, text_processor)
            memory = gpu_model.encode(images=kw["images"])[0]
            want = R.beam_search(oracle, first_tokens=kw["first_tokens"], tgt_langs=kw["tgt_langs"], pad_idx=kw["pad_idx"], max_len=256,
                                 image_embed=memory.float().cpu(), beam_width=4, max_len_a=1.3, max_len_b=5, len_penalty_ratio=0.8)
            want_s = [text_processor.tokenizer.decode(w[1:].numpy()) for w in want]
            assert out == want_s
            expect += "\n".join(p + "\t" + s for p, s in zip(cap_driver._paths(batch), out)) + "\n"     # src/caption.py:90-91
            k += len(out)
    assert k == 7 and written == expect


@pytest.mark.parametrize("dtype", ["fp32", "fp16"])
def test_fused_fc_location_gemm_matches_torch(dtype):
    """ImageCaptioning.encode(images=) on CUDA: fc + location_embedding as ONE GEMM of this library (location table as extra K
    columns) against the plain torch expression of src/image_model.py:39-41."""
    tp = R.FakeTextProcessor(100)
    torch.manual_seed(9)
    model = ImageCaptioning(tp, resnet_depth=3, lang_dec=False, enc_layer=1, dec_layer=1, use_obj=False).eval().to("cuda:0")
    if dtype == "fp16":
        model = model.half()
    im = model.image_model
    x = torch.randn(5, 3, 224, 224, device="cuda:0").to(im.fc.weight.dtype)
    with torch.no_grad():
        grid = im.grid(x)
        grid = grid / grid.float().std().to(grid.dtype)
        got = im._fc_location(grid)
        want = grid.float() @ im.fc.weight.float().t() + im.location_embedding.weight.float().unsqueeze(0)
    assert got.shape == (5, 49, 768)
    err = float((got.float() - want).abs().max()) / float(want.abs().max())
    assert err < (1e-4 if dtype == "fp32" else 4e-3), err
