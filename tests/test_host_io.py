"""Host side of the decode path against the reference's own classes (CPU; the comparisons with the reference run only in
the build container, where /root/reference exists): checkpoint format and state-dict compatibility of the product model
containers (src/seq2seq.py:183-213, src/image_model.py:267-306), the image batcher (src/dataset.py:426-476), the marshal
batch format (src/create_mt_batches.py:8-40, src/dataset.py:83-97) and the text processor surface (src/textprocessor.py)."""
import marshal
import os

import numpy as np
import pytest
import torch

from imagetranslate_b200 import dataset as D
from imagetranslate_b200.models import ImageCaptioning, MassSeq2Seq, Seq2Seq, SyntheticTextProcessor
from imagetranslate_b200.textprocessor import TextProcessor
from oracle import reference_loader, restated as R

needs_ref = pytest.mark.skipif(not reference_loader.available(), reason="/root/reference not present")
SMALL = dict(enc_layer=2, dec_layer=2, embed_dim=96, intermediate_dim=160)


def _randomize(model, seed):
    g = torch.Generator().manual_seed(seed)
    seen = set()
    with torch.no_grad():
        for p in model.parameters():
            if p.data_ptr() not in seen:
                seen.add(p.data_ptr())
                p.copy_(torch.randn(p.shape, generator=g))


@needs_ref
@pytest.mark.parametrize("lang_dec,tie,prop", [(False, False, False), (False, True, False), (True, False, False), (True, True, True),
                                               (False, False, True)])
def test_reference_written_checkpoint_loads_into_the_product_container(tmp_path, lang_dec, tie, prop):
    """A checkpoint written by the REAL reference's Seq2Seq.save, with every distinct parameter holding distinct random
    values (so aliasing mistakes show), loaded by the product's Seq2Seq.load: same keys, same tensors -- including the
    per-language decoders' own token_type tables (src/seq2seq.py:73-77 ties the ENCODER to the last decoder only)."""
    ns = reference_loader.load()
    tp = R.FakeTextProcessor(150, languages=("<en>", "<fa>", "<ar>"))
    ref = ns.seq2seq.Seq2Seq(tp, lang_dec=lang_dec, tie_embed=tie, use_proposals=prop, **SMALL)
    _randomize(ref, 5)
    ref.save(str(tmp_path))
    want = ref.state_dict()
    got_model = Seq2Seq.load(Seq2Seq, str(tmp_path), tok_dir=tp, map_location="cpu")
    got = got_model.state_dict()
    assert sorted(got.keys()) == sorted(want.keys())
    for k in want:
        assert torch.equal(got[k], want[k]), k
    if lang_dec:
        t = [d.embeddings.token_type_embeddings.weight for d in got_model.decoder]
        assert not torch.equal(t[0], t[1]) and not torch.equal(t[1], t[2])
        assert got_model.encoder.embeddings.token_type_embeddings.weight is t[-1]
    # and back: what the product saves, the reference loads
    out2 = tmp_path / "again"
    got_model.save(str(out2))
    sd = torch.load(os.path.join(str(out2), "mt_model.state_dict"))
    res = ref.load_state_dict(sd)
    assert not res.missing_keys and not res.unexpected_keys
    assert open(os.path.join(str(out2), "mt_config"), "rb").read() == open(os.path.join(str(tmp_path), "mt_config"), "rb").read()


@needs_ref
@pytest.mark.parametrize("depth,use_obj,lang_dec", [(1, False, False), (3, True, False), (2, True, True)])
def test_image_captioning_container_matches_the_reference_state_dict(depth, use_obj, lang_dec):
    ns = reference_loader.load()
    tp = R.FakeTextProcessor(120)
    kw = dict(resnet_depth=depth, enc_layer=1, dec_layer=1, embed_dim=96, intermediate_dim=128, use_obj=use_obj, lang_dec=lang_dec)
    ref = ns.image_model.ImageCaptioning(tp, **kw)
    mine = ImageCaptioning(tp, **kw)
    a, b = ref.state_dict(), mine.state_dict()
    assert sorted(a.keys()) == sorted(b.keys())
    assert all(a[k].shape == b[k].shape for k in a)


@needs_ref
def test_image_front_end_equals_the_reference_grid_features():
    """ImageCaptioning.encode(images=) on the same weights: ResNet grid -> fc + location_embedding (src/image_model.py:24-41),
    with the reference's always-on inference dropout switched off on both sides (dropout = 0)."""
    ns = reference_loader.load()
    tp = R.FakeTextProcessor(120)
    kw = dict(resnet_depth=1, enc_layer=1, dec_layer=1, embed_dim=768, intermediate_dim=128, use_obj=False)
    torch.manual_seed(3)
    ref = ns.image_model.ImageCaptioning(tp, **kw).eval()
    mine = ImageCaptioning(tp, **kw).eval()
    res = mine.load_state_dict(ref.state_dict())
    assert not res.missing_keys and not res.unexpected_keys
    ref.image_model.dropout = 0
    x = torch.randn(2, 3, 224, 224)
    with torch.no_grad():
        want, wobj = ref.encode(images=x)
        got, gobj = mine.encode(images=x)
    assert wobj is None and gobj is None and got.shape == (2, 49, 768)
    assert float((got - want).abs().max()) < 1e-4 * float(want.abs().max())
    # the reference's quirk is reproducible on request: dropout on the grid at inference
    mine.image_model.inference_dropout = True
    with torch.no_grad():
        noisy = mine.encode(images=x)[0]
    assert float((noisy - got).abs().max()) > 0


def _image_dir(tmp_path):
    from PIL import Image
    rng = np.random.RandomState(0)
    for i, (w, h) in enumerate([(300, 400), (640, 480), (224, 224), (500, 257), (100, 80), (257, 900)]):
        Image.fromarray(rng.randint(0, 255, (h, w, 3), dtype=np.uint8)).save(str(tmp_path / ("im%d.jpg" % i)))
    Image.fromarray(rng.randint(0, 255, (60, 70), dtype=np.uint8)).save(str(tmp_path / "gray.jpeg"))
    Image.fromarray(rng.randint(0, 255, (50, 50, 3), dtype=np.uint8)).save(str(tmp_path / "skipped.png"))
    (tmp_path / "broken.jpg").write_text("not an image")
    return str(tmp_path)


def test_image_dataset_batches(tmp_path):
    d = _image_dir(tmp_path)
    ds = D.ImageDataset(d, 3, target_lang=1, first_token=6)
    n_files = len([f for f in os.listdir(d) if not f.lower().endswith(".png")])
    assert sum(len(b) for b in ds.image_batches) == n_files == 8 and len(ds) == 3
    item = ds[0]
    assert item["images"].shape == (3, 3, 224, 224) and item["images"].dtype == torch.float32
    assert item["tgt_langs"].tolist() == [1, 1, 1] and item["first_tokens"].tolist() == [6, 6, 6]
    broken = [i for b in range(len(ds)) for i, p in zip(ds[b]["images"], ds[b]["paths"]) if p.endswith("broken.jpg")][0]
    black = (torch.zeros(3, 1, 1) - torch.tensor(D.ImageDataset.MEAN).view(3, 1, 1)) / torch.tensor(D.ImageDataset.STD).view(3, 1, 1)
    assert torch.allclose(broken, black.expand(3, 224, 224))


@needs_ref
def test_image_dataset_equals_the_reference(tmp_path):
    d = _image_dir(tmp_path)
    ns = reference_loader.load()
    ref = ns.dataset.ImageDataset(d, 3, target_lang=1, first_token=6)
    mine = D.ImageDataset(d, 3, target_lang=1, first_token=6)
    assert len(ref) == len(mine)
    for i in range(len(ref)):
        a, b = ref[i], mine[i]
        assert a["paths"] == b["paths"]
        assert torch.equal(a["images"], b["images"])                 # same resize / crop / normalisation, bit for bit
        assert torch.equal(a["tgt_langs"], b["tgt_langs"]) and torch.equal(a["first_tokens"], b["first_tokens"])


def test_marshal_batch_file_round_trip(tmp_path):
    sents = D.synthetic_sentences(40, 500, 4, 30, seed=3, sort=False)
    ex = [(s.tolist(), [6] + s.tolist()[1:(k % 7) + 2], 0, 1) for k, s in enumerate(sents)]
    path = str(tmp_path / "mt.batch")
    D.write_mt_examples(path, ex)
    stored = marshal.load(open(path, "rb"))
    assert [len(e[1]) for e in stored] == sorted(len(e[1]) for e in ex)            # sorted by target length (:30-36)
    a = D.MTDataset(max_batch_capacity=600, max_batch=300, pad_idx=0, max_seq_len=175, batch_pickle_dir=path)
    b = D.MTDataset(max_batch_capacity=600, max_batch=300, pad_idx=0, max_seq_len=175, examples=stored)
    assert len(a) == len(b) > 1
    for x, y in zip(a.batches, b.batches):
        assert all(torch.equal(x[k], y[k]) for k in x)


@needs_ref
def test_marshal_batch_file_read_by_the_reference_dataset(tmp_path):
    ns = reference_loader.load()
    sents = D.synthetic_sentences(30, 500, 4, 30, seed=4, sort=False)
    ex = [(s.tolist(), [6] + s.tolist()[1:(k % 5) + 2], 0, 1) for k, s in enumerate(sents)]
    path = str(tmp_path / "mt.batch")
    D.write_mt_examples(path, ex)
    ref = ns.dataset.MTDataset(batch_pickle_dir=path, max_batch_capacity=600, max_batch=300, pad_idx=0, max_seq_len=175,
                               keep_pad_idx=True)
    mine = D.MTDataset(max_batch_capacity=600, max_batch=300, pad_idx=0, max_seq_len=175, batch_pickle_dir=path)
    assert len(ref.batches) == len(mine.batches)
    for x, y in zip(ref.batches, mine.batches):
        for k in ("src_texts", "src_pad_mask", "dst_texts", "dst_pad_mask", "src_langs", "dst_langs", "pad_idx"):
            assert torch.equal(x[k], y[k]), k


def _train_tokenizer(tmp_path, ns=None):
    import tokenizers
    text = tmp_path / "train.txt"
    words = ["alpha", "beta", "gamma", "delta", "shab", "rooz", "ketab", "khane", "red", "green", "blue", "river"]
    rng = np.random.RandomState(1)
    text.write_text("\n".join(" ".join(rng.choice(words, size=rng.randint(3, 9))) for _ in range(400)))
    langs = {"<en>": 0, "<fa>": 1}
    tok = tokenizers.SentencePieceBPETokenizer()
    tok.train(files=[str(text)], vocab_size=120, min_frequency=2,
              special_tokens=["<pad>", "<s>", "<unk>", "<mask>", "</s>"] + list(langs))
    tp = TextProcessor()
    tp.tokenizer = tok
    tp.init_properties(langs)
    out = str(tmp_path / "tok")
    tp.save(out)
    return out


def test_text_processor_surface(tmp_path):
    tok_dir = _train_tokenizer(tmp_path)
    tp = TextProcessor(tok_dir)
    assert (tp.pad_token_id(), tp.bos_token_id(), tp.unk_token_id(), tp.mask_token_id(), tp.sep_token_id()) == (0, 1, 2, 3, 4)
    assert tp.token_id("<en>") == 5 and tp.token_id("<fa>") == 6 and tp.languages == {"<en>": 0, "<fa>": 1}
    ids = tp.tokenize_one_sentence("<fa> shab rooz ketab </s>")
    assert ids[0] == 6 and ids[-1] == 4 and len(ids) > 2
    assert tp.tokenizer.decode(np.array(ids[1:-1])).replace(" ", "") == "shabroozketab"
    assert tp.token_id("<nope>") == 0 and tp.is_lang(5) and not tp.is_lang(9) and tp.vocab_size() == tp.tokenizer.get_vocab_size()


@needs_ref
def test_text_processor_equals_the_reference(tmp_path):
    ns = reference_loader.load()
    tok_dir = _train_tokenizer(tmp_path)
    a, b = ns.TextProcessor(tok_dir), TextProcessor(tok_dir)
    for line in ("<fa> shab rooz ketab khane </s>", "<en> alpha beta river red </s>", "<en> unseenword beta </s>"):
        assert a.tokenize_one_sentence(line) == b.tokenize_one_sentence(line)
    assert a.tokenize_one_sentence_with_langid("green blue", 5) == b.tokenize_one_sentence_with_langid("green blue", 5)
    for f in ("pad_token_id", "bos_token_id", "sep_token_id", "mask_token_id", "unk_token_id", "vocab_size"):
        assert getattr(a, f)() == getattr(b, f)()


def test_product_seq2seq_and_mass_share_the_decode_surface():
    tp = SyntheticTextProcessor(100)
    m = MassSeq2Seq(tp, lang_dec=False, **SMALL)
    assert isinstance(m, Seq2Seq) and m.decoder.decoder.layer[0].attention is m.encoder.encoder.layer[0].attention


@needs_ref
def test_get_outputs_until_eos_equals_the_reference():
    """src/seq_gen.py:6-24: cut before the first EOS, else at size_limit[r] (or the row length); remove_first_token."""
    from imagetranslate_b200.seq_gen import get_outputs_until_eos
    ns = reference_loader.load()
    g = torch.Generator().manual_seed(5)
    out = torch.randint(5, 12, (9, 14), generator=g)
    out[2, 3] = 4
    out[2, 7] = 4
    out[5, 0] = 4
    out[7, 13] = 4
    lim = torch.randint(1, 15, (9,), generator=g)
    for kw in (dict(), dict(size_limit=lim), dict(size_limit=lim, remove_first_token=True), dict(remove_first_token=True)):
        a = ns.seq_gen.get_outputs_until_eos(4, out, **kw)
        b = get_outputs_until_eos(4, out, **kw)
        assert [x.tolist() for x in a] == [x.tolist() for x in b], kw
    a = ns.seq_gen.get_outputs_until_eos(4, out[2])
    b = get_outputs_until_eos(4, out[2])
    assert [x.tolist() for x in a] == [x.tolist() for x in b]
