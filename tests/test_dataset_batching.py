"""Token-budget batching (reference src/dataset.py:99-165 MTDataset.batch_examples): the host-side mirror in
imagetranslate_b200/dataset.py against the reference's own class (build container) and against its stated rule."""
import pytest
import torch

from imagetranslate_b200.dataset import MTDataset, synthetic_sentences
from oracle import reference_loader


def _examples(n, seed, lo=3, hi=60, dst_len=1):
    g = torch.Generator().manual_seed(seed)
    ex = []
    for _ in range(n):
        L = int(torch.randint(lo, hi + 1, (1,), generator=g))
        src = torch.randint(7, 500, (L,), generator=g).tolist()
        dst = torch.randint(7, 500, (dst_len,), generator=g).tolist()
        ex.append((src, dst, int(torch.randint(0, 2, (1,), generator=g)), 1))
    return ex


@pytest.mark.skipif(not reference_loader.available(), reason="/root/reference not present")
@pytest.mark.parametrize("n,max_batch,capacity,dst_len,ngpu", [(200, 400, 600, 1, 1), (150, 4000, 600, 1, 1), (120, 300, 1, 7, 1),
                                                              (90, 500, 600, 2, 3), (1, 4000, 600, 1, 1)])
def test_batches_equal_the_reference_class(n, max_batch, capacity, dst_len, ngpu):
    ref_cls = reference_loader.load().dataset.MTDataset
    ex = _examples(n, 5 + n, dst_len=dst_len)
    want = ref_cls(max_batch_capacity=capacity, max_batch=max_batch, pad_idx=0, max_seq_len=40, examples=ex, ngpu=ngpu).batches
    got = MTDataset(max_batch_capacity=capacity, max_batch=max_batch, pad_idx=0, max_seq_len=40, examples=ex, ngpu=ngpu).batches
    assert len(got) == len(want)
    for g, w in zip(got, want):
        for k in ("src_texts", "src_pad_mask", "dst_texts", "dst_pad_mask", "src_langs", "dst_langs", "pad_idx"):
            assert torch.equal(g[k], w[k]), k


def test_rule_order_budget_and_pad_index():
    sents = synthetic_sentences(500, 60000, 16, 128, seed=9)
    assert all(int(a.numel()) <= int(b.numel()) for a, b in zip(sents, sents[1:]))            # length-sorted
    assert all(int(s[0]) == 5 and int(s[-1]) == 4 and 16 <= s.numel() <= 128 for s in sents)
    batches = MTDataset(examples=[(s, [6], 0, 1) for s in sents], max_batch_capacity=600, max_batch=4000, pad_idx=0,
                        max_seq_len=10000).batches
    flat = [row[row != 0] for b in batches for row in b["src_texts"]]
    assert len(flat) == 500 and all(torch.equal(a, s) for a, s in zip(flat, sents))           # input order kept
    for b in batches:
        n, S = b["src_texts"].shape
        assert (S + 1) * n <= 4000                                                            # the budget that was not exceeded
        for r in range(n):
            row = b["src_texts"][r]
            pads = (row == 0).nonzero()
            assert int(b["pad_idx"][r]) == (int(pads[0]) if pads.numel() else S - 1)          # :157-163


def test_empty_and_ragged_inputs():
    assert MTDataset(max_batch_capacity=600, max_batch=4000, pad_idx=0, examples=[]).batches == []
    b = MTDataset(max_batch_capacity=600, max_batch=4000, pad_idx=0, examples=[([5, 9, 4], [6], 0, 1), ([5, 4], [6], 1, 1)]).batches
    assert len(b) == 1 and b[0]["src_texts"].tolist() == [[5, 9, 4], [5, 4, 0]] and b[0]["pad_idx"].tolist() == [2, 2]
