"""Token-budget batching for decoding, same rule and same batch dictionary as the reference's
`MTDataset.batch_examples` (reference src/dataset.py:99-165): examples are taken in input order and a batch
is flushed (without its last example) as soon as  (max_src + max_dst) * n > max_batch  or
(max_src^2 + max_dst^2) * n * max_dst > max_batch_capacity * 1e6.  `pad_idx` holds, per row, the index of the
first pad token, or S-1 when the row has none (:157-163) -- BeamDecoder uses it as `src_sizes`.
"""
from typing import List, Sequence, Tuple

import torch


def _pack(rows: List[torch.Tensor], pad: int) -> torch.Tensor:
    width = max(int(r.numel()) for r in rows)
    out = torch.full((len(rows), width), pad, dtype=torch.long)
    for i, r in enumerate(rows):
        out[i, :r.numel()] = r
    return out


def _entry(src, dst, sl, dl, pad):
    s, d = _pack(src, pad), _pack(dst, pad)
    is_pad = s == pad
    first_pad = torch.where(is_pad.any(1), is_pad.float().argmax(1), torch.full((s.size(0),), s.size(1) - 1))
    return {"src_texts": s, "src_pad_mask": s != pad, "dst_texts": d, "dst_pad_mask": d != pad,
            "src_langs": torch.tensor(sl, dtype=torch.long), "dst_langs": torch.tensor(dl, dtype=torch.long),
            "proposal": torch.tensor([pad], dtype=torch.long), "pad_idx": first_pad.long()}


class MTDataset(torch.utils.data.Dataset):
    def __init__(self, max_batch_capacity: int, max_batch: int, pad_idx: int, max_seq_len: int = 175,
                 examples: Sequence[Tuple[Sequence[int], Sequence[int], int, int]] = None, ngpu: int = 1):
        self.batches = []
        src, dst, sl, dl = [], [], [], []
        ms = md = 0
        for ex in examples:
            s = torch.as_tensor(ex[0][:max_seq_len], dtype=torch.long)
            d = torch.as_tensor(ex[1][:max_seq_len], dtype=torch.long)
            src.append(s); dst.append(d); sl.append(ex[2]); dl.append(ex[3])
            ms, md = max(ms, int(s.numel())), max(md, int(d.numel()))
            n = len(src)
            over = (ms + md) * n > max_batch or (ms * ms + md * md) * n * md > max_batch_capacity * 1000000
            if over and n - 1 >= ngpu and n > 1:
                self.batches.append(_entry(src[:-1], dst[:-1], sl[:-1], dl[:-1], pad_idx))
                src, dst, sl, dl = src[-1:], dst[-1:], sl[-1:], dl[-1:]
                ms, md = int(src[0].numel()), int(dst[0].numel())
        if len(src) > 0 and len(src) >= ngpu:
            self.batches.append(_entry(src, dst, sl, dl, pad_idx))

    def __len__(self):
        return len(self.batches)

    def __getitem__(self, item):
        return self.batches[item]


def synthetic_sentences(n, vocab, min_len=16, max_len=128, seed=1234, lang_token=5, eos=4, sort=True):
    """BASELINE.json configs[1]: lengths ~U{min..max} incl. the leading language token and trailing </s>,
    body tokens ~U{7..V-1}; length-sorted like the reference's training binariser (src/create_mt_batches.py:30-40)."""
    g = torch.Generator().manual_seed(seed)
    lens = torch.randint(min_len, max_len + 1, (n,), generator=g)
    if sort:
        lens, _ = torch.sort(lens)
    out = []
    for L in lens.tolist():
        body = torch.randint(7, vocab, (L,), generator=g)
        body[0], body[-1] = lang_token, eos
        out.append(body)
    return out
