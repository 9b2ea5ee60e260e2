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
                 examples: Sequence[Tuple[Sequence[int], Sequence[int], int, int]] = None, ngpu: int = 1,
                 batch_pickle_dir: str = None):
        if examples is None:
            # reference src/dataset.py:83-97: the on-disk batch format is ONE marshal-ed list of
            # (src ids, dst ids, src language index, dst language index), length-sorted by the binariser
            # (src/create_mt_batches.py:8-40)
            import marshal
            with open(batch_pickle_dir, "rb") as fr:
                examples = marshal.load(fr)
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


def synthetic_sentences(n, vocab, min_len=16, max_len=128, seed=1234, lang_token=5, eos=4, sort=True, body_seed=None):
    """BASELINE.json configs[1]: lengths ~U{min..max} incl. the leading language token and trailing </s>,
    body tokens ~U{7..V-1}; length-sorted like the reference's training binariser (src/create_mt_batches.py:30-40).
    `body_seed` draws the token bodies from a second generator, so that several ranks can decode DIFFERENT sentences of
    the SAME length multiset (equal work per rank: what a weak-scaling run should compare)."""
    g = torch.Generator().manual_seed(seed)
    lens = torch.randint(min_len, max_len + 1, (n,), generator=g)
    if sort:
        lens, _ = torch.sort(lens)
    gb = g if body_seed is None else torch.Generator().manual_seed(body_seed)
    out = []
    for L in lens.tolist():
        body = torch.randint(7, vocab, (L,), generator=gb)
        body[0], body[-1] = lang_token, eos
        out.append(body)
    return out


def write_mt_examples(path: str, examples, sort_by_target: bool = True):
    """The reference binariser's output format (src/create_mt_batches.py:30-40): a marshal-ed list of
    (src ids, dst ids, src lang, dst lang) sorted by target length; MTDataset(batch_pickle_dir=path) reads it back."""
    import marshal
    ex = [(list(map(int, e[0])), list(map(int, e[1])), int(e[2]), int(e[3])) for e in examples]
    if sort_by_target:
        ex = [ex[i] for i in sorted(range(len(ex)), key=lambda i: len(ex[i][1]))]       # stable, like sorted(lens.items())
    with open(path, "wb") as fw:
        marshal.dump(ex, fw)


class ImageDataset(torch.utils.data.Dataset):
    """Image batches for captioning (reference src/dataset.py:426-476): every file of `root_img_dir` in os.listdir order
    EXCEPT names ending in .png (:445-446), `max_img_per_batch` per batch; an item is
    {"images" (n,3,224,224) float, "tgt_langs" (n,), "first_tokens" (n,), "paths" [n]} with the reference's preprocessing:
    RGB, resize shorter side to 256 (bilinear), centre crop 224, scale to [0,1], ImageNet mean / std; an unreadable image
    becomes a black 224x224 one (:461-470)."""

    MEAN, STD = (0.485, 0.456, 0.406), (0.229, 0.224, 0.225)

    def __init__(self, root_img_dir: str, max_img_per_batch: int, target_lang: int, first_token: int):
        import os
        self.target_lang, self.first_token = target_lang, first_token
        self.image_batches, cur = [], []
        for name in os.listdir(root_img_dir):
            if name.lower().endswith(".png"):
                continue
            cur.append(os.path.join(root_img_dir, name))
            if len(cur) >= max_img_per_batch:
                self.image_batches.append(cur)
                cur = []
        if cur:
            self.image_batches.append(cur)

    def __len__(self):
        return len(self.image_batches)

    @staticmethod
    def get_img(path):
        from PIL import Image
        try:
            with Image.open(path) as im:
                img = im.convert("RGB")
            w, h = img.size
            if w <= h:                                    # torchvision Resize(256): shorter side -> 256, int() of the other
                nw, nh = 256, int(256 * h / w)
            else:
                nw, nh = int(256 * w / h), 256
            img = img.resize((nw, nh), Image.BILINEAR)
            left, top = int(round((nw - 224) / 2.0)), int(round((nh - 224) / 2.0))
            return img.crop((left, top, left + 224, top + 224))
        except Exception:
            print("Corrupted image", path)
            return Image.new("RGB", (224, 224))

    @classmethod
    def to_normalized_tensor(cls, img):
        import numpy as np
        a = torch.from_numpy(np.asarray(img, dtype=np.uint8).copy()).permute(2, 0, 1).to(torch.float32).div(255)
        mean = torch.tensor(cls.MEAN).view(3, 1, 1)
        std = torch.tensor(cls.STD).view(3, 1, 1)
        return (a - mean) / std

    def __getitem__(self, item):
        paths = self.image_batches[item]
        imgs = torch.stack([self.to_normalized_tensor(self.get_img(p)) for p in paths])
        n = len(paths)
        return {"images": imgs, "tgt_langs": torch.LongTensor([self.target_lang] * n),
                "first_tokens": torch.LongTensor([self.first_token] * n), "paths": paths}
