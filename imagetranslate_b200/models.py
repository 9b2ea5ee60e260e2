"""Model containers for the decode path: parameter-name compatible with the reference checkpoints
(reference src/seq2seq.py:183-213 `mt_model.state_dict`, key list in SURVEY.md Appendix A.10), so a
released ImageTranslate state dict loads with `load_state_dict`.

Only what decoding needs is implemented: parameter storage for the decoder(s) / output layer(s) (consumed
by engine.PackedDecoder), and the bidirectional encoder forward that produces the memory
(reference src/seq2seq.py:101-108 -> src/bert_seq2seq.py:103-144).  Training forwards are out of scope.
"""
import copy
import math
import os
import pickle

import torch
import torch.nn as nn
import torch.nn.functional as F


class BertConfig(object):
    def __init__(self, vocab_size, type_vocab_size, num_hidden_layers=6, hidden_size=768, intermediate_size=3072,
                 num_attention_heads=12, max_position_embeddings=512, pad_token_id=0, bos_token_id=1, eos_token_id=4,
                 layer_norm_eps=1e-12, initializer_range=0.02, is_decoder=False):
        self.__dict__.update({k: v for k, v in locals().items() if k != "self"})


class Embeddings(nn.Module):
    def __init__(self, c):
        super().__init__()
        self.word_embeddings = nn.Embedding(c.vocab_size, c.hidden_size, padding_idx=c.pad_token_id)
        self.position_embeddings = nn.Embedding(c.max_position_embeddings, c.hidden_size)
        self.token_type_embeddings = nn.Embedding(c.type_vocab_size, c.hidden_size)
        self.LayerNorm = nn.LayerNorm(c.hidden_size, eps=c.layer_norm_eps)

    def forward(self, ids, type_ids):
        pos = torch.arange(ids.size(1), device=ids.device)
        return self.LayerNorm(self.word_embeddings(ids) + self.position_embeddings(pos)[None] + self.token_type_embeddings(type_ids))


class SelfAttention(nn.Module):
    def __init__(self, c):
        super().__init__()
        self.num_attention_heads = c.num_attention_heads
        self.query = nn.Linear(c.hidden_size, c.hidden_size)
        self.key = nn.Linear(c.hidden_size, c.hidden_size)
        self.value = nn.Linear(c.hidden_size, c.hidden_size)

    def forward(self, x, add_mask):
        B, S, d = x.shape
        H = self.num_attention_heads
        sp = lambda t: t.view(B, S, H, d // H).transpose(1, 2)
        o = F.scaled_dot_product_attention(sp(self.query(x)), sp(self.key(x)), sp(self.value(x)), attn_mask=add_mask)
        return o.transpose(1, 2).reshape(B, S, d)


class DenseLN(nn.Module):
    def __init__(self, n_in, n_out, eps):
        super().__init__()
        self.dense = nn.Linear(n_in, n_out)
        self.LayerNorm = nn.LayerNorm(n_out, eps=eps)

    def forward(self, h, residual):
        return self.LayerNorm(self.dense(h) + residual)


class Attention(nn.Module):
    def __init__(self, c):
        super().__init__()
        self.self = SelfAttention(c)
        self.output = DenseLN(c.hidden_size, c.hidden_size, c.layer_norm_eps)

    def forward(self, x, add_mask):
        return self.output(self.self(x, add_mask), x)


class Intermediate(nn.Module):
    def __init__(self, c):
        super().__init__()
        self.dense = nn.Linear(c.hidden_size, c.intermediate_size)

    def forward(self, x):
        return F.gelu(self.dense(x))


class Layer(nn.Module):
    def __init__(self, c):
        super().__init__()
        self.attention = Attention(c)
        if c.is_decoder:
            self.crossattention = Attention(c)
        self.intermediate = Intermediate(c)
        self.output = DenseLN(c.intermediate_size, c.hidden_size, c.layer_norm_eps)

    def forward(self, x, add_mask):
        a = self.attention(x, add_mask)
        return self.output(self.intermediate(a), a)


class Stack(nn.Module):
    def __init__(self, c):
        super().__init__()
        self.layer = nn.ModuleList([Layer(c) for _ in range(c.num_hidden_layers)])


def _init_bert(module, std):
    """HF 2.9.0 init_weights."""
    for m in module.modules():
        if isinstance(m, (nn.Linear, nn.Embedding)):
            m.weight.data.normal_(0.0, std)
        if isinstance(m, nn.LayerNorm):
            m.weight.data.fill_(1.0)
            m.bias.data.zero_()
        if isinstance(m, nn.Linear) and m.bias is not None:
            m.bias.data.zero_()


class BertEncoderModel(nn.Module):
    """Bidirectional encoder (reference src/bert_seq2seq.py:94-144)."""

    def __init__(self, config):
        super().__init__()
        self.config = config
        self.embeddings = Embeddings(config)
        self.encoder = Stack(config)
        _init_bert(self, config.initializer_range)

    def forward(self, input_ids, attention_mask=None, token_type_ids=None):
        x = self.embeddings(input_ids, token_type_ids)
        add_mask = None
        if attention_mask is not None:
            add_mask = (1.0 - attention_mask[:, None, None, :].to(x.dtype)) * -10000.0
        for ly in self.encoder.layer:
            x = ly(x, add_mask)
        return x


class BertDecoderModel(nn.Module):
    """Parameter container of the causal decoder with cross-attention (reference src/bert_seq2seq.py:15-91).
    Its forward is the CUDA decode path (imagetranslate_b200.BeamDecoder)."""

    def __init__(self, config):
        super().__init__()
        self.config = copy.deepcopy(config)
        self.config.is_decoder = True
        self.embeddings = Embeddings(self.config)
        self.decoder = Stack(self.config)
        _init_bert(self, config.initializer_range)


class BertOutputLayer(nn.Module):
    """reference src/bert_seq2seq.py:6-12"""

    def __init__(self, config):
        super().__init__()
        self.layer = nn.Linear(config.hidden_size, config.vocab_size)

    def forward(self, x):
        return self.layer(x)


class Seq2Seq(nn.Module):
    """reference src/seq2seq.py:20-86 (structure + sharing), :101-108 (encode), :183-213 (save/load)."""

    def __init__(self, text_processor, lang_dec: bool = True, use_proposals=False, tie_embed=False, enc_layer: int = 6,
                 dec_layer: int = 3, embed_dim: int = 768, intermediate_dim: int = 3072, freeze_image: bool = False,
                 resnet_depth: int = 1, use_obj: bool = False):
        super().__init__()
        self.text_processor = text_processor
        langs = text_processor.languages
        self.config = BertConfig(vocab_size=text_processor.tokenizer.get_vocab_size(), type_vocab_size=len(langs),
                                 num_hidden_layers=enc_layer, hidden_size=embed_dim, intermediate_size=intermediate_dim,
                                 pad_token_id=text_processor.pad_token_id(), bos_token_id=text_processor.bos_token_id(),
                                 eos_token_id=text_processor.sep_token_id())
        dec_config = copy.deepcopy(self.config)
        dec_config.num_hidden_layers = dec_layer
        self.enc_layer, self.dec_layer, self.embed_dim, self.intermediate_dim = enc_layer, dec_layer, embed_dim, intermediate_dim
        self.lang_dec, self.tie_embed, self.use_proposals = lang_dec, tie_embed, use_proposals
        self.freeze_image, self.resnet_depth = freeze_image, resnet_depth
        self.encoder = BertEncoderModel(self.config)
        ee = self.encoder.embeddings
        if not lang_dec:
            self.decoder = BertDecoderModel(dec_config)
            de = self.decoder.embeddings
            de.position_embeddings.weight = ee.position_embeddings.weight
            de.token_type_embeddings.weight = ee.token_type_embeddings.weight
            de.word_embeddings.weight = ee.word_embeddings.weight
            if tie_embed:
                self.output_layer = BertOutputLayer(dec_config)
                # :56/:59 register the embedding as `output_layer.weight` on the WRAPPER; the projection
                # `output_layer.layer.weight` stays untied.  Kept so that state-dict keys match.
                self.output_layer.weight = de.word_embeddings.weight
            else:
                self.output_layer = nn.ModuleList([BertOutputLayer(dec_config) for _ in langs])
            if enc_layer == dec_layer:                               # :63-65 shared self-attention blocks
                for i in range(enc_layer):
                    self.decoder.decoder.layer[i].attention = self.encoder.encoder.layer[i].attention
        else:
            dec = BertDecoderModel(dec_config)
            self.decoder = nn.ModuleList([copy.deepcopy(dec) for _ in langs])
            self.output_layer = nn.ModuleList([BertOutputLayer(dec_config) for _ in langs])
            for i, dec in enumerate(self.decoder):
                if tie_embed:
                    dec.embeddings.position_embeddings = ee.position_embeddings
                self.output_layer[i].weight = dec.embeddings.word_embeddings.weight     # :75 (wrapper alias, see above)
                # :76-77 _tie_or_clone_weights(out=encoder..., inp=dec...) sets out.weight = inp.weight: the ENCODER's
                # table is re-pointed at each decoder's in turn, so it ends up shared with the LAST decoder only and
                # every other decoder keeps its own (deep-copied) token-type table.
                ee.token_type_embeddings.weight = dec.embeddings.token_type_embeddings.weight

        if use_proposals:                                            # :79-83; consumed by the attend_proposal kernel
            self.proposal_embedding = ee.word_embeddings
            self.lexical_gate = nn.Parameter(torch.zeros(1, embed_dim).fill_(0.1), requires_grad=True)
            self.lexical_layer_norm = nn.LayerNorm(embed_dim, eps=self.config.layer_norm_eps)

    def encode(self, src_inputs, src_mask, src_langs, images=None):
        dev = self.encoder.embeddings.word_embeddings.weight.device
        return (self.encoder(src_inputs.to(dev), attention_mask=src_mask.to(dev), token_type_ids=src_langs.to(dev)), None)

    def save(self, out_dir):
        os.makedirs(out_dir, exist_ok=True)
        with open(os.path.join(out_dir, "mt_config"), "wb") as fp:
            pickle.dump((self.lang_dec, self.use_proposals, self.enc_layer, self.dec_layer, self.embed_dim,
                         self.intermediate_dim, self.tie_embed, self.resnet_depth, self.freeze_image), fp)
        torch.save(self.state_dict(), os.path.join(out_dir, "mt_model.state_dict"))

    @staticmethod
    def load(cls, out_dir: str, tok_dir, use_obj: bool = False, map_location=None):
        """reference src/seq2seq.py:198-213: `Seq2Seq.load(Seq2Seq, model_dir, tok_dir=...)`.  `tok_dir` is the tokenizer
        directory (or an already built text processor); the two files are the reference's own checkpoint format."""
        from .textprocessor import TextProcessor
        text_processor = TextProcessor(tok_model_path=tok_dir) if isinstance(tok_dir, str) else tok_dir
        if map_location is None:
            map_location = torch.device("cuda" if torch.cuda.is_available() else "cpu")
        with open(os.path.join(out_dir, "mt_config"), "rb") as fp:
            lang_dec, use_proposals, enc_layer, dec_layer, embed_dim, intermediate_dim, tie_embed, resnet_depth, freeze_image = pickle.load(fp)
        m = cls(text_processor=text_processor, lang_dec=lang_dec, use_proposals=use_proposals, tie_embed=tie_embed,
                enc_layer=enc_layer, dec_layer=dec_layer, embed_dim=embed_dim, intermediate_dim=intermediate_dim,
                freeze_image=freeze_image, resnet_depth=resnet_depth, use_obj=use_obj)
        m.load_state_dict(torch.load(os.path.join(out_dir, "mt_model.state_dict"), map_location=map_location), strict=False)
        return m


class MassSeq2Seq(Seq2Seq):
    """reference src/mass_seq2seq.py -- identical decode surface."""


# ------------------------------------------------------------------------------------------------ image front-end
_RESNETS = {1: ("basic", [2, 2, 2, 2], {}), 2: ("basic", [3, 4, 6, 3], {}), 3: ("bottleneck", [3, 4, 6, 3], {}),
            4: ("bottleneck", [3, 4, 23, 3], {}), 5: ("bottleneck", [3, 8, 36, 3], {}),
            6: ("bottleneck", [3, 4, 23, 3], dict(groups=32, width_per_group=8))}        # src/image_model.py:85-97


def _make_grid_resnet_class():
    import torchvision.models.resnet as tvr

    class GridResnet(tvr.ResNet):
        """ResNet trunk whose 7x7 layer4 grid becomes the decoder's memory (reference src/image_model.py:14-82
        `ModifiedResnet`, built by `init_net` :85-124): (B,C,7,7) -> (B,49,C) -> fc (no bias) + location_embedding.
        Parameter names equal the reference's (torchvision ResNet keys + fc, object_feat_fc, location_embedding,
        object_embedding, layer_norm), so `image_model.*` of a released checkpoint loads.

        The convolution trunk runs on cuDNN through torchvision.  `fc + location_embedding` runs as ONE
        tcgen05 GEMM of this library on CUDA: the location table rides in the GEMM as 64 extra K columns driven by a one-hot
        position code, and the (B,49,d) memory leaves the GEMM in the decoder's activation dtype.

        Reference quirk, made explicit: `F.dropout(grid_hidden, p=self.dropout)` at :37-38 has no `training=` argument, so it
        is active at INFERENCE too (two eval calls differ).  `inference_dropout=False` (default) leaves it out --
        deterministic captions; set it True to reproduce the reference's noisy inference."""

        def setup(self, embed_dim, dropout, use_obj):
            self.dropout = dropout
            self.inference_dropout = False
            self.layer_norm = nn.LayerNorm(embed_dim, eps=1e-12)
            self.fc = nn.Linear(self.fc.in_features, embed_dim, bias=False)
            self.object_feat_fc = nn.Linear(1024 + 7 + embed_dim, embed_dim, bias=False)
            self.location_embedding = nn.Embedding(49, embed_dim)
            self.object_embedding = nn.Embedding(91, embed_dim)
            self.fcnn = None             # Faster R-CNN object detector (src/faster_rcnn_feats.py): supplied by the caller
            self.use_obj = use_obj
            self._packed = None
            return self

        def grid(self, x):
            x = self.maxpool(self.relu(self.bn1(self.conv1(x))))
            x = self.layer4(self.layer3(self.layer2(self.layer1(x))))
            return x.flatten(2).permute(0, 2, 1)                 # (B, 49, C)

        def _fc_location(self, grid):
            """grid (B,49,C) -> fc(grid) + location_embedding, (B,49,d)."""
            B, P, Cc = grid.shape
            w = self.fc.weight
            if not grid.is_cuda:
                return self.fc(grid) + self.location_embedding.weight.unsqueeze(0)
            from . import _lib
            code = _lib.F32 if w.dtype == torch.float32 else _lib.F16
            act = torch.float32 if code == _lib.F32 else torch.float16
            key = (w._version, w.data_ptr(), self.location_embedding.weight._version, self.location_embedding.weight.data_ptr(), code)
            if self._packed is None or self._packed[0] != key:
                d = w.shape[0]
                w_aug = torch.zeros(d, Cc + 64, dtype=act, device=w.device)
                w_aug[:, :Cc] = w.detach().to(act)
                w_aug[:, Cc:Cc + P] = self.location_embedding.weight.detach().t().to(act)
                self._packed = (key, w_aug, torch.zeros(d, dtype=torch.float32, device=w.device))
            _, w_aug, zero_bias = self._packed
            a = torch.zeros(B * P, Cc + 64, dtype=act, device=grid.device)
            a[:, :Cc] = grid.reshape(B * P, Cc)
            a.view(B, P, Cc + 64)[:, :, Cc:Cc + P] = torch.eye(P, dtype=act, device=grid.device)
            out = torch.empty(B * P, w_aug.shape[0], dtype=act, device=grid.device)
            import ctypes as C
            with torch.cuda.device(grid.device):
                stream = C.c_void_p(torch.cuda.current_stream(grid.device).cuda_stream)
                _lib.check(_lib.lib().itb_linear(code, _lib.GEMM_AUTO, _lib.EPI_BIAS_ACT, _lib.ptr(a), _lib.ptr(w_aug), _lib.ptr(zero_bias),
                                                 _lib.ptr(out), B * P, w_aug.shape[0], Cc + 64, w_aug.shape[0], stream), "fc+location")
            return out.view(B, P, -1)

        def object_features(self, x, ref):
            """reference src/image_model.py:44-75: detector outputs -> padded (B, n_obj, d) object memory."""
            if self.fcnn is None:
                return None
            with torch.no_grad():
                res = self.fcnn(x)
                n_max = max(r["boxes"].size(0) for r in res)
                if n_max == 0:
                    return None
                fdim = res[0]["features"].size(-1)
                feats = torch.full((len(res), n_max, fdim + 7), 1e-4, dtype=ref.dtype, device=ref.device)
                labels = torch.zeros((len(res), n_max), dtype=torch.long, device=ref.device)
                for i, r in enumerate(res):
                    b = r["boxes"] / 800
                    x1, y1, x2, y2 = b[:, 0], b[:, 1], b[:, 2], b[:, 3]
                    w_, h_ = x2 - x1, y2 - y1
                    locs = torch.stack([x1, x2, y1, y2, w_, h_, h_ * w_], dim=-1)
                    feats[i, :r["features"].size(0)] = torch.cat([r["features"], locs], dim=-1)
                    labels[i, :r["labels"].size(0)] = r["labels"]
            obj = torch.cat([self.object_embedding(labels), feats], dim=-1)
            obj[labels == 0] = 0
            return F.relu(self.object_feat_fc(obj))

        def forward(self, x):
            grid = self.grid(x)
            if self.dropout > 0 and (self.training or self.inference_dropout):
                grid = F.dropout(grid, p=self.dropout)                                  # :37-38 (see class docstring)
            out = self._fc_location(grid)
            obj = self.object_features(x, out)
            if self.dropout > 0 and self.training:                                      # :77-80
                out = F.dropout(out, p=self.dropout)
                if obj is not None:
                    obj = F.dropout(obj, p=self.dropout)
            return out, obj

    return GridResnet, tvr


def init_net(embed_dim: int, dropout: float = 0.1, freeze: bool = False, depth: int = 1, use_obj: bool = True):
    """reference src/image_model.py:85-124, random init (no download; a checkpoint supplies the weights)."""
    GridResnet, tvr = _make_grid_resnet_class()
    kind, layers, extra = _RESNETS[depth]
    net = GridResnet(tvr.BasicBlock if kind == "basic" else tvr.Bottleneck, layers, **extra).setup(embed_dim, dropout, use_obj)
    if freeze:
        for p in net.parameters():
            p.requires_grad = False
    return net


class ImageCaptioning(MassSeq2Seq):
    """reference src/image_model.py:267-306 (`ImageCaptioning(ImageMassSeq2Seq)`): text decoder over the ResNet grid memory.
    State-dict compatible with a reference checkpoint: image_model.*, multimodal_attention_gate, image_attention_w.*,
    encoder_attention_w.* (:139-143, unused in decoding but present in every checkpoint) and, with use_obj, obj_decoder.* and
    multistream_attention_gate (:279-296)."""

    def __init__(self, text_processor, freeze_image: bool = False, resnet_depth: int = 1, lang_dec: bool = False,
                 use_proposals: bool = False, tie_embed: bool = False, enc_layer: int = 6, dec_layer: int = 3,
                 embed_dim: int = 768, intermediate_dim: int = 3072, use_obj: bool = True):
        super().__init__(text_processor, lang_dec=lang_dec, use_proposals=use_proposals, tie_embed=tie_embed, enc_layer=enc_layer,
                         dec_layer=dec_layer, embed_dim=embed_dim, intermediate_dim=intermediate_dim, freeze_image=freeze_image,
                         resnet_depth=resnet_depth)
        self.image_model = init_net(embed_dim=embed_dim, dropout=0.1, freeze=freeze_image, depth=resnet_depth, use_obj=use_obj)
        self.multimodal_attention_gate = nn.Parameter(torch.zeros(1, embed_dim).fill_(0.1), requires_grad=True)
        self.image_attention_w = nn.Linear(embed_dim, 1)
        self.encoder_attention_w = nn.Linear(embed_dim, 1)
        if use_obj:
            obj_cfg = copy.deepcopy(self.config)           # :281 BertDecoderModel(self.config): ENCODER layer count
            if not lang_dec:
                self.obj_decoder = BertDecoderModel(obj_cfg)
            else:
                dec = BertDecoderModel(obj_cfg)
                self.obj_decoder = nn.ModuleList([copy.deepcopy(dec) for _ in text_processor.languages])
                for i, dec in enumerate(self.obj_decoder):
                    if tie_embed:
                        dec.embeddings.position_embeddings = self.encoder.embeddings.position_embeddings
                    self.output_layer[i].weight = dec.embeddings.word_embeddings.weight
                    self.encoder.embeddings.token_type_embeddings.weight = dec.embeddings.token_type_embeddings.weight
            self.multistream_attention_gate = nn.Parameter(torch.zeros(1, embed_dim).fill_(0.1), requires_grad=True)

    def encode(self, src_inputs=None, src_mask=None, src_langs=None, images=None):
        """reference src/image_model.py:298-309"""
        if images is not None:
            if isinstance(images, list):
                images = images[0]
            w = self.encoder.embeddings.word_embeddings.weight
            return self.image_model(images.to(device=w.device, dtype=w.dtype))
        return super().encode(src_inputs, src_mask, src_langs)


class SyntheticTextProcessor(object):
    """Minimal stand-in for reference src/textprocessor.py when no tokenizer files exist (benchmarks):
    pad 0, bos 1, unk 2, mask 3, </s> 4, language tokens from 5 (src/textprocessor.py:24-30)."""

    class _Tok(object):
        def __init__(self, v):
            self._v = v

        def get_vocab_size(self):
            return self._v

    def __init__(self, vocab_size, languages=("<en>", "<fa>")):
        self.tokenizer = SyntheticTextProcessor._Tok(vocab_size)
        self.languages = {l: i for i, l in enumerate(languages)}

    def pad_token_id(self):
        return 0

    def bos_token_id(self):
        return 1

    def sep_token_id(self):
        return 4

    def token_id(self, tok):
        return 5 + self.languages[tok]
