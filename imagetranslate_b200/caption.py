"""Captioning driver with the reference's command line (reference src/caption.py:15-29,76-93):

  python -m imagetranslate_b200.caption --input IMAGE_DIR --output O --target ar --tok T --model M [--batch 16] [--beam 4] [--fp16] [--obj]

Same helpers as the reference (`get_lm_option_parser`, `build_model`, `build_data_loader`, `caption_batch`).  The image
batches of `ImageDataset` (16 images each by default) are decoded a WINDOW at a time through `BeamDecoder.forward_many`, which
runs the ResNet front-end per batch and the beam search of many batches in lockstep (a 16-image batch alone is 64 beam rows).
`max_len=256` as in src/caption.py:38-39.  Output line = `<path>\\t<caption>` (:90).
"""
import datetime
import os
from optparse import OptionParser

import torch
import torch.utils.data as data_utils

from . import dataset
from .models import ImageCaptioning, Seq2Seq
from .seq_gen import BeamDecoder


def get_lm_option_parser():
    parser = OptionParser()
    parser.add_option("--input", dest="input_path", metavar="FILE", default=None)
    parser.add_option("--target", dest="target_lang", type="str", default=None)
    parser.add_option("--output", dest="output_path", metavar="FILE", default=None)
    parser.add_option("--batch", dest="batch", help="Batch size", type="int", default=16)
    parser.add_option("--tok", dest="tokenizer_path", help="Path to the tokenizer folder", metavar="FILE", default=None)
    parser.add_option("--model", dest="model_path", metavar="FILE", default=None)
    parser.add_option("--beam", dest="beam_width", type="int", default=4)
    parser.add_option("--max_len_a", dest="max_len_a", help="a for beam search (a*l+b)", type="float", default=1.3)
    parser.add_option("--max_len_b", dest="max_len_b", help="b for beam search (a*l+b)", type="int", default=5)
    parser.add_option("--len-penalty", dest="len_penalty_ratio", help="Length penalty", type="float", default=0.8)
    parser.add_option("--fp16", action="store_true", dest="fp16", default=False)
    parser.add_option("--obj", action="store_true", dest="obj", default=False)
    # additions of this driver
    parser.add_option("--rows", dest="rows", help="beam rows per super-batch", type="int", default=4096)
    parser.add_option("--lanes", dest="lanes", help="super-batches in flight per GPU", type="int", default=3)
    parser.add_option("--window", dest="window", help="image batches handed to the decoder at once", type="int", default=64)
    return parser


def batch_kwargs(batch, text_processor):
    """Keyword arguments of src/caption.py:38-39 for one DataLoader item."""
    images = batch["images"]
    if images.dim() == 5:                                       # DataLoader(batch_size=1) item
        images = images.squeeze(0)
    n = images.size(0)
    return dict(first_tokens=batch["first_tokens"].reshape(n), images=images, tgt_langs=batch["tgt_langs"].reshape(n),
                pad_idx=text_processor.pad_token_id(), max_len=256)


def _paths(batch):
    # DataLoader's default collate turns the list of n paths into a list of n 1-tuples (the reference prints x[0], :90)
    return [p[0] if isinstance(p, (tuple, list)) else p for p in batch["paths"]]


def caption_batch(batch, generator, text_processor):
    """reference src/caption.py:32-46"""
    outputs = generator(**batch_kwargs(batch, text_processor))
    return [text_processor.tokenizer.decode(x[1:].numpy()) for x in outputs], batch["paths"]


def caption_batches(batches, generator, text_processor, rows=4096, lanes=3):
    outs = generator.forward_many([batch_kwargs(b, text_processor) for b in batches], lanes=lanes, rows=rows)
    return [([text_processor.tokenizer.decode(x[1:].numpy()) for x in o], _paths(b)) for b, o in zip(batches, outs)]


def build_data_loader(options, text_processor):
    """reference src/caption.py:49-59"""
    print(datetime.datetime.now(), "Binarizing test data")
    assert options.target_lang is not None
    dst_lang = "<" + options.target_lang + ">"
    target_lang = text_processor.languages[dst_lang]
    first_token = text_processor.token_id(dst_lang)
    image_data = dataset.ImageDataset(options.input_path, options.batch, first_token=first_token, target_lang=target_lang)
    return data_utils.DataLoader(image_data, batch_size=1, shuffle=False, pin_memory=torch.cuda.is_available(),
                                 num_workers=min(4, os.cpu_count() or 1), prefetch_factor=2)


def build_model(options):
    """reference src/caption.py:62-73.  The reference never calls model.eval() here, which leaves every dropout of the
    decoder active while captioning; this driver evaluates deterministically (model.eval())."""
    model = Seq2Seq.load(ImageCaptioning, options.model_path, tok_dir=options.tokenizer_path, use_obj=options.obj)
    if not torch.cuda.is_available():
        raise RuntimeError("imagetranslate_b200.caption needs a CUDA device (no CPU fallback)")
    device = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")))
    model = model.to(device)
    if options.fp16:
        model = model.half()
    model.eval()
    generator = BeamDecoder(model, beam_width=options.beam_width, max_len_a=options.max_len_a,
                            max_len_b=options.max_len_b, len_penalty_ratio=options.len_penalty_ratio,
                            dtype="fp16" if options.fp16 else "fp32")
    return generator, model.text_processor


def run(options):
    generator, text_processor = build_model(options)
    test_loader = build_data_loader(options, text_processor)
    sen_count = 0
    with open(options.output_path, "w") as writer, torch.no_grad():
        window = []

        def flush():
            nonlocal sen_count
            for mt_output, paths in caption_batches(window, generator, text_processor, options.rows, options.lanes):
                sen_count += len(mt_output)
                writer.write("\n".join([x + "\t" + y for x, y in zip(paths, mt_output)]))
                writer.write("\n")
            print(datetime.datetime.now(), "Captioned", sen_count, "images!", end="\r")
            del window[:]

        for batch in test_loader:
            window.append(batch)
            if len(window) >= options.window:
                flush()
        if window:
            flush()
    print(datetime.datetime.now(), "Translated", sen_count, "sentences")
    print(datetime.datetime.now(), "Done!")
    return sen_count


if __name__ == "__main__":
    parser = get_lm_option_parser()
    (options, args) = parser.parse_args()
    run(options)
