"""Translation driver with the reference's command line (reference src/translate.py:14-31,101-123):

  python -m imagetranslate_b200.translate --tok T --model M --input I --output O --src fa --target en \\
         [--beam 4] [--batch 4000] [--capacity 600] [--max_len_a 1.3] [--max_len_b 5] [--len-penalty 0.8] [--fp16] [--verbose]

Same helper names and meaning as the reference (`get_lm_option_parser`, `build_model`, `build_data_loader`,
`translate_batch`), so a script written against src/translate.py keeps working.  What differs is how batches reach the GPU:
the reference calls the generator once per batch (src/translate.py:107-120), which leaves a B200 latency-bound (a batch at
--batch 4000 has 60-940 beam rows); `translate_batches` hands a WINDOW of consecutive batches to
`BeamDecoder.forward_many`, which decodes them in lockstep as super-batches of up to --rows beam rows on --lanes streams.
Each batch is still decoded exactly as the reference decodes it (its own loop bound, termination rule and memory), so the
output file equals the batch-by-batch one; input tensors come out of a pinned, prefetching DataLoader.

Multi-GPU: launch under torchrun (one process per GPU).  Whole batches are assigned to ranks by cost (parallel.shard_batches),
outputs return to rank 0 in ONE all-gather, rank 0 writes the file in input order (replaces DataParallelModel, src/parallel.py).
"""
import datetime
import os
from optparse import OptionParser

import torch
import torch.utils.data as data_utils

from . import dataset
from .models import Seq2Seq
from .seq_gen import BeamDecoder, get_outputs_until_eos


def get_lm_option_parser():
    parser = OptionParser()
    parser.add_option("--input", dest="input_path", metavar="FILE", default=None)
    parser.add_option("--src", dest="src_lang", type="str", default=None)
    parser.add_option("--target", dest="target_lang", type="str", default=None)
    parser.add_option("--output", dest="output_path", metavar="FILE", default=None)
    parser.add_option("--batch", dest="batch", help="Batch size", type="int", default=4000)
    parser.add_option("--tok", dest="tokenizer_path", help="Path to the tokenizer folder", metavar="FILE", default=None)
    parser.add_option("--cache_size", dest="cache_size", help="Number of blocks in cache", type="int", default=300)
    parser.add_option("--model", dest="model_path", metavar="FILE", default=None)
    parser.add_option("--verbose", action="store_true", dest="verbose", help="Include input!", default=False)
    parser.add_option("--beam", dest="beam_width", type="int", default=4)
    parser.add_option("--max_len_a", dest="max_len_a", help="a for beam search (a*l+b)", type="float", default=1.3)
    parser.add_option("--max_len_b", dest="max_len_b", help="b for beam search (a*l+b)", type="int", default=5)
    parser.add_option("--len-penalty", dest="len_penalty_ratio", help="Length penalty", type="float", default=0.8)
    parser.add_option("--capacity", dest="total_capacity", help="Batch capacity", type="int", default=600)
    parser.add_option("--fp16", action="store_true", dest="fp16", default=False)
    # additions of this driver (defaults keep the reference's results)
    parser.add_option("--rows", dest="rows", help="beam rows per super-batch", type="int", default=4096)
    parser.add_option("--lanes", dest="lanes", help="super-batches in flight per GPU", type="int", default=3)
    parser.add_option("--window", dest="window", help="batches handed to the decoder at once", type="int", default=256)
    return parser


def batch_kwargs(batch, text_processor):
    """The keyword arguments of src/translate.py:46-49 for one DataLoader item (batch_size=1 adds a leading dim)."""
    src_inputs, tgt_inputs = batch["src_texts"], batch["dst_texts"]
    if src_inputs.dim() == 3:                                   # DataLoader(batch_size=1) item
        src_inputs, tgt_inputs = src_inputs.squeeze(0), tgt_inputs.squeeze(0)
    B, S = src_inputs.shape
    return dict(src_inputs=src_inputs, src_sizes=batch["pad_idx"].reshape(B), first_tokens=tgt_inputs[:, 0],
                src_mask=batch["src_pad_mask"].reshape(B, S), src_langs=batch["src_langs"].reshape(B),
                tgt_langs=batch["dst_langs"].reshape(B), pad_idx=text_processor.pad_token_id())


def translate_batch(batch, generator, text_processor, verbose=False):
    """reference src/translate.py:34-56, one batch through `generator.forward`."""
    kw = batch_kwargs(batch, text_processor)
    src_text = None
    if verbose:
        src_ids = get_outputs_until_eos(text_processor.sep_token_id(), kw["src_inputs"], remove_first_token=True)
        src_text = [text_processor.tokenizer.decode(s.numpy()) for s in src_ids]
    outputs = generator(**kw)
    mt_output = [text_processor.tokenizer.decode(x[1:].numpy()) for x in outputs]
    return mt_output, src_text


def translate_batches(batches, generator, text_processor, verbose=False, rows=4096, lanes=3):
    """A window of DataLoader items through `forward_many`; returns [(mt_output, src_text)] per batch, in order."""
    kws = [batch_kwargs(b, text_processor) for b in batches]
    outs = generator.forward_many(kws, lanes=lanes, rows=rows)
    res = []
    for kw, o in zip(kws, outs):
        src_text = None
        if verbose:
            src_ids = get_outputs_until_eos(text_processor.sep_token_id(), kw["src_inputs"], remove_first_token=True)
            src_text = [text_processor.tokenizer.decode(s.numpy()) for s in src_ids]
        res.append(([text_processor.tokenizer.decode(x[1:].numpy()) for x in o], src_text))
    return res


def build_data_loader(options, text_processor):
    """reference src/translate.py:59-82: tokenise `<src> line </s>`, fixed first target token, token-budget batches."""
    print(datetime.datetime.now(), "Binarizing test data")
    assert options.src_lang is not None
    assert options.target_lang is not None
    src_lang = "<" + options.src_lang + ">"
    src_lang_id = text_processor.languages[src_lang]
    dst_lang = "<" + options.target_lang + ">"
    target_lang = text_processor.languages[dst_lang]
    fixed_output = [text_processor.token_id(dst_lang)]
    examples = []
    with open(options.input_path, "r") as s_fp:
        for src_line in s_fp:
            if len(src_line.strip()) == 0:
                continue
            src_line = " ".join([src_lang, src_line, "</s>"])
            src_tok_line = text_processor.tokenize_one_sentence(src_line.strip().replace(" </s> ", " "))
            examples.append((src_tok_line, fixed_output, src_lang_id, target_lang))
    print(datetime.datetime.now(), "Loaded %d examples" % len(examples))
    test_data = dataset.MTDataset(examples=examples, max_batch_capacity=options.total_capacity, max_batch=options.batch,
                                  pad_idx=text_processor.pad_token_id(), max_seq_len=10000)
    return data_utils.DataLoader(test_data, batch_size=1, shuffle=False, pin_memory=torch.cuda.is_available(),
                                 num_workers=0)


def build_model(options):
    """reference src/translate.py:85-98.  --fp16: the reference wraps the generator with apex O2 (fp16 weights and
    activations, fp32 scores); here it selects the fp16 tensor-core kernels.  Without it the fp32 parity kernels run."""
    model = Seq2Seq.load(Seq2Seq, options.model_path, tok_dir=options.tokenizer_path)
    if not torch.cuda.is_available():
        raise RuntimeError("imagetranslate_b200.translate needs a CUDA device (no CPU fallback)")
    device = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")))
    model = model.to(device)
    model.eval()
    generator = BeamDecoder(model, beam_width=options.beam_width, max_len_a=options.max_len_a,
                            max_len_b=options.max_len_b, len_penalty_ratio=options.len_penalty_ratio,
                            dtype="fp16" if options.fp16 else "fp32")
    generator.eval()
    return generator, model.text_processor


def run(options):
    from . import parallel
    world, rank = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0"))
    if world > 1 and not torch.distributed.is_initialized():
        torch.distributed.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0"))))
    generator, text_processor = build_model(options)
    test_loader = build_data_loader(options, text_processor)
    batches = test_loader.dataset.batches
    mine = list(range(len(batches)))
    if world > 1:
        costs = [parallel.batch_cost(int(b["src_texts"].size(0)), int(b["src_texts"].size(1)), options.beam_width,
                                     options.max_len_a, options.max_len_b) for b in batches]
        mine = parallel.shard_batches(costs, world)[rank]
    first_id, o = [], 0
    for b in batches:
        first_id.append(o)
        o += int(b["src_texts"].size(0))
    lines = {}                       # sentence index -> output line
    sen_count = 0
    with torch.no_grad():
        it = iter(test_loader) if world == 1 else None
        for w0 in range(0, len(mine), options.window):
            idx = mine[w0:w0 + options.window]
            window = [next(it) for _ in idx] if it is not None else [batches[i] for i in idx]
            try:
                res = translate_batches(window, generator, text_processor, options.verbose, options.rows, options.lanes)
            except RuntimeError as err:                  # reference src/translate.py:119-120: report and go on
                print("\n", repr(err))
                continue
            for bi, (mt_output, src_text) in zip(idx, res):
                for k, y in enumerate(mt_output):
                    lines[first_id[bi] + k] = y if not options.verbose else (y + " ||| " + src_text[k])
                sen_count += len(mt_output)
            print(datetime.datetime.now(), "Translated", sen_count, "sentences", end="\r")
    if world > 1:
        ids = sorted(lines)
        enc = [torch.tensor(list(lines[i].encode("utf-8")), dtype=torch.int64) for i in ids]
        got = parallel.gather_outputs(ids, enc, torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0"))))
        if rank != 0:
            return sen_count
        lines = {i: bytes(t.tolist()).decode("utf-8") for i, t in got.items()}
    with open(options.output_path, "w") as writer:
        for i in sorted(lines):
            writer.write(lines[i])
            writer.write("\n")
    print(datetime.datetime.now(), "Translated", len(lines), "sentences")
    print(datetime.datetime.now(), "Done!")
    return len(lines)


if __name__ == "__main__":
    parser = get_lm_option_parser()
    (options, args) = parser.parse_args()
    run(options)
