"""Kernel variants that the library picks by launch size or by an opt-in switch, forced through their environment switch
in a child process (the switches are read once per process) so the small parity cases cover them too:
 * ITB_ATTN_STAGE=32 -- the 32-position / 4-CTAs-per-SM build of the tensor-core attention kernel, which the library
   only selects for launches of more than ~2 waves (4096 beam rows); ITB_ATTN_STAGE=64 is the small-launch build;
 * ITB_VOCAB_PAIR=1 -- the CTA-pair (tcgen05 cta_group::2) launch of the fused vocabulary projection;
 * ITB_GEMM_PAIR=1 -- the CTA-pair launch of the layer GEMMs (gemm_tc2_kernel), opt-in like the one above;
 * ITB_ATTN_BULK=1 / 0 -- the bulk-copy producer / consumer attention kernel (csrc/attention_bulk.cu), which the library
   selects for launches of >= 96 work items (sentences x beam groups), forced on for every 16-bit attention launch / off;
 * ITB_ATTN_SPLIT=N -- the warp-per-head attention kernel shares an item's stages among 4 warps up to N items and among 2 up to
   2N (default N = 3 x SM count, so the small parity cases run the 4-way split): 20 puts the kernel tests (36 items and
   more) on the 2-way split, 0 on the unsplit kernel that large launches use."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(env, select, path="tests/test_gpu_kernels.py"):
    e = dict(os.environ)
    e.update(env)
    r = subprocess.run([sys.executable, "-m", "pytest", "-q", "-x", "-m", "gpu", "-p", "no:cacheprovider", path, "-k", select],
                       cwd=ROOT, env=e, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]
    assert " passed" in r.stdout and "no tests ran" not in r.stdout, r.stdout[-500:]


@pytest.mark.parametrize("stage", ["32", "64"])
def test_attention_kernels_in_both_staging_builds(stage):
    _run({"ITB_ATTN_STAGE": stage}, "attention")


@pytest.mark.parametrize("split", ["0", "20"])
def test_attention_kernels_unsplit_and_two_way_split(split):
    _run({"ITB_ATTN_SPLIT": split}, "attention")


def test_decode_parity_on_the_two_way_split_attention():
    _run({"ITB_ATTN_SPLIT": "40"}, "16bit or forward_many", path="tests/test_gpu_decode.py")       # 37..80 items: two warps per item


def test_decode_parity_with_the_many_wave_attention_build():
    _run({"ITB_ATTN_STAGE": "32"}, "16bit or forward_many or encoder", path="tests/test_gpu_decode.py")


def test_layer_gemms_on_cta_pairs():
    """ITB_GEMM_PAIR=1: every tcgen05 linear (all shapes of test_linear_*, ragged M and N included) through the
    cta_group::2 kernel with 256 x 256 tiles; then a whole decode on it"""
    _run({"ITB_GEMM_PAIR": "1"}, "linear")
    _run({"ITB_GEMM_PAIR": "1"}, "teacher_forced_logprob_error or free_running or forward_many", path="tests/test_gpu_decode.py")


def test_fused_vocab_projection_on_cta_pairs():
    _run({"ITB_VOCAB_PAIR": "1"}, "fused_vocab")


def test_attention_kernels_on_the_bulk_copy_kernel():
    _run({"ITB_ATTN_BULK": "1"}, "attention")


def test_decode_parity_on_the_bulk_copy_attention_kernel():
    _run({"ITB_ATTN_BULK": "1"}, "16bit or forward_many or encoder", path="tests/test_gpu_decode.py")
    _run({"ITB_ATTN_BULK": "1"}, "whole_vocabulary or free_running", path="tests/test_gpu_parity16.py")


def test_decode_variants_on_the_bulk_copy_attention_kernel():
    _run({"ITB_ATTN_BULK": "1"}, "text_plus_image", path="tests/test_gpu_decode_variants.py")


def test_full_size_superbatch_without_the_bulk_copy_kernel():
    """the warp-per-head kernel at the full launch size (what round 1 shipped), kept as the small-launch path"""
    _run({"ITB_ATTN_BULK": "0"}, "full_size", path="tests/test_gpu_decode.py")
