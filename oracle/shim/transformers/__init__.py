"""ORACLE / TEST INFRASTRUCTURE ONLY -- never imported by the product path.

Stand-in for the `transformers==2.9.0` package that the reference pins
(/root/reference/src/requirements.txt:8) and star-imports
(/root/reference/src/bert_seq2seq.py:3, src/lm.py:8, src/image_model.py:5).
That release cannot be installed offline, so its published BERT algorithm is
restated in `modeling_bert.py`.  This directory is put
first on sys.path ONLY inside the oracle process (oracle/reference_loader.py).
"""
from .modeling_bert import BertConfig, BertModel  # noqa: F401
