"""Text side of the decode drivers: the attribute surface of the reference's `TextProcessor`
(reference src/textprocessor.py:10-150) that translate.py / caption.py and BeamDecoder touch -- tokenizer files of a
trained ImageTranslate tokenizer directory, the pickled language table, the special-token ids, `tokenize_one_sentence`
and `tokenizer.decode`.  Tokenizer TRAINING and the long-document splitting helpers are outside the decode path.

A tokenizer directory holds `vocab.json` + `merges.txt` (what `tokenizers` 0.7 wrote for the reference,
src/textprocessor.py:42-45) and the pickle `langs`; directories written by newer `tokenizers` releases hold a single
`tokenizer.json` instead (a merge containing a newline cannot be re-read from merges.txt there).  Both load.
"""
import os
import pickle
from typing import Dict, List, Optional


class TextProcessor(object):
    def __init__(self, tok_model_path: Optional[str] = None):
        self.languages: Dict[str, int] = {}
        self.tokenizer = None
        if tok_model_path is not None:
            import tokenizers
            single = os.path.join(tok_model_path, "tokenizer.json")
            if os.path.isfile(single):
                self.tokenizer = tokenizers.Tokenizer.from_file(single)
            else:
                self.tokenizer = tokenizers.SentencePieceBPETokenizer.from_file(os.path.join(tok_model_path, "vocab.json"),
                                                                                os.path.join(tok_model_path, "merges.txt"))
            with open(os.path.join(tok_model_path, "langs"), "rb") as fp:
                self.languages = pickle.load(fp)
        self.init_properties(self.languages)

    def init_properties(self, languages: Dict[str, int] = None):
        """reference src/textprocessor.py:22-32"""
        self.max_len = 512
        self.pad_token, self.mask_token, self.unk_token, self.sep_token, self.bos = "<pad>", "<mask>", "<unk>", "</s>", "<s>"
        self.special_tokens = [self.pad_token, self.bos, self.unk_token, self.mask_token, self.sep_token] + list(languages.keys())
        self.languages = languages

    def save(self, directory: str):
        os.makedirs(directory, exist_ok=True)
        self.tokenizer.save(os.path.join(directory, "tokenizer.json"))
        with open(os.path.join(directory, "langs"), "wb") as fp:
            pickle.dump(self.languages, fp)

    def _tokenize(self, line):
        return self.tokenizer.encode(line)

    def tokenize_one_sentence(self, line) -> List[int]:
        """reference src/textprocessor.py:64-73: `<lang> words ... </s>` -> ids, cut at 512."""
        spl = line.strip().split(" ")
        lang_id, sen, eos = spl[0], " ".join(spl[1:-1]), spl[-1]
        return ([self.token_id(lang_id)] + self._tokenize(sen).ids + [self.token_id(eos)])[:512]

    def tokenize_one_sentence_with_langid(self, line, lang_id) -> List[int]:
        """reference src/textprocessor.py:75-77"""
        return ([lang_id] + self._tokenize(line).ids + [self.token_id("</s>")])[:512]

    def pad_token_id(self) -> int:
        return self.tokenizer.token_to_id(self.pad_token)

    def mask_token_id(self) -> int:
        return self.tokenizer.token_to_id(self.mask_token)

    def unk_token_id(self) -> int:
        return self.tokenizer.token_to_id(self.unk_token)

    def bos_token_id(self) -> int:
        return self.tokenizer.token_to_id(self.bos)

    def sep_token_id(self) -> int:
        return self.tokenizer.token_to_id(self.sep_token)

    def token_id(self, token: str) -> int:
        tok_id = self.tokenizer.token_to_id(token)
        return 0 if tok_id is None else tok_id

    def id2token(self, id: int) -> str:
        return self.tokenizer.id_to_token(id)

    def vocab_size(self) -> int:
        return self.tokenizer.get_vocab_size()

    def is_lang(self, id) -> bool:
        return self.tokenizer.id_to_token(id) in self.languages

    def lang_id(self, tok):
        return self.languages.get(tok, 0)
