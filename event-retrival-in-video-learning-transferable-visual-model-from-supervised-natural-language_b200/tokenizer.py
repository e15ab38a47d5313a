"""`clip.tokenize`: byte-level BPE of openai/CLIP (`clip/simple_tokenizer.py`, `clip/clip.py::tokenize`), restated.

Call site in the reference: `clip.tokenize([query]).to(device)` (Backend/services/embedding_service.py:169,
compare_models.py:1202).  The merge table `bpe_simple_vocab_16e6.txt.gz` ships inside the openai package, which is an
absent submodule here, and there is no network: the file is looked up at `bpe_path=`, $EVR_CLIP_ROOT or
~/.cache/clip/, and `tokenize` raises if it cannot be found (no approximate fallback).  The only deviation from the
original: `ftfy.fix_text` (mojibake repair, not installed) is skipped in `basic_clean`; html unescaping, whitespace
collapsing and lower-casing are kept.
"""
import gzip
import html
import os
from functools import lru_cache

import regex as re
import torch

BPE_FILE = "bpe_simple_vocab_16e6.txt.gz"


@lru_cache()
def bytes_to_unicode():
    """Reversible byte -> printable unicode character table of GPT-2 style BPE."""
    bs = list(range(ord("!"), ord("~") + 1)) + list(range(ord("¡"), ord("¬") + 1)) + list(range(ord("®"), ord("ÿ") + 1))
    cs = bs[:]
    n = 0
    for b in range(2 ** 8):
        if b not in bs:
            bs.append(b)
            cs.append(2 ** 8 + n)
            n += 1
    return dict(zip(bs, [chr(c) for c in cs]))


def get_pairs(word):
    return {(a, b) for a, b in zip(word[:-1], word[1:])}


def basic_clean(text):
    return html.unescape(html.unescape(text)).strip()


def whitespace_clean(text):
    return re.sub(r"\s+", " ", text).strip()


class SimpleTokenizer:
    def __init__(self, bpe_path):
        self.byte_encoder = bytes_to_unicode()
        self.byte_decoder = {v: k for k, v in self.byte_encoder.items()}
        with gzip.open(bpe_path) as f:
            merges = f.read().decode("utf-8").split("\n")
        merges = [tuple(m.split()) for m in merges[1:49152 - 256 - 2 + 1]]
        merges = [m for m in merges if len(m) == 2]
        vocab = list(bytes_to_unicode().values())
        vocab = vocab + [v + "</w>" for v in vocab]
        vocab += ["".join(m) for m in merges]
        vocab += ["<|startoftext|>", "<|endoftext|>"]
        self.encoder = dict(zip(vocab, range(len(vocab))))
        self.decoder = {v: k for k, v in self.encoder.items()}
        self.bpe_ranks = dict(zip(merges, range(len(merges))))
        self.cache = {"<|startoftext|>": "<|startoftext|>", "<|endoftext|>": "<|endoftext|>"}
        self.pat = re.compile(
            r"""<\|startoftext\|>|<\|endoftext\|>|'s|'t|'re|'ve|'m|'ll|'d|[\p{L}]+|[\p{N}]|[^\s\p{L}\p{N}]+""",
            re.IGNORECASE)

    def bpe(self, token):
        if token in self.cache:
            return self.cache[token]
        word = tuple(token[:-1]) + (token[-1] + "</w>",)
        pairs = get_pairs(word)
        if not pairs:
            return token + "</w>"
        while True:
            bigram = min(pairs, key=lambda p: self.bpe_ranks.get(p, float("inf")))
            if bigram not in self.bpe_ranks:
                break
            first, second = bigram
            new_word, i = [], 0
            while i < len(word):
                try:
                    j = word.index(first, i)
                except ValueError:
                    new_word.extend(word[i:])
                    break
                new_word.extend(word[i:j])
                i = j
                if word[i] == first and i < len(word) - 1 and word[i + 1] == second:
                    new_word.append(first + second)
                    i += 2
                else:
                    new_word.append(word[i])
                    i += 1
            word = tuple(new_word)
            if len(word) == 1:
                break
            pairs = get_pairs(word)
        out = " ".join(word)
        self.cache[token] = out
        return out

    def encode(self, text):
        ids = []
        text = whitespace_clean(basic_clean(text)).lower()
        for token in re.findall(self.pat, text):
            token = "".join(self.byte_encoder[b] for b in token.encode("utf-8"))
            ids.extend(self.encoder[t] for t in self.bpe(token).split(" "))
        return ids

    def decode(self, ids):
        text = "".join(self.decoder[int(t)] for t in ids)
        return bytearray(self.byte_decoder[c] for c in text).decode("utf-8", errors="replace").replace("</w>", " ")


_tokenizers = {}


def find_bpe(bpe_path=None):
    for cand in (bpe_path, os.path.join(os.environ.get("EVR_CLIP_ROOT", ""), BPE_FILE),
                 os.path.join(os.path.expanduser("~/.cache/clip"), BPE_FILE)):
        if cand and os.path.isfile(cand):
            return cand
    raise RuntimeError(f"clip.tokenize needs openai/CLIP's merge table {BPE_FILE}: it is not on this box (absent "
                       "submodule, no network); put it under $EVR_CLIP_ROOT or ~/.cache/clip/, or pass bpe_path=")


def get_tokenizer(bpe_path=None):
    path = find_bpe(bpe_path)
    if path not in _tokenizers:
        _tokenizers[path] = SimpleTokenizer(path)
    return _tokenizers[path]


def tokenize(texts, context_length=77, truncate=False, bpe_path=None):
    """clip.tokenize: str | list[str] -> LongTensor [n, context_length] = SOT, BPE ids, EOT, zero padding."""
    if isinstance(texts, str):
        texts = [texts]
    tok = get_tokenizer(bpe_path)
    sot, eot = tok.encoder["<|startoftext|>"], tok.encoder["<|endoftext|>"]
    result = torch.zeros(len(texts), context_length, dtype=torch.long)
    for i, text in enumerate(texts):
        ids = [sot] + tok.encode(text) + [eot]
        if len(ids) > context_length:
            if not truncate:
                raise RuntimeError(f"Input {texts[i]} is too long for context length {context_length}")
            ids = ids[:context_length]
            ids[-1] = eot
        result[i, :len(ids)] = torch.tensor(ids)
    return result
