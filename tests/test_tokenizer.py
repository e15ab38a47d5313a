"""`clip.tokenize` restatement (tokenizer.py) on a synthetic merge table: the real bpe_simple_vocab_16e6.txt.gz is
not on this box, so the BPE algorithm is exercised with hand-checkable merges (CPU test)."""
import gzip
import importlib

import pytest
import torch

from conftest import PKG


@pytest.fixture()
def bpe_file(tmp_path):
    merges = ["#version: synthetic", "h e", "l l", "he ll", "hell o</w>", "w o", "wo r", "l d</w>",
              "wor ld</w>", "! !</w>"]
    path = tmp_path / "bpe_simple_vocab_16e6.txt.gz"
    with gzip.open(path, "wb") as f:
        f.write("\n".join(merges).encode("utf-8"))
    return str(path)


def test_bpe_merges_and_layout(bpe_file):
    tk = importlib.import_module(PKG + ".tokenizer")
    tok = tk.SimpleTokenizer(bpe_file)
    n_merges = 9
    assert len(tok.encoder) == 512 + n_merges + 2
    sot, eot = tok.encoder["<|startoftext|>"], tok.encoder["<|endoftext|>"]
    assert (sot, eot) == (512 + n_merges, 512 + n_merges + 1)      # the two specials close the vocabulary
    # "hello" -> h e l l o</w> -> he ll o</w> -> hell o</w> -> hello</w>: one token, the 4th merge
    assert tok.encode("Hello") == [512 + 3]
    # "world" -> wo r l d</w> -> wor ld</w> -> world</w>
    assert tok.encode("  WORLD ") == [512 + 7]
    # unknown word falls back to byte-level symbols; the last carries the end-of-word mark
    be = tk.bytes_to_unicode()
    assert tok.encode("ab") == [tok.encoder[be[ord("a")]], tok.encoder[be[ord("b")] + "</w>"]]
    # punctuation run is one pre-token; digits are split one by one ([\p{N}])
    assert tok.encode("!!") == [512 + 8]
    assert len(tok.encode("2024")) == 4
    # non-ASCII goes through utf-8 bytes: "é" = 2 bytes -> 2 symbols
    assert len(tok.encode("é")) == 2
    assert tok.decode(tok.encode("hello world")) == "hello world "
    out = tk.tokenize(["hello world", "hello"], bpe_path=bpe_file)
    assert out.dtype == torch.long and out.shape == (2, 77)
    assert out[0, :4].tolist() == [sot, 512 + 3, 512 + 7, eot] and out[0, 4:].sum() == 0
    assert out.argmax(-1).tolist() == [3, 2]                        # EOT is the largest id of a row
    assert tk.tokenize("hello", context_length=8, bpe_path=bpe_file).shape == (1, 8)


def test_too_long_input_raises_or_truncates(bpe_file):
    tk = importlib.import_module(PKG + ".tokenizer")
    long_text = " ".join(["hello"] * 100)
    with pytest.raises(RuntimeError, match="too long for context length 77"):
        tk.tokenize([long_text], bpe_path=bpe_file)
    out = tk.tokenize([long_text], truncate=True, bpe_path=bpe_file)
    eot = tk.get_tokenizer(bpe_file).encoder["<|endoftext|>"]
    assert out.shape == (1, 77) and out[0, -1] == eot and (out[0] != 0).all()


def test_missing_merge_table_fails_loudly(tmp_path, monkeypatch):
    tk = importlib.import_module(PKG + ".tokenizer")
    monkeypatch.setenv("EVR_CLIP_ROOT", str(tmp_path))
    monkeypatch.setenv("HOME", str(tmp_path))
    with pytest.raises(RuntimeError, match="bpe_simple_vocab_16e6"):
        tk.tokenize(["a query"])
