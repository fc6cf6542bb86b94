import base64
import os
import sys

import pytest
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box via gpurun)")


def pytest_collection_modifyitems(config, items):
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def spm_model(tmp_path_factory):
    """Tiny sentencepiece model trained offline on synthetic text."""
    import sentencepiece as spm

    d = tmp_path_factory.mktemp("spm")
    corpus = d / "corpus.txt"
    words = ["the", "capital", "of", "germany", "is", "berlin", "city", "here", "my", "sonnet", "in",
             "style", "shakespeare", "about", "an", "artificial", "intelligence", "hello", "world"]
    lines = []
    for i in range(400):
        lines.append(" ".join(words[(i * 7 + j * 3) % len(words)] for j in range(3 + i % 9)).capitalize() + ".")
    corpus.write_text("\n".join(lines))
    prefix = str(d / "tok")
    spm.SentencePieceTrainer.train(input=str(corpus), model_prefix=prefix, vocab_size=300, model_type="bpe",
                                   bos_id=1, eos_id=2, unk_id=0, pad_id=-1, character_coverage=1.0,
                                   minloglevel=2)
    return prefix + ".model"


@pytest.fixture(scope="session")
def tiktoken_model(tmp_path_factory):
    """Synthetic tiktoken BPE file: 256 byte tokens plus a few merges."""
    d = tmp_path_factory.mktemp("tik")
    path = d / "tokenizer.model"
    ranks = [bytes([b]) for b in range(256)]
    merges = [b"th", b"he", b"in", b"er", b"an", b" t", b" a", b"the", b" the", b"ing", b"ll", b"lo", b"hel",
              b"hell", b"hello", b" w", b"or", b"ld", b"wor", b"world"]
    ranks += merges
    with open(path, "w") as f:
        for i, tok in enumerate(ranks):
            f.write(base64.b64encode(tok).decode() + f" {i}\n")
    return str(path)
