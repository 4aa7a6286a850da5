"""b2serve's byte-level BPE tokenizer (csrc/server/tokenizer.hpp) against the HF `tokenizers` library on the
same tokenizer.json (CPU, no GPU).  No real checkpoint's tokenizer file exists offline, so the test trains
small byte-level BPE tokenizers on this repository's own text with exactly the pre-tokenizer set-ups the
served models ship -- Qwen2/Qwen3 (Split on the Qwen2 regex + ByteLevel without its own regex, added special
tokens) and GPT-2/OPT (ByteLevel with its regex) -- writes them as tokenizer.json, and requires identical ids
and a loss-free round trip on text chosen to hit every branch of the two split patterns."""
import json
import os
import subprocess

import pytest

from _util import ROOT

tokenizers = pytest.importorskip("tokenizers")
from tokenizers import Regex, Tokenizer, decoders, models, pre_tokenizers, trainers  # noqa: E402

EXE = os.path.join(ROOT, "aws-k8s-ansible-provisioner_b200", "b2serve")
QWEN2_PATTERN = (r"(?i:'s|'t|'re|'ve|'m|'ll|'d)|[^\r\n\p{L}\p{N}]?\p{L}+|\p{N}| ?[^\s\p{L}\p{N}]+[\r\n]*|\s*[\r\n]+|\s+(?!\S)|\s+")

TEXTS = [
    "Who are you?",
    "Hello, world! It's 2026 and we're testing: don't, I'LL, you'Ve, he'd, 'tis.",
    "  leading spaces,   triple  spaces and a trailing one ",
    "tabs\tand\nnewlines\r\n\r\nmixed \n \n  with spaces\n",
    "numbers 1234567890 3.14159 1e-6 0xFF 12ab34",
    "punctuation!!! ... ---> <== [brackets] {braces} (parens) #hash @at $ % ^ & * ~ ` | \\ /",
    "code: for (int i = 0; i < n; ++i) { a[i] += b[i] * 2; }\n\treturn a;",
    "Ünïcödé letters: naïve café Straße Ελληνικά русский язык العربية עברית हिन्दी",
    "CJK: 你好，世界！これはテストです。한국어도 있습니다。",
    "emoji 😀👍🏽 and symbols ±×÷ → ∑ ∞ © ® ™ ½ ² ①",
    "<|im_start|>system\nBe brief.<|im_end|>\n<|im_start|>user\nWho are you?<|im_end|>\n<|im_start|>assistant\n",
    "text<|endoftext|>more text <|im_end|><|im_end|>",
    "a",
    " ",
    "\n",
    "x\u00a0y\u2003z\u3000w",            # non-ASCII white space
    "mixed١٢٣digits and Ⅻ roman",      # non-ASCII numbers (Nd, Nl)
    "'",
    "''s 'S' 's",
]


def corpus():
    out = []
    for name in ("SURVEY.md", "DESIGN.md", "INTEGRATION.md", "bench.py"):
        with open(os.path.join(ROOT, name), encoding="utf-8") as f:
            out.append(f.read())
    out.extend(TEXTS)
    return out


def build(kind: str, path: str) -> Tokenizer:
    tok = Tokenizer(models.BPE())
    if kind == "qwen2":
        tok.pre_tokenizer = pre_tokenizers.Sequence([
            pre_tokenizers.Split(Regex(QWEN2_PATTERN), behavior="isolated", invert=False),
            pre_tokenizers.ByteLevel(add_prefix_space=False, use_regex=False)])
        special = ["<|endoftext|>", "<|im_start|>", "<|im_end|>"]
    else:
        tok.pre_tokenizer = pre_tokenizers.ByteLevel(add_prefix_space=False, use_regex=True)
        special = ["</s>", "<pad>"]
    tok.decoder = decoders.ByteLevel()
    trainer = trainers.BpeTrainer(vocab_size=1500, special_tokens=special, initial_alphabet=pre_tokenizers.ByteLevel.alphabet(),
                                  show_progress=False)
    tok.train_from_iterator(corpus(), trainer)
    tok.save(path)
    return tok


def run_ours(tokenizer_json: str, texts):
    inp = "".join(json.dumps({"text": t}) + "\n" for t in texts)
    r = subprocess.run([EXE, "--tokenize", tokenizer_json], input=inp, capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr
    return [json.loads(line) for line in r.stdout.splitlines() if line.strip()]


@pytest.mark.parametrize("kind", ["qwen2", "gpt2"])
def test_bpe_matches_hf_tokenizers(tmp_path, kind):
    if not os.path.exists(EXE):
        pytest.skip("b2serve not built")
    path = str(tmp_path / "tokenizer.json")
    tok = build(kind, path)
    texts = TEXTS + corpus()[:2]  # the long documents too
    ours = run_ours(path, texts)
    assert len(ours) == len(texts)
    for t, o in zip(texts, ours):
        ref = tok.encode(t, add_special_tokens=False).ids
        assert o["ids"] == ref, "ids differ for %r:\n ours %s\n ref  %s" % (t[:80], o["ids"][:40], ref[:40])
        # round trip: specials are skipped on decode (vLLM's default skip_special_tokens=True)
        assert o["text"] == tok.decode(ref, skip_special_tokens=True)


def test_unsupported_tokenizer_is_refused(tmp_path):
    if not os.path.exists(EXE):
        pytest.skip("b2serve not built")
    tok = Tokenizer(models.BPE())
    tok.pre_tokenizer = pre_tokenizers.Metaspace()  # SentencePiece-style: not built
    trainer = trainers.BpeTrainer(vocab_size=300, show_progress=False)
    tok.train_from_iterator(TEXTS, trainer)
    path = str(tmp_path / "tokenizer.json")
    tok.save(path)
    r = subprocess.run([EXE, "--tokenize", path], input='{"text": "x"}\n', capture_output=True, text=True, timeout=60)
    assert r.returncode != 0 and "pre-tokenizer" in r.stderr
