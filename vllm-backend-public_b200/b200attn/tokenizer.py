"""Text <-> token ids for the OpenAI endpoints.

A deployment from the reference names a hub model (`VLLMConfig.model_name`, reference benchmark-vllm/models.py:7);
this box has no network, so unless `--model` (or `--tokenizer`) is a local directory holding `tokenizer.json` the
engine serves random-init weights and text is mapped by bytes: one UTF-8 byte = one token (ids 0..255), generated ids
fold back to bytes (`id % 256`).  Clients that need exact prompt lengths send token-id prompts (OpenAI's completions
API accepts `prompt: [int, ...]`), which is what the load generator does.
"""
from __future__ import annotations

from pathlib import Path
from typing import List, Optional


class ByteTokenizer:
    name = "bytes"
    eos_token_id: Optional[int] = None

    def __init__(self, vocab_size: int):
        if vocab_size < 256:
            raise ValueError("byte tokenizer needs a vocabulary of at least 256 ids")
        self.vocab_size = vocab_size

    def encode(self, text: str) -> List[int]:
        return list(text.encode("utf-8"))

    def decode(self, ids: List[int]) -> str:
        return bytes(i % 256 for i in ids).decode("utf-8", errors="replace")


class HFTokenizer:
    """`tokenizers` fast tokenizer from a local tokenizer.json."""

    def __init__(self, path: Path):
        from tokenizers import Tokenizer
        self._tok = Tokenizer.from_file(str(path / "tokenizer.json"))
        self.name = str(path)
        self.vocab_size = self._tok.get_vocab_size()
        self.eos_token_id = None
        cfg = path / "generation_config.json"
        if cfg.exists():
            import json
            eos = json.loads(cfg.read_text()).get("eos_token_id")
            self.eos_token_id = eos[0] if isinstance(eos, list) else eos

        # the model's own chat template (tokenizer_config.json), rendered with jinja2 as transformers does
        self._chat_template = None
        self._special = {}
        tc = path / "tokenizer_config.json"
        if tc.exists():
            import json
            c = json.loads(tc.read_text())
            t = c.get("chat_template")
            if isinstance(t, list):                      # named templates: take "default"
                t = next((x.get("template") for x in t if x.get("name") == "default"), None)
            self._chat_template = t if isinstance(t, str) else None
            for k in ("bos_token", "eos_token", "pad_token", "unk_token"):
                v = c.get(k)
                self._special[k] = v.get("content") if isinstance(v, dict) else v

    def encode(self, text: str) -> List[int]:
        return self._tok.encode(text).ids

    def encode_rendered(self, text: str) -> List[int]:
        """A rendered chat template already carries its special tokens (BOS ...): do not add them again."""
        return self._tok.encode(text, add_special_tokens=False).ids

    def decode(self, ids: List[int]) -> str:
        return self._tok.decode(ids, skip_special_tokens=True)

    def apply_chat_template(self, messages) -> Optional[str]:
        """The prompt text of a chat, by the model's own template; None when the tokenizer ships none."""
        if not self._chat_template:
            return None
        from jinja2.sandbox import ImmutableSandboxedEnvironment

        def raise_exception(msg):
            raise ValueError(msg)

        env = ImmutableSandboxedEnvironment(trim_blocks=True, lstrip_blocks=True)
        env.globals["raise_exception"] = raise_exception
        return env.from_string(self._chat_template).render(messages=messages, add_generation_prompt=True, **self._special)


def load_tokenizer(model: str, vocab_size: int, tokenizer: Optional[str] = None):
    for cand in (tokenizer, model):
        if cand and (Path(cand) / "tokenizer.json").exists():
            return HFTokenizer(Path(cand))
    return ByteTokenizer(vocab_size)
