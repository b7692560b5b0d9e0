"""Deterministic offline tokenizer shared by the B200 provider and the CPU oracle.

No tokenizer assets exist offline and the reference has no tokenizer at all —
it ships text to the gateway (/root/reference/src/llm/portkey.py:263, `m.to_dict()`). This one is
byte-level: UTF-8 byte b <-> id b (0..255), with four Llama-3-style control tokens at the top of the
model's vocabulary. Control tokens are only ever inserted by the chat template; text is never
scanned for their spellings, so user content cannot inject them.

Round trip: ``encode(decode(ids)) == ids`` for every id sequence that is valid UTF-8 without control
tokens — in particular for the sampler's default generable set (printable ASCII + newline), which is
what makes cached generated tokens a prefix of the re-templated history on the next turn.
"""
from __future__ import annotations

import codecs
from typing import Iterable, List


class ByteTokenizer:
    N_BYTES = 256

    def __init__(self, vocab_size: int):
        if vocab_size < self.N_BYTES + 4:
            raise ValueError(f"vocab_size {vocab_size} too small for the byte tokenizer (needs >= 260)")
        self.vocab_size = vocab_size
        self.cache_id = ("byte", vocab_size)   # identifies the vocabulary (template.RenderCache)
        self.bos_id = vocab_size - 4           # <|begin_of_text|>
        self.start_header_id = vocab_size - 3  # <|start_header_id|>
        self.end_header_id = vocab_size - 2    # <|end_header_id|>
        self.eot_id = vocab_size - 1           # <|eot_id|>
        self.special = {
            self.bos_id: "<|begin_of_text|>", self.start_header_id: "<|start_header_id|>",
            self.end_header_id: "<|end_header_id|>", self.eot_id: "<|eot_id|>",
        }

    def encode(self, text: str) -> List[int]:
        return list(text.encode("utf-8"))

    def decode(self, ids: Iterable[int], skip_special: bool = True) -> str:
        out: List[str] = []
        buf = bytearray()
        for i in ids:
            if 0 <= i < self.N_BYTES:
                buf.append(i)
                continue
            if buf:
                out.append(buf.decode("utf-8", errors="replace"))
                buf = bytearray()
            if i in self.special:
                if not skip_special:
                    out.append(self.special[i])
            else:
                out.append("�")  # ids between the bytes and the control tokens carry no text
        if buf:
            out.append(buf.decode("utf-8", errors="replace"))
        return "".join(out)

    def stream_decoder(self) -> "StreamDecoder":
        return StreamDecoder(self)

    # default candidate set of the sampler: printable ASCII [32,127) + newline + <|eot_id|>
    def generable_range(self):
        return 32, 127, 10, self.eot_id


class StreamDecoder:
    """Token-by-token detokeniser: emits text as soon as a UTF-8 sequence completes."""

    def __init__(self, tok: ByteTokenizer):
        self.tok = tok
        self._dec = codecs.getincrementaldecoder("utf-8")(errors="replace")

    def push(self, token_id: int) -> str:
        if 0 <= token_id < ByteTokenizer.N_BYTES:
            return self._dec.decode(bytes([token_id]))
        if token_id in self.tok.special:
            return ""
        return self._dec.decode(b"", final=False) + "�"

    def flush(self) -> str:
        return self._dec.decode(b"", final=True)


class HFTokenizer:
    """A checkpoint's own tokenizer (`tokenizer.json`, read with the `tokenizers` library) behind the interface the
    template, the provider and the sampler use. The four control tokens are looked up by their Llama-3 spellings;
    a tokenizer without them (Llama-2's sentencepiece vocabulary, say) is refused, since the chat template is built
    from them. As with the byte tokenizer, text is encoded WITHOUT parsing control-token spellings out of it."""

    CONTROL = {"bos_id": "<|begin_of_text|>", "start_header_id": "<|start_header_id|>",
               "end_header_id": "<|end_header_id|>", "eot_id": "<|eot_id|>"}

    def __init__(self, tokenizer_json: str, vocab_size: int):
        from tokenizers import Tokenizer
        self._tok = Tokenizer.from_file(tokenizer_json)
        self.vocab_size = vocab_size
        import os
        self.cache_id = ("hf", os.path.abspath(tokenizer_json), os.path.getmtime(tokenizer_json), vocab_size)
        missing = []
        for attr, text in self.CONTROL.items():
            i = self._tok.token_to_id(text)
            if i is None:
                missing.append(text)
            setattr(self, attr, i)
        if missing:
            raise ValueError(f"{tokenizer_json}: control tokens {missing} not in the vocabulary — the chat template "
                             "needs Llama-3's")
        if self._tok.get_vocab_size(with_added_tokens=True) > vocab_size:
            raise ValueError(f"{tokenizer_json}: {self._tok.get_vocab_size(with_added_tokens=True)} tokens, the model's "
                             f"embedding has {vocab_size} rows")
        added = self._tok.get_added_tokens_decoder()
        self.special = {int(i): t.content for i, t in added.items() if t.special}
        # ordinary tokens = everything below the first control / special id (Llama-3: 0 .. 127 999)
        self._first_special = min(self.special) if self.special else vocab_size
        # what every control spelling starts with ("<|" for Llama-3's 256): text without it needs no splitting, and
        # that one substring test replaces one per spelling on every message
        import os.path
        self._probe = os.path.commonprefix(list(self.special.values())) if self.special else ""

    def encode(self, text: str) -> List[int]:
        if not text:
            return []
        # split on control-token spellings so that none of them is ever produced from text: each spelling is
        # encoded piecewise ("<|", "eot_id", "|>" ...) like any other characters
        ids: List[int] = []
        for piece in self._split_specials(text):
            ids += self._tok.encode(piece, add_special_tokens=False).ids
        return ids

    def _split_specials(self, text: str) -> List[str]:
        if self._probe and self._probe not in text:
            return [text]
        pieces = [text]
        for spelling in set(self.special.values()):
            nxt: List[str] = []
            for p in pieces:
                if spelling not in p:
                    nxt.append(p)
                    continue
                parts = p.split(spelling)
                for k, part in enumerate(parts):
                    if part:
                        nxt.append(part)
                    if k + 1 < len(parts):        # the spelling itself, cut in two so that it cannot match
                        nxt += [spelling[:2], spelling[2:]]
            pieces = nxt
        return pieces

    def decode(self, ids: Iterable[int], skip_special: bool = True) -> str:
        ids = [int(i) for i in ids if 0 <= int(i) < self.vocab_size]
        return self._tok.decode(ids, skip_special_tokens=skip_special)

    def stream_decoder(self) -> "HFStreamDecoder":
        return HFStreamDecoder(self)

    def generable_range(self):
        """Every ordinary token, plus <|eot_id|>."""
        return 0, self._first_special, self.eot_id, -1


class HFStreamDecoder:
    """Token-by-token detokeniser for a BPE vocabulary: a token may end in the middle of a UTF-8 sequence, and the
    spacing of a token can depend on its left neighbour, so the text emitted for a new token is the difference
    between decoding the pending ids with and without it, held back while it ends in a replacement character."""

    def __init__(self, tok: HFTokenizer):
        self.tok = tok
        self._ids: List[int] = []
        self._emitted = ""

    def push(self, token_id: int) -> str:
        if token_id in self.tok.special:
            return ""
        self._ids.append(int(token_id))
        text = self.tok.decode(self._ids)
        if text.endswith("�"):
            return ""                              # an incomplete multi-byte character: wait for its remaining bytes
        new = text[len(self._emitted):]
        self._emitted = text
        if len(self._ids) > 64 and "\n" in new:    # keep the re-decoded window short: restart after a line break
            self._ids, self._emitted = [], ""
        return new

    def flush(self) -> str:
        text = self.tok.decode(self._ids)
        new = text[len(self._emitted):]
        self._ids, self._emitted = [], ""
        return new
