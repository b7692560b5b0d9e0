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
