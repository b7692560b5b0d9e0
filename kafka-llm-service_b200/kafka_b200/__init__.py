"""kafka_b200 — B200-native completion provider for the thread-based chat server
egrigokhan/kafka-llm-service (drop-in for its LLMProvider plugin API).

Layout:
  cabi.py         ctypes binding of libkafka_b200.so (include/kafka_b200.h)
  engine_core.py  one GPU's replica: weights, paged KV pools, workspace, prefill/decode calls
  model_config.py Llama-family shapes;  weights.py  seeded random-init weights
  kv_layout.py    host-side description of the paged KV layout
The CUDA sources live in ../csrc.
"""

__version__ = "0.1.0"
