"""Per-tile timeline of the tcgen05 prefill attention kernel (CTA 0): who waits for whom.
Needs a trace build:  make -C kafka-llm-service_b200/csrc clean all EXTRA=-DKB_ATTN_TRACE   (rebuild without EXTRA afterwards).

Slots per tile j (ns, relative to the MMA thread's start of tile 8):
  0 MMA thread begins iteration j | 1 S_{j+1} issued | 2 OT_j issued
  3 softmax begins iteration j    | 4 S_j ready      | 5 P_j written | 6 OT_{j-1} ready | 7 waits of issue_s(j+1) over
"""
import ctypes as C
import json
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "kafka-llm-service_b200"))
import torch
from kafka_b200 import cabi

dev = torch.device("cuda:0")
torch.cuda.set_device(dev)
lib = cabi.load()
lib.kb_debug_attn_trace.argtypes = [C.POINTER(C.c_ulonglong)]
n_seqs, ctx, new, n_heads, n_kv, hd = 64, 2048, 64, 32, 8, 128
bpt = (ctx + new + 63) // 64
g = torch.Generator(device="cpu").manual_seed(0)
kp = torch.randn(n_seqs * bpt, n_kv, 64 * hd, device=dev).half()
vp = torch.randn(n_seqs * bpt, n_kv, 64 * hd, device=dev).half()
bt = torch.randperm(n_seqs * bpt, generator=g).int().view(n_seqs, bpt).to(dev)
qkv = torch.randn(n_seqs * new, (n_heads + 2 * n_kv) * hd, device=dev).half()
cu = torch.arange(0, (n_seqs + 1) * new, new, dtype=torch.int32, device=dev)
ctxs = torch.full((n_seqs,), ctx, dtype=torch.int32, device=dev)
for _ in range(3):
    cabi.attn_prefill(qkv, kp, vp, cu, ctxs, bt, new, n_heads, n_kv, hd)
torch.cuda.synchronize()
t = (C.c_ulonglong * 512)()
assert lib.kb_debug_attn_trace(t) == 0
rows = [[int(t[j * 8 + s]) for s in range(8)] for j in range(34)]
base = rows[8][0]
for j in range(8, 16):
    print(j, [v - base if v else None for v in rows[j]])
per = (rows[30][3] - rows[10][3]) / 20
print(json.dumps({"ns_per_tile": per}))
