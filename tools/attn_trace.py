"""Per-tile timeline of the tcgen05 prefill attention kernel (CTA 0): who waits for whom.
Needs a trace build:  make -C kafka-llm-service_b200/csrc clean all EXTRA=-DKB_ATTN_TRACE   (rebuild without EXTRA afterwards).

Slots per 128-key tile j of the default kernel (attn_prefill_tc128_kernel), ns relative to slot 0 of tile 4:
  0 S thread: K/V of tile j landed   | 1 S thread: O += P V of tile j-2 seen retired | 2 S_j issued + committed
  3 softmax (warp 0): S_j seen done  | 4 softmax (warp 0): P_j published
  5 PV thread: P_j seen              | 6 PV thread: O += P_j V_j issued + committed
(The 64-key kernels, KB_ATTN_TC64=1, use their own slot meanings: see KB_ATRACE in csrc/kb_attn_tc.cu.)
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
rows = [[int(t[j * 8 + s]) for s in range(8)] for j in range(17)]
base = rows[4][0]
for j in range(4, 12):
    print(j, [v - base if v else None for v in rows[j][:7]])
d = lambda a, b: sum(rows[j][b] - rows[j][a] for j in range(4, 14)) / 10
print(json.dumps({"ns_per_tile": (rows[14][3] - rows[4][3]) / 10,
                  "S issue (1->2)": d(1, 2), "S committed -> softmax sees it (2->3)": d(2, 3),
                  "softmax (3->4)": d(3, 4), "P published -> PV thread sees it (4->5)": d(4, 5),
                  "PV issue (5->6)": d(5, 6),
                  "PV_j committed -> S thread sees it retired, tile j+2 (6->1)":
                      sum(rows[j + 2][1] - rows[j][6] for j in range(4, 12)) / 8}))
