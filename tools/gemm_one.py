"""One GEMM shape, a few launches — a target for `ncu -k regex:gemm_tcgen05 ...` (source-level stall reasons of the
epilogue). Usage: python tools/gemm_one.py M N K epilogue tile_n [in_place]"""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "kafka-llm-service_b200"))
import torch  # noqa: E402
from kafka_b200 import cabi  # noqa: E402

M, N, K, epi, tile = (int(v) for v in sys.argv[1:6])
dev = torch.device("cuda:0")
torch.cuda.set_device(dev)
a = torch.randn(M, K, device=dev).half()
w = (torch.randn(N, K, device=dev) * 0.02).half()
resid = torch.zeros(M, N, device=dev) if epi == cabi.EPI_RESID_F32 else None
for _ in range(6):
    cabi.gemm(a, w, epi, resid=resid, out=resid, tile_n=tile)
torch.cuda.synchronize()
print("ok")
