"""Where does a short GEMM spend its time?  Needs a trace build of the library:
    make -C kafka-llm-service_b200/csrc clean all EXTRA=-DKB_GEMM_TRACE
Prints, for the CTA-pair kernel's cluster 0, nanoseconds between: kernel entry -> setup done -> first
operands landed -> last operands landed -> accumulator complete -> rows stored -> cluster sync -> TMEM freed,
next to the CUDA-event time of the whole launch.  (Rebuild without EXTRA afterwards.)
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
lib.kb_debug_gemm_trace.argtypes = [C.POINTER(C.c_ulonglong)]
names = ["setup", "first_operands", "mainloop", "acc_complete", "store", "cluster_sync", "tmem_free"]
for M, N, K, epi, tile in [(512, 6144, 4096, cabi.EPI_BF16, -256), (512, 4096, 4096, cabi.EPI_RESID_F32, -128),
                           (512, 28672, 4096, cabi.EPI_SWIGLU_F16, -256), (512, 4096, 14336, cabi.EPI_RESID_F32, -128),
                           (512, 6144, 256, cabi.EPI_BF16, -256)]:
    ws = [(torch.randn(N, K, device=dev) * 0.02).half() for _ in range(max(2, int(300e6 // (N * K * 2)) + 1))]
    a = torch.randn(M, K, device=dev).half()
    resid = torch.zeros(M, N, device=dev) if epi == cabi.EPI_RESID_F32 else None
    o = cabi.gemm(a, ws[0], epi, resid=resid, out=resid, tile_n=tile)   # residual projections run in place
    big = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    rows = []
    for i in range(6):
        big.zero_()                                        # flush L2 and idle the launch queue
        if resid is not None:
            resid.add_(1.0)                                # ... but the residual stream is L2-warm in the real step
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        cabi.gemm(a, ws[i % len(ws)], epi, resid=resid, out=o, tile_n=tile)
        e1.record()
        torch.cuda.synchronize()
        t = (C.c_ulonglong * 16)()
        assert lib.kb_debug_gemm_trace(t) == 0
        d = {n: int(t[k + 1]) - int(t[k]) for k, n in enumerate(names)}
        d["kernel_total"] = int(t[7]) - int(t[0])
        d["event_us"] = round(e0.elapsed_time(e1) * 1e3, 1)
        rows.append(d)
    print(json.dumps({"M": M, "N": N, "K": K, "tile": tile, "ns": rows[2:]}))
