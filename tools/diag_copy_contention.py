"""Which kind of GPU load slows a pinned host->device copy?  60 MB H2D copies timed alone and under
(a) an HBM-streaming device copy, (b) an L2-resident device copy, (c) a bf16 GEMM, (d) an idle-SM sleep."""
import json
import torch

dev = torch.device("cuda:0")
NB = 60_211_200
host = torch.randint(0, 255, (NB,), dtype=torch.uint8).pin_memory()
d = torch.empty(NB, dtype=torch.uint8, device=dev)
s_c, s_k = torch.cuda.Stream(), torch.cuda.Stream()
big_a = torch.empty(1 << 30, dtype=torch.uint8, device=dev); big_b = torch.empty_like(big_a)
sm_a = torch.empty(24 << 20, dtype=torch.uint8, device=dev); sm_b = torch.empty_like(sm_a)
A = torch.randn(8192, 8192, device=dev, dtype=torch.bfloat16); B = torch.randn(8192, 8192, device=dev, dtype=torch.bfloat16)
loads = {
    "none": (lambda: None, 1),
    "hbm_stream_copy": (lambda: big_b.copy_(big_a), 80),
    "l2_resident_copy": (lambda: sm_b.copy_(sm_a), 3000),
    "bf16_gemm_8192": (lambda: torch.matmul(A, B), 30),
}
out = {}
for name, (fn, reps) in loads.items():
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ka, kb = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(s_k):
        ka.record()
        for _ in range(reps):
            fn()
        kb.record()
    with torch.cuda.stream(s_c):
        a.record()
        for _ in range(20):
            d.copy_(host, non_blocking=True)
        b.record()
    torch.cuda.synchronize()
    t = a.elapsed_time(b) / 20
    out[name] = {"copy_ms": round(t, 4), "copy_GBps": round(NB / t / 1e6, 1), "load_total_ms": round(ka.elapsed_time(kb), 2)}
print(json.dumps(out))
