"""Pinned host -> device copy rate of this box by chunk size (the ceiling of bench.py's e2e)."""
import json
import torch

dev = torch.device("cuda:0")
n = 1 << 30
src = torch.empty(n, dtype=torch.uint8).pin_memory()
dst = torch.empty(n, dtype=torch.uint8, device=dev)
out = {}
for chunk in (1 << 20, 4 << 20, 16 << 20, 64 << 20, 1 << 30):
    for rep in range(2):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for o in range(0, n, chunk):
            dst[o:o + chunk].copy_(src[o:o + chunk], non_blocking=True)
        e1.record()
        torch.cuda.synchronize()
    out[f"{chunk >> 20}MiB"] = round(n / e0.elapsed_time(e1) / 1e6, 2)
# two streams, 16 MiB chunks alternating
s = [torch.cuda.Stream(), torch.cuda.Stream()]
for rep in range(2):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for st in s:
        st.wait_stream(torch.cuda.current_stream())
    for i, o in enumerate(range(0, n, 16 << 20)):
        with torch.cuda.stream(s[i & 1]):
            dst[o:o + (16 << 20)].copy_(src[o:o + (16 << 20)], non_blocking=True)
    for st in s:
        torch.cuda.current_stream().wait_stream(st)
    e1.record()
    torch.cuda.synchronize()
out["16MiB_two_streams"] = round(n / e0.elapsed_time(e1) / 1e6, 2)
print(json.dumps({"h2d_GBps": out}))
