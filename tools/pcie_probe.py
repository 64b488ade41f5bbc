"""Development aid: pinned-memory copy bandwidth of the box (the ceiling of bench.py's e2e leg)."""
import torch
dev = "cuda:0"
n = 256 << 20
h_in, h_out = torch.empty(n, dtype=torch.uint8).pin_memory(), torch.empty(n, dtype=torch.uint8).pin_memory()
d_in, d_out = torch.empty(n, dtype=torch.uint8, device=dev), torch.empty(n, dtype=torch.uint8, device=dev)
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def t(fn, reps=5):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    s1.synchronize(); s2.synchronize(); e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


ms = t(lambda: d_in.copy_(h_in, non_blocking=True)); print(f"H2D alone  {n / ms / 1e6:.1f} GB/s")
ms = t(lambda: h_out.copy_(d_out, non_blocking=True)); print(f"D2H alone  {n / ms / 1e6:.1f} GB/s")


def both():
    with torch.cuda.stream(s1):
        d_in.copy_(h_in, non_blocking=True)
    with torch.cuda.stream(s2):
        h_out.copy_(d_out, non_blocking=True)
    torch.cuda.current_stream().wait_stream(s1); torch.cuda.current_stream().wait_stream(s2)


ms = t(both); print(f"H2D + D2H concurrently: {n / ms / 1e6:.1f} GB/s each way")
