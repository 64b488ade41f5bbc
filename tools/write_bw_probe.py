#!/usr/bin/env python
"""HBM bandwidth by direction on this GPU (torch library kernels, CUDA events, best of 10 over 2-GiB buffers):
write-only (fill_), read-only (sum) and copy (read + write).  K1's DRAM traffic is almost pure writes (its table is
L2-resident), so the write-only figure is the ceiling that matters for it.  Development aid."""
import json

import torch


def best(fn, n=10):
    fn(); torch.cuda.synchronize()
    ts = []
    for _ in range(n):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return min(ts)


x = torch.empty(1 << 29, device="cuda", dtype=torch.float32)  # 2 GiB
y = torch.empty_like(x)
by = x.numel() * 4
print(json.dumps({"write_only_gbs": by / best(lambda: x.fill_(1.0)) / 1e6,
                  "read_only_gbs": by / best(lambda: x.sum()) / 1e6,
                  "copy_gbs_read_plus_write": 2 * by / best(lambda: y.copy_(x)) / 1e6}))
