#!/usr/bin/env python
"""Development aid: per-warp timeline of the tile kernel (csrc/track_tile.cu TPROF stamps).  Run on a GPU with
PBHC_TILE_PROF=1:  python tools/tile_prof.py [envs].  Prints, per block 0..7, the mean cycles per tile each warp spends
waiting at the top barrier, in the per-env work, waiting at the end barrier and in the store / load-issue tail."""
import os
import sys
import tempfile

os.environ["PBHC_TILE_PROF"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from benchkit.workload import Workload  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
wl = Workload(n, "cuda:0", tempfile.mkdtemp(), seed=1)
for _ in range(3):
    wl.step(with_fk=False)
torch.cuda.synchronize()
wl.ts.partials.zero_()
wl.ts.launch()
torch.cuda.synchronize()
print(wl.ts.launch_info())
raw = wl.ts.partials.view(torch.int64).cpu()
for b in range(8):
    t = raw[600 + 16 * b: 600 + 16 * b + 12].reshape(-1)[:16 * 8 * 6].reshape(16, 8, 6).double()
    if t[0, 0, 0] == 0:
        continue
    t0 = t[:, :, 0].min()
    seg = torch.stack([t[:, :, 1] - t[:, :, 0], t[:, :, 2] - t[:, :, 1], t[:, :, 3] - t[:, :, 2], t[:, :, 4] - t[:, :, 3], t[:, :, 5] - t[:, :, 4]], -1)
    m = seg[1:15].mean(0)  # [warp, seg]
    print(f"block {b}: tile period {(t[15, 0, 0] - t[1, 0, 0]).item() / 14:.0f} cycles")
    for w in range(8):
        print(f"   warp {w}: top-wait {m[w, 0]:7.0f}  pre {m[w, 1]:6.0f}  env-work {m[w, 2]:7.0f}  end-wait {m[w, 3]:7.0f}  tail {m[w, 4]:6.0f}")
    if b == 0:
        print("   env-work per tile (cycles), warps 0..7:")
        for i in range(1, 15):
            print("     ", [int(v) for v in seg[i, :, 2]])
