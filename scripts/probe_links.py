"""Link probes that size the sampler: random small reads from pinned host memory (PCIe) and HBM."""
import ctypes as C
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from coala_gnn_b200._lib import lib, check  # noqa: E402

L = lib()
host = torch.empty(2 << 30, dtype=torch.uint8).pin_memory()
host.zero_()
dev = torch.empty(2 << 30, dtype=torch.uint8, device="cuda")
v = C.c_double()
res = {}
for name, buf in (("host", host), ("hbm", dev)):
    for warps in (1184, 4736, 9472):
        for span, mis in ((8, 0), (32, 0), (64, 0), (128, 0), (128, 1), (256, 0)):
            check(L.coala_probe_random_read(buf.data_ptr(), buf.numel(), span, mis, warps, 64, C.byref(v)))
            key = "%s_w%d_span%d%s" % (name, warps, span, "_mis" if mis else "")
            res[key] = round(v.value / 1e6, 1)
            print(key, "%.1f M locations/s  (%.1f GB/s useful)" % (v.value / 1e6, v.value * span / 1e9), flush=True)
print(json.dumps(res))
