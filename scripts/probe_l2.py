"""L2-gather roofline probe: useful GB/s of independent random loads from a table-sized resident buffer.
Run on the GPU box: python scripts/probe_l2.py  -> one JSON line per (buffer size, element size, pattern)."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nerf_navigation_b200 import _lib as ngp


def probe(buf_bytes, elem, pattern, blocks=148 * 16, iters=64, reps=5):
    L = ngp.lib()
    buf = torch.randint(0, 2 ** 31 - 1, (buf_bytes // 4,), device="cuda", dtype=torch.int32)
    sink = torch.zeros(4, device="cuda", dtype=torch.int32)
    st = torch.cuda.current_stream().cuda_stream
    run = lambda: ngp.check(L.ngp_probe_gather(buf.data_ptr(), buf_bytes, elem, iters, pattern, blocks, sink.data_ptr(), st))
    for _ in range(3):
        run()
    best = 1e9
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); run(); b.record(); torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b))
    loads = blocks * 256 * iters * 8
    return {"buf_MiB": buf_bytes / 2 ** 20, "elem_bytes": elem, "pattern": pattern, "ms": best,
            "Gloads_s": loads / best / 1e6, "useful_GBs": loads * elem / best / 1e6, "sector_GBs": loads * 32 / best / 1e6}


if __name__ == "__main__":
    for mib in (4.0, 24.15, 48.3, 96.0):
        for elem in (4, 8, 16):
            for pattern in (0, 1):
                print(json.dumps(probe(int(mib * 2 ** 20) // 16 * 16, elem, pattern)), flush=True)
