#!/usr/bin/env python
"""Times the two steps of the chained serial-order sum (pcp_serial_sum_begin / _end) on a 25 M-point range with a
carry, as the plan phase of the sharded pipeline runs them."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import pointcloudprocess_b200 as pcp
    ctx = pcp.Context(0)
    ctx.set_stream(torch.cuda.current_stream().cuda_stream)
    n = 25_000_000
    buf = torch.empty(n * 48, dtype=torch.uint8, device="cuda")
    ctx.synth_cloud("city", 200_000_000, first=3 * n, n=n, seed=5, out=buf.data_ptr())
    st = ctx.cloud_stats((buf.data_ptr(), n, 48))
    pred = st[6:9] * 3.0
    carry = torch.tensor(pred * (1 + 1e-9), dtype=torch.float64, device="cuda")
    for rep in range(4):
        c = carry.clone()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        h = ctx.serial_sum_begin((buf.data_ptr(), n, 48), pred)
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        ctx.serial_sum_end(h, c.data_ptr())
        torch.cuda.synchronize()
        t2 = time.perf_counter()
        print("begin %.3f ms, end %.3f ms, sum %r" % ((t1 - t0) * 1e3, (t2 - t1) * 1e3, c.cpu().numpy()))
    ctx.close()


if __name__ == "__main__":
    main()
