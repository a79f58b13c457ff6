#!/usr/bin/env python
"""L2-resident copy rate of the box's GPU (the denominator for kernels whose working set never
leaves L2, such as traversal of the cornell tree). Copies a buffer that fits in L2 back and forth and
prints the read + write rate; the same for a buffer far larger than L2 gives the HBM copy rate for comparison."""
import json

import torch


def rate(mbytes, reps=200):
    n = mbytes * (1 << 20) // 4
    a = torch.empty(n, dtype=torch.float32, device="cuda").normal_()
    b = torch.empty_like(a)
    for _ in range(10):
        b.copy_(a)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(reps):
        b.copy_(a)
    e1.record()
    torch.cuda.synchronize()
    return 2.0 * n * 4 * reps / (e0.elapsed_time(e1) * 1e-3) / 1e9


if __name__ == "__main__":
    out = {"l2_copy_gbs": {str(mb): round(rate(mb), 1) for mb in (8, 16, 32, 48)},
           "hbm_copy_gbs": {"2048": round(rate(2048, reps=10), 1)}}
    print(json.dumps(out))
