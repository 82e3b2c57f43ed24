#!/usr/bin/env python3
"""Time rg_raycast_device on the reference's own scene (512^3 volume of create_data, the grown mask, 512 x 512
pixels, camera 1000) and on larger pictures; CUDA events on the launching stream, L2 flushed between launches.
Prints one JSON line.  Run on the GPU box:  python tests/raycast_timing.py"""
import json, os, sys
import numpy as np
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "parallel-c-programs_b200"))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import rgb200
from conftest import Oracle


def main():
    torch.cuda.set_device(0)
    o = Oracle()
    vol = o.create_data_ref().reshape(512, 512, 512)
    d = torch.from_numpy(vol).cuda()
    ctx = rgb200.Context(0)
    seeds = torch.tensor([[50, 300, 300]], dtype=torch.int32, device="cuda")
    mask = ctx.grow(d, seeds, 1, torch.empty_like(d))
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    out = {"scene": "create_data 512^3 + grown mask, camera 1000", "region_voxels": int(mask.sum().item())}
    for I in (512, 1024, 2048):
        ts = []
        for it in range(8):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            img = rgb200.raycast_device(d, mask, I)
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        ts = sorted(ts[3:])
        out[f"ms_{I}"] = round(ts[len(ts) // 2], 4)
        out[f"Mrays_per_s_{I}"] = round(I * I / ts[len(ts) // 2] / 1e3, 1)
        out[f"lit_px_{I}"] = int((img > 0).sum().item())
    print(json.dumps(out))


if __name__ == "__main__":
    main()
