"""rg_grow_multi on the GPUs of this box: cold (first) call and warm calls, C4 geometry.  Tuning / evidence aid.
usage: python tests/multi_timing.py [n] [ndev]"""
import ctypes, json, os, sys, time
import numpy as np
import torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "parallel-c-programs_b200"))
import rgb200

n = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
ndev = int(sys.argv[2]) if len(sys.argv) > 2 else torch.cuda.device_count()
lib = rgb200.load_library()
ctx = rgb200.Context(0)
seed = 0x5EED0004
d = torch.empty((n, n, n), dtype=torch.uint8, device="cuda:0"); ctx.synth(4, d, n, seed)
h_img = torch.empty((n, n, n), dtype=torch.uint8, pin_memory=True); h_img.copy_(d)
seeds = rgb200.synth_seeds(4, n, seed)
want = torch.empty_like(d); ctx.grow(d, torch.from_numpy(seeds).cuda(), 1, want); torch.cuda.synchronize()
want_cnt = int(want.sum(dtype=torch.int64)); h_want = want.cpu()
del d, want; ctx.close(); torch.cuda.empty_cache()
h_out = torch.zeros((n, n, n), dtype=torch.uint8, pin_memory=True)
dev = (ctypes.c_int * ndev)(*range(ndev)); ms = (ctypes.c_float * 4)()
res = {"n": n, "ndev": ndev, "calls": []}
for i, flags in enumerate([1, 1, 1, 0, 0]):
    if flags:
        h_out.zero_()
    t0 = time.perf_counter()
    rc = lib.rg_grow_multi(h_img.data_ptr(), n, n, n, seeds.ctypes.data, len(seeds), 1, h_out.data_ptr(), dev, ndev, flags, ms)
    dt = time.perf_counter() - t0
    ok = rc == 0 and bool(torch.equal(h_out, h_want))
    res["calls"].append({"call": i, "cold": i == 0, "mask_zeroed": bool(flags), "ms": dt * 1e3, "phase_ms": [round(v, 2) for v in ms], "bit_exact_vs_single_gpu": ok,
                         "gvoxels_per_s": n ** 3 / dt / 1e9})
print(json.dumps(res))
