#!/usr/bin/env python3
"""Multi-GPU parity check (not a pytest file; run under torchrun on a box with >= 2 GPUs):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29700 tests/mgpu_check.py

Every rank labels its slab (z-slab of a volume, row slab of an image) with the CUDA engine, interface lists are all-gathered over NCCL, and the
assembled mask is compared byte for byte with the oracle on rank 0 (C4 geometry: the z-serpentine crosses
every slab interface 64 times; plus a random ragged volume)."""
import ctypes, os, sys
import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "parallel-c-programs_b200")); sys.path.insert(0, os.path.join(ROOT, "tests"))
import rgb200
from conftest import Oracle


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    lib = rgb200.load_library()
    ctx = rgb200.Context(local)
    ora = Oracle()
    cases = []
    img, seeds, T = ora.synth("c4", 256, 0x5EED0004); cases.append(("c4-256", img, seeds, T))
    rng = np.random.default_rng(5)
    img = rng.integers(0, 3, size=(61, 70, 208), dtype=np.uint8)
    seeds = np.stack([rng.integers(0, 208, 8), rng.integers(0, 70, 8), rng.integers(0, 61, 8)], 1).astype(np.int32)
    cases.append(("random-ragged", img, seeds, 1))
    img, seeds, T = ora.synth("c2", 4096, 0x5EED0002); cases.append(("c2-4096-rows", img, seeds, T))      # an image: row slabs
    img = rng.integers(0, 3, size=(1001, 1040), dtype=np.uint8)
    seeds = np.stack([rng.integers(0, 1040, 16), rng.integers(0, 1001, 16), np.zeros(16, np.int64)], 1).astype(np.int32)
    cases.append(("random-image-rows", img, seeds, 1))
    ok = True
    for name, img, seeds, T in cases:
        nz, ny, nx = ((1,) + img.shape) if img.ndim == 2 else img.shape
        na = ny if nz == 1 else nz
        g = rgb200.SlabGrower(lib, ctx.handle, nx, ny, nz, rank, world, dev)
        g.image.copy_(torch.from_numpy(img[g.z0:g.z1]).to(dev))
        g.exchange_image_halos()
        g.set_seeds(seeds)
        m = g.grow(T)                      # synchronous exchange (learns capacities)
        m = g.grow(T); m = g.validate(T)   # optimistic exchange
        assert g.optimistic_steps == 1
        torch.cuda.synchronize()
        st = ctx.status()
        parts = [torch.empty((rgb200.slab_bounds(na, r, world)[1] - rgb200.slab_bounds(na, r, world)[0],) + tuple(m.shape[1:]),
                             dtype=torch.uint8, device=dev) for r in range(world)]
        dist.all_gather(parts, m.contiguous())
        if rank == 0:
            got = torch.cat(parts).cpu().numpy()
            want = ora.grow(img, seeds, T)
            same = bool((got == want).all())
            ok = ok and same
            print(f"{name}: world={world} parity={'OK' if same else 'MISMATCH'} region={int(want.sum())} status={st}", flush=True)
    # a context on ANOTHER device than the current one: the call must run there and leave the current device alone
    other = (local + 1) % torch.cuda.device_count()
    if other != local:
        rng = np.random.default_rng(9)
        img = rng.integers(0, 3, size=(20, 40, 80), dtype=np.uint8)
        seeds = np.array([[3, 3, 3], [70, 30, 15]], np.int32)
        octx = rgb200.Context(other)
        d_img = torch.from_numpy(img).to(f"cuda:{other}"); d_seeds = torch.from_numpy(seeds).to(f"cuda:{other}")
        d_out = torch.empty_like(d_img)
        octx.grow(d_img, d_seeds, 1, d_out, stream=torch.cuda.current_stream(other).cuda_stream)
        torch.cuda.synchronize(other)
        same = bool((d_out.cpu().numpy() == ora.grow(img, seeds, 1)).all()) and torch.cuda.current_device() == local
        print(f"rank {rank}: context on cuda:{other} while current device is cuda:{local}: {'OK' if same else 'MISMATCH'}", flush=True)
        ok_t = torch.tensor([1 if same else 0], device=dev); dist.all_reduce(ok_t, op=dist.ReduceOp.MIN)
        ok = ok and bool(ok_t.item())
    dist.barrier()
    dist.destroy_process_group()
    if rank == 0 and not ok:
        sys.exit(1)


if __name__ == "__main__":
    main()
