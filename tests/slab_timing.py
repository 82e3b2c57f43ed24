import os, sys, json
sys.path.insert(0, "/root/repo/parallel-c-programs_b200")
import torch, torch.distributed as dist, rgb200
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local); dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
lib = rgb200.load_library(); ctx = rgb200.Context(local)
n = 2048; kind = 4; seed = 0x5EED0004
g = rgb200.SlabGrower(lib, ctx.handle, n, n, n, rank, world, dev)
plane = n * n
lo = g.z0 - (1 if g.has_lower else 0); hi = g.z1 + (1 if g.has_upper else 0)
ctx.synth(kind, g.buf[(lo - g.z0 + 1) * plane:], n, seed, z_begin=lo, z_end=hi)
g.set_seeds(rgb200.synth_seeds(kind, n, seed))
for _ in range(4): g.grow(1)
torch.cuda.synchronize(); dist.barrier()
os.environ["RG_SLAB_TIMING"] = "1"
g.grow(1)
rep = g.timing_report()
if rank in (0, world // 2): print(rank, json.dumps({k: {a: round(b, 3) for a, b in v.items()} for k, v in rep.items()}), flush=True)
dist.barrier(); dist.destroy_process_group()
