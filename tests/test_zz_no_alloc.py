"""rg_grow_device_noalloc: the grow for callers that may not allocate device memory while running
(a host whose own pool owns the card, a latency-critical loop that cannot take cudaMalloc's device-wide
synchronisation or a late out-of-memory error).  Its scratch is a fixed 64 MiB array inside the library
instead of a context's workspace, so the things to check are that it computes exactly what the
context-based entry point computes (and what the CPU oracle computes) -- on two calls in a row, because the
second call finds the first call's leftovers in the shared array --, that the amount of free device memory
does not move across a call, that its self-check can be read (rg_noalloc_status), and that it refuses sizes
the array cannot hold.

Collected last on purpose (file name): a failure here cannot hide the results of the other tests."""
import ctypes
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _setup():
    import torch
    if not torch.cuda.is_available():
        pytest.fail("needs a CUDA device; the engine has no CPU path")
    import rgb200
    lib = rgb200.load_library()
    torch.cuda.set_device(0)
    return torch, rgb200, lib


def _noalloc(torch, lib, d_img, d_seeds, T):
    out = torch.full_like(d_img, 5)
    nz, ny, nx = (1,) + tuple(d_img.shape) if d_img.dim() == 2 else tuple(d_img.shape)
    k = 0 if d_seeds is None else int(d_seeds.shape[0])
    rc = lib.rg_grow_device_noalloc(d_img.data_ptr(), nx, ny, nz, d_seeds.data_ptr() if k else None, k, T, out.data_ptr(),
                                    ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0, (rc, lib.rg_last_error())
    return out


def _status(torch, lib):
    st = (ctypes.c_uint32 * 4)()
    rc = lib.rg_noalloc_status(st, ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    return rc, list(st)


def test_noalloc_matches_context_entry_point_twice(oracle):
    torch, rgb200, lib = _setup()
    rng = np.random.default_rng(2024)
    ctx = rgb200.Context(0)
    # call 1: a volume with a solid (uniform tiles, mixed tiles, noise tiles); call 2: different size, seeds, threshold
    vol = rng.integers(0, 20, size=(96, 130, 192), dtype=np.uint8); vol[10:80, 20:120, 30:170] = 60
    img2 = rng.integers(0, 9, size=(700, 640), dtype=np.uint8)
    cases = [(vol, np.array([[100, 60, 40], [0, 0, 0], [191, 129, 95]], np.int32), 1),
             (img2, np.array([[5, 5, 0], [600, 650, 0], [320, 10, 0], [-4, 2, 0]], np.int32), 2)]
    assert lib.rg_workspace_bytes(192, 130, 96, 3) <= lib.rg_noalloc_scratch_bytes() == 64 << 20
    for img, seeds, T in cases:
        d_img = torch.from_numpy(img).cuda(); d_seeds = torch.from_numpy(seeds).cuda()
        want = torch.empty_like(d_img)
        ctx.grow(d_img, d_seeds, T, want)                          # the existing, context-based entry point
        _noalloc(torch, lib, d_img, d_seeds, T)                    # (the very first call on a device looks up the array's address)
        out = torch.full_like(d_img, 5)                            # the caller's output buffer exists before the measured call
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info()[0]
        nz, ny, nx = (1,) + tuple(d_img.shape) if d_img.dim() == 2 else tuple(d_img.shape)
        rc = lib.rg_grow_device_noalloc(d_img.data_ptr(), nx, ny, nz, d_seeds.data_ptr(), len(seeds), T, out.data_ptr(),
                                        ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
        torch.cuda.synchronize()
        assert rc == 0 and torch.cuda.mem_get_info()[0] == free_before      # the call took no device memory
        got = out
        assert ctx.status()["status"] == 0
        rc, st = _status(torch, lib)
        assert rc == 0 and st[1] == 0 and st[0] > 0                # self-check of the no-allocation path: readable and clean
        assert torch.equal(got, want)                              # integer path: bit-exact
        assert (got.cpu().numpy() == oracle.grow(img, seeds, T)).all()   # ... and equal to the CPU oracle, not just to the other CUDA path
        assert int(want.sum()) > 0
    # the first volume again: nothing of call 2 may leak into it
    d_img = torch.from_numpy(vol).cuda(); d_seeds = torch.from_numpy(cases[0][1]).cuda()
    want = torch.empty_like(d_img); ctx.grow(d_img, d_seeds, 1, want)
    assert torch.equal(_noalloc(torch, lib, d_img, d_seeds, 1), want)
    assert not bool(_noalloc(torch, lib, d_img, None, 1).any())    # no seeds: empty region, not stale data
    ctx.close()


def test_noalloc_rejects_what_its_scratch_cannot_hold():
    torch, rgb200, lib = _setup()
    big = torch.zeros((8, 8, 8), dtype=torch.uint8, device="cuda")
    seeds = torch.zeros((1, 3), dtype=torch.int32, device="cuda")
    out = torch.zeros_like(big)
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    # sizes are only arithmetic here: the call must refuse before touching the (too small) buffers
    rc = lib.rg_grow_device_noalloc(big.data_ptr(), 1024, 1024, 1024, seeds.data_ptr(), 1, 1, out.data_ptr(), st)
    assert rc == -5 and b"64 MiB" in lib.rg_last_error()
    rc = lib.rg_grow_device_noalloc(big.data_ptr(), 8, 8, 8, seeds.data_ptr(), 5000, 1, out.data_ptr(), st)
    assert rc == -5 and b"2048 seeds" in lib.rg_last_error()
    assert lib.rg_grow_device_noalloc(None, 8, 8, 8, seeds.data_ptr(), 1, 1, out.data_ptr(), st) == -1
    assert lib.rg_workspace_bytes(256, 256, 512, 16) <= lib.rg_noalloc_scratch_bytes()
    assert lib.rg_workspace_bytes(8192, 8192, 1, 64) <= lib.rg_noalloc_scratch_bytes()
