"""Parity of the CUDA engine with the oracle, through the C ABI (libb200rg.so).  Needs a B200.

Bit-exact everywhere (integer/byte path).  Sizes the oracle finishes in seconds are compared byte
for byte; the full-size configurations of BASELINE.json are checked through size-independent
properties (analytic region sizes, idempotence, slab-count independence)."""
import ctypes, hashlib
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def rg():
    import torch
    if not torch.cuda.is_available():
        pytest.fail("the -m gpu tests need a CUDA device; the engine has no CPU path")
    import rgb200
    rgb200.load_library()
    return rgb200


@pytest.fixture(scope="module")
def ctx(rg):
    c = rg.Context(0)
    yield c
    c.close()


def dev_grow(rg, ctx, img, seeds, T):
    import torch
    d_img = torch.from_numpy(np.ascontiguousarray(img)).cuda()
    s = np.ascontiguousarray(np.asarray(seeds, np.int32).reshape(-1, 3))
    d_seeds = torch.from_numpy(s).cuda() if len(s) else None
    out = torch.full_like(d_img, 9)
    ctx.grow(d_img, d_seeds, T, out)
    torch.cuda.synchronize()
    st = ctx.status()
    assert st["status"] == 0
    return out.cpu().numpy()


# ---------------------------------------------------------------- reference fixtures (C0, C1)
@pytest.mark.parametrize("k", [1, 2, 3, 4])
def test_region_c_pictures(rg, oracle, golden, k):
    """rg_grow_region_2d == unmodified region.c (4 corner seeds, |d|<2) incl. write_image's masking"""
    img = golden["arrays"][f"pic{k}"]
    mask, out = rg.grow_region_2d(img, apply=True)
    assert set(np.unique(mask)) <= {0, 1}
    assert hashlib.sha256(out.tobytes()).hexdigest() == golden["pics"][f"pic{k}"]["out_sha256"]
    assert (mask == oracle.grow(img, oracle.seeds_2d(512, 512), 2)).all()
    mask2, img2 = rg.grow_region_2d(img, apply=False)
    assert (mask2 == mask).all() and (img2 == img).all()


def test_c0_single_seed(rg, oracle, golden):
    img = golden["arrays"]["pic1"]
    m = rg.grow(img, [[4, 4, 0]], 2)
    assert int(m.sum()) == 24056 and (m == oracle.grow(img, [[4, 4, 0]], 2)).all()


def test_c1_reference_volume_all_three_names(rg, oracle, golden):
    """grow_region_gpu / grow_region_gpu_shared / grow_region_serial on the reference's own 512^3 volume"""
    vol = oracle.create_data_ref()
    assert oracle.fnv(vol) == golden["c1"]["volume_fnv1a64"]
    m = rg.grow_region_gpu_shared(vol)
    assert int(m.sum()) == 2197899
    assert oracle.fnv(m) == golden["c1"]["mask_fnv1a64"]
    assert set(np.unique(m)) == {0, 1}                         # never a 2 (raycast.cu:689-692 can leave them)
    assert (rg.grow_region_gpu(vol) == m).all()
    assert (rg._grow_512("grow_region_serial", vol) == m).all()


# ---------------------------------------------------------------- random grids, ragged sizes, edge cases
SHAPES3 = [(1, 1, 1), (3, 5, 7), (17, 33, 65), (16, 32, 64), (33, 65, 130), (40, 31, 200), (2, 40, 129), (70, 9, 16), (5, 100, 48)]


@pytest.mark.parametrize("shape", SHAPES3)
@pytest.mark.parametrize("levels,T", [(3, 1), (7, 2), (2, 1)])
def test_random_volumes(rg, ctx, oracle, shape, levels, T):
    rng = np.random.default_rng(abs(hash((shape, levels, T))) % 2**32)
    img = rng.integers(0, levels, size=shape, dtype=np.uint8)
    nz, ny, nx = shape
    seeds = np.stack([rng.integers(-1, nx + 1, 8), rng.integers(-1, ny + 1, 8), rng.integers(-1, nz + 1, 8)], 1)
    seeds[0] = 0
    want = oracle.grow(img, seeds, T)
    assert (dev_grow(rg, ctx, img, seeds, T) == want).all()
    assert (rg.grow(img, seeds, T) == want).all()              # host-pointer entry point


@pytest.mark.parametrize("shape", [(1, 1), (7, 5), (300, 200), (256, 128), (257, 129), (513, 700), (31, 1000), (1000, 17)])
@pytest.mark.parametrize("levels,T", [(3, 1), (9, 2)])
def test_random_images(rg, ctx, oracle, shape, levels, T):
    rng = np.random.default_rng(abs(hash((shape, levels, T))) % 2**32)
    img = rng.integers(0, levels, size=shape, dtype=np.uint8)
    ny, nx = shape
    seeds = np.stack([rng.integers(-1, nx + 1, 8), rng.integers(-1, ny + 1, 8), np.zeros(8, np.int64)], 1)
    seeds[0] = 0
    assert (dev_grow(rg, ctx, img, seeds, T) == oracle.grow(img, seeds, T)).all()


@pytest.mark.parametrize("T", [-5, 0, 1, 2, 128, 255, 256, 1000])
def test_threshold_range(rg, ctx, oracle, T):
    rng = np.random.default_rng(1000 + T)
    img = rng.integers(0, 256, size=(20, 40, 96), dtype=np.uint8)
    seeds = [[0, 0, 0], [95, 39, 19], [50, 20, 10]]
    assert (dev_grow(rg, ctx, img, seeds, T) == oracle.grow(img, seeds, T)).all()


def test_seed_edge_cases(rg, ctx, oracle):
    img = np.zeros((20, 40, 96), np.uint8); img[:, :, 48] = 9   # a wall splits the volume
    assert not dev_grow(rg, ctx, img, np.zeros((0, 3), np.int32), 1).any()          # no seeds
    assert not dev_grow(rg, ctx, img, [[-1, 0, 0], [96, 0, 0], [0, 40, 0], [0, 0, 20]], 1).any()   # all out of range
    dup = [[3, 3, 3]] * 5 + [[95, 39, 19]]                                              # duplicates, both halves
    got = dev_grow(rg, ctx, img, dup, 1)
    assert (got == oracle.grow(img, dup, 1)).all() and got.sum() == 20 * 40 * 95
    wall = dev_grow(rg, ctx, img, [[48, 0, 0]], 1)
    assert wall.sum() == 20 * 40 and wall[:, :, 48].all()


@pytest.mark.parametrize("shape", [(24, 48, 128), (300, 272)])
@pytest.mark.parametrize("off_img,off_mask", [(1, 0), (0, 3), (5, 9), (16, 32)])
def test_misaligned_device_pointers(rg, ctx, oracle, shape, off_img, off_mask):
    """rows that are a multiple of 16 bytes but buffers that do not start on a 16-byte boundary (a view into a larger
    allocation): the engine must leave its vector / TMA paths on its own"""
    import torch
    rng = np.random.default_rng(off_img * 100 + off_mask + len(shape))
    img = rng.integers(0, 3, size=shape, dtype=np.uint8)
    nvox = img.size
    dims = shape[::-1] if len(shape) == 3 else (shape[1], shape[0], 1)
    seeds = np.stack([rng.integers(0, d, 6) for d in dims], 1).astype(np.int32)
    big_i = torch.zeros(nvox + 64, dtype=torch.uint8, device="cuda")
    big_m = torch.full((nvox + 64,), 7, dtype=torch.uint8, device="cuda")
    d_img = big_i[off_img:off_img + nvox].view(*shape); d_img.copy_(torch.from_numpy(img))
    d_out = big_m[off_mask:off_mask + nvox].view(*shape)
    assert d_img.data_ptr() % 16 == off_img % 16 and d_out.data_ptr() % 16 == off_mask % 16
    ctx.grow(d_img, torch.from_numpy(seeds).cuda(), 1, d_out)
    torch.cuda.synchronize()
    assert ctx.status()["status"] == 0
    assert (d_out.cpu().numpy() == oracle.grow(img, seeds, 1)).all()
    assert bool((big_m[:off_mask] == 7).all()) and bool((big_m[off_mask + nvox:] == 7).all())      # nothing written outside the mask


def test_one_context_many_shapes_and_a_side_stream(rg, oracle):
    """a context is reused across sizes (its scratch grows), dimensionalities and streams"""
    import torch
    c = rg.Context(0)
    side = torch.cuda.Stream()
    rng = np.random.default_rng(77)
    for shape in [(8, 16, 64), (200, 144), (40, 70, 130), (1, 1, 1), (64, 64), (33, 65, 200), (17, 40, 96)]:
        img = rng.integers(0, 3, size=shape, dtype=np.uint8)
        dims = shape[::-1] if len(shape) == 3 else (shape[1], shape[0], 1)
        seeds = np.stack([rng.integers(0, d, 4) for d in dims], 1).astype(np.int32)
        d_img = torch.from_numpy(img).cuda(); d_seeds = torch.from_numpy(seeds).cuda(); out = torch.full_like(d_img, 9)
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            c.grow(d_img, d_seeds, 1, out)                          # rgb200 launches on torch's current stream
        side.synchronize()
        assert c.status()["status"] == 0
        assert (out.cpu().numpy() == oracle.grow(img, seeds, 1)).all(), shape
    c.close()


def test_seed_set_algebra_at_benchmark_scale(rg, ctx):
    """properties that need no oracle (C4 geometry at 1024^3, 1 Gvoxel): the region of a union of seed sets is the union
    of the regions; every in-range seed is in its region; growing again from a voxel of a region component returns
    that component; a second call on the same inputs returns the same bytes"""
    import torch
    n = 1024
    img = torch.empty((n, n, n), dtype=torch.uint8, device="cuda")
    ctx.synth(4, img, n, 0x5EED0004)
    seeds = rg.synth_seeds(4, n, 0x5EED0004)
    d_all = torch.from_numpy(seeds).cuda()
    a, b = seeds[:5].copy(), seeds[5:].copy()
    m_all = torch.empty_like(img); m_a = torch.empty_like(img); m_b = torch.empty_like(img); again = torch.empty_like(img)
    ctx.grow(img, d_all, 1, m_all)
    ctx.grow(img, torch.from_numpy(a).cuda(), 1, m_a)
    ctx.grow(img, torch.from_numpy(b).cuda(), 1, m_b)
    ctx.grow(img, d_all, 1, again)
    torch.cuda.synchronize()
    assert ctx.status()["status"] == 0
    assert torch.equal(m_all, m_a | m_b) and torch.equal(m_all, again)
    assert bool((m_all[d_all[:, 2].long(), d_all[:, 1].long(), d_all[:, 0].long()] == 1).all())
    # the component of the first seed (the big sphere), regrown from some other voxel of it
    ctx.grow(img, torch.from_numpy(seeds[:1].copy()).cuda(), 1, m_a)
    torch.cuda.synchronize()
    nz = torch.nonzero(m_a[n // 5])                                  # a plane that cuts the sphere
    assert len(nz) > 1000
    y, x = [int(v) for v in nz[len(nz) // 2]]
    other = torch.tensor([[x, y, n // 5]], dtype=torch.int32, device="cuda")
    ctx.grow(img, other, 1, m_b)
    torch.cuda.synchronize()
    assert torch.equal(m_a, m_b)


def test_many_seeds_radix_sort_path(rg, ctx, oracle):
    rng = np.random.default_rng(3)
    img = rng.integers(0, 4, size=(40, 70, 130), dtype=np.uint8)
    seeds = np.stack([rng.integers(0, 130, 5000), rng.integers(0, 70, 5000), rng.integers(0, 40, 5000)], 1)
    assert (dev_grow(rg, ctx, img, seeds, 1) == oracle.grow(img, seeds, 1)).all()


def test_all_similar_and_none_similar(rg, ctx, oracle):
    img = np.full((33, 65, 130), 7, np.uint8)
    assert dev_grow(rg, ctx, img, [[129, 64, 32]], 1).all()
    chk = (np.indices((33, 65, 130)).sum(0) % 2 * 50).astype(np.uint8)              # checkerboard: no edges at T<=50
    got = dev_grow(rg, ctx, chk, [[5, 5, 5], [0, 0, 0]], 50)
    assert got.sum() == 2 and got[5, 5, 5] == 1 and got[0, 0, 0] == 1


def test_tile_face_corridors(rg, ctx, oracle):
    """1-voxel corridors that cross tile faces, edges and corners of the 64x32x16 tiling"""
    img = np.zeros((48, 96, 192), np.uint8)
    path = np.zeros_like(img, bool)
    path[15, 31, :] = True; path[15:18, 31, 191] = True; path[17, 31:65, 191] = True     # along x, then z, then y
    path[17, 64, 60:192] = True; path[16:48, 64, 63] = True; path[47, 30:65, 63] = True
    path[0:48, 32, 64] = True
    img[path] = 100
    seeds = [[0, 31, 15]]
    got = dev_grow(rg, ctx, img, seeds, 1)
    assert (got == oracle.grow(img, seeds, 1)).all()
    assert got.sum() > 400


# ---------------------------------------------------------------- synthetic generators and benchmark shapes
@pytest.mark.parametrize("kind,name,n", [(1, "c1h", 96), (2, "c2", 320), (3, "c3", 80), (4, "c4", 128), (3, "c3", 33), (1, "c1h", 50)])
def test_device_generators_match_cpu_definitions(rg, ctx, oracle, kind, name, n):
    import torch
    want, seeds, T = oracle.synth(name, n, 0x5EED0000 + kind)
    out = torch.empty(want.shape, dtype=torch.uint8, device="cuda")
    ctx.synth(kind, out, n, 0x5EED0000 + kind)
    torch.cuda.synchronize()
    assert (out.cpu().numpy() == want).all()
    assert (rg.synth_seeds(kind, n, 0x5EED0000 + kind) == seeds).all()
    assert (dev_grow(rg, ctx, want, seeds, T) == oracle.grow(want, seeds, T)).all()


@pytest.mark.parametrize("kind,name,n", [(1, "c1h", 256), (3, "c3", 192), (4, "c4", 256), (2, "c2", 4096)])
def test_benchmark_shapes_mid_size(rg, ctx, oracle, kind, name, n):
    img, seeds, T = oracle.synth(name, n, 0x5EED0000 + kind)
    assert (dev_grow(rg, ctx, img, seeds, T) == oracle.grow(img, seeds, T)).all()


def test_c3_full_size_properties(rg, ctx, oracle):
    """C3 at BASELINE.json's full size (1024^3): region size is known in closed form; the mask equals
    'voxel is corridor'; growing the mask itself (T=1) from the same seed is idempotent."""
    import torch
    n = 1024
    img = torch.empty((n, n, n), dtype=torch.uint8, device="cuda")
    ctx.synth(3, img, n)
    seeds = torch.tensor([[0, 0, 0]], dtype=torch.int32, device="cuda")
    out = torch.empty_like(img)
    ctx.grow(img, seeds, 1, out)
    torch.cuda.synchronize()
    assert ctx.status()["status"] == 0
    assert int(out.sum(dtype=torch.int64)) == oracle.lib.oracle_c3_count(n)
    assert bool(((img == 200) == (out == 1)).all())
    out2 = torch.empty_like(img)
    ctx.grow(out, seeds, 1, out2)                                   # mask as an image: region = its 1-component of the seed
    torch.cuda.synchronize()
    assert bool((out2 == out).all())
    # a far-away seed reaches the same single region
    seeds2 = torch.tensor([[n - 2, n - 2, n - 2]], dtype=torch.int32, device="cuda")
    ctx.grow(img, seeds2, 1, out2)
    torch.cuda.synchronize()
    assert bool((out2 == out).all())


def test_c2_full_size_properties(rg, ctx, oracle):
    """C2 at full size (16384^2, 64 seeds): every solid's area is counted on the device and compared with
    the mask restricted to solids; noise seeds contribute nothing outside; sub-image parity vs oracle."""
    import torch
    n = 16384
    img = torch.empty((n, n), dtype=torch.uint8, device="cuda")
    ctx.synth(2, img, n, 0x5EED0002)
    seeds_np = rg.synth_seeds(2, n, 0x5EED0002)
    seeds = torch.from_numpy(seeds_np).cuda()
    out = torch.empty_like(img)
    ctx.grow(img, seeds, 2, out)
    torch.cuda.synchronize()
    assert ctx.status()["status"] == 0
    assert bool(((img >= 40) == (out == 1)).all())                  # all 64 solids hold a seed; noise (<20) never joins
    # first cell row x first two cells against the oracle (the bar across cells + its noise surroundings)
    sub = img[:2048, :4096].cpu().numpy()
    sub_seeds = [s for s in seeds_np.tolist() if s[0] < 4096 and s[1] < 2048]
    assert (oracle.grow(sub, sub_seeds, 2) == out[:2048, :4096].cpu().numpy()).all()


# ---------------------------------------------------------------- z-slab path on one device ("fake cluster")
def slab_grow_one_device(rg, img, seeds, T, G, return_parts=False):
    """run the multi-rank protocol for G logical ranks on one GPU by calling the C ABI phases in turn.
    img: numpy array, or a CUDA tensor (then nothing passes through the host; use return_parts=True to get the
    per-rank mask tensors instead of one assembled array)"""
    import torch
    lib = rg.load_library()
    on_device = hasattr(img, "is_cuda")
    rows = img.ndim == 2                                            # an image is cut into row slabs, a volume into z-slabs
    nz, ny, nx = ((1,) + img.shape) if rows else img.shape
    plane = nx if rows else nx * ny
    axis = 1 if rows else 2
    ranks = []
    for r in range(G):
        z0, z1 = rg.slab_bounds(ny if rows else nz, r, G)
        h = ctypes.c_void_p(); assert lib.rg_ctx_create(0, ctypes.byref(h)) == 0
        buf = torch.zeros((z1 - z0 + 2) * plane, dtype=torch.uint8, device="cuda")
        lo, hi = max(z0 - 1, 0), min(z1 + 1, ny if rows else nz)
        buf[(lo - z0 + 1) * plane:(hi - z0 + 1) * plane] = img[lo:hi].reshape(-1) if on_device else torch.from_numpy(img[lo:hi].reshape(-1)).cuda()
        mask = torch.empty((z1 - z0) * plane, dtype=torch.uint8, device="cuda")
        s = np.asarray(seeds, np.int32).reshape(-1, 3)
        s = s[(s[:, axis] >= z0) & (s[:, axis] < z1)].copy()
        s[:, axis] -= z0
        d_s = torch.from_numpy(np.ascontiguousarray(s)).cuda() if len(s) else None
        cap = int(lib.rg_slab_interface_capacity(nx, ny, 1 if rows else z1 - z0))
        lists = torch.zeros((2, cap), dtype=torch.int32, device="cuda")
        counts = torch.zeros(4, dtype=torch.int32, device="cuda")
        ny_l, nz_l = (z1 - z0, 1) if rows else (ny, z1 - z0)
        assert lib.rg_slab_label(h, buf.data_ptr() + plane, nx, ny_l, nz_l, int(r > 0), int(r < G - 1),
                                 d_s.data_ptr() if d_s is not None else None, len(s), T, mask.data_ptr(), None) == 0
        if r > 0:
            assert lib.rg_slab_export(h, 0, lists[0].data_ptr(), counts.data_ptr(), None) == 0
        if r < G - 1:
            assert lib.rg_slab_export(h, 1, lists[1].data_ptr(), counts.data_ptr(), None) == 0
        ranks.append(dict(h=h, buf=buf, mask=mask, lists=lists, counts=counts, z0=z0, z1=z1, seeds=d_s))
    torch.cuda.synchronize()
    cnt = np.stack([rk["counts"].cpu().numpy() for rk in ranks]).astype(np.int64)
    for r in range(G - 1):
        assert cnt[r, 1] == cnt[r + 1, 0]
    off = np.ones(G + 1, np.int64); off[1:] = 1 + np.cumsum(cnt[:, 2])
    for me in range(G):                                             # every rank resolves redundantly
        xpar = torch.empty(int(off[-1]) + 1, dtype=torch.int32, device="cuda")
        h = ranks[me]["h"]
        for r in range(G):
            for side in (0, 1):
                if cnt[r, side]:
                    assert lib.rg_slab_resolve_init(h, xpar.data_ptr(), ranks[r]["lists"][side].data_ptr(), int(cnt[r, side]), int(off[r]), None) == 0
        for r in range(G - 1):
            if cnt[r, 1]:
                assert lib.rg_slab_resolve_union(h, xpar.data_ptr(), ranks[r]["lists"][1].data_ptr(), int(off[r]),
                                                 ranks[r + 1]["lists"][0].data_ptr(), int(off[r + 1]), int(cnt[r, 1]), None) == 0
        for side in (0, 1):
            if cnt[me, side]:
                assert lib.rg_slab_resolve_apply(h, xpar.data_ptr(), ranks[me]["lists"][side].data_ptr(), int(cnt[me, side]), int(off[me]), None) == 0
        assert lib.rg_slab_finish(h, None) == 0
        torch.cuda.synchronize()
    for rk in ranks:
        st = (ctypes.c_uint32 * 4)()
        assert lib.rg_ctx_status(rk["h"], st) == 0
        lib.rg_ctx_destroy(rk["h"])
    if return_parts:
        return [(rk["z0"], rk["z1"], rk["mask"]) for rk in ranks]
    return np.concatenate([rk["mask"].cpu().numpy() for rk in ranks]).reshape(tuple(img.shape))


@pytest.mark.parametrize("G", [2, 4, 8])
def test_slabs_equal_single(rg, ctx, oracle, G):
    img, seeds, T = oracle.synth("c4", 128, 0x5EED0004)             # z-serpentine: 64 crossings per interface
    want = oracle.grow(img, seeds, T)
    assert (slab_grow_one_device(rg, img, seeds, T, G) == want).all()


@pytest.mark.parametrize("G", [2, 3, 5])
def test_slabs_random_ragged(rg, ctx, oracle, G):
    rng = np.random.default_rng(40 + G)
    img = rng.integers(0, 3, size=(37, 45, 150), dtype=np.uint8)
    seeds = np.stack([rng.integers(0, 150, 6), rng.integers(0, 45, 6), rng.integers(0, 37, 6)], 1)
    assert (slab_grow_one_device(rg, img, seeds, 1, G) == oracle.grow(img, seeds, 1)).all()


def test_node_budget_overflow_is_reported(rg, oracle, monkeypatch):
    """volumes beyond ~17 Gvoxel (and RG_NODE_BUDGET runs) size the node space for an average number of nodes per
    tile; an input that needs more must fail loudly through the status word, and succeed again with room"""
    import torch
    # every (x, y) column is its own component and runs through all planes: 2048 nodes per tile, twice the fixed block of
    # node ids a tile owns, so every tile takes the rest from the shared space that RG_NODE_BUDGET shrinks
    z, y, x = np.meshgrid(np.arange(64), np.arange(96), np.arange(192), indexing="ij")
    img = (((x + 2 * y) % 5) * 40).astype(np.uint8)
    seeds, T = np.array([[0, 0, 0], [100, 50, 63], [191, 95, 10]], np.int32), 1
    want = oracle.grow(img, seeds, T)
    assert int(want.sum()) == 3 * 64
    d_img = torch.from_numpy(img).cuda(); d_seeds = torch.from_numpy(seeds).cuda(); out = torch.empty_like(d_img)
    monkeypatch.setenv("RG_NODE_BUDGET", "4")
    c1 = rg.Context(0)
    c1.grow(d_img, d_seeds, T, out)
    torch.cuda.synchronize()
    st = (ctypes.c_uint32 * 4)()
    assert c1.lib.rg_ctx_status(c1.handle, st) != 0 and (st[1] & 1)
    assert b"node ids" in c1.lib.rg_last_error()
    monkeypatch.delenv("RG_NODE_BUDGET")
    c2 = rg.Context(0)
    c2.grow(d_img, d_seeds, T, out)
    torch.cuda.synchronize()
    assert c2.status()["status"] == 0 and (out.cpu().numpy() == want).all()


def test_volume_beyond_the_worst_case_node_space(rg, oracle):
    """2048 x 2048 x 4608 (19.3 Gvoxel, 2.25 x the benchmark volume): the worst-case node count no longer fits 32 bits,
    so the workspace is sized by budget.  Checked against the same volume grown as three z-slabs (each small enough for
    worst-case sizing) joined by the interface exchange -- a different code path over the same data."""
    import torch
    free, _ = torch.cuda.mem_get_info()
    if free < 150e9:
        pytest.skip("needs ~135 GB of device memory")
    n, nz = 2048, 4608
    lib = rg.load_library()
    ctx = rg.Context(0)
    img = torch.empty((nz, n, n), dtype=torch.uint8, device="cuda")
    for z0 in range(0, nz, n):                                     # copies of the C4 cube stacked along z
        z1 = min(z0 + n, nz)
        ctx.synth(4, img[z0:z1], n, 0x5EED0004, z_begin=0, z_end=z1 - z0)
    base = rg.synth_seeds(4, n, 0x5EED0004)
    seeds = np.concatenate([base + [0, 0, k * n] for k in range(3)]).astype(np.int32)
    seeds = seeds[seeds[:, 2] < nz]
    d_seeds = torch.from_numpy(np.ascontiguousarray(seeds)).cuda()
    assert lib.rg_workspace_bytes(n, n, nz, len(seeds)) < 2.5 * n * n * nz      # budgeted, not 2 B/voxel + the worst-case nodes
    out = torch.empty_like(img)
    ctx.grow(img, d_seeds, 1, out)
    torch.cuda.synchronize()
    st = ctx.status()
    assert st["status"] == 0, st
    region = sum(int(torch.count_nonzero(out[z:z + 256])) for z in range(0, nz, 256))     # no full-size temporaries
    assert region > 2 * 572079655                                    # at least the solids of two whole copies
    ctx.close()
    torch.cuda.empty_cache()
    for z0, z1, m in slab_grow_one_device(rg, img, seeds, 1, 3, return_parts=True):
        assert torch.equal(m.view(z1 - z0, n, n), out[z0:z1]), (z0, z1)


@pytest.mark.parametrize("G", [2, 3, 8])
@pytest.mark.parametrize("k", [1, 2, 3, 4])
def test_row_slabs_reference_pictures(rg, ctx, oracle, golden, G, k):
    """the reference's own use case (region.c: one picture, several ranks): 512 rows over G ranks do not divide into
    126-row tiles, so every rank has a ragged last tile row whose YP faces form the interface"""
    img = golden["arrays"][f"pic{k}"]
    seeds = oracle.seeds_2d(512, 512)
    got = slab_grow_one_device(rg, img, seeds, 2, G)
    assert hashlib.sha256((img * (got == 0)).tobytes()).hexdigest() == golden["pics"][f"pic{k}"]["out_sha256"]


@pytest.mark.parametrize("G,shape", [(2, (300, 272)), (5, (1000, 1040)), (4, (37, 48)), (3, (255, 130))])
def test_row_slabs_random(rg, ctx, oracle, G, shape):
    """aligned widths (TMA staging) and unaligned ones (generic staging), rows per rank below / above one tile"""
    rng = np.random.default_rng(60 + G)
    img = rng.integers(0, 3, size=shape, dtype=np.uint8)
    h, w = shape
    seeds = np.stack([rng.integers(0, w, 12), rng.integers(0, h, 12), np.zeros(12, np.int64)], 1)
    assert (slab_grow_one_device(rg, img, seeds, 1, G) == oracle.grow(img, seeds, 1)).all()


def test_row_slabs_c2_snake(rg, ctx, oracle):
    """C2 geometry (discs and bars on a lattice, some touching) at 2048^2 over 8 row slabs, and a snake that crosses
    every interface many times"""
    img, seeds, T = oracle.synth("c2", 2048, 0x5EED0002)
    assert (slab_grow_one_device(rg, img, seeds, T, 8) == oracle.grow(img, seeds, T)).all()
    h, w = 512, 768
    img = np.zeros((h, w), np.uint8)
    img[:, 1::4] = 200
    img[0, 1::8] = 0; img[h - 1, 5::8] = 0
    for k, x in enumerate(range(1, w - 4, 4)):
        img[h - 1 if k % 2 == 0 else 0, x:x + 5] = 200
    want = oracle.grow(img, [[1, 1, 0]], 1)
    assert want.sum() > 0.2 * h * w
    assert (slab_grow_one_device(rg, img, [[1, 1, 0]], 1, 7) == want).all()


def test_apply_mask_device(rg, ctx, oracle, golden):
    import torch
    img = golden["arrays"]["pic3"]
    d = torch.from_numpy(img.copy()).cuda()
    seeds = torch.from_numpy(oracle.seeds_2d(512, 512)).cuda()
    m = torch.empty_like(d)
    ctx.grow(d, seeds, 2, m)
    ctx.apply_mask(d, m)
    torch.cuda.synchronize()
    assert hashlib.sha256(d.cpu().numpy().tobytes()).hexdigest() == golden["pics"]["pic3"]["out_sha256"]


@pytest.mark.parametrize("kind,n", [("c4", 768), ("c3", 512)])
def test_bench_sample_sizes_vs_oracle(rg, ctx, oracle, kind, n):
    """the sizes bench.py's CPU legs grow anyway (C4 geometry at 768^3, C3 at 512^3), byte for byte against the oracle --
    both through the default path and with every non-uniform tile forced down the dense labelling path"""
    img, seeds, T = oracle.synth(kind, n, 0x5EED0000 + (4 if kind == "c4" else 3))
    want = oracle.grow(img, seeds, T)
    got = dev_grow(rg, ctx, img, seeds, T)
    assert (got == want).all() and int(want.sum()) > 1000


def _host_stats(lib):
    st = (ctypes.c_int64 * 4)()
    assert lib.rg_host_stats(st) == 0
    return list(st)


def test_sparse_output_and_staged_pageable_copies(rg, oracle):
    """rg_grow_sparse copies back only the boxes of tiles that hold region voxels into a mask the caller has zeroed, and a
    pageable (numpy / malloc) buffer of 32 MiB or more travels through the library's pinned bounce buffers on its copy threads:
    same bytes as the oracle either way, ragged sizes included"""
    lib = rg.load_library()
    rng = np.random.default_rng(31)
    vol = rng.integers(0, 20, size=(300, 352, 336), dtype=np.uint8)            # 35.5 MB: above the staging threshold
    vol[40:200, 100:300, 50:250] = 90
    vol[250:290, 10:60, 300:336] = 91
    seeds = np.array([[100, 150, 100], [320, 30, 270], [0, 0, 0]], np.int32)
    want = oracle.grow(vol, seeds, 1)
    # full copy, pageable in and out
    got = rg.grow(vol, seeds, 1)
    st = _host_stats(lib)
    assert (got == want).all() and st[0] == vol.size and st[1] == vol.size and st[3] == 1
    # sparse copy into a zeroed buffer
    out = np.zeros_like(vol)
    rc = lib.rg_grow_sparse(vol.ctypes.data, 336, 352, 300, seeds.ctypes.data, len(seeds), 1, out.ctypes.data)
    st = _host_stats(lib)
    assert rc == 0 and (out == want).all()
    assert 0 < st[1] < vol.size // 2 and st[2] >= 2                              # far fewer bytes came back
    # nothing outside the copied boxes is written: a buffer that is NOT zero keeps its bytes where no region tile is
    out2 = np.full_like(vol, 7)
    assert lib.rg_grow_sparse(vol.ctypes.data, 336, 352, 300, seeds.ctypes.data, len(seeds), 1, out2.ctypes.data) == 0
    assert ((out2 == want) | (out2 == 7)).all() and (out2[want == 1] == 1).all() and int((out2 == 7).sum()) > vol.size // 2
    # an image, and an empty region
    img = rng.integers(0, 6, size=(700, 1000), dtype=np.uint8); img[100:300, 200:900] = 50
    s2 = np.array([[500, 200, 0]], np.int32)
    o2 = np.zeros_like(img)
    assert lib.rg_grow_sparse(img.ctypes.data, 1000, 700, 1, s2.ctypes.data, 1, 2, o2.ctypes.data) == 0
    assert (o2 == oracle.grow(img, s2, 2)).all()
    o3 = np.zeros_like(img)
    assert lib.rg_grow_sparse(img.ctypes.data, 1000, 700, 1, None, 0, 2, o3.ctypes.data) == 0 and not o3.any()
    assert _host_stats(lib)[1] == 0


def _grow_multi(lib, img, seeds, T, devices, zeroed=False):
    img = np.ascontiguousarray(img, np.uint8)
    nz, ny, nx = (1,) + img.shape if img.ndim == 2 else img.shape
    seeds = np.ascontiguousarray(np.asarray(seeds, np.int32).reshape(-1, 3))
    out = np.zeros(img.shape, np.uint8) if zeroed else np.full(img.shape, 9, np.uint8)
    dev = (ctypes.c_int * len(devices))(*devices)
    ms = (ctypes.c_float * 4)()
    rc = lib.rg_grow_multi(img.ctypes.data, nx, ny, nz, seeds.ctypes.data if len(seeds) else None, len(seeds), T, out.ctypes.data,
                           dev, len(devices), 1 if zeroed else 0, ms)
    assert rc == 0, lib.rg_last_error()
    return out, list(ms)


@pytest.mark.parametrize("G", [1, 2, 3, 5])
def test_grow_multi_from_c_abi(rg, oracle, G):
    """rg_grow_multi: one host image scattered over G slabs (here: G slabs on the GPUs this box has, round-robin -- one GPU
    simply carries several slabs), one interface exchange by peer copies, the mask gathered into the caller's buffer:
    region.c's distribute -> grow -> gather (region.c:591-597) from one C call.  Volume, image, serpentine; plain and
    zeroed-mask (sparse copy back) forms."""
    import torch
    lib = rg.load_library()
    ndev = torch.cuda.device_count()
    devices = [g % ndev for g in range(G)]
    rng = np.random.default_rng(50 + G)
    vol = rng.integers(0, 4, size=(70, 90, 200), dtype=np.uint8); vol[5:60, 20:70, 30:170] = 40
    seeds = [[100, 40, 30], [0, 0, 0], [199, 89, 69], [50, 50, 35]]
    want = oracle.grow(vol, seeds, 1)
    got, ms = _grow_multi(lib, vol, seeds, 1, devices)
    assert (got == want).all() and int(ms[3]) == G
    st = (ctypes.c_int64 * 4)()
    assert lib.rg_multi_stats(st) == 0
    plane = vol.shape[1] * vol.shape[2]
    assert st[0] == vol.size + 2 * (G - 1) * plane + 12 * len(seeds)          # every slab with its neighbours' boundary planes, each seed to its slab
    assert st[1] == vol.size and st[2] >= G and st[3] == G                  # every byte of the mask comes back, at least one copy per slab
    got, _ = _grow_multi(lib, vol, seeds, 1, devices, zeroed=True)
    assert (got == want).all()
    assert lib.rg_multi_stats(st) == 0
    assert int(want.sum()) <= st[1] < vol.size and st[2] >= 1               # zeroed mask: only the boxes that hold region voxels
    img, s2, T2 = oracle.synth("c2", 1024, 0x5EED0002)
    got, _ = _grow_multi(lib, img, s2, T2, devices)
    assert (got == oracle.grow(img, s2, T2)).all()
    v3, s3, T3 = oracle.synth("c4", 160, 0x5EED0004)            # the z-serpentine crosses every slab interface many times
    got, _ = _grow_multi(lib, v3, s3, T3, devices, zeroed=True)
    assert (got == oracle.grow(v3, s3, T3)).all()
    got, _ = _grow_multi(lib, v3, np.zeros((0, 3), np.int32), T3, devices)
    assert not got.any()


def test_row_tiles_on_gpu(rg, ctx, oracle, monkeypatch):
    """tiles of constant rows (rg_rows.h) next to noise, uniform and ragged tiles, seeds in every kind; same bytes as the
    oracle, and as the run with the row path switched off"""
    rng = np.random.default_rng(8)
    nz, ny, nx = 50, 100, 320
    rowval = rng.integers(0, 3, size=(nz, ny, 1), dtype=np.uint8) * 70
    img = np.broadcast_to(rowval, (nz, ny, nx)).copy()
    img[:, :, 128:192] = rng.integers(0, 5, size=(nz, ny, 64))
    img[5:40, 10:90, 200:320] = 9
    seeds = [[0, 0, 0], [319, 99, 49], [150, 50, 25], [250, 50, 20], [10, 70, 33]]
    for T in (1, 71, 256):
        want = oracle.grow(img, seeds, T)
        assert (dev_grow(rg, ctx, img, seeds, T) == want).all(), T
    img3, s3, T3 = oracle.synth("c3", 256)
    want3 = oracle.grow(img3, s3, T3)
    got3 = dev_grow(rg, ctx, img3, s3, T3)
    assert (got3 == want3).all()
    # the output passes store only the ones of a tile the labelling pass has cleared; with that clear switched off they and the
    # flag pass have to write every byte (dev_grow hands out a mask full of garbage either way)
    monkeypatch.setenv("RG_NO_PREZERO", "1")
    assert (dev_grow(rg, ctx, img3, s3, T3) == want3).all()
    assert (dev_grow(rg, ctx, img, seeds, 1) == oracle.grow(img, seeds, 1)).all()


@pytest.mark.parametrize("R,Cn", [(1, 2), (2, 2), (3, 4), (2, 1)])
def test_grow_multi_grid_from_c_abi(rg, oracle, golden, R, Cn):
    """rg_grow_multi_grid: one host image over a rows x columns grid of blocks (the reference's Cartesian process grid,
    region.c:537-548; blocks dealt round-robin to the GPUs this box has), open seams, one exchange that compares the pixel
    pairs across every seam, the mask gathered block by block -- against the oracle on the four reference pictures' kind of
    input, a noise image with lines through every seam, and C2's geometry"""
    import torch
    lib = rg.load_library()
    ndev = torch.cuda.device_count()
    devices = (ctypes.c_int * (R * Cn))(*[b % ndev for b in range(R * Cn)])

    def run(img, seeds, T):
        img = np.ascontiguousarray(img, np.uint8); h, w = img.shape
        seeds = np.ascontiguousarray(np.asarray(seeds, np.int32).reshape(-1, 3))
        out = np.full(img.shape, 9, np.uint8); ms = (ctypes.c_float * 4)()
        rc = lib.rg_grow_multi_grid(img.ctypes.data, w, h, seeds.ctypes.data if len(seeds) else None, len(seeds), T, out.ctypes.data,
                                    devices, R, Cn, 0, ms)
        assert rc == 0, lib.rg_last_error()
        return out, int(ms[3])

    pic = golden["arrays"]["pic1"]; s = oracle.seeds_2d(512, 512)
    got, used = run(pic, s, 2)
    assert (got == oracle.grow(pic, s, 2)).all() and used == 1000 * R + min(Cn, 4)
    rng = np.random.default_rng(60 + R * 10 + Cn)
    img = rng.integers(0, 4, size=(700, 1111), dtype=np.uint8); img[233, :] = 1; img[:, 554:556] = 1; img[300:500, 100:900] = 77
    seeds = np.stack([rng.integers(0, 1111, 8), rng.integers(0, 700, 8), np.zeros(8, np.int64)], 1).astype(np.int32); seeds[0] = [555, 233, 0]
    for T in (1, 2, 256):
        got, _ = run(img, seeds, T)
        assert (got == oracle.grow(img, seeds, T)).all(), T
    img2, s2, T2 = oracle.synth("c2", 2048, 0x5EED0002)
    got, _ = run(img2, s2, T2)
    assert (got == oracle.grow(img2, s2, T2)).all()
    got, _ = run(img2, np.zeros((0, 3), np.int32), T2)
    assert not got.any()


@pytest.mark.parametrize("block", range(4))
def test_random_mixtures_on_gpu(rg, ctx, oracle, block):
    """the emulator's random mixtures (noise, bands of constant rows, solids, thin lines) at the real tile sizes: shapes ragged in every
    axis or whole tiles, thresholds <= 0 and > 255 included, seeds outside the volume and duplicates included -- always the oracle's bytes"""
    from test_emu_logic import _mixed_volume
    rng = np.random.default_rng(0xB200 + block)
    for case in range(12):
        two_d = rng.random() < 0.3
        nx = int(rng.integers(1, 400)) if rng.random() < 0.5 else 64 * int(rng.integers(1, 7))
        shape = (int(rng.integers(1, 300)), nx) if two_d else (int(rng.integers(1, 50)), int(rng.integers(1, 100)), nx)
        img = _mixed_volume(rng, shape)
        dims = (shape[1], shape[0], 1) if two_d else shape[::-1]
        k = int(rng.integers(0, 7))
        seeds = np.stack([rng.integers(-1, d + 1, size=k) for d in dims], 1).astype(np.int32) if k else np.zeros((0, 3), np.int32)
        if two_d and k:
            seeds[:, 2] = 0
        if k > 1:
            seeds[-1] = seeds[0]
        T = int(rng.choice([-3, 0, 1, 1, 2, 5, 41, 255, 256, 1000]))
        want = oracle.grow(img, seeds, T)
        got = dev_grow(rg, ctx, img, seeds, T)
        assert (got == want).all(), (block, case, shape, T, seeds.tolist())
