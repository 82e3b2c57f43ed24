"""Engine LOGIC on the CPU: the kernel bodies of rg_core.h / rg_slab_core.h, walked sequentially by
tests/emu (test code, not part of the product), against the oracle.  Tiny tile shapes put many tile
faces, partial tiles and slab interfaces inside small random grids."""
import ctypes, os, subprocess
import numpy as np
import pytest
from conftest import ROOT

u8p = ctypes.c_void_p


@pytest.fixture(scope="module")
def emu():
    d = os.path.join(ROOT, "tests", "emu")
    subprocess.check_call(["make", "-s", "-C", d])
    L = ctypes.CDLL(os.path.join(d, "librg_emu.so"))
    L.emu_grow.argtypes = [ctypes.c_int, u8p, ctypes.c_int64, ctypes.c_int64, ctypes.c_int64, ctypes.c_void_p, ctypes.c_int,
                           ctypes.c_int, u8p, ctypes.c_void_p]
    L.emu_grow_slabs.argtypes = [ctypes.c_int, ctypes.c_int, u8p, ctypes.c_int64, ctypes.c_int64, ctypes.c_int64, ctypes.c_void_p,
                                 ctypes.c_int, ctypes.c_int, u8p, ctypes.c_void_p]
    return L


def run_emu(L, cfg, img, seeds, T, slabs=0):
    img = np.ascontiguousarray(img, np.uint8)
    nz, ny, nx = (1,) + img.shape if img.ndim == 2 else img.shape
    seeds = np.ascontiguousarray(np.asarray(seeds, np.int32).reshape(-1, 3))
    mask = np.full(img.shape, 7, np.uint8)
    stats = np.zeros(4, np.uint32)
    if slabs:
        rc = L.emu_grow_slabs(cfg, slabs, img.ctypes.data, nx, ny, nz, seeds.ctypes.data, len(seeds), T, mask.ctypes.data, stats.ctypes.data)
    else:
        rc = L.emu_grow(cfg, img.ctypes.data, nx, ny, nz, seeds.ctypes.data, len(seeds), T, mask.ctypes.data, stats.ctypes.data)
    assert rc == 0, (rc, stats)
    return mask, stats


def rand_case(rng, shape, levels, nseeds):
    img = rng.integers(0, levels, size=shape, dtype=np.uint8)
    dims = shape[::-1] if len(shape) == 3 else (shape[1], shape[0], 1)
    seeds = np.stack([rng.integers(-1, d + 1, size=nseeds) for d in dims], 1).astype(np.int32)
    if len(shape) == 2:
        seeds[:, 2] = 0
    return img, seeds


@pytest.mark.parametrize("cfg", [1, 2])
@pytest.mark.parametrize("shape", [(9, 10, 70), (8, 8, 64), (5, 13, 33), (16, 4, 32), (1, 1, 1), (2, 9, 100), (12, 12, 12)])
@pytest.mark.parametrize("levels,T", [(3, 1), (6, 2), (2, 1), (40, 3)])
def test_emu_3d_tiny_tiles(emu, oracle, cfg, shape, levels, T):
    rng = np.random.default_rng(abs(hash((cfg, shape, levels, T))) % 2**32)
    img, seeds = rand_case(rng, shape, levels, 6)
    got, _ = run_emu(emu, cfg, img, seeds, T)
    assert (got == oracle.grow(img, seeds, T)).all()


@pytest.mark.parametrize("shape", [(20, 70), (8, 32), (33, 33), (1, 5), (17, 128), (64, 31)])
@pytest.mark.parametrize("levels,T", [(3, 1), (6, 2), (2, 1)])
def test_emu_2d_tiny_tiles(emu, oracle, shape, levels, T):
    rng = np.random.default_rng(abs(hash((shape, levels, T))) % 2**32)
    img, seeds = rand_case(rng, shape, levels, 5)
    got, _ = run_emu(emu, 1, img, seeds, T)
    assert (got == oracle.grow(img, seeds, T)).all()


@pytest.mark.parametrize("T", [0, 1, 255, 256, 300, -3])
def test_emu_threshold_extremes(emu, oracle, T):
    rng = np.random.default_rng(5 + abs(T))
    img = rng.integers(0, 256, size=(6, 9, 40), dtype=np.uint8)
    seeds = [[0, 0, 0], [39, 8, 5]]
    got, _ = run_emu(emu, 1, img, seeds, T)
    assert (got == oracle.grow(img, seeds, T)).all()


def test_emu_product_tile_shapes(emu, oracle):
    """the real 64x32x16 and 128x256 tiles, on grids a little larger than one tile"""
    rng = np.random.default_rng(11)
    img = rng.integers(0, 3, size=(20, 40, 100), dtype=np.uint8)
    seeds = [[1, 1, 1], [99, 39, 19], [64, 32, 16], [63, 31, 15]]
    got, st = run_emu(emu, 0, img, seeds, 1)
    assert (got == oracle.grow(img, seeds, 1)).all()
    img2 = rng.integers(0, 4, size=(300, 200), dtype=np.uint8)
    seeds2 = [[4, 4, 0], [194, 4, 0], [194, 294, 0], [4, 294, 0]]
    got2, _ = run_emu(emu, 0, img2, seeds2, 2)
    assert (got2 == oracle.grow(img2, seeds2, 2)).all()


def test_emu_no_seeds_and_all_dropped(emu, oracle):
    img = np.zeros((4, 4, 32), np.uint8)
    got, _ = run_emu(emu, 1, img, np.zeros((0, 3), np.int32), 1)
    assert not got.any()
    got, _ = run_emu(emu, 1, img, [[-1, 0, 0], [32, 0, 0], [0, 4, 0], [0, 0, 4]], 1)
    assert not got.any()
    got, _ = run_emu(emu, 1, img, [[31, 3, 3]], 1)
    assert got.all()


@pytest.mark.parametrize("kind,n", [("c3", 24), ("c3", 33), ("c1h", 64), ("c4", 64), ("c2", 128)])
def test_emu_synthetic_shapes(emu, oracle, kind, n):
    img, seeds, T = oracle.synth(kind, n, 0x5EED0000)
    got, _ = run_emu(emu, 1, img, seeds, T)
    assert (got == oracle.grow(img, seeds, T)).all()


@pytest.mark.parametrize("G", [2, 3, 4, 8])
@pytest.mark.parametrize("cfg", [1, 2])
def test_emu_slabs_random(emu, oracle, G, cfg):
    rng = np.random.default_rng(100 * G + cfg)
    img, seeds = rand_case(rng, (19, 10, 70), 3, 6)
    got, st = run_emu(emu, cfg, img, seeds, 1, slabs=G)
    assert (got == oracle.grow(img, seeds, 1)).all()


@pytest.mark.parametrize("G", [2, 3, 5, 8])
@pytest.mark.parametrize("cfg,shape", [(1, (37, 70)), (1, (64, 96)), (1, (16, 33)), (0, (300, 272)), (0, (255, 130))])
@pytest.mark.parametrize("levels,T", [(3, 1), (5, 2)])
def test_emu_row_slabs_of_an_image(emu, oracle, G, cfg, shape, levels, T):
    """2-D inputs are cut into ROW slabs (the reference's region.c gives each MPI rank a block of the picture,
    region.c:106-143): the interface is the YM / YP face row of the first / last tile row, ragged last tile rows
    (ny % TY != 0) included, one exchange"""
    rng = np.random.default_rng(abs(hash((G, cfg, shape, levels))) % 2**32)
    img, seeds = rand_case(rng, shape, levels, 6)
    got, st = run_emu(emu, cfg, img, seeds, T, slabs=G)
    assert (got == oracle.grow(img, seeds, T)).all()
    assert int(st[3]) == G


@pytest.mark.parametrize("G", [2, 4, 7])
def test_emu_row_slabs_snake(emu, oracle, G):
    """a 1-pixel snake that crosses every row interface many times + the reference picture with its four seeds"""
    h, w = 64, 96
    img = np.zeros((h, w), np.uint8)
    img[:, 1::4] = 200                                    # vertical corridors every 4 columns ...
    img[0, 1::8] = 0; img[h - 1, 5::8] = 0                # ... alternately closed at the top / bottom
    for k, x in enumerate(range(1, w - 4, 4)):
        img[h - 1 if k % 2 == 0 else 0, x:x + 5] = 200    # joined into one snake by turns in the end rows
    seeds = [[1, 1, 0]]
    want = oracle.grow(img, seeds, 1)
    assert want.sum() > 0.2 * h * w
    got, _ = run_emu(emu, 1, img, seeds, 1, slabs=G)
    assert (got == want).all()


def test_emu_row_slabs_reference_picture(emu, oracle, golden):
    img = golden["arrays"]["pic1"]
    seeds = oracle.seeds_2d(512, 512)
    want = oracle.grow(img, seeds, 2)
    for G in (2, 4):
        got, _ = run_emu(emu, 0, img, seeds, 2, slabs=G)
        assert (got == want).all()


@pytest.mark.parametrize("G", [2, 4, 8, 11])
def test_emu_slabs_serpentine(emu, oracle, G):
    """C4's z-serpentine crosses every slab interface many times: one exchange must be enough"""
    img, seeds, T = oracle.synth("c4", 64, 0x5EED0004)
    got, _ = run_emu(emu, 1, img, seeds, T, slabs=G)
    assert (got == oracle.grow(img, seeds, T)).all()
    img, seeds, T = oracle.synth("c3", 20)
    got, _ = run_emu(emu, 2, img, seeds, T, slabs=G)
    assert (got == oracle.grow(img, seeds, T)).all()


def test_emu_fast_and_full_paths_mix(emu, oracle):
    """sparse noise tiles take the fast (walker) pass, solid tiles the full pass; neighbours on different
    paths must still agree on their shared face lists"""
    img, seeds, T = oracle.synth("c4", 96, 0x5EED0004)
    want = oracle.grow(img, seeds, T)
    got, st = run_emu(emu, 1, img, seeds, T)
    assert (got == want).all()
    ncomplex, ntiles = int(st[2]), int(st[3])
    assert 0 < ncomplex < ntiles, (ncomplex, ntiles)
    emu.emu_force_full(1)
    try:
        got2, st2 = run_emu(emu, 1, img, seeds, T)
    finally:
        emu.emu_force_full(0)
    assert int(st2[2]) == ntiles and (got2 == want).all()
    assert int(st2[0]) == int(st[0])                      # same number of coarse nodes whichever path labelled the tiles


def test_emu_saved_link_planes_do_not_change_the_result(emu, oracle):
    """the full pass keeps the link planes of the tiles it labels and pass C floods from them instead of staging the
    tile again; the emulator keeps room for every second such tile, so both branches run -- and with no room at all
    the result is the same"""
    img, seeds, T = oracle.synth("c3", 40)
    want = oracle.grow(img, seeds, T)
    got, st = run_emu(emu, 1, img, seeds, T)
    assert (got == want).all() and int(st[2]) > 2                   # several tiles went through the full pass
    emu.emu_no_planes(1)
    try:
        got2, _ = run_emu(emu, 1, img, seeds, T)
    finally:
        emu.emu_no_planes(0)
    assert (got2 == want).all()


@pytest.mark.parametrize("levels", [2, 4, 12, 40])
def test_emu_walk_limit_boundary(emu, oracle, levels):
    """component sizes around WALK_MAX: some walks overflow, some do not"""
    rng = np.random.default_rng(levels)
    img = rng.integers(0, levels, size=(12, 20, 96), dtype=np.uint8)
    seeds = np.stack([rng.integers(0, 96, 10), rng.integers(0, 20, 10), rng.integers(0, 12, 10)], 1)
    got, st = run_emu(emu, 1, img, seeds, 1)
    assert (got == oracle.grow(img, seeds, 1)).all()
    got, st = run_emu(emu, 0, img, seeds, 1)
    assert (got == oracle.grow(img, seeds, 1)).all()


@pytest.mark.parametrize("cfg", [0, 1, 2])
def test_emu_uniform_tiles(emu, oracle, cfg):
    """tiles lying wholly inside a solid take the uniform shortcut (one component, representative 0,
    pass C fills ones without reading the image); their neighbours are mixed / noise tiles"""
    rng = np.random.default_rng(9 + cfg)
    img = rng.integers(0, 20, size=(48, 96, 192), dtype=np.uint8)
    img[4:44, 10:90, 20:180] = 77                       # a solid spanning many whole tiles
    img[10:12, 30:40, 64:128] = 3                       # a hole touching whole-tile boundaries
    seeds = [[100, 50, 20], [0, 0, 0], [191, 95, 47]]
    want = oracle.grow(img, seeds, 1)
    got, st = run_emu(emu, cfg, img, seeds, 1)
    assert (got == want).all()
    assert want[20, 50, 100] == 1 and want.sum() > 40 * 80 * 160 - 2 * 10 * 64 - 10
    # same image, seed only in the noise: the solid stays out although its tiles are uniform
    got2, _ = run_emu(emu, cfg, img, [[0, 0, 0]], 1)
    assert (got2 == oracle.grow(img, [[0, 0, 0]], 1)).all() and got2[20, 50, 100] == 0
    # threshold 0: no edges at all, the shortcut must not fire
    got3, _ = run_emu(emu, cfg, img, seeds, 0)
    assert got3.sum() == 3


def test_emu_random_stress(emu, oracle):
    """many small random cases in one go: shapes around the tile sizes of the three emulator configurations (ragged in
    every axis), 2 to 40 grey levels, thresholds 0 to 300, seeds inside, outside and duplicated, single- and multi-slab"""
    rng = np.random.default_rng(20240607)
    for case in range(120):
        cfg = int(rng.integers(1, 3))
        if rng.random() < 0.3:
            shape = (int(rng.integers(1, 40)), int(rng.integers(1, 140)))
        else:
            shape = (int(rng.integers(1, 14)), int(rng.integers(1, 22)), int(rng.integers(1, 140)))
        levels = int(rng.choice([2, 3, 5, 12, 40]))
        T = int(rng.choice([0, 1, 1, 2, 3, 17, 255, 256, 300]))
        img, seeds = rand_case(rng, shape, levels, int(rng.integers(0, 7)))
        if len(seeds) > 1 and rng.random() < 0.3:
            seeds[-1] = seeds[0]
        want = oracle.grow(img, seeds, T)
        got, _ = run_emu(emu, cfg, img, seeds, T)
        assert (got == want).all(), (case, cfg, shape, levels, T)
        axis_len = shape[0]
        G = int(rng.integers(2, 5))
        if axis_len >= 2 * G:                                       # every logical rank needs two planes / rows
            got, _ = run_emu(emu, cfg, img, seeds, T, slabs=G)
            assert (got == want).all(), ("slabs", case, cfg, shape, levels, T, G)


def test_emu_dense_path_labels_every_tile(emu, oracle):
    """the dense path of the fast pass (rg_dense.h: link words, ranked runs, union-find on run ids) forced onto every
    non-uniform tile: random grids in all emulator tile shapes, ragged sizes, 2-D and 3-D, multi-slab too"""
    emu.emu_last_dense.restype = ctypes.c_uint32; emu.emu_last_full.restype = ctypes.c_uint32
    rng = np.random.default_rng(77)
    emu.emu_force_dense(1)
    try:
        for case in range(80):
            cfg = int(rng.integers(0, 3))
            if rng.random() < 0.3:
                shape = (int(rng.integers(1, 150)), int(rng.integers(1, 300)))
            else:
                shape = (int(rng.integers(1, 20)), int(rng.integers(1, 40)), int(rng.integers(1, 140)))
            levels = int(rng.choice([2, 3, 5, 12, 40]))
            T = int(rng.choice([0, 1, 1, 2, 3, 17, 129, 200, 255, 256]))
            img, seeds = rand_case(rng, shape, levels, int(rng.integers(0, 7)))
            want = oracle.grow(img, seeds, T)
            got, st = run_emu(emu, cfg, img, seeds, T)
            assert (got == want).all(), (case, cfg, shape, levels, T)
            assert emu.emu_last_full() == 0, (case, cfg, shape)
            G = int(rng.integers(2, 4))
            if cfg != 0 and shape[0] >= 2 * G:
                got, _ = run_emu(emu, cfg, img, seeds, T, slabs=G)
                assert (got == want).all(), ("slabs", case, cfg, shape, levels, T, G)
        for kind, n in (("c3", 40), ("c4", 96), ("c1h", 64), ("c2", 256)):
            img, seeds, T = oracle.synth(kind, n, 0x5EED0000)
            for cfg in (0, 1):
                got, _ = run_emu(emu, cfg, img, seeds, T)
                assert (got == oracle.grow(img, seeds, T)).all(), (kind, cfg)
                assert emu.emu_last_dense() > 0
    finally:
        emu.emu_force_dense(0)


def test_emu_dense_path_overflow_goes_to_the_full_pass(emu, oracle):
    """a product-size tile with more ranked runs than the dense path's parent table holds (every voxel its own run,
    all of them linked in y) is handed to the full labelling pass; its neighbours stay on the dense path"""
    emu.emu_last_dense.restype = ctypes.c_uint32; emu.emu_last_full.restype = ctypes.c_uint32
    img = np.zeros((16, 64, 128), np.uint8)
    img[:, :, :64] = (np.arange(64) % 2 * 100).astype(np.uint8)[None, None, :]      # columns 0/100: runs of one voxel, linked in y and z
    img[:, :, 64:] = np.random.default_rng(3).integers(0, 12, size=(16, 64, 64))
    seeds = [[0, 0, 0], [1, 5, 5], [100, 40, 3]]
    want = oracle.grow(img, seeds, 1)
    emu.emu_force_dense(1)
    try:
        got, st = run_emu(emu, 0, img, seeds, 1)
    finally:
        emu.emu_force_dense(0)
    assert (got == want).all()
    assert emu.emu_last_full() >= 1 and emu.emu_last_dense() >= 1, (emu.emu_last_full(), emu.emu_last_dense())


@pytest.mark.parametrize("cfg", [0, 1, 2])
def test_emu_row_tiles(emu, oracle, cfg):
    """tiles whose rows are constant are labelled as a grid of rows (rg_rows.h) and flooded row-wise in pass C; their
    neighbours are noise tiles, uniform tiles and ragged tiles on the ordinary paths; with no link-plane slot left (the
    emulator keeps room for every second tile) a row tile is labelled like any other"""
    emu.emu_last_rows.restype = ctypes.c_uint32
    rng = np.random.default_rng(123 + cfg)
    nz, ny, nx = (40, 70, 256) if cfg == 0 else (14, 26, 208)
    rowval = rng.integers(0, 3, size=(nz, ny, 1), dtype=np.uint8) * 50           # every row constant, neighbours often equal
    img = np.broadcast_to(rowval, (nz, ny, nx)).copy()
    img[:, :, nx // 2: nx // 2 + 40] = rng.integers(0, 4, size=(nz, ny, 40))      # a noise band: those tiles are not row tiles
    img[2:8, 3:9, :] = 200                                                        # a slab of equal rows
    seeds = [[0, 0, 0], [nx - 1, ny - 1, nz - 1], [5, 5, 5], [nx // 2 + 3, 7, 3], [10, ny // 2, nz // 2]]
    for T in (1, 60, 300):
        want = oracle.grow(img, seeds, T)
        got, st = run_emu(emu, cfg, img, seeds, T)
        assert (got == want).all(), (cfg, T)
        if T == 1:
            assert emu.emu_last_rows() > 0
        emu.emu_no_rows(1)
        try:
            got2, _ = run_emu(emu, cfg, img, seeds, T)
        finally:
            emu.emu_no_rows(0)
        assert (got2 == want).all() and emu.emu_last_rows() == 0
        if nz >= 6:
            got3, _ = run_emu(emu, cfg, img, seeds, T, slabs=3)
            assert (got3 == want).all(), ("slabs", cfg, T)
    # the serpentine: every tile column but the two with the turns is made of constant rows
    img3, s3, T3 = oracle.synth("c3", 192 if cfg != 1 else 96)
    got, _ = run_emu(emu, cfg, img3, s3, T3)
    assert (got == oracle.grow(img3, s3, T3)).all() and emu.emu_last_rows() > 0


@pytest.mark.parametrize("cfg", [1, 2])
def test_emu_output_with_and_without_a_cleared_mask(emu, oracle, cfg):
    """The labelling pass clears each tile's part of the mask, so the output passes store only the chunks / rows that hold a one;
    with the clear switched off (the library's RG_NO_PREZERO) they and the flag pass must write every byte themselves.  Both ways
    start from a mask full of garbage (run_emu fills it with 7) over every kind of tile: noise, uniform, rows, dense, ragged."""
    rng = np.random.default_rng(77 + cfg)
    nz, ny, nx = 22, 45, 208
    rowval = rng.integers(0, 3, size=(nz, ny, 1), dtype=np.uint8) * 50
    img = np.broadcast_to(rowval, (nz, ny, nx)).copy()                            # row tiles ...
    img[:, :, 96:150] = rng.integers(0, 4, size=(nz, ny, 54))                     # ... a noise band ...
    img[:, 30:, :64] = 120                                                        # ... a solid (uniform tiles and their borders)
    img[4:9, 2:20, 10:70] = 200
    seeds = [[0, 0, 0], [5, 35, 3], [12, 5, 5], [100, 7, 3], [nx - 1, ny - 1, nz - 1]]
    c3, s3, T3 = oracle.synth("c3", 96)
    for (vol, sd, T) in ((img, seeds, 1), (img, seeds, 60), (c3, s3, T3)):
        want = oracle.grow(vol, sd, T)
        got, _ = run_emu(emu, cfg, vol, sd, T)
        emu.emu_no_prezero(1)
        try:
            got2, _ = run_emu(emu, cfg, vol, sd, T)
        finally:
            emu.emu_no_prezero(0)
        assert (got == want).all() and (got2 == want).all(), (cfg, T)


@pytest.mark.parametrize("R,Cn", [(1, 2), (2, 1), (2, 2), (3, 4), (4, 3)])
@pytest.mark.parametrize("cfg,shape", [(1, (40, 160)), (1, (33, 131)), (0, (260, 530)), (1, (9, 64))])
def test_emu_image_over_a_grid_of_blocks(emu, oracle, golden, R, Cn, cfg, shape):
    """an image cut into rows x columns blocks (the reference's Cartesian process grid, region.c:537-548, 106-143): every block
    labels itself with open seams, the exchange compares the pixel pairs across each seam; ragged blocks, seeds on seams,
    snakes that cross seams many times"""
    emu.emu_grow_grid.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_int, u8p, ctypes.c_int64, ctypes.c_int64, ctypes.c_void_p, ctypes.c_int,
                                  ctypes.c_int, u8p, ctypes.c_void_p]
    rng = np.random.default_rng(abs(hash((R, Cn, cfg, shape))) % 2**32)
    h, w = shape
    for levels, T in ((3, 1), (6, 2), (2, 1), (5, 300), (4, 0)):
        img = rng.integers(0, levels, size=shape, dtype=np.uint8)
        img[h // 3, :] = 1                                             # a line through every column of blocks
        img[:, w // 2 - 1: w // 2 + 1] = 1                             # ... and through every row of blocks
        seeds = np.stack([rng.integers(0, w, 6), rng.integers(0, h, 6), np.zeros(6, np.int64)], 1).astype(np.int32)
        seeds[0] = [w // 2, h // 3, 0]
        want = oracle.grow(img, seeds, T)
        got = np.full(shape, 9, np.uint8); st = np.zeros(4, np.uint32)
        rc = emu.emu_grow_grid(cfg, R, Cn, img.ctypes.data, w, h, seeds.ctypes.data, len(seeds), T, got.ctypes.data, st.ctypes.data)
        assert rc == 0 and (got == want).all(), (levels, T, list(st))
    if cfg == 0 and R * Cn >= 4:
        pic = golden["arrays"]["pic1"]
        s = oracle.seeds_2d(512, 512)
        got = np.empty_like(pic); st = np.zeros(4, np.uint32)
        assert emu.emu_grow_grid(0, R, Cn, pic.ctypes.data, 512, 512, s.ctypes.data, 4, 2, got.ctypes.data, st.ctypes.data) == 0
        assert (got == oracle.grow(pic, s, 2)).all() and int(st[2]) == R and int(st[3]) == min(Cn, 4)


def _mixed_volume(rng, shape):
    """noise, constant rows, a solid and a thin line in one volume / image: every labelling path of the fast pass next to every other"""
    img = rng.integers(0, int(rng.integers(2, 6)), size=shape, dtype=np.uint8)
    nd = len(shape)
    if rng.random() < 0.7:                                       # a band of constant rows (row tiles where the band covers whole tiles)
        rowval = (rng.integers(0, 3, size=shape[:-1] + (1,), dtype=np.uint8) * 40).astype(np.uint8)
        if rng.random() < 0.5:
            lo = int(rng.integers(0, shape[-1])); hi = int(rng.integers(lo, shape[-1] + 1))
        else:                                                    # ... cut on multiples of 32 columns: whole tiles of the tiny tile shapes
            lo = 32 * int(rng.integers(0, shape[-1] // 32 + 1)); hi = max(lo, 32 * int(rng.integers(0, shape[-1] // 32 + 1)))
            if rng.random() < 0.5:
                lo, hi = 0, shape[-1]
        img[..., lo:hi] = np.broadcast_to(rowval, shape)[..., lo:hi]
    if rng.random() < 0.7:                                       # a solid box
        sl = tuple(slice(int(a), int(a) + int(rng.integers(1, max(2, s)))) for s in shape for a in [rng.integers(0, s)])
        img[sl] = int(rng.integers(100, 255))
    if rng.random() < 0.5:                                       # a one-voxel line along the first axis
        idx = tuple(int(rng.integers(0, s)) for s in shape[1:])
        img[(slice(None),) + idx] = 250
    return img


def test_emu_property_random_mixtures(emu, oracle):
    """hypothesis over shapes (ragged in every axis), contents, thresholds (<= 0 and > 255 included), seeds (outside the volume and
    duplicates included), tile shapes, slab counts and the two output modes: always the oracle's bytes"""
    from hypothesis import given, settings, strategies as st, HealthCheck

    @settings(max_examples=400, deadline=None, suppress_health_check=list(HealthCheck), derandomize=True)
    @given(st.integers(0, 2**31 - 1), st.sampled_from([1, 2]), st.sampled_from([0, 0, 2, 3, 5]), st.booleans(), st.booleans(),
           st.sampled_from([-3, 0, 1, 1, 2, 5, 41, 255, 256, 1000]))
    def prop(seed, cfg, slabs, two_d, no_prezero, T):
        rng = np.random.default_rng(seed)
        nx = int(rng.integers(1, 150)) if rng.random() < 0.5 else 32 * int(rng.integers(1, 6))       # ragged, or whole tiles and 16-byte rows
        shape = (int(rng.integers(1, 40)), nx) if two_d else (int(rng.integers(1, 20)), int(rng.integers(1, 30)), nx)
        img = _mixed_volume(rng, shape)
        dims = (shape[1], shape[0], 1) if two_d else shape[::-1]
        k = int(rng.integers(0, 7))
        seeds = np.stack([rng.integers(-1, d + 1, size=k) for d in dims], 1).astype(np.int32) if k else np.zeros((0, 3), np.int32)
        if two_d and k:
            seeds[:, 2] = 0
        if k > 1:
            seeds[-1] = seeds[0]
        want = oracle.grow(img, seeds, T)
        nslab = slabs if slabs and shape[0] >= slabs else 0
        emu.emu_no_prezero(1 if no_prezero else 0)
        try:
            got, _ = run_emu(emu, cfg, img, seeds, T, slabs=nslab)
        finally:
            emu.emu_no_prezero(0)
        assert (got == want).all(), (seed, cfg, nslab, shape, T, no_prezero)
        if not nslab:
            hit["rows"] += emu.emu_last_rows(); hit["dense"] += emu.emu_last_dense(); hit["runs"] += 1

    hit = {"rows": 0, "dense": 0, "runs": 0}
    emu.emu_last_rows.restype = ctypes.c_uint32; emu.emu_last_dense.restype = ctypes.c_uint32
    prop()
    assert hit["runs"] > 50 and hit["rows"] > 0 and hit["dense"] > 0, hit      # the mixtures did reach the row and the dense labelling paths
