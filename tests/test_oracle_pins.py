"""The oracle (oracle/grow_cpu.c) against the reference's own outputs (tests/golden, made by
tests/golden/make_golden.py from the UNMODIFIED reference) and an independent scipy model."""
import hashlib
import numpy as np
import pytest
from conftest import cc_model


@pytest.mark.parametrize("k", [1, 2, 3, 4])
def test_region_c_payloads(oracle, golden, k):
    """region.c (4 corner seeds, |d|<2, 4-conn; region.c:76-83,450-533,572-580) on pic1..pic4."""
    img = golden["arrays"][f"pic{k}"]
    assert hashlib.sha256(img.tobytes()).hexdigest() == golden["pics"][f"pic{k}"]["in_sha256"]
    seeds = oracle.seeds_2d(512, 512)
    assert seeds.tolist() == [[4, 4, 0], [506, 4, 0], [506, 506, 0], [4, 506, 0]]
    mask = oracle.grow(img, seeds, 2)
    assert set(np.unique(mask)) <= {0, 1}
    out = img * (mask == 0)
    assert hashlib.sha256(out.tobytes()).hexdigest() == golden["pics"][f"pic{k}"]["out_sha256"]
    assert (out == golden["arrays"][f"out{k}"]).all()


def test_committed_out_bmp(oracle, golden):
    """The only known-answer artefact the reference ships: 2-mpi-region-growing/out.bmp (64 420 px)."""
    img = golden["arrays"]["pic1"]
    mask = oracle.grow(img, oracle.seeds_2d(512, 512), 2)
    assert int(mask.sum()) == 64420
    out = img * (mask == 0)
    assert hashlib.sha256(out.tobytes()).hexdigest() == golden["committed_out_bmp_payload_sha256"]


def test_pic1_single_seeds_and_thresholds(oracle, golden):
    img = golden["arrays"]["pic1"]
    sizes = [int(oracle.grow(img, [s], 2).sum()) for s in oracle.seeds_2d(512, 512)]
    assert sizes == [24056, 15446, 12082, 12836]
    s4 = oracle.seeds_2d(512, 512)
    assert [int(oracle.grow(img, s4, t).sum()) for t in (1, 2, 3, 5)] == [38169, 64420, 64420, 68348]


def test_c1_reference_volume(oracle, golden):
    """raycast.cu:114-158 + 281-318 on the reference's own 512^3 volume."""
    vol = oracle.create_data_ref()
    assert oracle.fnv(vol) == golden["c1"]["volume_fnv1a64"]
    mask = oracle.grow(vol, [golden["c1"]["seed"]], golden["c1"]["threshold"])
    assert int(mask.sum()) == golden["c1"]["region_voxels"] == 2197899
    assert oracle.fnv(mask) == golden["c1"]["mask_fnv1a64"]
    assert set(np.unique(mask)) == {0, 1}


@pytest.mark.parametrize("k", [1, 3])
def test_scipy_model_agrees(oracle, golden, k):
    img = golden["arrays"][f"pic{k}"]
    seeds = oracle.seeds_2d(512, 512)
    assert (oracle.grow(img, seeds, 2) == cc_model(img, seeds, 2)).all()


@pytest.mark.parametrize("shape,T", [((1, 1, 1), 1), ((1, 1, 37), 1), ((5, 7, 9), 2), ((16, 16, 16), 1),
                                     ((33, 20), 3), ((1, 64), 2), ((40, 3, 2), 0), ((6, 6, 6), 255), ((6, 6, 6), 256)])
def test_small_random_vs_model(oracle, shape, T):
    rng = np.random.default_rng(hash(shape) % 2**32 + T)
    img = rng.integers(0, 6, size=shape, dtype=np.uint8)
    if T >= 255:
        img = rng.integers(0, 256, size=shape, dtype=np.uint8)
    dims = shape[::-1] if len(shape) == 3 else (shape[1], shape[0], 1)
    seeds = np.stack([rng.integers(-1, d + 1, size=5) for d in dims], 1).astype(np.int32)
    if len(shape) == 2:
        seeds[:, 2] = rng.integers(0, 2, size=5)
    seeds[0] = 0
    assert (oracle.grow(img, seeds, T) == cc_model(img, seeds, T)).all()


@pytest.mark.parametrize("kind,n,nthreads", [("c1h", 64, 4), ("c3", 32, 3), ("c4", 64, 8), ("c2", 256, 4), ("c3", 17, 5)])
def test_parallel_port_equals_serial(oracle, kind, n, nthreads):
    """oracle_grow_par (region.c:493-533 round structure on threads) == oracle_grow."""
    img, seeds, T = oracle.synth(kind, n, 0x5EED0000)
    a = oracle.grow(img, seeds, T)
    b, rounds = oracle.grow_par(img, seeds, T, nthreads)
    assert (a == b).all() and rounds >= 1


def test_c3_count_formula(oracle):
    for n in (8, 17, 32):
        img, seeds, T = oracle.synth("c3", n)
        m = oracle.grow(img, seeds, T)
        assert int(m.sum()) == int((img == 200).sum()) == oracle.lib.oracle_c3_count(n)


def test_raycast_restatement_equals_reference(oracle, golden):
    """oracle_raycast (raycast.cu:183-267) at IMAGE_DIM 64 on the reference volume + mask == the unmodified
    raycast_serial (golden ray64, produced by tests/golden/make_golden.py)"""
    vol = oracle.create_data_ref()
    mask = oracle.grow(vol, [golden["c1"]["seed"]], 1)
    img = oracle.raycast(vol, mask, 64)
    assert hashlib.sha256(img.tobytes()).hexdigest() == golden["raycast64_sha256"]
    assert (img == golden["arrays"]["ray64"]).all()
    assert int((img > 0).sum()) == golden["raycast64_nonzero_px"] > 500
