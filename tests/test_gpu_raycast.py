"""The CUDA ray-caster (csrc/rg_raycast.cu) against the oracle's restatement of raycast_serial
(oracle/raycast_cpu.c, pinned to the unmodified reference by tests/test_oracle_pins.py) and against the
committed picture the unmodified reference rendered (tests/golden: ray64).

Floating point, so the tolerance is stated here once: ZERO grey levels on every pixel.  The kernel is written with
explicit round-to-nearest operations (no FMA contraction) and the same double accumulation as the reference, every
operation it uses is correctly rounded on both machines, and on the B200 every case below came out byte-identical
to the x86-64 result (including the 512 x 512 picture of the reference's own volume), so the test holds it to that.
TOL_LEVELS / TOL_PIXELS are the two knobs, should a future compiler force them open.
"""
import ctypes
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

TOL_LEVELS = 0      # a pixel may differ from the oracle's by at most this many grey levels (of 255)
TOL_PIXELS = 0.0    # and at most this fraction of the pixels may differ at all


def _setup():
    import torch
    if not torch.cuda.is_available():
        pytest.fail("needs a CUDA device; the ray-caster has no CPU path")
    import rgb200
    torch.cuda.set_device(0)
    return torch, rgb200


def _close(got, want, what):
    got = np.asarray(got, np.int16)
    want = np.asarray(want, np.int16)
    assert got.shape == want.shape
    diff = np.abs(got - want)
    nbad = int((diff > 0).sum())
    print(f"{what}: {nbad} of {diff.size} pixels differ, max |diff| = {int(diff.max())}")
    assert int(diff.max()) <= TOL_LEVELS, what
    assert nbad <= TOL_PIXELS * diff.size, what


def _device_render(torch, rgb200, vol, mask, image_dim, cam=1000.0):
    d = torch.from_numpy(np.ascontiguousarray(vol)).cuda()
    r = torch.from_numpy(np.ascontiguousarray(mask)).cuda()
    img = rgb200.raycast_device(d, r, image_dim, cam)
    torch.cuda.synchronize()
    return img.cpu().numpy()


def test_reference_picture_64(oracle, golden):
    """grow on the GPU, render on the GPU, compare with what the unmodified raycast_serial drew (IMAGE_DIM 64)"""
    torch, rgb200 = _setup()
    vol = oracle.create_data_ref().reshape(512, 512, 512)
    mask = rgb200.grow(vol, [golden["c1"]["seed"]], golden["c1"]["threshold"])
    assert int(mask.sum()) == golden["c1"]["region_voxels"]
    img = _device_render(torch, rgb200, vol, mask, 64)
    assert int((img > 0).sum()) > 500
    _close(img, golden["arrays"]["ray64"], "ray64 vs unmodified reference")


@pytest.mark.parametrize("D,I,cam,seed", [(48, 32, 100.0, 1), (64, 64, 90.0, 2), (33, 16, 40.0, 3), (96, 48, 1000.0, 4),
                                          (16, 8, 30.0, 5), (40, 64, 20.0, 6)])
def test_random_volumes_vs_oracle(oracle, D, I, cam, seed):
    """smooth-ish random data and a random blob as region, several cube / image sizes and camera distances;
    (40, 64, 20.0) puts the camera INSIDE the volume, (96, 48, 1000.0) makes the cube a few pixels wide"""
    torch, rgb200 = _setup()
    rng = np.random.default_rng(seed)
    vol = rng.integers(0, 40, size=(D, D, D), dtype=np.uint8)
    zz, yy, xx = np.mgrid[0:D, 0:D, 0:D]
    c = rng.uniform(D * 0.3, D * 0.7, size=3)
    blob = ((zz - c[0]) ** 2 + (yy - c[1]) ** 2 + (xx - c[2]) ** 2) < (D * 0.25) ** 2
    vol[blob] = rng.integers(100, 256, size=int(blob.sum()), dtype=np.uint8)
    mask = (blob & (rng.random((D, D, D)) < 0.9)).astype(np.uint8)
    want = oracle.raycast(vol, mask, I, cam)
    got = _device_render(torch, rgb200, vol, mask, I, cam)
    _close(got, want, f"D={D} I={I} cam={cam}")
    if cam < 500:
        assert int((want > 0).sum()) > 0


def test_region_values_above_one_and_empty_region(oracle):
    """the region sample multiplies the colour ((int) of its trilinear value): masks holding 2s, as the reference's own
    GPU growers leave behind (raycast.cu:637,693), and an all-zero mask (only the 0.01 haze remains)"""
    torch, rgb200 = _setup()
    rng = np.random.default_rng(11)
    D, I, cam = 56, 40, 80.0
    vol = rng.integers(0, 256, size=(D, D, D), dtype=np.uint8)
    mask = rng.integers(0, 3, size=(D, D, D), dtype=np.uint8)
    _close(_device_render(torch, rgb200, vol, mask, I, cam), oracle.raycast(vol, mask, I, cam), "mask with 2s")
    zero = np.zeros_like(mask)
    want = oracle.raycast(vol, zero, I, cam)
    assert 0 < int(want.max())
    _close(_device_render(torch, rgb200, vol, zero, I, cam), want, "empty region")
    _close(_device_render(torch, rgb200, np.zeros_like(vol), mask, I, cam), np.zeros((I, I), np.uint8), "black volume")


def test_reference_shaped_entry_points_512(oracle, golden):
    """raycast_gpu / raycast_gpu_texture: host 512^3 volume + mask in, 512 x 512 picture out, as main() uses them
    (raycast.cu:842); checked against the oracle at the reference's full IMAGE_DIM"""
    torch, rgb200 = _setup()
    vol = oracle.create_data_ref().reshape(512, 512, 512)
    mask = rgb200.grow_region_gpu_shared(vol)
    want = oracle.raycast(vol, mask, 512)
    assert int((want > 0).sum()) > 50000
    a = rgb200.raycast_gpu(vol, mask)
    b = rgb200.raycast_gpu_texture(vol, mask)
    assert (a == b).all()
    _close(a, want, "raycast_gpu 512")
    # a picture is worth a file: the reference's main() ends in write_bmp(image, 512, 512) (raycast.cu:844)
    assert a.shape == (512, 512) and a.dtype == np.uint8


def test_argument_errors():
    torch, rgb200 = _setup()
    lib = rgb200.load_library()
    t = torch.zeros(8, dtype=torch.uint8, device="cuda")
    s = ctypes.c_void_p(0)
    assert lib.rg_raycast_device(None, t.data_ptr(), 2, 2, 1000.0, t.data_ptr(), s) == -1
    assert lib.rg_raycast_device(t.data_ptr(), t.data_ptr(), 1, 2, 1000.0, t.data_ptr(), s) == -1
    assert lib.rg_raycast_device(t.data_ptr(), t.data_ptr(), 2, 0, 1000.0, t.data_ptr(), s) == -1
    assert lib.raycast_gpu(None, None) is None
