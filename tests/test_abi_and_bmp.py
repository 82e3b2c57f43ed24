"""CPU-only checks of the drop-in boundary: the library loads, exports every symbol include/*.h
declares, and the BMP I/O keeps the reference's on-disk format (bmp.c:6-71)."""
import ctypes, os, re, struct, subprocess
import numpy as np
import pytest
from conftest import ROOT, PKG


@pytest.fixture(scope="module")
def lib():
    so = os.path.join(PKG, "libb200rg.so")
    if not os.path.exists(so):
        subprocess.check_call(["make", "-C", PKG], env=dict(os.environ, PATH="/usr/local/cuda/bin:" + os.environ["PATH"]))
    import rgb200
    return rgb200.load_library()


def declared_symbols():
    names = set()
    for h in ("rg.h", "bmp.h"):
        src = open(os.path.join(ROOT, "include", h)).read()
        src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
        names |= set(re.findall(r"\b((?:rg_|grow_region_|raycast_gpu|read_bmp|write_bmp)\w*)\s*\(", src))
    return names


def test_every_declared_symbol_is_exported_and_bound(lib):
    import rgb200
    decl = declared_symbols()
    assert len(decl) >= 25
    for n in decl:
        assert hasattr(lib, n), f"{n} declared in include/ but not exported"
    assert decl == set(rgb200.SIGNATURES), decl ^ set(rgb200.SIGNATURES)


def test_no_cuda_means_error_not_fallback(lib):
    """without a GPU the engine must fail loudly (there is no CPU path)"""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    img = np.zeros((4, 4, 4), np.uint8); out = np.zeros_like(img)
    seeds = np.zeros((1, 3), np.int32)
    rc = lib.rg_grow(img.ctypes.data, 4, 4, 4, seeds.ctypes.data, 1, 1, out.ctypes.data)
    assert rc == -2 and b"no CPU path" in lib.rg_last_error() or rc == -2


def test_argument_errors(lib):
    img = np.zeros((4, 4, 4), np.uint8); out = np.zeros_like(img)
    assert lib.rg_grow(None, 4, 4, 4, None, 0, 1, out.ctypes.data) == -1
    assert lib.rg_grow(img.ctypes.data, 0, 4, 4, None, 0, 1, out.ctypes.data) == -1
    assert lib.rg_grow(img.ctypes.data, 4, 4, 4, None, -1, 1, out.ctypes.data) == -1
    assert lib.rg_grow(img.ctypes.data, 4, 4, 4, None, 0, 1, None) == -1
    assert lib.rg_workspace_bytes(512, 512, 512, 1) > 0
    assert lib.rg_workspace_bytes(-1, 512, 512, 1) == -1


def test_bmp_layout_and_roundtrip(lib, tmp_path, golden):
    import rgb200
    img = golden["arrays"]["pic1"]
    p = str(tmp_path / "a.bmp")
    rgb200.write_bmp(p, img)
    b = open(p, "rb").read()
    assert len(b) == 512 * 512 + 1078 + 2                      # bmp.c:41-45
    assert b[:2] == b"BM"
    assert struct.unpack_from("<I", b, 2)[0] == 512 * 512 + 56  # file_size = w*h+54+2, bmp.c:12
    assert struct.unpack_from("<I", b, 10)[0] == 1078           # pixel_offset, bmp.c:13
    assert b[6:10] == b"\0\0\0\0" and b[-2:] == b"\0\0"       # deterministic where the reference is not
    assert struct.unpack_from("<IiiHHIIiiII", b, 14) == (40, 512, 512, 1, 8, 0, 512 * 512, 0, 0, 256, 0)
    pal = np.frombuffer(b, np.uint8, 1024, 54).reshape(256, 4)
    assert (pal[:, 0] == np.arange(256)).all() and (pal[:, 1] == pal[:, 0]).all() and (pal[:, 3] == 0).all()
    assert b[1078:1078 + 512 * 512] == img.tobytes()
    back = rgb200.read_bmp(p)
    assert back.shape == (512, 512) and (back == img).all()


def test_write_bmp_fixed_name_and_read_bmp_symbol(lib, tmp_path, golden):
    """the reference-shaped symbols themselves: write_bmp always writes ./out.bmp (bmp.c:39)"""
    img = np.ascontiguousarray(golden["arrays"]["pic2"])
    cwd = os.getcwd()
    os.chdir(tmp_path)
    try:
        lib.write_bmp(img.ctypes.data, 512, 512)
        assert os.path.getsize("out.bmp") == 512 * 512 + 1080
        ptr = lib.read_bmp(b"out.bmp")
        assert ptr
        back = np.ctypeslib.as_array(ctypes.cast(ptr, ctypes.POINTER(ctypes.c_uint8)), (512, 512)).copy()
        ctypes.CDLL(None).free(ctypes.c_void_p(ptr))
        assert (back == img).all()
        assert not lib.read_bmp(b"does-not-exist.bmp")          # NULL, not a crash
    finally:
        os.chdir(cwd)


def test_read_bmp_uses_the_files_pixel_offset(lib, tmp_path):
    """pic*.bmp carry a 124-byte V5 header (offset 1162); read_bmp must honour bytes 10..13 (bmp.c:57-60)"""
    w, h, off = 8, 4, 1162
    payload = np.arange(w * h, dtype=np.uint8)
    raw = bytearray(off + w * h)
    raw[0:2] = b"BM"; struct.pack_into("<I", raw, 10, off); struct.pack_into("<ii", raw, 18, w, h)
    raw[off:] = payload.tobytes()
    p = tmp_path / "v5.bmp"; p.write_bytes(bytes(raw))
    import rgb200
    assert (rgb200.read_bmp(str(p)).ravel() == payload).all()
