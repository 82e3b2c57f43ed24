"""A compiled C host on the drop-in boundary: the reference's two main() programs (region.c:582-604,
raycast.cu:824-854) rewritten as INTEGRATION.md sections 1-2 show, built with gcc against include/*.h and
-lb200rg -- no Python between the caller and the library.  The CPU part checks that the headers compile as C99
and that the programs link; the GPU part runs them and compares what they write with the unmodified reference's
output (tests/golden)."""
import hashlib, os, re, subprocess
import numpy as np
import pytest
from conftest import ROOT, PKG

HOSTS = os.path.join(ROOT, "tests", "c_host")


def build(tmp, name):
    exe = os.path.join(tmp, name)
    subprocess.check_call(["gcc", "-std=gnu99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), "-o", exe,
                           os.path.join(HOSTS, name + ".c"), "-L", PKG, "-lb200rg", f"-Wl,-rpath,{PKG}"])
    return exe


def write_bmp_file(path, payload, w, h):
    """an 8-bit BMP the reference's reader accepts: it only looks at offset (10), width (18) and height (22)"""
    head = bytearray(1078)
    head[0:2] = b"BM"
    head[2:6] = (w * h + 1078).to_bytes(4, "little"); head[10:14] = (1078).to_bytes(4, "little")
    head[14:18] = (40).to_bytes(4, "little"); head[18:22] = w.to_bytes(4, "little"); head[22:26] = h.to_bytes(4, "little")
    head[26:28] = (1).to_bytes(2, "little"); head[28:30] = (8).to_bytes(2, "little")
    with open(path, "wb") as f:
        f.write(head); f.write(payload.tobytes())


def test_c_hosts_compile_and_link(tmp_path):
    for name in ("region_host", "raycast_host"):
        exe = build(str(tmp_path), name)
        assert os.access(exe, os.X_OK)


@pytest.mark.gpu
@pytest.mark.parametrize("k", [1, 3])
def test_region_host_writes_the_reference_picture(tmp_path, golden, oracle, k):
    exe = build(str(tmp_path), "region_host")
    pic = str(tmp_path / f"pic{k}.bmp")
    write_bmp_file(pic, golden["arrays"][f"pic{k}"], 512, 512)
    out = subprocess.run([exe, pic], cwd=str(tmp_path), capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr
    want = oracle.grow(golden["arrays"][f"pic{k}"], oracle.seeds_2d(512, 512), 2)
    assert int(re.search(r"region pixels: (\d+)", out.stdout).group(1)) == int(want.sum())
    data = open(tmp_path / "out.bmp", "rb").read()
    assert len(data) == 1078 + 512 * 512 + 2
    assert hashlib.sha256(data[1078:1078 + 512 * 512]).hexdigest() == golden["pics"][f"pic{k}"]["out_sha256"]


@pytest.mark.gpu
def test_raycast_host_on_the_reference_volume(tmp_path, oracle, golden):
    exe = build(str(tmp_path), "raycast_host")
    vol = oracle.create_data_ref()
    raw = str(tmp_path / "vol.raw")
    vol.tofile(raw)
    out = subprocess.run([exe, raw], cwd=str(tmp_path), capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr
    m = re.search(r"region voxels: (\d+) fnv1a64: ([0-9a-f]+)", out.stdout)
    assert int(m.group(1)) == golden["c1"]["region_voxels"] and m.group(2) == golden["c1"]["mask_fnv1a64"]
    data = open(tmp_path / "out.bmp", "rb").read()
    pic = np.frombuffer(data[1078:1078 + 512 * 512], np.uint8)
    assert len(data) == 1078 + 512 * 512 + 2 and int((pic > 0).sum()) > 1000
