#!/usr/bin/env python3
"""Regenerate tests/golden/ from the UNMODIFIED reference (run in the build container only).

Needs /root/reference and `make -C oracle` (which builds oracle/_ref/region from the reference's
own region.c + bmp.c against oracle/mpi_shim, and oracle/_ref/libref_serial.so from the
reference's plain-C grow_region_serial/create_data).  Writes:

  pics.npz        the four 512x512 input payloads exactly as the reference's read_bmp returns
                  them (2-mpi-region-growing/bmp.c:49-71: width*height bytes from the pixel offset)
                  and the payloads region.c wrote for them at np=4 (out.bmp bytes 1078..+262144)
  golden.json     sha256 / counts of the above, the committed out.bmp payload hash, and the
                  C1 pins (reference 512^3 volume + grow_region_serial mask: count, fnv1a-64)

The GPU box has no /root/reference: tests read only these files.
"""
import ctypes, hashlib, json, os, struct, subprocess, sys, tempfile
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))


def read_bmp_like_reference(path):
    b = open(path, "rb").read()
    w, = struct.unpack_from("<i", b, 18)
    h, = struct.unpack_from("<i", b, 22)
    off, = struct.unpack_from("<i", b, 10)
    return np.frombuffer(b, np.uint8, w * h, off).reshape(h, w).copy(), off


def fnv1a64(a):
    lib = ctypes.CDLL(os.path.join(ROOT, "oracle", "liboracle.so"))
    lib.oracle_fnv1a64.restype = ctypes.c_uint64
    lib.oracle_fnv1a64.argtypes = [ctypes.c_void_p, ctypes.c_int64]
    a = np.ascontiguousarray(a)
    return "%016x" % lib.oracle_fnv1a64(a.ctypes.data, a.size)


def main():
    subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle")])
    region = os.path.join(ROOT, "oracle", "_ref", "region")
    gold = {"source": "unmodified reference built by oracle/Makefile", "pics": {}}
    arrays = {}
    for k in (1, 2, 3, 4):
        src = f"{REF}/2-mpi-region-growing/pic{k}.bmp"
        img, off = read_bmp_like_reference(src)
        outs = {}
        for np_ in (1, 4, 16):
            with tempfile.TemporaryDirectory() as d:
                subprocess.check_call([region, src], cwd=d, env=dict(os.environ, SHIM_NP=str(np_)))
                ob = open(os.path.join(d, "out.bmp"), "rb").read()
            assert len(ob) == 512 * 512 + 1078 + 2
            outs[np_] = np.frombuffer(ob, np.uint8, 512 * 512, 1078).reshape(512, 512).copy()
        assert (outs[1] == outs[4]).all() and (outs[4] == outs[16]).all()
        out = outs[4]
        arrays[f"pic{k}"] = img
        arrays[f"out{k}"] = out
        gold["pics"][f"pic{k}"] = {
            "pixel_offset": off,
            "in_sha256": hashlib.sha256(img.tobytes()).hexdigest(),
            "out_sha256": hashlib.sha256(out.tobytes()).hexdigest(),
            "zeroed_px": int((out == 0).sum()),
            "in_zero_px": int((img == 0).sum()),
        }
    committed = open(f"{REF}/2-mpi-region-growing/out.bmp", "rb").read()
    gold["committed_out_bmp_payload_sha256"] = hashlib.sha256(committed[1078:1078 + 262144]).hexdigest()
    assert gold["committed_out_bmp_payload_sha256"] == gold["pics"]["pic1"]["out_sha256"]
    np.savez_compressed(os.path.join(HERE, "pics.npz"), **arrays)

    # C1: the reference's own volume and serial grower
    ref = ctypes.CDLL(os.path.join(ROOT, "oracle", "_ref", "libref_serial.so"))
    ref.ref_create_data.restype = ctypes.c_void_p
    ref.ref_grow_region_serial.restype = ctypes.c_void_p
    ref.ref_grow_region_serial.argtypes = [ctypes.c_void_p]
    ref.ref_free.argtypes = [ctypes.c_void_p]
    n = 512 ** 3
    dptr = ref.ref_create_data()
    vol = np.ctypeslib.as_array(ctypes.cast(dptr, ctypes.POINTER(ctypes.c_uint8)), (n,))
    mptr = ref.ref_grow_region_serial(dptr)
    msk = np.ctypeslib.as_array(ctypes.cast(mptr, ctypes.POINTER(ctypes.c_uint8)), (n,))
    zz, yy, xx = np.nonzero(msk.reshape(512, 512, 512))
    gold["c1"] = {
        "volume_fnv1a64": fnv1a64(vol), "mask_fnv1a64": fnv1a64(msk), "region_voxels": int(msk.sum()),
        "bbox_xyz": [int(xx.min()), int(xx.max()), int(yy.min()), int(yy.max()), int(zz.min()), int(zz.max())],
        "mask_values": sorted(int(v) for v in np.unique(msk)),
        "seed": [50, 300, 300], "threshold": 1,
    }
    # the ray-caster that consumes the mask (raycast_serial, IMAGE_DIM = 64 as in 6-opencl-region-growing/raycast.c:16)
    ref.ref_raycast_serial.restype = ctypes.c_void_p
    ref.ref_raycast_serial.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
    iptr = ref.ref_raycast_serial(dptr, mptr)
    ray = np.ctypeslib.as_array(ctypes.cast(iptr, ctypes.POINTER(ctypes.c_uint8)), (64, 64)).copy()
    gold["raycast64_sha256"] = hashlib.sha256(ray.tobytes()).hexdigest()
    gold["raycast64_nonzero_px"] = int((ray > 0).sum())
    arrays["ray64"] = ray
    np.savez_compressed(os.path.join(HERE, "pics.npz"), **arrays)
    ref.ref_free(iptr)
    ref.ref_free(mptr); ref.ref_free(dptr)
    json.dump(gold, open(os.path.join(HERE, "golden.json"), "w"), indent=1, sort_keys=True)
    print(json.dumps(gold, indent=1, sort_keys=True))


if __name__ == "__main__":
    sys.exit(main())
