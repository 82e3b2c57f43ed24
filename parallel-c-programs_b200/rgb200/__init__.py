"""rgb200 -- thin host-side mirror of the C ABI of libb200rg.so (include/rg.h, include/bmp.h).

The product is the CUDA library; this package only loads it with ctypes, mirrors the reference's
entry points by name (read_bmp / write_bmp / grow_region_gpu / grow_region_gpu_shared, and the
parametrised grow()) and drives the multi-rank z-slab protocol with torch.distributed
(SlabGrower).  There is no CPU fallback: if the library is missing or no CUDA device is present,
calls raise.
"""
import ctypes
import os

import numpy as np

_PKG_DIR = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB_PATH = os.environ.get("RG_LIB") or os.path.join(_PKG_DIR, "libb200rg.so")   # RG_LIB: tuning experiments only

RG_OK = 0
_c = ctypes
_vp, _i64, _i32, _u32, _u64 = _c.c_void_p, _c.c_int64, _c.c_int, _c.c_uint32, _c.c_uint64

# name -> (restype, argtypes); every symbol include/rg.h and include/bmp.h declare
SIGNATURES = {
    "write_bmp": (None, [_vp, _i32, _i32]),
    "read_bmp": (_vp, [_c.c_char_p]),
    "rg_read_bmp_dims": (_vp, [_c.c_char_p, _c.POINTER(_i32), _c.POINTER(_i32)]),
    "rg_write_bmp_path": (_i32, [_c.c_char_p, _vp, _i32, _i32]),
    "grow_region_gpu": (_vp, [_vp]),
    "grow_region_gpu_shared": (_vp, [_vp]),
    "grow_region_serial": (_vp, [_vp]),
    "raycast_gpu": (_vp, [_vp, _vp]),
    "raycast_gpu_texture": (_vp, [_vp, _vp]),
    "rg_grow_region_2d": (_i32, [_vp, _i32, _i32, _vp, _i32]),
    "rg_grow": (_i32, [_vp, _i64, _i64, _i64, _vp, _i32, _i32, _vp]),
    "rg_grow_sparse": (_i32, [_vp, _i64, _i64, _i64, _vp, _i32, _i32, _vp]),
    "rg_host_stats": (_i32, [_vp]),
    "rg_multi_stats": (_i32, [_vp]),
    "rg_grow_multi_grid": (_i32, [_vp, _i64, _i64, _vp, _i32, _i32, _vp, _vp, _i32, _i32, _i32, _vp]),
    "rg_block_label": (_i32, [_vp, _vp, _i64, _i64, _i32, _vp, _i32, _i32, _vp, _vp]),
    "rg_block_export": (_i32, [_vp, _i32, _vp, _vp, _vp]),
    "rg_block_resolve_union": (_i32, [_vp, _vp, _vp, _u32, _vp, _vp, _u32, _vp, _u32, _i32, _vp]),
    "rg_grow_multi": (_i32, [_vp, _i64, _i64, _i64, _vp, _i32, _i32, _vp, _vp, _i32, _i32, _vp]),
    "rg_ctx_create": (_i32, [_i32, _c.POINTER(_vp)]),
    "rg_ctx_destroy": (None, [_vp]),
    "rg_workspace_bytes": (_i64, [_i64, _i64, _i64, _i32]),
    "rg_ctx_reserve": (_i32, [_vp, _i64, _i64, _i64, _i32]),
    "rg_grow_device": (_i32, [_vp, _vp, _i64, _i64, _i64, _vp, _i32, _i32, _vp, _vp]),
    "rg_grow_device_noalloc": (_i32, [_vp, _i64, _i64, _i64, _vp, _i32, _i32, _vp, _vp]),
    "rg_noalloc_scratch_bytes": (_i64, []),
    "rg_noalloc_status": (_i32, [_vp, _vp]),
    "rg_ctx_status": (_i32, [_vp, _vp]),
    "rg_ctx_profile": (_i32, [_vp, _i32]),
    "rg_ctx_profile_read": (_i32, [_vp, _vp]),
    "rg_apply_mask_device": (_i32, [_vp, _vp, _vp, _i64, _vp]),
    "rg_raycast_device": (_i32, [_vp, _vp, _i32, _i32, _c.c_float, _vp, _vp]),
    "rg_last_error": (_c.c_char_p, []),
    "rg_slab_label": (_i32, [_vp, _vp, _i64, _i64, _i64, _i32, _i32, _vp, _i32, _i32, _vp, _vp]),
    "rg_slab_interface_capacity": (_i64, [_i64, _i64, _i64]),
    "rg_slab_export": (_i32, [_vp, _i32, _vp, _vp, _vp]),
    "rg_slab_resolve_init": (_i32, [_vp, _vp, _vp, _u32, _u32, _vp]),
    "rg_slab_resolve_union": (_i32, [_vp, _vp, _vp, _u32, _vp, _u32, _u32, _vp]),
    "rg_slab_resolve_apply": (_i32, [_vp, _vp, _vp, _u32, _u32, _vp]),
    "rg_slab_resolve": (_i32, [_vp, _vp, _vp, _u32, _vp, _i32, _i32, _vp]),
    "rg_slab_export_packed": (_i32, [_vp, _vp, _u32, _vp]),
    "rg_slab_resolve_packed": (_i32, [_vp, _vp, _u32, _vp, _u32, _i32, _i32, _vp]),
    "rg_slab_finish": (_i32, [_vp, _vp]),
    "rg_slab_resolve_peers": (_i32, [_vp, _vp, _u32, _vp, _u32, _i32, _i32, _vp]),
    "rg_ipc_alloc": (_i32, [_i64, _c.POINTER(_vp), _vp]),
    "rg_ipc_open": (_i32, [_vp, _c.POINTER(_vp)]),
    "rg_ipc_close": (_i32, [_vp]),
    "rg_ipc_free": (_i32, [_vp]),
    "rg_synth_device": (_i32, [_i32, _vp, _i64, _u64, _vp]),
    "rg_synth_slab_device": (_i32, [_i32, _vp, _i64, _i64, _i64, _u64, _vp]),
    "rg_synth_seeds": (_i32, [_i32, _i64, _u64, _vp, _i32]),
}


class RGError(RuntimeError):
    pass


def bind(lib, names=None):
    """attach the C signatures to a loaded library (the product, or the test emulator)"""
    for name, (res, args) in SIGNATURES.items():
        if names is not None and name not in names:
            continue
        fn = getattr(lib, name, None)
        if fn is None:
            if os.environ.get("RG_LIB"):                     # tuning: an older build may lack the newest entry points
                continue
            raise RGError(f"{LIB_PATH} does not export {name}")
        fn.restype = res
        fn.argtypes = args
    return lib


_lib = None


def load_library(path=None):
    """Load libb200rg.so.  Fails loudly: there is no other implementation to fall back to."""
    global _lib
    if path is None and _lib is not None:
        return _lib
    p = path or LIB_PATH
    if not os.path.exists(p):
        raise RGError(f"{p} not found: build it with `make -C {_PKG_DIR}` (or __graft_entry__.build()); "
                      "rgb200 has no CPU fallback")
    lib = bind(ctypes.CDLL(p))
    if path is None:
        _lib = lib
    return lib


def _check(lib, rc, what):
    if rc != RG_OK:
        msg = lib.rg_last_error()
        raise RGError(f"{what} failed with {rc}: {msg.decode() if msg else ''}")


def _dims(a):
    if a.ndim == 2:
        return a.shape[1], a.shape[0], 1
    if a.ndim == 3:
        return a.shape[2], a.shape[1], a.shape[0]
    raise ValueError("image must be (ny,nx) or (nz,ny,nx) uint8")


def _seeds_np(seeds):
    s = np.ascontiguousarray(np.asarray(seeds, np.int32).reshape(-1, 3))
    return s


# ---------------------------------------------------------------- reference-shaped host calls
def grow(image, seeds, threshold):
    """grow(image, seeds, threshold) -> mask.  Host arrays in and out (rg_grow): the call the
    reference's main()s make, with size / seeds / threshold as arguments instead of #defines."""
    lib = load_library()
    img = np.ascontiguousarray(image, np.uint8)
    nx, ny, nz = _dims(img)
    s = _seeds_np(seeds)
    mask = np.empty(img.shape, np.uint8)
    _check(lib, lib.rg_grow(img.ctypes.data, nx, ny, nz, s.ctypes.data if len(s) else None, len(s), int(threshold),
                            mask.ctypes.data), "rg_grow")
    return mask


def _grow_512(fn_name, data):
    lib = load_library()
    d = np.ascontiguousarray(data, np.uint8)
    if d.size != 512 ** 3:
        raise ValueError("the reference-shaped entry points take a 512^3 volume (raycast.cu:11)")
    ptr = getattr(lib, fn_name)(d.ctypes.data)
    if not ptr:
        raise RGError(f"{fn_name} failed: {lib.rg_last_error().decode()}")
    out = np.ctypeslib.as_array(ctypes.cast(ptr, ctypes.POINTER(ctypes.c_uint8)), (512, 512, 512)).copy()
    ctypes.CDLL(None).free(ctypes.c_void_p(ptr))
    return out


def grow_region_gpu(data):
    """raycast.cu:702-756"""
    return _grow_512("grow_region_gpu", data)


def grow_region_gpu_shared(data):
    """raycast.cu:759-822 (what the reference's main() calls, :833)"""
    return _grow_512("grow_region_gpu_shared", data)


def _raycast_512(fn_name, data, region):
    lib = load_library()
    d = np.ascontiguousarray(data, np.uint8)
    r = np.ascontiguousarray(region, np.uint8)
    if d.size != 512 ** 3 or r.size != 512 ** 3:
        raise ValueError("the reference-shaped entry points take 512^3 volumes (raycast.cu:11)")
    ptr = getattr(lib, fn_name)(d.ctypes.data, r.ctypes.data)
    if not ptr:
        raise RGError(f"{fn_name} failed")
    out = np.ctypeslib.as_array(ctypes.cast(ptr, ctypes.POINTER(ctypes.c_uint8)), (512, 512)).copy()
    ctypes.CDLL(None).free(ctypes.c_void_p(ptr))
    return out


def raycast_gpu(data, region):
    """raycast.cu:436-470"""
    return _raycast_512("raycast_gpu", data, region)


def raycast_gpu_texture(data, region):
    """raycast.cu:472-531 (what the reference's main() calls, :842)"""
    return _raycast_512("raycast_gpu_texture", data, region)


def raycast_device(d_data, d_region, image_dim, camera=1000.0, stream=None, lib=None):
    """rg_raycast_device on torch CUDA uint8 tensors (data_dim^3 each); returns the image tensor."""
    import torch
    lib = lib or load_library()
    D = d_data.shape[0]
    assert d_data.is_cuda and d_region.is_cuda and d_data.dtype == torch.uint8 and d_region.dtype == torch.uint8
    assert tuple(d_data.shape) == (D, D, D) and tuple(d_region.shape) == (D, D, D)
    assert d_data.is_contiguous() and d_region.is_contiguous()
    img = torch.empty((image_dim, image_dim), dtype=torch.uint8, device=d_data.device)
    with torch.cuda.device(d_data.device):
        _check(lib, lib.rg_raycast_device(d_data.data_ptr(), d_region.data_ptr(), D, image_dim, float(camera),
                                          img.data_ptr(), _stream_ptr(stream)), "rg_raycast_device")
    return img


def grow_region_2d(image, apply=False):
    """region.c:493-533 (+ write_image's masking, :572-580, when apply=True) on a whole image.
    Returns (mask, image_out)."""
    lib = load_library()
    img = np.ascontiguousarray(image, np.uint8).copy()
    h, w = img.shape
    mask = np.empty_like(img)
    _check(lib, lib.rg_grow_region_2d(img.ctypes.data, w, h, mask.ctypes.data, 1 if apply else 0), "rg_grow_region_2d")
    return mask, img


def read_bmp(path):
    """bmp.c:49-71 -> (h, w) uint8 array"""
    lib = load_library()
    w, h = ctypes.c_int(0), ctypes.c_int(0)
    ptr = lib.rg_read_bmp_dims(os.fsencode(path), ctypes.byref(w), ctypes.byref(h))
    if not ptr:
        raise RGError(f"cannot read {path}")
    out = np.ctypeslib.as_array(ctypes.cast(ptr, ctypes.POINTER(ctypes.c_uint8)), (h.value, w.value)).copy()
    ctypes.CDLL(None).free(ctypes.c_void_p(ptr))
    return out


def write_bmp(path, data):
    """bmp.c:6-47 with a path (write_bmp itself always writes ./out.bmp)"""
    lib = load_library()
    d = np.ascontiguousarray(data, np.uint8)
    if lib.rg_write_bmp_path(os.fsencode(path), d.ctypes.data, d.shape[1], d.shape[0]) != 0:
        raise RGError(f"cannot write {path}")


def synth_seeds(kind, n, seed=0, lib=None):
    lib = lib or load_library()
    buf = np.zeros((64, 3), np.int32)
    cnt = lib.rg_synth_seeds(kind, n, seed, buf.ctypes.data, 64)
    if cnt < 0:
        raise RGError("rg_synth_seeds")
    return buf[:cnt].copy()


# ---------------------------------------------------------------- device-resident context (torch tensors)
def _stream_ptr(stream):
    if stream is None:
        import torch
        return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    return ctypes.c_void_p(stream)


class Context:
    """Per-device engine context.  Tensors are torch CUDA uint8 / int32 tensors; all work is
    enqueued on torch's current stream (or `stream`), nothing synchronises."""

    def __init__(self, device=0, lib=None):
        self.lib = lib or load_library()
        h = ctypes.c_void_p()
        _check(self.lib, self.lib.rg_ctx_create(int(device), ctypes.byref(h)), "rg_ctx_create")
        self.handle = h
        self.device = int(device)

    def close(self):
        if self.handle:
            self.lib.rg_ctx_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def reserve(self, nx, ny, nz, nseeds=64):
        _check(self.lib, self.lib.rg_ctx_reserve(self.handle, nx, ny, nz, nseeds), "rg_ctx_reserve")

    def grow(self, image, seeds_dev, threshold, out, stream=None):
        """image: CUDA uint8 (nz,ny,nx) or (ny,nx); seeds_dev: CUDA int32 (k,3); out: CUDA uint8 like image"""
        nx, ny, nz = _dims(image)
        k = 0 if seeds_dev is None else int(seeds_dev.shape[0])
        self._check_tensors(image, out, seeds_dev)
        _check(self.lib, self.lib.rg_grow_device(self.handle, image.data_ptr(), nx, ny, nz,
                                                 seeds_dev.data_ptr() if k else None, k, int(threshold), out.data_ptr(),
                                                 _stream_ptr(stream)), "rg_grow_device")
        return out

    def _check_tensors(self, image, out, seeds_dev=None):
        """the C ABI takes raw device pointers: a sliced, mistyped or misplaced tensor would be read / written out of bounds"""
        import torch
        for name, t in (("image", image), ("out", out)):
            if t is None:
                continue
            if not (t.is_cuda and t.device.index == self.device and t.dtype == torch.uint8 and t.is_contiguous()):
                raise ValueError(f"{name} must be a contiguous uint8 CUDA tensor on device {self.device}")
        if out is not None and tuple(out.shape) != tuple(image.shape):
            raise ValueError("out must have the image's shape")
        if seeds_dev is not None and int(seeds_dev.shape[0]):
            if not (seeds_dev.is_cuda and seeds_dev.device.index == self.device and seeds_dev.dtype == torch.int32
                    and seeds_dev.is_contiguous() and seeds_dev.dim() == 2 and seeds_dev.shape[1] == 3):
                raise ValueError(f"seeds must be a contiguous int32 (k, 3) CUDA tensor on device {self.device}")

    def status(self):
        st = (ctypes.c_uint32 * 4)()
        rc = self.lib.rg_ctx_status(self.handle, st)
        _check(self.lib, rc, "rg_ctx_status")
        return {"nodes": st[0], "status": st[1], "complex_tiles": st[2], "region_tiles": st[3]}

    def profile(self, enable=True):
        _check(self.lib, self.lib.rg_ctx_profile(self.handle, 1 if enable else 0), "rg_ctx_profile")

    def profile_read(self):
        ms = (ctypes.c_float * 6)()
        _check(self.lib, self.lib.rg_ctx_profile_read(self.handle, ms), "rg_ctx_profile_read")
        return {"seed_sort": ms[0], "pass_a_fast": ms[1], "pass_a_full": ms[2], "merge": ms[3], "flag_tiles": ms[4], "pass_c": ms[5]}

    def apply_mask(self, image, mask, stream=None):
        self._check_tensors(image, mask)
        _check(self.lib, self.lib.rg_apply_mask_device(self.handle, image.data_ptr(), mask.data_ptr(), image.numel(),
                                                       _stream_ptr(stream)), "rg_apply_mask_device")

    def synth(self, kind, out, n, seed=0, z_begin=None, z_end=None, stream=None):
        if z_begin is None:
            rc = self.lib.rg_synth_device(kind, out.data_ptr(), n, seed, _stream_ptr(stream))
        else:
            rc = self.lib.rg_synth_slab_device(kind, out.data_ptr(), n, z_begin, z_end, seed, _stream_ptr(stream))
        _check(self.lib, rc, "rg_synth_device")
        return out


from .slab import SlabGrower, slab_bounds  # noqa: E402,F401
