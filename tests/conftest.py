"""pytest configuration: markers, paths, and ctypes handles to the oracle (checker) library."""
import ctypes, json, os, subprocess, sys
import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "parallel-c-programs_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


u8p = ctypes.c_void_p


class Oracle:
    """ctypes view of oracle/liboracle.so -- the CPU checker.  Test infrastructure only."""

    def __init__(self):
        so = os.path.join(ROOT, "oracle", "liboracle.so")
        if not os.path.exists(so):
            subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "liboracle.so"])
        L = self.lib = ctypes.CDLL(so)
        L.oracle_grow.restype = ctypes.c_int64
        L.oracle_grow.argtypes = [u8p, ctypes.c_int64, ctypes.c_int64, ctypes.c_int64, ctypes.c_void_p,
                                  ctypes.c_int, ctypes.c_int, ctypes.c_int, u8p]
        L.oracle_grow_par.restype = ctypes.c_int64
        L.oracle_grow_par.argtypes = [u8p, ctypes.c_int64, ctypes.c_int64, ctypes.c_int64, ctypes.c_void_p,
                                      ctypes.c_int, ctypes.c_int, ctypes.c_int, u8p, ctypes.c_void_p]
        L.oracle_fnv1a64.restype = ctypes.c_uint64
        L.oracle_fnv1a64.argtypes = [u8p, ctypes.c_int64]
        L.oracle_create_data_ref.argtypes = [u8p]
        L.oracle_apply_mask.argtypes = [u8p, u8p, ctypes.c_int64]
        L.oracle_add_seeds_2d.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
        L.oracle_synth_exercise.argtypes = [u8p, ctypes.c_int64, ctypes.c_uint64]
        L.oracle_synth_c4.argtypes = [u8p, ctypes.c_int64, ctypes.c_uint64]
        L.oracle_seeds_c4.argtypes = [ctypes.c_int64, ctypes.c_uint64, ctypes.c_void_p]
        L.oracle_synth_c2.argtypes = [u8p, ctypes.c_int64, ctypes.c_uint64]
        L.oracle_seeds_c2.argtypes = [ctypes.c_int64, ctypes.c_void_p]
        L.oracle_synth_c3.argtypes = [u8p, ctypes.c_int64]
        L.oracle_raycast.argtypes = [u8p, u8p, ctypes.c_int, ctypes.c_int, ctypes.c_float, u8p]
        L.oracle_c3_count.restype = ctypes.c_int64
        L.oracle_c3_count.argtypes = [ctypes.c_int64]

    def grow(self, image, seeds, threshold, conn=None):
        image = np.ascontiguousarray(image, np.uint8)
        if image.ndim == 2:
            nz, (ny, nx) = 1, image.shape
        else:
            nz, ny, nx = image.shape
        if conn is None:
            conn = 4 if image.ndim == 2 else 6
        seeds = np.ascontiguousarray(np.asarray(seeds, np.int32).reshape(-1, 3))
        mask = np.empty(image.shape, np.uint8)
        n = self.lib.oracle_grow(image.ctypes.data, nx, ny, nz, seeds.ctypes.data, len(seeds), int(threshold), conn,
                                 mask.ctypes.data)
        assert n >= 0
        return mask

    def grow_par(self, image, seeds, threshold, nthreads):
        image = np.ascontiguousarray(image, np.uint8)
        if image.ndim == 2:
            nz, (ny, nx) = 1, image.shape
        else:
            nz, ny, nx = image.shape
        seeds = np.ascontiguousarray(np.asarray(seeds, np.int32).reshape(-1, 3))
        mask = np.empty(image.shape, np.uint8)
        rounds = ctypes.c_int(0)
        n = self.lib.oracle_grow_par(image.ctypes.data, nx, ny, nz, seeds.ctypes.data, len(seeds), int(threshold),
                                     nthreads, mask.ctypes.data, ctypes.byref(rounds))
        assert n >= 0
        return mask, rounds.value

    def raycast(self, data, region, image_dim, cam=1000.0):
        data = np.ascontiguousarray(data, np.uint8); region = np.ascontiguousarray(region, np.uint8)
        D = data.shape[0]
        out = np.empty((image_dim, image_dim), np.uint8)
        self.lib.oracle_raycast(data.ctypes.data, region.ctypes.data, D, image_dim, cam, out.ctypes.data)
        return out

    def fnv(self, a):
        a = np.ascontiguousarray(a)
        return "%016x" % self.lib.oracle_fnv1a64(a.ctypes.data, a.size)

    def seeds_2d(self, w, h):
        s = np.zeros((4, 3), np.int32)
        self.lib.oracle_add_seeds_2d(w, h, s.ctypes.data)
        return s

    def create_data_ref(self):
        v = np.empty((512, 512, 512), np.uint8)
        self.lib.oracle_create_data_ref(v.ctypes.data)
        return v

    def synth(self, kind, n, seed=0):
        if kind == "c2":
            a = np.empty((n, n), np.uint8); self.lib.oracle_synth_c2(a.ctypes.data, n, seed)
            s = np.zeros((64, 3), np.int32); self.lib.oracle_seeds_c2(n, s.ctypes.data)
            return a, s, 2
        a = np.empty((n, n, n), np.uint8)
        if kind == "c1h":
            self.lib.oracle_synth_exercise(a.ctypes.data, n, seed)
            s = np.array([[50 * n // 512, 300 * n // 512, 300 * n // 512]], np.int32)
        elif kind == "c3":
            self.lib.oracle_synth_c3(a.ctypes.data, n)
            s = np.array([[0, 0, 0]], np.int32)
        elif kind == "c4":
            self.lib.oracle_synth_c4(a.ctypes.data, n, seed)
            s = np.zeros((16, 3), np.int32); self.lib.oracle_seeds_c4(n, seed, s.ctypes.data)
        else:
            raise ValueError(kind)
        return a, s, 1


@pytest.fixture(scope="session")
def oracle():
    return Oracle()


@pytest.fixture(scope="session")
def golden():
    g = json.load(open(os.path.join(GOLDEN, "golden.json")))
    g["arrays"] = np.load(os.path.join(GOLDEN, "pics.npz"))
    return g


def cc_model(image, seeds, threshold):
    """Independent model: scipy connected components on the |delta|<T neighbour graph."""
    import scipy.sparse as sp
    from scipy.sparse.csgraph import connected_components
    a = np.asarray(image, np.int16)
    shape = a.shape
    idx = np.arange(a.size).reshape(shape)
    rows, cols = [], []
    for ax in range(a.ndim):
        sl_lo = [slice(None)] * a.ndim; sl_hi = [slice(None)] * a.ndim
        sl_lo[ax] = slice(0, -1); sl_hi[ax] = slice(1, None)
        e = np.abs(a[tuple(sl_lo)] - a[tuple(sl_hi)]) < threshold
        rows.append(idx[tuple(sl_lo)][e]); cols.append(idx[tuple(sl_hi)][e])
    r = np.concatenate(rows); c = np.concatenate(cols)
    g = sp.coo_matrix((np.ones(len(r), np.uint8), (r, c)), shape=(a.size, a.size))
    _, lab = connected_components(g, directed=False)
    lab = lab.reshape(shape)
    keep = set()
    for s in np.asarray(seeds).reshape(-1, 3):
        x, y, z = (int(v) for v in s)
        if a.ndim == 2:
            if 0 <= x < shape[1] and 0 <= y < shape[0] and z == 0:
                keep.add(lab[y, x])
        elif 0 <= x < shape[2] and 0 <= y < shape[1] and 0 <= z < shape[0]:
            keep.add(lab[z, y, x])
    return np.isin(lab, list(keep)).astype(np.uint8)
