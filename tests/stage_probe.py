"""tuning aid: pageable host -> device staging throughput of rg_grow_sparse for a few copy-thread counts (run on the GPU box)"""
import os, subprocess, sys, time
if len(sys.argv) > 1:
    import numpy as np, ctypes
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "parallel-c-programs_b200"))
    import rgb200
    lib = rgb200.load_library()
    n = 1536
    img = np.zeros((n, n, n), np.uint8); img[::7] = 3
    out = np.zeros_like(img)
    seeds = np.array([[1, 1, 1]], np.int32)
    for i in range(3):
        t0 = time.perf_counter()
        rc = lib.rg_grow_sparse(img.ctypes.data, n, n, n, seeds.ctypes.data, 1, 1, out.ctypes.data)
        dt = time.perf_counter() - t0
    print(sys.argv[1], "threads:", rc, "%.1f ms" % (dt * 1e3), "%.1f GB/s in" % (img.size / dt / 1e9))
else:
    for t in ("0", "2", "4", "8", "12", "16"):
        env = dict(os.environ, RG_STAGE_THREADS=t)
        if t == "0":
            env["RG_NO_STAGING"] = "1"
        subprocess.run([sys.executable, __file__, t], env=env)
