"""Multi-rank slab driver: one process per GPU, torch.distributed for the plumbing.  A volume is cut into z-slabs,
an image (nz == 1) into row slabs -- either way each rank owns a contiguous range of the slowest axis.

Mirrors the distributed half of the reference's 2-D MPI program (2-mpi-region-growing/region.c):
  distribute_image_halo (:145-246)  -> exchange_image_halos(): one static boundary-plane swap
  exchange + add_halo_to_stack + finished (:250-387, :435-440), repeated until no rank progresses
                                    -> ONE all-gather of interface root lists + a redundant
                                       union-find on every rank (rg_slab_* in include/rg.h)
The per-rank compute is the same CUDA engine as the single-GPU path.  `lib` is libb200rg.so; the
CPU logic tests pass the sequential emulator (tests/emu) with CPU tensors and the gloo backend --
this module never picks a CPU implementation on its own.
"""
import ctypes

import numpy as np


def slab_bounds(nz, rank, world):
    """slab k = z in [floor(k*nz/G), floor((k+1)*nz/G))  """
    return nz * rank // world, nz * (rank + 1) // world


class SlabGrower:
    def __init__(self, lib, ctx_handle, nx, ny, nz, rank, world, device, group=None, use_cuda_stream=True):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.lib, self.ctx = lib, ctx_handle
        self.nx, self.ny, self.nz = int(nx), int(ny), int(nz)
        self.rank, self.world, self.group = int(rank), int(world), group
        self.device = device
        self.use_cuda_stream = use_cuda_stream
        # the slab axis: z for a volume, y (rows) for an image; z0 / z1 / nzl are positions along THAT axis and
        # "plane" below is one unit of it (a plane of nx*ny voxels, or a row of nx pixels)
        self.rows = self.nz == 1
        self.axis = 1 if self.rows else 2
        self.plane = self.nx if self.rows else self.nx * self.ny
        self.z0, self.z1 = slab_bounds(self.ny if self.rows else self.nz, self.rank, self.world)
        self.nzl = self.z1 - self.z0
        if self.world > 1 and self.nzl < 2:
            raise ValueError("each rank needs at least 2 planes (rows)")
        self.has_lower = self.rank > 0
        self.has_upper = self.rank < self.world - 1
        plane = self.plane
        shape = (self.nzl, self.nx) if self.rows else (self.nzl, self.ny, self.nx)
        # slab with one halo plane on each side (present even when unused: keeps offsets simple)
        self.buf = torch.zeros((self.nzl + 2) * plane, dtype=torch.uint8, device=device)
        self.image = self.buf[plane:(self.nzl + 1) * plane].view(*shape)
        self.mask = torch.empty(shape, dtype=torch.uint8, device=device)
        if self.world > 1:
            cap = int(lib.rg_slab_interface_capacity(self.nx, self.ny, 1 if self.rows else self.nzl))
            self.cap = cap
            self.lists = torch.zeros((2, cap), dtype=torch.int32, device=device)
            self.counts = torch.zeros(4, dtype=torch.int32, device=device)
            self.all_counts = torch.zeros(self.world * 4, dtype=torch.int32, device=device)   # flat: gloo wants 1-D
        self.seeds_dev, self.nseeds = None, 0
        self.rounds = 0
        self._timing = None
        self._cap = None            # capacities learnt from the last synchronous exchange (optimistic mode)
        self._peers, self._peer_handles, self._peer_turn = None, [], 0     # export buffers mapped by the peers (CUDA IPC), see _map_peer_buffers
        self._xpar_cap = 0
        self.optimistic_steps = 0

    # ---- helpers
    def _stream(self):
        if self.use_cuda_stream:
            return ctypes.c_void_p(self.torch.cuda.current_stream().cuda_stream)
        return None

    def _ck(self, rc, what):
        if rc != 0:
            msg = self.lib.rg_last_error()
            raise RuntimeError(f"{what} failed with {rc}: {msg.decode() if msg else ''}")

    def set_seeds(self, seeds_global):
        """keep the seeds that fall into this slab, in slab-local coordinates (add_seeds, region.c:450-490:
        only the owning rank takes a seed)"""
        s = np.asarray(seeds_global, np.int64).reshape(-1, 3)
        a = self.axis
        keep = (s[:, a] >= self.z0) & (s[:, a] < self.z1)
        loc = s[keep].copy()
        loc[:, a] -= self.z0
        self.nseeds = int(len(loc))
        self.seeds_dev = self.torch.tensor(loc.astype(np.int32).reshape(-1, 3), dtype=self.torch.int32,
                                           device=self.device).contiguous() if self.nseeds else None

    def exchange_image_halos(self):
        """static, once per image: my top plane -> upper rank's lower halo, my bottom plane -> lower
        rank's upper halo (distribute_image_halo, region.c:145-246)"""
        if self.world == 1:
            return
        dist, plane = self.dist, self.plane
        ops = []
        lo_halo = self.buf[:plane]
        hi_halo = self.buf[(self.nzl + 1) * plane:]
        first = self.buf[plane:2 * plane]
        last = self.buf[self.nzl * plane:(self.nzl + 1) * plane]
        if self.has_upper:
            ops.append(dist.P2POp(dist.isend, last, self.rank + 1, self.group))
            ops.append(dist.P2POp(dist.irecv, hi_halo, self.rank + 1, self.group))
        if self.has_lower:
            ops.append(dist.P2POp(dist.isend, first, self.rank - 1, self.group))
            ops.append(dist.P2POp(dist.irecv, lo_halo, self.rank - 1, self.group))
        for w in dist.batch_isend_irecv(ops):
            w.wait()

    def _release_peer_buffers(self):
        for own, table, opened in getattr(self, "_peer_handles", []):
            for p in opened:
                self.lib.rg_ipc_close(ctypes.c_void_p(p))
            self.lib.rg_ipc_free(ctypes.c_void_p(own))
        self._peer_handles, self._peers = [], None

    def _map_peer_buffers(self):
        """(collective) two export buffers per rank that the other ranks of the node map through CUDA IPC, and for each a device
        table of every rank's buffer address in THIS process.  Falls back to the all-gather exchange where the library or the
        platform has no IPC (the CPU emulator of the tests, RG_NO_PEER_EXCHANGE=1)."""
        import os
        torch, dist, lib = self.torch, self.dist, self.lib
        self._release_peer_buffers()
        self._peer_turn = 0
        ok = hasattr(lib, "rg_ipc_alloc") and self.use_cuda_stream and not os.environ.get("RG_NO_PEER_EXCHANGE")
        nbytes = 4 * (4 + 2 * self._cap)
        mine = []
        if ok:
            try:
                for _ in range(2):
                    ptr = ctypes.c_void_p(); h = (ctypes.c_ubyte * 64)()
                    if lib.rg_ipc_alloc(nbytes, ctypes.byref(ptr), h) != 0:
                        raise RuntimeError("rg_ipc_alloc")
                    mine.append((ptr.value, bytes(h)))
            except Exception:
                ok = False
        everyone = [None] * self.world
        dist.all_gather_object(everyone, mine if ok else None, group=self.group)
        if not all(e is not None and len(e) == 2 for e in everyone):
            for ptr, _ in mine:
                lib.rg_ipc_free(ctypes.c_void_p(ptr))
            return
        peers, handles = [], []
        try:
            for k in range(2):
                addrs, opened = [], []
                for r in range(self.world):
                    if r == self.rank:
                        addrs.append(mine[k][0]); continue
                    p = ctypes.c_void_p(); hb = (ctypes.c_ubyte * 64).from_buffer_copy(everyone[r][k][1])
                    if lib.rg_ipc_open(hb, ctypes.byref(p)) != 0:
                        raise RuntimeError("rg_ipc_open")
                    addrs.append(p.value); opened.append(p.value)
                table = torch.tensor(addrs, dtype=torch.int64, device=self.device)
                peers.append((ctypes.c_void_p(mine[k][0]), table)); handles.append((mine[k][0], table, opened))
            good = 1
        except Exception:
            good = 0
        flag = torch.tensor([good], dtype=torch.int32, device=self.device)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=self.group)
        self._peer_handles = handles
        if int(flag.item()) == 1:
            self._peers = peers
            self._meet = torch.zeros(1, dtype=torch.int32, device=self.device)
        else:
            self._release_peer_buffers()

    # ---- the grow
    def _mark(self, name):
        """optional segment timing (RG_SLAB_TIMING=1): CUDA events on the current stream + host clock"""
        if self._timing is None:
            return
        import time
        ev = self.torch.cuda.Event(enable_timing=True); ev.record()
        self._timing.append((name, ev, time.perf_counter()))

    def timing_report(self):
        if not self._timing:
            return {}
        self.torch.cuda.synchronize()
        out = {}
        for (n0, e0, t0), (n1, e1, t1) in zip(self._timing[:-1], self._timing[1:]):
            out[n1] = {"gpu_ms": e0.elapsed_time(e1), "host_ms": (t1 - t0) * 1e3}
        return out

    def validate(self, threshold):
        """after optimistic steps: synchronise and check the device-side verdict; if a guessed capacity was too small
        the last grow is repeated with the synchronous exchange (which also refreshes the guesses)"""
        st = (ctypes.c_uint32 * 4)()
        rc = self.lib.rg_ctx_status(self.ctx, st)
        if rc != 0 and (st[1] & 4):
            self._cap = None
            self.grow(threshold)
            rc = self.lib.rg_ctx_status(self.ctx, st)
        if rc != 0:
            raise RuntimeError(f"slab grow failed: status=0x{st[1]:x}")
        return self.mask

    def grow(self, threshold, optimistic=True):
        """enqueue label -> exchange -> resolve -> output; returns the slab's mask tensor.  With optimistic=True and
        capacities known from an earlier grow, nothing waits on the host in mid-step (call validate() afterwards)."""
        import os
        self._timing = [] if (os.environ.get("RG_SLAB_TIMING") and self.use_cuda_stream) else None
        self._mark("start")
        lib, st, plane = self.lib, self._stream(), self.plane
        slab_ptr = self.buf.data_ptr() + plane
        ny_l, nz_l = (self.nzl, 1) if self.rows else (self.ny, self.nzl)
        self._ck(lib.rg_slab_label(self.ctx, slab_ptr, self.nx, ny_l, nz_l, int(self.has_lower), int(self.has_upper),
                                   self.seeds_dev.data_ptr() if self.nseeds else None, self.nseeds, int(threshold),
                                   self.mask.data_ptr(), st), "rg_slab_label")
        self._mark("label")
        if self.world > 1 and optimistic and self._cap and self._peers:
            # no gather: every rank exports into a buffer its peers have mapped (CUDA IPC), one tiny all-reduce is the barrier
            # ("all exports are complete"), and the resolve kernels read the other ranks' lists over NVLink where they lie
            torch, dist = self.torch, self.dist
            cap = self._cap
            own, table = self._peers[self._peer_turn]
            self._peer_turn ^= 1                             # two buffers: a fast rank may export step k+1 while a slow one reads step k
            self._ck(lib.rg_slab_export_packed(self.ctx, own, cap, st), "rg_slab_export_packed")
            self._mark("export")
            dist.all_reduce(self._meet, group=self.group)
            self._mark("gather_lists")
            self._ck(lib.rg_slab_resolve_peers(self.ctx, self._xpar.data_ptr(), self._xpar_cap, table.data_ptr(), cap,
                                               self.rank, self.world, st), "rg_slab_resolve_peers")
            self._mark("resolve")
            self.optimistic_steps += 1
        elif self.world > 1 and optimistic and self._cap:
            torch, dist = self.torch, self.dist
            cap = self._cap
            self._ck(lib.rg_slab_export_packed(self.ctx, self._packed.data_ptr(), cap, st), "rg_slab_export_packed")
            self._mark("export")
            dist.all_gather_into_tensor(self._allpacked, self._packed, group=self.group)
            self._mark("gather_lists")
            self._ck(lib.rg_slab_resolve_packed(self.ctx, self._xpar.data_ptr(), self._xpar_cap, self._allpacked.data_ptr(), cap,
                                                self.rank, self.world, st), "rg_slab_resolve_packed")
            self._mark("resolve")
            self.optimistic_steps += 1
        elif self.world > 1:
            torch, dist = self.torch, self.dist
            self.counts.zero_()
            if self.has_lower:
                self._ck(lib.rg_slab_export(self.ctx, 0, self.lists[0].data_ptr(), self.counts.data_ptr(), st), "rg_slab_export")
            if self.has_upper:
                self._ck(lib.rg_slab_export(self.ctx, 1, self.lists[1].data_ptr(), self.counts.data_ptr(), st), "rg_slab_export")
            if not (self.has_lower or self.has_upper):
                raise RuntimeError("unreachable")
            self._mark("export")
            dist.all_gather_into_tensor(self.all_counts, self.counts, group=self.group)
            self._mark("gather_counts")
            cnt = self.all_counts.cpu().numpy().astype(np.int64).reshape(self.world, 4)          # the one host sync of the exchange
            for r in range(self.world - 1):
                if cnt[r, 1] != cnt[r + 1, 0]:
                    raise RuntimeError(f"interface {r}/{r + 1}: list lengths differ ({cnt[r, 1]} vs {cnt[r + 1, 0]})")
            maxn = int(max(cnt[:, 0].max(), cnt[:, 1].max()))
            self._mark("host_sync")
            if maxn > 0:
                mine = self.lists[:, :maxn].contiguous()
                flat = torch.empty(self.world * 2 * maxn, dtype=torch.int32, device=self.device)
                dist.all_gather_into_tensor(flat, mine.view(-1), group=self.group)
                gathered = flat.view(self.world, 2, maxn)
                self._mark("gather_lists")
                total_nodes = int(cnt[:, 2].sum())
                xpar = torch.empty(total_nodes + 2, dtype=torch.int32, device=self.device)
                cnt64 = np.ascontiguousarray(cnt, np.int64)
                self._ck(lib.rg_slab_resolve(self.ctx, xpar.data_ptr(), flat.data_ptr(), maxn, cnt64.ctypes.data, self.rank, self.world, st),
                         "rg_slab_resolve")
                self._mark("resolve")
                self._keepalive = (gathered, xpar, mine)
                # what the next, optimistic step may assume (25 % headroom)
                self._cap = int(maxn * 1.25) + 1024
                self._xpar_cap = int(total_nodes * 1.25) + 4096
                self._packed = torch.zeros(4 + 2 * self._cap, dtype=torch.int32, device=self.device)
                self._allpacked = torch.zeros(self.world * (4 + 2 * self._cap), dtype=torch.int32, device=self.device)
                self._xpar = torch.empty(self._xpar_cap, dtype=torch.int32, device=self.device)
                self._map_peer_buffers()
            self.rounds = 1
        self._ck(lib.rg_slab_finish(self.ctx, st), "rg_slab_finish")
        self._mark("finish")
        return self.mask
