/*
 * rg_engine.cu -- CUDA kernels (sm_100a) + context + C ABI of the region-growing engine.
 *
 * Kernel bodies live in rg_core.h (see the pass A / B / C description there); this file wraps
 * them in __global__ functions, owns the per-device scratch ("context") and implements the
 * entry points declared in include/rg.h.  There is deliberately no CPU path in this library.
 */
#include <cuda_runtime.h>
#include <cuda.h>                 /* CUtensorMap types only; the encoder is fetched with cudaGetDriverEntryPoint */
#include <cub/device/device_radix_sort.cuh>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <mutex>
#include <thread>
#include <vector>
#include <functional>
#include <algorithm>
#include <chrono>

#include "rg_core.h"
#include "rg_fast.h"
#include "rg_slab_core.h"
#include "../../include/rg.h"

using namespace rg;

/* ------------------------------------------------------------------ Exec for a CUDA block */
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

template <int NT>
struct DevExec {
    bool grid3d = false;                    /* tile coordinates come straight from blockIdx */
    const CUtensorMap *tmap = nullptr;      /* image descriptor (kernel parameter) */
    uint32_t phase = 0;                     /* parity of the next mbarrier completion */
    bool tma_pending = false;

    const CUtensorMap *tmap_mask = nullptr; /* mask descriptor for the bulk stores that clear a tile (fast labelling pass) */
    bool store_pending = false;
    /* barrier after generic-proxy writes to shared memory that the copy engine is going to read */
    __device__ __forceinline__ void sync_async_smem()
    {
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncthreads();
    }
    /* Clear this tile's part of the mask with bulk tensor stores from `zeros` (shared memory, all zero, `bytes` long):
       boxes of TX x BR x BZ bytes, one instruction each, issued by one thread; the hardware clips them at the volume's
       border.  A box may reach into the next tile's rows (images: 126-row tiles, 32-row boxes) -- harmless, every CTA of
       this kernel writes nothing but zeros to the mask.  The shared-memory source must stay zero until store_wait(). */
    template <class C>
    __device__ __forceinline__ bool tma_clear_tile(const Params &P, const void *zeros, int bytes, int64_t x0, int64_t y0, int64_t z0)
    {
        if (!P.use_tma || !tmap_mask) return false;
        constexpr int BR = C::TY < 32 ? C::TY : 32;
        const int BZ = bytes / (C::TX * BR) < C::TZ ? bytes / (C::TX * BR) : C::TZ;
        store_pending = true;
        if (threadIdx.x == 0) {
            const uint32_t src = smem_u32(zeros);
            for (int z = 0; z < C::TZ; z += BZ)
                for (int y = 0; y < C::TY; y += BR)
                    asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.tile.bulk_group [%0, {%2, %3, %4}], [%1];"
                                 :: "l"((uint64_t)tmap_mask), "r"(src), "r"((int)x0), "r"((int)y0 + y), "r"((int)z0 + z) : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
        return true;
    }
    /* the bulk stores have read their shared-memory source (called by every thread; only the issuing thread waits) */
    __device__ __forceinline__ void store_wait()
    {
        if (store_pending && threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        store_pending = false;
    }
    /* one thread, once per kernel */
    __device__ __forceinline__ void mbar_init(uint64_t *bar)
    {
        if (threadIdx.x == 0) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(smem_u32(bar)) : "memory");
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        __syncthreads();
    }
    /* Stage the haloed tile [x0-16, x0+TX+16) x [y0-1, y0+TY+1) x [z0-1, z0+TZ+1) with the tensor-memory
       accelerator: one elected thread, hardware zero-fill outside the volume, completion on the mbarrier. */
    template <class C>
    __device__ __forceinline__ bool tma_tile(const Params &P, uint8_t *img, uint64_t *bar, int64_t x0, int64_t y0, int64_t z0)
    {
        if (!P.use_tma) return false;
        tma_pending = true;
        if (threadIdx.x == 0) {
            const uint32_t b = smem_u32(bar);
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");     /* earlier generic-proxy accesses to this buffer */
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(b), "r"((uint32_t)C::IMG_BYTES) : "memory");
            const int cx = (int)x0 - C::XO, cz = (int)z0 - C::HZ + ext_zlo<C>(P), ylo = ext_ylo<C>(P);    /* the map starts at the lower halo plane / row */
#pragma unroll
            for (int k = 0; k < C::TMA_NY; k++) {
                const int cy = (int)y0 - 1 + ylo + k * C::BOXY;
                asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                             :: "r"(smem_u32(img + k * C::BOXY * C::RS)), "l"((uint64_t)tmap), "r"(b), "r"(cx), "r"(cy), "r"(cz) : "memory");
            }
        }
        return true;
    }
    __device__ __forceinline__ void tile_wait(uint64_t *bar)
    {
        store_wait();
        cp_async_wait_all();
        if (tma_pending) {
            const uint32_t b = smem_u32(bar);
            uint32_t done;
            do {
                asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                             : "=r"(done) : "r"(b), "r"(phase) : "memory");
            } while (!done);
            phase ^= 1u;
            tma_pending = false;
        }
    }
    static constexpr int NTHREADS = NT;
    __device__ __forceinline__ int tid() const { return (int)threadIdx.x; }
    template <class F> __device__ __forceinline__ void pfor(int n, F f)
    {
        for (int i = (int)threadIdx.x; i < n; i += NT) f(i);
    }
    /* the same split, f(i, k) with k = 0, 1, ... counting this thread's items: two pfor_k over the same n hand a
       thread the same items, so per-thread arrays indexed by k carry values across a barrier in registers */
    template <class F> __device__ __forceinline__ void pfor_k(int n, F f)
    {
#pragma unroll
        for (int k = 0; k < (n + NT - 1) / NT; k++) { const int i = (int)threadIdx.x + k * NT; if (i < n) f(i, k); }
    }
    /* one warp per item: pred(item, lane) is OR-ed over the warp and handed to body(item, lane, any) */
    template <class Pr, class Bo> __device__ __forceinline__ void wfor(uint32_t n, Pr pred, Bo body)
    {
        const uint32_t lane = threadIdx.x & 31u, wpb = NT / 32;
        for (uint32_t item = blockIdx.x * wpb + (threadIdx.x >> 5); item < n; item += gridDim.x * wpb) {
            const bool any = __any_sync(0xFFFFFFFFu, pred(item, lane));
            body(item, lane, any);
        }
    }
    template <class F> __device__ __forceinline__ void gpfor(int64_t n, F f)
    {
        for (int64_t i = (int64_t)blockIdx.x * NT + threadIdx.x; i < n; i += (int64_t)gridDim.x * NT) f(i);
    }
    __device__ __forceinline__ void sync() { __syncthreads(); }
    __device__ __forceinline__ bool sync_or(bool p) { return __syncthreads_or(p ? 1 : 0) != 0; }
    template <class C> __device__ __forceinline__ void origin(const Params &P, uint32_t tile, int &tx, int &ty, int &tz) const
    {
        if (grid3d) { tx = (int)blockIdx.x; ty = (int)blockIdx.y; tz = (int)blockIdx.z; } else tile_origin<C>(P, tile, tx, ty, tz);
    }
    __device__ __forceinline__ uint32_t bid() const { return blockIdx.x; }
    __device__ __forceinline__ uint32_t nblk() const { return gridDim.x; }
    /* block-wide exclusive scan of count(i), i < n, into out[i]; returns the total to every thread.
       Each thread owns a contiguous run of ceil(n / blockDim) items, so one shuffle scan covers the block. */
    template <class T, class F> __device__ __forceinline__ uint32_t scan(int n, F count, T *out, uint32_t *scratch)
    {
        const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nwarp = NT >> 5;
        const int per = (n + NT - 1) / NT;
        const int i0 = (int)threadIdx.x * per;
        uint32_t local = 0;
        for (int k = 0; k < per; k++) if (i0 + k < n) local += count(i0 + k);
        uint32_t incl = local;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, d); if (lane >= d) incl += t; }
        if (lane == 31) scratch[wid] = incl;
        __syncthreads();
        if (wid == 0) {
            uint32_t v = lane < nwarp ? scratch[lane] : 0u, inc2 = v;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) { uint32_t t = __shfl_up_sync(0xFFFFFFFFu, inc2, d); if (lane >= d) inc2 += t; }
            scratch[lane] = inc2 - v;
            if (lane == 31) scratch[32] = inc2;
        }
        __syncthreads();
        uint32_t run = scratch[wid] + incl - local;
        for (int k = 0; k < per; k++) if (i0 + k < n) { out[i0 + k] = (T)run; run += count(i0 + k); }
        const uint32_t total = scratch[32];
        __syncthreads();
        return total;
    }
    template <class F> __device__ __forceinline__ uint32_t scan16(int n, F count, uint16_t *out, uint32_t *scratch) { return scan<uint16_t>(n, count, out, scratch); }
    template <class F> __device__ __forceinline__ uint32_t scan32(int n, F count, uint32_t *out, uint32_t *scratch) { return scan<uint32_t>(n, count, out, scratch); }
};

/* ------------------------------------------------------------------ kernels */
#define RG_THREADS 512          /* full labelling pass: 2 CTAs/SM (104 KB smem each) */
#ifndef RG_FAST_THREADS
#define RG_FAST_THREADS 256     /* fast labelling pass: 4 CTAs/SM (56 KB smem each, lean staging layout) */
#endif
#ifndef RG_FAST_MINB
#define RG_FAST_MINB 4
#endif
#define RG_FLAG_THREADS 256     /* flag / zero-fill: no smem, 8 tiles per CTA */
#define RG_PERSISTENT_CTAS (148 * 2)
#define RG_PASSC_THREADS 384    /* pass C: 3 CTAs/SM (66 KB smem each) */
#define RG_PASSC_CTAS (148 * 3)

/* pass A, fast path: every tile.  C = the tile shape; the kernel stages it in the lean layout (LeanOf<C>) */
template <class C>
__global__ void __launch_bounds__(RG_FAST_THREADS, RG_FAST_MINB) k_pass_a_fast(const Params P, const __grid_constant__ CUtensorMap tmap,
                                                                               const __grid_constant__ CUtensorMap tmap_mask)
{
    typedef typename LeanOf<C>::type L;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    FastSmem<L> &s = *reinterpret_cast<FastSmem<L> *>(smem_raw);
    DevExec<RG_FAST_THREADS> ex; ex.tmap = &tmap;
    ex.grid3d = gridDim.y * gridDim.z > 1 || (P.nty == 1 && P.ntz == 1);
    ex.tmap_mask = &tmap_mask;
    if (P.use_tma && threadIdx.x == 0) {                    /* the barrier that publishes this is the first one of the tile pass */
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(smem_u32(&s.mbar)) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    const uint32_t tile = ex.grid3d ? (blockIdx.z * (uint32_t)P.nty + blockIdx.y) * (uint32_t)P.ntx + blockIdx.x : blockIdx.x;
    tile_pass_a_fast<L>(ex, P, tile, s);
}

static_assert(sizeof(TileSmem<Cfg3D>) + 1024 <= 228 * 1024 / 2, "the full labelling pass is sized for two CTAs per SM");
static_assert(3 * (sizeof(FloodSmem<LeanOf<Cfg3D>::type>) + 1024) <= 228 * 1024, "pass C is sized for three CTAs per SM");
static_assert(RG_FAST_MINB * (sizeof(FastSmem<LeanOf<Cfg3D>::type>) + 1024) <= 228 * 1024, "the fast labelling pass is sized for RG_FAST_MINB CTAs per SM");

/* pass A, full path: the tiles the fast path listed (persistent CTAs) */
#ifdef RG_PARENT32
#define RG_FULL_MINB 1
#else
#define RG_FULL_MINB 2
#endif
template <class C>
__global__ void __launch_bounds__(RG_THREADS, RG_FULL_MINB) k_pass_a_full(const Params P, const __grid_constant__ CUtensorMap tmap)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    TileSmem<C> &s = *reinterpret_cast<TileSmem<C> *>(smem_raw);
    DevExec<RG_THREADS> ex; ex.tmap = &tmap;
    if (P.use_tma) ex.mbar_init(&s.mbar);
    pass_a_full_list<C>(ex, P, s);
}

/* pass C, part 1: zero-fill tiles without region voxels, list the others (one warp per tile) */
template <class C>
__global__ void __launch_bounds__(RG_FLAG_THREADS) k_flag_tiles(const Params P)
{
    DevExec<RG_FLAG_THREADS> ex;
    flag_tiles_all<C>(ex, P);
}

/* pass C, part 2: the listed tiles (persistent CTAs, 3 per SM: the tile is staged in the lean layout and the
   stride planes of the flood reuse its bytes) */
template <class C>
__global__ void __launch_bounds__(RG_PASSC_THREADS, 3) k_pass_c(const Params P, const __grid_constant__ CUtensorMap tmap)
{
    typedef typename LeanOf<C>::type L;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    FloodSmem<L> &s = *reinterpret_cast<FloodSmem<L> *>(smem_raw);
    DevExec<RG_PASSC_THREADS> ex; ex.tmap = &tmap;
    if (P.use_tma) ex.mbar_init(&s.mbar);
    pass_c_list<L>(ex, P, s);
}

/* pass C for tiles of constant rows: no staging, a few words of shared memory.  A tile is a chain of dependent loads (list, record,
   link words, finds) and then a few stores: as many tiles in flight as an SM takes -- 32 CTAs of two warps (C3, row output +
   pass C: 0.350 ms with 12 x 128 threads, 0.326 with 16 x 128, 0.308 with 32 x 64) */
#ifndef RG_ROWS_THREADS
#define RG_ROWS_THREADS 64
#endif
#ifndef RG_ROWS_PER_SM
#define RG_ROWS_PER_SM 32
#endif
#define RG_ROWS_CTAS (148 * RG_ROWS_PER_SM)
template <class C>
__global__ void __launch_bounds__(RG_ROWS_THREADS) k_rows_out(const Params P)
{
    __shared__ RowsOutSmem<C> s;
    DevExec<RG_ROWS_THREADS> ex;
    rows_out_list<C>(ex, P, s);
}

template <class C>
__global__ void __launch_bounds__(256) k_region_boxes(const Params P, int32_t *out)
{
    DevExec<256> ex;
    region_boxes<C>(ex, P, out);
}

template <class C>
__global__ void __launch_bounds__(256, 8) k_merge(const Params P)      /* latency bound: all 64 warps of an SM */
{
    DevExec<256> ex;
    merge_faces<C>(ex, P);
}

template <class C>
__global__ void __launch_bounds__(256) k_seed_keys(const Params P, const int32_t *seeds_xyz, uint64_t *keys)
{
    DevExec<256> ex;
    if (blockIdx.x == 0 && threadIdx.x == 0) reset_counters(P);
    make_seed_keys<C>(ex, P, seeds_xyz, keys);
}

/* bitonic sort of up to 2048 keys in one block (the usual case: a handful of seeds) */
__global__ void __launch_bounds__(1024) k_sort_small(uint64_t *keys, int n)
{
    __shared__ uint64_t sk[2048];
    for (int i = threadIdx.x; i < 2048; i += blockDim.x) sk[i] = i < n ? keys[i] : ~0ull;
    __syncthreads();
    for (int k = 2; k <= 2048; k <<= 1)
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = threadIdx.x; i < 2048; i += blockDim.x) {
                int p = i ^ j;
                if (p > i) {
                    uint64_t a = sk[i], b = sk[p];
                    bool up = (i & k) == 0;
                    if ((a > b) == up) { sk[i] = b; sk[p] = a; }
                }
            }
            __syncthreads();
        }
    for (int i = threadIdx.x; i < n; i += blockDim.x) keys[i] = sk[i];
}

__global__ void __launch_bounds__(256) k_apply_mask(uint8_t *image, const uint8_t *mask, int64_t n)
{
    /* region.c:576  image[i] *= (region[i] == 0) */
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) if (mask[i]) image[i] = 0;
}

/* multi-GPU (z-slab) path, bodies in rg_slab_core.h */
template <class C>
__global__ void __launch_bounds__(1024) k_export_offsets(const Params P, int side, uint32_t *tile_off, uint32_t *count_out)
{
    __shared__ uint32_t scratch[40];
    DevExec<1024> ex;
    export_offsets<C>(ex, P, side, tile_off, count_out, scratch);
}
template <class C>
__global__ void __launch_bounds__(128) k_export_entries(const Params P, int side, const uint32_t *tile_off, uint32_t *out_roots, uint32_t *counts3, uint32_t cap)
{
    DevExec<128> ex;
    if (blockIdx.x == 0 && threadIdx.x == 0) counts3[2] = P.counters[CTR_NODES];
    export_entries<C>(ex, P, side, tile_off, out_roots, cap);
}
template <class C>
__global__ void __launch_bounds__(256) k_export_border(const Params P, int side, uint32_t *roots, uint8_t *vals)
{
    DevExec<256> ex;
    export_border<C>(ex, P, side, roots, vals);
}
__global__ void __launch_bounds__(256) k_resolve_union_seam(uint32_t *xpar, const uint32_t *a, uint32_t offa, const uint8_t *va,
                                                            const uint32_t *b, uint32_t offb, const uint8_t *vb, uint32_t n, uint32_t thr)
{
    DevExec<256> ex;
    resolve_union_seam(ex, xpar, a, offa, va, b, offb, vb, n, thr);
}
__global__ void k_resolve_prepare(const Params P, const uint32_t *packed_all, uint32_t stride_words, uint32_t cap, uint32_t xpar_cap,
                                  uint32_t rank, uint32_t world, ResolveArgs *A, uint32_t *xpar, const uint32_t *const *peer_records)
{
    if (blockIdx.x == 0 && threadIdx.x == 0) { xpar[0] = 0; resolve_prepare(P, packed_all, stride_words, cap, xpar_cap, rank, world, A, peer_records); }
}
__global__ void __launch_bounds__(256) k_resolve_init_packed(uint32_t *xpar, const uint32_t *packed_all, uint32_t cap, const ResolveArgs *A)
{
    DevExec<256> ex; resolve_init_packed(ex, xpar, packed_all, cap, *A);
}
__global__ void __launch_bounds__(256) k_resolve_union_packed(uint32_t *xpar, const uint32_t *packed_all, uint32_t cap, const ResolveArgs *A)
{
    DevExec<256> ex; resolve_union_packed(ex, xpar, packed_all, cap, *A);
}
__global__ void __launch_bounds__(256) k_resolve_apply_packed(const Params P, uint32_t *xpar, const uint32_t *packed_all, uint32_t cap, const ResolveArgs *A)
{
    DevExec<256> ex; resolve_apply_packed(ex, P, xpar, packed_all, cap, *A);
}
__global__ void __launch_bounds__(256) k_resolve_init(uint32_t *xpar, const uint32_t *roots, uint32_t n, uint32_t off)
{
    DevExec<256> ex;
    if (blockIdx.x == 0 && threadIdx.x == 0) xpar[0] = 0;
    resolve_init(ex, xpar, roots, n, off);
}
__global__ void __launch_bounds__(256) k_resolve_union(uint32_t *xpar, const uint32_t *a, uint32_t offa, const uint32_t *b, uint32_t offb, uint32_t n)
{
    DevExec<256> ex;
    resolve_union(ex, xpar, a, offa, b, offb, n);
}
__global__ void __launch_bounds__(256) k_resolve_apply(const Params P, uint32_t *xpar, const uint32_t *roots, uint32_t n, uint32_t off)
{
    DevExec<256> ex;
    resolve_apply(ex, P, xpar, roots, n, off);
}

__global__ void __launch_bounds__(256) k_resolve_init_all(uint32_t *xpar, const uint32_t *lists, const ResolveArgs A)
{
    DevExec<256> ex;
    if (blockIdx.x == 0 && threadIdx.x == 0) xpar[0] = 0;
    resolve_init_all(ex, xpar, lists, A);
}
__global__ void __launch_bounds__(256) k_resolve_union_all(uint32_t *xpar, const uint32_t *lists, const ResolveArgs A)
{
    DevExec<256> ex;
    resolve_union_all(ex, xpar, lists, A);
}
__global__ void __launch_bounds__(256) k_resolve_apply_all(const Params P, uint32_t *xpar, const uint32_t *lists, const ResolveArgs A)
{
    DevExec<256> ex;
    resolve_apply_all(ex, P, xpar, lists, A);
}

/* ------------------------------------------------------------------ errors */
static thread_local char g_err[512] = "";
static int fail(int code, const char *what, cudaError_t e = cudaSuccess)
{
    if (e != cudaSuccess) snprintf(g_err, sizeof g_err, "%s: %s", what, cudaGetErrorString(e));
    else snprintf(g_err, sizeof g_err, "%s", what);
    return code;
}
#define CK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail(RG_ERR_CUDA, #call, e_); } while (0)

/* The entry points run on the context's device and leave the caller's current device as they found it
   (a host with several contexts in one thread must not find itself on another GPU after a call). */
struct DeviceScope {
    int prev; cudaError_t err;
    explicit DeviceScope(int dev) : prev(-1)
    {
        int cur = -1;
        err = cudaGetDevice(&cur);
        if (err == cudaSuccess && cur != dev) { err = cudaSetDevice(dev); if (err == cudaSuccess) prev = cur; }
    }
    ~DeviceScope() { if (prev >= 0) cudaSetDevice(prev); }
    DeviceScope(const DeviceScope &) = delete;
    DeviceScope &operator=(const DeviceScope &) = delete;
};
#define ON_DEVICE(dev) DeviceScope device_scope_(dev); CK(device_scope_.err)

extern "C" const char *rg_last_error(void) { return g_err; }

/* ------------------------------------------------------------------ context */
struct rg_ctx {
    int device;
    void *ws; size_t ws_bytes;                 /* engine scratch */
    uint8_t *d_image, *d_mask; size_t io_bytes; /* staging for the host-pointer entry points */
    int32_t *d_seeds; size_t seeds_cap;
    uint32_t *h_status;                        /* pinned, CTR_WORDS words */
    uint32_t *last_counters;                   /* device counters of the last call */
    cudaStream_t last_stream;
    bool attr_set;
    uint32_t next_open;                        /* Params::open of the next label pass (set by rg_block_label, consumed once) */
    bool may_grow;                             /* the context owns its scratch and may enlarge it (not the noalloc view) */
    bool worst_case_layout;                    /* size the node space for the worst case whatever RG_NODE_BUDGET says (the noalloc view) */
    size_t tail_refused;                       /* smallest size for which the optional link-plane tail did not fit (0 = never failed) */
    Params last_P; int last_is2d; uint32_t last_ntiles; bool have_last;
    uint32_t *if_off[2];                       /* per-tile offsets of the two slab interfaces */
    ResolveArgs *d_rargs;                      /* device copy of the exchange bookkeeping (optimistic exchange) */
    int32_t *d_boxes;                          /* per row of tiles: x-extent of the region (output-sensitive copy of the mask) */
    int32_t *h_boxes; size_t h_boxes_cap;      /* pinned host copy of it */
    uint8_t *bounce; size_t bounce_bytes;      /* pinned staging for pageable host buffers */
    int64_t last_h2d_bytes, last_d2h_bytes; int last_copies, last_staged;   /* what the last host-pointer call moved (rg_host_stats) */
    dim3 last_grid;
    CUtensorMap last_tmap;                     /* image descriptor of the last label pass */
    CUtensorMap last_tmap_fast;                /* the same image with the lean box of the fast labelling pass */
    CUtensorMap last_tmap_mask;                /* the mask, box = one clearing store of the fast labelling pass */
    bool profile; cudaEvent_t ev[7]; bool ev_made;  /* optional per-kernel timing (rg_ctx_profile) */
    uint32_t last_full_tiles;                      /* tiles of the last call that overflowed the dense path (set by rg_ctx_status) */
};

/* cuTensorMapEncodeTiled through the runtime, so the library does not link against libcuda */
typedef CUresult (*tmap_encode_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                   const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                   CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static tmap_encode_fn lookup_tmap_encoder()
{
    void *p = nullptr; cudaDriverEntryPointQueryResult q;
    if (getenv("RG_NO_TMA")) return nullptr;                /* debugging aid: force the cp.async staging path */
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
        return (tmap_encode_fn)p;
    return nullptr;
}
static tmap_encode_fn get_tmap_encoder()
{
    static const tmap_encode_fn fn = lookup_tmap_encoder(); /* function-local static: initialised once, also with several host threads */
    return fn;
}

/* uint8 volume [nz_total][ny][nx] starting at `base`; box = one staged tile (TMA_NY boxes stacked in y) */
template <class C>
static bool make_tmap(CUtensorMap *tm, const uint8_t *base, int64_t nx, int64_t ny, int64_t nz_total)
{
    tmap_encode_fn enc = get_tmap_encoder();
    if (!enc) return false;
    cuuint64_t dims[3] = { (cuuint64_t)nx, (cuuint64_t)ny, (cuuint64_t)nz_total };
    cuuint64_t strides[2] = { (cuuint64_t)nx, (cuuint64_t)nx * (cuuint64_t)ny };
    cuuint32_t box[3] = { (cuuint32_t)C::RS, (cuuint32_t)C::BOXY, (cuuint32_t)C::PZ };
    cuuint32_t es[3] = { 1, 1, 1 };
    if (strides[1] >= (1ull << 40)) return false;
    return enc(tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, (void *)base, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
               CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

/* uint8 mask [nz][ny][nx]; box = what one bulk store of the fast labelling pass clears (see DevExec::tma_clear_tile) */
template <class C>
static bool make_tmap_mask(CUtensorMap *tm, uint8_t *base, int64_t nx, int64_t ny, int64_t nz, int zero_bytes)
{
    tmap_encode_fn enc = get_tmap_encoder();
    if (!enc) return false;
    const int BR = C::TY < 32 ? C::TY : 32;
    const int BZ = zero_bytes / (C::TX * BR) < C::TZ ? zero_bytes / (C::TX * BR) : C::TZ;
    if (BZ < 1) return false;
    cuuint64_t dims[3] = { (cuuint64_t)nx, (cuuint64_t)ny, (cuuint64_t)nz };
    cuuint64_t strides[2] = { (cuuint64_t)nx, (cuuint64_t)nx * (cuuint64_t)ny };
    cuuint32_t box[3] = { (cuuint32_t)C::TX, (cuuint32_t)BR, (cuuint32_t)BZ };
    cuuint32_t es[3] = { 1, 1, 1 };
    if (strides[1] >= (1ull << 40)) return false;
    return enc(tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, (void *)base, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
               CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

struct Layout {
    size_t off_rec, off_ent, off_npar, off_nrep, off_ctr, off_keys, off_keys2, off_cub, off_if0, off_if1, off_clist, off_flist, off_rargs, off_boxes, total;
    size_t off_planes, total_with_planes; uint32_t plane_cap;      /* optional tail: link planes the full pass keeps for pass C */
    uint32_t ntiles, node_cap; int ntx, nty, ntz; size_t cub_bytes;
};

static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

template <class C>
static int make_layout(int64_t nx, int64_t ny, int64_t nz, int nseeds, Layout &L, bool honour_env = true)
{
    int64_t ntx = (nx + C::TX - 1) / C::TX, nty = (ny + C::TY - 1) / C::TY, ntz = (nz + C::TZ - 1) / C::TZ;
    int64_t ntiles = ntx * nty * ntz;
    /* Node ids: one fixed block per tile for the fast pass, then room for the full pass.  The worst case (every tile
       through the full pass with a node per face entry) stops fitting 32 bits at ~17 Gvoxel, long before a 180 GB card
       is full; beyond that -- or when RG_NODE_BUDGET says so -- the full pass gets an average of RG_NODE_BUDGET (2048)
       nodes per tile.  Running out is detected on the device (status bit 1, rg_ctx_status fails), never silent. */
    int64_t budget = (int64_t)C::ENT_MAX;
    if (honour_env) { const char *e = getenv("RG_NODE_BUDGET"); if (e && atoll(e) > 0 && atoll(e) < budget) budget = atoll(e); }
    int64_t node_cap = 1 + ntiles * (budget + (int64_t)FAST_NODE_SLOTS) + (int64_t)nseeds;
    if (node_cap >= (1ll << 32)) node_cap = 1 + ntiles * ((budget < 2048 ? budget : 2048) + (int64_t)FAST_NODE_SLOTS) + (int64_t)nseeds;
    if (ntiles >= (1ll << 31) || node_cap >= (1ll << 32)) return fail(RG_ERR_TOO_LARGE, "volume too large for 32-bit node ids");
    L.ntx = (int)ntx; L.nty = (int)nty; L.ntz = (int)ntz; L.ntiles = (uint32_t)ntiles; L.node_cap = (uint32_t)node_cap;
    size_t o = 0;
    L.off_rec = o;  o = align_up(o + (size_t)ntiles * sizeof(TileRec), 256);
    L.off_ent = o;  o = align_up(o + (size_t)ntiles * C::ENT_MAX * sizeof(uint16_t), 256);
    L.off_npar = o; o = align_up(o + (size_t)node_cap * sizeof(uint32_t), 256);
    L.off_nrep = o; o = align_up(o + (size_t)node_cap * sizeof(uint16_t), 256);
    L.off_ctr = o;  o = align_up(o + CTR_WORDS * sizeof(uint32_t), 256);
    size_t nk = (size_t)(nseeds > 0 ? nseeds : 1);
    L.off_keys = o;  o = align_up(o + nk * 8, 256);
    L.off_keys2 = o; o = align_up(o + nk * 8, 256);
    L.cub_bytes = 0;
    if (nseeds > 2048) {
        size_t tb = 0;
        cub::DeviceRadixSort::SortKeys(nullptr, tb, (const uint64_t *)nullptr, (uint64_t *)nullptr, nseeds);
        L.cub_bytes = tb;
    }
    L.off_cub = o; o = align_up(o + L.cub_bytes, 256);
    L.off_if0 = o; o = align_up(o + ((size_t)ntx * nty + 1) * 4, 256);
    L.off_if1 = o; o = align_up(o + ((size_t)ntx * nty + 1) * 4, 256);
    L.off_clist = o; o = align_up(o + (size_t)ntiles * 4, 256);
    L.off_flist = o; o = align_up(o + (size_t)ntiles * 4, 256);
    L.off_rargs = o; o = align_up(o + sizeof(ResolveArgs), 256);
    L.off_boxes = o; o = align_up(o + (size_t)nty * ntz * 2 * sizeof(int32_t), 256);
    L.total = o;
    /* optional: 12 KB of link planes for each of the first plane_cap tiles of the full pass's list (all of them up to
       64 Ki tiles).  A context takes it if the allocation succeeds; rg_workspace_bytes does not count it. */
    L.plane_cap = (uint32_t)(ntiles < 65536 ? ntiles : 65536);
    L.off_planes = o; o = align_up(o + (size_t)L.plane_cap * 3 * C::NW * sizeof(uint32_t), 256);
    L.total_with_planes = o;
    return RG_OK;
}

static int layout_for(int64_t nx, int64_t ny, int64_t nz, int nseeds, Layout &L, bool honour_env = true)
{
    return nz == 1 ? make_layout<Cfg2D>(nx, ny, nz, nseeds, L, honour_env) : make_layout<Cfg3D>(nx, ny, nz, nseeds, L, honour_env);
}

extern "C" int64_t rg_workspace_bytes(int64_t nx, int64_t ny, int64_t nz, int nseeds)
{
    if (nx <= 0 || ny <= 0 || nz <= 0 || nseeds < 0) return fail(RG_ERR_ARG, "bad size");
    Layout L; int rc = layout_for(nx, ny, nz, nseeds, L);
    return rc ? rc : (int64_t)L.total;
}

extern "C" int rg_ctx_create(int device, rg_ctx **out)
{
    if (!out) return fail(RG_ERR_ARG, "out == NULL");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) return fail(RG_ERR_CUDA, "no CUDA device (this library has no CPU path)", e);
    if (device < 0 || device >= ndev) return fail(RG_ERR_ARG, "bad device index");
    ON_DEVICE(device);
    rg_ctx *c = (rg_ctx *)calloc(1, sizeof(rg_ctx));
    if (!c) return fail(RG_ERR_NOMEM, "calloc");
    c->device = device;
    c->may_grow = true;
    e = cudaHostAlloc((void **)&c->h_status, CTR_WORDS * sizeof(uint32_t), cudaHostAllocDefault);
    if (e != cudaSuccess) { free(c); return fail(RG_ERR_CUDA, "cudaHostAlloc", e); }
    *out = c;
    return RG_OK;
}

extern "C" void rg_ctx_destroy(rg_ctx *c)
{
    if (!c) return;
    DeviceScope scope(c->device);
    cudaFree(c->ws); cudaFree(c->d_image); cudaFree(c->d_mask); cudaFree(c->d_seeds);
    cudaFreeHost(c->h_status); cudaFreeHost(c->h_boxes); cudaFreeHost(c->bounce);
    if (c->ev_made) for (int i = 0; i < 7; i++) cudaEventDestroy(c->ev[i]);
    free(c);
}

/* Grow the context's scratch to `bytes`.  The larger block is allocated FIRST and the old one freed only when that
   worked, so a failed attempt leaves the context as it was (nothing to synchronise, nothing lost).  `required` = false
   is the optional link-plane tail: its failure is remembered per size and not retried on every call. */
static int ensure_ws(rg_ctx *c, size_t bytes, bool required = true)
{
    if (bytes <= c->ws_bytes) return RG_OK;
    if (!c->may_grow) return fail(RG_ERR_NOMEM, "the scratch of this context is fixed and too small");
    if (!required && c->tail_refused && bytes >= c->tail_refused) return RG_ERR_NOMEM;
    ON_DEVICE(c->device);
    void *bigger = nullptr;
    cudaError_t e = cudaMalloc(&bigger, bytes);
    if (e != cudaSuccess && c->ws && required) {
        /* not enough room for both: give the old block back and try once more */
        cudaGetLastError();
        CK(cudaDeviceSynchronize()); cudaFree(c->ws); c->ws = nullptr; c->ws_bytes = 0;
        e = cudaMalloc(&bigger, bytes);
    }
    if (e != cudaSuccess) {
        cudaGetLastError();
        if (!required) { c->tail_refused = bytes; return RG_ERR_NOMEM; }
        return fail(RG_ERR_NOMEM, "cudaMalloc(workspace)", e);
    }
    if (c->ws) { CK(cudaDeviceSynchronize()); cudaFree(c->ws); }   /* kernels of an earlier call may still use the old block */
    c->ws = bigger; c->ws_bytes = bytes;
    return RG_OK;
}

extern "C" int rg_ctx_reserve(rg_ctx *c, int64_t nx, int64_t ny, int64_t nz, int nseeds)
{
    if (!c || nx <= 0 || ny <= 0 || nz <= 0 || nseeds < 0) return fail(RG_ERR_ARG, "bad argument");
    Layout L; int rc = layout_for(nx, ny, nz, nseeds, L);
    if (rc) return rc;
    if (ensure_ws(c, L.total_with_planes, false) == RG_OK) return RG_OK;      /* with the optional link-plane tail if it fits */
    return ensure_ws(c, L.total);
}

template <class C>
static int set_attrs()
{
    CK(cudaFuncSetAttribute(k_pass_a_fast<C>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(FastSmem<typename LeanOf<C>::type>)));
    CK(cudaFuncSetAttribute(k_pass_a_full<C>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(TileSmem<C>)));
    CK(cudaFuncSetAttribute(k_pass_c<C>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(FloodSmem<typename LeanOf<C>::type>)));
    return RG_OK;
}

/* ------------------------------------------------------------------ the launch sequence */
static unsigned flag_blocks(uint32_t ntiles)        /* one warp per tile, grid-stride beyond 32 CTAs per SM */
{
    const unsigned wpb = RG_FLAG_THREADS / 32, b = (ntiles + wpb - 1) / wpb;
    return b < 1 ? 1 : b > 148u * 32u ? 148u * 32u : b;
}

template <class C>
static int launch_label(rg_ctx *c, const Layout &L, const uint8_t *d_image, int64_t nx, int64_t ny, int64_t nz,
                       int lo_ext, int hi_ext, const int32_t *d_seeds, int nseeds, int threshold,
                       uint8_t *d_mask, cudaStream_t st)
{
    char *ws = (char *)c->ws;
    Params P;
    memset(&P, 0, sizeof P);
    P.image = d_image; P.mask = d_mask; P.nx = nx; P.ny = ny; P.nz = nz;
    P.ntx = L.ntx; P.nty = L.nty; P.ntz = L.ntz;
    P.lo_ext = lo_ext; P.hi_ext = hi_ext;
    P.open = c->next_open; c->next_open = 0;
    P.tmode = threshold <= 0 ? 0 : threshold >= 256 ? 2 : 1;
    P.t4 = (uint32_t)(threshold & 0xFF) * 0x01010101u;
    P.thr = threshold <= 0 ? 0u : threshold >= 256 ? 256u : (uint32_t)threshold;
    P.fast16 = (nx % 16 == 0) && (((uintptr_t)d_image & 15) == 0) && (((uintptr_t)d_mask & 15) == 0);
    P.nseeds = nseeds;
    { static int dbg = -1; if (dbg < 0) { const char *e = getenv("RG_DEBUG_STOP"); dbg = e ? atoi(e) : 0; } P.debug_stop = dbg; }
    { static int nr = -1; if (nr < 0) { const char *e = getenv("RG_NO_ROWS"); nr = e ? atoi(e) : 0; } P.no_rows = nr; }
    { static int fd = -1; if (fd < 0) { const char *e = getenv("RG_FORCE_DENSE"); fd = e ? atoi(e) : 0; } P.force_dense = fd; }   /* testing: every non-uniform tile takes the dense path */
    memset(&c->last_tmap, 0, sizeof c->last_tmap);
    /* the tensor map covers the neighbours' halo planes (volumes) or halo rows (images) as well */
    const int64_t ey = ext_ylo<C>(P) + ext_yhi<C>(P), ez = ext_zlo<C>(P) + ext_zhi<C>(P);
    const uint8_t *map_base = d_image - (size_t)ext_zlo<C>(P) * nx * ny - (size_t)ext_ylo<C>(P) * nx;
    P.use_tma = (P.fast16 && make_tmap<C>(&c->last_tmap, map_base, nx, ny + ey, nz + ez)
                 && make_tmap<typename LeanOf<C>::type>(&c->last_tmap_fast, map_base, nx, ny + ey, nz + ez)
                 && make_tmap_mask<C>(&c->last_tmap_mask, d_mask, nx, ny, nz, (int)sizeof(((FastSmem<typename LeanOf<C>::type> *)0)->tab))) ? 1 : 0;
    P.rec = (TileRec *)(ws + L.off_rec); P.ent = (uint16_t *)(ws + L.off_ent);
    P.nparent = (uint32_t *)(ws + L.off_npar); P.nrep = (uint16_t *)(ws + L.off_nrep);
    P.counters = (uint32_t *)(ws + L.off_ctr); P.node_cap = L.node_cap;
    P.complex_list = (uint32_t *)(ws + L.off_clist); P.flag_list = (uint32_t *)(ws + L.off_flist);
    const bool have_planes = c->ws_bytes >= L.total_with_planes && !getenv("RG_NO_PLANES");
    P.planes = have_planes ? (uint32_t *)(ws + L.off_planes) : nullptr; P.plane_cap = have_planes ? L.plane_cap : 0u;
    uint64_t *keys = (uint64_t *)(ws + L.off_keys), *keys2 = (uint64_t *)(ws + L.off_keys2);
    P.seed_keys = keys;

    P.mask_prezeroed = getenv("RG_NO_PREZERO") ? 0 : 1;          /* the labelling CTAs clear their own mask tile while their load is in flight */
    int kb = nseeds > 0 ? (nseeds + 255) / 256 : 1; if (kb > 1024) kb = 1024;
    if (c->profile) cudaEventRecord(c->ev[0], st);
    k_seed_keys<C><<<kb, 256, 0, st>>>(P, d_seeds, keys);
    if (nseeds > 2048) {
        size_t tb = L.cub_bytes;
        CK(cub::DeviceRadixSort::SortKeys(ws + L.off_cub, tb, keys, keys2, nseeds, 0, 64, st));
        P.seed_keys = keys2;
    } else if (nseeds > 1) {
        k_sort_small<<<1, 1024, 0, st>>>(keys, nseeds);
    }
    if (c->profile) cudaEventRecord(c->ev[1], st);
    const bool g3 = L.nty <= 65535 && L.ntz <= 65535;
    const dim3 tile_grid = g3 ? dim3((unsigned)L.ntx, (unsigned)L.nty, (unsigned)L.ntz) : dim3(L.ntiles, 1, 1);
    c->last_grid = tile_grid;
    {
        static int pad = -1;                                        /* tuning: RG_FAST_PAD_SMEM=bytes lowers the CTAs per SM */
        if (pad < 0) { const char *e = getenv("RG_FAST_PAD_SMEM"); pad = e ? atoi(e) : 0;
                       if (pad) cudaFuncSetAttribute(k_pass_a_fast<C>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(FastSmem<typename LeanOf<C>::type>) + pad); }
        k_pass_a_fast<C><<<tile_grid, RG_FAST_THREADS, sizeof(FastSmem<typename LeanOf<C>::type>) + pad, st>>>(P, c->last_tmap_fast, c->last_tmap_mask);
    }
    if (c->profile) cudaEventRecord(c->ev[2], st);
    if (!P.debug_stop || P.debug_stop >= 10) k_pass_a_full<C><<<RG_PERSISTENT_CTAS, RG_THREADS, sizeof(TileSmem<C>), st>>>(P, c->last_tmap);
    if (c->profile) cudaEventRecord(c->ev[3], st);
    int64_t items = (int64_t)L.ntiles * MERGE_LANES;                 /* one warp per tile */
    int mb = (int)((items + 255) / 256); if (mb > 148 * 64) mb = 148 * 64;
    if (!P.debug_stop || P.debug_stop >= 20) k_merge<C><<<mb, 256, 0, st>>>(P);
    if (c->profile) cudaEventRecord(c->ev[4], st);
    CK(cudaGetLastError());
    c->last_counters = P.counters; c->last_stream = st;
    c->last_P = P; c->last_is2d = (C::TZ == 1); c->last_ntiles = L.ntiles; c->have_last = true;
    c->if_off[0] = (uint32_t *)(ws + L.off_if0); c->if_off[1] = (uint32_t *)(ws + L.off_if1);
    c->d_rargs = (ResolveArgs *)(ws + L.off_rargs);
    c->d_boxes = (int32_t *)(ws + L.off_boxes);
    return RG_OK;
}

static int launch_finish(rg_ctx *c, cudaStream_t st)
{
    if (!c->have_last) return fail(RG_ERR_ARG, "finish without a label pass");
    if (c->last_P.debug_stop && c->last_P.debug_stop < 20) { if (c->profile) { cudaEventRecord(c->ev[5], st); cudaEventRecord(c->ev[6], st); } return RG_OK; }   /* tuning runs time the fast pass only */
    if (c->last_is2d) {
        k_flag_tiles<Cfg2D><<<flag_blocks(c->last_ntiles), RG_FLAG_THREADS, 0, st>>>(c->last_P);
        if (c->profile) cudaEventRecord(c->ev[5], st);
        k_pass_c<Cfg2D><<<RG_PASSC_CTAS, RG_PASSC_THREADS, sizeof(FloodSmem<LeanOf<Cfg2D>::type>), st>>>(c->last_P, c->last_tmap_fast);
    } else {
        k_flag_tiles<Cfg3D><<<flag_blocks(c->last_ntiles), RG_FLAG_THREADS, 0, st>>>(c->last_P);
        if (c->profile) cudaEventRecord(c->ev[5], st);
        if (c->last_P.planes) k_rows_out<Cfg3D><<<RG_ROWS_CTAS, RG_ROWS_THREADS, 0, st>>>(c->last_P);      /* row tiles exist only with link-plane slots */
        k_pass_c<Cfg3D><<<RG_PASSC_CTAS, RG_PASSC_THREADS, sizeof(FloodSmem<LeanOf<Cfg3D>::type>), st>>>(c->last_P, c->last_tmap_fast);
    }
    if (c->profile) cudaEventRecord(c->ev[6], st);
    CK(cudaGetLastError());
    c->last_stream = st;
    return RG_OK;
}

static int label_device_impl(rg_ctx *c, const uint8_t *d_image, int64_t nx, int64_t ny, int64_t nz, int lo_ext, int hi_ext,
                            const int32_t *d_seeds, int nseeds, int threshold, uint8_t *d_mask, void *stream)
{
    if (!c || !d_image || !d_mask || nx <= 0 || ny <= 0 || nz <= 0 || nseeds < 0 || (nseeds > 0 && !d_seeds))
        return fail(RG_ERR_ARG, "bad argument");
    ON_DEVICE(c->device);
    if (!c->attr_set) {
        int rc = set_attrs<Cfg3D>(); if (rc) return rc;
        rc = set_attrs<Cfg2D>(); if (rc) return rc;
        c->attr_set = true;
    }
    Layout L; int rc = layout_for(nx, ny, nz, nseeds, L, !c->worst_case_layout);
    if (rc) return rc;
    /* A context whose scratch already holds this grow (rg_ctx_reserve, or an earlier call) is not touched: no allocation, no
       synchronisation inside the asynchronous path.  Only a context that has to grow anyway tries for the optional tail too. */
    if (c->ws_bytes < L.total) {
        if (c->may_grow) ensure_ws(c, L.total_with_planes, false);
        rc = ensure_ws(c, L.total); if (rc) return rc;
    }
    cudaStream_t st = (cudaStream_t)stream;
    return nz == 1 ? launch_label<Cfg2D>(c, L, d_image, nx, ny, nz, lo_ext, hi_ext, d_seeds, nseeds, threshold, d_mask, st)
                   : launch_label<Cfg3D>(c, L, d_image, nx, ny, nz, lo_ext, hi_ext, d_seeds, nseeds, threshold, d_mask, st);
}

static int grow_device_impl(rg_ctx *c, const uint8_t *d_image, int64_t nx, int64_t ny, int64_t nz, int lo_ext, int hi_ext,
                            const int32_t *d_seeds, int nseeds, int threshold, uint8_t *d_mask, void *stream)
{
    if (!c) return fail(RG_ERR_ARG, "ctx == NULL");
    ON_DEVICE(c->device);                                   /* both halves run on the context's device */
    int rc = label_device_impl(c, d_image, nx, ny, nz, lo_ext, hi_ext, d_seeds, nseeds, threshold, d_mask, stream);
    if (rc) return rc;
    return launch_finish(c, (cudaStream_t)stream);
}

/* ------------------------------------------------------------------ slab (multi-GPU) phases: z-slabs of a volume, row slabs of an image */
extern "C" int rg_slab_label(rg_ctx *c, const uint8_t *d_slab, int64_t nx, int64_t ny, int64_t nz_local, int has_lower, int has_upper,
                             const int32_t *d_seeds_xyz, int nseeds, int threshold, uint8_t *d_mask_slab, void *stream)
{
    if (nz_local < 1 || ny < 1) return fail(RG_ERR_ARG, "empty slab");
    /* nz_local == 1 is an image cut into row slabs (ny = this rank's rows); a neighbour needs two rows / planes to face */
    if ((has_lower || has_upper) && (nz_local == 1 ? ny : nz_local) < 2)
        return fail(RG_ERR_ARG, "a slab with neighbours needs at least 2 planes (volume) or 2 rows (image) per rank");
    return label_device_impl(c, d_slab, nx, ny, nz_local, has_lower ? 1 : 0, has_upper ? 1 : 0, d_seeds_xyz, nseeds, threshold, d_mask_slab, stream);
}

extern "C" int64_t rg_slab_interface_capacity(int64_t nx, int64_t ny, int64_t nz_local)
{
    if (nz_local == 1) return (nx + Cfg2D::TX - 1) / Cfg2D::TX * (int64_t)Cfg2D::FYN;      /* one tile row's y faces */
    int64_t ntx = (nx + Cfg3D::TX - 1) / Cfg3D::TX, nty = (ny + Cfg3D::TY - 1) / Cfg3D::TY;
    return ntx * nty * (int64_t)Cfg3D::FZN;
}

/* one interface of the last labelled slab: entry count to *d_count, local roots to d_roots (at most cap of them) */
static void export_side(rg_ctx *c, int side, uint32_t *d_count, uint32_t *d_roots, uint32_t *d_counts_base, uint32_t cap, cudaStream_t st)
{
    const Params &P = c->last_P;
    if (c->last_is2d) {
        const int b = (iface_tiles<Cfg2D>(P) * EXPORT_LANES + 127) / 128;
        k_export_offsets<Cfg2D><<<1, 1024, 0, st>>>(P, side, c->if_off[side], d_count);
        k_export_entries<Cfg2D><<<b, 128, 0, st>>>(P, side, c->if_off[side], d_roots, d_counts_base, cap);
    } else {
        const int b = (iface_tiles<Cfg3D>(P) * EXPORT_LANES + 127) / 128;
        k_export_offsets<Cfg3D><<<1, 1024, 0, st>>>(P, side, c->if_off[side], d_count);
        k_export_entries<Cfg3D><<<b, 128, 0, st>>>(P, side, c->if_off[side], d_roots, d_counts_base, cap);
    }
}

extern "C" int rg_slab_export(rg_ctx *c, int side, uint32_t *d_roots_out, uint32_t *d_counts3, void *stream)
{
    if (!c || !c->have_last || !d_roots_out || !d_counts3 || side < 0 || side > 1) return fail(RG_ERR_ARG, "bad argument");
    ON_DEVICE(c->device);
    export_side(c, side, d_counts3 + side, d_roots_out, d_counts3, 0xFFFFFFFFu, (cudaStream_t)stream);
    CK(cudaGetLastError());
    return RG_OK;
}

static int blocks_for(uint32_t n) { int b = (int)((n + 255u) / 256u); return b < 1 ? 1 : b > 148 * 16 ? 148 * 16 : b; }

extern "C" int rg_slab_resolve_init(rg_ctx *c, uint32_t *d_xpar, const uint32_t *d_roots, uint32_t n, uint32_t vertex_offset, void *stream)
{
    if (!c || !d_xpar || (n && !d_roots)) return fail(RG_ERR_ARG, "bad argument");
    ON_DEVICE(c->device);
    k_resolve_init<<<blocks_for(n), 256, 0, (cudaStream_t)stream>>>(d_xpar, d_roots, n, vertex_offset);
    CK(cudaGetLastError());
    return RG_OK;
}

extern "C" int rg_slab_resolve_union(rg_ctx *c, uint32_t *d_xpar, const uint32_t *d_upper_of_lower_rank, uint32_t off_lower_rank,
                                     const uint32_t *d_lower_of_upper_rank, uint32_t off_upper_rank, uint32_t n, void *stream)
{
    if (!c || !d_xpar || (n && (!d_upper_of_lower_rank || !d_lower_of_upper_rank))) return fail(RG_ERR_ARG, "bad argument");
    ON_DEVICE(c->device);
    if (!n) return RG_OK;
    k_resolve_union<<<blocks_for(n), 256, 0, (cudaStream_t)stream>>>(d_xpar, d_upper_of_lower_rank, off_lower_rank, d_lower_of_upper_rank, off_upper_rank, n);
    CK(cudaGetLastError());
    return RG_OK;
}

extern "C" int rg_slab_resolve_apply(rg_ctx *c, uint32_t *d_xpar, const uint32_t *d_own_roots, uint32_t n, uint32_t own_vertex_offset, void *stream)
{
    if (!c || !c->have_last || !d_xpar || (n && !d_own_roots)) return fail(RG_ERR_ARG, "bad argument");
    ON_DEVICE(c->device);
    if (!n) return RG_OK;
    k_resolve_apply<<<blocks_for(n), 256, 0, (cudaStream_t)stream>>>(c->last_P, d_xpar, d_own_roots, n, own_vertex_offset);
    CK(cudaGetLastError());
    return RG_OK;
}

extern "C" int rg_slab_resolve(rg_ctx *c, uint32_t *d_xpar, const uint32_t *d_lists, uint32_t stride, const int64_t *counts4,
                               int rank, int world, void *stream)
{
    if (!c || !c->have_last || !d_xpar || !d_lists || !counts4 || world < 2 || world > RESOLVE_MAX_RANKS || rank < 0 || rank >= world || !stride)
        return fail(RG_ERR_ARG, "bad argument");
    ON_DEVICE(c->device);
    ResolveArgs A; memset(&A, 0, sizeof A);
    A.world = (uint32_t)world; A.rank = (uint32_t)rank; A.stride = stride;
    uint64_t off = 1;
    for (int r = 0; r < world; r++) {
        A.cnt[r][0] = (uint32_t)counts4[4 * r]; A.cnt[r][1] = (uint32_t)counts4[4 * r + 1];
        if (A.cnt[r][0] > stride || A.cnt[r][1] > stride) return fail(RG_ERR_ARG, "interface list longer than the stride");
        A.off[r] = (uint32_t)off; off += (uint64_t)counts4[4 * r + 2];
        if (off >= (1ull << 32)) return fail(RG_ERR_TOO_LARGE, "too many nodes across ranks for 32-bit vertex ids");
    }
    for (int r = 0; r + 1 < world; r++)
        if (A.cnt[r][1] != A.cnt[r + 1][0]) return fail(RG_ERR_INTERNAL, "interface lists of neighbouring ranks differ in length");
    cudaStream_t st = (cudaStream_t)stream;
    k_resolve_init_all<<<blocks_for((uint32_t)world * 2u * stride), 256, 0, st>>>(d_xpar, d_lists, A);
    k_resolve_union_all<<<blocks_for((uint32_t)(world - 1) * stride), 256, 0, st>>>(d_xpar, d_lists, A);
    k_resolve_apply_all<<<blocks_for(2u * stride), 256, 0, st>>>(c->last_P, d_xpar, d_lists, A);
    CK(cudaGetLastError());
    return RG_OK;
}

extern "C" int rg_slab_export_packed(rg_ctx *c, uint32_t *d_packed, uint32_t cap, void *stream)
{
    if (!c || !c->have_last || !d_packed || !cap) return fail(RG_ERR_ARG, "bad argument");
    ON_DEVICE(c->device);
    cudaStream_t st = (cudaStream_t)stream;
    const Params &P = c->last_P;
    CK(cudaMemsetAsync(d_packed, 0, 16, st));
    for (int side = 0; side < 2; side++) {
        if (!(side ? P.hi_ext : P.lo_ext)) continue;
        export_side(c, side, d_packed + side, d_packed + 4 + (size_t)side * cap, d_packed, cap, st);
    }
    CK(cudaGetLastError());
    return RG_OK;
}

extern "C" int rg_slab_resolve_packed(rg_ctx *c, uint32_t *d_xpar, uint32_t xpar_cap, const uint32_t *d_packed_all, uint32_t cap,
                                      int rank, int world, void *stream)
{
    if (!c || !c->have_last || !d_xpar || !d_packed_all || !cap || world < 2 || world > RESOLVE_MAX_RANKS || rank < 0 || rank >= world)
        return fail(RG_ERR_ARG, "bad argument");
    ON_DEVICE(c->device);
    cudaStream_t st = (cudaStream_t)stream;
    const uint32_t stride = 4u + 2u * cap;
    k_resolve_prepare<<<1, 32, 0, st>>>(c->last_P, d_packed_all, stride, cap, xpar_cap, (uint32_t)rank, (uint32_t)world, c->d_rargs, d_xpar, nullptr);
    k_resolve_init_packed<<<blocks_for((uint32_t)world * 2u * cap), 256, 0, st>>>(d_xpar, d_packed_all, cap, c->d_rargs);
    k_resolve_union_packed<<<blocks_for((uint32_t)(world - 1) * cap), 256, 0, st>>>(d_xpar, d_packed_all, cap, c->d_rargs);
    k_resolve_apply_packed<<<blocks_for(2u * cap), 256, 0, st>>>(c->last_P, d_xpar, d_packed_all, cap, c->d_rargs);
    CK(cudaGetLastError());
    return RG_OK;
}

/* ---- blocks of an image (rows x columns) ---- */
extern "C" int rg_block_label(rg_ctx *c, const uint8_t *d_block, int64_t w, int64_t h, int open_sides, const int32_t *d_seeds_xyz, int nseeds,
                              int threshold, uint8_t *d_mask_block, void *stream)
{
    if (!c || w < 1 || h < 1 || (open_sides & ~15)) return fail(RG_ERR_ARG, "bad argument");
    if ((open_sides & 2) && w % Cfg2D::TX) return fail(RG_ERR_ARG, "a block with a seam on its +x border must be a multiple of 128 pixels wide");
    c->next_open = (uint32_t)open_sides;
    int rc = label_device_impl(c, d_block, w, h, 1, 0, 0, d_seeds_xyz, nseeds, threshold, d_mask_block, stream);
    c->next_open = 0;
    return rc;
}

extern "C" int rg_block_export(rg_ctx *c, int side, uint32_t *d_roots, uint8_t *d_values, void *stream)
{
    if (!c || !c->have_last || !c->last_is2d || side < 0 || side > 3 || !d_roots || !d_values) return fail(RG_ERR_ARG, "bad argument");
    if (!(c->last_P.open & (1u << side))) return fail(RG_ERR_ARG, "that border of the block was not labelled as a seam");
    ON_DEVICE(c->device);
    const int64_t n = side < 2 ? c->last_P.ny : c->last_P.nx;
    k_export_border<Cfg2D><<<blocks_for((uint32_t)n), 256, 0, (cudaStream_t)stream>>>(c->last_P, side, d_roots, d_values);
    CK(cudaGetLastError());
    return RG_OK;
}

extern "C" int rg_block_resolve_union(rg_ctx *c, uint32_t *d_xpar, const uint32_t *d_roots_a, uint32_t off_a, const uint8_t *d_values_a,
                                      const uint32_t *d_roots_b, uint32_t off_b, const uint8_t *d_values_b, uint32_t n, int threshold, void *stream)
{
    if (!c || !d_xpar || (n && (!d_roots_a || !d_roots_b || !d_values_a || !d_values_b))) return fail(RG_ERR_ARG, "bad argument");
    ON_DEVICE(c->device);
    if (!n) return RG_OK;
    const uint32_t thr = threshold <= 0 ? 0u : threshold >= 256 ? 256u : (uint32_t)threshold;
    k_resolve_union_seam<<<blocks_for(n), 256, 0, (cudaStream_t)stream>>>(d_xpar, d_roots_a, off_a, d_values_a, d_roots_b, off_b, d_values_b, n, thr);
    CK(cudaGetLastError());
    return RG_OK;
}

/* the same resolve, reading every rank's packed record where that rank exported it (d_peer_records: device array of `world`
   pointers, entry r = rank r's export buffer as mapped into this process): no gather step, the lists cross NVLink inside the
   resolve kernels */
extern "C" int rg_slab_resolve_peers(rg_ctx *c, uint32_t *d_xpar, uint32_t xpar_cap, const uint32_t *const *d_peer_records, uint32_t cap,
                                     int rank, int world, void *stream)
{
    if (!c || !c->have_last || !d_xpar || !d_peer_records || !cap || world < 2 || world > RESOLVE_MAX_RANKS || rank < 0 || rank >= world)
        return fail(RG_ERR_ARG, "bad argument");
    ON_DEVICE(c->device);
    cudaStream_t st = (cudaStream_t)stream;
    k_resolve_prepare<<<1, 32, 0, st>>>(c->last_P, nullptr, 4u + 2u * cap, cap, xpar_cap, (uint32_t)rank, (uint32_t)world, c->d_rargs, d_xpar, d_peer_records);
    k_resolve_init_packed<<<blocks_for((uint32_t)world * 2u * cap), 256, 0, st>>>(d_xpar, nullptr, cap, c->d_rargs);
    k_resolve_union_packed<<<blocks_for((uint32_t)(world - 1) * cap), 256, 0, st>>>(d_xpar, nullptr, cap, c->d_rargs);
    k_resolve_apply_packed<<<blocks_for(2u * cap), 256, 0, st>>>(c->last_P, d_xpar, nullptr, cap, c->d_rargs);
    CK(cudaGetLastError());
    return RG_OK;
}

/* Buffers other processes on this node can map: plain cudaMalloc blocks (an allocation base, which is what an IPC handle names)
   and the 64-byte handles to pass around (any transport: torch.distributed, MPI, a pipe). */
extern "C" int rg_ipc_alloc(int64_t bytes, void **d_ptr_out, unsigned char handle_out[64])
{
    if (bytes <= 0 || !d_ptr_out || !handle_out) return fail(RG_ERR_ARG, "bad argument");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handles are 64 bytes");
    CK(cudaMalloc(d_ptr_out, (size_t)bytes));
    cudaIpcMemHandle_t h;
    cudaError_t e = cudaIpcGetMemHandle(&h, *d_ptr_out);
    if (e != cudaSuccess) { cudaFree(*d_ptr_out); *d_ptr_out = nullptr; return fail(RG_ERR_CUDA, "cudaIpcGetMemHandle", e); }
    memcpy(handle_out, &h, 64);
    return RG_OK;
}
extern "C" int rg_ipc_open(const unsigned char handle[64], void **d_ptr_out)
{
    if (!handle || !d_ptr_out) return fail(RG_ERR_ARG, "bad argument");
    cudaIpcMemHandle_t h; memcpy(&h, handle, 64);
    CK(cudaIpcOpenMemHandle(d_ptr_out, h, cudaIpcMemLazyEnablePeerAccess));
    return RG_OK;
}
extern "C" int rg_ipc_close(void *d_ptr) { if (d_ptr) CK(cudaIpcCloseMemHandle(d_ptr)); return RG_OK; }
extern "C" int rg_ipc_free(void *d_ptr) { if (d_ptr) CK(cudaFree(d_ptr)); return RG_OK; }

extern "C" int rg_slab_finish(rg_ctx *c, void *stream)
{
    if (!c) return fail(RG_ERR_ARG, "ctx == NULL");
    ON_DEVICE(c->device);
    return launch_finish(c, (cudaStream_t)stream);
}

extern "C" int rg_grow_device(rg_ctx *c, const uint8_t *d_image, int64_t nx, int64_t ny, int64_t nz,
                              const int32_t *d_seeds_xyz, int nseeds, int threshold, uint8_t *d_mask, void *stream)
{
    return grow_device_impl(c, d_image, nx, ny, nz, 0, 0, d_seeds_xyz, nseeds, threshold, d_mask, stream);
}

extern "C" int rg_ctx_status(rg_ctx *c, uint32_t stats[4])
{
    if (!c) return fail(RG_ERR_ARG, "ctx == NULL");
    ON_DEVICE(c->device);
    if (!c->last_counters) { if (stats) memset(stats, 0, 16); return RG_OK; }
    CK(cudaMemcpyAsync(c->h_status, c->last_counters, CTR_WORDS * sizeof(uint32_t), cudaMemcpyDeviceToHost, c->last_stream));
    CK(cudaStreamSynchronize(c->last_stream));
    /* stats[2]: tiles the walkers did not get through = labelled by the dense path + handed to the full pass */
    if (stats) { stats[0] = c->h_status[CTR_NODES]; stats[1] = c->h_status[CTR_STATUS]; stats[2] = c->h_status[CTR_NCOMPLEX] + c->h_status[CTR_NDENSE] + c->h_status[CTR_NROWTILES]; stats[3] = c->h_status[CTR_NFLAGGED] + c->h_status[CTR_NROWOUT]; }
    c->last_full_tiles = c->h_status[CTR_NCOMPLEX];
    if (c->h_status[CTR_STATUS]) {
        if (c->h_status[CTR_STATUS] & ST_NODE_OVERFLOW)
            snprintf(g_err, sizeof g_err, "out of node ids (status=0x%x): this input needs more nodes per tile than the workspace holds "
                     "(volumes beyond ~17 Gvoxel and RG_NODE_BUDGET runs size it for an average, not the worst case)", c->h_status[CTR_STATUS]);
        else
            snprintf(g_err, sizeof g_err, "engine self-check failed: status=0x%x mismatching faces=%u", c->h_status[CTR_STATUS], c->h_status[CTR_FACE_MISMATCH]);
        return RG_ERR_INTERNAL;
    }
    return RG_OK;
}

extern "C" int rg_ctx_profile(rg_ctx *c, int enable)
{
    if (!c) return fail(RG_ERR_ARG, "ctx == NULL");
    ON_DEVICE(c->device);
    if (enable && !c->ev_made) { for (int i = 0; i < 7; i++) CK(cudaEventCreate(&c->ev[i])); c->ev_made = true; }
    c->profile = enable != 0;
    return RG_OK;
}

extern "C" int rg_ctx_profile_read(rg_ctx *c, float ms[6])
{
    if (!c || !ms || !c->ev_made) return fail(RG_ERR_ARG, "profiling was not enabled");
    ON_DEVICE(c->device);
    CK(cudaEventSynchronize(c->ev[6]));
    for (int i = 0; i < 6; i++) CK(cudaEventElapsedTime(&ms[i], c->ev[i], c->ev[i + 1]));
    return RG_OK;
}

extern "C" int rg_apply_mask_device(rg_ctx *c, uint8_t *d_image, const uint8_t *d_mask, int64_t n, void *stream)
{
    if (!c || !d_image || !d_mask || n < 0) return fail(RG_ERR_ARG, "bad argument");
    ON_DEVICE(c->device);
    if (n == 0) return RG_OK;
    int64_t b = (n + 255) / 256; if (b > 148 * 32) b = 148 * 32;
    k_apply_mask<<<(int)b, 256, 0, (cudaStream_t)stream>>>(d_image, d_mask, n);
    CK(cudaGetLastError());
    return RG_OK;
}

/* ------------------------------------------------------------------ host-pointer entry points */
static std::mutex g_mu;
static rg_ctx *g_default[64];

static int default_ctx(rg_ctx **out)
{
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return fail(RG_ERR_CUDA, "no CUDA device (this library has no CPU path)", e);
    if (dev < 0 || dev >= 64) return fail(RG_ERR_ARG, "device index");
    if (!g_default[dev]) { int rc = rg_ctx_create(dev, &g_default[dev]); if (rc) return rc; }
    *out = g_default[dev];
    return RG_OK;
}

static int ensure_io(rg_ctx *c, size_t bytes, int nseeds)
{
    if (bytes > c->io_bytes) {
        cudaFree(c->d_image); cudaFree(c->d_mask); c->d_image = c->d_mask = nullptr; c->io_bytes = 0;
        cudaError_t e = cudaMalloc((void **)&c->d_image, bytes);
        if (e != cudaSuccess) return fail(RG_ERR_NOMEM, "cudaMalloc(image)", e);
        e = cudaMalloc((void **)&c->d_mask, bytes);
        if (e != cudaSuccess) return fail(RG_ERR_NOMEM, "cudaMalloc(mask)", e);
        c->io_bytes = bytes;
    }
    if ((size_t)nseeds > c->seeds_cap) {
        cudaFree(c->d_seeds); c->d_seeds = nullptr; c->seeds_cap = 0;
        size_t cap = (size_t)nseeds < 64 ? 64 : (size_t)nseeds;
        cudaError_t e = cudaMalloc((void **)&c->d_seeds, cap * 12);
        if (e != cudaSuccess) return fail(RG_ERR_NOMEM, "cudaMalloc(seeds)", e);
        c->seeds_cap = cap;
    }
    return RG_OK;
}

/* ---- moving host buffers.  Pinned (page-locked) buffers go to the copy engine as they are.  A pageable buffer -- what
   malloc / calloc return, i.e. what the reference's callers pass (raycast.cu:147, 705) -- would make the driver stage the whole
   copy through one internal buffer on the calling thread; here a few host threads copy chunks into pinned bounce buffers and
   start the DMA of each chunk at once, so the host copies run in parallel with each other and with the DMA. */
enum { STAGE_THREADS_MAX = 16, STAGE_CHUNK = 4 << 20, STAGE_SLOTS = 2 };
static int stage_threads()
{
    static const int n = [] {
        const char *e = getenv("RG_STAGE_THREADS");             /* tuning aid */
        int v = e ? atoi(e) : (int)(std::thread::hardware_concurrency() / 2);
        return v < 1 ? 1 : v > STAGE_THREADS_MAX ? STAGE_THREADS_MAX : v;
    }();
    return n;
}

static bool is_pageable(const void *p)
{
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return true; }
    return a.type == cudaMemoryTypeUnregistered;
}

static int ensure_bounce(rg_ctx *c)
{
    const size_t need = (size_t)STAGE_THREADS_MAX * STAGE_SLOTS * STAGE_CHUNK;
    if (c->bounce_bytes >= need) return RG_OK;
    CK(cudaHostAlloc((void **)&c->bounce, need, cudaHostAllocDefault));
    c->bounce_bytes = need;
    return RG_OK;
}

/* dir = 0: host -> device, 1: device -> host; returns when the data has arrived */
static int staged_copy(rg_ctx *c, int dir, uint8_t *host, uint8_t *dev, size_t bytes)
{
    int rc = ensure_bounce(c); if (rc) return rc;
    const size_t nchunks = (bytes + STAGE_CHUNK - 1) / STAGE_CHUNK;
    const int nthreads = stage_threads();
    int err[STAGE_THREADS_MAX] = {0};
    auto worker = [&](int t) {
        cudaStream_t st; cudaEvent_t ev[STAGE_SLOTS];
        if (cudaSetDevice(c->device) != cudaSuccess || cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking) != cudaSuccess) { err[t] = 1; return; }
        for (int k = 0; k < STAGE_SLOTS; k++) cudaEventCreateWithFlags(&ev[k], cudaEventDisableTiming);
        uint8_t *mine = c->bounce + (size_t)t * STAGE_SLOTS * STAGE_CHUNK;
        size_t pending_off[STAGE_SLOTS], pending_len[STAGE_SLOTS]; bool pending[STAGE_SLOTS] = {false};
        int slot = 0;
        for (size_t i = (size_t)t; i < nchunks && !err[t]; i += (size_t)nthreads, slot = (slot + 1) % STAGE_SLOTS) {
            const size_t off = i * STAGE_CHUNK, len = bytes - off < (size_t)STAGE_CHUNK ? bytes - off : (size_t)STAGE_CHUNK;
            uint8_t *buf = mine + (size_t)slot * STAGE_CHUNK;
            if (pending[slot]) {                            /* the slot's previous chunk: DMA done? (dir 1: then hand it to the caller) */
                if (cudaEventSynchronize(ev[slot]) != cudaSuccess) { err[t] = 1; break; }
                if (dir == 1) memcpy(host + pending_off[slot], buf, pending_len[slot]);
                pending[slot] = false;
            }
            if (dir == 0) {
                memcpy(buf, host + off, len);
                if (cudaMemcpyAsync(dev + off, buf, len, cudaMemcpyHostToDevice, st) != cudaSuccess) err[t] = 1;
            } else if (cudaMemcpyAsync(buf, dev + off, len, cudaMemcpyDeviceToHost, st) != cudaSuccess) err[t] = 1;
            cudaEventRecord(ev[slot], st);
            pending[slot] = true; pending_off[slot] = off; pending_len[slot] = len;
        }
        for (int k = 0; k < STAGE_SLOTS; k++) {
            const int sl = (slot + k) % STAGE_SLOTS;        /* oldest first */
            if (pending[sl]) {
                if (cudaEventSynchronize(ev[sl]) != cudaSuccess) err[t] = 1;
                else if (dir == 1) memcpy(host + pending_off[sl], mine + (size_t)sl * STAGE_CHUNK, pending_len[sl]);
            }
        }
        for (int k = 0; k < STAGE_SLOTS; k++) cudaEventDestroy(ev[k]);
        cudaStreamDestroy(st);
    };
    std::thread th[STAGE_THREADS_MAX];
    for (int t = 1; t < nthreads; t++) th[t] = std::thread(worker, t);
    worker(0);
    for (int t = 1; t < nthreads; t++) th[t].join();
    for (int t = 0; t < nthreads; t++) if (err[t]) return fail(RG_ERR_CUDA, "staged host copy failed", cudaGetLastError());
    c->last_staged = 1;
    return RG_OK;
}

static int copy_in(rg_ctx *c, uint8_t *dev, const uint8_t *host, size_t bytes, cudaStream_t st)
{
    c->last_h2d_bytes += (int64_t)bytes;
    if (bytes >= (size_t)(32 << 20) && !getenv("RG_NO_STAGING") && is_pageable(host)) {
        CK(cudaStreamSynchronize(st));
        return staged_copy(c, 0, const_cast<uint8_t *>(host), dev, bytes);
    }
    CK(cudaMemcpyAsync(dev, host, bytes, cudaMemcpyHostToDevice, st));
    return RG_OK;
}

static int copy_out_full(rg_ctx *c, uint8_t *host, const uint8_t *dev, size_t bytes, cudaStream_t st)
{
    c->last_d2h_bytes += (int64_t)bytes; c->last_copies++;
    if (bytes >= (size_t)(32 << 20) && !getenv("RG_NO_STAGING") && is_pageable(host)) {
        CK(cudaStreamSynchronize(st));
        return staged_copy(c, 1, host, const_cast<uint8_t *>(dev), bytes);
    }
    CK(cudaMemcpyAsync(host, dev, bytes, cudaMemcpyDeviceToHost, st));
    return RG_OK;
}

/* Output-sensitive copy of the mask: `host` is known to be all zeros (a fresh calloc, or the caller's promise), so only the
   boxes that hold region voxels are copied -- per plane of tiles, rows of tiles with the same x-extent merged into one
   three-dimensional copy.  C1's region is 1.6 % of its volume, C4's 7 %. */
static int copy_out_sparse(rg_ctx *c, uint8_t *host, int64_t nx, int64_t ny, int64_t nz, cudaStream_t st)
{
    const Params &P = c->last_P;
    const int TX = c->last_is2d ? Cfg2D::TX : Cfg3D::TX, TY = c->last_is2d ? Cfg2D::TY : Cfg3D::TY, TZ = c->last_is2d ? Cfg2D::TZ : Cfg3D::TZ;
    const size_t rows = (size_t)P.nty * P.ntz;
    if (c->h_boxes_cap < rows * 2) {
        cudaFreeHost(c->h_boxes); c->h_boxes = nullptr; c->h_boxes_cap = 0;
        CK(cudaHostAlloc((void **)&c->h_boxes, rows * 2 * sizeof(int32_t), cudaHostAllocDefault));
        c->h_boxes_cap = rows * 2;
    }
    const int blocks = (int)((rows + 255) / 256);
    if (c->last_is2d) k_region_boxes<Cfg2D><<<blocks, 256, 0, st>>>(P, c->d_boxes);
    else k_region_boxes<Cfg3D><<<blocks, 256, 0, st>>>(P, c->d_boxes);
    CK(cudaMemcpyAsync(c->h_boxes, c->d_boxes, rows * 2 * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    cudaMemcpy3DParms q;
    memset(&q, 0, sizeof q);
    q.srcPtr = make_cudaPitchedPtr(P.mask, (size_t)nx, (size_t)nx, (size_t)ny);
    q.dstPtr = make_cudaPitchedPtr(host, (size_t)nx, (size_t)nx, (size_t)ny);
    q.kind = cudaMemcpyDeviceToHost;
    for (int tz = 0; tz < P.ntz; tz++) {
        const int32_t *b = c->h_boxes + (size_t)tz * P.nty * 2;
        for (int ty = 0; ty < P.nty; ) {
            if (!b[2 * ty + 1]) { ty++; continue; }
            int t1 = ty + 1;
            while (t1 < P.nty && b[2 * t1] == b[2 * ty] && b[2 * t1 + 1] == b[2 * ty + 1]) t1++;
            const int64_t x0 = (int64_t)b[2 * ty] * TX, x1 = (int64_t)b[2 * ty + 1] * TX < nx ? (int64_t)b[2 * ty + 1] * TX : nx;
            const int64_t y0 = (int64_t)ty * TY, y1 = (int64_t)t1 * TY < ny ? (int64_t)t1 * TY : ny;
            const int64_t z0 = (int64_t)tz * TZ, z1 = (int64_t)(tz + 1) * TZ < nz ? (int64_t)(tz + 1) * TZ : nz;
            q.srcPos = make_cudaPos((size_t)x0, (size_t)y0, (size_t)z0); q.dstPos = q.srcPos;
            q.extent = make_cudaExtent((size_t)(x1 - x0), (size_t)(y1 - y0), (size_t)(z1 - z0));
            CK(cudaMemcpy3DAsync(&q, st));
            c->last_d2h_bytes += (x1 - x0) * (y1 - y0) * (z1 - z0); c->last_copies++;
            ty = t1;
        }
    }
    return RG_OK;
}

enum { OUT_FULL = 0, OUT_SPARSE = 1 };

static int grow_host(const uint8_t *image, int64_t nx, int64_t ny, int64_t nz, const int32_t *seeds, int nseeds,
                     int threshold, uint8_t *mask_out, uint8_t *image_inout_apply, int out_mode = OUT_FULL)
{
    if (!image || nx <= 0 || ny <= 0 || nz <= 0 || nseeds < 0 || (nseeds > 0 && !seeds)) return fail(RG_ERR_ARG, "bad argument");
    std::lock_guard<std::mutex> lk(g_mu);
    rg_ctx *c; int rc = default_ctx(&c); if (rc) return rc;
    ON_DEVICE(c->device);
    size_t bytes = (size_t)nx * ny * nz;
    rc = ensure_io(c, bytes, nseeds); if (rc) return rc;
    cudaStream_t st = 0;
    c->last_h2d_bytes = c->last_d2h_bytes = 0; c->last_copies = 0; c->last_staged = 0;
    rc = copy_in(c, c->d_image, image, bytes, st); if (rc) return rc;
    if (nseeds) CK(cudaMemcpyAsync(c->d_seeds, seeds, (size_t)nseeds * 12, cudaMemcpyHostToDevice, st));
    rc = grow_device_impl(c, c->d_image, nx, ny, nz, 0, 0, c->d_seeds, nseeds, threshold, c->d_mask, st);
    if (rc) return rc;
    if (mask_out) {
        rc = out_mode == OUT_SPARSE ? copy_out_sparse(c, mask_out, nx, ny, nz, st) : copy_out_full(c, mask_out, c->d_mask, bytes, st);
        if (rc) return rc;
    }
    if (image_inout_apply) {
        rc = rg_apply_mask_device(c, c->d_image, c->d_mask, (int64_t)bytes, st); if (rc) return rc;
        rc = copy_out_full(c, image_inout_apply, c->d_image, bytes, st); if (rc) return rc;
    }
    return rg_ctx_status(c, nullptr);
}

extern "C" int rg_grow(const uint8_t *image, int64_t nx, int64_t ny, int64_t nz,
                       const int32_t *seeds_xyz, int nseeds, int threshold, uint8_t *mask_out)
{
    if (!mask_out) return fail(RG_ERR_ARG, "mask_out == NULL");
    return grow_host(image, nx, ny, nz, seeds_xyz, nseeds, threshold, mask_out, nullptr);
}

extern "C" int rg_grow_sparse(const uint8_t *image, int64_t nx, int64_t ny, int64_t nz,
                              const int32_t *seeds_xyz, int nseeds, int threshold, uint8_t *mask_zeroed)
{
    if (!mask_zeroed) return fail(RG_ERR_ARG, "mask == NULL");
    return grow_host(image, nx, ny, nz, seeds_xyz, nseeds, threshold, mask_zeroed, nullptr, OUT_SPARSE);
}

extern "C" int rg_host_stats(int64_t out[4])
{
    if (!out) return fail(RG_ERR_ARG, "out == NULL");
    std::lock_guard<std::mutex> lk(g_mu);
    rg_ctx *c; int rc = default_ctx(&c); if (rc) return rc;
    out[0] = c->last_h2d_bytes; out[1] = c->last_d2h_bytes; out[2] = c->last_copies; out[3] = c->last_staged;
    return RG_OK;
}

static unsigned char *grow_512(unsigned char *data)
{
    /* raycast.cu:11 DATA_DIM 512; :286/:775 seed; :275 threshold */
    const int64_t D = 512;
    const int32_t seed[3] = {50, 300, 300};
    if (!data) { fail(RG_ERR_ARG, "data == NULL"); return nullptr; }
    unsigned char *region = (unsigned char *)calloc(sizeof(unsigned char), (size_t)(D * D * D));
    if (!region) { fail(RG_ERR_NOMEM, "calloc"); return nullptr; }
    /* the mask comes from calloc two lines up: only the boxes that hold region voxels are copied into it */
    if (grow_host(data, D, D, D, seed, 1, 1, region, nullptr, OUT_SPARSE) != RG_OK) { free(region); return nullptr; }
    return region;
}
extern "C" unsigned char *grow_region_gpu(unsigned char *data) { return grow_512(data); }
extern "C" unsigned char *grow_region_gpu_shared(unsigned char *data) { return grow_512(data); }
extern "C" unsigned char *grow_region_serial(unsigned char *data) { return grow_512(data); }

extern "C" int rg_grow_region_2d(unsigned char *image, int width, int height, unsigned char *mask_out, int apply)
{
    if (!image || width <= 0 || height <= 0) return fail(RG_ERR_ARG, "bad argument");
    /* region.c:450-490 in global 0-based pixels */
    const int32_t seeds[12] = { 4, 4, 0, width - 6, 4, 0, width - 6, height - 6, 0, 4, height - 6, 0 };
    return grow_host(image, width, height, 1, seeds, 4, 2 /* region.c:81 */, mask_out, apply ? image : nullptr);
}

/* ------------------------------------------------------------------ grow without any allocation
 *
 * Some hosts may not allocate device memory once they are running: an application that has handed the
 * whole card to its own memory pool at start-up and refuses later driver allocations, a service on a
 * latency-critical path that cannot afford cudaMalloc's implicit device-wide synchronisation or a late
 * out-of-memory failure, a plugin that must keep a fixed footprint.  For them this entry point runs the
 * same kernels as rg_grow_device, but every byte of engine scratch (tile records, face lists, coarse
 * union-find, work lists, seed keys) lives in ONE array that is part of the library image, so nothing is
 * allocated, freed or grown -- not here and not in any routine it calls.
 *
 * The price: the array has a fixed size (RG_NOALLOC_SCRATCH_BYTES, 64 MiB), it occupies that much memory on
 * every GPU that loads the library until the process exits, and every call on a device uses the same
 * array.  Two calls whose GPU work overlaps (different streams, different host threads) would overwrite
 * each other's scratch: run them one after the other.  The array keeps whatever the previous call left
 * in it; the first kernel of each call resets the counters and the "holds a seed" node, and everything
 * else is written before it is read, so no call depends on the previous one.
 */
#define RG_NOALLOC_SCRATCH_BYTES (64u << 20)
#define RG_NOALLOC_MAX_SEEDS 2048
static __device__ __align__(256) unsigned char g_noalloc_scratch[RG_NOALLOC_SCRATCH_BYTES];
/* `g_noalloc_scratch` names a device object; in host code it is only a handle.  The address the kernels
   need is looked up once per device, with that device current, and kept here (all of it under g_mu). */
static void *g_noalloc_addr[64];
static bool g_noalloc_attr[64];
static uint32_t *g_noalloc_counters[64];     /* where the last call on the device keeps its counters / status word */

extern "C" int64_t rg_noalloc_scratch_bytes(void) { return (int64_t)RG_NOALLOC_SCRATCH_BYTES; }

extern "C" int rg_grow_device_noalloc(const uint8_t *d_image, int64_t nx, int64_t ny, int64_t nz,
                                      const int32_t *d_seeds_xyz, int nseeds, int threshold, uint8_t *d_mask, void *stream)
{
    if (!d_image || !d_mask || nx <= 0 || ny <= 0 || nz <= 0 || nseeds < 0 || (nseeds > 0 && !d_seeds_xyz))
        return fail(RG_ERR_ARG, "bad argument");
    if (nseeds > RG_NOALLOC_MAX_SEEDS)
        return fail(RG_ERR_TOO_LARGE, "rg_grow_device_noalloc takes at most 2048 seeds (larger lists need the sorting scratch of rg_grow_device)");
    /* always the worst-case layout: RG_NODE_BUDGET (a tuning aid of the context path) must not squeeze a larger volume into
       the fixed array with average-case node space */
    Layout L; int rc = layout_for(nx, ny, nz, nseeds, L, false);
    if (rc) return rc;
    if (L.total > (size_t)RG_NOALLOC_SCRATCH_BYTES) {
        snprintf(g_err, sizeof g_err, "rg_grow_device_noalloc: this size needs %zu bytes of scratch but the built-in scratch holds %u "
                 "(64 MiB, e.g. up to 256x256x512 voxels); use rg_grow_device with a context instead", L.total, RG_NOALLOC_SCRATCH_BYTES);
        return RG_ERR_TOO_LARGE;
    }
    int dev = 0;
    CK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return fail(RG_ERR_ARG, "device index");
    rg_ctx view;                                            /* a view of the built-in scratch; owns nothing */
    memset(&view, 0, sizeof view);
    {
        std::lock_guard<std::mutex> lk(g_mu);
        if (!g_noalloc_addr[dev]) {                         /* first call on this device */
            void *p = nullptr;
            CK(cudaGetSymbolAddress(&p, g_noalloc_scratch));
            if (!p) return fail(RG_ERR_CUDA, "cudaGetSymbolAddress returned a null address");
            g_noalloc_addr[dev] = p;
        }
        view.ws = g_noalloc_addr[dev];
        view.attr_set = g_noalloc_attr[dev];
    }
    view.device = dev;
    view.ws_bytes = RG_NOALLOC_SCRATCH_BYTES;
    view.worst_case_layout = true;                          /* may_grow stays false: ensure_ws can only refuse */
    rc = grow_device_impl(&view, d_image, nx, ny, nz, 0, 0, d_seeds_xyz, nseeds, threshold, d_mask, stream);
    {
        std::lock_guard<std::mutex> lk(g_mu);
        g_noalloc_attr[dev] = g_noalloc_attr[dev] || view.attr_set;
        if (rc == RG_OK) g_noalloc_counters[dev] = view.last_counters;
    }
    return rc;
}

/* The engine's self-check of the last rg_grow_device_noalloc on the current device (the context path reads the same
   words with rg_ctx_status).  Waits for `stream`, copies 32 bytes into a local array -- no allocation here either. */
extern "C" int rg_noalloc_status(uint32_t stats[4], void *stream)
{
    int dev = 0;
    CK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return fail(RG_ERR_ARG, "device index");
    uint32_t *src;
    { std::lock_guard<std::mutex> lk(g_mu); src = g_noalloc_counters[dev]; }
    if (!src) { if (stats) memset(stats, 0, 16); return RG_OK; }
    uint32_t h[CTR_WORDS];
    CK(cudaMemcpyAsync(h, src, sizeof h, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    CK(cudaStreamSynchronize((cudaStream_t)stream));
    if (stats) { stats[0] = h[CTR_NODES]; stats[1] = h[CTR_STATUS]; stats[2] = h[CTR_NCOMPLEX] + h[CTR_NDENSE] + h[CTR_NROWTILES]; stats[3] = h[CTR_NFLAGGED] + h[CTR_NROWOUT]; }
    if (h[CTR_STATUS]) {
        snprintf(g_err, sizeof g_err, "engine self-check failed on the no-allocation path: status=0x%x mismatching faces=%u", h[CTR_STATUS], h[CTR_FACE_MISMATCH]);
        return RG_ERR_INTERNAL;
    }
    return RG_OK;
}

/* ------------------------------------------------------------------ one host image over several GPUs, from C
 *
 * rg_grow_multi is the whole of region.c's main() (region.c:582-604) for a C host with several GPUs in ONE process:
 * distribute_image + distribute_image_halo (region.c:106-246: every slab and its two boundary planes come straight from the
 * host image), grow per slab, one interface exchange in place of the exchange()/finished() rounds (region.c:250-353,
 * 435-440; lists move between the devices with cudaMemcpyPeerAsync), gather_region (region.c:391-432: every slab's mask
 * goes back into the caller's mask).  Volumes are cut along z, images (nz == 1) along y.  One host thread per slab drives
 * its device; the only host-side meeting point is the one where the interface list lengths become known.
 * A device may be named more than once (several slabs on one GPU) -- that is how the single-GPU tests exercise the exchange.
 */
struct MultiSlab {
    int device; rg_ctx *ctx; cudaStream_t st;
    uint8_t *buf; size_t buf_bytes; uint8_t *mask; size_t mask_bytes;
    int32_t *seeds; size_t seeds_cap;
    uint32_t *lists; size_t cap;                /* [2][cap] own interface lists */
    uint32_t *counts_d; uint32_t *counts_h;     /* {lower entries, upper entries, node id space, -} */
    uint32_t *gathered; size_t gathered_words; uint32_t *xpar; size_t xpar_words;
    int rc; char err[256];
};
static MultiSlab g_multi[RESOLVE_MAX_RANKS];
static std::mutex g_multi_mu;
static int64_t g_multi_last[4];                /* last rg_grow_multi: host -> device bytes, device -> host bytes, device -> host copies, slabs */

static int multi_fail(MultiSlab &m, int code) { m.rc = code; snprintf(m.err, sizeof m.err, "%s", g_err); return code; }
#define MCK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { fail(RG_ERR_CUDA, #call, e_); return multi_fail(m, RG_ERR_CUDA); } } while (0)

template <class T> static int multi_room(MultiSlab &m, T *&p, size_t &have, size_t want)
{
    if (have >= want) return RG_OK;
    if (p) { cudaStreamSynchronize(m.st); cudaFree(p); p = nullptr; have = 0; }
    MCK(cudaMalloc((void **)&p, want * sizeof(T)));
    have = want;
    return RG_OK;
}

extern "C" int rg_grow_multi(const uint8_t *image, int64_t nx, int64_t ny, int64_t nz, const int32_t *seeds_xyz, int nseeds,
                             int threshold, uint8_t *mask_out, const int *devices, int ndev, int flags, float *phase_ms)
{
    if (!image || !mask_out || nx <= 0 || ny <= 0 || nz <= 0 || nseeds < 0 || (nseeds > 0 && !seeds_xyz) || !devices || ndev < 1)
        return fail(RG_ERR_ARG, "bad argument");
    int have = 0;
    if (cudaGetDeviceCount(&have) != cudaSuccess || have == 0) return fail(RG_ERR_CUDA, "no CUDA device (this library has no CPU path)");
    for (int g = 0; g < ndev; g++) if (devices[g] < 0 || devices[g] >= have) return fail(RG_ERR_ARG, "bad device index");
    const bool rows = nz == 1;
    const int64_t na = rows ? ny : nz, plane = rows ? nx : nx * ny;
    int G = ndev > RESOLVE_MAX_RANKS ? RESOLVE_MAX_RANKS : ndev;
    while (G > 1 && na / G < 2) G--;                          /* every slab with a neighbour needs two planes (rows) */
    std::lock_guard<std::mutex> lk(g_multi_mu);
    int prev_dev = -1; cudaGetDevice(&prev_dev);
    const auto t_start = std::chrono::steady_clock::now();
    auto since = [&](std::chrono::steady_clock::time_point t) { return std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t).count(); };

    /* phase 1 per slab: scatter + label + export + counts */
    auto phase1 = [&](int g) -> int {
        MultiSlab &m = g_multi[g];
        m.rc = RG_OK; m.err[0] = 0;
        const int64_t z0 = na * g / G, z1 = na * (g + 1) / G, nzl = z1 - z0;
        const int lower = g > 0, upper = g < G - 1;
        if (m.ctx && m.device != devices[g]) {               /* the slot served another device before: start over */
            cudaSetDevice(m.device);
            cudaFree(m.buf); cudaFree(m.mask); cudaFree(m.seeds); cudaFree(m.lists); cudaFree(m.counts_d); cudaFree(m.gathered); cudaFree(m.xpar);
            cudaFreeHost(m.counts_h); cudaStreamDestroy(m.st); rg_ctx_destroy(m.ctx);
            memset(&m, 0, sizeof m);
        }
        MCK(cudaSetDevice(devices[g]));
        if (!m.ctx) {
            m.device = devices[g];
            if (rg_ctx_create(devices[g], &m.ctx) != RG_OK) return multi_fail(m, RG_ERR_CUDA);
            MCK(cudaStreamCreateWithFlags(&m.st, cudaStreamNonBlocking));
            MCK(cudaMalloc((void **)&m.counts_d, 16));
            MCK(cudaHostAlloc((void **)&m.counts_h, 16, cudaHostAllocDefault));
            for (int o = 0; o < G; o++)                       /* peer copies go over NVLink where the devices can reach each other */
                if (devices[o] != devices[g]) { int can = 0; cudaDeviceCanAccessPeer(&can, devices[g], devices[o]); if (can) cudaDeviceEnablePeerAccess(devices[o], 0); cudaGetLastError(); }
        }
        int rc;
        if ((rc = multi_room(m, m.buf, m.buf_bytes, (size_t)((nzl + 2) * plane)))) return rc;
        if ((rc = multi_room(m, m.mask, m.mask_bytes, (size_t)(nzl * plane)))) return rc;
        const size_t cap = G > 1 ? (size_t)rg_slab_interface_capacity(nx, rows ? nzl : ny, rows ? 1 : nzl) : 1;
        if (m.cap < cap) { size_t words = 2 * m.cap; if ((rc = multi_room(m, m.lists, words, 2 * cap))) return rc; m.cap = cap; }
        /* distribute_image + distribute_image_halo: own planes plus the neighbours' boundary planes, from the host image */
        const int64_t lo = z0 - lower, hi = z1 + upper;
        if (copy_in(m.ctx, m.buf + (size_t)((lo - z0 + 1) * plane), image + (size_t)(lo * plane), (size_t)((hi - lo) * plane), m.st)) return multi_fail(m, RG_ERR_CUDA);
        /* add_seeds (region.c:450-490): a seed belongs to the slab that owns its plane (row) */
        std::vector<int32_t> loc;
        for (int i = 0; i < nseeds; i++) {
            const int64_t a = seeds_xyz[3 * i + (rows ? 1 : 2)];
            if (a >= z0 && a < z1) { int32_t t[3] = { seeds_xyz[3 * i], seeds_xyz[3 * i + 1], seeds_xyz[3 * i + 2] }; t[rows ? 1 : 2] = (int32_t)(a - z0); loc.insert(loc.end(), t, t + 3); }
        }
        const int nloc = (int)loc.size() / 3;
        if ((rc = multi_room(m, m.seeds, m.seeds_cap, (size_t)(nloc > 0 ? 3 * nloc : 3)))) return rc;
        if (nloc) { MCK(cudaMemcpyAsync(m.seeds, loc.data(), loc.size() * 4, cudaMemcpyHostToDevice, m.st)); MCK(cudaStreamSynchronize(m.st)); }   /* loc dies with this scope */
        if (rg_slab_label(m.ctx, m.buf + plane, nx, rows ? nzl : ny, rows ? 1 : nzl, lower, upper, nloc ? m.seeds : nullptr, nloc, threshold, m.mask, m.st) != RG_OK)
            return multi_fail(m, RG_ERR_INTERNAL);
        MCK(cudaMemsetAsync(m.counts_d, 0, 16, m.st));
        if (G > 1) {
            if (lower && rg_slab_export(m.ctx, 0, m.lists, m.counts_d, m.st) != RG_OK) return multi_fail(m, RG_ERR_INTERNAL);
            if (upper && rg_slab_export(m.ctx, 1, m.lists + m.cap, m.counts_d, m.st) != RG_OK) return multi_fail(m, RG_ERR_INTERNAL);
        }
        MCK(cudaMemcpyAsync(m.counts_h, m.counts_d, 16, cudaMemcpyDeviceToHost, m.st));
        MCK(cudaStreamSynchronize(m.st));
        return RG_OK;
    };
    /* phase 2 per slab: gather every rank's lists, solve the interface union-find redundantly, output pass, gather_region */
    int64_t counts4[RESOLVE_MAX_RANKS * 4];
    uint32_t maxn = 0; uint64_t total_nodes = 0;
    auto phase2 = [&](int g) -> int {
        MultiSlab &m = g_multi[g];
        const int64_t z0 = na * g / G, z1 = na * (g + 1) / G, nzl = z1 - z0;
        MCK(cudaSetDevice(devices[g]));
        int rc;
        if (G > 1 && maxn > 0) {
            if ((rc = multi_room(m, m.gathered, m.gathered_words, (size_t)G * 2 * maxn))) return rc;
            if ((rc = multi_room(m, m.xpar, m.xpar_words, (size_t)(total_nodes + 2)))) return rc;
            for (int r = 0; r < G; r++)
                for (int side = 0; side < 2; side++) {
                    const size_t n = (size_t)counts4[4 * r + side];
                    if (!n) continue;
                    MCK(cudaMemcpyPeerAsync(m.gathered + ((size_t)r * 2 + side) * maxn, devices[g], g_multi[r].lists + (size_t)side * g_multi[r].cap, devices[r], n * 4, m.st));
                }
            if (rg_slab_resolve(m.ctx, m.xpar, m.gathered, maxn, counts4, g, G, m.st) != RG_OK) return multi_fail(m, RG_ERR_INTERNAL);
        }
        if (rg_slab_finish(m.ctx, m.st) != RG_OK) return multi_fail(m, RG_ERR_INTERNAL);
        uint8_t *dst = mask_out + (size_t)(z0 * plane);
        m.ctx->last_h2d_bytes = m.ctx->last_d2h_bytes = 0;
        if (flags & RG_MULTI_MASK_ZEROED) rc = copy_out_sparse(m.ctx, dst, nx, rows ? nzl : ny, rows ? 1 : nzl, m.st);
        else rc = copy_out_full(m.ctx, dst, m.mask, (size_t)(nzl * plane), m.st);
        if (rc) return multi_fail(m, rc);
        MCK(cudaStreamSynchronize(m.st));
        if (rg_ctx_status(m.ctx, nullptr) != RG_OK) return multi_fail(m, RG_ERR_INTERNAL);
        return RG_OK;
    };
    auto run_all = [&](const std::function<int(int)> &fn) {
        std::vector<std::thread> th;
        for (int g = 1; g < G; g++) th.emplace_back([&, g] { fn(g); });
        fn(0);
        for (auto &t : th) t.join();
        for (int g = 0; g < G; g++) if (g_multi[g].rc) { snprintf(g_err, sizeof g_err, "slab %d (device %d): %s", g, devices[g], g_multi[g].err); return g_multi[g].rc; }
        return (int)RG_OK;
    };
    int rc = run_all(phase1);
    const float t1 = since(t_start);
    if (rc == RG_OK) {
        for (int g = 0; g < G; g++) {
            const uint32_t *h = g_multi[g].counts_h;
            counts4[4 * g] = h[0]; counts4[4 * g + 1] = h[1]; counts4[4 * g + 2] = h[2]; counts4[4 * g + 3] = 0;
            if (h[0] > maxn) maxn = h[0];
            if (h[1] > maxn) maxn = h[1];
            total_nodes += h[2];
        }
        for (int g = 0; g + 1 < G; g++)
            if (counts4[4 * g + 1] != counts4[4 * (g + 1)]) rc = fail(RG_ERR_INTERNAL, "interface lists of neighbouring slabs differ in length");
        if (total_nodes + 2 >= (1ull << 32)) rc = fail(RG_ERR_TOO_LARGE, "too many nodes across slabs for 32-bit vertex ids");
    }
    if (rc == RG_OK) rc = run_all(phase2);
    if (phase_ms) { phase_ms[0] = since(t_start); phase_ms[1] = t1; phase_ms[2] = phase_ms[0] - t1; phase_ms[3] = (float)G; }
    g_multi_last[0] = g_multi_last[1] = g_multi_last[2] = 0; g_multi_last[3] = G;
    for (int i = 0; i < nseeds && rc == RG_OK; i++) {                                   /* a seed goes to the slab that owns its plane (row) */
        const int64_t a = seeds_xyz[3 * i + (rows ? 1 : 2)];
        if (a >= 0 && a < na) g_multi_last[0] += 12;
    }
    for (int g = 0; g < G && rc == RG_OK; g++) {
        const int64_t z0 = na * g / G, z1 = na * (g + 1) / G;
        const int64_t lo = z0 > 0 ? 1 : 0, hi = z1 < na ? 1 : 0;                        /* the slab and its neighbours' boundary planes */
        g_multi_last[0] += (z1 - z0 + lo + hi) * plane;
        g_multi_last[1] += g_multi[g].ctx->last_d2h_bytes; g_multi_last[2] += g_multi[g].ctx->last_copies;
    }
    if (prev_dev >= 0) cudaSetDevice(prev_dev);
    return rc;
}

extern "C" int rg_multi_stats(int64_t out[4])
{
    if (!out) return fail(RG_ERR_ARG, "out == NULL");
    std::lock_guard<std::mutex> lk(g_multi_mu);
    for (int i = 0; i < 4; i++) out[i] = g_multi_last[i];
    return RG_OK;
}

/* ------------------------------------------------------------------ one host IMAGE over a rows x columns grid of GPUs
 *
 * The reference spreads its picture over a Cartesian grid of MPI ranks (region.c:537-548 MPI_Dims_create / Cart_create;
 * distribute_image :106-143; exchange of the four borders :250-353; gather_region :391-432).  rg_grow_multi_grid does
 * that from one C call: block (i, j) = rows [h*i/R, h*(i+1)/R) x columns [x_j, x_(j+1)) with the column cuts on multiples
 * of 128 pixels, one block per entry of devices[] (row-major; a device may be named several times).  No halo is exchanged:
 * a block labels itself with its seams "open" (every pixel of a seam is a cross-tile edge, Params::open), exports for each
 * seam the local root and the value of every border pixel, and the exchange unites two roots where the two pixels facing
 * each other across the seam are similar -- the comparison the reference's ranks make on their halo copies.  Every block
 * resolves the whole seam graph redundantly (as the slabs do) and writes its part of the mask.
 */
struct GridBlock {
    int device; rg_ctx *ctx; cudaStream_t st;
    uint8_t *img, *mask; size_t img_bytes, mask_bytes;
    int32_t *seeds; size_t seeds_cap;
    uint32_t *roots; uint8_t *vals; size_t list_cap;         /* own four border lists: [4][list_cap] */
    uint32_t *counts_h;                                     /* pinned: {node id space} */
    uint32_t *all_roots; uint8_t *all_vals; size_t all_cap;  /* every block's lists, gathered */
    uint32_t *xpar; size_t xpar_words;
    int64_t x0, x1, y0, y1;
    int rc; char err[256];
};
static GridBlock g_grid[RESOLVE_MAX_RANKS];

extern "C" int rg_grow_multi_grid(const uint8_t *image, int64_t w, int64_t h, const int32_t *seeds_xyz, int nseeds, int threshold,
                                  uint8_t *mask_out, const int *devices, int rows, int cols, int flags, float *phase_ms)
{
    if (!image || !mask_out || w <= 0 || h <= 0 || nseeds < 0 || (nseeds > 0 && !seeds_xyz) || !devices || rows < 1 || cols < 1 || rows * cols > RESOLVE_MAX_RANKS)
        return fail(RG_ERR_ARG, "bad argument");
    int have = 0;
    if (cudaGetDeviceCount(&have) != cudaSuccess || have == 0) return fail(RG_ERR_CUDA, "no CUDA device (this library has no CPU path)");
    for (int g = 0; g < rows * cols; g++) if (devices[g] < 0 || devices[g] >= have) return fail(RG_ERR_ARG, "bad device index");
    (void)flags;
    /* cuts: rows anywhere (at least one row per block), columns on multiples of the tile width */
    const int64_t TXW = Cfg2D::TX;
    int R = rows, Cn = cols;
    while (R > 1 && h / R < 1) R--;
    while (Cn > 1 && (w / Cn) / TXW < 1) Cn--;
    std::vector<int64_t> xc((size_t)Cn + 1), yc((size_t)R + 1);
    for (int j = 0; j <= Cn; j++) xc[(size_t)j] = j == Cn ? w : (w * j / Cn) / TXW * TXW;
    for (int i = 0; i <= R; i++) yc[(size_t)i] = h * i / R;
    const int B = R * Cn;
    std::lock_guard<std::mutex> lk(g_multi_mu);
    int prev_dev = -1; cudaGetDevice(&prev_dev);
    const auto t_start = std::chrono::steady_clock::now();
    auto since = [&](std::chrono::steady_clock::time_point t) { return std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t).count(); };
    size_t list_cap = 0;
    for (int j = 0; j < Cn; j++) list_cap = std::max(list_cap, (size_t)(xc[(size_t)j + 1] - xc[(size_t)j]));
    for (int i = 0; i < R; i++) list_cap = std::max(list_cap, (size_t)(yc[(size_t)i + 1] - yc[(size_t)i]));
    auto dev_of = [&](int b) { return devices[(b / Cn) * cols + (b % Cn)]; };       /* the caller's grid is rows x cols; we may use fewer */

    auto phase1 = [&](int b) -> int {
        GridBlock &m = g_grid[b];
        m.rc = RG_OK; m.err[0] = 0;
        const int i = b / Cn, j = b % Cn;
        m.x0 = xc[(size_t)j]; m.x1 = xc[(size_t)j + 1]; m.y0 = yc[(size_t)i]; m.y1 = yc[(size_t)i + 1];
        const int64_t bw = m.x1 - m.x0, bh = m.y1 - m.y0;
        auto bad = [&](int code) { m.rc = code; snprintf(m.err, sizeof m.err, "%s", g_err); return code; };
#define GCK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { fail(RG_ERR_CUDA, #call, e_); return bad(RG_ERR_CUDA); } } while (0)
        if (m.ctx && m.device != dev_of(b)) {
            cudaSetDevice(m.device);
            cudaFree(m.img); cudaFree(m.mask); cudaFree(m.seeds); cudaFree(m.roots); cudaFree(m.vals); cudaFree(m.all_roots); cudaFree(m.all_vals); cudaFree(m.xpar);
            cudaFreeHost(m.counts_h); cudaStreamDestroy(m.st); rg_ctx_destroy(m.ctx);
            memset(&m, 0, sizeof m);
            m.x0 = xc[(size_t)j]; m.x1 = xc[(size_t)j + 1]; m.y0 = yc[(size_t)i]; m.y1 = yc[(size_t)i + 1];
        }
        GCK(cudaSetDevice(dev_of(b)));
        if (!m.ctx) {
            m.device = dev_of(b);
            if (rg_ctx_create(m.device, &m.ctx) != RG_OK) return bad(RG_ERR_CUDA);
            GCK(cudaStreamCreateWithFlags(&m.st, cudaStreamNonBlocking));
            GCK(cudaHostAlloc((void **)&m.counts_h, 16, cudaHostAllocDefault));
            for (int o = 0; o < B; o++)
                if (dev_of(o) != m.device) { int can = 0; cudaDeviceCanAccessPeer(&can, m.device, dev_of(o)); if (can) cudaDeviceEnablePeerAccess(dev_of(o), 0); cudaGetLastError(); }
        }
        auto room = [&](void **p, size_t &have_b, size_t want) -> int {
            if (have_b >= want) return RG_OK;
            if (*p) { cudaStreamSynchronize(m.st); cudaFree(*p); *p = nullptr; have_b = 0; }
            GCK(cudaMalloc(p, want)); have_b = want; return RG_OK;
        };
        int rc;
        if ((rc = room((void **)&m.img, m.img_bytes, (size_t)(bw * bh)))) return rc;
        if ((rc = room((void **)&m.mask, m.mask_bytes, (size_t)(bw * bh)))) return rc;
        if (m.list_cap < list_cap) {
            size_t a = m.list_cap * 16, c = m.list_cap * 4;
            if ((rc = room((void **)&m.roots, a, list_cap * 16))) return rc;
            if ((rc = room((void **)&m.vals, c, list_cap * 4))) return rc;
            m.list_cap = list_cap;
        }
        /* distribute_image (region.c:106-143): the block's rows out of the picture */
        GCK(cudaMemcpy2DAsync(m.img, (size_t)bw, image + m.y0 * w + m.x0, (size_t)w, (size_t)bw, (size_t)bh, cudaMemcpyHostToDevice, m.st));
        std::vector<int32_t> loc;
        for (int k = 0; k < nseeds; k++) {
            const int64_t sx = seeds_xyz[3 * k], sy = seeds_xyz[3 * k + 1];
            if (sx >= m.x0 && sx < m.x1 && sy >= m.y0 && sy < m.y1 && seeds_xyz[3 * k + 2] == 0) { loc.push_back((int32_t)(sx - m.x0)); loc.push_back((int32_t)(sy - m.y0)); loc.push_back(0); }
        }
        const int nloc = (int)loc.size() / 3;
        { size_t sb = m.seeds_cap * 4; if ((rc = room((void **)&m.seeds, sb, (size_t)(nloc > 0 ? 12 * nloc : 12)))) return rc; m.seeds_cap = sb / 4; }
        if (nloc) { GCK(cudaMemcpyAsync(m.seeds, loc.data(), loc.size() * 4, cudaMemcpyHostToDevice, m.st)); GCK(cudaStreamSynchronize(m.st)); }
        const int open = (j > 0 ? 1 : 0) | (j < Cn - 1 ? 2 : 0) | (i > 0 ? 4 : 0) | (i < R - 1 ? 8 : 0);
        if (rg_block_label(m.ctx, m.img, bw, bh, open, nloc ? m.seeds : nullptr, nloc, threshold, m.mask, m.st) != RG_OK) return bad(RG_ERR_INTERNAL);
        for (int side = 0; side < 4; side++)
            if ((open >> side) & 1)
                if (rg_block_export(m.ctx, side, m.roots + (size_t)side * m.list_cap, m.vals + (size_t)side * m.list_cap, m.st) != RG_OK) return bad(RG_ERR_INTERNAL);
        GCK(cudaMemcpyAsync(m.counts_h, m.ctx->last_counters + CTR_NODES, 4, cudaMemcpyDeviceToHost, m.st));
        GCK(cudaStreamSynchronize(m.st));
        return RG_OK;
    };
    std::vector<uint32_t> off((size_t)B + 1, 1u);
    uint64_t total = 1;
    auto phase2 = [&](int b) -> int {
        GridBlock &m = g_grid[b];
        const int64_t bw = m.x1 - m.x0, bh = m.y1 - m.y0;
        auto bad = [&](int code) { m.rc = code; snprintf(m.err, sizeof m.err, "%s", g_err); return code; };
        GCK(cudaSetDevice(m.device));
        if (B > 1) {
            auto room = [&](void **p, size_t &have_b, size_t want) -> int {
                if (have_b >= want) return RG_OK;
                if (*p) { cudaStreamSynchronize(m.st); cudaFree(*p); *p = nullptr; have_b = 0; }
                GCK(cudaMalloc(p, want)); have_b = want; return RG_OK;
            };
            int rc;
            if (m.all_cap < (size_t)B * 4 * list_cap) {
                size_t a = m.all_cap * 4, c = m.all_cap;
                if ((rc = room((void **)&m.all_roots, a, (size_t)B * 16 * list_cap))) return rc;
                if ((rc = room((void **)&m.all_vals, c, (size_t)B * 4 * list_cap))) return rc;
                m.all_cap = (size_t)B * 4 * list_cap;
            }
            { size_t xb = m.xpar_words * 4; if ((rc = room((void **)&m.xpar, xb, (size_t)(total + 2) * 4))) return rc; m.xpar_words = xb / 4; }
            auto len = [&](int o, int side) { const GridBlock &q = g_grid[o]; return (size_t)(side < 2 ? q.y1 - q.y0 : q.x1 - q.x0); };
            auto has = [&](int o, int side) { const int oi = o / Cn, oj = o % Cn; return side == 0 ? oj > 0 : side == 1 ? oj < Cn - 1 : side == 2 ? oi > 0 : oi < R - 1; };
            auto R_ = [&](int o, int side) { return m.all_roots + ((size_t)o * 4 + side) * list_cap; };
            auto V_ = [&](int o, int side) { return m.all_vals + ((size_t)o * 4 + side) * list_cap; };
            for (int o = 0; o < B; o++)
                for (int side = 0; side < 4; side++) {
                    if (!has(o, side)) continue;
                    GCK(cudaMemcpyPeerAsync(R_(o, side), m.device, g_grid[o].roots + (size_t)side * g_grid[o].list_cap, g_grid[o].device, len(o, side) * 4, m.st));
                    GCK(cudaMemcpyPeerAsync(V_(o, side), m.device, g_grid[o].vals + (size_t)side * g_grid[o].list_cap, g_grid[o].device, len(o, side), m.st));
                }
            GCK(cudaMemsetAsync(m.xpar, 0, 4, m.st));
            for (int o = 0; o < B; o++)
                for (int side = 0; side < 4; side++)
                    if (has(o, side) && rg_slab_resolve_init(m.ctx, m.xpar, R_(o, side), (uint32_t)len(o, side), off[(size_t)o], m.st) != RG_OK) return bad(RG_ERR_INTERNAL);
            for (int o = 0; o < B; o++) {
                const int oi = o / Cn, oj = o % Cn;
                if (oj < Cn - 1 && rg_block_resolve_union(m.ctx, m.xpar, R_(o, 1), off[(size_t)o], V_(o, 1), R_(o + 1, 0), off[(size_t)o + 1], V_(o + 1, 0), (uint32_t)len(o, 1), threshold, m.st) != RG_OK) return bad(RG_ERR_INTERNAL);
                if (oi < R - 1 && rg_block_resolve_union(m.ctx, m.xpar, R_(o, 3), off[(size_t)o], V_(o, 3), R_(o + Cn, 2), off[(size_t)o + Cn], V_(o + Cn, 2), (uint32_t)len(o, 3), threshold, m.st) != RG_OK) return bad(RG_ERR_INTERNAL);
            }
            for (int side = 0; side < 4; side++)
                if (has(b, side) && rg_slab_resolve_apply(m.ctx, m.xpar, R_(b, side), (uint32_t)len(b, side), off[(size_t)b], m.st) != RG_OK) return bad(RG_ERR_INTERNAL);
        }
        if (rg_slab_finish(m.ctx, m.st) != RG_OK) return bad(RG_ERR_INTERNAL);
        /* gather_region (region.c:391-432) */
        GCK(cudaMemcpy2DAsync(mask_out + m.y0 * w + m.x0, (size_t)w, m.mask, (size_t)bw, (size_t)bw, (size_t)bh, cudaMemcpyDeviceToHost, m.st));
        GCK(cudaStreamSynchronize(m.st));
        if (rg_ctx_status(m.ctx, nullptr) != RG_OK) return bad(RG_ERR_INTERNAL);
        return RG_OK;
#undef GCK
    };
    auto run_all = [&](const std::function<int(int)> &fn) {
        std::vector<std::thread> th;
        for (int b = 1; b < B; b++) th.emplace_back([&, b] { fn(b); });
        fn(0);
        for (auto &t : th) t.join();
        for (int b = 0; b < B; b++) if (g_grid[b].rc) { snprintf(g_err, sizeof g_err, "block %d (device %d): %s", b, g_grid[b].device, g_grid[b].err); return g_grid[b].rc; }
        return (int)RG_OK;
    };
    int rc = run_all(phase1);
    const float t1 = since(t_start);
    if (rc == RG_OK) {
        for (int b = 0; b < B; b++) { off[(size_t)b] = (uint32_t)total; total += g_grid[b].counts_h[0]; }
        if (total + 2 >= (1ull << 32)) rc = fail(RG_ERR_TOO_LARGE, "too many nodes across blocks for 32-bit vertex ids");
    }
    if (rc == RG_OK) rc = run_all(phase2);
    if (phase_ms) { phase_ms[0] = since(t_start); phase_ms[1] = t1; phase_ms[2] = phase_ms[0] - t1; phase_ms[3] = (float)(R * 1000 + Cn); }
    if (prev_dev >= 0) cudaSetDevice(prev_dev);
    return rc;
}
