/*
 * include/rg.h -- C ABI of the B200 seeded region-growing engine (libb200rg.so).
 *
 * Plain C: pointers, sizes, ints.  No torch / C++ types.  Every entry point cites the
 * reference interface it replaces (paths relative to the reference repository
 * anonyomous4/parallel-c-programs).  Semantics frozen from the reference:
 *   region = union of the connected components, under "grid neighbours with
 *   abs((int)a-(int)b) < threshold", that contain a seed;  4-connectivity when nz == 1
 *   (region.c:506), 6-connectivity otherwise (raycast.cu:290-292);  x fastest, then y, then z
 *   (raycast.cu:178-180);  seeds are (x,y,z) triples, out-of-range seeds are ignored
 *   (region.c:485), duplicates are harmless;  mask bytes are exactly 0 or 1.
 *
 * There is no CPU fallback: every call fails with RG_ERR_CUDA when no CUDA device is usable.
 */
#ifndef B200RG_RG_H
#define B200RG_RG_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum {
    RG_OK = 0,
    RG_ERR_ARG = -1,        /* NULL pointer, non-positive size, nseeds < 0 ... */
    RG_ERR_CUDA = -2,       /* CUDA runtime error (rg_last_error() has the text) */
    RG_ERR_NOMEM = -3,      /* host or device allocation failed */
    RG_ERR_INTERNAL = -4,   /* engine self-check failed (status word non-zero) */
    RG_ERR_TOO_LARGE = -5   /* volume exceeds what this entry point supports */
};

typedef struct rg_ctx rg_ctx;

/* ---- reference-shaped entry points -------------------------------------------------------- */

/* replaces grow_region_gpu()        5-cuda-region-growing/raycast.cu:702-756
 *          grow_region_gpu_shared() 5-cuda-region-growing/raycast.cu:759-822  (what main() runs, :833)
 *          grow_region_serial()     5-cuda-region-growing/raycast.cu:281-318  (same mask, computed on the GPU)
 * and the OpenCL twin grow_region_gpu() 6-opencl-region-growing/raycast.c:308-377.
 * data = host 512^3 uint8, x fastest.  Seed (50,300,300), threshold 1, 6-connectivity
 * (raycast.cu:275,286,775).  Returns a calloc-compatible 512^3 mask the caller free()s
 * (raycast.cu:852), NULL on failure.  Unlike the reference kernels the mask never contains 2s. */
unsigned char *grow_region_gpu(unsigned char *data);
unsigned char *grow_region_gpu_shared(unsigned char *data);
unsigned char *grow_region_serial(unsigned char *data);

/* replaces raycast_gpu()         5-cuda-region-growing/raycast.cu:436-470 (kernel :321-371)
 *          raycast_gpu_texture() 5-cuda-region-growing/raycast.cu:472-531 (kernel :374-433; what main() runs, :842)
 *          raycast_serial()      5-cuda-region-growing/raycast.cu:216-267 (same picture, computed on the GPU)
 * and the OpenCL twin raycast_gpu() 6-opencl-region-growing/raycast.c:379-436 / raycast.cl:93-137.
 * data, region = host 512^3 uint8 (region = the grower's mask).  Camera (1000,1000,1000) looking along
 * (-1,-1,-1), 512 x 512 pixels, as the #defines of raycast.cu:14-17 say.  Returns a malloc'd 512 x 512 image
 * the caller free()s (raycast.cu:853), NULL on failure. */
unsigned char *raycast_gpu(unsigned char *data, unsigned char *region);
unsigned char *raycast_gpu_texture(unsigned char *data, unsigned char *region);

/* replaces the grow_region() + write_image() pair of 2-mpi-region-growing/region.c:493-533,572-580
 * for one whole image: seeds = the four corner seeds of add_seeds() (region.c:450-490) expressed
 * in global 0-based pixels (4,4),(w-6,4),(w-6,h-6),(4,h-6); threshold 2 (region.c:81).
 * mask_out (w*h, may be NULL) receives the region; if apply != 0, image[i] *= (mask[i]==0) is
 * applied in place on the device before the copy back (region.c:576). */
int rg_grow_region_2d(unsigned char *image, int width, int height, unsigned char *mask_out, int apply);

/* ---- parametrised entry points ------------------------------------------------------------- */

/* Host buffers in, host mask out (H2D + grow + D2H on device 0 / the current device).
 * Generalises the three growers above to any size, seed list and threshold. */
int rg_grow(const uint8_t *image, int64_t nx, int64_t ny, int64_t nz,
            const int32_t *seeds_xyz, int nseeds, int threshold, uint8_t *mask_out);

/* Like rg_grow for a caller whose mask buffer is ALL ZEROS on entry -- fresh from calloc (the reference's convention:
 * its growers return a calloc'd region, raycast.cu:705,762), or cleared by the caller.  Only the boxes of the mask that hold
 * region voxels are copied back (whole tiles; C1's region is 1.6 % of its volume), the rest of the buffer is not touched.
 * The reference-shaped growers above use this path with their own calloc.  No counterpart in the reference. */
int rg_grow_sparse(const uint8_t *image, int64_t nx, int64_t ny, int64_t nz,
                   const int32_t *seeds_xyz, int nseeds, int threshold, uint8_t *mask_zeroed);

/* What the last host-pointer call on the current device moved: {bytes host->device, bytes device->host, device->host copies
 * issued, 1 if a pageable buffer was staged through the library's pinned bounce buffers by its copy threads}.
 * Pageable (malloc'd) buffers of 32 MiB and more are staged that way; pinned buffers go to the copy engine directly.
 * No counterpart in the reference (bench.py reports these as h2d/d2h bytes per step). */
int rg_host_stats(int64_t out[4]);

/* Context = per-device workspace (scratch for tile records, face lists, coarse union-find). */
int rg_ctx_create(int device, rg_ctx **out);
void rg_ctx_destroy(rg_ctx *ctx);

/* Bytes of device scratch a grow of this size needs: the worst case, so that no call can overflow, up to ~17 Gvoxel;
 * beyond that the node space is sized for an average of 2048 nodes per tile (the worst case would not fit 32-bit node
 * ids) and an input that needs more is refused through the status word (rg_ctx_status fails with "out of node ids"). */
int64_t rg_workspace_bytes(int64_t nx, int64_t ny, int64_t nz, int nseeds);

/* Pre-size the context's scratch so that later calls do not allocate. */
int rg_ctx_reserve(rg_ctx *ctx, int64_t nx, int64_t ny, int64_t nz, int nseeds);

/* Device-resident grow: d_image, d_seeds_xyz and d_mask are DEVICE pointers on the context's
 * device; everything is enqueued on `stream` (a cudaStream_t, NULL = default stream) and the
 * call returns without synchronising.  This is the timed region of bench.py ("value").
 * The device analogue of the do{kernel}while(flag) loops at raycast.cu:737-745,804-811 -- with
 * no host round trip: convergence is structural (one union-find pass), not iterative. */
int rg_grow_device(rg_ctx *ctx, const uint8_t *d_image, int64_t nx, int64_t ny, int64_t nz,
                   const int32_t *d_seeds_xyz, int nseeds, int threshold, uint8_t *d_mask, void *stream);

/* No counterpart in the reference (its growers cudaMalloc and cudaFree on every call, raycast.cu:711-713,
 * 751-753).  Same computation and same arguments as rg_grow_device, for callers that must not allocate
 * device memory while they run -- a host that gave the whole card to its own pool at start-up, a
 * latency-critical loop that cannot take cudaMalloc's device-wide synchronisation or a late out-of-memory
 * error, a plugin with a fixed footprint.  No context: all engine scratch is one 64 MiB array built into the
 * library (rg_noalloc_scratch_bytes()), resident on every GPU that loads the library until the process
 * exits.  The call allocates and frees nothing, on the host or the device.
 *   - Limits: rg_workspace_bytes(nx, ny, nz, nseeds) <= 64 MiB (for instance 256 x 256 x 512 voxels, or a
 *     8192 x 8192 image) and nseeds <= 2048; anything larger returns RG_ERR_TOO_LARGE with a message.
 *   - The first call on each device looks up the array's device address (cudaGetSymbolAddress) and sets
 *     the kernels' shared-memory attributes; make that first call before any stream capture begins.
 *   - All calls on a device share the array.  Calls whose GPU work can overlap (two streams, two host
 *     threads) corrupt each other: issue them back to back on one stream, or serialise them yourself.
 *     Nothing carries over between calls: each call's first kernel resets the counters it relies on.
 * d_image, d_seeds_xyz, d_mask are device pointers on the current device; work is enqueued on `stream`. */
int rg_grow_device_noalloc(const uint8_t *d_image, int64_t nx, int64_t ny, int64_t nz,
                           const int32_t *d_seeds_xyz, int nseeds, int threshold, uint8_t *d_mask, void *stream);
int64_t rg_noalloc_scratch_bytes(void);
/* No counterpart in the reference.  The engine's self-check for the last rg_grow_device_noalloc on the current device:
 * waits for `stream` (pass the one the grow ran on), then returns RG_OK or RG_ERR_INTERNAL like rg_ctx_status and
 * fills stats (may be NULL) the same way.  Allocates nothing.  The scratch of that entry point is always sized for the
 * worst case (it ignores RG_NODE_BUDGET), so "out of node ids" cannot happen there; what this reports is the
 * face-list cross-check of the labelling passes. */
int rg_noalloc_status(uint32_t stats[4], void *stream);

/* Synchronises the stream last used and returns RG_OK or RG_ERR_INTERNAL / RG_ERR_CUDA;
 * stats (may be NULL) receives {coarse nodes, status bits, tiles the bounded walkers did not get through (labelled by the
 * dense path or the full pass), tiles holding region voxels}. */
int rg_ctx_status(rg_ctx *ctx, uint32_t stats[4]);

/* Optional per-kernel timing with CUDA events on the launching stream (used by bench.py for the
 * roofline of the dominant kernel).  ms = {seed sort, fast labelling pass, full labelling pass, coarse
 * merge, flag / zero-fill, pass C} of the last grow (multi-GPU: the interface exchange falls into `flag`).
 * Stands in for the gettimeofday()/print_time() pair around the grow call, raycast.cu:23-27,832-836. */
int rg_ctx_profile(rg_ctx *ctx, int enable);
int rg_ctx_profile_read(rg_ctx *ctx, float ms[6]);

/* replaces write_image()'s loop, region.c:575-577, on the device: image[i] *= (mask[i] == 0). */
int rg_apply_mask_device(rg_ctx *ctx, uint8_t *d_image, const uint8_t *d_mask, int64_t n, void *stream);

/* The ray-caster on device-resident buffers, any cube side and image side (generalises raycast.cu:14-17):
 * d_data, d_region = data_dim^3 bytes, d_image = image_dim^2 bytes, camera = the coordinate the reference
 * puts on all three axes (1000).  Asynchronous on `stream`. */
int rg_raycast_device(const uint8_t *d_data, const uint8_t *d_region, int data_dim, int image_dim, float camera,
                      uint8_t *d_image, void *stream);

const char *rg_last_error(void);

/* ---- one host image over several GPUs, one process -------------------------------------------- */
/* Replaces the whole of region.c's main() sequence distribute_image -> distribute_image_halo -> grow_region (with its
 * exchange()/finished() rounds) -> gather_region (region.c:106-246, 250-353, 391-440, 591-597) for a C host that owns
 * several GPUs: the volume is cut into one z-slab per entry of `devices` (an image, nz == 1, into row slabs), every slab
 * and its neighbours' boundary planes are copied from `image`, each device labels its slab, the interface lists are moved
 * between the devices with peer copies and resolved in ONE exchange, and every slab's mask lands in `mask_out`.
 * Same result as rg_grow, whatever the number of devices.  devices[] may name a GPU more than once (several slabs on one
 * GPU).  flags: RG_MULTI_MASK_ZEROED = mask_out is all zeros on entry (calloc), copy back only the boxes that hold region
 * voxels.  phase_ms (may be NULL) receives {whole call, scatter + label + export, exchange + output + gather, slabs used}
 * in host milliseconds.  The first call allocates per-device buffers and contexts (kept for later calls of the same or a
 * smaller size): that cold call is what a one-shot program pays. */
#define RG_MULTI_MASK_ZEROED 1
int rg_grow_multi(const uint8_t *image, int64_t nx, int64_t ny, int64_t nz, const int32_t *seeds_xyz, int nseeds,
                  int threshold, uint8_t *mask_out, const int *devices, int ndev, int flags, float *phase_ms);
/* What the last rg_grow_multi of this process moved over PCIe: {host -> device bytes (slabs, boundary planes, seeds), device -> host
 * bytes, device -> host copies, slabs used} -- the counterpart of rg_host_stats for the multi-device call (region.c:106-246 scatter,
 * :391-432 gather). */
int rg_multi_stats(int64_t out[4]);

/* The same for an IMAGE over a rows x cols grid of blocks, the reference's Cartesian process grid (region.c:537-548: MPI_Dims_create /
 * MPI_Cart_create; distribute_image :106-143; the four-border exchange() :250-353; gather_region :391-432): block (i, j) =
 * rows [h*i/rows, h*(i+1)/rows) x columns [x_j, x_(j+1)), column cuts on multiples of 128 pixels, devices[i * cols + j] its GPU
 * (a GPU may be named several times).  No halo travels: each block labels itself with its seams open (every seam pixel becomes an
 * interface entry), and the exchange unites two components where the two pixels facing each other across a seam are similar.
 * Same result as rg_grow.  An image too narrow for `cols` blocks of at least 128 columns gets fewer; phase_ms[3] = 1000 * rows used +
 * columns used.  flags: reserved (0). */
int rg_grow_multi_grid(const uint8_t *image, int64_t width, int64_t height, const int32_t *seeds_xyz, int nseeds, int threshold,
                       uint8_t *mask_out, const int *devices, int rows, int cols, int flags, float *phase_ms);

/* The phases rg_grow_multi_grid is built from, for a host with one process per GPU (MPI ranks on a Cartesian grid, as region.c):
 * rg_block_label labels a device-resident block whose borders in `open_sides` (bit 0 / 1 / 2 / 3 = -x / +x / -y / +y) are seams
 * to other blocks -- a block with a +x seam must be a multiple of 128 pixels wide --; rg_block_export lists, for every pixel of
 * one such border in order, the local root of its component and its value; after moving the lists (MPI / NCCL / peer copies) every
 * rank runs rg_slab_resolve_init on each list with its block's vertex offset, rg_block_resolve_union on each seam (two lists that
 * face each other: entry i against entry i, united where |value_a - value_b| < threshold), rg_slab_resolve_apply on its own lists,
 * then rg_slab_finish.  Replaces exchange() / add_halo_to_stack() / finished() of region.c:250-387, 435-440 for a 2-D grid. */
int rg_block_label(rg_ctx *ctx, const uint8_t *d_block, int64_t width, int64_t height, int open_sides, const int32_t *d_seeds_xyz, int nseeds,
                   int threshold, uint8_t *d_mask_block, void *stream);
int rg_block_export(rg_ctx *ctx, int side, uint32_t *d_roots, uint8_t *d_values, void *stream);
int rg_block_resolve_union(rg_ctx *ctx, uint32_t *d_xpar, const uint32_t *d_roots_a, uint32_t off_a, const uint8_t *d_values_a,
                           const uint32_t *d_roots_b, uint32_t off_b, const uint8_t *d_values_b, uint32_t n, int threshold, void *stream);

/* ---- z-slab (multi-GPU) phases ------------------------------------------------------------- */
/* Together these replace the distributed half of 2-mpi-region-growing/region.c -- exchange()
 * (:250-353), add_halo_to_stack() (:355-387), finished() (:435-440) and the static image-halo
 * scatter distribute_image_halo() (:145-246) -- for one rank that owns planes [z0, z0+nz_local)
 * of the volume.  One process per GPU; the host layer (rgb200.SlabGrower, torch.distributed over
 * NCCL) moves the interface lists between ranks.  Sequence per rank:
 *   rg_slab_label -> rg_slab_export(side 0/1) -> [all-gather counts, lists] ->
 *   rg_slab_resolve_init (every list) -> rg_slab_resolve_union (every interface) ->
 *   rg_slab_resolve_apply (own two lists) -> rg_slab_finish.
 * One exchange, independent of how often a region crosses slab boundaries (the reference needs a
 * round per crossing). */

/* d_slab points at the slab's first own plane; if has_lower / has_upper the image planes just
 * below / above it (the neighbour's boundary planes, cf. distribute_image_halo) must be readable
 * at d_slab - nx*ny and d_slab + nz_local*nx*ny.  Seeds are slab-local (z - z0) and only those
 * inside the slab.  Enqueues seed sort + pass A + pass B.
 * An IMAGE is cut into row slabs the same way (2-D inputs of region.c, whose MPI ranks each own a block of the picture,
 * region.c:106-143): nz_local = 1, ny = the rank's own rows, seeds slab-local (y - y0), and the neighbours' boundary
 * ROWS must be readable at d_slab - nx and d_slab + ny*nx.  A rank with a neighbour needs >= 2 planes (rows). */
int rg_slab_label(rg_ctx *ctx, const uint8_t *d_slab, int64_t nx, int64_t ny, int64_t nz_local,
                  int has_lower, int has_upper, const int32_t *d_seeds_xyz, int nseeds, int threshold,
                  uint8_t *d_mask_slab, void *stream);
/* worst-case number of entries of one interface list (size the export buffers with it); nz_local == 1 means
 * row slabs of an image (the interface is a row), anything else z-slabs of a volume (the interface is a plane) */
int64_t rg_slab_interface_capacity(int64_t nx, int64_t ny, int64_t nz_local);
/* side 0 = interface with the lower rank, 1 = with the upper rank.  d_roots_out receives, for every
 * cross-slab edge in canonical order, the local root node on this side; d_counts3[side] the
 * number of entries and d_counts3[2] the rank's node count. */
int rg_slab_export(rg_ctx *ctx, int side, uint32_t *d_roots_out, uint32_t *d_counts3, void *stream);
/* union-find over (rank, root) vertices: vertex = root ? vertex_offset(rank) + root : 0 */
int rg_slab_resolve_init(rg_ctx *ctx, uint32_t *d_xpar, const uint32_t *d_roots, uint32_t n, uint32_t vertex_offset, void *stream);
int rg_slab_resolve_union(rg_ctx *ctx, uint32_t *d_xpar, const uint32_t *d_upper_of_lower_rank, uint32_t off_lower_rank,
                          const uint32_t *d_lower_of_upper_rank, uint32_t off_upper_rank, uint32_t n, void *stream);
int rg_slab_resolve_apply(rg_ctx *ctx, uint32_t *d_xpar, const uint32_t *d_own_roots, uint32_t n, uint32_t own_vertex_offset, void *stream);
/* The three resolve steps for all interfaces of all ranks in one call (three launches): d_lists is the
 * all-gathered buffer [world][2][stride] of interface lists, counts4 the all-gathered HOST array
 * [world][4] = {entries towards lower, entries towards upper, node count, unused}.  Vertex offsets are
 * the running sum of the node counts.  Equivalent to the init / union / apply calls above. */
int rg_slab_resolve(rg_ctx *ctx, uint32_t *d_xpar, const uint32_t *d_lists, uint32_t stride, const int64_t *counts4,
                    int rank, int world, void *stream);
/* Optimistic exchange (no host wait in mid-step).  The caller guesses capacities -- typically from the previous
 * grow of the same volume: `cap` entries per interface list and `xpar_cap` vertices.  rg_slab_export_packed writes
 * this rank's block {entries towards lower, entries towards upper, node id space, 0, lower list[cap], upper list[cap]}
 * (4 + 2*cap words); the blocks of all ranks are all-gathered into d_packed_all; rg_slab_resolve_packed validates the
 * counts ON THE DEVICE and resolves.  If a guess was too small nothing is written out of bounds: the status word gets
 * bit 4 (rg_ctx_status then returns RG_ERR_INTERNAL) and the caller repeats the grow with rg_slab_export / rg_slab_resolve. */
int rg_slab_export_packed(rg_ctx *ctx, uint32_t *d_packed, uint32_t cap, void *stream);
int rg_slab_resolve_packed(rg_ctx *ctx, uint32_t *d_xpar, uint32_t xpar_cap, const uint32_t *d_packed_all, uint32_t cap,
                           int rank, int world, void *stream);
/* pass C on the slab labelled last */
int rg_slab_finish(rg_ctx *ctx, void *stream);

/* The packed exchange without a gather: rank r exports into a buffer the other processes of the node have mapped (rg_ipc_alloc /
 * rg_ipc_open below), the ranks meet at one small barrier (any all-reduce on the stream), and rg_slab_resolve_peers reads every
 * rank's record where it lies -- d_peer_records is a DEVICE array of `world` pointers, entry r = rank r's record in this process's
 * address space (its own buffer for r == rank).  The lists cross NVLink inside the resolve kernels (plain peer loads).  Alternate
 * two export buffers from step to step: a fast rank may export its next step while a slow one still reads this one.
 * No counterpart in the reference (its exchange() is four MPI_Send / MPI_Recv pairs per round, region.c:250-353). */
int rg_slab_resolve_peers(rg_ctx *ctx, uint32_t *d_xpar, uint32_t xpar_cap, const uint32_t *const *d_peer_records, uint32_t cap,
                          int rank, int world, void *stream);
/* device buffers that other processes on the node can map: rg_ipc_alloc = cudaMalloc + the 64-byte handle to hand to the peers
 * (torch.distributed, MPI, a pipe ...), rg_ipc_open maps a peer's buffer, rg_ipc_close unmaps it, rg_ipc_free frees one's own. */
int rg_ipc_alloc(int64_t bytes, void **d_ptr_out, unsigned char handle_out[64]);
int rg_ipc_open(const unsigned char handle[64], void **d_ptr_out);
int rg_ipc_close(void *d_ptr);
int rg_ipc_free(void *d_ptr);

/* ---- synthetic inputs on the device (replaces func()/create_data(), raycast.cu:114-158) ---- */
/* kind: 1 = exercise geometry with hash noise (C1h), 2 = C2 image (n x n), 3 = C3 serpentine,
 * 4 = C4 (exercise geometry + z-serpentine).  Cube / square of side n; seed = noise seed. */
int rg_synth_device(int kind, uint8_t *d_out, int64_t n, uint64_t seed, void *stream);
/* slab version: writes planes [z_begin, z_end) of the n^3 volume to d_out (slab-local); for kind 2 (an image)
 * the range is a ROW range and d_out points at row z_begin. */
int rg_synth_slab_device(int kind, uint8_t *d_out, int64_t n, int64_t z_begin, int64_t z_end, uint64_t seed, void *stream);
/* seed lists that go with the synthetic inputs; out = max_seeds*(x,y,z) host ints; returns count */
int rg_synth_seeds(int kind, int64_t n, uint64_t seed, int32_t *out_xyz, int max_seeds);

#ifdef __cplusplus
}
#endif
#endif
