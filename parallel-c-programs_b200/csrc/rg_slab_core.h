/*
 * rg_slab_core.h -- kernel bodies of the multi-GPU slab path (z-slabs of a volume, row slabs of an image).
 *
 * Replaces the distributed half of 2-mpi-region-growing/region.c: exchange() (:250-353),
 * add_halo_to_stack() (:355-387) and finished() (:435-440).  The reference swaps mask borders
 * and re-floods until an all-reduce reports no progress, so its round count grows with the
 * number of times a region crosses a partition boundary.  Here every rank labels its slab once
 * (pass A + B of rg_core.h, with the slab's outer z faces treated as tile faces), then
 *   export_interface   lists, for each cross-slab edge of a boundary plane, the LOCAL ROOT of the
 *                      component on this side (both sides enumerate the same edges in the same
 *                      order, exactly like two tiles sharing a face);
 *   [all-gather of the lists: torch.distributed / NCCL over NVLink, done by the host layer]
 *   resolve_*          every rank solves the small union-find over (rank, root) vertices
 *                      redundantly -- vertex 0 is the shared "holds a seed" vertex -- and hooks
 *                      its own roots that reach vertex 0 to its local node 0;
 * and pass C runs unchanged.  One exchange, whatever the number of crossings.
 */
#pragma once
#include "rg_core.h"

namespace rg {

/* The interface of a slab: volumes are cut along z (the tile layers tz = 0 / ntz-1 with their ZM / ZP faces),
 * images along y (the tile rows ty = 0 / nty-1 with their YM / YP faces; nz = ntz = 1).  Tiles of a layer / row
 * are consecutive in tile order either way. */
template <class C> RG_HDFN int iface_tiles(const Params &P) { return C::HZ ? P.ntx * P.nty : P.ntx; }
template <class C> RG_HDFN uint32_t iface_first_tile(const Params &P, int side)
{
    if (!side) return 0u;
    return C::HZ ? (uint32_t)(P.ntz - 1) * (uint32_t)(P.ntx * P.nty) : (uint32_t)(P.nty - 1) * (uint32_t)P.ntx;
}
template <class C> RG_HDFN int iface_face(int side) { return C::HZ ? (side ? F_ZP : F_ZM) : (side ? F_YP : F_YM); }

/* side 0: lower faces of the bottom tile layer; side 1: upper faces of the top tile layer.
 * tile_off[0..iface_tiles] = exclusive prefix of the per-tile entry counts (scratch, written here). */
template <class C, class Exec>
RG_FN void export_offsets(Exec &ex, const Params &P, int side, uint32_t *tile_off, uint32_t *count_out, uint32_t *scratch)
{
    const int nt = iface_tiles<C>(P);
    const uint32_t layer = iface_first_tile<C>(P, side);
    const int f = iface_face<C>(side);
    uint32_t total = ex.scan32(nt, [&](int i) { return (uint32_t)P.rec[layer + (uint32_t)i].cnt[f]; }, tile_off, scratch);
    if (ex.tid() == 0) { tile_off[nt] = total; *count_out = total; }
}

enum { EXPORT_LANES = 32 };     /* one warp per boundary tile */
template <class C, class Exec>
RG_FN void export_entries(Exec &ex, const Params &P, int side, const uint32_t *tile_off, uint32_t *out_roots, uint32_t cap = 0xFFFFFFFFu)
{
    const int nt = iface_tiles<C>(P);
    const uint32_t layer = iface_first_tile<C>(P, side);
    const int f = iface_face<C>(side);
    ex.gpfor((int64_t)nt * EXPORT_LANES, [&](int64_t it) {
        const uint32_t lane = (uint32_t)(it % EXPORT_LANES);
        const int64_t i = it / EXPORT_LANES;
        uint32_t tile = layer + (uint32_t)i;
        const TileRec &rc = P.rec[tile];
        uint32_t cnt = rc.cnt[f], off = 0;
        for (int k = 0; k < f; k++) off += rc.cnt[k];
        const uint16_t *e = P.ent + (size_t)tile * C::ENT_MAX + off;
        uint32_t *o = out_roots + tile_off[i];
        const uint32_t first = tile_off[i];
        for (uint32_t k = lane; k < cnt; k += EXPORT_LANES)
            if (first + k < cap) o[k] = find32(P.nparent, rc.node_base + (e[k] & (ENT_LEAF - 1u)));      /* a list longer than the buffer is reported, not written */
    });
}

#if !RG_DEVICE_BUILD
/* sequential convenience used by the emulator */
template <class C, class Exec>
RG_FN void export_interface(Exec &ex, const Params &P, int side, uint32_t *out_roots, uint32_t *count_out)
{
    const int nt = iface_tiles<C>(P);
    uint32_t *tile_off = new uint32_t[(size_t)nt + 1];
    uint32_t scratch[40];
    export_offsets<C>(ex, P, side, tile_off, count_out, scratch);
    export_entries<C>(ex, P, side, tile_off, out_roots);
    delete[] tile_off;
}
#endif

/* ---- blocks of an image (rows x columns): the four borders of a block labelled with Params::open ----
   side 0 / 1 / 2 / 3 = -x / +x / -y / +y.  One entry per border pixel, in pixel order: the local root of its component and
   the pixel's value (the exchange compares the values across the seam: no halo was read when the block was labelled). */
template <class C> RG_HDFN int64_t border_len(const Params &P, int side) { return side < 2 ? P.ny : P.nx; }
template <class C, class Exec>
RG_FN void export_border(Exec &ex, const Params &P, int side, uint32_t *out_roots, uint8_t *out_vals)
{
    const int f = side;                                           /* F_XM, F_XP, F_YM, F_YP */
    ex.gpfor(border_len<C>(P, side), [&](int64_t i) {
        const int64_t x = side == 0 ? 0 : side == 1 ? P.nx - 1 : i, y = side == 2 ? 0 : side == 3 ? P.ny - 1 : i;
        const uint32_t tile = (uint32_t)(y / C::TY) * (uint32_t)P.ntx + (uint32_t)(x / C::TX);
        const TileRec &rc = P.rec[tile];
        uint32_t off = 0;
        for (int k = 0; k < f; k++) off += rc.cnt[k];
        const uint32_t idx = (uint32_t)(side < 2 ? y % C::TY : x % C::TX);      /* every valid border pixel is an entry, in order */
        uint32_t r = 1u;                                            /* (a valid vertex, should the self-check below ever fire) */
        if (P.tmode == 0) r = 0u;                                   /* threshold <= 0: no pixel is similar to any other, nothing to exchange */
        else if (idx < rc.cnt[f]) r = find32(P.nparent, rc.node_base + (P.ent[(size_t)tile * C::ENT_MAX + off + idx] & (ENT_LEAF - 1u)));
        else atomic_or(&P.counters[CTR_STATUS], ST_FACE_MISMATCH);
        out_roots[i] = r;
        out_vals[i] = P.image[y * P.nx + x];
    });
}

RG_FN uint32_t xvertex(uint32_t root, uint32_t off) { return root ? off + root : 0u; }

/* a seam between two blocks: entry i of both lists faces entry i of the other; united where the two pixels are similar */
template <class Exec>
RG_FN void resolve_union_seam(Exec &ex, uint32_t *xpar, const uint32_t *a, uint32_t offa, const uint8_t *va,
                              const uint32_t *b, uint32_t offb, const uint8_t *vb, uint32_t n, uint32_t thr)
{
    ex.gpfor((int64_t)n, [&](int64_t i) {
        if (absdiff(va[i], vb[i]) < thr) unite32(xpar, xvertex(a[i], offa), xvertex(b[i], offb));
    });
}

/* lazily initialise only the vertices that appear on an interface */
template <class Exec>
RG_FN void resolve_init(Exec &ex, uint32_t *xpar, const uint32_t *roots, uint32_t n, uint32_t off)
{
    ex.gpfor((int64_t)n, [&](int64_t i) { uint32_t v = xvertex(roots[i], off); xpar[v] = v; });
}

template <class Exec>
RG_FN void resolve_union(Exec &ex, uint32_t *xpar, const uint32_t *a, uint32_t offa, const uint32_t *b, uint32_t offb, uint32_t n)
{
    ex.gpfor((int64_t)n, [&](int64_t i) { unite32(xpar, xvertex(a[i], offa), xvertex(b[i], offb)); });
}

/* hook this rank's roots that reached the shared seeded vertex to the local node 0 */
template <class Exec>
RG_FN void resolve_apply(Exec &ex, const Params &P, uint32_t *xpar, const uint32_t *roots, uint32_t n, uint32_t off)
{
    ex.gpfor((int64_t)n, [&](int64_t i) {
        uint32_t r = roots[i];
        if (r && find32(xpar, off + r) == 0u) unite32(P.nparent, r, 0u);
    });
}

/* ---- all interfaces of all ranks in three sweeps (what rg_slab_resolve launches) ----
   lists[(r * 2 + side) * stride + i]: rank r's list towards its lower (side 0) / upper (side 1) neighbour */
enum { RESOLVE_MAX_RANKS = 64 };
struct ResolveArgs {
    uint32_t world, rank, stride, pad;
    uint32_t cnt[RESOLVE_MAX_RANKS][2];
    uint32_t off[RESOLVE_MAX_RANKS];
    const uint32_t *base[RESOLVE_MAX_RANKS];   /* packed exchange: where rank r's record lives -- a slice of the gathered buffer, or that
                                                  rank's own export buffer mapped into this process (peer reads over NVLink, no gather) */
};

template <class Exec>
RG_FN void resolve_init_all(Exec &ex, uint32_t *xpar, const uint32_t *lists, const ResolveArgs &A)
{
    ex.gpfor((int64_t)A.world * 2 * A.stride, [&](int64_t it) {
        const uint32_t i = (uint32_t)(it % A.stride), rs = (uint32_t)(it / A.stride), r = rs >> 1, side = rs & 1u;
        if (i < A.cnt[r][side]) { uint32_t v = xvertex(lists[(size_t)rs * A.stride + i], A.off[r]); xpar[v] = v; }
    });
}
template <class Exec>
RG_FN void resolve_union_all(Exec &ex, uint32_t *xpar, const uint32_t *lists, const ResolveArgs &A)
{
    ex.gpfor((int64_t)(A.world - 1) * A.stride, [&](int64_t it) {
        const uint32_t i = (uint32_t)(it % A.stride), r = (uint32_t)(it / A.stride);
        if (i < A.cnt[r][1])
            unite32(xpar, xvertex(lists[(size_t)(r * 2 + 1) * A.stride + i], A.off[r]),
                          xvertex(lists[(size_t)((r + 1) * 2) * A.stride + i], A.off[r + 1]));
    });
}
template <class Exec>
RG_FN void resolve_apply_all(Exec &ex, const Params &P, uint32_t *xpar, const uint32_t *lists, const ResolveArgs &A)
{
    ex.gpfor((int64_t)2 * A.stride, [&](int64_t it) {
        const uint32_t i = (uint32_t)(it % A.stride), side = (uint32_t)(it / A.stride);
        if (i >= A.cnt[A.rank][side]) return;
        const uint32_t r = lists[(size_t)(A.rank * 2 + side) * A.stride + i];
        if (r && find32(xpar, A.off[A.rank] + r) == 0u) unite32(P.nparent, r, 0u);
    });
}

/* ---- optimistic exchange: capacities come from the previous step, so the host never waits in mid-step ----
   packed[r] = { entries towards lower, entries towards upper, node id space, unused, lower list[cap], upper list[cap] }.
   resolve_prepare turns the gathered counts into ResolveArgs on the device and refuses (status bit, empty
   work) if a list outgrew `cap` or the vertex space outgrew `xpar_cap`. */
enum { ST_EXCHANGE_OVERFLOW = 4u };
RG_FN void resolve_prepare(const Params &P, const uint32_t *packed_all, uint32_t stride_words, uint32_t cap, uint32_t xpar_cap,
                           uint32_t rank, uint32_t world, ResolveArgs *A, const uint32_t *const *peer_records = nullptr)
{
    bool ok = true;
    uint64_t off = 1;
    A->world = world; A->rank = rank; A->stride = stride_words; A->pad = 0;
    for (uint32_t r = 0; r < world; r++) {
        const uint32_t *h = peer_records ? peer_records[r] : packed_all + (size_t)r * stride_words;
        A->base[r] = h;
        A->cnt[r][0] = h[0]; A->cnt[r][1] = h[1];
        if (h[0] > cap || h[1] > cap) ok = false;
        A->off[r] = (uint32_t)off; off += h[2];
    }
    for (uint32_t r = 0; r + 1 < world; r++) if (A->cnt[r][1] != A->cnt[r + 1][0]) ok = false;
    if (off + 1 > (uint64_t)xpar_cap) ok = false;
    if (!ok) {
        for (uint32_t r = 0; r < world; r++) A->cnt[r][0] = A->cnt[r][1] = 0;
        atomic_or(&P.counters[CTR_STATUS], ST_EXCHANGE_OVERFLOW);
    }
}
/* list of (rank r, side) inside the packed buffer: header of 4 words, then the two lists */
RG_FN const uint32_t *packed_list(const uint32_t *, const ResolveArgs &A, uint32_t cap, uint32_t r, uint32_t side)
{
    return A.base[r] + 4 + (size_t)side * cap;
}
template <class Exec>
RG_FN void resolve_init_packed(Exec &ex, uint32_t *xpar, const uint32_t *packed_all, uint32_t cap, const ResolveArgs &A)
{
    ex.gpfor((int64_t)A.world * 2 * cap, [&](int64_t it) {
        const uint32_t i = (uint32_t)(it % cap), rs = (uint32_t)(it / cap), r = rs >> 1, side = rs & 1u;
        if (i < A.cnt[r][side]) { uint32_t v = xvertex(packed_list(packed_all, A, cap, r, side)[i], A.off[r]); xpar[v] = v; }
    });
}
template <class Exec>
RG_FN void resolve_union_packed(Exec &ex, uint32_t *xpar, const uint32_t *packed_all, uint32_t cap, const ResolveArgs &A)
{
    ex.gpfor((int64_t)(A.world - 1) * cap, [&](int64_t it) {
        const uint32_t i = (uint32_t)(it % cap), r = (uint32_t)(it / cap);
        if (i < A.cnt[r][1])
            unite32(xpar, xvertex(packed_list(packed_all, A, cap, r, 1)[i], A.off[r]), xvertex(packed_list(packed_all, A, cap, r + 1, 0)[i], A.off[r + 1]));
    });
}
template <class Exec>
RG_FN void resolve_apply_packed(Exec &ex, const Params &P, uint32_t *xpar, const uint32_t *packed_all, uint32_t cap, const ResolveArgs &A)
{
    ex.gpfor((int64_t)2 * cap, [&](int64_t it) {
        const uint32_t i = (uint32_t)(it % cap), side = (uint32_t)(it / cap);
        if (i >= A.cnt[A.rank][side]) return;
        const uint32_t r = packed_list(packed_all, A, cap, A.rank, side)[i];
        if (r && find32(xpar, A.off[A.rank] + r) == 0u) unite32(P.nparent, r, 0u);
    });
}

} // namespace rg
