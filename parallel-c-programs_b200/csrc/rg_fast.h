/*
 * rg_fast.h -- the cheap paths in front of the full tile passes of rg_core.h.
 *
 * Measured on B200 (profiles/r01_v1_ncu_summary.md) the full labelling pass is instruction bound:
 * it builds three edge bit planes and runs a union-find over every run of every tile, although in
 * the inputs the reference generates (rand()%20 noise around a few solids, raycast.cu:114-144)
 * almost every component is a handful of voxels and only ~1 % of them reach a tile face.  So:
 *
 *   tile_pass_a_fast  one CTA per tile, 57 KB smem (4 CTAs/SM, lean staging layout): stage the tile with TMA and, while
 *                     the load is in flight, clear the tile's part of the mask with bulk stores; compute edge words ONLY
 *                     on the six face layers, then let one thread per cross-tile edge (and per seed) walk its
 *                     component directly on the staged bytes, at most WALK_MAX voxels.  The walk returns
 *                     the component's smallest voxel index -- the same representative the union-find paths
 *                     produce -- and a shared-memory hash table turns representatives
 *                     into dense node ids, so nodes and face lists come out as in rg_core.h.  A tile whose voxels are all
 *                     equal is one component; a tile of constant rows is labelled as a grid of rows (rg_rows.h); a tile
 *                     in which a walk overflows (a solid's surface, a corridor ...) is labelled on the spot by the dense
 *                     path (rg_dense.h).  Only what overflows even that goes to the work list of the full pass.
 *   flag_tiles_all    one warp per tile, no smem: decides whether any node of the tile ended up in the
 *                     region (root == node 0).  Uniform region tiles are filled with ones, the other region
 *                     tiles are appended to the pass C work list; empty tiles were cleared by the fast pass.
 *
 * Same reference semantics as rg_core.h (region.c:493-533, raycast.cu:281-318); results are
 * identical whichever path a tile takes.
 */
#pragma once
#include "rg_core.h"
#include "rg_dense.h"
#include "rg_rows.h"

namespace rg {

#ifndef RG_WALK_MAX
#define RG_WALK_MAX 16
#endif
enum { WALK_MAX = RG_WALK_MAX };                     /* voxels a walk may visit before the tile is called complex */
enum { FAST_ENT_MAX = 768 };                  /* cross-tile edges a non-uniform tile may have on the fast path */
enum { HASH_BITS = 10, HASH_N = 1 << HASH_BITS, HASH_MAX_KEYS = 704 };
static_assert((int)HASH_N == (int)FAST_NODE_SLOTS, "one node slot per hash slot");   /* representative -> dense node id */

template <class C>
struct alignas(128) FastSmem {
    union alignas(128) U {
        uint8_t img[C::IMG_BYTES];                /* the staged tile */
        DenseSmem<C> d;                           /* ... and, once its link words are in registers, the state of the dense labelling path (rg_dense.h) */
    } u;
    /* open addressing: representative + 1, 0 = empty (32-bit: shared-memory CAS is native).  All zeros until the walks start:
       the source of the bulk stores that clear the tile's part of the mask (128-byte aligned for them). */
    alignas(128) uint32_t tab[HASH_N];
    uint64_t mbar;
    /* Storage shared over time.  The cross-tile edge words and their prefix sums are dead once the edge list exists and the
       list of the longer walks takes their place; that list and the x = TX halo column (lean layout: one byte per (y, z),
       used by the face masks only) are dead once the walks are done, and the dense node ids of the slots go there. */
    union {
        struct {
            union {
                struct { uint32_t fmask[C::FW_TOTAL]; uint16_t fprefix[C::FW_TOTAL + 2]; } f;
                uint16_t slow[FAST_ENT_MAX];      /* edges whose start voxel has a similar neighbour inside the tile */
            } u;
            uint8_t xr[C::XR ? 16 : C::TY * C::TZ];
        } e;
        uint16_t slotid[HASH_N];                  /* dense node id of each occupied slot (scan after the walks) */
    } v;
    uint16_t ent_root[FAST_ENT_MAX];              /* per edge: start voxel, later the hash slot of its component */
    alignas(16) uint32_t seedtab[HASH_N / 32];    /* bit per slot: the component holds a seed */
    uint32_t scal[24];                            /* SC_* and, from SC_CNT0 on, the six face-list lengths */
    uint32_t scan_scratch[40];
    static_assert(C::NV <= 32768, "walk queue entries are 15-bit voxel indices + a valid bit");
};

/* find or insert representative `root`; returns its slot.  Dense node ids are handed out afterwards by one
   scan over the slots (a shared counter bumped by every inserting thread serialises the whole phase). */
RG_FN uint32_t hash_slot(uint32_t *tab, uint32_t *scal, uint32_t root)
{
    const uint32_t key = root + 1u;
    uint32_t slot = (root * 2654435761u) >> (32 - HASH_BITS);
    for (int probe = 0; probe < HASH_N; probe++) {
        uint32_t cur = ldv32(&tab[slot]);
        if (cur == 0u) cur = cas32(&tab[slot], 0u, key);
        if (cur == 0u || cur == key) return slot;
        slot = (slot + 1u) & (HASH_N - 1);
    }
    scal[SC_COMPLEX] = 1;
    return 0;
}

/* ---- a tiny queue of 16-bit entries packed in four 64-bit registers ---- */
struct WalkQ { uint64_t a, b, c, d; };
RG_FN void wq_clear(WalkQ &w) { w.a = w.b = w.c = w.d = 0; }
RG_FN void wq_set(WalkQ &w, int i, uint32_t v)
{
    const uint64_t e = (uint64_t)(v | 0x8000u) << (16 * (i & 3));
    const int k = i >> 2;
    w.a |= (k == 0) ? e : 0; w.b |= (k == 1) ? e : 0; w.c |= (k == 2) ? e : 0; w.d |= (k == 3) ? e : 0;
}
RG_FN uint32_t wq_get(const WalkQ &w, int i)
{
    const int k = i >> 2;
    const uint64_t x = (k == 0) ? w.a : (k == 1) ? w.b : (k == 2) ? w.c : w.d;
    return (uint32_t)(x >> (16 * (i & 3))) & 0x7FFFu;
}
RG_FN bool wq_zero16(uint64_t x) { return ((x - 0x0001000100010001ULL) & ~x & 0x8000800080008000ULL) != 0; }
RG_FN bool wq_has(const WalkQ &w, uint32_t v, int n)
{
    const uint64_t pat = (uint64_t)(v | 0x8000u) * 0x0001000100010001ULL;   /* a zero half-word of q ^ pat <=> v is queued */
    bool hit = wq_zero16(w.a ^ pat);
    if (n > 4) {                                                            /* rarely: most components have <= 4 voxels */
        hit = hit || wq_zero16(w.b ^ pat);
        if (n > 8) hit = hit || wq_zero16(w.c ^ pat) || wq_zero16(w.d ^ pat);
    }
    return hit;
}

/* which in-tile neighbours of voxel v are similar to it: one bit per direction (-x +x -y +y -z +z) */
template <class C>
RG_FN uint32_t neighbour_mask(const uint8_t *img, int vx, int vy, int vz, uint32_t thr, uint32_t v)
{
    const int x = (int)(v % C::TX), y = (int)((v / C::TX) % C::TY), z = (int)(v / (C::TX * C::TY));
    const uint8_t *p = img + img_off<C>(x, y, z);
    const uint32_t a = *p;
    uint32_t nb = (uint32_t)(x > 0 && absdiff(a, p[-1]) < thr) | ((uint32_t)(x + 1 < vx && absdiff(a, p[1]) < thr) << 1)
                | ((uint32_t)(y > 0 && absdiff(a, p[-C::RS]) < thr) << 2) | ((uint32_t)(y + 1 < vy && absdiff(a, p[C::RS]) < thr) << 3);
    if (C::HZ) nb |= ((uint32_t)(z > 0 && absdiff(a, p[-C::PS]) < thr) << 4) | ((uint32_t)(z + 1 < vz && absdiff(a, p[C::PS]) < thr) << 5);
    return nb;
}

/* Walk the component of voxel `start` on the staged bytes (6- or 4-connectivity inside the tile's
 * valid extent); nb = neighbour_mask(start).  Returns false if the component has more than WALK_MAX
 * voxels; else root = smallest voxel index, n = size, and the visited voxels are left in wq. */
template <class C>
RG_FN bool walk_component(const uint8_t *img, int vx, int vy, int vz, uint32_t thr, uint32_t start, uint32_t nb,
                          WalkQ &wq, int &n, uint32_t &root)
{
    n = 1; root = start;
    wq_clear(wq);
    wq_set(wq, 0, start);
    int head = 0;
    uint32_t v = start;
    for (;;) {
        while (nb) {                                                /* usually one bit */
            const int d = ffs1(nb) - 1; nb &= nb - 1;
            const int step = (d < 2) ? 1 : (d < 4) ? C::TX : C::TX * C::TY;
            const uint32_t u = (d & 1) ? v + step : v - step;
            if (wq_has(wq, u, n)) continue;
            if (n == WALK_MAX) return false;
            wq_set(wq, n++, u);
            if (u < root) root = u;
        }
        if (++head >= n) break;
        v = wq_get(wq, head);
        nb = neighbour_mask<C>(img, vx, vy, vz, thr, v);
    }
    return true;
}

/* cross-tile edge masks of the six faces, from the staged bytes (the same words build_edges<FACES>
 * produces, without touching the interior) */
template <class C, class Exec>
RG_FN void build_face_masks(Exec &ex, const Params &P, const uint8_t *img, const uint8_t *xr, uint32_t *fmask, int64_t x0, int64_t y0, int64_t z0)
{
    const int64_t nx = P.nx, ny = P.ny, nz = P.nz;
    const int tmode = P.tmode; const uint32_t t4 = P.t4;
    const int vy = (int)((ny - y0) < C::TY ? (ny - y0) : C::TY);
    const int vz = (int)((nz - z0) < C::TZ ? (nz - z0) : C::TZ);
    const bool has_yp = (y0 + C::TY < ny) || (y0 + vy == ny && ext_yhi<C>(P)), has_ym = y0 > 0 || ext_ylo<C>(P);
    const bool has_zp = C::HZ && ((z0 + C::TZ < nz) || (z0 + vz == nz && ext_zhi<C>(P)));
    const bool has_zm = C::HZ && (z0 > 0 || ext_zlo<C>(P));
    const bool has_xp = x0 + C::TX < nx, has_xm = x0 > 0;
    /* seams of a block decomposition (Params::open): all valid border pixels, no comparison, no dedupe */
    const bool live = tmode != 0;
    const bool open_xm = live && (P.open & 1u) && x0 == 0, open_xp = live && (P.open & 2u) && x0 + C::TX == nx;
    const bool open_ym = live && (P.open & 4u) && y0 == 0, open_yp = live && (P.open & 8u) && y0 + vy == ny;
    /* y faces: item = (side, z, wx); z faces: item = (side, y, wx) */
    ex.pfor(2 * C::FYW + 2 * C::FZW, [&](int it) {
        const bool zface = it >= 2 * C::FYW;
        int r = zface ? it - 2 * C::FYW : it;
        const int nper = zface ? C::FZW : C::FYW;
        const int side = r / nper; r %= nper;
        const int wx = r % C::WX, o = r / C::WX;                   /* o = z for y faces, y for z faces */
        int y, z, step, f;
        if (!zface) {
            const bool seam = side ? open_yp : open_ym;
            if (!seam && (side ? !has_yp : !has_ym)) return;
            z = o; y = side ? vy - 1 : 0; step = side ? C::RS : -C::RS; f = side ? F_YP : F_YM;
            if (z >= vz) return;
            if (seam) { fmask[C::fw_off(f) + o * C::WX + wx] = valid_bits(nx - (x0 + 32 * wx)); return; }
        } else {
            if (side ? !has_zp : !has_zm) return;
            y = o; z = side ? vz - 1 : 0; step = side ? C::PS : -C::PS; f = side ? F_ZP : F_ZM;
            if (y >= vy) return;
        }
        const uint32_t vmask = valid_bits(nx - (x0 + 32 * wx));
        if (!vmask) return;
        const uint8_t *rowp = img + img_off<C>(32 * wx, y, z);
        uint32_t a[8], b[8];
        load8(rowp, a); load8(rowp + step, b);
        const uint32_t m = edge_word(a, b, tmode, t4) & vmask;
        if (m) fmask[C::fw_off(f) + o * C::WX + wx] = dedupe(m, xedges_within(a, tmode, t4), xedges_within(b, tmode, t4));
    });
    /* x faces: one bit per (y, z) */
#if RG_DEVICE_BUILD
    if (C::TY == 32) {
        /* a warp per (side, plane): lane = row, the 32 verdicts are one mask word -- no index arithmetic, no atomics */
        const int lane = ex.tid() & 31, nwarp = Exec::NTHREADS / 32;
        for (int wi = ex.tid() >> 5; wi < 2 * C::TZ; wi += nwarp) {
            const int side = wi / C::TZ, z = wi % C::TZ, y = lane, q = z * C::TY + y;
            const bool seam = side ? open_xp : open_xm;
            if (!seam && (side ? !has_xp : !has_xm)) continue;               /* the same for the whole warp */
            bool e = false;
            if (y < vy && z < vz) {
                const uint8_t *o = img + img_off<C>(side ? C::TX - 1 : 0, y, z);
                const uint32_t own = o[0], other = side ? (C::XR ? o[1] : xr[q]) : o[-1];
                e = seam || absdiff(own, other) < P.thr;
            }
            const uint32_t m = __ballot_sync(0xFFFFFFFFu, e);
            if (lane == 0 && m) fmask[C::fw_off(side ? F_XP : F_XM) + z] = m;
        }
        return;
    }
#endif
    ex.pfor(2 * C::TY * C::TZ, [&](int it) {
        const int side = it / (C::TY * C::TZ), q = it % (C::TY * C::TZ);
        const int y = q % C::TY, z = q / C::TY;
        const bool seam = side ? open_xp : open_xm;
        if ((!seam && (side ? !has_xp : !has_xm)) || y >= vy || z >= vz) return;
        const uint8_t *o = img + img_off<C>(side ? C::TX - 1 : 0, y, z);
        const uint32_t own = o[0], other = side ? (C::XR ? o[1] : xr[q]) : o[-1];       /* lean layout: the x = TX column is in xr */
        bool e = seam || absdiff(own, other) < P.thr;
        if (!C::HZ && e && !seam && y > 0) {
            /* images (tiles 126 rows high): one entry per pair of overlapping y-runs, the way the y faces keep one per pair of x-runs --
               redundant if the row above has an edge too and both columns continue from there to here; both tiles see the same four
               pixels.  (For volumes the same rule costs the labelling pass more than it saves the merge: measured, left out.) */
            const uint32_t own1 = o[-C::RS], other1 = side ? o[1 - C::RS] : o[-1 - C::RS];
            if (absdiff(own1, other1) < P.thr && absdiff(own, own1) < P.thr && absdiff(other, other1) < P.thr) e = false;
        }
        if (e) atomic_or(&fmask[C::fw_off(side ? F_XP : F_XM) + (q >> 5)], 1u << (q & 31));
    });
}

/* this tile's slice [lo, hi) of the sorted seed keys */
RG_FN void seed_range(const Params &P, uint32_t tile, uint32_t &lo_out, uint32_t &hi_out)
{
    uint64_t lo_key = (uint64_t)tile << 16, hi_key = ((uint64_t)tile + 1) << 16;
    int lo = 0, hi = P.nseeds;
    while (lo < hi) { int mid = (lo + hi) >> 1; if (P.seed_keys[mid] < lo_key) lo = mid + 1; else hi = mid; }
    int a = lo; hi = P.nseeds;
    while (lo < hi) { int mid = (lo + hi) >> 1; if (P.seed_keys[mid] < hi_key) lo = mid + 1; else hi = mid; }
    lo_out = (uint32_t)a; hi_out = (uint32_t)lo;
}

/* ================================================================ fast labelling pass */
RG_FN uint32_t fast_node_base(uint32_t tile) { return 1u + tile * (uint32_t)HASH_N; }   /* fixed slots: no global atomic */

template <class C, class Exec>
RG_FN void fast_give_up(Exec &ex, const Params &P, uint32_t tile)
{
    if (ex.tid() == 0) P.complex_list[atomic_add(&P.counters[CTR_NCOMPLEX], 1u)] = tile;
}

template <class C, class Exec>
RG_FN void tile_pass_a_fast(Exec &ex, const Params &P, uint32_t tile, FastSmem<C> &s)
{
    int tx, ty, tz; ex.template origin<C>(P, tile, tx, ty, tz);
    const int64_t x0 = (int64_t)tx * C::TX, y0 = (int64_t)ty * C::TY, z0 = (int64_t)tz * C::TZ;
    const int vx = (int)((P.nx - x0) < C::TX ? (P.nx - x0) : C::TX);
    const int vy = (int)((P.ny - y0) < C::TY ? (P.ny - y0) : C::TY);
    const int vz = (int)((P.nz - z0) < C::TZ ? (P.nz - z0) : C::TZ);

    ex.pfor(HASH_N / 4, [&](int i) { st128(&s.tab[4 * i], u32x4{0, 0, 0, 0}); });
    ex.sync_async_smem();                                           /* the zeros are visible to the copy engine; the load barrier is initialised */
    if (!ex.template tma_tile<C>(P, s.u.img, &s.mbar, x0, y0, z0)) load_tile<C>(ex, P, s.u.img, x0, y0, z0, true);   /* copies are in flight from here */
    /* the tile's part of the mask is cleared while the load is in flight: bulk stores of the (still empty) hash table,
       one instruction each, or -- unaligned buffers, no TMA -- fire-and-forget vector stores */
    if (P.mask_prezeroed && !ex.template tma_clear_tile<C>(P, s.tab, (int)sizeof(s.tab), x0, y0, z0)) zero_tile<C>(ex, P, x0, y0, z0);
    ex.pfor(C::FW_TOTAL, [&](int i) { s.v.e.u.f.fmask[i] = 0; });
    auto load_xr = [&]() {
        if (C::XR) return;
        /* lean layout: the x = TX column is not in the staged box; one byte per (y, z) straight from the image (the
           neighbouring tile's CTA reads the same lines, so these come from L2) */
        const bool has_col = x0 + C::TX < P.nx;
        ex.pfor(C::TY * C::TZ, [&](int q) {
            const int y = q % C::TY, z = q / C::TY;
            s.v.e.xr[q] = (has_col && y < vy && z < vz) ? P.image[((z0 + z) * P.ny + (y0 + y)) * P.nx + x0 + C::TX] : (uint8_t)0;
        });
    };
    load_xr();
    ex.pfor(HASH_N / 32, [&](int i) { s.seedtab[i] = 0; });
    if (ex.tid() == 0) {
        uint32_t lo, hi; seed_range(P, tile, lo, hi);
        s.scal[SC_SEED_LO] = lo; s.scal[SC_SEED_HI] = hi; s.scal[SC_NNODES] = 0; s.scal[SC_COMPLEX] = 0; s.scal[SC_NSLOW] = 0;
    }
    ex.tile_wait(&s.mbar);
    ex.sync();
    if (P.debug_stop == 1) return;
    /* a tile whose own voxels are all equal is one component with representative 0 (solid interiors) */
    /* ... and a tile whose ROWS are constant is a small grid of rows (rg_rows.h) */
    bool uniform = false, rows = false, structured = false;
    if (P.fast16 && P.tmode >= 1) {
        const uint32_t pat = (uint32_t)s.u.img[img_off<C>(0, 0, 0)] * 0x01010101u;
        u32x4 q0 = ld128(&s.u.img[img_off<C>(0, 0, 0)]);
        /* the first 16 voxels decide for almost every tile; all threads read the same bytes, so this branch is uniform */
        if (vx >= 16 && ((q0.x ^ pat) | (q0.y ^ pat) | (q0.z ^ pat) | (q0.w ^ pat)) == 0u) {
            bool differs = false, row_differs = false;
            /* volumes: eight neighbouring lanes read the same 16-byte column of eight consecutive rows -- with rows 80 bytes apart those
               are eight different groups of four banks, where the four chunks of two neighbouring rows overlap in one */
            constexpr bool BY_COLUMN = C::CX == 4 && (C::TY * C::TZ) % 8 == 0;
            ex.pfor(C::TZ * C::TY * C::CX, [&](int it) {
                const int c = BY_COLUMN ? (it >> 3) & 3 : it % C::CX, r = BY_COLUMN ? (it & 7) | ((it >> 5) << 3) : it / C::CX;
                const int y = r % C::TY, z = r / C::TY;
                if (y >= vy || z >= vz || 16 * c >= vx) return;
                u32x4 q = ld128(&s.u.img[img_off<C>(16 * c, y, z)]);
                if ((q.x ^ pat) | (q.y ^ pat) | (q.z ^ pat) | (q.w ^ pat)) {
                    differs = true;
                    const uint32_t rp = (uint32_t)s.u.img[img_off<C>(0, y, z)] * 0x01010101u;
                    if ((q.x ^ rp) | (q.y ^ rp) | (q.z ^ rp) | (q.w ^ rp)) row_differs = true;
                }
            });
            if (!ex.sync_or(row_differs)) {
                uniform = !ex.sync_or(differs);
                rows = !uniform && RowsCfg<C>::OK && vx == C::TX && P.planes && !P.no_rows && !P.force_dense;
            }
            /* sixteen equal voxels in a row do not happen in noise: the tile starts inside a solid or a corridor and is neither
               uniform nor made of rows, i.e. it holds a surface or a turn -- the walkers would give up on it, so it goes straight
               to the dense path */
            structured = !uniform && !rows;
        }
    }
    build_face_masks<C>(ex, P, s.u.img, s.v.e.xr, s.v.e.u.f.fmask, x0, y0, z0);
    ex.sync();
    if (P.debug_stop == 3) return;
    const uint32_t etot = ex.scan16(C::FW_TOTAL, [&](int i) { return (uint32_t)popc(s.v.e.u.f.fmask[i]); }, s.v.e.u.f.fprefix, s.scan_scratch);
    if (ex.tid() == 0) {
        s.v.e.u.f.fprefix[C::FW_TOTAL] = (uint16_t)etot;
        for (int f = 0; f < 6; f++) s.scal[SC_CNT0 + f] = (uint32_t)(s.v.e.u.f.fprefix[C::fw_off(f + 1)] - s.v.e.u.f.fprefix[C::fw_off(f)]);   /* the prefix sums do not live to the end */
    }
    const bool seeded = s.scal[SC_SEED_HI] > s.scal[SC_SEED_LO];
    const uint32_t base = fast_node_base(tile);
    uint16_t *ent = P.ent + (size_t)tile * C::ENT_MAX;
    uint32_t nnodes;
    if (rows) {
        /* needs a link-plane slot for its row links; without one (more such tiles than slots) it is labelled like any other tile */
        if (plane_slot_is_tile(P)) { if (ex.tid() == 0) s.scal[SC_PLANE1] = tile + 1u; }       /* read after later barriers only */
        else {
            if (ex.tid() == 0) s.scal[SC_PLANE1] = plane_slot1(P, tile);
            ex.sync();
            rows = s.scal[SC_PLANE1] != 0;
        }
    }
    if (rows) {
        if (RowsCfg<C>::OK) tile_label_rows<C>(ex, P, tile, s, vy, vz);
        return;
    }
    if (uniform) {
        ex.pfor((int)etot, [&](int e) { ent[e] = 0; });
        nnodes = (etot || seeded) ? 1u : 0u;
        if (ex.tid() == 0 && nnodes) { P.nparent[base] = seeded ? 0u : base; P.nrep[base] = 0; }
    } else {
        if (etot > (uint32_t)FAST_ENT_MAX || structured || P.force_dense) {              /* same decision in every thread */
            /* too many cross-tile edges for the walkers: label the tile from the staged bytes (rg_dense.h) */
            if (!tile_label_dense<C>(ex, P, tile, s, x0, vy, vz)) fast_give_up<C>(ex, P, tile);
            return;
        }
        /* start voxel of every cross-tile edge, in list order ... */
        ex.pfor(C::FW_TOTAL, [&](int fw) {
            uint32_t m = s.v.e.u.f.fmask[fw];
            if (!m) return;
            const int f = (fw >= C::fw_off(1)) + (fw >= C::fw_off(2)) + (fw >= C::fw_off(3)) + (fw >= C::fw_off(4)) + (fw >= C::fw_off(5));
            const int idx = fw - C::fw_off(f);
            uint32_t slot = s.v.e.u.f.fprefix[fw];
            while (m) { int b = ffs1(m) - 1; m &= m - 1; s.ent_root[slot++] = (uint16_t)face_voxel<C>(f, idx, b, vy, vz); }
        });
        ex.sync();
        /* ... then one thread per edge.  Most start voxels have no similar neighbour inside the tile and are their
           own component; the others go to a compact list so that the longer walks fill whole warps. */
        ex.pfor((int)etot, [&](int e) {
            const uint32_t v = s.ent_root[e];
            if (neighbour_mask<C>(s.u.img, vx, vy, vz, P.thr, v)) s.v.e.u.slow[atomic_add(&s.scal[SC_NSLOW], 1u)] = (uint16_t)e;
            else {
                /* a voxel with no similar neighbour inside the tile that lies on ONE face has exactly this one edge: if its partner on
                   the other side is of the same kind the two form a closed pair, and the merge need not look at them (ENT_LEAF) */
                const int x = (int)(v % C::TX), y = (int)((v / C::TX) % C::TY), z = (int)(v / (C::TX * C::TY));
                const int onfaces = (x == 0) + (x == C::TX - 1) + (y == 0) + (y == vy - 1) + (C::HZ ? (z == 0) + (z == vz - 1) : 0);
                s.ent_root[e] = (uint16_t)(hash_slot(s.tab, s.scal, v) | (onfaces == 1 ? (uint32_t)ENT_LEAF : 0u));
            }
        });
        {
            const int lo = (int)s.scal[SC_SEED_LO], hi = (int)s.scal[SC_SEED_HI];
            ex.pfor(hi - lo, [&](int i) {
                const uint32_t v = (uint32_t)(P.seed_keys[lo + i] & 0xFFFFu);
                WalkQ wq; int n; uint32_t root;
                if (!walk_component<C>(s.u.img, vx, vy, vz, P.thr, v, neighbour_mask<C>(s.u.img, vx, vy, vz, P.thr, v), wq, n, root)) { s.scal[SC_COMPLEX] = 1; return; }
                const uint32_t slot = hash_slot(s.tab, s.scal, root);
                atomic_or(&s.seedtab[slot >> 5], 1u << (slot & 31));
            });
        }
        ex.sync();
        if (P.debug_stop == 6) return;
        ex.pfor((int)s.scal[SC_NSLOW], [&](int i) {
            if (ldv32(&s.scal[SC_COMPLEX])) return;                 /* a walk has overflowed already: the tile goes to the dense path, the rest is moot */
            const int e = s.v.e.u.slow[i];
            const uint32_t v = s.ent_root[e];
            WalkQ wq; int n; uint32_t root;
            const uint32_t nb0 = neighbour_mask<C>(s.u.img, vx, vy, vz, P.thr, v);
            if (P.debug_stop == 61) { s.ent_root[e] = (uint16_t)nb0; return; }
            if (!walk_component<C>(s.u.img, vx, vy, vz, P.thr, v, nb0, wq, n, root)) { s.scal[SC_COMPLEX] = 1; return; }
            if (P.debug_stop == 62) { s.ent_root[e] = (uint16_t)root; return; }
            s.ent_root[e] = (uint16_t)hash_slot(s.tab, s.scal, root);
        });
        ex.sync();
        if (P.debug_stop == 7 || P.debug_stop == 61 || P.debug_stop == 62) return;
        /* a walk that did not end, or more components than node slots: the dense path labels the tile from the bytes
           that are still staged (the walkers have used the storage of the edge masks: rebuild them first) */
        bool dense = s.scal[SC_COMPLEX] != 0;
        nnodes = 0;
        if (!dense) {
            /* dense ids in slot order (this overwrites the halo column and the edge masks too: rebuilt below if it comes to that) */
            nnodes = ex.scan16(HASH_N, [&](int i) { return s.tab[i] ? 1u : 0u; }, s.v.slotid, s.scan_scratch);
            dense = nnodes > (uint32_t)HASH_MAX_KEYS;
        }
        if (dense) {
            ex.pfor(C::FW_TOTAL, [&](int i) { s.v.e.u.f.fmask[i] = 0; });
            load_xr();
            ex.sync();
            build_face_masks<C>(ex, P, s.u.img, s.v.e.xr, s.v.e.u.f.fmask, x0, y0, z0);
            ex.sync();
            const uint32_t etot2 = ex.scan16(C::FW_TOTAL, [&](int i) { return (uint32_t)popc(s.v.e.u.f.fmask[i]); }, s.v.e.u.f.fprefix, s.scan_scratch);
            if (ex.tid() == 0) s.v.e.u.f.fprefix[C::FW_TOTAL] = (uint16_t)etot2;
            ex.sync();
            if (!tile_label_dense<C>(ex, P, tile, s, x0, vy, vz)) fast_give_up<C>(ex, P, tile);
            return;
        }
        /* face lists and nodes go out in one sweep */
        ex.pfor((int)etot, [&](int e) {
            const uint32_t r = s.ent_root[e], slot = r & (ENT_LEAF - 1u);
            uint32_t id = s.v.slotid[slot];
            if ((r & ENT_LEAF) && !((s.seedtab[slot >> 5] >> (slot & 31)) & 1u)) id |= ENT_LEAF;      /* a seeded voxel is never a closed pair */
            ent[e] = (uint16_t)id;
        });
        ex.pfor(HASH_N, [&](int slot) {
            const uint32_t cur = s.tab[slot];
            if (!cur) return;
            const uint32_t g = base + s.v.slotid[slot];
            P.nparent[g] = ((s.seedtab[slot >> 5] >> (slot & 31)) & 1u) ? 0u : g;      /* seeded components start hooked to node 0 */
            P.nrep[g] = (uint16_t)(cur - 1u);
        });
    }
    if (ex.tid() == 0) {
        TileRec rc;
        rc.node_base = base; rc.nnodes = nnodes;
        for (int f = 0; f < 6; f++) rc.cnt[f] = (uint16_t)s.scal[SC_CNT0 + f];
        rc.uniform = uniform ? 1 : 0; rc.has_region = 0; rc.plane_slot1 = 0; rc.pad3 = 0;
        P.rec[tile] = rc;
    }
}

/* the full pass over the tiles the fast pass gave up on (persistent CTAs) */
template <class C, class Exec>
RG_FN void pass_a_full_list(Exec &ex, const Params &P, TileSmem<C> &s)
{
    const uint32_t n = P.counters[CTR_NCOMPLEX];
    for (uint32_t i = ex.bid(); i < n; i += ex.nblk()) {
        tile_pass_a<C>(ex, P, P.complex_list[i], s);
        ex.sync();
    }
}

/* x-extent of the region in every row of tiles: one item per (ty, tz), out[2 * item] = first tile column that holds region
   voxels, out[2 * item + 1] = one past the last (both 0: none).  The host copies back only those boxes of the mask. */
template <class C, class Exec>
RG_FN void region_boxes(Exec &ex, const Params &P, int32_t *out)
{
    ex.gpfor((int64_t)P.nty * P.ntz, [&](int64_t it) {
        const TileRec *row = P.rec + it * P.ntx;
        int lo = P.ntx, hi = 0;
        for (int tx = 0; tx < P.ntx; tx++) if (row[tx].has_region) { if (tx < lo) lo = tx; hi = tx + 1; }
        out[2 * it] = hi ? lo : 0; out[2 * it + 1] = hi;
    });
}

/* ================================================================ region tiles vs. empty tiles */
/* One warp per tile (a tile has a few hundred nodes at most; a CTA per tile spends its time being scheduled):
 * does any node of the tile hang under node 0?  Then a uniform tile is filled with ones on the spot and any other
 * tile goes to the list of pass C; a tile without region voxels is cleared here unless the labelling pass did that. */
enum { FLAG_LANES = 32 };
template <class C, class Exec>
RG_FN void flag_tiles_all(Exec &ex, const Params &P)
{
    const uint32_t ntiles = (uint32_t)P.ntx * (uint32_t)P.nty * (uint32_t)P.ntz;
    ex.wfor(ntiles,
        [&](uint32_t tile, uint32_t lane) -> bool {
            if (P.rec[tile].uniform == 2) return false;              /* row tiles: decided and written by the row output kernel */
            const uint32_t nn = P.rec[tile].nnodes, nb = P.rec[tile].node_base;
            /* tuning: 40 = the pass without its finds, 41 = with one plain load per node instead.  Both take 0.025 ms on C4 where the
               pass takes 0.31 (0.07 of it the ones of the uniform region tiles): the time is the SECOND step of the finds, a scattered
               sector per node that has a parent.  Four first steps in flight per lane, cached or gpu-scope loads instead of
               volatile ones, finds without path compression: each was measured and none is faster. */
            if (P.debug_stop == 40) return false;
            if (P.debug_stop == 41) { uint32_t acc = 1u; for (uint32_t i = lane; i < nn; i += FLAG_LANES) acc &= P.nparent[nb + i]; return acc == 0xFFFFFFFFu; }
            for (uint32_t i = lane; i < nn; i += FLAG_LANES) if (find32(P.nparent, nb + i) == 0u) return true;
            return false;
        },
        [&](uint32_t tile, uint32_t lane, bool mine) {
            if (P.rec[tile].uniform == 2) return;
            if (lane == 0) P.rec[tile].has_region = mine ? 1 : 0;
            if (mine && P.rec[tile].uniform != 1) {
                if (lane == 0) P.flag_list[atomic_add(&P.counters[CTR_NFLAGGED], 1u)] = tile;
                return;
            }
            if (!mine && P.mask_prezeroed) return;
            int tx, ty, tz; tile_origin<C>(P, tile, tx, ty, tz);
            const int64_t x0 = (int64_t)tx * C::TX, y0 = (int64_t)ty * C::TY, z0 = (int64_t)tz * C::TZ;
            for (int it = (int)lane; it < C::TZ * C::TY * C::CX; it += FLAG_LANES) fill_chunk<C>(P, x0, y0, z0, mine ? 1u : 0u, it);
        });
}

/* the region tiles of constant rows: 2 x TZ link words in, whole rows of mask bytes out */
template <class C, class Exec>
RG_FN void rows_out_list(Exec &ex, const Params &P, RowsOutSmem<C> &s)
{
    const uint32_t n = P.counters[CTR_NROWTILES], ntiles = (uint32_t)P.ntx * (uint32_t)P.nty * (uint32_t)P.ntz;
    for (uint32_t i = ex.bid(); i < n; i += ex.nblk()) {
        const uint32_t tile = P.flag_list[ntiles - 1u - i];
        int tx, ty, tz; tile_origin<C>(P, tile, tx, ty, tz);
        tile_pass_c_rows<C>(ex, P, s, tile, P.rec[tile], (int64_t)tx * C::TX, (int64_t)ty * C::TY, (int64_t)tz * C::TZ);
        ex.sync();
    }
}

template <class C, class Exec>
RG_FN void pass_c_list(Exec &ex, const Params &P, FloodSmem<C> &s)
{
    const uint32_t n = P.counters[CTR_NFLAGGED];
    for (uint32_t i = ex.bid(); i < n; i += ex.nblk()) {
        tile_pass_c<C>(ex, P, P.flag_list[i], s);
        ex.sync();
    }
}

} // namespace rg
