/*
 * rg_dense.h -- labelling of a "complex" tile inside the CTA that has already staged it.
 *
 * A tile is complex when the bounded walkers of rg_fast.h do not get through it: a solid's surface, a
 * corridor, a background that hangs together -- every tile of the 1024^3 serpentine (BASELINE config C3).
 * Round 1 sent such a tile to a second kernel that staged it again and ran a union-find over all 32 K voxel
 * slots at two CTAs per SM (3.1 of C3's 5.4 ms, plus 1.0 ms for the fast pass's wasted attempt).  Here the
 * CTA of the fast labelling pass keeps going on the bytes it already holds:
 *
 *   1. every thread turns its share of the staged bytes into link words (bit i: voxel i is similar to its
 *      +x / +y / +z neighbour) IN REGISTERS; after a barrier the words are written over the staged bytes,
 *      which are dead by then, and to the slot pass C will read them from.  So the labelling state lives in
 *      the 48 KB the tile was staged in and the kernel stays at four CTAs per SM.
 *   2. runs (maximal x-segments of linked voxels inside one 32-voxel word) are ranked: a run gets an id only
 *      if it has a link to another row, plane or word, a cross-tile edge or a seed -- on noise that is a fifth
 *      of them -- by a prefix sum over the words.  Ids grow with the voxel index.
 *   3. lock-free union-find on 16-bit parents indexed by run id (after ECL-CC, Jaiganesh & Burtscher 2018;
 *      roots are minima, so a component's root is its smallest voxel, as everywhere else in the engine);
 *      one union per pair of overlapping runs (dedupe).
 *   4. roots of the cross-tile edges and seeds become nodes; face lists, node records and the tile record go
 *      out exactly as the other labelling passes write them (rg_core.h emit_nodes).
 *
 * A tile with more ranked runs than the parent table holds (DenseCfg::CAP, ~16 K for volumes) is handed to the
 * old full pass (rg_core.h tile_pass_a), which has a parent slot for every voxel.
 *
 * Reference semantics: region.c:493-533 / raycast.cu:281-318 (see rg_core.h); this file changes how a tile
 * is labelled, not what a component is.
 */
#pragma once
#include "rg_core.h"

namespace rg {

template <class C>
struct DenseCfg {
    /* volumes: whatever the staged tile leaves after the three link planes and the run-start plane;
       images: their staged tile is small (20 KB), so the table gets a slot for every pixel */
    static constexpr int ROOM = (C::IMG_BYTES - 16 * C::NW) / 2;
    static constexpr int CAP = C::HZ ? (ROOM < C::NV ? (ROOM > 64 ? ROOM / 8 * 8 : 64) : C::NV) : C::NV;
    static constexpr int CAPW = (CAP + 31) / 32;             /* words of a bit set over run ids */
    static_assert(CAP < (int)ENT_LEAF, "node ids in the face lists stay below the ENT_LEAF bit");
    /* once the unions are done the +y / +z planes are dead: the open / seeded bit sets and the node prefix
       (4 + 4 + 2 bytes per 32 ids) move in if they fit */
    static constexpr bool ALIAS_POST = 10 * CAPW <= 8 * C::NW;
};

template <class C>
struct DenseSmem {
    alignas(16) uint32_t ex[C::NW];
    alignas(16) uint32_t eyz[2 * C::NW];                     /* +y plane, then +z plane; after the unions: open / seeded bits, node prefix */
    alignas(16) uint32_t ls[C::NW];                          /* bit i: a ranked run starts at voxel i */
    alignas(16) uint16_t parent[DenseCfg<C>::CAP];
    uint32_t post_bits[DenseCfg<C>::ALIAS_POST ? 1 : 2 * DenseCfg<C>::CAPW];
    uint16_t post_prefix[DenseCfg<C>::ALIAS_POST ? 1 : DenseCfg<C>::CAPW];
};

/* item -> word.  Eight consecutive lanes take eight consecutive rows of the same word column: with the lean
   staging stride of 80 bytes those rows start in eight different 16-byte bank groups, so a quarter warp's
   16-byte loads do not collide; the 32 lanes of a warp still cover 32 consecutive words. */
template <class C>
RG_FN int dense_word(int i)
{
    if (C::TY % 8 != 0) return i;
    const int per_plane = C::WX * C::TY;
    const int z = i / per_plane, j = i % per_plane;
    const int yl = j & 7, wx = (j >> 3) % C::WX, yb = j / (8 * C::WX);
    return (z * C::TY + yb * 8 + yl) * C::WX + wx;
}

/* link words of word w from the staged bytes: the in-tile part of build_edges (no face masks) */
template <class C>
RG_FN void link_words(const Params &P, const uint8_t *img, int w, int64_t x0, int vy, int vz,
                      uint32_t &exw, uint32_t &eyw, uint32_t &ezw)
{
    const int wx = w % C::WX, y = (w / C::WX) % C::TY, z = w / (C::WX * C::TY);
    const int64_t gx0 = x0 + 32 * wx;
    const uint32_t vmask = (y < vy && z < vz) ? valid_bits(P.nx - gx0) : 0u;
    exw = 0; eyw = 0; ezw = 0;
    if (!vmask) return;
    const int tmode = P.tmode; const uint32_t t4 = P.t4;
    const uint8_t *rowp = &img[img_off<C>(32 * wx, y, z)];
    uint32_t a[8], b[8];
    load8(rowp, a);
    shift8(a, wx < C::WX - 1 ? (uint32_t)rowp[32] : 0u, b);
    const uint32_t nextvalid = (gx0 + 32 < P.nx) ? 0x80000000u : 0u;
    uint32_t e = edge_word(a, b, tmode, t4) & vmask & ((vmask >> 1) | nextvalid);
    if (wx == C::WX - 1) e &= 0x7FFFFFFFu;                  /* the link into the +x tile is a face entry, not a run link */
    exw = e;
    if (y < vy - 1) { load8(rowp + C::RS, b); eyw = edge_word(a, b, tmode, t4) & vmask; }
    if (C::HZ && z < vz - 1) { load8(rowp + C::PS, b); ezw = edge_word(a, b, tmode, t4) & vmask; }
}

/* id of the run that holds voxel (w, bit): ranked run starts below it */
RG_FN uint32_t run_id(uint32_t exw, uint32_t lsw, uint32_t pre, int bit)
{
    const uint32_t S = ~(exw << 1);                          /* bit 0 always starts a run */
    const int sb = 31 - clz(S & (0xFFFFFFFFu >> (31 - bit)));
    return pre + (uint32_t)popc(lsw & ((1u << sb) - 1u));
}

/* the rest of the dense path: union-find over `nruns` ranked runs on parents of type PT, nodes, face lists, tile record */
template <class C, class PT, class S, class Exec>
RG_FN void dense_label_runs(Exec &ex, const Params &P, uint32_t tile, S &s, PT *par, uint32_t nruns, int vy, int vz)
{
    typedef DenseCfg<C> D;
    DenseSmem<C> &d = s.u.d;
    uint32_t *const pey = d.eyz, *const pez = d.eyz + C::NW;
    const uint32_t *fmask = s.v.e.u.f.fmask;
    const uint16_t *fprefix = s.v.e.u.f.fprefix;
    const uint16_t *prefix = reinterpret_cast<const uint16_t *>(s.tab);
    const int slo = (int)s.scal[SC_SEED_LO], shi = (int)s.scal[SC_SEED_HI];
    /* 3. union-find over the ranked runs */
    if (sizeof(PT) == 2) ex.pfor((int)((nruns + 7u) / 8u), [&](int it) {
        const uint32_t v = (uint32_t)it * 8;
        st128(&par[v], u32x4{ v | ((v + 1) << 16), (v + 2) | ((v + 3) << 16), (v + 4) | ((v + 5) << 16), (v + 6) | ((v + 7) << 16) });
    });
    else ex.pfor((int)((nruns + 3u) / 4u), [&](int it) {
        const uint32_t v = (uint32_t)it * 4;
        st128(&par[v], u32x4{ v, v + 1, v + 2, v + 3 });
    });
    ex.sync();
    /* Order matters for speed, not for the result.  If every word united with its neighbours at once, the rows of a solid
       (or of C3's background) would all hook under the row above in the same instant: chains as long as the tile is high,
       which every later find walks.  A thread that works through SEG consecutive rows (then planes) finds the tree it has
       just extended already flat, so chains grow by one per SEG rows instead of one per row. */
#ifndef RG_DENSE_SEG
#define RG_DENSE_SEG 4
#endif
    constexpr int SEG = RG_DENSE_SEG;
    constexpr int NYSEG = (C::TY + SEG - 1) / SEG, NZSEG = (C::TZ + SEG - 1) / SEG;
    ex.pfor(C::WX * NYSEG * C::TZ, [&](int it) {
        const int wx = it % C::WX, ys = (it / C::WX) % NYSEG, z = it / (C::WX * NYSEG);
        for (int j = 0; j < SEG; j++) {
            const int y = ys * SEG + j;
            if (y >= C::TY) break;
            const int w = (z * C::TY + y) * C::WX + wx;
            const uint32_t lsw = d.ls[w];
            if (!lsw) continue;                              /* no ranked run: no link leaves this word */
            const uint32_t exw = d.ex[w], pre = prefix[w];
            if (wx < C::WX - 1 && (exw >> 31))
                unite16(par, run_id(exw, lsw, pre, 31), (uint32_t)prefix[w + 1]);   /* bit 0 of the next word starts its first run */
            const uint32_t m = pey[w];
            if (m) {
                const int up = w + C::WX; const uint32_t exu = d.ex[up], lsu = d.ls[up], preu = prefix[up];
                uint32_t n = dedupe(m, exw, exu);
                if (wx > 0 && (n & 1u)) n &= ~((pey[w - 1] & d.ex[w - 1] & d.ex[up - 1]) >> 31);   /* the pair of runs reaches over from the word before: united there */
                while (n) { const int i = ffs1(n) - 1; n &= n - 1; unite16(par, run_id(exw, lsw, pre, i), run_id(exu, lsu, preu, i)); }
            }
        }
    });
    if (C::HZ) {
        ex.sync();
        ex.pfor(C::WX * C::TY * NZSEG, [&](int it) {
            const int r = it % (C::WX * C::TY), zs = it / (C::WX * C::TY);
            for (int j = 0; j < SEG; j++) {
                const int z = zs * SEG + j;
                if (z >= C::TZ - 1) break;
                const int w = z * C::WX * C::TY + r;
                const uint32_t m = pez[w];
                if (!m) continue;
                const int up = w + C::WX * C::TY;
                const uint32_t exw = d.ex[w], lsw = d.ls[w], pre = prefix[w], exu = d.ex[up], lsu = d.ls[up], preu = prefix[up];
                uint32_t n = dedupe(m, exw, exu);
                if (r % C::WX > 0 && (n & 1u)) n &= ~((pez[w - 1] & d.ex[w - 1] & d.ex[up - 1]) >> 31);
                while (n) { const int i = ffs1(n) - 1; n &= n - 1; unite16(par, run_id(exw, lsw, pre, i), run_id(exu, lsu, preu, i)); }
            }
        });
    }
    ex.sync();
    /* flatten: every run looks its root up once, all lanes busy; the lookups of the cross-tile edges below -- a handful of
       threads with eight edges each -- then take one step instead of walking chains on their own */
    ex.pfor((int)nruns, [&](int i) { const uint32_t r = find16(par, (uint32_t)i); if (r != (uint32_t)i) par[i] = (PT)r; });
    ex.sync();

    /* 4. roots that own a cross-tile edge or a seed become nodes */
    uint32_t *open = D::ALIAS_POST ? d.eyz : d.post_bits;
    uint32_t *seeded = open + D::CAPW;
    uint16_t *nprefix = D::ALIAS_POST ? reinterpret_cast<uint16_t *>(d.eyz + 2 * D::CAPW) : d.post_prefix;
    const int idw = (int)((nruns + 31u) / 32u);
    ex.pfor(idw, [&](int i) { open[i] = 0; seeded[i] = 0; });
    ex.sync();
    uint16_t *ent = P.ent + (size_t)tile * C::ENT_MAX;
    /* a quarter of a face word per item: the root of each edge's run goes to the face list for now */
    ex.pfor(4 * C::FW_TOTAL, [&](int it) {
        const int fw = it >> 2, q8 = (it & 3) * 8;
        const uint32_t all = fmask[fw];
        uint32_t m = all & (0xFFu << q8);
        if (!m) return;
        const int f = (fw >= C::fw_off(1)) + (fw >= C::fw_off(2)) + (fw >= C::fw_off(3)) + (fw >= C::fw_off(4)) + (fw >= C::fw_off(5));
        const int idx = fw - C::fw_off(f);
        uint32_t slot = (uint32_t)fprefix[fw] + (uint32_t)popc(all & ((1u << q8) - 1u));
        while (m) {
            const int b = ffs1(m) - 1; m &= m - 1;
            const uint32_t v = face_voxel<C>(f, idx, b, vy, vz);
            const int w = (int)(v >> 5);
            const uint32_t r = find16(par, run_id(d.ex[w], d.ls[w], prefix[w], (int)(v & 31)));
            atomic_or(&open[r >> 5], 1u << (r & 31));
            ent[slot++] = (uint16_t)r;
        }
    });
    ex.pfor(shi - slo, [&](int i) {
        const uint32_t v = (uint32_t)(P.seed_keys[slo + i] & 0xFFFFu);
        const int w = (int)(v >> 5);
        const uint32_t r = find16(par, run_id(d.ex[w], d.ls[w], prefix[w], (int)(v & 31)));
        atomic_or(&open[r >> 5], 1u << (r & 31));
        atomic_or(&seeded[r >> 5], 1u << (r & 31));
    });
    ex.sync();
    const uint32_t k = ex.scan16(idw, [&](int i) { return (uint32_t)popc(open[i]); }, nprefix, s.scan_scratch);
    if (ex.tid() == 0) {
        /* the tile's fixed block of node ids if it is enough (no global atomic), else ids from behind the blocks */
        uint32_t base = 1u + tile * (uint32_t)FAST_NODE_SLOTS, bad = 0;
        if (k > (uint32_t)FAST_NODE_SLOTS) {
            base = atomic_add(&P.counters[CTR_NODES], k);
            if (base + k > P.node_cap || base + k < base) { atomic_or(&P.counters[CTR_STATUS], ST_NODE_OVERFLOW); bad = 1; }
        }
        s.scal[SC_NODE_BASE] = base; s.scal[SC_OVERFLOW] = bad;
    }
    ex.sync();
    const uint32_t base = s.scal[SC_NODE_BASE];
    const bool ok = !s.scal[SC_OVERFLOW];
    if (ok) {
        ex.pfor(C::NW, [&](int w) {
            uint32_t m = d.ls[w];
            if (!m) return;
            uint32_t id = prefix[w];
            while (m) {
                const int b = ffs1(m) - 1; m &= m - 1;
                const uint32_t ow = open[id >> 5], bit = 1u << (id & 31);
                if (ow & bit) {
                    const uint32_t g = base + nprefix[id >> 5] + (uint32_t)popc(ow & (bit - 1u));
                    P.nparent[g] = (seeded[id >> 5] & bit) ? 0u : g;       /* seeded components start hooked to node 0 */
                    P.nrep[g] = (uint16_t)(w * 32 + b);
                }
                id++;
            }
        });
        /* root ids -> node ids, each thread over the entries it wrote itself */
        ex.pfor(4 * C::FW_TOTAL, [&](int it) {
            const int fw = it >> 2, q8 = (it & 3) * 8;
            const uint32_t all = fmask[fw];
            const uint32_t m = all & (0xFFu << q8);
            if (!m) return;
            uint32_t slot = (uint32_t)fprefix[fw] + (uint32_t)popc(all & ((1u << q8) - 1u));
            for (int n = popc(m); n > 0; n--, slot++) {
                const uint32_t r = ent[slot];
                ent[slot] = (uint16_t)(nprefix[r >> 5] + (uint32_t)popc(open[r >> 5] & ((1u << (r & 31)) - 1u)));
            }
        });
    }
    if (ex.tid() == 0) {
        TileRec rc;
        rc.node_base = base; rc.nnodes = ok ? k : 0u;
        for (int f = 0; f < 6; f++) rc.cnt[f] = ok ? (uint16_t)s.scal[SC_CNT0 + f] : (uint16_t)0;
        rc.uniform = 0; rc.has_region = 0; rc.plane_slot1 = ok ? s.scal[SC_PLANE1] : 0u; rc.pad3 = 0;
        P.rec[tile] = rc;
        atomic_add(&P.counters[CTR_NDENSE], 1u);
    }
}

/* S = FastSmem<C> (rg_fast.h).  On entry the tile is staged in s.u.img and the cross-tile edge masks, their
   prefix sums (s.v.e.u.f) and the face counts (s.scal[SC_CNT0..]) are those of this tile.  Returns false if the
   tile has more ranked runs than the table holds: the caller hands it to the full pass. */
template <class C, class S, class Exec>
RG_FN bool tile_label_dense(Exec &ex, const Params &P, uint32_t tile, S &s, int64_t x0, int vy, int vz)
{
    typedef DenseCfg<C> D;
    DenseSmem<C> &d = s.u.d;
    uint32_t *const pey = d.eyz, *const pez = d.eyz + C::NW;
    const uint32_t *fmask = s.v.e.u.f.fmask;
    uint16_t *prefix = reinterpret_cast<uint16_t *>(s.tab);  /* ranked runs before word w (the hash table is not needed here) */
    static_assert(sizeof(s.tab) >= 2 * C::NW, "the run prefix sums live in the hash table's storage");

    /* 1. link words in registers, while the staged bytes are still there */
    constexpr int NPT = (C::NW + Exec::NTHREADS - 1) / Exec::NTHREADS;
    uint32_t rx[NPT], ry[NPT], rz[NPT];
    ex.pfor_k(C::NW, [&](int i, int k) { link_words<C>(P, s.u.img, dense_word<C>(i), x0, vy, vz, rx[k], ry[k], rz[k]); });
    if (ex.tid() == 0) {
        s.scal[SC_PLANE1] = plane_slot1(P, tile);
    }
    ex.sync();                                               /* every thread is done with the staged bytes */
    {
        const uint32_t p1 = s.scal[SC_PLANE1];
        uint32_t *keep = p1 ? P.planes + (size_t)(p1 - 1u) * 3 * C::NW : (uint32_t *)0;
        ex.pfor_k(C::NW, [&](int i, int k) {
            const int w = dense_word<C>(i);
            d.ex[w] = rx[k]; pey[w] = ry[k]; pez[w] = rz[k];
            if (keep) { keep[w] = rx[k]; keep[C::NW + w] = ry[k]; keep[2 * C::NW + w] = rz[k]; }   /* pass C floods from these */
        });
    }
    ex.sync();

    /* 2. which voxels matter: any link out of the row, out of the word, out of the tile -- or a seed */
    ex.pfor(C::NW, [&](int w) {
        const int wx = w % C::WX, y = (w / C::WX) % C::TY, z = w / (C::WX * C::TY);
        const uint32_t exw = d.ex[w];
        uint32_t l = pey[w] | pez[w];
        if (y > 0) l |= pey[w - C::WX];
        if (C::HZ && z > 0) l |= pez[w - C::WX * C::TY];
        if (wx < C::WX - 1) l |= exw & 0x80000000u;
        if (wx > 0) l |= d.ex[w - 1] >> 31;
        if (y == 0) l |= fmask[C::fw_off(F_YM) + z * C::WX + wx];
        if (y == vy - 1) l |= fmask[C::fw_off(F_YP) + z * C::WX + wx];
        if (C::HZ) {
            if (z == 0) l |= fmask[C::fw_off(F_ZM) + y * C::WX + wx];
            if (z == vz - 1) l |= fmask[C::fw_off(F_ZP) + y * C::WX + wx];
        }
        const int q = z * C::TY + y;
        if (wx == 0) l |= (fmask[C::fw_off(F_XM) + (q >> 5)] >> (q & 31)) & 1u;
        if (wx == C::WX - 1) l |= ((fmask[C::fw_off(F_XP) + (q >> 5)] >> (q & 31)) & 1u) << 31;
        d.ls[w] = l;
    });
    const int slo = (int)s.scal[SC_SEED_LO], shi = (int)s.scal[SC_SEED_HI];
    if (shi > slo) {
        ex.sync();
        ex.pfor(shi - slo, [&](int i) {
            const uint32_t v = (uint32_t)(P.seed_keys[slo + i] & 0xFFFFu);
            atomic_or(&d.ls[v >> 5], 1u << (v & 31));
        });
    }
    ex.sync();
    /* ... reduced to the first voxel of their runs; ranks by a prefix sum over the words */
    ex.pfor(C::NW, [&](int w) {
        const uint32_t l = d.ls[w];
        if (l) { const uint32_t exw = d.ex[w]; d.ls[w] = fill_word(l, exw) & ~(exw << 1); }
    });
    ex.sync();
    const uint32_t nruns = ex.scan16(C::NW, [&](int i) { return (uint32_t)popc(d.ls[i]); }, prefix, s.scan_scratch);
    if (nruns > (uint32_t)D::CAP) return false;               /* same total in every thread */

    /* 3./4. in dense_label_runs.  Up to CAP / 2 runs -- the usual case -- the parents are 32-bit words: shared-memory
       compare-and-swap is native at that width, while the 16-bit form is a retry loop on the enclosing word in which
       neighbouring runs disturb each other */
    if (nruns <= (uint32_t)D::CAP / 2) dense_label_runs<C>(ex, P, tile, s, reinterpret_cast<uint32_t *>(d.parent), nruns, vy, vz);
    else dense_label_runs<C>(ex, P, tile, s, d.parent, nruns, vy, vz);
    return true;
}

} // namespace rg
