/*
 * rg_rows.h -- labelling of a tile whose rows are constant ("row tiles").
 *
 * rg_fast.h already recognises a tile whose voxels are all equal (a solid's interior): one component, no labelling.  The
 * same staged bytes often have the weaker property that every ROW of TX voxels is constant -- the corridor and the walls
 * of the 1024^3 serpentine (BASELINE config C3: 7 of its 8 tile columns), layered or banded volumes, the flat faces of
 * boxes.  Such a tile is a TY x TZ grid of rows: whether two neighbouring rows are linked is one comparison of their values,
 * so the links of the whole tile are 2 x TZ words (bit y of word z: row (y, z) is similar to row (y+1, z) / (y, z+1)),
 * and labelling is a union-find over at most TY * TZ rows instead of 32 K voxels.  The cross-tile edge masks are the ordinary
 * ones (rg_fast.h build_face_masks), so neighbouring tiles on any other path see the same face lists.
 *
 * What pass C needs to flood such a tile is those 2 x TZ link words: they go to the tile's link-plane slot and the tile record
 * says uniform = 2.
 *
 * Reference semantics as everywhere (region.c:493-533, raycast.cu:281-318): this changes how a tile is labelled, not what a
 * component is.
 */
#pragma once
#include "rg_core.h"

namespace rg {

template <class C> struct RowsCfg {
    static constexpr bool OK = C::HZ && C::TY <= 32 && C::TY * C::TZ <= 1024;      /* one word of row bits per plane */
    static constexpr int NROWS = C::TY * C::TZ;
};

/* layout inside the (free) hash table: words */
template <class C> struct RowsLay {
    static constexpr int VAL = 0;                                  /* uint8 val[NROWS] */
    static constexpr int LY = VAL + (RowsCfg<C>::NROWS + 3) / 4;   /* TZ words */
    static constexpr int LZ = LY + C::TZ;
    static constexpr int OPEN = LZ + C::TZ;
    static constexpr int SEEDED = OPEN + C::TZ;
    static constexpr int NPRE = SEEDED + C::TZ;                    /* uint16 nprefix[TZ] */
    static constexpr int PAR = NPRE + (C::TZ + 1) / 2;             /* uint16 parent[NROWS] */
    static constexpr int END = PAR + (RowsCfg<C>::NROWS + 1) / 2;
};

/* S = FastSmem<C>.  On entry: tile staged, every valid row constant over the whole tile width, edge masks / prefix sums / face
   counts of this tile in place, s.scal[SC_PLANE1] = the tile's link-plane slot + 1 (non-zero). */
template <class C, class S, class Exec>
RG_FN void tile_label_rows(Exec &ex, const Params &P, uint32_t tile, S &s, int vy, int vz)
{
    typedef RowsLay<C> L;
    constexpr int NROWS = RowsCfg<C>::NROWS;
    static_assert(L::END * 4 <= (int)sizeof(s.tab), "the row state lives in the hash table's storage");
    uint32_t *t = s.tab;
    uint8_t *val = reinterpret_cast<uint8_t *>(t + L::VAL);
    uint32_t *ly = t + L::LY, *lz = t + L::LZ, *open = t + L::OPEN, *seeded = t + L::SEEDED;
    uint16_t *nprefix = reinterpret_cast<uint16_t *>(t + L::NPRE), *par = reinterpret_cast<uint16_t *>(t + L::PAR);
    const uint32_t *fmask = s.v.e.u.f.fmask;
    const uint16_t *fprefix = s.v.e.u.f.fprefix;
    const int tmode = P.tmode; const uint32_t t4 = P.t4;

    ex.pfor(NROWS, [&](int r) {
        const int y = r % C::TY, z = r / C::TY;
        val[r] = (y < vy && z < vz) ? s.u.img[img_off<C>(0, y, z)] : (uint8_t)0;
        par[r] = (uint16_t)r;
        if (r < C::TZ) { ly[r] = 0; lz[r] = 0; open[r] = 0; seeded[r] = 0; }
    });
    ex.sync();
#if RG_DEVICE_BUILD
    if (C::TY == 32) {
        /* a warp per plane: lane = row, the 32 verdicts of a plane are one ballot (32 lanes OR-ing bits into one shared word is a
           31-way conflict: with the plane words of the open bits below, half of this path's excess shared-memory wavefronts on C3) */
        const int lane = ex.tid() & 31, nwarp = Exec::NTHREADS / 32;
        for (int z = ex.tid() >> 5; z < C::TZ; z += nwarp) {
            const int r = z * C::TY + lane;
            const bool in = lane < vy && z < vz;
            const bool ey = in && lane + 1 < vy && similar1(val[r], val[r + 1], tmode, t4);
            const bool ez = in && z + 1 < vz && similar1(val[r], val[r + C::TY], tmode, t4);
            const uint32_t my = __ballot_sync(0xFFFFFFFFu, ey), mz = __ballot_sync(0xFFFFFFFFu, ez);
            if (lane == 0) { ly[z] = my; lz[z] = mz; }
        }
    } else
#endif
    ex.pfor(NROWS, [&](int r) {
        const int y = r % C::TY, z = r / C::TY;
        if (y >= vy || z >= vz) return;
        if (y + 1 < vy && similar1(val[r], val[r + 1], tmode, t4)) atomic_or(&ly[z], 1u << y);
        if (z + 1 < vz && similar1(val[r], val[r + C::TY], tmode, t4)) atomic_or(&lz[z], 1u << y);
    });
    ex.sync();
    /* Rows of a plane that are linked in y form runs, exactly like voxels of a word linked in x: a run is named by its first
       row (bit tricks, no union), and planes are joined by one union per pair of overlapping runs -- all of them at once. */
    auto run_row = [&](int z, int y) { return (uint32_t)(z * C::TY) + (run_start(ly[z], 0, y) & 31u); };
    ex.pfor(NROWS, [&](int r) {
        const int y = r % C::TY, z = r / C::TY;
        if (z + 1 >= C::TZ) return;
        const uint32_t n = dedupe(lz[z], ly[z], ly[z + 1]);
        if ((n >> y) & 1u) unite16(par, run_row(z, y), run_row(z + 1, y));
    });
    ex.sync();
    /* root of every row, flat */
    ex.pfor(NROWS, [&](int r) {
        const int y = r % C::TY, z = r / C::TY;
        const uint32_t q = find16(par, run_row(z, y));
        if (q != (uint32_t)r) par[r] = (uint16_t)q;
    });
    /* the link words are all pass C needs */
    {
        uint32_t *keep = P.planes + (size_t)(s.scal[SC_PLANE1] - 1u) * 3 * C::NW;
        ex.pfor(2 * C::TZ, [&](int i) { keep[i] = i < C::TZ ? ly[i] : lz[i - C::TZ]; });
    }
    ex.sync();
    uint16_t *ent = P.ent + (size_t)tile * C::ENT_MAX;
    ex.pfor(4 * C::FW_TOTAL, [&](int it) {
        const int fw = it >> 2, q8 = (it & 3) * 8;
        const uint32_t all = fmask[fw];
        uint32_t m = all & (0xFFu << q8);
        if (!m) return;
        const int f = (fw >= C::fw_off(1)) + (fw >= C::fw_off(2)) + (fw >= C::fw_off(3)) + (fw >= C::fw_off(4)) + (fw >= C::fw_off(5));
        const int idx = fw - C::fw_off(f);
        uint32_t slot = (uint32_t)fprefix[fw] + (uint32_t)popc(all & ((1u << q8) - 1u));
        /* neighbouring entries mostly share a root or at least its plane: their bits are collected and OR-ed into the plane's word
           once, and only if one of them is new (one atomic per entry was a 16-way conflict on every instruction) */
        uint32_t accw = 0, acc = 0;
        while (m) {
            const int b = ffs1(m) - 1; m &= m - 1;
            const uint32_t r = find16(par, face_voxel<C>(f, idx, b, vy, vz) / (uint32_t)C::TX);
            const uint32_t w = r / C::TY;
            if (w != accw) { if (acc & ~ldv32(&open[accw])) atomic_or(&open[accw], acc); accw = w; acc = 0; }
            acc |= 1u << (r % C::TY);
            ent[slot++] = (uint16_t)r;
        }
        if (acc & ~ldv32(&open[accw])) atomic_or(&open[accw], acc);
    });
    const int slo = (int)s.scal[SC_SEED_LO], shi = (int)s.scal[SC_SEED_HI];
    ex.pfor(shi - slo, [&](int i) {
        const uint32_t v = (uint32_t)(P.seed_keys[slo + i] & 0xFFFFu);
        const uint32_t r = find16(par, v / (uint32_t)C::TX);
        atomic_or(&open[r / C::TY], 1u << (r % C::TY));
        atomic_or(&seeded[r / C::TY], 1u << (r % C::TY));
    });
    ex.sync();
    const uint32_t k = ex.scan16(C::TZ, [&](int i) { return (uint32_t)popc(open[i]); }, nprefix, s.scan_scratch);
    const uint32_t base = 1u + tile * (uint32_t)FAST_NODE_SLOTS;          /* at most TY * TZ nodes: the tile's fixed block holds them */
    static_assert(NROWS <= (int)FAST_NODE_SLOTS, "a row tile's nodes fit its fixed block of node ids");
    ex.pfor(NROWS, [&](int r) {
        const uint32_t ow = open[r / C::TY], bit = 1u << (r % C::TY);
        if (!(ow & bit)) return;
        const uint32_t g = base + nprefix[r / C::TY] + (uint32_t)popc(ow & (bit - 1u));
        P.nparent[g] = (seeded[r / C::TY] & bit) ? 0u : g;
        P.nrep[g] = (uint16_t)(r * C::TX);
    });
    ex.pfor(4 * C::FW_TOTAL, [&](int it) {
        const int fw = it >> 2, q8 = (it & 3) * 8;
        const uint32_t all = fmask[fw];
        const uint32_t m = all & (0xFFu << q8);
        if (!m) return;
        uint32_t slot = (uint32_t)fprefix[fw] + (uint32_t)popc(all & ((1u << q8) - 1u));
        for (int n = popc(m); n > 0; n--, slot++) {
            const uint32_t r = ent[slot];
            ent[slot] = (uint16_t)(nprefix[r / C::TY] + (uint32_t)popc(open[r / C::TY] & ((1u << (r % C::TY)) - 1u)));
        }
    });
    if (ex.tid() == 0) {
        TileRec rc;
        rc.node_base = base; rc.nnodes = k;
        for (int f = 0; f < 6; f++) rc.cnt[f] = (uint16_t)s.scal[SC_CNT0 + f];
        rc.uniform = 2; rc.has_region = 0; rc.plane_slot1 = s.scal[SC_PLANE1]; rc.pad3 = 0;
        P.rec[tile] = rc;
        /* row tiles get their own work list (from the end of the pass C list downwards): the row output kernel decides by itself
           whether the tile holds region voxels, the flag pass skips them */
        const uint32_t ntiles = (uint32_t)P.ntx * (uint32_t)P.nty * (uint32_t)P.ntz;
        P.flag_list[ntiles - 1u - atomic_add(&P.counters[CTR_NROWTILES], 1u)] = tile;
    }
}

} // namespace rg
