/*
 * tests/emu/rg_emu.cpp -- sequential walk of the engine's kernel bodies.  TEST CODE ONLY.
 *
 * Compiles parallel-c-programs_b200/csrc/rg_core.h with g++ (RG_FN = inline, plain memory ops
 * instead of atomics) and runs pass A / B / C one "thread" at a time, one tile after another.
 * It lets the `not gpu` tests check the engine's LOGIC (edge planes, face-list alignment,
 * node bookkeeping, slab interfaces) against the oracle without a GPU, including with tiny tile
 * shapes that put many tile faces inside small random grids.  It is not part of libb200rg.so
 * and nothing in the product can reach it.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <algorithm>
#include <vector>
#include "../../parallel-c-programs_b200/csrc/rg_core.h"
#include "../../parallel-c-programs_b200/csrc/rg_fast.h"
#include "../../parallel-c-programs_b200/csrc/rg_slab_core.h"

using namespace rg;

struct HostExec {
    static constexpr int NTHREADS = 1;
    int tid() const { return 0; }
    template <class F> void pfor(int n, F f) { for (int i = 0; i < n; i++) f(i); }
    template <class F> void pfor_k(int n, F f) { for (int i = 0; i < n; i++) f(i, i); }
    template <class F> void gpfor(int64_t n, F f) { for (int64_t i = 0; i < n; i++) f(i); }
    template <class Pr, class Bo> void wfor(uint32_t n, Pr pred, Bo body)
    {
        for (uint32_t item = 0; item < n; item++) {
            bool any = false;
            for (uint32_t lane = 0; lane < 32; lane++) any = pred(item, lane) || any;
            for (uint32_t lane = 0; lane < 32; lane++) body(item, lane, any);
        }
    }
    void sync() {}
    void sync_async_smem() {}
    template <class C> bool tma_clear_tile(const Params &, const void *, int, int64_t, int64_t, int64_t) { return false; }
    bool sync_or(bool p) { return p; }
    template <class C> bool tma_tile(const Params &, uint8_t *, uint64_t *, int64_t, int64_t, int64_t) { return false; }
    void tile_wait(uint64_t *) {}
    template <class C> void origin(const Params &P, uint32_t tile, int &tx, int &ty, int &tz) const { tile_origin<C>(P, tile, tx, ty, tz); }
    uint32_t bid() const { return 0; }
    uint32_t nblk() const { return 1; }
    template <class F> uint32_t scan16(int n, F count, uint16_t *out, uint32_t *)
    {
        uint32_t run = 0;
        for (int i = 0; i < n; i++) { out[i] = (uint16_t)run; run += count(i); }
        return run;
    }
    template <class F> uint32_t scan32(int n, F count, uint32_t *out, uint32_t *)
    {
        uint32_t run = 0;
        for (int i = 0; i < n; i++) { out[i] = run; run += count(i); }
        return run;
    }
};

static int g_force_full = 0;
static int g_no_planes = 0;      /* 1: label every tile with the full pass (fast pass off) */
static int g_force_dense = 0;    /* 1: every non-uniform tile takes the dense path of the fast pass */
static int g_no_rows = 0;        /* 1: tiles of constant rows are labelled like any other tile */
static uint32_t g_last_dense = 0, g_last_full = 0, g_last_rows = 0;
static int g_no_prezero = 0;     /* 1: the labelling pass leaves the mask alone (the library's RG_NO_PREZERO): the later passes write every byte */
extern "C" void emu_no_prezero(int v) { g_no_prezero = v; }
extern "C" void emu_no_rows(int v) { g_no_rows = v; }
extern "C" uint32_t emu_last_rows(void) { return g_last_rows; }
extern "C" void emu_force_dense(int v) { g_force_dense = v; }
extern "C" uint32_t emu_last_dense(void) { return g_last_dense; }
extern "C" uint32_t emu_last_full(void) { return g_last_full; }
extern "C" void emu_force_full(int v) { g_force_full = v; }
extern "C" void emu_no_planes(int v) { g_no_planes = v; }

struct EmuState {
    std::vector<TileRec> rec; std::vector<uint16_t> ent; std::vector<uint32_t> npar; std::vector<uint16_t> nrep;
    std::vector<uint64_t> keys; std::vector<uint32_t> clist, flist, planes; uint32_t counters[CTR_WORDS];
    Params P;
};

template <class C>
static void emu_setup(EmuState &E, const uint8_t *image, int64_t nx, int64_t ny, int64_t nz, int zlo, int zhi,
                      const int32_t *seeds, int nseeds, int T, uint8_t *mask)
{
    Params &P = E.P; memset(&P, 0, sizeof P);
    P.image = image; P.mask = mask; P.nx = nx; P.ny = ny; P.nz = nz;
    P.ntx = (int)((nx + C::TX - 1) / C::TX); P.nty = (int)((ny + C::TY - 1) / C::TY); P.ntz = (int)((nz + C::TZ - 1) / C::TZ);
    P.lo_ext = zlo; P.hi_ext = zhi;
    P.tmode = T <= 0 ? 0 : T >= 256 ? 2 : 1; P.t4 = (uint32_t)(T & 0xFF) * 0x01010101u; P.thr = T <= 0 ? 0u : T >= 256 ? 256u : (uint32_t)T;
    P.mask_prezeroed = (g_force_full || g_no_prezero) ? 0 : 1;
    P.force_dense = g_force_dense;
    P.no_rows = g_no_rows;
    P.fast16 = (nx % 16 == 0) && (((uintptr_t)image & 15) == 0) && (((uintptr_t)mask & 15) == 0);
    size_t ntiles = (size_t)P.ntx * P.nty * P.ntz;
    size_t cap = 1 + ntiles * ((size_t)C::ENT_MAX + FAST_NODE_SLOTS) + (size_t)nseeds;
    E.rec.assign(ntiles, TileRec()); E.ent.assign(ntiles * C::ENT_MAX, 0); E.npar.assign(cap, 0xDEADBEEF); E.nrep.assign(cap, 0);
    E.keys.assign((size_t)std::max(nseeds, 1), ~0ull);
    E.clist.assign(ntiles, 0); E.flist.assign(ntiles, 0);
    P.complex_list = E.clist.data(); P.flag_list = E.flist.data();
    /* link planes kept by the full pass for pass C: room for every second tile, so that both branches of pass C run */
    P.plane_cap = g_no_planes ? 0u : (uint32_t)((ntiles + 1) / 2);
    E.planes.assign((size_t)P.plane_cap * 3 * C::NW + 4, 0xDEADBEEFu);
    P.planes = P.plane_cap ? E.planes.data() : nullptr;
    P.rec = E.rec.data(); P.ent = E.ent.data(); P.nparent = E.npar.data(); P.nrep = E.nrep.data();
    P.counters = E.counters; P.node_cap = (uint32_t)cap; P.nseeds = nseeds; P.seed_keys = E.keys.data();
    HostExec ex;
    reset_counters(P);
    make_seed_keys<C>(ex, P, seeds, E.keys.data());
    std::sort(E.keys.begin(), E.keys.begin() + nseeds);
}

template <class C> static void emu_label(EmuState &E)
{
    HostExec ex; const Params &P = E.P;
    uint32_t ntiles = (uint32_t)P.ntx * P.nty * P.ntz;
    TileSmem<C> *s = (TileSmem<C> *)aligned_alloc(128, (sizeof(TileSmem<C>) + 127) / 128 * 128);
    typedef typename LeanOf<C>::type LC;                             /* the fast pass stages in the lean layout, as the kernel does */
    FastSmem<LC> *f = (FastSmem<LC> *)aligned_alloc(128, (sizeof(FastSmem<LC>) + 127) / 128 * 128);
    if (g_force_full) { for (uint32_t t = 0; t < ntiles; t++) E.clist[t] = t; E.counters[CTR_NCOMPLEX] = ntiles; }
    else for (uint32_t t = 0; t < ntiles; t++) { memset(f, 0xA5, sizeof *f); tile_pass_a_fast<LC>(ex, P, t, *f); }
    memset(s, 0xA5, sizeof *s);
    pass_a_full_list<C>(ex, P, *s);
    merge_faces<C>(ex, P);
    free(s); free(f);
}
template <class C> static void emu_finish(EmuState &E)
{
    HostExec ex; const Params &P = E.P;
    typedef typename LeanOf<C>::type LC;                             /* pass C stages in the lean layout, as the kernel does */
    FloodSmem<LC> *s = (FloodSmem<LC> *)aligned_alloc(128, (sizeof(FloodSmem<LC>) + 127) / 128 * 128);
    flag_tiles_all<C>(ex, P);
    RowsOutSmem<LC> rs; memset(&rs, 0x5A, sizeof rs);
    rows_out_list<LC>(ex, P, rs);
    memset(s, 0x5A, sizeof *s);
    pass_c_list<LC>(ex, P, *s);
    free(s);
}

template <class C>
static int emu_grow_t(const uint8_t *image, int64_t nx, int64_t ny, int64_t nz, const int32_t *seeds, int nseeds, int T,
                      uint8_t *mask, uint32_t *stats)
{
    EmuState E;
    emu_setup<C>(E, image, nx, ny, nz, 0, 0, seeds, nseeds, T, mask);
    emu_label<C>(E);
    emu_finish<C>(E);
    uint32_t real_nodes = 0; for (const TileRec &r : E.rec) real_nodes += r.nnodes;
    if (stats) { stats[0] = real_nodes; stats[1] = E.counters[CTR_STATUS]; stats[2] = E.counters[CTR_NCOMPLEX] + E.counters[CTR_NDENSE] + E.counters[CTR_NROWTILES]; stats[3] = (uint32_t)E.rec.size(); }
    g_last_dense = E.counters[CTR_NDENSE]; g_last_full = E.counters[CTR_NCOMPLEX]; g_last_rows = E.counters[CTR_NROWTILES];
    return E.counters[CTR_STATUS] ? -4 : 0;
}

typedef TileCfg<32, 4, 4> CfgS3;      /* tiny tiles: many faces inside small grids */
typedef TileCfg<64, 8, 2> CfgT3;
typedef TileCfg<32, 8, 1> CfgS2;

/* cfg: 0 = product shapes (Cfg3D / Cfg2D by nz), 1 = tiny shapes, 2 = another 3-D shape */
extern "C" int emu_grow(int cfg, const uint8_t *image, int64_t nx, int64_t ny, int64_t nz, const int32_t *seeds, int nseeds,
                        int T, uint8_t *mask, uint32_t *stats)
{
    if (nz == 1) return cfg == 0 ? emu_grow_t<Cfg2D>(image, nx, ny, nz, seeds, nseeds, T, mask, stats)
                                 : emu_grow_t<CfgS2>(image, nx, ny, nz, seeds, nseeds, T, mask, stats);
    if (cfg == 0) return emu_grow_t<Cfg3D>(image, nx, ny, nz, seeds, nseeds, T, mask, stats);
    if (cfg == 1) return emu_grow_t<CfgS3>(image, nx, ny, nz, seeds, nseeds, T, mask, stats);
    return emu_grow_t<CfgT3>(image, nx, ny, nz, seeds, nseeds, T, mask, stats);
}

/* ---- multi-slab emulation: G logical ranks on one host, the same protocol as rgb200.SlabGrower ---- */
template <class C>
static int emu_grow_slabs_t(int G, const uint8_t *image, int64_t nx, int64_t ny, int64_t nz, const int32_t *seeds, int nseeds,
                            int T, uint8_t *mask, uint32_t *stats)
{
    std::vector<EmuState> E((size_t)G);
    std::vector<int64_t> zb((size_t)G + 1);
    /* volumes are cut along z, images (tiles without a z halo) along y */
    const bool rows = (C::HZ == 0);
    const int64_t na = rows ? ny : nz, unit = rows ? nx : nx * ny;
    const int ci = rows ? 1 : 2;
    for (int g = 0; g <= G; g++) zb[(size_t)g] = na * g / G;
    std::vector<std::vector<int32_t>> sl((size_t)G);
    HostExec ex;
    /* each rank: local seeds (shifted to slab coordinates), pass A + B on its slab */
    for (int g = 0; g < G; g++) {
        int64_t z0 = zb[(size_t)g], z1 = zb[(size_t)g + 1];
        if (z1 <= z0) continue;
        for (int i = 0; i < nseeds; i++) {
            int32_t t[3] = { seeds[3 * i], seeds[3 * i + 1], seeds[3 * i + 2] };
            if (t[ci] >= z0 && t[ci] < z1) { t[ci] = (int32_t)(t[ci] - z0); sl[(size_t)g].insert(sl[(size_t)g].end(), t, t + 3); }
        }
        int zlo = 0, zhi = 0;
        for (int h = 0; h < g; h++) if (zb[(size_t)h + 1] > zb[(size_t)h]) zlo = 1;
        for (int h = g + 1; h < G; h++) if (zb[(size_t)h + 1] > zb[(size_t)h]) zhi = 1;
        emu_setup<C>(E[(size_t)g], image + z0 * unit, nx, rows ? z1 - z0 : ny, rows ? 1 : z1 - z0, zlo, zhi, sl[(size_t)g].data(),
                     (int)sl[(size_t)g].size() / 3, T, mask + z0 * unit);
        emu_label<C>(E[(size_t)g]);
    }
    /* export interface lists (roots), "all-gather", resolve redundantly, then pass C */
    std::vector<std::vector<uint32_t>> lo((size_t)G), hi((size_t)G);
    std::vector<int> live;
    for (int g = 0; g < G; g++) if (zb[(size_t)g + 1] > zb[(size_t)g]) live.push_back(g);
    for (int g : live) {
        const Params &P = E[(size_t)g].P;
        size_t cap = (size_t)iface_tiles<C>(P) * (C::HZ ? C::FZN : C::FYN) + 1;
        lo[(size_t)g].assign(cap, 0); hi[(size_t)g].assign(cap, 0);
        uint32_t nlo = 0, nhi = 0;
        if (P.lo_ext) export_interface<C>(ex, P, 0, lo[(size_t)g].data(), &nlo);
        if (P.hi_ext) export_interface<C>(ex, P, 1, hi[(size_t)g].data(), &nhi);
        lo[(size_t)g].resize(nlo); hi[(size_t)g].resize(nhi);
    }
    int rc = 0;
    for (size_t li = 0; li + 1 < live.size(); li++)
        if (hi[(size_t)live[li]].size() != lo[(size_t)live[li + 1]].size()) rc = -4;
    if (!rc) {
        /* global vertex ids: offset[g] + local root; vertex 0 of every rank is the shared "seeded" vertex */
        std::vector<uint32_t> off(live.size() + 1, 1);
        for (size_t li = 0; li < live.size(); li++) off[li + 1] = off[li] + E[(size_t)live[li]].counters[CTR_NODES];
        std::vector<uint32_t> xpar(off.back(), 0);
        for (size_t li = 0; li + 1 < live.size(); li++) {
            const std::vector<uint32_t> &a = hi[(size_t)live[li]], &b = lo[(size_t)live[li + 1]];
            xpar[0] = 0;
            resolve_init(ex, xpar.data(), a.data(), (uint32_t)a.size(), off[li]);
            resolve_init(ex, xpar.data(), b.data(), (uint32_t)b.size(), off[li + 1]);
        }
        for (size_t li = 0; li + 1 < live.size(); li++) {
            const std::vector<uint32_t> &a = hi[(size_t)live[li]], &b = lo[(size_t)live[li + 1]];
            resolve_union(ex, xpar.data(), a.data(), off[li], b.data(), off[li + 1], (uint32_t)a.size());
        }
        for (size_t li = 0; li < live.size(); li++) {
            int g = live[li];
            resolve_apply(ex, E[(size_t)g].P, xpar.data(), lo[(size_t)g].data(), (uint32_t)lo[(size_t)g].size(), off[li]);
            resolve_apply(ex, E[(size_t)g].P, xpar.data(), hi[(size_t)g].data(), (uint32_t)hi[(size_t)g].size(), off[li]);
        }
    }
    uint32_t st = 0, nodes = 0;
    for (int g : live) { emu_finish<C>(E[(size_t)g]); st |= E[(size_t)g].counters[CTR_STATUS]; nodes += E[(size_t)g].counters[CTR_NODES]; }
    if (stats) { stats[0] = nodes; stats[1] = st; stats[2] = 0; stats[3] = (uint32_t)live.size(); }
    return (rc || st) ? -4 : 0;
}

extern "C" int emu_grow_slabs(int cfg, int G, const uint8_t *image, int64_t nx, int64_t ny, int64_t nz, const int32_t *seeds, int nseeds,
                              int T, uint8_t *mask, uint32_t *stats)
{
    if (G < 1) return -1;
    if (nz == 1) return cfg == 0 ? emu_grow_slabs_t<Cfg2D>(G, image, nx, ny, nz, seeds, nseeds, T, mask, stats)
                                 : emu_grow_slabs_t<CfgS2>(G, image, nx, ny, nz, seeds, nseeds, T, mask, stats);
    if (cfg == 0) return emu_grow_slabs_t<Cfg3D>(G, image, nx, ny, nz, seeds, nseeds, T, mask, stats);
    if (cfg == 1) return emu_grow_slabs_t<CfgS3>(G, image, nx, ny, nz, seeds, nseeds, T, mask, stats);
    return emu_grow_slabs_t<CfgT3>(G, image, nx, ny, nz, seeds, nseeds, T, mask, stats);
}

/* ---- an image over a rows x columns grid of blocks: the protocol of rg_grow_multi_grid on host memory ---- */
template <class C>
static int emu_grow_grid_t(int R, int Cn, const uint8_t *image, int64_t w, int64_t h, const int32_t *seeds, int nseeds, int T, uint8_t *mask, uint32_t *stats)
{
    while (R > 1 && h / R < 1) R--;
    while (Cn > 1 && (w / Cn) / C::TX < 1) Cn--;
    std::vector<int64_t> xc((size_t)Cn + 1), yc((size_t)R + 1);
    for (int j = 0; j <= Cn; j++) xc[(size_t)j] = j == Cn ? w : (w * j / Cn) / C::TX * C::TX;
    for (int i = 0; i <= R; i++) yc[(size_t)i] = h * i / R;
    const int B = R * Cn;
    std::vector<EmuState> E((size_t)B);
    std::vector<std::vector<uint8_t>> img((size_t)B), msk((size_t)B);
    std::vector<std::vector<int32_t>> sl((size_t)B);
    std::vector<std::vector<uint32_t>> roots((size_t)B * 4);
    std::vector<std::vector<uint8_t>> vals((size_t)B * 4);
    HostExec ex;
    const uint32_t thr = T <= 0 ? 0u : T >= 256 ? 256u : (uint32_t)T;
    for (int b = 0; b < B; b++) {
        const int i = b / Cn, j = b % Cn;
        const int64_t x0 = xc[(size_t)j], x1 = xc[(size_t)j + 1], y0 = yc[(size_t)i], y1 = yc[(size_t)i + 1], bw = x1 - x0, bh = y1 - y0;
        img[(size_t)b].resize((size_t)(bw * bh) + 64); msk[(size_t)b].assign((size_t)(bw * bh), 7);
        uint8_t *base = (uint8_t *)(((uintptr_t)img[(size_t)b].data() + 15) & ~(uintptr_t)15);
        for (int64_t y = 0; y < bh; y++) memcpy(base + y * bw, image + (y0 + y) * w + x0, (size_t)bw);
        for (int k = 0; k < nseeds; k++) {
            const int64_t sx = seeds[3 * k], sy = seeds[3 * k + 1];
            if (sx >= x0 && sx < x1 && sy >= y0 && sy < y1 && seeds[3 * k + 2] == 0) { sl[(size_t)b].push_back((int32_t)(sx - x0)); sl[(size_t)b].push_back((int32_t)(sy - y0)); sl[(size_t)b].push_back(0); }
        }
        emu_setup<C>(E[(size_t)b], base, bw, bh, 1, 0, 0, sl[(size_t)b].data(), (int)sl[(size_t)b].size() / 3, T, msk[(size_t)b].data());
        E[(size_t)b].P.open = (j > 0 ? 1u : 0u) | (j < Cn - 1 ? 2u : 0u) | (i > 0 ? 4u : 0u) | (i < R - 1 ? 8u : 0u);
        emu_label<C>(E[(size_t)b]);
        for (int side = 0; side < 4; side++)
            if ((E[(size_t)b].P.open >> side) & 1u) {
                const size_t n = (size_t)(side < 2 ? bh : bw);
                roots[(size_t)b * 4 + side].assign(n, 0); vals[(size_t)b * 4 + side].assign(n, 0);
                export_border<C>(ex, E[(size_t)b].P, side, roots[(size_t)b * 4 + side].data(), vals[(size_t)b * 4 + side].data());
            }
    }
    std::vector<uint32_t> off((size_t)B + 1, 1u);
    for (int b = 0; b < B; b++) off[(size_t)b + 1] = off[(size_t)b] + E[(size_t)b].counters[CTR_NODES];
    std::vector<uint32_t> xpar(off[(size_t)B] + 2, 0);
    for (int b = 0; b < B; b++)
        for (int side = 0; side < 4; side++) { auto &r = roots[(size_t)b * 4 + side]; if (!r.empty()) resolve_init(ex, xpar.data(), r.data(), (uint32_t)r.size(), off[(size_t)b]); }
    xpar[0] = 0;
    for (int b = 0; b < B; b++) {
        const int i = b / Cn, j = b % Cn;
        if (j < Cn - 1) { auto &a = roots[(size_t)b * 4 + 1]; auto &c2 = roots[(size_t)(b + 1) * 4 + 0];
            resolve_union_seam(ex, xpar.data(), a.data(), off[(size_t)b], vals[(size_t)b * 4 + 1].data(), c2.data(), off[(size_t)b + 1], vals[(size_t)(b + 1) * 4].data(), (uint32_t)a.size(), thr); }
        if (i < R - 1) { auto &a = roots[(size_t)b * 4 + 3]; auto &c2 = roots[(size_t)(b + Cn) * 4 + 2];
            resolve_union_seam(ex, xpar.data(), a.data(), off[(size_t)b], vals[(size_t)b * 4 + 3].data(), c2.data(), off[(size_t)b + Cn], vals[(size_t)(b + Cn) * 4 + 2].data(), (uint32_t)a.size(), thr); }
    }
    uint32_t st = 0;
    for (int b = 0; b < B; b++) {
        for (int side = 0; side < 4; side++) { auto &r = roots[(size_t)b * 4 + side]; if (!r.empty()) resolve_apply(ex, E[(size_t)b].P, xpar.data(), r.data(), (uint32_t)r.size(), off[(size_t)b]); }
        emu_finish<C>(E[(size_t)b]);
        st |= E[(size_t)b].counters[CTR_STATUS];
        const int i = b / Cn, j = b % Cn;
        const int64_t x0 = xc[(size_t)j], bw = xc[(size_t)j + 1] - x0, y0 = yc[(size_t)i], bh = yc[(size_t)i + 1] - y0;
        for (int64_t y = 0; y < bh; y++) memcpy(mask + (y0 + y) * w + x0, msk[(size_t)b].data() + y * bw, (size_t)bw);
    }
    if (stats) { stats[0] = off[(size_t)B]; stats[1] = st; stats[2] = (uint32_t)R; stats[3] = (uint32_t)Cn; }
    return st ? -4 : 0;
}
extern "C" int emu_grow_grid(int cfg, int R, int Cn, const uint8_t *image, int64_t w, int64_t h, const int32_t *seeds, int nseeds, int T, uint8_t *mask, uint32_t *stats)
{
    return cfg == 0 ? emu_grow_grid_t<Cfg2D>(R, Cn, image, w, h, seeds, nseeds, T, mask, stats)
                    : emu_grow_grid_t<CfgS2>(R, Cn, image, w, h, seeds, nseeds, T, mask, stats);
}

/* ---- the slab phases of include/rg.h on HOST memory, same names and signatures, so that the
 * `not gpu` tests can drive rgb200.SlabGrower over the gloo backend with CPU tensors.  TEST ONLY. ---- */
static int g_cfg = 1;
extern "C" void emu_set_cfg(int cfg) { g_cfg = cfg; }
struct EmuCtx { EmuState E; bool have; int cfg; std::vector<int32_t> seeds; };

/* 0 / 1 / 2 = the three volume shapes; 3 / 4 = the two image shapes (chosen by emu_cfg_for when nz_local == 1) */
#define DISPATCH(cfg, CALL) do { if ((cfg) == 0) { typedef Cfg3D C; CALL; } else if ((cfg) == 1) { typedef CfgS3 C; CALL; } \
    else if ((cfg) == 2) { typedef CfgT3 C; CALL; } else if ((cfg) == 3) { typedef Cfg2D C; CALL; } else { typedef CfgS2 C; CALL; } } while (0)
static int emu_cfg_for(int64_t nz_local) { return nz_local == 1 ? (g_cfg == 0 ? 3 : 4) : g_cfg; }

extern "C" int rg_ctx_create(int, void **out) { *out = new EmuCtx(); ((EmuCtx *)*out)->have = false; return 0; }
extern "C" void rg_ctx_destroy(void *c) { delete (EmuCtx *)c; }
extern "C" const char *rg_last_error(void) { return "emulator"; }
extern "C" int rg_slab_label(void *cv, const uint8_t *slab, int64_t nx, int64_t ny, int64_t nzl, int has_lower, int has_upper,
                             const int32_t *seeds, int nseeds, int T, uint8_t *mask, void *)
{
    EmuCtx *c = (EmuCtx *)cv; c->cfg = emu_cfg_for(nzl);
    c->seeds.assign(seeds, seeds + 3 * (size_t)nseeds);
    DISPATCH(c->cfg, (emu_setup<C>(c->E, slab, nx, ny, nzl, has_lower, has_upper, c->seeds.data(), nseeds, T, mask), emu_label<C>(c->E)));
    c->have = true;
    return 0;
}
extern "C" int64_t rg_slab_interface_capacity(int64_t nx, int64_t ny, int64_t nz_local)
{
    int64_t r = 0;
    DISPATCH(emu_cfg_for(nz_local), (r = C::HZ ? ((nx + C::TX - 1) / C::TX) * ((ny + C::TY - 1) / C::TY) * (int64_t)C::FZN
                               : ((nx + C::TX - 1) / C::TX) * (int64_t)C::FYN));
    return r;
}
extern "C" int rg_slab_export(void *cv, int side, uint32_t *roots, uint32_t *counts3, void *)
{
    EmuCtx *c = (EmuCtx *)cv; HostExec ex;
    DISPATCH(c->cfg, (export_interface<C>(ex, c->E.P, side, roots, counts3 + side)));
    counts3[2] = c->E.counters[CTR_NODES];
    return 0;
}
extern "C" int rg_slab_resolve_init(void *, uint32_t *xpar, const uint32_t *roots, uint32_t n, uint32_t off, void *)
{
    HostExec ex; xpar[0] = 0; resolve_init(ex, xpar, roots, n, off); return 0;
}
extern "C" int rg_slab_resolve_union(void *, uint32_t *xpar, const uint32_t *a, uint32_t offa, const uint32_t *b, uint32_t offb, uint32_t n, void *)
{
    HostExec ex; resolve_union(ex, xpar, a, offa, b, offb, n); return 0;
}
extern "C" int rg_slab_resolve_apply(void *cv, uint32_t *xpar, const uint32_t *roots, uint32_t n, uint32_t off, void *)
{
    EmuCtx *c = (EmuCtx *)cv; HostExec ex; resolve_apply(ex, c->E.P, xpar, roots, n, off); return 0;
}
extern "C" int rg_slab_resolve(void *cv, uint32_t *xpar, const uint32_t *lists, uint32_t stride, const int64_t *counts4, int rank, int world, void *)
{
    EmuCtx *c = (EmuCtx *)cv; HostExec ex;
    ResolveArgs A; memset(&A, 0, sizeof A);
    A.world = (uint32_t)world; A.rank = (uint32_t)rank; A.stride = stride;
    uint32_t off = 1;
    for (int r = 0; r < world; r++) { A.cnt[r][0] = (uint32_t)counts4[4 * r]; A.cnt[r][1] = (uint32_t)counts4[4 * r + 1]; A.off[r] = off; off += (uint32_t)counts4[4 * r + 2]; }
    xpar[0] = 0;
    resolve_init_all(ex, xpar, lists, A);
    resolve_union_all(ex, xpar, lists, A);
    resolve_apply_all(ex, c->E.P, xpar, lists, A);
    return 0;
}
extern "C" int rg_slab_export_packed(void *cv, uint32_t *packed, uint32_t cap, void *)
{
    EmuCtx *c = (EmuCtx *)cv; HostExec ex;
    const Params &P = c->E.P;
    packed[0] = packed[1] = packed[2] = packed[3] = 0;
    for (int side = 0; side < 2; side++) {
        if (!(side ? P.hi_ext : P.lo_ext)) continue;
        std::vector<uint32_t> off((size_t)P.ntx * P.nty + 1); uint32_t scratch[40];     /* >= iface_tiles + 1 */
        DISPATCH(c->cfg, (export_offsets<C>(ex, P, side, off.data(), packed + side, scratch),
                          export_entries<C>(ex, P, side, off.data(), packed + 4 + (size_t)side * cap, cap)));
    }
    packed[2] = c->E.counters[CTR_NODES];
    return 0;
}
extern "C" int rg_slab_resolve_packed(void *cv, uint32_t *xpar, uint32_t xpar_cap, const uint32_t *packed_all, uint32_t cap, int rank, int world, void *)
{
    EmuCtx *c = (EmuCtx *)cv; HostExec ex;
    ResolveArgs A; memset(&A, 0, sizeof A);
    xpar[0] = 0;
    resolve_prepare(c->E.P, packed_all, 4u + 2u * cap, cap, xpar_cap, (uint32_t)rank, (uint32_t)world, &A);
    resolve_init_packed(ex, xpar, packed_all, cap, A);
    resolve_union_packed(ex, xpar, packed_all, cap, A);
    resolve_apply_packed(ex, c->E.P, xpar, packed_all, cap, A);
    return 0;
}
extern "C" int rg_ctx_status(void *cv, uint32_t *stats)
{
    EmuCtx *c = (EmuCtx *)cv;
    if (stats) { stats[0] = c->E.counters[CTR_NODES]; stats[1] = c->E.counters[CTR_STATUS]; stats[2] = c->E.counters[CTR_NCOMPLEX] + c->E.counters[CTR_NDENSE] + c->E.counters[CTR_NROWTILES]; stats[3] = c->E.counters[CTR_NFLAGGED] + c->E.counters[CTR_NROWOUT]; }
    return c->E.counters[CTR_STATUS] ? -4 : 0;
}
extern "C" int rg_slab_finish(void *cv, void *)
{
    EmuCtx *c = (EmuCtx *)cv;
    DISPATCH(c->cfg, (emu_finish<C>(c->E)));
    return 0;     /* like the product: errors surface in rg_ctx_status */
}
