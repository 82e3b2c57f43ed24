/*
 * rg_core.h -- kernel bodies of the B200 seeded region-growing engine.
 *
 * Replaces the grow loops of the reference (citations relative to /root/reference):
 *   2-mpi-region-growing/region.c:493-533   grow_region()          (2-D, 4-conn, |d| < T)
 *   5-cuda-region-growing/raycast.cu:281-318 grow_region_serial()  (3-D, 6-conn)
 *   5-cuda-region-growing/raycast.cu:534-574, 589-699, 702-822     the two CUDA growers
 * It is NOT a translation of those.  The reference result equals "the union of the connected
 * components of the graph {grid neighbours with |delta| < T} that contain a seed" (SURVEY.md
 * section 0), so the engine labels components instead of sweeping a frontier:
 *
 *   pass A  (tile_pass_a)   one CTA per tile: stage the uint8 tile + 1-voxel halo in shared
 *           memory, build x/y/z "similar" bit planes with byte-SIMD, run-based union-find on
 *           16-bit parents in shared memory, give dense ids ("nodes") only to components that
 *           own a cross-tile edge or a seed, emit per-face lists of node ids.  Reads 1 B/voxel.
 *   pass B  (merge_faces)   union-find over the coarse node graph with lock-free hooking; seeds
 *           are pre-hooked to node 0, so "in the region" == "root is node 0".  One launch, no
 *           global iteration and no host round trip, whatever the region's diameter.
 *   pass C  (tile_pass_c)   tiles with no selected node need no image read (their mask bytes are
 *           zeros); the others re-stage, rebuild the link planes and flood the region from the selected
 *           representatives with bit-parallel sweeps (stride planes carry it 1, 2, 4, ... rows and planes
 *           per sweep), then write 0/1 mask bytes with 16-byte stores.  Writes 1 B/voxel, reads <= 1 B/voxel.
 *
 * Every body is written against an `Exec` (thread id, parallel-for, barrier, scan) so that the
 * same source runs as a CUDA block (rg_engine.cu) and, for the CPU logic tests only, as a
 * sequential loop (tests/emu).  See rg_port.h.
 */
#pragma once
#include "rg_port.h"

namespace rg {

enum { F_XM = 0, F_XP = 1, F_YM = 2, F_YP = 3, F_ZM = 4, F_ZP = 5 };
enum { CTR_NODES = 0, CTR_STATUS = 1, CTR_FACE_MISMATCH = 2, CTR_NCOMPLEX = 3, CTR_NFLAGGED = 4, CTR_NPLANES = 5, CTR_NDENSE = 6, CTR_NROWTILES = 7, CTR_NROWOUT = 8, CTR_WORDS = 16 };
/* CTR_NCOMPLEX: tiles listed for the full labelling pass; CTR_NDENSE: tiles the in-CTA dense path labelled (rg_dense.h);
   CTR_NPLANES: link-plane slots handed out (both of those passes keep a tile's planes for pass C);
   CTR_NROWTILES: tiles of constant rows labelled at row granularity (rg_rows.h); CTR_NROWOUT: those of them that hold region
   voxels (statistics; the row tiles themselves are listed from the END of flag_list downwards for the row output kernel) */
enum { ST_NODE_OVERFLOW = 1u, ST_FACE_MISMATCH = 2u };
/* top bit of a face-list entry: the node is a single voxel with this one cross-tile edge and no seed.  Two such nodes facing each
   other are a closed component of two voxels: never part of the region, so the merge skips the pair.  Node ids are below 2^15. */
enum { ENT_LEAF = 0x8000u };

/* XR_: bytes staged to the right of the tile.  16 (the default) brings the x = TX halo byte along in the TMA box; 0 is
   the lean layout of the fast labelling pass, which fetches that one column by itself and gains a fourth CTA per SM. */
template <int TX_, int TY_, int TZ_, int XR_ = 16>
struct TileCfg {
    static constexpr int TX = TX_, TY = TY_, TZ = TZ_, XR = XR_;
    static constexpr int WX = TX / 32;             /* 32-voxel words per row */
    static constexpr int NW = WX * TY * TZ;        /* words per tile */
    static constexpr int NV = TX * TY * TZ;        /* voxels per tile */
    static constexpr int HZ = TZ > 1 ? 1 : 0;      /* z halo planes (none for images) */
    static constexpr int PZ = TZ + 2 * HZ, PY = TY + 2;
    static constexpr int CX = TX / 16;             /* 16-byte chunks per row (own voxels) */
    static constexpr int XO = 16;                  /* staged rows start 16 bytes left of the tile: x = -1 is byte 15 */
    static constexpr int RS = TX + XO + XR;        /* staged row stride: [x0-16, x0+TX+XR) */
    static_assert(XR == 0 || XR == 16, "the right margin is one 16-byte chunk or nothing");
    static constexpr int CXH = RS / 16;            /* 16-byte chunks per staged row */
    static constexpr int PS = PY * RS;             /* staged plane stride */
    static constexpr int IMG_BYTES = PZ * PS;
    static constexpr int TMA_NY = PY <= 256 ? 1 : 2;   /* TMA boxes are at most 256 rows high */
    static constexpr int BOXY = PY / TMA_NY;
    static_assert(PY % TMA_NY == 0, "split rows evenly");
    static_assert(TMA_NY == 1 || (BOXY * RS) % 128 == 0, "every TMA box must land 128-byte aligned");
    static constexpr int FXN = TY * TZ, FYN = TX * TZ, FZN = HZ ? TX * TY : 0;
    static constexpr int FXW = (FXN + 31) / 32, FYW = WX * TZ, FZW = HZ ? WX * TY : 0;
    static constexpr int FW_TOTAL = 2 * (FXW + FYW + FZW);
    static constexpr int ENT_MAX = 2 * (FXN + FYN + FZN);   /* face-list entries per tile, worst case */
    static_assert(TX % 32 == 0, "rows are whole words");
    static_assert(NV <= 65536, "16-bit tile-local indices");
    static_assert(NW % 4 == 0, "bit sets are cleared 16 bytes at a time");
    RG_MFN static constexpr int fw_off(int f)               /* first mask word of face f: branch-free, f is often a run-time value */
    {
        return (f > 0) * FXW + (f > 1) * FXW + (f > 2) * FYW + (f > 3) * FYW + (f > 4) * FZW + (f > 5) * FZW;
    }
};

/* tile shapes: ~32 K voxels either way.  2-D: 126 rows (staged tile = 128 rows, one TMA box; the flood of pass C advances one row per sweep) */
#ifndef RG_TX3
#define RG_TX3 64
#define RG_TY3 32
#define RG_TZ3 16
#endif
typedef TileCfg<RG_TX3, RG_TY3, RG_TZ3> Cfg3D;
#ifndef RG_TY2
#define RG_TY2 126
#endif
typedef TileCfg<128, RG_TY2, 1> Cfg2D;
/* the staging layout the fast labelling pass uses for a tile shape: lean for volumes, unchanged for images (their
   staged tile is 20 KB anyway) */
template <class C> struct LeanOf { typedef TileCfg<C::TX, C::TY, C::TZ, C::HZ ? 0 : C::XR> type; };

struct TileRec {                 /* 32 bytes per tile, written by pass A */
    uint32_t node_base;          /* first global node id of this tile */
    uint32_t nnodes;
    uint16_t cnt[6];             /* face-list lengths, order XM XP YM YP ZM ZP */
    uint16_t uniform;            /* 1: all voxels of the tile are one component (pass C needs no image read); 2: a tile of constant rows,
                                    its 2 x TZ row-link words are in its link-plane slot (rg_rows.h) */
    uint16_t has_region;         /* written by the flag pass: 1 = the tile holds region voxels (output-sensitive copies of the mask) */
    uint32_t plane_slot1;        /* 1 + index of the tile's saved link planes (Params::planes), 0 = none */
    uint32_t pad3;
};

struct Params {
    const uint8_t *image;        /* plane z = 0 of this slab; planes -1 / nz readable if z*_ext */
    uint8_t *mask;
    int64_t nx, ny, nz;
    int32_t ntx, nty, ntz;
    int32_t lo_ext, hi_ext;      /* slab has a neighbour slab below / above (multi-GPU) along the slab axis:
                                    z for volumes (tiles with a z halo), y for images -- see ext_ylo() ... */
    uint32_t open;               /* images cut into blocks (rows x columns over several GPUs): bit 0 / 1 / 2 / 3 = the block's -x / +x / -y / +y
                                    border is a seam to another block.  Every pixel on such a border becomes a cross-tile edge (no halo is
                                    read, nothing is compared): the exchange itself tests the pixel pairs across the seam (rg_slab_core.h) */
    int32_t tmode;               /* 0: T <= 0 (no edges)  1: 1..255  2: T >= 256 (all edges) */
    uint32_t t4;                 /* T replicated into 4 bytes */
    uint32_t thr;                /* scalar form: similar <=> |a-b| < thr  (0, T or 256) */
    int32_t fast16;              /* nx % 16 == 0 and both buffers 16-byte aligned */
    int32_t use_tma;             /* tiles are staged by one cp.async.bulk.tensor (needs fast16) */
    const uint64_t *seed_keys;   /* sorted (tile << 16 | voxel-in-tile); ~0 = dropped */
    int32_t nseeds;
    TileRec *rec;
    uint16_t *ent;               /* [ntiles][ENT_MAX] node ids local to the tile */
    uint32_t *nparent;           /* coarse union-find; node 0 = "holds a seed" */
    uint16_t *nrep;              /* node -> root voxel index inside its tile */
    uint32_t *counters;          /* CTR_* */
    uint32_t node_cap;
    uint32_t *complex_list;      /* tiles the fast labelling pass handed to the full one */
    uint32_t *flag_list;         /* tiles that hold part of the region (pass C work list) */
    uint32_t *planes;            /* [plane_cap][3][NW] link planes (ex, ey, ez) the full pass saves for pass C; may be null */
    uint32_t plane_cap;          /* how many tiles' planes fit */
    int32_t mask_prezeroed;      /* every CTA of the fast labelling pass clears its mask tile while its load is in flight; flag_tiles_all then stores no zeros */
    int32_t force_dense;         /* testing aid (env RG_FORCE_DENSE): the fast labelling pass sends every non-uniform tile down its dense path */
    int32_t no_rows;             /* testing aid (env RG_NO_ROWS): tiles of constant rows are labelled like any other tile */
    int32_t debug_stop;          /* tuning aid (env RG_DEBUG_STOP): fast pass returns after phase k; results are then invalid */
};

/* Link-plane slot of a tile (+ 1; 0 = none).  With a slot for every tile -- volumes up to 64 Ki tiles -- the slot is the tile index:
   no counter, nothing to wait for.  Otherwise slots are handed out by a counter to the tiles that ask (one thread per tile calls this). */
RG_FN uint32_t plane_slot1(const Params &P, uint32_t tile)
{
    if (!P.planes) return 0u;
    if (P.plane_cap >= (uint32_t)P.ntx * (uint32_t)P.nty * (uint32_t)P.ntz) return tile + 1u;
    const uint32_t slot = atomic_add(&P.counters[CTR_NPLANES], 1u);
    return slot < P.plane_cap ? slot + 1u : 0u;
}
RG_FN bool plane_slot_is_tile(const Params &P) { return P.planes && P.plane_cap >= (uint32_t)P.ntx * (uint32_t)P.nty * (uint32_t)P.ntz; }

/* which axis a multi-GPU slab is cut along: z for volumes, y (rows) for images (tiles without a z halo) */
template <class C> RG_HDFN int ext_zlo(const Params &P) { return C::HZ ? P.lo_ext : 0; }
template <class C> RG_HDFN int ext_zhi(const Params &P) { return C::HZ ? P.hi_ext : 0; }
template <class C> RG_HDFN int ext_ylo(const Params &P) { return C::HZ ? 0 : P.lo_ext; }
template <class C> RG_HDFN int ext_yhi(const Params &P) { return C::HZ ? 0 : P.hi_ext; }

#ifdef RG_PARENT32
typedef uint32_t parent_t;     /* experiment: native 32-bit CAS, 1 CTA/SM */
#else
typedef uint16_t parent_t;
#endif

/* ---------------------------------------------------------------- shared-memory tile state */
constexpr int ceil_log2(int n) { return n <= 1 ? 0 : 1 + ceil_log2((n + 1) / 2); }

/* longer strides for the flood: eyk[k-1] bit i = voxel i is linked through 2^k consecutive +y steps (k >= 1;
   k = 0 is the ey plane itself), ezk likewise in z -- a sweep then carries a set 1, 2, 4, ... rows / planes at once */
template <class C>
struct StridePlanes {
    static constexpr int NLY = ceil_log2(C::TY), NLZ = ceil_log2(C::TZ);
    alignas(16) uint32_t eyk[NLY > 1 ? NLY - 1 : 1][C::NW];
    alignas(16) uint32_t ezk[NLZ > 1 ? NLZ - 1 : 1][C::NW];
};

template <class C>
struct alignas(128) TileSmem {
    union alignas(128) U {
        struct {
            uint8_t img[C::IMG_BYTES];             /* tile + 1-voxel halo, rows of RS bytes (see img_off) */
        } i;
        parent_t parent[C::NV];                    /* run union-find (image is dead by then) */
    } u;
    uint64_t mbar;                                 /* TMA completion barrier */
    alignas(16) uint32_t ex[C::NW];                /* bit i: voxel i similar to its +x/+y/+z neighbour (in-tile); 16-byte */
    alignas(16) uint32_t ey[C::NW];                /* aligned: they are copied out 16 bytes at a time for pass C */
    alignas(16) uint32_t ez[C::NW];
    uint32_t fmask[C::FW_TOTAL];                   /* cross-tile edges per face, de-duplicated */
    uint16_t fprefix[C::FW_TOTAL + 2];
    uint16_t ent_root[C::ENT_MAX];
    uint32_t openbits[C::NW], seedbits[C::NW];
    uint16_t wprefix[C::NW];
    uint32_t scal[16];
    uint32_t scan_scratch[40];
};
enum { SC_SEED_LO = 0, SC_SEED_HI = 1, SC_NODE_BASE = 2, SC_NNODES = 3, SC_ETOT = 4, SC_OVERFLOW = 5, SC_ANY = 6, SC_COMPLEX = 7, SC_UNIFORM = 8, SC_NSLOW = 9, SC_PLANE1 = 10, SC_CNT0 = 16 };

/* ---------------------------------------------------------------- small helpers */
RG_FN uint32_t valid_bits(int64_t n) { return n >= 32 ? 0xFFFFFFFFu : n <= 0 ? 0u : ((1u << n) - 1u); }

RG_FN void load8(const uint8_t *p, uint32_t a[8])
{
    u32x4 lo = ld128(p), hi = ld128(p + 16);
    a[0] = lo.x; a[1] = lo.y; a[2] = lo.z; a[3] = lo.w; a[4] = hi.x; a[5] = hi.y; a[6] = hi.z; a[7] = hi.w;
}
RG_FN void shift8(const uint32_t a[8], uint32_t next_byte, uint32_t b[8])
{
#if RG_DEVICE_BUILD
#pragma unroll
#endif
    for (int j = 0; j < 7; j++) b[j] = (a[j] >> 8) | (a[j + 1] << 24);
    b[7] = (a[7] >> 8) | (next_byte << 24);
}
RG_FN uint32_t edge_word(const uint32_t a[8], const uint32_t b[8], int tmode, uint32_t t4)
{
    if (tmode == 0) return 0u;
    if (tmode == 2) return 0xFFFFFFFFu;
    const ByteCmp k = byte_cmp_for(t4 & 0xFFu);                /* depends on the launch only: hoisted out of every loop */
    uint32_t w = 0;
#if RG_DEVICE_BUILD
#pragma unroll
#endif
    for (int j = 7; j >= 0; j--) w = funnel_l(lt4_top_nibble(a[j], b[j], k), w, 4);   /* (w << 4) | nibble, one SHF each */
    return w;
}
RG_FN bool similar1(uint32_t a, uint32_t b, int tmode, uint32_t t4)
{
    if (tmode == 0) return false;
    if (tmode == 2) return true;
    int d = (int)a - (int)b; if (d < 0) d = -d;
    return d < (int)(t4 & 0xFF);
}
/* x-edges inside one 32-voxel word (bit 31 is meaningless) */
RG_FN uint32_t xedges_within(const uint32_t a[8], int tmode, uint32_t t4)
{
    uint32_t b[8]; shift8(a, 0, b);
    return edge_word(a, b, tmode, t4);
}
/* keep one cross edge per pair of overlapping x-runs: bit i is redundant if bit i-1 is a cross
   edge too and both rows continue from i-1 to i.  Both sides of a face evaluate this on the
   same voxel values, so their lists stay aligned entry by entry. */
RG_FN uint32_t dedupe(uint32_t m, uint32_t ex_a, uint32_t ex_b) { return m & ~((m & ex_a & ex_b) << 1); }

RG_FN uint32_t run_start(uint32_t exw, int w, int bit)
{
    uint32_t S = ~(exw << 1);                                   /* bit 0 always starts a run */
    return (uint32_t)(w * 32 + 31 - clz(S & (0xFFFFFFFFu >> (31 - bit))));
}

/* lock-free union-find, roots are minima (after ECL-CC, Jaiganesh & Burtscher 2018) */
RG_FN uint32_t find16(uint16_t *par, uint32_t x)
{
    uint32_t curr = ldv16(par + x);
    if (curr != x) {
        uint32_t prev = x, next;
        while (curr > (next = ldv16(par + curr))) { stv16(par + prev, (uint16_t)next); prev = curr; curr = next; }
    }
    return curr;
}
RG_FN void unite16(uint16_t *par, uint32_t a, uint32_t b)
{
    a = find16(par, a); b = find16(par, b);
    while (a != b) {
        if (a < b) { uint32_t t = a; a = b; b = t; }
        uint32_t old = cas16(par + a, (uint16_t)a, (uint16_t)b);
        if (old == a) break;
        a = old;
    }
}
RG_FN uint32_t find32(uint32_t *par, uint32_t x)
{
    uint32_t curr = ldv32(par + x);
    if (curr != x) {
        uint32_t prev = x, next;
        while (curr > (next = ldv32(par + curr))) { stv32(par + prev, next); prev = curr; curr = next; }
    }
    return curr;
}
RG_FN void unite32(uint32_t *par, uint32_t a, uint32_t b)
{
    a = find32(par, a); b = find32(par, b);
    while (a != b) {
        if (a < b) { uint32_t t = a; a = b; b = t; }
        uint32_t old = cas32(par + a, a, b);
        if (old == a) break;
        a = old;
    }
}

RG_FN uint32_t find16(uint32_t *par, uint32_t x) { return find32(par, x); }
RG_FN void unite16(uint32_t *par, uint32_t a, uint32_t b) { unite32(par, a, b); }

template <class C>
RG_FN void tile_origin(const Params &P, uint32_t tile, int &tx, int &ty, int &tz)
{
    tx = (int)(tile % (uint32_t)P.ntx);
    uint32_t r = tile / (uint32_t)P.ntx;
    ty = (int)(r % (uint32_t)P.nty);
    tz = (int)(r / (uint32_t)P.nty);
}

/* staged position of tile-local voxel (x,y,z), -1 <= x <= TX, -1 <= y <= TY, -HZ <= z <= TZ-1+HZ */
template <class C> RG_FN int img_off(int x, int y, int z) { return ((z + C::HZ) * C::PY + (y + 1)) * C::RS + C::XO + x; }

/* voxel (tile-local linear index) on this side of cross-tile edge `b` of face word (f, idx) */
template <class C>
RG_FN uint32_t face_voxel(int f, int idx, int b, int vy, int vz)
{
    int x, y, z;
    if (f < 2) { int bi = idx * 32 + b; z = bi / C::TY; y = bi % C::TY; x = (f == F_XP) ? C::TX - 1 : 0; }
    else if (f < 4) { z = idx / C::WX; x = (idx % C::WX) * 32 + b; y = (f == F_YP) ? vy - 1 : 0; }
    else { y = idx / C::WX; x = (idx % C::WX) * 32 + b; z = (f == F_ZP) ? vz - 1 : 0; }
    return (uint32_t)((z * C::TY + y) * C::TX + x);
}

/* ---------------------------------------------------------------- stage a tile in shared memory */
template <class C, class Exec>
RG_FN void load_tile(Exec &ex, const Params &P, uint8_t *img, int64_t x0, int64_t y0, int64_t z0, bool halo)
{
    /* Generic staging (used when TMA is not available: unaligned rows, or the CPU walk of this code).
       With 16-byte aligned rows every chunk is one asynchronous copy, so a thread keeps all its chunks
       in flight; the caller runs ex.tile_wait() + a barrier before reading.  The TMA path of the CUDA
       Exec (rg_engine.cu) produces the same bytes with one instruction per tile. */
    const int64_t nx = P.nx, ny = P.ny, nz = P.nz;
    const int64_t zmin = -(int64_t)ext_zlo<C>(P), zmax = nz + ext_zhi<C>(P);    /* readable plane range */
    const int64_t ymin = -(int64_t)ext_ylo<C>(P), ymax = ny + ext_yhi<C>(P);    /* readable row range (images cut into row slabs) */
    const bool fast = P.fast16 != 0;
    ex.pfor(C::PZ * C::PY * C::CXH, [&](int it) {
        int c = it % C::CXH, r = it / C::CXH;
        int py = r % C::PY, pz = r / C::PY;
        int ly = py - 1, lz = pz - C::HZ;
        bool yh = (ly < 0 || ly >= C::TY), zh = (lz < 0 || lz >= C::TZ), xh = (c == 0 || (C::XR > 0 && c == C::CXH - 1));
        bool want = !(yh && zh) && !(xh && (yh || zh)) && (halo || !(yh || zh || xh));
        if (!want) return;                                      /* never read: edge / corner columns, halo of pass C */
        int64_t gx = x0 - C::XO + 16 * c, gy = y0 + ly, gz = z0 + lz;
        bool inside = gy >= ymin && gy < ymax && gz >= zmin && gz < zmax && gx >= 0 && gx < nx;
        if (xh) inside = inside && gz >= 0 && gz < nz && gy >= 0 && gy < ny;   /* x halo columns only matter for own rows */
        uint8_t *dst = &img[(pz * C::PY + py) * C::RS + 16 * c];
        if (fast) cp_async16(dst, P.image + (inside ? (gz * ny + gy) * nx + gx : 0), inside);
        else {
            uint32_t t[4] = {0, 0, 0, 0};
            if (gy >= ymin && gy < ymax && gz >= zmin && gz < zmax) {
                const uint8_t *g = P.image + (gz * ny + gy) * nx;
                for (int b = 0; b < 16; b++) if (gx + b >= 0 && gx + b < nx) t[b >> 2] |= (uint32_t)g[gx + b] << (8 * (b & 3));
            }
            st128(dst, u32x4{t[0], t[1], t[2], t[3]});
        }
    });
}

/* ---------------------------------------------------------------- edge bit planes (+ face masks if FACES) */
template <class C, bool FACES, class Exec>
RG_FN void build_edges(Exec &ex, const Params &P, const uint8_t *img, uint32_t *pex, uint32_t *pey, uint32_t *pez, uint32_t *fmask,
                        int64_t x0, int64_t y0, int64_t z0)
{
    const int64_t nx = P.nx, ny = P.ny, nz = P.nz;
    const int tmode = P.tmode; const uint32_t t4 = P.t4;
    const int vy = (int)((ny - y0) < C::TY ? (ny - y0) : C::TY);
    const int vz = (int)((nz - z0) < C::TZ ? (nz - z0) : C::TZ);
    const bool has_yp = (y0 + C::TY < ny) || (y0 + vy == ny && ext_yhi<C>(P)), has_ym = y0 > 0 || ext_ylo<C>(P);
    const bool has_zp = C::HZ && ((z0 + C::TZ < nz) || (z0 + vz == nz && ext_zhi<C>(P)));
    const bool has_zm = C::HZ && (z0 > 0 || ext_zlo<C>(P));
    ex.pfor(C::NW, [&](int w) {
        int wx = w % C::WX, y = (w / C::WX) % C::TY, z = w / (C::WX * C::TY);
        int64_t gx0 = x0 + 32 * wx;
        uint32_t vmask = (y < vy && z < vz) ? valid_bits(nx - gx0) : 0u;
        uint32_t exw = 0, eyw = 0, ezw = 0;
        if (vmask) {
            const uint8_t *rowp = &img[img_off<C>(32 * wx, y, z)];
            uint32_t a[8], b[8];
            load8(rowp, a);
            /* +x */
            uint32_t nextb = 0;
            if (wx < C::WX - 1) nextb = rowp[32];
            else if (FACES) nextb = rowp[32];                        /* x = TX halo byte */
            shift8(a, nextb, b);
            uint32_t nextvalid = (gx0 + 32 < nx) ? 0x80000000u : 0u;
            uint32_t e = edge_word(a, b, tmode, t4) & vmask & ((vmask >> 1) | nextvalid);
            if (wx == C::WX - 1) {
                if (FACES && (e >> 31)) {
                    /* images: one entry per pair of overlapping y-runs (the same rule as rg_fast.h build_face_masks) */
                    const uint8_t *pc = rowp + 31;
                    const bool dup = !C::HZ && y > 0 && similar1(pc[-C::RS], pc[1 - C::RS], tmode, t4) && similar1(pc[0], pc[-C::RS], tmode, t4)
                                     && similar1(pc[1], pc[1 - C::RS], tmode, t4);
                    if (!dup) { int bi = z * C::TY + y; atomic_or(&fmask[C::fw_off(F_XP) + (bi >> 5)], 1u << (bi & 31)); }
                }
                e &= 0x7FFFFFFFu;
            }
            exw = e;
            if (FACES && wx == 0 && x0 > 0 && (vmask & 1u)) {
                if (similar1(rowp[-1], a[0] & 0xFFu, tmode, t4)) {
                    const bool dup = !C::HZ && y > 0 && similar1(rowp[-1 - C::RS], rowp[-C::RS], tmode, t4) && similar1(rowp[0], rowp[-C::RS], tmode, t4)
                                     && similar1(rowp[-1], rowp[-1 - C::RS], tmode, t4);
                    if (!dup) { int bi = z * C::TY + y; atomic_or(&fmask[C::fw_off(F_XM) + (bi >> 5)], 1u << (bi & 31)); }
                }
            }
            /* +y */
            if (y < vy - 1) {
                load8(rowp + C::RS, b);
                eyw = edge_word(a, b, tmode, t4) & vmask;
            } else if (FACES && has_yp) {                       /* y == TY-1: edges into the +y tile */
                load8(rowp + C::RS, b);
                uint32_t m = edge_word(a, b, tmode, t4) & vmask;
                fmask[C::fw_off(F_YP) + z * C::WX + wx] = dedupe(m, e, xedges_within(b, tmode, t4));
            }
            if (FACES && y == 0 && has_ym) {
                load8(rowp - C::RS, b);
                uint32_t m = edge_word(b, a, tmode, t4) & vmask;
                fmask[C::fw_off(F_YM) + z * C::WX + wx] = dedupe(m, xedges_within(b, tmode, t4), e);
            }
            /* +z */
            if (C::HZ) {
                const int zs = C::PS;
                if (z < vz - 1) {
                    load8(rowp + zs, b);
                    ezw = edge_word(a, b, tmode, t4) & vmask;
                } else if (FACES && has_zp) {                   /* top valid plane: edges into the +z tile / slab */
                    load8(rowp + zs, b);
                    uint32_t m = edge_word(a, b, tmode, t4) & vmask;
                    fmask[C::fw_off(F_ZP) + y * C::WX + wx] = dedupe(m, e, xedges_within(b, tmode, t4));
                }
                if (FACES && z == 0 && has_zm) {
                    load8(rowp - zs, b);
                    uint32_t m = edge_word(b, a, tmode, t4) & vmask;
                    fmask[C::fw_off(F_ZM) + y * C::WX + wx] = dedupe(m, xedges_within(b, tmode, t4), e);
                }
            }
        }
        pex[w] = exw; pey[w] = eyw; pez[w] = ezw;
    });
}

/* ---------------------------------------------------------------- bit-parallel flood inside a tile
   (pass C floods the region from its representatives) */
#ifndef RG_FLOOD_PLAIN
#define RG_FLOOD_PLAIN 3
#endif
enum { FLOOD_PLAIN_SWEEPS = RG_FLOOD_PLAIN };     /* pass C: sweeps with unit stride before the stride planes are built */

/* all voxels reachable from the set bits of s along the links of A (bit i: voxel i linked to i+1), inside
   one 32-voxel word: a carry runs from each seed to the end of its run (upwards), and the same on the
   bit-reversed word (downwards) */
RG_FN uint32_t fill_up(uint32_t s, uint32_t A) { return ((A + (s & A)) ^ A) | s; }
#if RG_DEVICE_BUILD
RG_FN uint32_t brev32(uint32_t x) { return __brev(x); }
#else
RG_FN uint32_t brev32(uint32_t x) { uint32_t r = 0; for (int i = 0; i < 32; i++) if ((x >> i) & 1u) r |= 1u << (31 - i); return r; }
#endif
RG_FN uint32_t fill_word(uint32_t s, uint32_t A)
{
    A &= 0x7FFFFFFFu;
    return fill_up(s, A) | brev32(fill_up(brev32(s), brev32(A) >> 1));
}

/* one level per barrier: linked over 2h steps = linked over h steps from here and from h further on */
template <class C, class Exec>
RG_FN void build_stride_planes(Exec &ex, const uint32_t *pey, const uint32_t *pez, StridePlanes<C> &sp)
{
    constexpr int NLY = StridePlanes<C>::NLY, NLZ = StridePlanes<C>::NLZ, NLMAX = NLY > NLZ ? NLY : NLZ;
    for (int k = 1; k < NLMAX; k++) {
        const int h = 1 << (k - 1);
        ex.pfor(C::NW, [&](int w) {
            const int y = (w / C::WX) % C::TY, z = w / (C::WX * C::TY);
            if (k < NLY) {
                const uint32_t *lo = k == 1 ? pey : sp.eyk[k - 2];
                sp.eyk[k - 1][w] = (y + h < C::TY) ? (lo[w] & lo[w + h * C::WX]) : 0u;
            }
            if (k < NLZ) {
                const uint32_t *lo = k == 1 ? pez : sp.ezk[k - 2];
                sp.ezk[k - 1][w] = (z + h < C::TZ) ? (lo[w] & lo[w + h * C::WX * C::TY]) : 0u;
            }
        });
        ex.sync();
    }
}

/* One sweep over the words of M (bit i: voxel i belongs to the set): take what the neighbouring words offer through
   the link planes -- with `strided` also the words 2, 4, ... rows and planes away -- then close the word's own
   x-runs.  The words race benignly: bits are only ever added, and a sweep that added one reports it, so the caller
   sweeps again.  Ends with a barrier; returns whether any word changed. */
template <class C, class Exec>
RG_FN bool flood_sweep(Exec &ex, uint32_t *M, const uint32_t *pex, const uint32_t *pey, const uint32_t *pez,
                       const StridePlanes<C> &sp, const bool strided)
{
    constexpr int NLY = StridePlanes<C>::NLY, NLZ = StridePlanes<C>::NLZ;
    bool changed = false;
    ex.pfor(C::NW, [&](int w) {
        const int wx = w % C::WX, y = (w / C::WX) % C::TY, z = w / (C::WX * C::TY);
        const uint32_t m = M[w], exw = pex[w];
        uint32_t in = m;
        if (y > 0) in |= M[w - C::WX] & pey[w - C::WX];
        if (y < C::TY - 1) in |= M[w + C::WX] & pey[w];
        if (C::HZ) {
            if (z > 0) in |= M[w - C::WX * C::TY] & pez[w - C::WX * C::TY];
            if (z < C::TZ - 1) in |= M[w + C::WX * C::TY] & pez[w];
        }
        if (strided) {
#if RG_DEVICE_BUILD
#pragma unroll
#endif
            for (int k = 1; k < NLY; k++) {
                const int d = 1 << k, o = d * C::WX;
                if (y >= d) in |= M[w - o] & sp.eyk[k - 1][w - o];
                if (y + d < C::TY) in |= M[w + o] & sp.eyk[k - 1][w];
            }
#if RG_DEVICE_BUILD
#pragma unroll
#endif
            for (int k = 1; k < NLZ; k++) {
                const int d = 1 << k, o = d * C::WX * C::TY;
                if (z >= d) in |= M[w - o] & sp.ezk[k - 1][w - o];
                if (z + d < C::TZ) in |= M[w + o] & sp.ezk[k - 1][w];
            }
        }
        if (wx > 0) in |= (M[w - 1] & pex[w - 1]) >> 31;
        if (wx < C::WX - 1) in |= ((M[w + 1] & 1u) << 31) & exw;
        if (in) {
            const uint32_t out = fill_word(in, exw);
            if (out != m) { M[w] = out; changed = true; }
        }
    });
    return ex.sync_or(changed);
}

/* ---------------------------------------------------------------- run-based union-find inside the tile */
template <class C, class Exec>
RG_FN void label_runs(Exec &ex, TileSmem<C> &s)
{
    /* parent[v] = v, eight 16-bit entries per 16-byte store */
    if (sizeof(parent_t) == 2) ex.pfor(C::NV / 8, [&](int it) {
        uint32_t v = (uint32_t)it * 8;
        u32x4 q = { v | ((v + 1) << 16), (v + 2) | ((v + 3) << 16), (v + 4) | ((v + 5) << 16), (v + 6) | ((v + 7) << 16) };
        st128(&s.u.parent[v], q);
    });
    else ex.pfor(C::NV / 4, [&](int it) {
        uint32_t v = (uint32_t)it * 4;
        st128(&s.u.parent[v], u32x4{v, v + 1, v + 2, v + 3});
    });
    ex.sync();
    /* three sweeps: links inside rows first, then between rows, then between planes -- each sweep starts from
       trees the previous one has already flattened along one axis, which keeps the find paths short when a
       tile holds one large component */
    ex.pfor(C::NW, [&](int w) {
        if (w % C::WX == 0) return;
        uint32_t exl = s.ex[w - 1];
        if (exl >> 31) unite16(s.u.parent, run_start(exl, w - 1, 31), (uint32_t)w * 32);
    });
    ex.sync();
    ex.pfor(C::NW, [&](int w) {
        uint32_t m = s.ey[w];
        if (!m) return;
        uint32_t exw = s.ex[w];
        int up = w + C::WX; uint32_t exu = s.ex[up];
        uint32_t n = dedupe(m, exw, exu);
        while (n) { int i = ffs1(n) - 1; n &= n - 1; unite16(s.u.parent, run_start(exw, w, i), run_start(exu, up, i)); }
    });
    if (C::HZ) {
        ex.sync();
        ex.pfor(C::NW, [&](int w) {
            uint32_t m = s.ez[w];
            if (!m) return;
            uint32_t exw = s.ex[w];
            int up = w + C::WX * C::TY; uint32_t exu = s.ex[up];
            uint32_t n = dedupe(m, exw, exu);
            while (n) { int i = ffs1(n) - 1; n &= n - 1; unite16(s.u.parent, run_start(exw, w, i), run_start(exu, up, i)); }
        });
    }
    ex.sync();
}

/* ---------------------------------------------------------------- nodes, face lists and the tile record
   (shared by the fast and the full labelling pass; S = TileSmem<C> or FastSmem<C>) */
template <class C, class S, class Exec>
RG_FN void emit_nodes(Exec &ex, const Params &P, uint32_t tile, S &s)
{
    /* dense node ids in root order */
    uint32_t k = ex.scan16(C::NW, [&](int i) { return (uint32_t)popc(s.openbits[i]); }, s.wprefix, s.scan_scratch);
    if (ex.tid() == 0) {
        uint32_t base = k ? atomic_add(&P.counters[CTR_NODES], k) : 0u;
        if (k && (base + k > P.node_cap || base + k < base)) { atomic_or(&P.counters[CTR_STATUS], ST_NODE_OVERFLOW); s.scal[SC_OVERFLOW] = 1; }
        s.scal[SC_NODE_BASE] = base; s.scal[SC_NNODES] = k;
    }
    ex.sync();
    const uint32_t base = s.scal[SC_NODE_BASE];
    const bool ok = !s.scal[SC_OVERFLOW];
    if (ok) {
        ex.pfor(C::NW, [&](int wi) {
            uint32_t m = s.openbits[wi];
            if (!m) return;
            uint32_t id = s.wprefix[wi], sb = s.seedbits[wi];
            while (m) {
                int b = ffs1(m) - 1; m &= m - 1;
                uint32_t g = base + id++;
                P.nparent[g] = ((sb >> b) & 1u) ? 0u : g;       /* seeded components start hooked to node 0 */
                P.nrep[g] = (uint16_t)(wi * 32 + b);
            }
        });
        uint16_t *ent = P.ent + (size_t)tile * C::ENT_MAX;
        ex.pfor((int)s.scal[SC_ETOT], [&](int e) {
            uint32_t r = s.ent_root[e];
            ent[e] = (uint16_t)(s.wprefix[r >> 5] + popc(s.openbits[r >> 5] & ((1u << (r & 31)) - 1u)));
        });
    }
    if (ex.tid() == 0) {
        TileRec rc;
        rc.node_base = base; rc.nnodes = ok ? k : 0u;
        for (int f = 0; f < 6; f++) rc.cnt[f] = ok ? (uint16_t)(s.fprefix[C::fw_off(f + 1)] - s.fprefix[C::fw_off(f)]) : (uint16_t)0;
        rc.uniform = (uint16_t)(ok ? s.scal[SC_UNIFORM] : 0u); rc.has_region = 0; rc.plane_slot1 = ok ? s.scal[SC_PLANE1] : 0u; rc.pad3 = 0;
        P.rec[tile] = rc;
    }
}

/* ================================================================ pass A */
template <class C, class Exec>
RG_FN void tile_pass_a(Exec &ex, const Params &P, uint32_t tile, TileSmem<C> &s)
{
    int tx, ty, tz; tile_origin<C>(P, tile, tx, ty, tz);
    const int64_t x0 = (int64_t)tx * C::TX, y0 = (int64_t)ty * C::TY, z0 = (int64_t)tz * C::TZ;
    const int vz = (int)((P.nz - z0) < C::TZ ? (P.nz - z0) : C::TZ);
    const int vy = (int)((P.ny - y0) < C::TY ? (P.ny - y0) : C::TY);

    ex.pfor(C::FW_TOTAL, [&](int i) { s.fmask[i] = 0; });
    ex.pfor(C::NW, [&](int i) { s.openbits[i] = 0; s.seedbits[i] = 0; });
    if (ex.tid() == 0) {
        /* this tile's slice of the sorted seed keys */
        uint64_t lo_key = (uint64_t)tile << 16, hi_key = ((uint64_t)tile + 1) << 16;
        int lo = 0, hi = P.nseeds;
        while (lo < hi) { int mid = (lo + hi) >> 1; if (P.seed_keys[mid] < lo_key) lo = mid + 1; else hi = mid; }
        int a = lo; hi = P.nseeds;
        while (lo < hi) { int mid = (lo + hi) >> 1; if (P.seed_keys[mid] < hi_key) lo = mid + 1; else hi = mid; }
        s.scal[SC_SEED_LO] = (uint32_t)a; s.scal[SC_SEED_HI] = (uint32_t)lo; s.scal[SC_OVERFLOW] = 0; s.scal[SC_UNIFORM] = 0;
        s.scal[SC_PLANE1] = plane_slot1(P, tile);              /* a slot for this tile's link planes (shared with the dense path) */
    }
    if (!ex.template tma_tile<C>(P, s.u.i.img, &s.mbar, x0, y0, z0)) load_tile<C>(ex, P, s.u.i.img, x0, y0, z0, true);
    ex.tile_wait(&s.mbar);
    ex.sync();
    if (P.debug_stop == 10) return;
    build_edges<C, true>(ex, P, s.u.i.img, s.ex, s.ey, s.ez, s.fmask, x0, y0, z0);
    ex.sync();
    if (P.debug_stop == 11) return;
    if (s.scal[SC_PLANE1]) {
        /* pass C needs exactly these three planes if the tile turns out to hold region voxels: 12 KB kept in place of
           a second 49 KB stage + the byte-SIMD comparison of every voxel pair */
        uint32_t *dst = P.planes + (size_t)(s.scal[SC_PLANE1] - 1u) * 3 * C::NW;
        ex.pfor(3 * C::NW / 4, [&](int i) {
            const int pl = i / (C::NW / 4), k = i % (C::NW / 4);
            const uint32_t *src = pl == 0 ? s.ex : pl == 1 ? s.ey : s.ez;
            st128(dst + (size_t)pl * C::NW + 4 * k, ld128(src + 4 * k));
        });
    }
    label_runs<C>(ex, s);
    if (P.debug_stop == 12) return;

    /* positions of the face-list entries */
    uint32_t etot = ex.scan16(C::FW_TOTAL, [&](int i) { return (uint32_t)popc(s.fmask[i]); }, s.fprefix, s.scan_scratch);
    if (ex.tid() == 0) { s.fprefix[C::FW_TOTAL] = (uint16_t)etot; s.scal[SC_ETOT] = etot; }
    ex.sync();

    /* start voxel of every cross-tile edge, in list order ... */
    ex.pfor(C::FW_TOTAL, [&](int fw) {
        uint32_t m = s.fmask[fw];
        if (!m) return;
        int f = 0; while (fw >= C::fw_off(f + 1)) f++;
        int idx = fw - C::fw_off(f);
        uint32_t slot = s.fprefix[fw];
        while (m) {
            int b = ffs1(m) - 1; m &= m - 1;
            int x, y, z;
            if (f < 2) { int bi = idx * 32 + b; z = bi / C::TY; y = bi % C::TY; x = (f == F_XP) ? C::TX - 1 : 0; }
            else if (f < 4) { z = idx / C::WX; x = (idx % C::WX) * 32 + b; y = (f == F_YP) ? vy - 1 : 0; }
            else { y = idx / C::WX; x = (idx % C::WX) * 32 + b; z = (f == F_ZP) ? vz - 1 : 0; }
            s.ent_root[slot++] = (uint16_t)((z * C::TY + y) * C::TX + x);
        }
    });
    ex.sync();
    /* ... then the root of each edge's component, one thread per edge (a face can hold 32 edges per word) */
    ex.pfor((int)etot, [&](int e) {
        const uint32_t v = s.ent_root[e];
        const int w = (int)(v >> 5);
        uint32_t r = find16(s.u.parent, run_start(s.ex[w], w, (int)(v & 31)));
        atomic_or(&s.openbits[r >> 5], 1u << (r & 31));
        s.ent_root[e] = (uint16_t)r;
    });
    if (P.debug_stop == 13) return;
    /* ... or a seed */
    {
        int lo = (int)s.scal[SC_SEED_LO], hi = (int)s.scal[SC_SEED_HI];
        ex.pfor(hi - lo, [&](int i) {
            uint32_t v = (uint32_t)(P.seed_keys[lo + i] & 0xFFFFu);
            int w = (int)(v >> 5);
            uint32_t r = find16(s.u.parent, run_start(s.ex[w], w, (int)(v & 31)));
            atomic_or(&s.openbits[r >> 5], 1u << (r & 31));
            atomic_or(&s.seedbits[r >> 5], 1u << (r & 31));
        });
    }
    ex.sync();

    emit_nodes<C>(ex, P, tile, s);
}

/* ================================================================ pass B: one (tile, +face) per item */
enum { MERGE_LANES = 32 };     /* one warp per tile: lanes share the tile's three + face lists */
#ifndef RG_MERGE_RUN
#define RG_MERGE_RUN 4
#endif
enum { MERGE_RUN = RG_MERGE_RUN };   /* entries of a list a lane reads ahead */
/* One lane's share of a pair of face lists.  A lane reads MERGE_RUN entries of each block of 32 x MERGE_RUN before it unites any:
   the entry loads of a block are in flight together instead of two per round trip to L2.  consec: the lane's entries are
   neighbours in the list, else 32 apart. */
RG_FN void merge_list(const Params &P, const uint16_t *ea, const uint16_t *eb, uint32_t cnt, uint32_t lane, bool consec, uint32_t base_a, uint32_t base_b)
{
    /* faces inside a solid list the same pair of nodes hundreds of times: a lane skips a pair it has just united.  A pair is one
       word here (entry of this tile | entry of the neighbour << 16); the last two are kept: C3's faces alternate corridor / background */
    uint32_t p = 0xFFFFFFFFu, q = 0xFFFFFFFFu;
    const uint32_t step = consec ? 1u : (uint32_t)MERGE_LANES;
    for (uint32_t i0 = consec ? lane * MERGE_RUN : lane; i0 < cnt; i0 += MERGE_LANES * MERGE_RUN) {
        uint32_t key[MERGE_RUN];
#if RG_DEVICE_BUILD
#pragma unroll
#endif
        for (uint32_t k = 0; k < (uint32_t)MERGE_RUN; k++) {
            const uint32_t i = i0 + k * step;
            key[k] = i < cnt ? (uint32_t)ea[i] | (uint32_t)eb[i] << 16 : 0xFFFFFFFFu;
        }
#if RG_DEVICE_BUILD
#pragma unroll
#endif
        for (uint32_t k = 0; k < (uint32_t)MERGE_RUN; k++) {
            const uint32_t w = key[k];
            if ((w & (w >> 16) & ENT_LEAF) || w == p || w == q) continue;
            unite32(P.nparent, base_a + (w & (ENT_LEAF - 1u)), base_b + ((w >> 16) & (ENT_LEAF - 1u)));
            q = p; p = w;
        }
    }
}

template <class C, class Exec>
RG_FN void merge_faces(Exec &ex, const Params &P)
{
    const uint32_t ntiles = (uint32_t)P.ntx * (uint32_t)P.nty * (uint32_t)P.ntz;
    ex.gpfor((int64_t)ntiles * MERGE_LANES, [&](int64_t it0) {
        const uint32_t lane = (uint32_t)(it0 % MERGE_LANES);
        const uint32_t tile = (uint32_t)(it0 / MERGE_LANES);
        const TileRec ra = P.rec[tile];
        if (!(ra.cnt[F_XP] | ra.cnt[F_YP] | ra.cnt[F_ZP])) return;
        int tx, ty, tz; tile_origin<C>(P, tile, tx, ty, tz);
        /* Which entries a lane takes.  A tile of constant rows (rg_rows.h) touches its neighbours over whole rows: neighbouring
           entries repeat a pair, and a lane that reads consecutive entries drops the repeats with its pair cache (C3 merge
           0.38 -> 0.26 ms).  Elsewhere edges are unrelated (C4's noise, the surface of a sphere): every 32nd entry, so that the
           lanes of a warp walk neighbouring parent words together (consecutive entries cost C4 0.47 -> 0.63 ms, C1 0.03 -> 0.05). */
        const bool consec = ra.uniform == 2;
        uint32_t offa = 0;                                       /* this tile's lists lie in face order: -x +x -y +y -z +z */
        for (int axis = 0; axis < 3; axis++) {
            const int fa = 2 * axis + 1, fb = 2 * axis;
            offa += ra.cnt[fb];
            const uint32_t cnt = ra.cnt[fa];
            const uint32_t here = offa; offa += cnt;
            if (!cnt) continue;
            uint32_t nb;
            if (axis == 0) { if (tx + 1 >= P.ntx) continue; nb = tile + 1; }
            else if (axis == 1) { if (ty + 1 >= P.nty) continue; nb = tile + (uint32_t)P.ntx; }
            else { if (tz + 1 >= P.ntz) continue; nb = tile + (uint32_t)P.ntx * (uint32_t)P.nty; }
            const TileRec rb = P.rec[nb];
            if (rb.cnt[fb] != cnt) {
                if (lane == 0) { atomic_or(&P.counters[CTR_STATUS], ST_FACE_MISMATCH); atomic_add(&P.counters[CTR_FACE_MISMATCH], 1u); }
                continue;
            }
            uint32_t offb = 0;
            for (int f = 0; f < fb; f++) offb += rb.cnt[f];
            merge_list(P, P.ent + (size_t)tile * C::ENT_MAX + here, P.ent + (size_t)nb * C::ENT_MAX + offb, cnt, lane, consec, ra.node_base, rb.node_base);
        }
    });
}

/* ================================================================ pass C */
/* one 16-byte chunk of the tile's part of the mask (item = chunk index inside the tile) */
template <class C>
RG_FN void fill_chunk(const Params &P, int64_t x0, int64_t y0, int64_t z0, uint32_t byte01, int it)
{
    const uint32_t w = byte01 * 0x01010101u;
    int c = it % C::CX, r = it / C::CX, y = r % C::TY, z = r / C::TY;
    int64_t gx = x0 + 16 * c, gy = y0 + y, gz = z0 + z;
    if (gy >= P.ny || gz >= P.nz || gx >= P.nx) return;
    uint8_t *g = P.mask + (gz * P.ny + gy) * P.nx + gx;
    if (P.fast16) st128_stream(g, u32x4{w, w, w, w});
    else for (int b = 0; b < 16 && gx + b < P.nx; b++) g[b] = (uint8_t)byte01;
}
template <class C, class Exec>
RG_FN void fill_tile(Exec &ex, const Params &P, int64_t x0, int64_t y0, int64_t z0, uint32_t byte01)
{
    ex.pfor(C::TZ * C::TY * C::CX, [&](int it) { fill_chunk<C>(P, x0, y0, z0, byte01, it); });
}
template <class C, class Exec>
RG_FN void zero_tile(Exec &ex, const Params &P, int64_t x0, int64_t y0, int64_t z0) { fill_tile<C>(ex, P, x0, y0, z0, 0u); }

/* shared memory of pass C: staged tile, edge planes, region bits */
template <class C>
struct alignas(128) FloodSmem {
    union alignas(128) U {
        uint8_t img[C::IMG_BYTES];                 /* staged tile: dead once the link planes exist */
        StridePlanes<C> sp;                        /* built (if at all) after the first unit-stride sweeps */
    } u;
    uint64_t mbar;
    alignas(16) uint32_t ex[C::NW];
    alignas(16) uint32_t ey[C::NW];
    alignas(16) uint32_t ez[C::NW];
    alignas(16) uint32_t M[C::NW];                 /* bit i: voxel i is in the region */
    uint32_t scal[16];
};

/* all the shared memory the output of a row tile needs (k_rows_out: many light CTAs per SM) */
template <class C>
struct RowsOutSmem {
    uint32_t ex[2 * C::TZ + 2];
    uint32_t M[C::TZ + 1];
    uint32_t scal[16];
};

/* pass C for a tile of constant rows (rg_rows.h): the flood runs over TY x TZ row bits -- word z holds the rows of plane z,
   its links along y are the bits of ly[z] (closed with the same carry trick as a voxel word), lz[z] joins planes z and z+1 */
template <class C, class S, class Exec>
RG_FN void tile_pass_c_rows(Exec &ex, const Params &P, S &s, uint32_t tile, const TileRec &rc, int64_t x0, int64_t y0, int64_t z0)
{
    uint32_t *ly = s.ex, *lz = s.ex + C::TZ, *M = s.M;
    const uint32_t *src = P.planes + (size_t)(rc.plane_slot1 - 1u) * 3 * C::NW;
    ex.pfor(2 * C::TZ, [&](int i) { ly[i] = src[i]; if (i < C::TZ) M[i] = 0; });
    if (ex.tid() == 0) s.scal[SC_ANY] = 0;
    ex.sync();
    ex.pfor((int)rc.nnodes, [&](int i) {
        const uint32_t g = rc.node_base + (uint32_t)i;
        if ((P.debug_stop == 31 || P.debug_stop == 32) ? (g & 1u) != 0u : find32(P.nparent, g) == 0u) {      /* tuning: 31 / 32 time the output without the finds */
            const uint32_t r = (uint32_t)P.nrep[g] / (uint32_t)C::TX;
            atomic_or(&M[r / C::TY], 1u << (r % C::TY));
            s.scal[SC_ANY] = 1;
        }
    });
    ex.sync();
    if (ex.tid() == 0) {                                          /* the flag pass leaves row tiles to this function */
        P.rec[tile].has_region = s.scal[SC_ANY] ? 1 : 0;
        if (s.scal[SC_ANY]) atomic_add(&P.counters[CTR_NROWOUT], 1u);
    }
    if (!s.scal[SC_ANY]) { if (!P.mask_prezeroed) zero_tile<C>(ex, P, x0, y0, z0); return; }
    for (;;) {
        bool changed = false;
        ex.pfor(C::TZ, [&](int z) {
            const uint32_t m = M[z];
            uint32_t in = m;
            if (z > 0) in |= M[z - 1] & lz[z - 1];
            if (z < C::TZ - 1) in |= M[z + 1] & lz[z];
            if (in) { const uint32_t out = fill_word(in, ly[z]); if (out != m) { M[z] = out; changed = true; } }
        });
        if (!ex.sync_or(changed)) break;
    }
    if (P.debug_stop == 30 || P.debug_stop == 32) return;              /* tuning: ... and without the stores */
    /* Stores alone drain at 3.9 TB/s on this part (a copy moves 6.6: half of it reads), and the labelling pass has cleared the
       tile already: only the rows of ones are written (C3: a quarter of the bytes, row output 0.35 -> ... ms). */
    const bool ones_only = P.mask_prezeroed != 0;
    ex.pfor(C::TZ * C::TY * C::CX, [&](int it) {
        const int r = it / C::CX, y = r % C::TY, z = r / C::TY;
        const uint32_t bit = (M[z] >> y) & 1u;
        if (bit | (uint32_t)!ones_only) fill_chunk<C>(P, x0, y0, z0, bit, it);
    });
}

template <class C, class Exec>
RG_FN void tile_pass_c(Exec &ex, const Params &P, uint32_t tile, FloodSmem<C> &s)
{
    /* Pass C needs no labels, only which voxels hang together with a selected component: start from the
       representatives of the selected nodes and flood along the edge bit planes, 32 voxels per operation,
       until nothing changes.  (The first version re-ran the union-find here; on tiles with one huge
       component -- C3's background -- its hooks contend for a single root.) */
    int tx, ty, tz; tile_origin<C>(P, tile, tx, ty, tz);
    const int64_t x0 = (int64_t)tx * C::TX, y0 = (int64_t)ty * C::TY, z0 = (int64_t)tz * C::TZ;
    const TileRec rc = P.rec[tile];

    if (rc.nnodes == 0) { zero_tile<C>(ex, P, x0, y0, z0); return; }
    if (rc.uniform == 2) { tile_pass_c_rows<C>(ex, P, s, tile, rc, x0, y0, z0); return; }
    const bool saved = P.planes && rc.plane_slot1;                  /* the full pass kept this tile's link planes */
    if (saved) {
        const uint32_t *src = P.planes + (size_t)(rc.plane_slot1 - 1u) * 3 * C::NW;
        ex.pfor(3 * C::NW / 4, [&](int i) {
            const int pl = i / (C::NW / 4), k = i % (C::NW / 4);
            uint32_t *dst = pl == 0 ? s.ex : pl == 1 ? s.ey : s.ez;
            st128(dst + 4 * k, ld128(src + (size_t)pl * C::NW + 4 * k));
        });
    } else if (!ex.template tma_tile<C>(P, s.u.img, &s.mbar, x0, y0, z0)) load_tile<C>(ex, P, s.u.img, x0, y0, z0, false);
    ex.pfor(C::NW / 4, [&](int i) { st128(&s.M[4 * i], u32x4{0, 0, 0, 0}); });
    if (ex.tid() == 0) s.scal[SC_ANY] = 0;
    ex.sync();
    ex.pfor((int)rc.nnodes, [&](int i) {
        uint32_t g = rc.node_base + (uint32_t)i;
        if (find32(P.nparent, g) == 0u) {
            uint32_t r = P.nrep[g];
            atomic_or(&s.M[r >> 5], 1u << (r & 31));
            s.scal[SC_ANY] = 1;
        }
    });
    ex.tile_wait(&s.mbar);
    ex.sync();
    if (!s.scal[SC_ANY]) { zero_tile<C>(ex, P, x0, y0, z0); return; }
    if (P.debug_stop == 20) return;
    if (!saved) {
        build_edges<C, false>(ex, P, s.u.img, s.ex, s.ey, s.ez, (uint32_t *)0, x0, y0, z0);
        ex.sync();
    }
    if (P.debug_stop == 21) return;
    /* A region advances a whole run per sweep in x but one row / plane per sweep in y and z, so a tile that is still
       changing after a few sweeps (a solid: up to TY + TZ of them) gets stride planes and continues 1, 2, 4, ...
       rows and planes at a time. */
    bool more = true;
    for (int i = 0; i < FLOOD_PLAIN_SWEEPS && more; i++) more = flood_sweep<C>(ex, s.M, s.ex, s.ey, s.ez, s.u.sp, false);
    if (more) {
        build_stride_planes<C>(ex, s.ey, s.ez, s.u.sp);
        while (flood_sweep<C>(ex, s.M, s.ex, s.ey, s.ez, s.u.sp, true)) { }
    }
    if (P.debug_stop == 22) return;
    ex.pfor(C::TZ * C::TY * C::CX, [&](int it) {
        int c = it % C::CX, r = it / C::CX, y = r % C::TY, z = r / C::TY;
        int64_t gx = x0 + 16 * c, gy = y0 + y, gz = z0 + z;
        if (gy >= P.ny || gz >= P.nz || gx >= P.nx) return;
        uint32_t bits = (s.M[(z * C::TY + y) * C::WX + (c >> 1)] >> ((c & 1) * 16)) & 0xFFFFu;
        if (!bits && P.mask_prezeroed) return;                       /* cleared by the labelling pass: zeros are not stored twice */
        u32x4 v = { expand4(bits & 15u), expand4((bits >> 4) & 15u), expand4((bits >> 8) & 15u), expand4(bits >> 12) };
        uint8_t *g = P.mask + (gz * P.ny + gy) * P.nx + gx;
        if (P.fast16) st128_stream(g, v);
        else {
            uint32_t t[4] = { v.x, v.y, v.z, v.w };
            for (int b = 0; b < 16 && gx + b < P.nx; b++) g[b] = (uint8_t)((t[b >> 2] >> (8 * (b & 3))) & 0xFFu);
        }
    });
}

/* ================================================================ seeds -> sortable keys */
template <class C, class Exec>
RG_FN void make_seed_keys(Exec &ex, const Params &P, const int32_t *seeds_xyz, uint64_t *keys)
{
    ex.gpfor((int64_t)P.nseeds, [&](int64_t i) {
        int64_t x = seeds_xyz[3 * i], y = seeds_xyz[3 * i + 1], z = seeds_xyz[3 * i + 2];
        uint64_t key = ~0ull;                                   /* out of range: dropped (region.c:485) */
        if (x >= 0 && x < P.nx && y >= 0 && y < P.ny && z >= 0 && z < P.nz) {
            int64_t tx = x / C::TX, ty = y / C::TY, tz = z / C::TZ;
            uint64_t tile = (uint64_t)((tz * P.nty + ty) * P.ntx + tx);
            uint32_t v = (uint32_t)(((z - tz * C::TZ) * C::TY + (y - ty * C::TY)) * C::TX + (x - tx * C::TX));
            key = (tile << 16) | v;
        }
        keys[i] = key;
    });
}

/* resets the per-call counters; node 0 is the "holds a seed" super node */
enum { FAST_NODE_SLOTS = 1024 };    /* node ids 1 + tile * 1024 .. belong to the tile's fast pass (rg_fast.h) */
RG_FN void reset_counters(const Params &P)
{
    for (int i = 0; i < CTR_WORDS; i++) P.counters[i] = 0;
    /* node 0, then one fixed block per tile for the fast pass; the full pass allocates behind them */
    P.counters[CTR_NODES] = 1u + (uint32_t)P.ntx * (uint32_t)P.nty * (uint32_t)P.ntz * (uint32_t)FAST_NODE_SLOTS;
    P.nparent[0] = 0;
}

} // namespace rg
