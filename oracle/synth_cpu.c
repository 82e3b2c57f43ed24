/*
 * oracle/synth_cpu.c -- CPU definitions of the synthetic benchmark inputs (C1h, C2, C3, C4).
 * TEST INFRASTRUCTURE ONLY.  The product has its own on-device generators
 * (parallel-c-programs_b200/csrc/rg_synth.cu); tests compare the two byte for byte.
 *
 * Geometry of C1/C4 follows the reference's func() (5-cuda-region-growing/raycast.cu:114-144):
 * noise 0..19, sphere r100 @(300,400,100)=30, sphere r50 @(100,200,400)=50,
 * box (200..300,300..500,200..300)=45, box (0..100,250..400,250..400)=35 (exclusive
 * bounds), all scaled by n/512.  The noise source is splitmix64 of the linear voxel
 * index instead of glibc rand(), so that it can be evaluated
 * in parallel.  Squared integer distances replace sqrt(): d < r  <=>  d*d < r*r.
 */
#include <stdint.h>
#include <stdlib.h>

static inline uint64_t splitmix64(uint64_t x)
{
    uint64_t z = x + 0x9e3779b97f4a7c15ULL;
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ULL;
    z = (z ^ (z >> 27)) * 0x94d049bb133111ebULL;
    return z ^ (z >> 31);
}
static inline uint8_t noise20(uint64_t seed, uint64_t idx)
{
    uint64_t h = splitmix64(seed + idx);
    return (uint8_t)(((h >> 32) * 20ULL) >> 32);          /* uniform 0..19 */
}

static inline int in_sphere(int64_t x, int64_t y, int64_t z, int64_t cx, int64_t cy, int64_t cz, int64_t r)
{
    int64_t dx = x - cx, dy = y - cy, dz = z - cz;
    return dx * dx + dy * dy + dz * dz < r * r;
}

/* exercise geometry on an n^3 cube; returns 0 if the voxel is noise */
static inline uint8_t exercise_solid(int64_t x, int64_t y, int64_t z, int64_t n)
{
#define S(v) ((int64_t)(v) * n / 512)
    uint8_t value = 0;
    if (in_sphere(x, y, z, S(300), S(400), S(100), S(100))) value = 30;
    if (in_sphere(x, y, z, S(100), S(200), S(400), S(50))) value = 50;
    if (x > S(200) && x < S(300) && y > S(300) && y < S(500) && z > S(200) && z < S(300)) value = 45;
    if (x > S(0) && x < S(100) && y > S(250) && y < S(400) && z > S(250) && z < S(400)) value = 35;
#undef S
    return value;
}

/* C1h: exercise geometry with hash noise */
void oracle_synth_exercise(uint8_t *out, int64_t n, uint64_t seed)
{
    for (int64_t z = 0; z < n; z++)
        for (int64_t y = 0; y < n; y++)
            for (int64_t x = 0; x < n; x++) {
                uint64_t i = (uint64_t)((z * n + y) * n + x);
                uint8_t s = exercise_solid(x, y, z, n);
                out[i] = s ? s : noise20(seed, i);
            }
}

/*
 * C4: exercise geometry + a z-serpentine of value 200 in the plane y = Y0 = n/64:
 * NCOL columns x = X0 + 2c (X0 = n/64) run the full z range; consecutive columns are
 * joined by one voxel at x = X0 + 2c + 1, alternately at z = n-1 and z = 0.
 * It crosses every z-slab interface NCOL times.
 */
#define C4_NCOL 64
static inline int c4_serpentine(int64_t x, int64_t y, int64_t z, int64_t n)
{
    int64_t Y0 = n / 64, X0 = n / 64;
    if (y != Y0) return 0;
    int64_t d = x - X0;
    if (d < 0 || d > 2 * (C4_NCOL - 1)) return 0;
    if ((d & 1) == 0) return 1;
    int64_t j = (d - 1) / 2;                               /* connector between column j and j+1 */
    return (j & 1) == 0 ? (z == n - 1) : (z == 0);
}

void oracle_synth_c4(uint8_t *out, int64_t n, uint64_t seed)
{
    for (int64_t z = 0; z < n; z++)
        for (int64_t y = 0; y < n; y++)
            for (int64_t x = 0; x < n; x++) {
                uint64_t i = (uint64_t)((z * n + y) * n + x);
                uint8_t s = exercise_solid(x, y, z, n);
                if (c4_serpentine(x, y, z, n)) s = 200;
                out[i] = s ? s : noise20(seed, i);
            }
}

/* 16 seeds of C4: 8 in solids / the serpentine, 8 at hashed noise positions. out = 16*(x,y,z) */
void oracle_seeds_c4(int64_t n, uint64_t seed, int32_t *out)
{
#define S(v) ((int32_t)((int64_t)(v) * n / 512))
    int32_t s[8][3] = {
        { S(300), S(400), S(100) }, { S(300), S(350), S(60) },      /* sphere 30 (twice) */
        { S(100), S(200), S(400) }, { S(110), S(190), S(410) },     /* sphere 50 (twice) */
        { S(250), S(400), S(250) }, { S(210), S(480), S(290) },     /* box 45 (twice)    */
        { S(50), S(300), S(300) },                                   /* box 35 = the exercise seed */
        { (int32_t)(n / 64), (int32_t)(n / 64), 0 },                 /* start of the serpentine */
    };
#undef S
    for (int k = 0; k < 8; k++) { out[3 * k] = s[k][0]; out[3 * k + 1] = s[k][1]; out[3 * k + 2] = s[k][2]; }
    for (int k = 8; k < 16; k++) {                                    /* noise seeds: upper y half is solid-free below y=150/512 */
        uint64_t h = splitmix64(seed ^ (0xABCD0000ULL + (uint64_t)k));
        out[3 * k]     = (int32_t)(h % (uint64_t)n);
        out[3 * k + 1] = (int32_t)((n / 32) + (h >> 20) % (uint64_t)(n / 8));   /* y in [n/32, n/32+n/8): no solids, not Y0 */
        out[3 * k + 2] = (int32_t)((h >> 40) % (uint64_t)n);
    }
}

/*
 * C2: n x n image, cells of c = n/8.  Rows of cells with j % 4 == 0 hold a horizontal bar
 * |dy| < c/4 of value 40 spanning the whole cell width (the 8 bars of a row touch and form ONE
 * region holding 8 seeds); other cells hold a disc 25(dx^2+dy^2) < 4c^2 (radius 0.4c) of value
 * 50 + 5*((i+j)%3).  Everything else is hash noise 0..19.
 */
static inline uint8_t c2_solid(int64_t x, int64_t y, int64_t n)
{
    int64_t c = n / 8; if (c < 1) c = 1;
    int64_t i = x / c, j = y / c;
    if (i > 7 || j > 7) return 0;
    int64_t dx = x - i * c - c / 2, dy = y - j * c - c / 2;
    if (j % 4 == 0) { int64_t a = dy < 0 ? -dy : dy; return (4 * a < c) ? 40 : 0; }
    return (25 * (dx * dx + dy * dy) < 4 * c * c) ? (uint8_t)(50 + 5 * ((i + j) % 3)) : 0;
}

void oracle_synth_c2(uint8_t *out, int64_t n, uint64_t seed)
{
    for (int64_t y = 0; y < n; y++)
        for (int64_t x = 0; x < n; x++) {
            uint64_t i = (uint64_t)(y * n + x);
            uint8_t s = c2_solid(x, y, n);
            out[i] = s ? s : noise20(seed, i);
        }
}

void oracle_seeds_c2(int64_t n, int32_t *out /* 64*(x,y,z) */)
{
    int64_t c = n / 8;
    for (int j = 0; j < 8; j++)
        for (int i = 0; i < 8; i++) {
            int k = j * 8 + i;
            out[3 * k] = (int32_t)(i * c + c / 2); out[3 * k + 1] = (int32_t)(j * c + c / 2); out[3 * k + 2] = 0;
        }
}

/*
 * C3: n^3 serpentine.  Background 0, corridor 200, one voxel wide with one-voxel walls:
 * even z, even y: whole x row; even z, odd y (between two rows): one joint at x = n-1 or 0
 * alternating; odd z (between two planes): one joint alternating between the end of the
 * last row and (0,0).  The corridor is a single path of about n^3/4 voxels.
 */
static inline int c3_corridor(int64_t x, int64_t y, int64_t z, int64_t n)
{
    int64_t R = (n + 1) / 2;                               /* rows / planes used: 0,2,..,2(R-1) */
    int64_t last = 2 * (R - 1);
    if ((z & 1) == 0) {
        if ((y & 1) == 0) return 1;
        if (y + 1 > last) return 0;
        return x == ((((y - 1) / 2) & 1) == 0 ? n - 1 : 0);
    }
    if (z + 1 > last) return 0;
    int64_t ex = (((R - 1) & 1) == 0) ? n - 1 : 0, ey = last;   /* far end of the last row */
    if ((((z - 1) / 2) & 1) == 0) return x == ex && y == ey;
    return x == 0 && y == 0;
}

void oracle_synth_c3(uint8_t *out, int64_t n)
{
    for (int64_t z = 0; z < n; z++)
        for (int64_t y = 0; y < n; y++)
            for (int64_t x = 0; x < n; x++)
                out[(z * n + y) * n + x] = c3_corridor(x, y, z, n) ? 200 : 0;
}

/* number of corridor voxels of C3, computed without building the volume */
int64_t oracle_c3_count(int64_t n)
{
    int64_t R = (n + 1) / 2;
    int64_t per_plane = R * n + (R - 1);                  /* rows + joints */
    return R * per_plane + (R - 1);                        /* planes + inter-plane joints */
}
