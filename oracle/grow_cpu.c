/*
 * oracle/grow_cpu.c -- CPU restatement of the reference's seeded region growing.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is linked into, loaded by or
 * called from the product library (libb200rg.so).  Only tests/, bench.py's
 * cpu_baseline / --impl reference legs and __graft_entry__.smoke() may use it,
 * and only as the checker / reported CPU baseline.
 *
 * Parity status: PINNED.  tests/test_oracle_pins.py checks this file against
 *   - the committed golden output of the reference (2-mpi-region-growing/out.bmp
 *     payload, 64 420 px) and pic1..pic4 masks produced by the UNMODIFIED
 *     region.c built in oracle/_ref (see oracle/Makefile, oracle/mpi_shim/),
 *   - the UNMODIFIED grow_region_serial()/create_data() extracted at build time
 *     from 6-opencl-region-growing/raycast.c (plain-C twin of raycast.cu:281-318),
 *     2 197 899 voxels on the reference's own 512^3 volume.
 *
 * What it restates (reference file:line, relative to /root/reference):
 *   oracle_similar     <- region.c:76-83  (T=2)   raycast.cu:271-277 (T=1)
 *   oracle_grow (3-D)  <- raycast.cu:281-318      (6-conn, order -x,+x,-y,+y,-z,+z,
 *                                                  seed pre-marked, mark on push)
 *   oracle_grow (2-D)  <- region.c:493-527        (4-conn, order +y,-y,+x,-x,
 *                                                  mark on pop and on push)
 *   oracle_inside      <- raycast.cu:169-175, region.c:444-446
 *   stack              <- raycast.cu:29-60, region.c:12-71
 *   oracle_apply_mask  <- region.c:572-580  (image[i] *= (region[i]==0))
 *   oracle_add_seeds_2d<- region.c:450-490  (global 0-based corner seeds)
 *   oracle_create_data_ref <- raycast.cu:114-158  (glibc rand(), unseeded)
 * Differences from the reference, on purpose: sizes, seeds and threshold are
 * run-time parameters; linear indices are 64-bit (reference `int index()`
 * overflows at 2048^3, raycast.cu:178-180); out-of-range seeds are dropped
 * (region.c:485 `if(inside(seed))`); mask bytes are exactly 0/1.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>

typedef struct { int32_t x, y, z; } ovox_t;

typedef struct {
    int64_t size, cap;
    ovox_t *v;
} ostack_t;

static void st_init(ostack_t *s) {
    s->size = 0; s->cap = 1024;                       /* raycast.cu:37-38 */
    s->v = (ovox_t *)malloc(sizeof(ovox_t) * (size_t)s->cap);
}
static void st_push(ostack_t *s, ovox_t p) {
    if (s->size == s->cap) {                          /* doubling, raycast.cu:45-51 */
        s->cap *= 2;
        s->v = (ovox_t *)realloc(s->v, sizeof(ovox_t) * (size_t)s->cap);
    }
    s->v[s->size++] = p;
}
static ovox_t st_pop(ostack_t *s) { return s->v[--s->size]; }

static inline int oracle_inside(ovox_t p, int64_t nx, int64_t ny, int64_t nz) {
    return p.x >= 0 && p.x < nx && p.y >= 0 && p.y < ny && p.z >= 0 && p.z < nz;
}
static inline int64_t oidx(ovox_t p, int64_t nx, int64_t ny) {
    return ((int64_t)p.z * ny + p.y) * nx + p.x;       /* index(z,y,x), raycast.cu:178 */
}
static inline int oracle_similar(const uint8_t *im, int64_t a, int64_t b, int T) {
    int va = im[a], vb = im[b];
    return abs(va - vb) < T;                          /* strict <, promoted ints */
}

/*
 * conn = 6: 3-D order of raycast.cu:290-292.  conn = 4: 2-D order of region.c:506
 * (nz must be 1).  The result is order-independent; the order
 * is kept so the restatement reads like the reference.
 * Returns the number of region voxels, or -1 on bad arguments.
 */
int64_t oracle_grow(const uint8_t *image, int64_t nx, int64_t ny, int64_t nz,
                    const int32_t *seeds_xyz, int nseeds, int threshold, int conn,
                    uint8_t *mask)
{
    static const int dx6[6] = {-1, 1, 0, 0, 0, 0};
    static const int dy6[6] = {0, 0, -1, 1, 0, 0};
    static const int dz6[6] = {0, 0, 0, 0, -1, 1};
    static const int dx4[4] = {0, 0, 1, -1};
    static const int dy4[4] = {1, -1, 0, 0};
    if (nx <= 0 || ny <= 0 || nz <= 0 || !mask || !image) return -1;
    if (conn != 4 && conn != 6) return -1;
    if (conn == 4 && nz != 1) return -1;
    memset(mask, 0, (size_t)(nx * ny * nz));

    ostack_t st; st_init(&st);
    int64_t count = 0;
    for (int s = 0; s < nseeds; s++) {
        ovox_t p = { seeds_xyz[3 * s], seeds_xyz[3 * s + 1], seeds_xyz[3 * s + 2] };
        if (!oracle_inside(p, nx, ny, nz)) continue;   /* region.c:485 */
        int64_t i = oidx(p, nx, ny);
        if (mask[i]) continue;                         /* duplicate seed */
        mask[i] = 1; count++;                          /* raycast.cu:288 */
        st_push(&st, p);
    }
    const int nn = conn;
    while (st.size > 0) {
        ovox_t p = st_pop(&st);
        int64_t ip = oidx(p, nx, ny);
        for (int n = 0; n < nn; n++) {
            ovox_t c = p;
            if (conn == 6) { c.x += dx6[n]; c.y += dy6[n]; c.z += dz6[n]; }
            else           { c.x += dx4[n]; c.y += dy4[n]; }
            if (!oracle_inside(c, nx, ny, nz)) continue;
            int64_t ic = oidx(c, nx, ny);
            if (mask[ic]) continue;
            if (oracle_similar(image, ip, ic, threshold)) {
                mask[ic] = 1; count++;
                st_push(&st, c);
            }
        }
    }
    free(st.v);
    return count;
}

/* region.c:572-580 */
void oracle_apply_mask(uint8_t *image, const uint8_t *mask, int64_t n)
{
    for (int64_t i = 0; i < n; i++) image[i] *= (mask[i] == 0);
}

/*
 * region.c:450-490 -- the four corner seeds, translated to global 0-based (x,y)
 * of a w x h image: local 1-based (5,5),(ly-5,5),(ly-5,lx-5),(5,lx-5) on the four
 * corner ranks = global (4,4),(w-6,4),(w-6,h-6),(4,h-6).  out = 4*(x,y,z).
 */
void oracle_add_seeds_2d(int w, int h, int32_t *out)
{
    int32_t s[12] = { 4, 4, 0,  w - 6, 4, 0,  w - 6, h - 6, 0,  4, h - 6, 0 };
    memcpy(out, s, sizeof s);
}

/* raycast.cu:114-144, with dim 512 and rand() drawn once per voxel in z,y,x loop order */
static unsigned char ref_func(int x, int y, int z)
{
    unsigned char value = rand() % 20;
    int x1 = 300, y1 = 400, z1 = 100;
    float dist = sqrt((x - x1) * (x - x1) + (y - y1) * (y - y1) + (z - z1) * (z - z1));
    if (dist < 100) value = 30;
    x1 = 100; y1 = 200; z1 = 400;
    dist = sqrt((x - x1) * (x - x1) + (y - y1) * (y - y1) + (z - z1) * (z - z1));
    if (dist < 50) value = 50;
    if (x > 200 && x < 300 && y > 300 && y < 500 && z > 200 && z < 300) value = 45;
    if (x > 0 && x < 100 && y > 250 && y < 400 && z > 250 && z < 400) value = 35;
    return value;
}

/* raycast.cu:146-158.  Calls srand(1) first: the state an unseeded program starts in. */
void oracle_create_data_ref(uint8_t *data /* 512^3 */)
{
    const int D = 512;
    srand(1);
    for (int i = 0; i < D; i++)
        for (int j = 0; j < D; j++)
            for (int k = 0; k < D; k++)
                data[(size_t)i * D * D + (size_t)j * D + k] = ref_func(k, j, i);
}

uint64_t oracle_fnv1a64(const uint8_t *p, int64_t n)
{
    uint64_t h = 0xcbf29ce484222325ULL;
    for (int64_t i = 0; i < n; i++) { h ^= p[i]; h *= 0x100000001b3ULL; }
    return h;
}

int64_t oracle_count(const uint8_t *p, int64_t n)
{
    int64_t c = 0;
    for (int64_t i = 0; i < n; i++) c += (p[i] != 0);
    return c;
}
