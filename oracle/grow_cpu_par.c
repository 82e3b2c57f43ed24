/*
 * oracle/grow_cpu_par.c -- multi-threaded CPU port of the reference's DISTRIBUTED grower.
 * TEST INFRASTRUCTURE ONLY (see oracle/grow_cpu.c header); used as the
 * "all host cores" CPU baseline in bench.py (cpu_baseline.kind = "port").
 *
 * Follows the round structure of 2-mpi-region-growing/region.c:493-533:
 *   while(!finished(local_finish)) {            region.c:497  Allreduce(MIN)   -> barrier + shared flag
 *       local stack flood-fill to fixpoint      region.c:499-527
 *       exchange(stack)                         region.c:250-353 (mask borders) -> neighbours' boundary planes are
 *                                                                                 read directly from shared memory
 *       add_halo_to_stack(stack)                region.c:355-387 -> every set neighbour-border voxel re-seeds the local fill
 *       MPI_Barrier                             region.c:531
 *   }
 * with the 2-D Cartesian rank grid replaced by slabs along the slowest axis
 * (z for volumes, y for images) and MPI ranks by pthreads.  Result is identical
 * to oracle_grow (checked in tests/test_oracle_pins.py).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <pthread.h>

typedef struct { int32_t x, y, z; } pvox_t;
typedef struct { int64_t size, cap; pvox_t *v; } pstack_t;

static void ps_push(pstack_t *s, pvox_t p) {
    if (s->size == s->cap) { s->cap *= 2; s->v = (pvox_t *)realloc(s->v, sizeof(pvox_t) * (size_t)s->cap); }
    s->v[s->size++] = p;
}

typedef struct {
    const uint8_t *image; uint8_t *mask;
    int64_t nx, ny, nz;          /* nz == 1: slabs along y */
    const int32_t *seeds; int nseeds; int T;
    int nthreads;
    pthread_barrier_t bar;
    volatile int progress[2];    /* double-buffered "somebody progressed" word = finished() */
    int64_t *counts;
    int rounds;
} pjob_t;

typedef struct { pjob_t *job; int tid; } parg_t;

static inline int psimilar(const uint8_t *im, int64_t a, int64_t b, int T) {
    int d = (int)im[a] - (int)im[b]; if (d < 0) d = -d; return d < T;
}

static void *worker(void *argp)
{
    parg_t *arg = (parg_t *)argp; pjob_t *J = arg->job; const int t = arg->tid;
    const int64_t nx = J->nx, ny = J->ny, nz = J->nz;
    const int along_y = (nz == 1);
    const int64_t L = along_y ? ny : nz;                 /* slab axis length */
    const int64_t lo = L * t / J->nthreads, hi = L * (t + 1) / J->nthreads;
    const int64_t sx = 1, sy = nx, sz = nx * ny;
    const uint8_t *im = J->image; uint8_t *mk = J->mask;
    pstack_t st = { 0, 1024, (pvox_t *)malloc(sizeof(pvox_t) * 1024) };
    int64_t count = 0;

    /* add_seeds, region.c:450-490: only the owning rank takes a seed */
    for (int s = 0; s < J->nseeds; s++) {
        pvox_t p = { J->seeds[3 * s], J->seeds[3 * s + 1], J->seeds[3 * s + 2] };
        if (p.x < 0 || p.x >= nx || p.y < 0 || p.y >= ny || p.z < 0 || p.z >= nz) continue;
        int64_t c = along_y ? p.y : p.z;
        if (c < lo || c >= hi) continue;
        int64_t i = p.z * sz + p.y * sy + p.x * sx;
        if (!mk[i]) { mk[i] = 1; count++; ps_push(&st, p); }
    }
    static const int dx[6] = {-1, 1, 0, 0, 0, 0}, dy[6] = {0, 0, -1, 1, 0, 0}, dz[6] = {0, 0, 0, 0, -1, 1};
    const int nn = along_y ? 4 : 6;
    int round = 0;
    for (;;) {
        int local_progress = 0;
        /* local fill to fixpoint, region.c:499-527 */
        while (st.size > 0) {
            pvox_t p = st.v[--st.size];
            int64_t ip = p.z * sz + p.y * sy + p.x;
            for (int n = 0; n < nn; n++) {
                pvox_t c = { p.x + dx[n], p.y + dy[n], p.z + dz[n] };
                if (c.x < 0 || c.x >= nx || c.y < 0 || c.y >= ny || c.z < 0 || c.z >= nz) continue;
                int64_t cs = along_y ? c.y : c.z;
                if (cs < lo || cs >= hi) continue;          /* inside(): own slab only */
                int64_t ic = c.z * sz + c.y * sy + c.x;
                if (mk[ic]) continue;
                if (psimilar(im, ip, ic, J->T)) { mk[ic] = 1; count++; ps_push(&st, c); local_progress = 1; }
            }
        }
        pthread_barrier_wait(&J->bar);                      /* all local fills done: borders are stable */
        /* exchange + add_halo_to_stack: every set voxel of the neighbour's border plane
           tries to grow across the interface into this slab */
        for (int side = 0; side < 2; side++) {
            int64_t nb = side ? hi : lo - 1;                /* neighbour's border coordinate */
            int64_t me = side ? hi - 1 : lo;                /* my border coordinate */
            if (nb < 0 || nb >= L || hi <= lo) continue;
            int64_t plane = along_y ? nx : nx * ny;
            for (int64_t q = 0; q < plane; q++) {
                int64_t inb = along_y ? nb * sy + q : nb * sz + q;
                if (!mk[inb]) continue;
                int64_t ime = along_y ? me * sy + q : me * sz + q;
                if (mk[ime]) continue;
                if (psimilar(im, inb, ime, J->T)) {
                    mk[ime] = 1; count++; local_progress = 1;
                    pvox_t p;
                    if (along_y) { p.x = (int32_t)q; p.y = (int32_t)me; p.z = 0; }
                    else { p.x = (int32_t)(q % nx); p.y = (int32_t)(q / nx); p.z = (int32_t)me; }
                    ps_push(&st, p);
                }
            }
        }
        if (local_progress) J->progress[round & 1] = 1;
        pthread_barrier_wait(&J->bar);                      /* finished(): Allreduce */
        int any = J->progress[round & 1];
        J->progress[(round + 1) & 1] = 0;
        pthread_barrier_wait(&J->bar);                      /* region.c:531 */
        round++;
        if (!any) break;
    }
    J->counts[t] = count;
    if (t == 0) J->rounds = round;
    free(st.v);
    return NULL;
}

/* Returns region size (or -1); *rounds_out = number of outer rounds (region.c:497 loop trips). */
int64_t oracle_grow_par(const uint8_t *image, int64_t nx, int64_t ny, int64_t nz,
                        const int32_t *seeds_xyz, int nseeds, int threshold,
                        int nthreads, uint8_t *mask, int *rounds_out)
{
    if (nx <= 0 || ny <= 0 || nz <= 0 || nthreads < 1 || nthreads > 1024) return -1;
    memset(mask, 0, (size_t)(nx * ny * nz));
    pjob_t J; memset(&J, 0, sizeof J);
    J.image = image; J.mask = mask; J.nx = nx; J.ny = ny; J.nz = nz;
    J.seeds = seeds_xyz; J.nseeds = nseeds; J.T = threshold; J.nthreads = nthreads;
    J.counts = (int64_t *)calloc((size_t)nthreads, sizeof(int64_t));
    pthread_barrier_init(&J.bar, NULL, (unsigned)nthreads);
    pthread_t *th = (pthread_t *)malloc(sizeof(pthread_t) * (size_t)nthreads);
    parg_t *args = (parg_t *)malloc(sizeof(parg_t) * (size_t)nthreads);
    for (int t = 0; t < nthreads; t++) { args[t].job = &J; args[t].tid = t; pthread_create(&th[t], NULL, worker, &args[t]); }
    int64_t total = 0;
    for (int t = 0; t < nthreads; t++) { pthread_join(th[t], NULL); total += J.counts[t]; }
    if (rounds_out) *rounds_out = J.rounds;
    pthread_barrier_destroy(&J.bar);
    free(th); free(args); free(J.counts);
    return total;
}
