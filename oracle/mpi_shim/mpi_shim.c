/*
 * oracle/mpi_shim/mpi_shim.c -- fork + shared-memory implementation of mpi.h.
 * TEST INFRASTRUCTURE ONLY.  See mpi.h.
 *
 * Model: MPI_Init reads SHIM_NP (default 1), maps one anonymous shared segment
 * holding an (np x np) matrix of FIFO byte rings plus barrier / all-reduce
 * words, then fork()s np-1 children.  Sends are buffered (ring capacity 1 MiB),
 * which region.c relies on: rank 0 sends 256 rows to itself before receiving
 * (region.c:108-141) and every rank sends its mask rows before rank 0 starts
 * receiving (region.c:393-429).  Tags are ignored (region.c uses one, :18).
 */
#define _GNU_SOURCE
#include "mpi.h"
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <stdint.h>
#include <unistd.h>
#include <sched.h>
#include <sys/mman.h>
#include <sys/wait.h>

#define RING_BYTES (1u << 20)
#define MAX_NP 64
#define MAX_TYPES 64

typedef struct {
    volatile uint64_t head;            /* bytes written  (producer only) */
    volatile uint64_t tail;            /* bytes consumed (consumer only) */
    unsigned char data[RING_BYTES];
} ring_t;

typedef struct {
    volatile int bar_count;
    volatile int bar_sense;
    volatile int slots[MAX_NP];
} ctrl_t;

typedef struct { int count, blocklen, stride, elem; } vtype_t;

static int g_np = 1, g_rank = 0;
static ring_t *g_rings;               /* [src * np + dst] */
static ctrl_t *g_ctrl;
static int g_local_sense = 0;
static int g_dims[2] = {1, 1};
static vtype_t g_types[MAX_TYPES];
static int g_ntypes = 3;              /* 0 unused, 1 uchar, 2 int */
static pid_t g_children[MAX_NP];

static int elem_size(MPI_Datatype t) { return t == MPI_INT ? 4 : 1; }

int MPI_Init(int *argc, char ***argv)
{
    (void)argc; (void)argv;
    const char *e = getenv("SHIM_NP");
    g_np = e ? atoi(e) : 1;
    if (g_np < 1 || g_np > MAX_NP) { fprintf(stderr, "mpi_shim: bad SHIM_NP\n"); exit(2); }
    size_t bytes = sizeof(ctrl_t) + sizeof(ring_t) * (size_t)g_np * g_np;
    void *m = mmap(NULL, bytes, PROT_READ | PROT_WRITE, MAP_SHARED | MAP_ANONYMOUS, -1, 0);
    if (m == MAP_FAILED) { perror("mpi_shim mmap"); exit(2); }
    g_ctrl = (ctrl_t *)m;
    g_rings = (ring_t *)((char *)m + sizeof(ctrl_t));
    memset(g_ctrl, 0, sizeof *g_ctrl);
    for (int i = 0; i < g_np * g_np; i++) { g_rings[i].head = 0; g_rings[i].tail = 0; }
    fflush(NULL);
    g_rank = 0;
    for (int r = 1; r < g_np; r++) {
        pid_t p = fork();
        if (p < 0) { perror("mpi_shim fork"); exit(2); }
        if (p == 0) { g_rank = r; break; }
        g_children[r] = p;
    }
    return MPI_SUCCESS;
}

int MPI_Comm_size(MPI_Comm c, int *size) { (void)c; *size = g_np; return 0; }
int MPI_Comm_rank(MPI_Comm c, int *rank) { (void)c; *rank = g_rank; return 0; }

/* most-square factorisation, larger factor first; only ndims == 2 is needed */
int MPI_Dims_create(int nnodes, int ndims, int *dims)
{
    if (ndims != 2) { fprintf(stderr, "mpi_shim: ndims != 2\n"); exit(2); }
    int a = 1;
    for (int f = 1; f * f <= nnodes; f++) if (nnodes % f == 0) a = f;
    dims[0] = nnodes / a; dims[1] = a;
    return 0;
}

int MPI_Cart_create(MPI_Comm old, int ndims, int *dims, int *periods, int reorder, MPI_Comm *cart)
{
    (void)old; (void)ndims; (void)periods; (void)reorder;
    g_dims[0] = dims[0]; g_dims[1] = dims[1];
    *cart = 1;
    return 0;
}

int MPI_Cart_coords(MPI_Comm c, int rank, int maxdims, int *coords)
{
    (void)c; (void)maxdims;           /* region.c:413 passes maxdims = size; write 2 only */
    coords[0] = rank / g_dims[1];
    coords[1] = rank % g_dims[1];
    return 0;
}

int MPI_Cart_rank(MPI_Comm c, int *coords, int *rank)
{
    (void)c; *rank = coords[0] * g_dims[1] + coords[1]; return 0;
}

int MPI_Cart_shift(MPI_Comm c, int direction, int disp, int *src, int *dst)
{
    (void)c;
    int co[2] = { g_rank / g_dims[1], g_rank % g_dims[1] };
    int lo[2] = { co[0], co[1] }, hi[2] = { co[0], co[1] };
    lo[direction] -= disp; hi[direction] += disp;
    *src = (lo[direction] < 0 || lo[direction] >= g_dims[direction]) ? MPI_PROC_NULL
                                                                    : lo[0] * g_dims[1] + lo[1];
    *dst = (hi[direction] < 0 || hi[direction] >= g_dims[direction]) ? MPI_PROC_NULL
                                                                    : hi[0] * g_dims[1] + hi[1];
    return 0;
}

int MPI_Type_vector(int count, int blocklen, int stride, MPI_Datatype old, MPI_Datatype *newtype)
{
    if (g_ntypes >= MAX_TYPES) { fprintf(stderr, "mpi_shim: too many types\n"); exit(2); }
    g_types[g_ntypes].count = count;
    g_types[g_ntypes].blocklen = blocklen;
    g_types[g_ntypes].stride = stride;
    g_types[g_ntypes].elem = elem_size(old);
    *newtype = g_ntypes++;
    return 0;
}
int MPI_Type_commit(MPI_Datatype *t) { (void)t; return 0; }

static void ring_write(ring_t *r, const unsigned char *p, size_t n)
{
    while (n) {
        uint64_t h = r->head, t = r->tail;
        size_t space = RING_BYTES - (size_t)(h - t);
        if (!space) { sched_yield(); continue; }
        size_t k = n < space ? n : space;
        size_t off = (size_t)(h % RING_BYTES);
        size_t first = k < RING_BYTES - off ? k : RING_BYTES - off;
        memcpy(r->data + off, p, first);
        memcpy(r->data, p + first, k - first);
        __sync_synchronize();
        r->head = h + k;
        p += k; n -= k;
    }
}

static void ring_read(ring_t *r, unsigned char *p, size_t n)
{
    while (n) {
        uint64_t h = r->head, t = r->tail;
        size_t avail = (size_t)(h - t);
        if (!avail) { sched_yield(); continue; }
        __sync_synchronize();
        size_t k = n < avail ? n : avail;
        size_t off = (size_t)(t % RING_BYTES);
        size_t first = k < RING_BYTES - off ? k : RING_BYTES - off;
        memcpy(p, r->data + off, first);
        memcpy(p + first, r->data, k - first);
        __sync_synchronize();
        r->tail = t + k;
        p += k; n -= k;
    }
}

/* walk `count` items of `type` starting at buf, calling the ring op per contiguous block */
static void xfer(int send, ring_t *r, unsigned char *buf, int count, MPI_Datatype type)
{
    if (type == MPI_UNSIGNED_CHAR || type == MPI_INT) {
        size_t n = (size_t)count * elem_size(type);
        if (send) ring_write(r, buf, n); else ring_read(r, buf, n);
        return;
    }
    vtype_t *v = &g_types[type];
    size_t extent = ((size_t)(v->count - 1) * v->stride + v->blocklen) * v->elem;
    for (int c = 0; c < count; c++) {
        unsigned char *base = buf + (size_t)c * extent;
        for (int b = 0; b < v->count; b++) {
            unsigned char *p = base + (size_t)b * v->stride * v->elem;
            size_t n = (size_t)v->blocklen * v->elem;
            if (send) ring_write(r, p, n); else ring_read(r, p, n);
        }
    }
}

int MPI_Send(void *buf, int count, MPI_Datatype type, int dest, int tag, MPI_Comm c)
{
    (void)tag; (void)c;
    if (dest < 0) return 0;
    xfer(1, &g_rings[g_rank * g_np + dest], (unsigned char *)buf, count, type);
    return 0;
}

int MPI_Recv(void *buf, int count, MPI_Datatype type, int src, int tag, MPI_Comm c, MPI_Status *st)
{
    (void)tag; (void)c; (void)st;
    if (src < 0) return 0;
    xfer(0, &g_rings[src * g_np + g_rank], (unsigned char *)buf, count, type);
    return 0;
}

int MPI_Barrier(MPI_Comm c)
{
    (void)c;
    g_local_sense = !g_local_sense;
    if (__sync_add_and_fetch(&g_ctrl->bar_count, 1) == g_np) {
        g_ctrl->bar_count = 0;
        __sync_synchronize();
        g_ctrl->bar_sense = g_local_sense;
    } else {
        while (g_ctrl->bar_sense != g_local_sense) sched_yield();
    }
    __sync_synchronize();
    return 0;
}

int MPI_Allreduce(void *sendbuf, void *recvbuf, int count, MPI_Datatype type, MPI_Op op, MPI_Comm c)
{
    if (count != 1 || type != MPI_INT || op != MPI_MIN) {
        fprintf(stderr, "mpi_shim: only Allreduce(1, INT, MIN) is implemented\n"); exit(2);
    }
    g_ctrl->slots[g_rank] = *(int *)sendbuf;
    MPI_Barrier(c);
    int m = g_ctrl->slots[0];
    for (int r = 1; r < g_np; r++) if (g_ctrl->slots[r] < m) m = g_ctrl->slots[r];
    MPI_Barrier(c);
    *(int *)recvbuf = m;
    return 0;
}

int MPI_Finalize(void)
{
    MPI_Barrier(0);
    fflush(NULL);
    if (g_rank == 0) {
        for (int r = 1; r < g_np; r++) { int s; waitpid(g_children[r], &s, 0); }
    }
    return 0;
}
