/*
 * oracle/ref_cuda_main.cu -- times the REFERENCE's own CUDA growers on this GPU.  TEST / BENCH INFRASTRUCTURE ONLY.
 *
 * 5-cuda-region-growing/raycast.cu does not compile with CUDA >= 12 as a whole (legacy texture
 * references, raycast.cu:20-21, 374-433, 472-531).  oracle/Makefile extracts the rest of the file, untouched, into the
 * git-ignored oracle/_ref/raycast_cuda_extract.inc (sed at build time; no reference source is copied into the
 * repository) and this file includes it: create_data (raycast.cu:146-158), grow_region_serial (:281-318), the naive
 * grower (:534-574, 702-756) and the shared-memory grower that the reference's main() runs (:576-699, 759-822).
 *
 * The only interposition is a counting wrapper around cudaMemcpy: the reference's host loops copy ONE byte of their
 * "unfinished" flag to the device before every kernel launch (raycast.cu:739, 806), so the number of 1-byte
 * host-to-device copies is the number of launches.  Arithmetic and launch shapes are the reference's.
 *
 * Runs as its own process (bench.py starts it): the reference kernels read outside their allocations
 *  and a fault must not take the benchmark's CUDA context with it.
 * Output: one JSON line.
 */
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <stdint.h>

static int g_flag_copies = 0;
static cudaError_t counting_memcpy(void *dst, const void *src, size_t n, cudaMemcpyKind kind)
{
    if (n == 1 && kind == cudaMemcpyHostToDevice) g_flag_copies++;
    return cudaMemcpy(dst, src, n, kind);
}
#define cudaMemcpy counting_memcpy
#include "_ref/raycast_cuda_extract.inc"
#undef cudaMemcpy

static double now_s(void) { struct timeval t; gettimeofday(&t, NULL); return t.tv_sec + 1e-6 * t.tv_usec; }

struct MaskStats { long ones, twos, others, differs; };
static MaskStats compare(const unsigned char *got, const unsigned char *want)
{
    MaskStats m = {0, 0, 0, 0};
    for (long i = 0; i < (long)DATA_SIZE; i++) {
        unsigned char g = got[i];
        if (g == 1) m.ones++; else if (g == 2) m.twos++; else if (g) m.others++;
        if ((g != 0) != (want[i] != 0)) m.differs++;
    }
    return m;
}

static void run(const char *name, unsigned char *(*grow)(unsigned char *), unsigned char *data, const unsigned char *want, int reps, int last)
{
    double best = 1e30; int launches = 0; MaskStats st = {0, 0, 0, 0};
    for (int r = 0; r < reps + 1; r++) {                    /* first call = warm-up (context, first touch) */
        g_flag_copies = 0;
        double t0 = now_s();
        unsigned char *m = grow(data);
        cudaDeviceSynchronize();
        double dt = now_s() - t0;
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) { printf("\"%s\": {\"error\": \"%s\"}%s", name, cudaGetErrorString(e), last ? "" : ", "); free(m); return; }
        if (r > 0 && dt < best) best = dt;
        launches = g_flag_copies;
        if (r == reps) st = compare(m, want);
        free(m);
    }
    printf("\"%s\": {\"ms\": %.3f, \"launches\": %d, \"region_voxels\": %ld, \"mask_has_2s\": %ld, \"voxels_differing_from_serial\": %ld, "
           "\"timer\": \"host wall clock around the reference's host function: cudaMalloc x3, 2 x 128 MiB H2D, launch loop with two 1-byte copies per launch, 128 MiB D2H, cudaFree x3 -- what its main() times (raycast.cu:832-835)\"}%s",
           name, best * 1e3, launches, st.ones, st.twos, st.differs, last ? "" : ", ");
}

int main(int argc, char **argv)
{
    int reps = argc > 1 ? atoi(argv[1]) : 3;
    int with_naive = argc > 2 ? atoi(argv[2]) : 0;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { printf("{\"error\": \"no CUDA device\"}\n"); return 0; }
    unsigned char *data = create_data();                    /* raycast.cu:146-158: glibc rand(), unseeded */
    double t0 = now_s();
    unsigned char *want = grow_region_serial(data);         /* raycast.cu:281-318 on one host core */
    double t_serial = now_s() - t0;
    long region = 0; for (long i = 0; i < (long)DATA_SIZE; i++) region += want[i] != 0;
    printf("{\"volume\": \"reference create_data() 512^3, seed (50,300,300), T=1\", \"serial_ms\": %.3f, \"serial_region_voxels\": %ld, ", t_serial * 1e3, region);
    run("grow_region_gpu_shared", grow_region_gpu_shared, data, want, reps, !with_naive);
    if (with_naive) run("grow_region_gpu", grow_region_gpu, data, want, reps, 1);
    printf("}\n");
    free(want); free(data);
    return 0;
}
