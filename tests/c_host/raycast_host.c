/*
 * tests/c_host/raycast_host.c -- the reference's 5-cuda-region-growing main() (raycast.cu:824-854) as a plain C host
 * of libb200rg.so.  TEST CODE.
 *
 * The volume comes from a file instead of create_data() (4-5 s of rand() per run); everything after it is the
 * reference's sequence: grow_region_gpu_shared(data) with the caller's pageable malloc'd buffer, raycast_gpu_texture,
 * write_bmp(image, 512, 512), three free()s.  Prints the region size, an FNV-1a hash of the mask and the wall-clock
 * time of the grow call (what the reference prints as "Grow time", raycast.cu:832-836); with a second argument N the
 * grow is repeated N times and the best time printed (bench.py's same-config end-to-end number).
 *
 * usage: raycast_host <volume512.raw> [repeats]      (writes ./out.bmp)
 */
#include <stdio.h>
#include <stdlib.h>
#include <stdint.h>
#include <sys/time.h>
#include "bmp.h"
#include "rg.h"

static double now_s(void) { struct timeval t; gettimeofday(&t, NULL); return t.tv_sec + 1e-6 * t.tv_usec; }

int main(int argc, char **argv)
{
    const size_t n = (size_t)512 * 512 * 512;
    if (argc < 2) { fprintf(stderr, "usage: %s volume512.raw [repeats]\n", argv[0]); return 2; }
    const int reps = argc > 2 ? atoi(argv[2]) : 1;
    unsigned char *data = (unsigned char *)malloc(n);                            /* raycast.cu:147 */
    FILE *f = fopen(argv[1], "rb");
    if (!data || !f || fread(data, 1, n, f) != n) { fprintf(stderr, "cannot read the volume\n"); return 1; }
    fclose(f);
    unsigned char *region = NULL;
    double best = 1e30, first = 0;
    for (int r = 0; r < reps; r++) {
        free(region);
        double t0 = now_s();
        region = grow_region_gpu_shared(data);                                   /* raycast.cu:833 */
        double dt = now_s() - t0;
        if (!region) { fprintf(stderr, "grow_region_gpu_shared: %s\n", rg_last_error()); return 1; }
        if (r == 0) first = dt;
        if (r > 0 && dt < best) best = dt;
    }
    if (reps == 1) best = first;
    uint64_t h = 0xcbf29ce484222325ull; long cnt = 0;
    for (size_t i = 0; i < n; i++) { h = (h ^ region[i]) * 0x100000001b3ull; cnt += region[i]; }
    printf("region voxels: %ld fnv1a64: %016llx\n", cnt, (unsigned long long)h);
    printf("grow seconds: first %.6f best %.6f\n", first, best);
    unsigned char *image = raycast_gpu_texture(data, region);                    /* raycast.cu:842 */
    if (!image) { fprintf(stderr, "raycast_gpu_texture: %s\n", rg_last_error()); return 1; }
    write_bmp(image, 512, 512);                                                  /* raycast.cu:849 */
    free(data); free(region); free(image);                                       /* raycast.cu:851-853 */
    return 0;
}
