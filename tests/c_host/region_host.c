/*
 * tests/c_host/region_host.c -- a compiled C host on the drop-in boundary.  TEST CODE.
 *
 * What is left of the reference's 2-mpi-region-growing/region.c main() (region.c:582-604) when it links against
 * libb200rg.so: read_bmp -> one call that replaces init_mpi ... gather_region (region.c:586-597) -> write_bmp.
 * Compiled by tests/test_c_host.py with `gcc -I include ... -lb200rg`, run on a 512 x 512 picture; the payload of the
 * ./out.bmp it writes must equal the payload the unmodified region.c wrote (tests/golden).
 *
 * usage: region_host <picture.bmp>      (writes ./out.bmp like the reference, region.c:578 / bmp.c:39)
 */
#include <stdio.h>
#include <stdlib.h>
#include "bmp.h"
#include "rg.h"

int main(int argc, char **argv)
{
    if (argc != 2) { printf("Useage: %s image\n", argv[0]); return 2; }          /* region.c:552-555 */
    unsigned char *image = read_bmp(argv[1]);                                   /* region.c:558 */
    if (!image) { fprintf(stderr, "cannot read %s\n", argv[1]); return 1; }
    unsigned char *region = (unsigned char *)calloc(1, 512 * 512);              /* region.c:559 */
    if (rg_grow_region_2d(image, 512, 512, region, 1) != RG_OK) {               /* grow_region + write_image's masking */
        fprintf(stderr, "rg_grow_region_2d: %s\n", rg_last_error());
        return 1;
    }
    long n = 0;
    for (int i = 0; i < 512 * 512; i++) n += region[i];
    printf("region pixels: %ld\n", n);
    write_bmp(image, 512, 512);                                                 /* region.c:578 */
    /* the header records of bmp.h are part of the surface too */
    struct bmp_id id = { 'B', 'M' }; struct bmp_header h = { 512 * 512 + 56, 0, 0, 1078 }; struct bmp_dib_header d; d.width = 512;
    if (sizeof id != 2 || sizeof h != 12 || sizeof d != 40 || id.magic1 != 'B' || h.pixel_offset != 1078 || d.width != 512) return 3;
    free(region); free(image);
    return 0;
}
