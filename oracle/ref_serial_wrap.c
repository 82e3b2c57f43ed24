/*
 * oracle/ref_serial_wrap.c -- thin exports around the UNMODIFIED 3-D serial grower.
 * TEST INFRASTRUCTURE ONLY.
 *
 * oracle/Makefile generates oracle/_ref/serial_extract.inc at build time with
 *   sed -n '20,101p;103,168p;170,306p' /root/reference/6-opencl-region-growing/raycast.c
 * (types + stack + float3 helpers, func/create_data/inside/index, value_at/raycast_serial,
 * similar/grow_region_serial: the plain-C twin of 5-cuda-region-growing/raycast.cu:29-97,
 * 114-180,183-267,271-318).  The
 * extract is never committed (oracle/_ref/ is git-ignored); this wrapper only
 * re-exports its symbols under names that do not collide with libc (`index`).
 * Must be built with -std=c99 (strict), see SURVEY.md section 8c.
 */
#include <stdio.h>
#include <stdlib.h>
#include <math.h>
#include <string.h>

#define DATA_DIM 512
#define IMAGE_DIM 64      /* the value of 6-opencl-region-growing/raycast.c:16 */
#include "_ref/serial_extract.inc"

unsigned char *ref_create_data(void) { return create_data(); }
unsigned char *ref_grow_region_serial(unsigned char *data) { return grow_region_serial(data); }
unsigned char *ref_raycast_serial(unsigned char *data, unsigned char *region) { return raycast_serial(data, region); }
void ref_free(void *p) { free(p); }
