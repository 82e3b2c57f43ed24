/*
 * include/bmp.h -- image I/O boundary of the drop-in library (libb200rg.so).
 *
 * Same two symbols, same signatures and same on-disk bytes as the reference's
 * bmp.h:32-33 / bmp.c:6-71 (identical copies in 2-mpi-region-growing/, 4-.../,
 * 5-cuda-region-growing/bmp.cpp and 6-opencl-region-growing/), so region.c,
 * raycast.cu and raycast.c link against this library unchanged.
 */
#ifndef B200RG_BMP_H
#define B200RG_BMP_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* replaces bmp.c:6-47.  Always writes ./out.bmp: "BM", 12-byte header (file_size = w*h+56,
 * pixel_offset = 1078), 40-byte DIB (8 bpp, 256 colours), 1024-byte grey palette, w*h payload,
 * 2 padding bytes.  The 6 bytes the reference leaves uninitialised (creator1/2 at offsets 6-9
 * and the 2 trailing padding bytes, bmp.c:11-13,28) are written as zeros. */
void write_bmp(unsigned char *data, int width, int height);

/* replaces bmp.c:49-71.  Returns malloc'd width*height bytes read from the file's pixel offset
 * (palette, bit depth, row padding and row order are ignored exactly like the reference).
 * Unlike the reference it returns NULL instead of crashing when the file cannot be opened. */
unsigned char *read_bmp(char *filename);

/* new: same as read_bmp but also reports the dimensions the reference makes callers hard-code
 * (region.c:27).  No counterpart in the reference. */
unsigned char *rg_read_bmp_dims(const char *filename, int *width, int *height);

/* new: write_bmp to an arbitrary path.  No counterpart in the reference. */
int rg_write_bmp_path(const char *path, const unsigned char *data, int width, int height);

#ifdef __cplusplus
}
#endif
#endif
