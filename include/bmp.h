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

/* The three on-disk header records of an 8-bit BMP, under the names and field names the reference's header
 * exports (bmp.h:6-30), so that a caller that fills or inspects them keeps compiling.  The library's own reader
 * and writer (csrc/bmp.c) do not use them -- they address the file by byte offset -- and neither do region.c,
 * raycast.cu or raycast.c.  Sizes: 2, 12 and 40 bytes; a file is bmp_id, bmp_header, bmp_dib_header, palette, pixels. */
struct bmp_id {                    /* the two magic bytes "BM" */
    uint8_t magic1, magic2;
};
struct bmp_header {                /* file header after the magic: offsets 2..13 */
    uint32_t file_size;
    uint16_t creator1, creator2;   /* application-defined; the reference leaves them uninitialised, this library writes 0 */
    uint32_t pixel_offset;         /* where the pixel payload starts (1078 for an 8-bit image with a 256-entry palette) */
};
struct bmp_dib_header {            /* BITMAPINFOHEADER: offsets 14..53 */
    uint32_t header_size;          /* 40 */
    int32_t  width, height;
    uint16_t num_planes;           /* 1 */
    uint16_t bit_pr_pixel;         /* 8 */
    uint32_t compress_type;        /* 0 = uncompressed */
    uint32_t data_size;            /* width * height */
    int32_t  hres, vres;
    uint32_t num_colors;           /* 256 */
    uint32_t num_imp_colors;
};

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
