/*
 * bmp.c -- image I/O of the drop-in library: same symbols and file format as the reference's
 * bmp.c:6-71 (see include/bmp.h).  Written from the format, not from the reference source:
 * one packed little-endian header block is assembled in memory and written with the payload.
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "../../include/bmp.h"

static void put32(unsigned char *p, uint32_t v) { p[0] = v & 255; p[1] = (v >> 8) & 255; p[2] = (v >> 16) & 255; p[3] = (v >> 24) & 255; }
static void put16(unsigned char *p, uint32_t v) { p[0] = v & 255; p[1] = (v >> 8) & 255; }
static uint32_t get32(const unsigned char *p) { return (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24); }

enum { HDR = 14, DIB = 40, PAL = 1024, PIXEL_OFFSET = HDR + DIB + PAL /* 1078 */ };

int rg_write_bmp_path(const char *path, const unsigned char *data, int width, int height)
{
    if (!path || !data || width <= 0 || height <= 0) return -1;
    unsigned char head[PIXEL_OFFSET];
    memset(head, 0, sizeof head);
    size_t npix = (size_t)width * (size_t)height;
    head[0] = 'B'; head[1] = 'M';
    put32(head + 2, (uint32_t)(npix + 54 + 2));      /* the reference's file_size, bmp.c:12 */
    /* bytes 6..9 (creator1/2): zeros where the reference writes uninitialised stack */
    put32(head + 10, PIXEL_OFFSET);
    put32(head + 14, DIB);
    put32(head + 18, (uint32_t)width);
    put32(head + 22, (uint32_t)height);
    put16(head + 26, 1);                             /* planes */
    put16(head + 28, 8);                             /* bits per pixel */
    put32(head + 30, 0);                             /* no compression */
    put32(head + 34, (uint32_t)npix);
    put32(head + 38, 0); put32(head + 42, 0);        /* resolution */
    put32(head + 46, 256); put32(head + 50, 0);      /* colours */
    for (int c = 0; c < 256; c++) { unsigned char *e = head + HDR + DIB + 4 * c; e[0] = e[1] = e[2] = (unsigned char)c; e[3] = 0; }
    FILE *fp = fopen(path, "wb");
    if (!fp) return -1;
    const unsigned char tail[2] = {0, 0};
    int ok = fwrite(head, 1, sizeof head, fp) == sizeof head && fwrite(data, 1, npix, fp) == npix && fwrite(tail, 1, 2, fp) == 2;
    if (fclose(fp) != 0) ok = 0;
    return ok ? 0 : -1;
}

void write_bmp(unsigned char *data, int width, int height)
{
    (void)rg_write_bmp_path("out.bmp", data, width, height);      /* fixed name, bmp.c:39 */
}

unsigned char *rg_read_bmp_dims(const char *filename, int *width, int *height)
{
    if (!filename) return NULL;
    FILE *fp = fopen(filename, "rb");
    if (!fp) return NULL;
    unsigned char h[26];
    if (fread(h, 1, sizeof h, fp) != sizeof h) { fclose(fp); return NULL; }
    int32_t w = (int32_t)get32(h + 18), ht = (int32_t)get32(h + 22);
    uint32_t off = get32(h + 10);
    if (w <= 0 || ht <= 0) { fclose(fp); return NULL; }
    size_t n = (size_t)w * (size_t)ht;
    unsigned char *data = (unsigned char *)malloc(n);
    if (!data) { fclose(fp); return NULL; }
    /* rows are taken as stored: no padding, palette or orientation handling (bmp.c:65-66) */
    if (fseek(fp, (long)off, SEEK_SET) != 0) { free(data); fclose(fp); return NULL; }
    size_t got = fread(data, 1, n, fp);
    if (got < n) memset(data + got, 0, n - got);
    fclose(fp);
    if (width) *width = w;
    if (height) *height = ht;
    return data;
}

unsigned char *read_bmp(char *filename) { return rg_read_bmp_dims(filename, NULL, NULL); }
