/*
 * rg_synth.cu -- on-device synthetic inputs for the benchmark configurations.
 * Replaces func()/create_data() of 5-cuda-region-growing/raycast.cu:114-158 (4-5 s per 512^3 on a
 * CPU core; minutes at 2048^3) with one kernel: the same solids scaled by n/512, squared integer
 * distances instead of sqrt(), and a counter-based hash (splitmix64 of the voxel index) instead of
 * the sequential rand().  C2/C3/C4 add the shapes described in SURVEY.md section 8d.
 * tests/test_synth.py compares every kind byte for byte with the CPU definitions in oracle/synth_cpu.c.
 */
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/rg.h"

namespace {

__host__ __device__ inline uint64_t splitmix64(uint64_t x)
{
    uint64_t z = x + 0x9e3779b97f4a7c15ULL;
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ULL;
    z = (z ^ (z >> 27)) * 0x94d049bb133111ebULL;
    return z ^ (z >> 31);
}
__device__ inline uint8_t noise20(uint64_t seed, uint64_t idx)
{
    return (uint8_t)(((splitmix64(seed + idx) >> 32) * 20ULL) >> 32);
}
__host__ __device__ inline int64_t sc(int64_t v, int64_t n) { return v * n / 512; }
__device__ inline bool sphere(int64_t x, int64_t y, int64_t z, int64_t cx, int64_t cy, int64_t cz, int64_t r)
{
    int64_t dx = x - cx, dy = y - cy, dz = z - cz;
    return dx * dx + dy * dy + dz * dz < r * r;
}
__device__ inline uint8_t exercise_solid(int64_t x, int64_t y, int64_t z, int64_t n)
{
    uint8_t v = 0;
    if (sphere(x, y, z, sc(300, n), sc(400, n), sc(100, n), sc(100, n))) v = 30;
    if (sphere(x, y, z, sc(100, n), sc(200, n), sc(400, n), sc(50, n))) v = 50;
    if (x > sc(200, n) && x < sc(300, n) && y > sc(300, n) && y < sc(500, n) && z > sc(200, n) && z < sc(300, n)) v = 45;
    if (x > 0 && x < sc(100, n) && y > sc(250, n) && y < sc(400, n) && z > sc(250, n) && z < sc(400, n)) v = 35;
    return v;
}
__device__ inline bool c4_serpentine(int64_t x, int64_t y, int64_t z, int64_t n)
{
    const int ncol = 64;
    int64_t y0 = n / 64, x0 = n / 64;
    if (y != y0) return false;
    int64_t d = x - x0;
    if (d < 0 || d > 2 * (ncol - 1)) return false;
    if ((d & 1) == 0) return true;
    int64_t j = (d - 1) / 2;
    return (j & 1) == 0 ? (z == n - 1) : (z == 0);
}
__device__ inline uint8_t c2_solid(int64_t x, int64_t y, int64_t n)
{
    int64_t c = n / 8; if (c < 1) c = 1;
    int64_t i = x / c, j = y / c;
    if (i > 7 || j > 7) return 0;
    int64_t dx = x - i * c - c / 2, dy = y - j * c - c / 2;
    if (j % 4 == 0) { int64_t a = dy < 0 ? -dy : dy; return (4 * a < c) ? 40 : 0; }
    return (25 * (dx * dx + dy * dy) < 4 * c * c) ? (uint8_t)(50 + 5 * ((i + j) % 3)) : 0;
}
__device__ inline bool c3_corridor(int64_t x, int64_t y, int64_t z, int64_t n)
{
    int64_t R = (n + 1) / 2, last = 2 * (R - 1);
    if ((z & 1) == 0) {
        if ((y & 1) == 0) return true;
        if (y + 1 > last) return false;
        return x == ((((y - 1) / 2) & 1) == 0 ? n - 1 : 0);
    }
    if (z + 1 > last) return false;
    int64_t ex = (((R - 1) & 1) == 0) ? n - 1 : 0, ey = last;
    if ((((z - 1) / 2) & 1) == 0) return x == ex && y == ey;
    return x == 0 && y == 0;
}

__device__ inline uint8_t voxel(int kind, int64_t x, int64_t y, int64_t z, int64_t n, uint64_t seed)
{
    uint64_t idx = (uint64_t)((z * n + y) * n + x);
    if (kind == 3) return c3_corridor(x, y, z, n) ? 200 : 0;
    uint8_t s;
    if (kind == 2) s = c2_solid(x, y, n);
    else { s = exercise_solid(x, y, z, n); if (kind == 4 && c4_serpentine(x, y, z, n)) s = 200; }
    return s ? s : noise20(seed, idx);
}

/* one thread = 16 consecutive voxels of a row -> one 16-byte store */
__global__ void __launch_bounds__(256) k_synth(int kind, uint8_t *out, int64_t n, int64_t z_begin, int64_t nzl, uint64_t seed)
{
    const int64_t cpr = (n + 15) / 16;                           /* chunks per row */
    const int64_t rows = (kind == 2 ? 1 : nzl) * n;
    const int64_t total = rows * cpr;
    const bool vec = (n % 16 == 0) && (((uintptr_t)out & 15) == 0);
    for (int64_t it = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; it < total; it += (int64_t)gridDim.x * blockDim.x) {
        int64_t c = it % cpr, row = it / cpr;
        int64_t y = row % n, zl = row / n, z = z_begin + zl, x0 = c * 16;
        uint32_t w[4] = {0, 0, 0, 0};
        for (int b = 0; b < 16; b++) {
            int64_t x = x0 + b;
            if (x < n) w[b >> 2] |= (uint32_t)voxel(kind, x, y, z, n, seed) << (8 * (b & 3));
        }
        uint8_t *g = out + (zl * n + y) * n + x0;
        if (vec) *(uint4 *)g = make_uint4(w[0], w[1], w[2], w[3]);
        else for (int b = 0; b < 16 && x0 + b < n; b++) g[b] = (uint8_t)(w[b >> 2] >> (8 * (b & 3)));
    }
}

} // namespace

extern "C" int rg_synth_slab_device(int kind, uint8_t *d_out, int64_t n, int64_t z_begin, int64_t z_end, uint64_t seed, void *stream)
{
    if (!d_out || n <= 0 || kind < 1 || kind > 4) return RG_ERR_ARG;
    if (kind == 2) { z_begin = 0; z_end = 1; }
    if (z_begin < 0 || z_end > n || z_end <= z_begin) return RG_ERR_ARG;
    int64_t nzl = z_end - z_begin;
    int64_t total = nzl * n * ((n + 15) / 16);
    int64_t blocks = (total + 255) / 256; if (blocks > 148 * 64) blocks = 148 * 64;
    k_synth<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(kind, d_out, n, z_begin, nzl, seed);
    return cudaGetLastError() == cudaSuccess ? RG_OK : RG_ERR_CUDA;
}

extern "C" int rg_synth_device(int kind, uint8_t *d_out, int64_t n, uint64_t seed, void *stream)
{
    return rg_synth_slab_device(kind, d_out, n, 0, kind == 2 ? 1 : n, seed, stream);
}

extern "C" int rg_synth_seeds(int kind, int64_t n, uint64_t seed, int32_t *out, int max_seeds)
{
    if (!out || n <= 0) return RG_ERR_ARG;
    int cnt = 0;
#define PUT(X, Y, Z) do { if (cnt < max_seeds) { out[3 * cnt] = (int32_t)(X); out[3 * cnt + 1] = (int32_t)(Y); out[3 * cnt + 2] = (int32_t)(Z); } cnt++; } while (0)
    if (kind == 1) PUT(sc(50, n), sc(300, n), sc(300, n));       /* raycast.cu:286 scaled */
    else if (kind == 3) PUT(0, 0, 0);
    else if (kind == 2) {
        int64_t c = n / 8;
        for (int j = 0; j < 8; j++) for (int i = 0; i < 8; i++) PUT(i * c + c / 2, j * c + c / 2, 0);
    } else if (kind == 4) {
        PUT(sc(300, n), sc(400, n), sc(100, n)); PUT(sc(300, n), sc(350, n), sc(60, n));
        PUT(sc(100, n), sc(200, n), sc(400, n)); PUT(sc(110, n), sc(190, n), sc(410, n));
        PUT(sc(250, n), sc(400, n), sc(250, n)); PUT(sc(210, n), sc(480, n), sc(290, n));
        PUT(sc(50, n), sc(300, n), sc(300, n));
        PUT(n / 64, n / 64, 0);
        for (int k = 8; k < 16; k++) {
            uint64_t h = splitmix64(seed ^ (0xABCD0000ULL + (uint64_t)k));
            PUT(h % (uint64_t)n, (n / 32) + (h >> 20) % (uint64_t)(n / 8), (h >> 40) % (uint64_t)n);
        }
    } else return RG_ERR_ARG;
#undef PUT
    return cnt;
}
