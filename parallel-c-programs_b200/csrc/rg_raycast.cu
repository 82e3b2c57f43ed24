/*
 * rg_raycast.cu -- the volume ray-caster that consumes the region mask.
 *
 * Replaces raycast_kernel / raycast_gpu (5-cuda-region-growing/raycast.cu:321-371, 436-470), the legacy-texture
 * variant raycast_kernel_texture / raycast_gpu_texture (:374-433, 472-531; texture references were removed in
 * CUDA 12, so the reference file no longer compiles) and the OpenCL twin (6-opencl-region-growing/raycast.cl:93-137).
 * Same image as raycast_serial (raycast.cu:216-267) up to float rounding: one ray per pixel from the camera at
 * (cam,cam,cam) towards (-1,-1,-1), at most 5000 steps of 0.5, colour += trilinear(data) * (0.01 + (int)trilinear(region)),
 * stop at colour >= 255.  The reference's quirks are kept on purpose (they decide the picture): inside() accepts
 * 0 <= c < D-1 only, and the interpolation weights are swapped (rx * value[floor] + (1-rx) * value[ceil]).
 *
 * Not a translation of the reference kernel: one thread per pixel in 16x4 pixel blocks (measured best of five shapes) (coherent rays share cache
 * lines; the reference launches 8-thread blocks), the position is advanced exactly as the reference does (iterated
 * float adds, so the sample positions are bit-identical) but samples outside the volume cost only the bounds test
 * and a ray that has left the volume for good stops marching,
 * all indexing is 64-bit, sizes and camera are arguments.
 *
 * Floating point.  Every device operation is written with the round-to-nearest intrinsics (__fmul_rn, __fadd_rn,
 * __fdiv_rn, __fsqrt_rn, and the double ones for the "(0.01 + r)" accumulation, which the reference evaluates in
 * double), so nvcc cannot contract a*b+c into an FMA and each operation rounds as the reference's x86-64 SSE code
 * does.  tests/test_gpu_raycast.py states the tolerance it holds the image to.
 */
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include "../../include/rg.h"

#ifndef RG_RAY_BW
#define RG_RAY_BW 16     /* pixel block = RG_RAY_BW x RG_RAY_BH threads; a warp covers RG_RAY_BW x (32 / RG_RAY_BW) pixels */
#define RG_RAY_BH 4
#endif

namespace {

struct F3 { float x, y, z; };
/* host side (camera basis, computed once per call; g++ on x86-64 does not contract) */
inline F3 h_cross(F3 a, F3 b) { return F3{ a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x }; }
inline F3 h_norm(F3 v) { float l = sqrt(v.x * v.x + v.y * v.y + v.z * v.z); return F3{ v.x / l, v.y / l, v.z / l }; }
/* device side: explicit round-to-nearest operations, never fused */
__device__ __forceinline__ float mul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float add(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ F3 f3add(F3 a, F3 b) { return F3{ add(a.x, b.x), add(a.y, b.y), add(a.z, b.z) }; }
__device__ __forceinline__ F3 f3scale(F3 a, float s) { return F3{ mul(a.x, s), mul(a.y, s), mul(a.z, s) }; }
__device__ __forceinline__ F3 f3norm(F3 v)
{
    const float l = __fsqrt_rn(add(add(mul(v.x, v.x), mul(v.y, v.y)), mul(v.z, v.z)));
    return F3{ __fdiv_rn(v.x, l), __fdiv_rn(v.y, l), __fdiv_rn(v.z, l) };
}
/* one interpolation step with the reference's (swapped) weights: w * lo + (1 - w) * hi */
__device__ __forceinline__ float lerp_ref(float w, float lo, float hi) { return add(mul(w, lo), mul(__fsub_rn(1.0f, w), hi)); }

/* trilinear sample with the reference's weights (raycast.cu:183-212); the caller has done the inside() test */
__device__ __forceinline__ float value_at(const uint8_t *__restrict__ v, size_t D, F3 p)
{
    const float fx = floorf(p.x), fy = floorf(p.y), fz = floorf(p.z);
    const size_t x = (size_t)fx, y = (size_t)fy, z = (size_t)fz;
    const size_t xu = (size_t)ceilf(p.x), yu = (size_t)ceilf(p.y), zu = (size_t)ceilf(p.z);
    const float rx = __fsub_rn(p.x, fx), ry = __fsub_rn(p.y, fy), rz = __fsub_rn(p.z, fz);
    const uint8_t *r00 = v + (z * D + y) * D, *r01 = v + (z * D + yu) * D, *r10 = v + (zu * D + y) * D, *r11 = v + (zu * D + yu) * D;
    const float a0 = lerp_ref(rx, (float)__ldg(r00 + x), (float)__ldg(r00 + xu));
    const float a1 = lerp_ref(rx, (float)__ldg(r01 + x), (float)__ldg(r01 + xu));
    const float a2 = lerp_ref(rx, (float)__ldg(r10 + x), (float)__ldg(r10 + xu));
    const float a3 = lerp_ref(rx, (float)__ldg(r11 + x), (float)__ldg(r11 + xu));
    return lerp_ref(rz, lerp_ref(ry, a0, a1), lerp_ref(ry, a2, a3));
}

__global__ void __launch_bounds__(RG_RAY_BW * RG_RAY_BH) k_raycast(const uint8_t *__restrict__ data, const uint8_t *__restrict__ region, int D, int I,
                                                 float cam, float pixel_width, F3 forward, F3 right, F3 up, uint8_t *__restrict__ image)
{
    const int px = blockIdx.x * RG_RAY_BW + (threadIdx.x % RG_RAY_BW), py = blockIdx.y * RG_RAY_BH + (threadIdx.x / RG_RAY_BW);
    if (px >= I || py >= I) return;
    const int x = px - I / 2, y = py - I / 2;
    const F3 camera = { cam, cam, cam };
    F3 ray = f3add(f3add(f3add(camera, forward), f3scale(right, mul((float)x, pixel_width))), f3scale(up, mul((float)y, pixel_width)));
    ray = f3norm(f3add(ray, f3scale(camera, -1.0f)));
    const F3 step = f3scale(ray, 0.5f);
    const float lim = (float)(D - 1);
    F3 pos = camera;
    float color = 0.0f;
    for (int i = 0; i < 5000 && color < 255.0f; i++) {
        pos = f3add(pos, step);
        /* outside the volume both samples are 0 and the colour does not move (raycast.cu:184-186) */
        if (pos.x >= 0 && pos.x < lim && pos.y >= 0 && pos.y < lim && pos.z >= 0 && pos.z < lim) {
            const int r = (int)value_at(region, (size_t)D, pos);
            /* "color += value * (0.01 + r)": double product and sum, rounded to float once (raycast.cu:260) */
            const double w = __dadd_rn(0.01, (double)r);
            color = (float)__dadd_rn((double)color, __dmul_rn((double)value_at(data, (size_t)D, pos), w));
        } else if ((pos.x < 0 && step.x <= 0) || (pos.x >= lim && step.x >= 0) || (pos.y < 0 && step.y <= 0) ||
                   (pos.y >= lim && step.y >= 0) || (pos.z < 0 && step.z <= 0) || (pos.z >= lim && step.z >= 0)) {
            break;      /* past the volume and moving away (rounded adds are monotone): no later sample can be inside */
        }
    }
    image[(size_t)py * I + px] = color > 255.0f ? 255 : (uint8_t)color;
}

} // namespace

extern "C" int rg_raycast_device(const uint8_t *d_data, const uint8_t *d_region, int data_dim, int image_dim, float camera,
                                 uint8_t *d_image, void *stream)
{
    if (!d_data || !d_region || !d_image || data_dim < 2 || image_dim < 2) return RG_ERR_ARG;
    /* the constant part of the camera set-up, in the reference's order of operations (raycast.cu:324-339) */
    F3 forward = { -1, -1, -1 }, z_axis = { 0, 0, 1 };
    F3 right = h_cross(forward, z_axis);
    F3 up = h_cross(right, forward);
    forward = h_norm(forward); right = h_norm(right); up = h_norm(up);
    const float fov = 3.14 / 4;
    const float pixel_width = tan(fov / 2.0) / (image_dim / 2);
    dim3 grid((image_dim + RG_RAY_BW - 1) / RG_RAY_BW, (image_dim + RG_RAY_BH - 1) / RG_RAY_BH);
    k_raycast<<<grid, RG_RAY_BW * RG_RAY_BH, 0, (cudaStream_t)stream>>>(d_data, d_region, data_dim, image_dim, camera, pixel_width, forward, right, up, d_image);
    return cudaGetLastError() == cudaSuccess ? RG_OK : RG_ERR_CUDA;
}

/* reference-shaped: host 512^3 data + region in, malloc'd 512 x 512 image out (the caller free()s it, raycast.cu:853) */
static unsigned char *raycast_512(unsigned char *data, unsigned char *region)
{
    const int D = 512, I = 512;
    const size_t nvox = (size_t)D * D * D;
    if (!data || !region) return nullptr;
    uint8_t *d_data = nullptr, *d_region = nullptr, *d_image = nullptr;
    unsigned char *image = (unsigned char *)malloc((size_t)I * I);
    bool ok = image != nullptr;
    ok = ok && cudaMalloc((void **)&d_data, nvox) == cudaSuccess;
    ok = ok && cudaMalloc((void **)&d_region, nvox) == cudaSuccess;
    ok = ok && cudaMalloc((void **)&d_image, (size_t)I * I) == cudaSuccess;
    ok = ok && cudaMemcpy(d_data, data, nvox, cudaMemcpyHostToDevice) == cudaSuccess;
    ok = ok && cudaMemcpy(d_region, region, nvox, cudaMemcpyHostToDevice) == cudaSuccess;
    ok = ok && rg_raycast_device(d_data, d_region, D, I, 1000.0f, d_image, nullptr) == RG_OK;
    ok = ok && cudaMemcpy(image, d_image, (size_t)I * I, cudaMemcpyDeviceToHost) == cudaSuccess;
    cudaFree(d_data); cudaFree(d_region); cudaFree(d_image);
    if (!ok) { free(image); return nullptr; }
    return image;
}
extern "C" unsigned char *raycast_gpu(unsigned char *data, unsigned char *region) { return raycast_512(data, region); }
extern "C" unsigned char *raycast_gpu_texture(unsigned char *data, unsigned char *region) { return raycast_512(data, region); }
