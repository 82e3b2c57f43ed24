/*
 * oracle/raycast_cpu.c -- CPU restatement of the reference's volume ray-caster.  TEST INFRASTRUCTURE ONLY.
 *
 * Follows raycast_serial() and value_at() of 5-cuda-region-growing/raycast.cu:183-267 (plain-C twin:
 * 6-opencl-region-growing/raycast.c:171-255) expression by expression, including their quirks:
 *   - inside(float3) accepts 0 <= c < D-1 only (raycast.cu:161-167);
 *   - the interpolation weights are swapped: rx*data[floor] + (1-rx)*data[ceil] (raycast.cu:200-203);
 *   - the region sample is truncated to int before use (raycast.cu:259), the constant 0.01 is a double;
 *   - fov = 3.14/4, pixel width tan(fov/2.0)/(I/2) in double, step 0.5, at most 5000 steps, stop at colour >= 255.
 * Parameters instead of #defines: D = volume side (DATA_DIM), I = image side (IMAGE_DIM), cam = camera
 * coordinate on all three axes (1000 in the reference).
 * Pinned in tests/test_oracle_pins.py against the unmodified raycast_serial built into oracle/_ref.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>

typedef struct { float x, y, z; } f3;

static f3 f3cross(f3 a, f3 b) { f3 c; c.x = a.y * b.z - a.z * b.y; c.y = a.z * b.x - a.x * b.z; c.z = a.x * b.y - a.y * b.x; return c; }
static f3 f3norm(f3 v) { float l = sqrt(v.x * v.x + v.y * v.y + v.z * v.z); v.x /= l; v.y /= l; v.z /= l; return v; }
static f3 f3add(f3 a, f3 b) { a.x += b.x; a.y += b.y; a.z += b.z; return a; }
static f3 f3scale(f3 a, float b) { a.x *= b; a.y *= b; a.z *= b; return a; }

static int inside_f(f3 p, int D)
{
    int x = (p.x >= 0 && p.x < D - 1), y = (p.y >= 0 && p.y < D - 1), z = (p.z >= 0 && p.z < D - 1);
    return x && y && z;
}

static float value_at(f3 pos, const uint8_t *data, int D)
{
    if (!inside_f(pos, D)) return 0;
    int x = floor(pos.x), y = floor(pos.y), z = floor(pos.z);
    int xu = ceil(pos.x), yu = ceil(pos.y), zu = ceil(pos.z);
    float rx = pos.x - x, ry = pos.y - y, rz = pos.z - z;
#define IDX(Z, Y, X) ((size_t)(Z) * D * D + (size_t)(Y) * D + (X))
    float a0 = rx * data[IDX(z, y, x)] + (1 - rx) * data[IDX(z, y, xu)];
    float a1 = rx * data[IDX(z, yu, x)] + (1 - rx) * data[IDX(z, yu, xu)];
    float a2 = rx * data[IDX(zu, y, x)] + (1 - rx) * data[IDX(zu, y, xu)];
    float a3 = rx * data[IDX(zu, yu, x)] + (1 - rx) * data[IDX(zu, yu, xu)];
#undef IDX
    float b0 = ry * a0 + (1 - ry) * a1;
    float b1 = ry * a2 + (1 - ry) * a3;
    float c0 = rz * b0 + (1 - rz) * b1;
    return c0;
}

void oracle_raycast(const uint8_t *data, const uint8_t *region, int D, int I, float cam, uint8_t *image)
{
    f3 camera = { cam, cam, cam };
    f3 forward = { -1, -1, -1 };
    f3 z_axis = { 0, 0, 1 };
    f3 right = f3cross(forward, z_axis);
    f3 up = f3cross(right, forward);
    forward = f3norm(forward); right = f3norm(right); up = f3norm(up);
    float fov = 3.14 / 4;
    float pixel_width = tan(fov / 2.0) / (I / 2);
    float step_size = 0.5;
    for (int y = -(I / 2); y < (I / 2); y++)
        for (int x = -(I / 2); x < (I / 2); x++) {
            f3 screen_center = f3add(camera, forward);
            f3 ray = f3add(f3add(screen_center, f3scale(right, x * pixel_width)), f3scale(up, y * pixel_width));
            ray = f3add(ray, f3scale(camera, -1));
            ray = f3norm(ray);
            f3 pos = camera;
            int i = 0;
            float color = 0;
            while (color < 255 && i < 5000) {
                i++;
                pos = f3add(pos, f3scale(ray, step_size));
                int r = value_at(pos, region, D);
                color += value_at(pos, data, D) * (0.01 + r);
            }
            image[(y + (I / 2)) * I + (x + (I / 2))] = color > 255 ? 255 : color;
        }
}
