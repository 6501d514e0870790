// rt_scenes.cpp — deterministic procedural inputs for the bench and the parity
// tests.  They emit the arrays the reference's glTF loader would
// hand to BvhBuilder / BufferManager (src/scene.rs:43-69): vertices, indexed
// triangles, materials, lights, camera.  No file I/O, no CUDA.
#include <math.h>
#include <string.h>

#include "../../include/rtb200_host.h"

namespace {
// 32-bit integer hash -> [0,1)
inline float hash01(uint32_t seed, uint32_t ix, uint32_t iz) {
    uint32_t h = seed * 0x9E3779B1u ^ (ix * 0x85EBCA77u) ^ (iz * 0xC2B2AE3Du);
    h ^= h >> 16; h *= 0x7FEB352Du; h ^= h >> 15; h *= 0x846CA68Bu; h ^= h >> 16;
    return (float)(h >> 8) * (1.0f / 16777216.0f);
}
inline void set3(float* d, float x, float y, float z) { d[0] = x; d[1] = y; d[2] = z; }
} // namespace

extern "C" {

// Terrain(nx, nz, seed): regular XZ grid, spacing 100/nx, centred on the origin;
// y = 4 sin(0.15 x) cos(0.2 z) + 0.08 hash.  Triangles row-major, two per quad:
// (v00, v01, v11) and (v00, v11, v10) so geometric normals point +y.
void rt_scene_terrain(uint32_t nx, uint32_t nz, uint32_t seed, RtVertex* vertices, RtTriangle* triangles,
                      RtMaterial materials[3], RtLight lights[2], RtCamera* camera) {
    const float s = 100.0f / (float)nx;
    const float width = 100.0f, depth = s * (float)nz;
    for (uint32_t iz = 0; iz <= nz; iz++)
        for (uint32_t ix = 0; ix <= nx; ix++) {
            float x = (float)ix * s - 0.5f * width, z = (float)iz * s - 0.5f * depth;
            double y = 4.0 * sin(0.15 * (double)x) * cos(0.2 * (double)z) + 0.08 * (double)hash01(seed, ix, iz);
            set3(vertices[(size_t)iz * (nx + 1) + ix].position, x, (float)y, z);
        }
    size_t t = 0;
    for (uint32_t iz = 0; iz < nz; iz++)
        for (uint32_t ix = 0; ix < nx; ix++) {
            uint32_t v00 = iz * (nx + 1) + ix, v10 = v00 + 1, v01 = v00 + (nx + 1), v11 = v01 + 1;
            uint32_t mat = (iz / 8u) % 3u; // stripes: diffuse, metallic, glass
            triangles[t].v0_index = v00; triangles[t].v1_index = v01; triangles[t].v2_index = v11; triangles[t].material_id = mat; t++;
            triangles[t].v0_index = v00; triangles[t].v1_index = v11; triangles[t].v2_index = v10; triangles[t].material_id = mat; t++;
        }
    const float green[3] = {0.35f, 0.6f, 0.3f}, steel[3] = {0.75f, 0.75f, 0.8f}, aqua[3] = {0.3f, 0.6f, 0.9f};
    rt_material_diffuse(&materials[0], green);
    rt_material_metallic(&materials[1], steel, 0.2f);
    rt_material_glass(&materials[2], aqua, 1.5f, 0.6f);
    const float lp[3] = {0.0f, 0.8f * depth, 0.0f}, white[3] = {1.0f, 1.0f, 1.0f};
    rt_light_point(&lights[0], lp, white, 20.0f, INFINITY);
    const float ld[3] = {-0.3f, -1.0f, -0.2f}, warm[3] = {1.0f, 0.95f, 0.9f};
    rt_light_directional(&lights[1], ld, warm, 0.8f);
    // elevated camera looking 45 degrees down at the terrain centre; the terrain fills the frame
    set3(camera->position, 0.0f, 0.4f * depth, 0.56f * depth);
    set3(camera->direction, 0.0f, -0.70710678f, -0.70710678f);
    set3(camera->up, 0.0f, 1.0f, 0.0f);
    camera->fov = 45.0f;
}

// Field(n_tori, seg, ring, gnx, gnz, seed): a scene that is NOT a height field (tree-quality study off the terrain):
// n_tori closed, non-convex meshes (tori of seg x ring quads = 2 seg ring triangles each) at log-uniform scales
// 0.3 .. 6 with random orientations, scattered through the volume [-50,50] x [1,26] x [-50,50] so that they overlap
// in every projection, above a flat ground grid of gnx x gnz quads.  Vertices: n_tori seg ring + (gnx+1)(gnz+1);
// triangles: 2 n_tori seg ring + 2 gnx gnz.  Field(1996, 25, 10, 40, 25) = exactly 1 000 000 triangles.
void rt_scene_field(uint32_t n_tori, uint32_t seg, uint32_t ring, uint32_t gnx, uint32_t gnz, uint32_t seed,
                    RtVertex* vertices, RtTriangle* triangles, RtMaterial materials[3], RtLight lights[2], RtCamera* camera) {
    size_t nv = 0, nt = 0;
    const double two_pi = 6.283185307179586;
    for (uint32_t o = 0; o < n_tori; o++) {
        const double R = 0.3 * pow(20.0, (double)hash01(seed, o, 1u)), r = 0.35 * R;
        const double cx = -50.0 + 100.0 * hash01(seed, o, 2u), cy = 1.0 + R + (25.0 - R) * hash01(seed, o, 3u),
                     cz = -50.0 + 100.0 * hash01(seed, o, 4u);
        // orientation: rotate the torus axis (0,1,0) by angle a about x, then b about y
        const double a = two_pi * hash01(seed, o, 5u), b = two_pi * hash01(seed, o, 6u);
        const double ca = cos(a), sa = sin(a), cb = cos(b), sb = sin(b);
        const size_t base = nv;
        for (uint32_t i = 0; i < seg; i++)
            for (uint32_t j = 0; j < ring; j++) {
                const double u = two_pi * i / seg, v = two_pi * j / ring;
                const double x = (R + r * cos(v)) * cos(u), y = r * sin(v), z = (R + r * cos(v)) * sin(u);
                const double y1 = ca * y - sa * z, z1 = sa * y + ca * z;  // about x
                const double x2 = cb * x + sb * z1, z2 = -sb * x + cb * z1; // about y
                set3(vertices[nv++].position, (float)(cx + x2), (float)(cy + y1), (float)(cz + z2));
            }
        const uint32_t mat = o % 3u;
        for (uint32_t i = 0; i < seg; i++)
            for (uint32_t j = 0; j < ring; j++) {
                const uint32_t i1 = (i + 1) % seg, j1 = (j + 1) % ring;
                const uint32_t v00 = (uint32_t)(base + i * ring + j), v01 = (uint32_t)(base + i * ring + j1),
                               v10 = (uint32_t)(base + i1 * ring + j), v11 = (uint32_t)(base + i1 * ring + j1);
                triangles[nt].v0_index = v00; triangles[nt].v1_index = v01; triangles[nt].v2_index = v11; triangles[nt].material_id = mat; nt++;
                triangles[nt].v0_index = v00; triangles[nt].v1_index = v11; triangles[nt].v2_index = v10; triangles[nt].material_id = mat; nt++;
            }
    }
    const size_t gbase = nv;
    for (uint32_t iz = 0; iz <= gnz; iz++)
        for (uint32_t ix = 0; ix <= gnx; ix++)
            set3(vertices[nv++].position, -60.0f + 120.0f * (float)ix / (float)gnx, 0.0f, -60.0f + 120.0f * (float)iz / (float)gnz);
    for (uint32_t iz = 0; iz < gnz; iz++)
        for (uint32_t ix = 0; ix < gnx; ix++) {
            const uint32_t v00 = (uint32_t)(gbase + iz * (gnx + 1) + ix), v10 = v00 + 1, v01 = v00 + (gnx + 1), v11 = v01 + 1;
            triangles[nt].v0_index = v00; triangles[nt].v1_index = v01; triangles[nt].v2_index = v11; triangles[nt].material_id = 0; nt++;
            triangles[nt].v0_index = v00; triangles[nt].v1_index = v11; triangles[nt].v2_index = v10; triangles[nt].material_id = 0; nt++;
        }
    const float sand[3] = {0.7f, 0.65f, 0.5f}, steel[3] = {0.75f, 0.75f, 0.8f}, aqua[3] = {0.3f, 0.6f, 0.9f};
    rt_material_diffuse(&materials[0], sand);
    rt_material_metallic(&materials[1], steel, 0.2f);
    rt_material_glass(&materials[2], aqua, 1.5f, 0.6f);
    const float lp[3] = {0.0f, 60.0f, 20.0f}, white[3] = {1.0f, 1.0f, 1.0f};
    rt_light_point(&lights[0], lp, white, 30.0f, INFINITY);
    const float ld[3] = {-0.3f, -1.0f, -0.2f}, warm[3] = {1.0f, 0.95f, 0.9f};
    rt_light_directional(&lights[1], ld, warm, 0.8f);
    set3(camera->position, 0.0f, 32.0f, 98.0f);
    set3(camera->direction, 0.0f, -0.24253563f, -0.97014250f);
    set3(camera->up, 0.0f, 1.0f, 0.0f);
    camera->fov = 45.0f;
}

namespace {
struct Builder {
    RtVertex* v; RtTriangle* t; uint32_t nv = 0, nt = 0;
    uint32_t vert(float x, float y, float z) { set3(v[nv].position, x, y, z); return nv++; }
    void tri(uint32_t a, uint32_t b, uint32_t c, uint32_t m) { t[nt].v0_index = a; t[nt].v1_index = b; t[nt].v2_index = c; t[nt].material_id = m; nt++; }
    void quad(const float p[4][3], uint32_t m) {
        uint32_t i0 = vert(p[0][0], p[0][1], p[0][2]), i1 = vert(p[1][0], p[1][1], p[1][2]),
                 i2 = vert(p[2][0], p[2][1], p[2][2]), i3 = vert(p[3][0], p[3][1], p[3][2]);
        tri(i0, i1, i2, m); tri(i0, i2, i3, m);
    }
    void box(float cx, float cz, float hw, float h, float angle, uint32_t m) {
        float c = (float)cos((double)angle), s = (float)sin((double)angle);
        uint32_t id[8];
        for (int k = 0; k < 8; k++) {
            float lx = (k & 1) ? hw : -hw, lz = (k & 2) ? hw : -hw, y = (k & 4) ? -1.0f + h : -1.0f;
            id[k] = vert(cx + c * lx - s * lz, y, cz + s * lx + c * lz);
        }
        const int f[6][4] = {{0, 1, 3, 2}, {4, 6, 7, 5}, {0, 4, 5, 1}, {2, 3, 7, 6}, {0, 2, 6, 4}, {1, 5, 7, 3}};
        for (int q = 0; q < 6; q++) { tri(id[f[q][0]], id[f[q][1]], id[f[q][2]], m); tri(id[f[q][0]], id[f[q][2]], id[f[q][3]], m); }
    }
};
} // namespace

// Cornell box in [-1,1]^3, open towards +z: floor, ceiling, back (white), left (red),
// right (green), a ceiling light quad (emissive) and two white boxes: 36 triangles.
void rt_scene_cornell(RtVertex* vertices, uint32_t* n_vertices, RtTriangle triangles[36], RtMaterial materials[4],
                      RtLight lights[1], RtCamera* camera) {
    Builder b;
    b.v = vertices; b.t = triangles;
    const float fl[4][3] = {{-1, -1, -1}, {-1, -1, 1}, {1, -1, 1}, {1, -1, -1}};
    const float ce[4][3] = {{-1, 1, -1}, {1, 1, -1}, {1, 1, 1}, {-1, 1, 1}};
    const float bk[4][3] = {{-1, -1, -1}, {1, -1, -1}, {1, 1, -1}, {-1, 1, -1}};
    const float lf[4][3] = {{-1, -1, -1}, {-1, 1, -1}, {-1, 1, 1}, {-1, -1, 1}};
    const float rt[4][3] = {{1, -1, -1}, {1, -1, 1}, {1, 1, 1}, {1, 1, -1}};
    const float lq[4][3] = {{-0.3f, 0.995f, -0.3f}, {0.3f, 0.995f, -0.3f}, {0.3f, 0.995f, 0.3f}, {-0.3f, 0.995f, 0.3f}};
    b.quad(fl, 0); b.quad(ce, 0); b.quad(bk, 0); b.quad(lf, 1); b.quad(rt, 2); b.quad(lq, 3);
    b.box(-0.35f, -0.3f, 0.3f, 1.2f, 0.3f, 0);
    b.box(0.4f, 0.3f, 0.28f, 0.6f, -0.25f, 0);
    *n_vertices = b.nv;
    const float white[3] = {0.73f, 0.73f, 0.73f}, red[3] = {0.65f, 0.05f, 0.05f}, green[3] = {0.12f, 0.45f, 0.15f},
                lw[3] = {1.0f, 1.0f, 1.0f}, le[3] = {8.0f, 8.0f, 8.0f};
    rt_material_diffuse(&materials[0], white);
    rt_material_diffuse(&materials[1], red);
    rt_material_diffuse(&materials[2], green);
    rt_material_emissive(&materials[3], lw, le);
    const float lp[3] = {0.0f, 0.9f, 0.0f};
    rt_light_point(&lights[0], lp, lw, 1.0f, INFINITY);
    set3(camera->position, 0.0f, 0.0f, 3.4f);
    set3(camera->direction, 0.0f, 0.0f, -1.0f);
    set3(camera->up, 0.0f, 1.0f, 0.0f);
    camera->fov = 45.0f;
}

} // extern "C"
