// scene.cpp — procedural scenes, light list construction and noise tiles (host only, no CUDA).
//
// Caller-side of the hot path: everything -initRaytracing does before it uploads buffers
// (reference: renderer/Renderer.mm:257-448) — material table, vertex/index/TriangleReference arrays, the
// emissive-triangle list with area pdf / exclusive cdf / zero-area sentinel.  Scene CONTENT is ours (the
// reference loads OBJ files through SceneKit; see obj_loader.cpp); all scenes sit in the reference's
// hard-wired camera frustum: eye (0,1,2.35) looking down -z, 90 degree horizontal FOV (Shaders.metal:84-97).
#include "scene.h"
#include <cmath>
#include <cstring>
#include <map>
#include <random>

namespace mrt {

// ---- deterministic generator for scene layout (splitmix64; identical on every platform) ----------------
struct Rng {
    uint64_t s;
    explicit Rng(uint64_t seed) : s(seed * 0x9E3779B97F4A7C15ull + 0x1234567ull) {}
    uint64_t next() { uint64_t z = (s += 0x9E3779B97F4A7C15ull); z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
                      z = (z ^ (z >> 27)) * 0x94D049BB133111EBull; return z ^ (z >> 31); }
    float uniform() { return float(next() >> 40) * (1.0f / 16777216.0f); }          // [0,1)
    float range(float a, float b) { return a + (b - a) * uniform(); }
};

uint32_t add_vertex(mrt_scene& s, float x, float y, float z, float nx, float ny, float nz) {
    s.vertices.push_back(mrt_vertex{{x, y, z}, {nx, ny, nz}});
    return uint32_t(s.vertices.size() - 1);
}
uint32_t add_material(mrt_scene& s, float r, float g, float b, float er, float eg, float eb, float ior, uint32_t type) {
    mrt_material m{};
    m.diffuse.values[0] = r; m.diffuse.values[1] = g; m.diffuse.values[2] = b;
    m.emissive.values[0] = er; m.emissive.values[1] = eg; m.emissive.values[2] = eb;
    m.ior = ior; m.materialType = type;
    s.materials.push_back(m);
    return uint32_t(s.materials.size() - 1);
}
void add_triangle(mrt_scene& s, uint32_t a, uint32_t b, uint32_t c, uint32_t mat) {
    s.indices.push_back(a); s.indices.push_back(b); s.indices.push_back(c);
    mrt_triangle_reference r{};
    r.tri[0] = a; r.tri[1] = b; r.tri[2] = c; r.materialIndex = mat; r.lightTriangleIndex = MRT_NOT_A_LIGHT;
    s.references.push_back(r);
}

// Renderer.mm:373-448: a triangle is a light iff its material has any emissive component > 0 (:378-381);
// area = 0.5*|cross(p2-p1, p3-p1)| (:408); pdf = area/totalArea, cdf = exclusive prefix sum (:435-442);
// one sentinel {cdf = sum, pdf = 1, area = 0} is appended and NOT counted (:444-448).  All in float.
void finalize_lights(mrt_scene& s) {
    s.lights.clear();
    float totalArea = 0.0f;
    for (size_t t = 0; t < s.references.size(); ++t) {
        mrt_triangle_reference& ref = s.references[t];
        const mrt_material& m = s.materials[ref.materialIndex];
        bool isEmitter = m.emissive.values[0] > 0.0f || m.emissive.values[1] > 0.0f || m.emissive.values[2] > 0.0f;
        ref.lightTriangleIndex = MRT_NOT_A_LIGHT;
        if (!isEmitter) continue;
        const mrt_vertex& v1 = s.vertices[ref.tri[0]];
        const mrt_vertex& v2 = s.vertices[ref.tri[1]];
        const mrt_vertex& v3 = s.vertices[ref.tri[2]];
        float ax = v2.v[0] - v1.v[0], ay = v2.v[1] - v1.v[1], az = v2.v[2] - v1.v[2];
        float bx = v3.v[0] - v1.v[0], by = v3.v[1] - v1.v[1], bz = v3.v[2] - v1.v[2];
        float cx = ay * bz - az * by, cy = az * bx - ax * bz, cz = ax * by - ay * bx;
        mrt_light_triangle lt{};
        lt.index = uint32_t(t);
        lt.v1 = v1; lt.v2 = v2; lt.v3 = v3;
        lt.area = 0.5f * std::sqrt(cx * cx + cy * cy + cz * cz);
        lt.emissive = m.emissive;
        ref.lightTriangleIndex = uint32_t(s.lights.size());
        s.lights.push_back(lt);
        totalArea += lt.area;
    }
    float cdf = 0.0f;
    for (mrt_light_triangle& lt : s.lights) { lt.pdf = lt.area / totalArea; lt.cdf = cdf; cdf += lt.pdf; }
    s.light_count = uint32_t(s.lights.size());
    mrt_light_triangle sentinel{};
    sentinel.cdf = cdf; sentinel.pdf = 1.0f; sentinel.area = 0.0f;
    s.lights.push_back(sentinel);
}

// ---- primitives ----------------------------------------------------------------------------------------
// Quad p0,p1,p2,p3 (counter-clockwise seen from the side the normal points to), flat normal n.
static void add_quad(mrt_scene& s, const float p[4][3], const float n[3], uint32_t mat) {
    uint32_t v[4];
    for (int i = 0; i < 4; ++i) v[i] = add_vertex(s, p[i][0], p[i][1], p[i][2], n[0], n[1], n[2]);
    add_triangle(s, v[0], v[1], v[2], mat);
    add_triangle(s, v[2], v[3], v[0], mat);
}

// Axis-aligned box rotated by `angle` about +y around its centre; outward normals; 12 triangles.
static void add_box(mrt_scene& s, float cx, float cy, float cz, float hx, float hy, float hz, float angle, uint32_t mat) {
    float ca = std::cos(angle), sa = std::sin(angle);
    auto P = [&](float x, float y, float z, float* o) { o[0] = cx + ca * x + sa * z; o[1] = cy + y; o[2] = cz - sa * x + ca * z; };
    auto N = [&](float x, float y, float z, float* o) { o[0] = ca * x + sa * z; o[1] = y; o[2] = -sa * x + ca * z; };
    const float f[6][4][3] = {
        {{+1, -1, -1}, {+1, +1, -1}, {+1, +1, +1}, {+1, -1, +1}},   // +x
        {{-1, -1, +1}, {-1, +1, +1}, {-1, +1, -1}, {-1, -1, -1}},   // -x
        {{-1, +1, -1}, {-1, +1, +1}, {+1, +1, +1}, {+1, +1, -1}},   // +y
        {{-1, -1, +1}, {-1, -1, -1}, {+1, -1, -1}, {+1, -1, +1}},   // -y
        {{-1, -1, +1}, {+1, -1, +1}, {+1, +1, +1}, {-1, +1, +1}},   // +z
        {{+1, -1, -1}, {-1, -1, -1}, {-1, +1, -1}, {+1, +1, -1}}};  // -z
    const float fn[6][3] = {{1, 0, 0}, {-1, 0, 0}, {0, 1, 0}, {0, -1, 0}, {0, 0, 1}, {0, 0, -1}};
    for (int k = 0; k < 6; ++k) {
        float p[4][3], n[3];
        for (int i = 0; i < 4; ++i) P(f[k][i][0] * hx, f[k][i][1] * hy, f[k][i][2] * hz, p[i]);
        N(fn[k][0], fn[k][1], fn[k][2], n);
        add_quad(s, p, n, mat);
    }
}

// Unit icosphere template (shared vertices, outward winding): 20*4^subdiv triangles.
struct Icosphere {
    std::vector<float> pos;        // unit vectors
    std::vector<uint32_t> tri;
    explicit Icosphere(int subdiv) {
        const double t = (1.0 + std::sqrt(5.0)) / 2.0;
        const double v[12][3] = {{-1, t, 0}, {1, t, 0}, {-1, -t, 0}, {1, -t, 0}, {0, -1, t}, {0, 1, t},
                                 {0, -1, -t}, {0, 1, -t}, {t, 0, -1}, {t, 0, 1}, {-t, 0, -1}, {-t, 0, 1}};
        for (auto& p : v) push(p[0], p[1], p[2]);
        const uint32_t f[20][3] = {{0, 11, 5}, {0, 5, 1}, {0, 1, 7}, {0, 7, 10}, {0, 10, 11}, {1, 5, 9}, {5, 11, 4},
                                   {11, 10, 2}, {10, 7, 6}, {7, 1, 8}, {3, 9, 4}, {3, 4, 2}, {3, 2, 6}, {3, 6, 8},
                                   {3, 8, 9}, {4, 9, 5}, {2, 4, 11}, {6, 2, 10}, {8, 6, 7}, {9, 8, 1}};
        for (auto& q : f) { tri.push_back(q[0]); tri.push_back(q[1]); tri.push_back(q[2]); }
        for (int level = 0; level < subdiv; ++level) {
            std::map<uint64_t, uint32_t> mid;
            std::vector<uint32_t> next;
            auto midpoint = [&](uint32_t a, uint32_t b) {
                uint64_t key = a < b ? (uint64_t(a) << 32) | b : (uint64_t(b) << 32) | a;
                auto it = mid.find(key);
                if (it != mid.end()) return it->second;
                uint32_t id = push(0.5 * (double(pos[3 * a]) + pos[3 * b]), 0.5 * (double(pos[3 * a + 1]) + pos[3 * b + 1]),
                                   0.5 * (double(pos[3 * a + 2]) + pos[3 * b + 2]));
                mid[key] = id;
                return id;
            };
            for (size_t k = 0; k < tri.size(); k += 3) {
                uint32_t a = tri[k], b = tri[k + 1], c = tri[k + 2];
                uint32_t ab = midpoint(a, b), bc = midpoint(b, c), ca = midpoint(c, a);
                const uint32_t q[12] = {a, ab, ca, b, bc, ab, c, ca, bc, ab, bc, ca};
                next.insert(next.end(), q, q + 12);
            }
            tri.swap(next);
        }
    }
    uint32_t push(double x, double y, double z) {
        double l = std::sqrt(x * x + y * y + z * z);
        pos.push_back(float(x / l)); pos.push_back(float(y / l)); pos.push_back(float(z / l));
        return uint32_t(pos.size() / 3 - 1);
    }
};

static void add_sphere(mrt_scene& s, const Icosphere& ico, float cx, float cy, float cz, float radius, uint32_t mat) {
    uint32_t base = uint32_t(s.vertices.size());
    for (size_t k = 0; k < ico.pos.size(); k += 3) {
        float nx = ico.pos[k], ny = ico.pos[k + 1], nz = ico.pos[k + 2];
        add_vertex(s, cx + radius * nx, cy + radius * ny, cz + radius * nz, nx, ny, nz);
    }
    for (size_t k = 0; k < ico.tri.size(); k += 3) add_triangle(s, base + ico.tri[k], base + ico.tri[k + 1], base + ico.tri[k + 2], mat);
}

// Cornell shell: floor, ceiling, back, left, right (10 triangles, normals facing inward), box [-1,1]x[0,2]x[-1,1],
// open toward the camera (+z).
static void add_cornell_shell(mrt_scene& s, uint32_t white, uint32_t left, uint32_t right) {
    const float floor_[4][3] = {{-1, 0, 1}, {1, 0, 1}, {1, 0, -1}, {-1, 0, -1}};    const float nf[3] = {0, 1, 0};
    const float ceil_[4][3] = {{-1, 2, -1}, {1, 2, -1}, {1, 2, 1}, {-1, 2, 1}};     const float nc[3] = {0, -1, 0};
    const float back_[4][3] = {{-1, 0, -1}, {1, 0, -1}, {1, 2, -1}, {-1, 2, -1}};   const float nb[3] = {0, 0, 1};
    const float left_[4][3] = {{-1, 0, 1}, {-1, 0, -1}, {-1, 2, -1}, {-1, 2, 1}};   const float nl[3] = {1, 0, 0};
    const float right_[4][3] = {{1, 0, -1}, {1, 0, 1}, {1, 2, 1}, {1, 2, -1}};      const float nr[3] = {-1, 0, 0};
    add_quad(s, floor_, nf, white); add_quad(s, ceil_, nc, white); add_quad(s, back_, nb, white);
    add_quad(s, left_, nl, left); add_quad(s, right_, nr, right);
}

// Emissive quad just under the ceiling, facing down: 2 light triangles.
static void add_ceiling_light(mrt_scene& s, float x0, float z0, float x1, float z1, float y, uint32_t mat) {
    const float q[4][3] = {{x0, y, z0}, {x1, y, z0}, {x1, y, z1}, {x0, y, z1}};
    const float n[3] = {0, -1, 0};
    add_quad(s, q, n, mat);
}

// ---- the five benchmark scenes (SURVEY.md §8d / BASELINE.md §3) -----------------------------------------
enum { DIFFUSE = MRT_MATERIAL_DIFFUSE, MIRROR = MRT_MATERIAL_MIRROR, PLASTIC = MRT_MATERIAL_SMOOTH_PLASTIC,
       DIELECTRIC = MRT_MATERIAL_SMOOTH_DIELECTRIC };

// config 1: Cornell box, 36 triangles: 10 shell + 2 light (Le = 5,4,3 as Media/cornellbox.mtl's `light`) + 2 boxes.
static void scene_cornell(mrt_scene& s) {
    uint32_t white = add_material(s, 0.73f, 0.73f, 0.73f, 0, 0, 0, 1.0f, DIFFUSE);
    uint32_t red = add_material(s, 0.65f, 0.05f, 0.05f, 0, 0, 0, 1.0f, DIFFUSE);
    uint32_t green = add_material(s, 0.12f, 0.45f, 0.15f, 0, 0, 0, 1.0f, DIFFUSE);
    uint32_t light = add_material(s, 1.0f, 1.0f, 1.0f, 5.0f, 4.0f, 3.0f, 1.0f, DIFFUSE);
    add_cornell_shell(s, white, red, green);
    add_ceiling_light(s, -0.25f, -0.25f, 0.25f, 0.25f, 1.98f, light);
    add_box(s, 0.33f, 0.30f, 0.35f, 0.30f, 0.30f, 0.30f, -0.30f, white);    // short box
    add_box(s, -0.35f, 0.60f, -0.30f, 0.30f, 0.60f, 0.30f, 0.31f, white);   // tall box
}

// config 2: procedural equivalent of Media/CornellBox-Water-plastic (14,176 triangles there; 14,124 here):
// shell + light (Le = 10) + a plastic sphere (ior 1.5) + a mirror sphere + a rippled water surface of
// 76 x 76 quads with analytic smooth normals, plastic ior 1.33333.
static void scene_cornell_water(mrt_scene& s) {
    uint32_t white = add_material(s, 0.725f, 0.71f, 0.68f, 0, 0, 0, 1.0f, DIFFUSE);
    uint32_t left = add_material(s, 0.63f, 0.065f, 0.05f, 0, 0, 0, 1.0f, DIFFUSE);
    uint32_t right = add_material(s, 0.161f, 0.133f, 0.427f, 0, 0, 0, 1.0f, DIFFUSE);
    uint32_t light = add_material(s, 0.725f, 0.71f, 0.68f, 10.0f, 10.0f, 10.0f, 1.0f, DIFFUSE);
    uint32_t plastic = add_material(s, 1.0f, 1.0f, 1.0f, 0, 0, 0, 1.5f, PLASTIC);
    uint32_t mirror = add_material(s, 1.0f, 1.0f, 1.0f, 0, 0, 0, 0.0f, MIRROR);
    uint32_t water = add_material(s, 1.0f, 1.0f, 1.0f, 0, 0, 0, 1.33333f, PLASTIC);
    add_cornell_shell(s, white, left, right);
    add_ceiling_light(s, -0.24f, -0.22f, 0.23f, 0.16f, 1.98f, light);
    Icosphere ico(3);
    add_sphere(s, ico, -0.42f, 0.95f, -0.25f, 0.33f, plastic);
    add_sphere(s, ico, 0.45f, 0.80f, 0.30f, 0.33f, mirror);
    const int N = 76;
    const float y0 = 0.42f, amp = 0.018f, kx = 9.0f, kz = 7.0f;
    auto h = [&](float x, float z) { return y0 + amp * (std::sin(kx * x) * std::cos(kz * z) + 0.5f * std::sin(2.3f * kx * x + 1.7f * kz * z)); };
    uint32_t base = uint32_t(s.vertices.size());
    for (int j = 0; j <= N; ++j)
        for (int i = 0; i <= N; ++i) {
            float x = -1.0f + 2.0f * float(i) / N, z = -1.0f + 2.0f * float(j) / N;
            float dhdx = amp * (kx * std::cos(kx * x) * std::cos(kz * z) + 0.5f * 2.3f * kx * std::cos(2.3f * kx * x + 1.7f * kz * z));
            float dhdz = amp * (-kz * std::sin(kx * x) * std::sin(kz * z) + 0.5f * 1.7f * kz * std::cos(2.3f * kx * x + 1.7f * kz * z));
            float l = std::sqrt(dhdx * dhdx + 1.0f + dhdz * dhdz);
            add_vertex(s, x, h(x, z), z, -dhdx / l, 1.0f / l, -dhdz / l);
        }
    for (int j = 0; j < N; ++j)
        for (int i = 0; i < N; ++i) {
            uint32_t a = base + j * (N + 1) + i, b = a + 1, c = a + (N + 1) + 1, d = a + (N + 1);
            add_triangle(s, a, d, c, water);
            add_triangle(s, c, b, a, water);
        }
}

// configs 3 and 4: Cornell shell (10 wall + 2 light triangles) + `count` flattened copies of an icosphere on a
// jittered nx x ny x nz lattice inside [-0.9,0.9] x [0.05,1.55] x [-0.9,0.9]; materials round-robin over the four
// material kinds (ior 1.5).  "Instanced" = generated by replication, stored flat (the reference has no instancing).
static void scene_spheres(mrt_scene& s, uint64_t seed, int subdiv, int count, int nx, int ny, int nz, float radius) {
    uint32_t white = add_material(s, 0.73f, 0.73f, 0.73f, 0, 0, 0, 1.0f, DIFFUSE);
    uint32_t red = add_material(s, 0.65f, 0.05f, 0.05f, 0, 0, 0, 1.0f, DIFFUSE);
    uint32_t green = add_material(s, 0.12f, 0.45f, 0.15f, 0, 0, 0, 1.0f, DIFFUSE);
    uint32_t light = add_material(s, 1.0f, 1.0f, 1.0f, 12.0f, 12.0f, 12.0f, 1.0f, DIFFUSE);
    uint32_t kinds[4];
    kinds[0] = add_material(s, 0.80f, 0.55f, 0.25f, 0, 0, 0, 1.0f, DIFFUSE);
    kinds[1] = add_material(s, 0.95f, 0.95f, 0.95f, 0, 0, 0, 0.0f, MIRROR);
    kinds[2] = add_material(s, 0.30f, 0.55f, 0.85f, 0, 0, 0, 1.5f, PLASTIC);
    kinds[3] = add_material(s, 1.0f, 1.0f, 1.0f, 0, 0, 0, 1.5f, DIELECTRIC);
    add_cornell_shell(s, white, red, green);
    add_ceiling_light(s, -0.5f, -0.5f, 0.5f, 0.5f, 1.98f, light);
    Icosphere ico(subdiv);
    Rng rng(seed);
    s.vertices.reserve(s.vertices.size() + size_t(count) * ico.pos.size() / 3);
    s.indices.reserve(s.indices.size() + size_t(count) * ico.tri.size());
    s.references.reserve(s.references.size() + size_t(count) * ico.tri.size() / 3);
    const float sx = 1.8f / nx, sy = 1.5f / ny, sz = 1.8f / nz;
    for (int k = 0; k < count; ++k) {
        int ix = k % nx, iy = (k / nx) % ny, iz = k / (nx * ny);
        float jx = rng.range(-0.5f, 0.5f) * (sx - 2.0f * radius), jy = rng.range(-0.5f, 0.5f) * (sy - 2.0f * radius),
              jz = rng.range(-0.5f, 0.5f) * (sz - 2.0f * radius);
        float cx = -0.9f + (ix + 0.5f) * sx + jx, cy = 0.05f + (iy + 0.5f) * sy + jy, cz = -0.9f + (iz + 0.5f) * sz + jz;
        add_sphere(s, ico, cx, cy, cz, radius, kinds[k & 3]);
    }
}

// config 5: Cornell shell + 32 x 32 emissive ceiling quads (2,048 light triangles, Le in U[1,20]^3, one material
// each) + 64 plastic + 64 dielectric spheres (ior 1.5, 320 triangles each).
static void scene_many_lights(mrt_scene& s, uint64_t seed) {
    uint32_t white = add_material(s, 0.73f, 0.73f, 0.73f, 0, 0, 0, 1.0f, DIFFUSE);
    uint32_t red = add_material(s, 0.65f, 0.05f, 0.05f, 0, 0, 0, 1.0f, DIFFUSE);
    uint32_t green = add_material(s, 0.12f, 0.45f, 0.15f, 0, 0, 0, 1.0f, DIFFUSE);
    uint32_t plastic = add_material(s, 0.85f, 0.35f, 0.30f, 0, 0, 0, 1.5f, PLASTIC);
    uint32_t glass = add_material(s, 1.0f, 1.0f, 1.0f, 0, 0, 0, 1.5f, DIELECTRIC);
    add_cornell_shell(s, white, red, green);
    Rng rng(seed);
    const int G = 32;
    for (int j = 0; j < G; ++j)
        for (int i = 0; i < G; ++i) {
            float er = rng.range(1.0f, 20.0f), eg = rng.range(1.0f, 20.0f), eb = rng.range(1.0f, 20.0f);
            uint32_t m = add_material(s, 1.0f, 1.0f, 1.0f, er, eg, eb, 1.0f, DIFFUSE);
            float cell = 1.9f / G, x0 = -0.95f + i * cell, z0 = -0.95f + j * cell;
            add_ceiling_light(s, x0 + 0.2f * cell, z0 + 0.2f * cell, x0 + 0.8f * cell, z0 + 0.8f * cell, 1.99f, m);
        }
    Icosphere ico(2);
    for (int k = 0; k < 128; ++k) {
        int ix = k % 8, iz = (k / 8) % 8, iy = k / 64;
        float cx = -0.84f + ix * 0.24f + rng.range(-0.03f, 0.03f), cz = -0.84f + iz * 0.24f + rng.range(-0.03f, 0.03f);
        float cy = 0.12f + iy * 0.55f + rng.range(0.0f, 0.25f);
        add_sphere(s, ico, cx, cy, cz, 0.085f, (k & 1) ? glass : plastic);
    }
}

// white-box: the all-white furnace-like scene of Media/white-box.obj (12 triangles): shell + pi-radiance light.
static void scene_white_box(mrt_scene& s) {
    uint32_t white = add_material(s, 1.0f, 1.0f, 1.0f, 0, 0, 0, 1.0f, DIFFUSE);
    uint32_t light = add_material(s, 1.0f, 1.0f, 1.0f, 3.1415926536f, 3.1415926536f, 3.1415926536f, 1.0f, DIFFUSE);
    add_cornell_shell(s, white, white, white);
    add_ceiling_light(s, -0.25f, -0.25f, 0.25f, 0.25f, 1.98f, light);
}

bool make_procedural(const std::string& name, uint64_t seed, mrt_scene& s, std::string& err) {
    s = mrt_scene();
    s.name = name;
    if (name == "cornell") scene_cornell(s);
    else if (name == "white-box") scene_white_box(s);
    else if (name == "cornell-water") scene_cornell_water(s);
    else if (name == "spheres-1m") scene_spheres(s, seed, 3, 800, 10, 8, 10, 0.06f);        // 1,024,012 triangles
    else if (name == "spheres-10m") scene_spheres(s, seed, 4, 1953, 13, 12, 13, 0.05f);     // 9,999,372 triangles
    else if (name == "spheres-64k") scene_spheres(s, seed, 2, 200, 6, 6, 6, 0.09f);         // 64,012 (test size)
    else if (name == "many-lights") scene_many_lights(s, seed);
    else { err = "unknown procedural scene '" + name + "'"; return false; }
    finalize_lights(s);
    return true;
}

// Noise tile for (seed, frame): 64*64 float4 of (mt19937_64(seed ^ frame)() >> 40) * 2^-24, uniform on [0,1).
// Stands in for the wall-clock-seeded std::uniform_real_distribution fills of Renderer.mm:109-126, 486-495
// (those are neither reproducible nor portable across standard libraries; mt19937_64 itself is fully specified).
void noise_fill(uint64_t seed, uint32_t frame, float* tile) {
    std::mt19937_64 rng(seed ^ uint64_t(frame));
    for (int i = 0; i < MRT_NOISE_TILE_FLOATS; ++i) tile[i] = float(rng() >> 40) * (1.0f / 16777216.0f);
}

}  // namespace mrt
