// ORACLE — TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product; only tests/,
// __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may build or call it.
//
// orc_abi.h — independent restatement of the reference's shared buffer structs
// (reference: renderer/Raytracing.h:35-123, renderer/Spectrum.h:3-21).  Deliberately NOT shared with
// include/mrt_abi.h: the two are pinned against each other (and against the layouts produced by
// compiling the reference header itself, oracle/ref_probe.cpp -> tests/golden/abi_layout.json) by tests.
#pragma once
#include <cstddef>
#include <cstdint>

namespace orc {

constexpr int   SPECTRUM_SAMPLES = 3;                 // Spectrum.h:3
constexpr float DISTANCE_EPSILON = 0.0001f;           // Raytracing.h:16
constexpr float ANGLE_EPSILON    = 0.00003807693583f; // Raytracing.h:17
constexpr float PI               = 3.1415926f;        // Raytracing.h:18
constexpr float INV_PI           = 1.0f / PI;         // "(1.0 / PI)" KernelHelpers.h:78,108,141,171
constexpr uint32_t NOISE_DIMENSIONS = 64;             // Raytracing.h:19
constexpr uint32_t NOISE_SLOTS   = 3;                 // MaxBuffersInFlight Renderer.mm:16

enum : uint32_t {                                     // Raytracing.h:35-43
    MATERIAL_DIFFUSE, MATERIAL_MIRROR, MATERIAL_SMOOTH_PLASTIC, MATERIAL_SMOOTH_DIELECTRIC, MATERIAL_COUNT
};

struct Spectrum { float values[SPECTRUM_SAMPLES]; };  // Spectrum.h:6-21

struct SharedData {                                   // Raytracing.h:47-52
    uint32_t frameIndex;
    uint32_t lightTrianglesCount;
    float time;
};

struct alignas(16) Ray {                              // Raytracing.h:54-69 (80 bytes, not 68)
    float origin[3];
    float minDistance;
    float direction[3];
    float maxDistance;
    Spectrum throughput;
    Spectrum radiance;
    float pad_[2];
    float params[4];   // x: material pdf, y: "light pdf" flag (prev bounce diffuse), z: bounce, w: ior
};

struct LightSamplingRay {                             // Raytracing.h:71-83
    float origin[3];
    float minDistance;
    float direction[3];
    float maxDistance;
    Spectrum throughput;
    uint32_t targetIndex;
};

struct Intersection {                                 // Raytracing.h:85-90
    float distance;
    uint32_t triangleIndex;
    float coordinates[2];
};

struct Vertex { float v[3]; float n[3]; };            // Raytracing.h:92-96

struct Material {                                     // Raytracing.h:98-104
    Spectrum diffuse;
    Spectrum emissive;
    float ior;
    uint32_t materialType;
};

struct TriangleReference {                            // Raytracing.h:106-111
    uint32_t tri[3];
    uint32_t materialIndex;
    uint32_t lightTriangleIndex;
};

struct LightTriangle {                                // Raytracing.h:113-123
    Spectrum emissive;
    Vertex v1, v2, v3;
    float area, pdf, cdf;
    uint32_t index;
};

static_assert(sizeof(Ray) == 80 && offsetof(Ray, params) == 64, "Ray");
static_assert(sizeof(LightSamplingRay) == 48, "LightSamplingRay");
static_assert(sizeof(Intersection) == 16, "Intersection");
static_assert(sizeof(Vertex) == 24 && sizeof(Material) == 32, "Vertex/Material");
static_assert(sizeof(TriangleReference) == 20 && sizeof(LightTriangle) == 100, "refs/lights");

}  // namespace orc
