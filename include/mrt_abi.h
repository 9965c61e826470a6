/*
 * mrt_abi.h — buffer ABI of the path-tracing hot path (plain C, no CUDA/torch types).
 *
 * These are the layouts the reference shares between its host code and its Metal kernels
 * (reference: renderer/Raytracing.h:35-123, renderer/Spectrum.h:3-21).  Sizes and offsets are the
 * REAL ones produced by compiling that header (Ray is 80 bytes / align 16, not the 68 bytes its
 * comment claims, Raytracing.h:67-68); they are pinned by the static asserts below and by
 * tests/golden/abi_layout.json, which is generated from the reference header itself
 * (oracle/ref_probe.cpp).
 */
#ifndef MRT_ABI_H
#define MRT_ABI_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Tunables; defaults of mrt_config equal these (reference: renderer/Raytracing.h:11-25). */
#define MRT_SPECTRUM_SAMPLES   3            /* Spectrum.h:3 */
#define MRT_DISTANCE_EPSILON   0.0001f      /* Raytracing.h:16 */
#define MRT_ANGLE_EPSILON      0.00003807693583f /* Raytracing.h:17 */
#define MRT_PI                 3.1415926f   /* Raytracing.h:18 (sic: 8 digits) */
#define MRT_NOISE_DIMENSIONS   64           /* Raytracing.h:19 */
#define MRT_NOISE_SLOTS        3            /* MaxBuffersInFlight, Renderer.mm:16 */
#define MRT_MAX_PATH_LENGTH    8            /* Raytracing.h:23 (default; runtime parameter here) */
#define MRT_NOISE_TILE_FLOATS  (4 * MRT_NOISE_DIMENSIONS * MRT_NOISE_DIMENSIONS)
#define MRT_NOT_A_LIGHT        0xFFFFFFFFu  /* uint32_t(-1), Renderer.mm:393 */
/* comparison modes of the presentation pass, Raytracing.h:27-31 */
#define MRT_COMPARE_DISABLED        0
#define MRT_COMPARE_ABSOLUTE_VALUE  1          /* |color - ref| */
#define MRT_COMPARE_REF_TO_COLOR    2          /* visible where the output is darker than the reference */
#define MRT_COMPARE_COLOR_TO_REF    3          /* visible where the reference is darker than the output */
#define MRT_COMPARE_LUMINANCE       4          /* red: output brighter, green: reference brighter */
#define MRT_COMPARISON_SCALE        10.0f      /* Raytracing.h:33 */

/* Material kinds (Raytracing.h:35-43). */
enum {
    MRT_MATERIAL_DIFFUSE = 0,
    MRT_MATERIAL_MIRROR = 1,
    MRT_MATERIAL_SMOOTH_PLASTIC = 2,
    MRT_MATERIAL_SMOOTH_DIELECTRIC = 3,
    MRT_MATERIAL_COUNT = 4
};

/* Spectrum.h:6-21 — three plain floats (RGB), no wavelength sampling in the reference. */
typedef struct mrt_spectrum { float values[MRT_SPECTRUM_SAMPLES]; } mrt_spectrum;

/* Raytracing.h:47-52 */
typedef struct mrt_shared_data {
    uint32_t frameIndex;
    uint32_t lightTrianglesCount;
    float    time;
} mrt_shared_data;

/* Raytracing.h:54-69.  params = (bsdf pdf, previous-bounce-was-diffuse flag, bounce, ior). */
typedef struct
#if defined(__cplusplus) || defined(__GNUC__)
    __attribute__((aligned(16)))
#endif
mrt_ray {
    float origin[3];
    float minDistance;
    float direction[3];
    float maxDistance;
    mrt_spectrum throughput;
    mrt_spectrum radiance;
    float _pad[2];          /* implicit in the reference: vector_float4 is 16-aligned */
    float params[4];
} mrt_ray;

/* Raytracing.h:71-83 */
typedef struct mrt_light_sampling_ray {
    float origin[3];
    float minDistance;
    float direction[3];
    float maxDistance;
    mrt_spectrum throughput;
    uint32_t targetIndex;
} mrt_light_sampling_ray;

/* Raytracing.h:85-90.  coordinates = barycentric weights of triangle vertices 0 and 1
 * (MPS convention, consumed by KernelHelpers.h:39-42).  Miss: distance < 0. */
typedef struct mrt_intersection {
    float    distance;
    uint32_t triangleIndex;
    float    coordinates[2];
} mrt_intersection;

/* Raytracing.h:92-96 */
typedef struct mrt_vertex { float v[3]; float n[3]; } mrt_vertex;

/* Raytracing.h:98-104 */
typedef struct mrt_material {
    mrt_spectrum diffuse;
    mrt_spectrum emissive;
    float    ior;
    uint32_t materialType;
} mrt_material;

/* Raytracing.h:106-111 */
typedef struct mrt_triangle_reference {
    uint32_t tri[3];
    uint32_t materialIndex;
    uint32_t lightTriangleIndex;   /* MRT_NOT_A_LIGHT if the triangle does not emit */
} mrt_triangle_reference;

/* Raytracing.h:113-123.  The light array carries ONE extra sentinel entry
 * {cdf = sum of pdfs, pdf = 1, area = 0, rest 0} (Renderer.mm:444-448). */
typedef struct mrt_light_triangle {
    mrt_spectrum emissive;
    mrt_vertex v1, v2, v3;
    float    area;
    float    pdf;
    float    cdf;
    uint32_t index;                /* index into the TriangleReference array */
} mrt_light_triangle;

#if defined(__cplusplus)
#define MRT_SA(c, m) static_assert(c, m)
#elif defined(__STDC_VERSION__) && __STDC_VERSION__ >= 201112L
#define MRT_SA(c, m) _Static_assert(c, m)
#else
#define MRT_SA(c, m)
#endif
MRT_SA(sizeof(mrt_spectrum) == 12, "Spectrum");
MRT_SA(sizeof(mrt_shared_data) == 12, "SharedData");
MRT_SA(sizeof(mrt_ray) == 80, "Ray is 80 bytes");
MRT_SA(offsetof(mrt_ray, direction) == 16 && offsetof(mrt_ray, maxDistance) == 28, "Ray");
MRT_SA(offsetof(mrt_ray, throughput) == 32 && offsetof(mrt_ray, radiance) == 44, "Ray");
MRT_SA(offsetof(mrt_ray, params) == 64, "Ray.params at 64");
MRT_SA(sizeof(mrt_light_sampling_ray) == 48, "LightSamplingRay");
MRT_SA(offsetof(mrt_light_sampling_ray, throughput) == 32, "LightSamplingRay");
MRT_SA(offsetof(mrt_light_sampling_ray, targetIndex) == 44, "LightSamplingRay");
MRT_SA(sizeof(mrt_intersection) == 16, "Intersection");
MRT_SA(sizeof(mrt_vertex) == 24, "Vertex");
MRT_SA(sizeof(mrt_material) == 32, "Material");
MRT_SA(offsetof(mrt_material, ior) == 24 && offsetof(mrt_material, materialType) == 28, "Material");
MRT_SA(sizeof(mrt_triangle_reference) == 20, "TriangleReference");
MRT_SA(sizeof(mrt_light_triangle) == 100, "LightTriangle");
MRT_SA(offsetof(mrt_light_triangle, v1) == 12 && offsetof(mrt_light_triangle, v2) == 36, "LightTriangle");
MRT_SA(offsetof(mrt_light_triangle, v3) == 60 && offsetof(mrt_light_triangle, area) == 84, "LightTriangle");
MRT_SA(offsetof(mrt_light_triangle, pdf) == 88 && offsetof(mrt_light_triangle, cdf) == 92, "LightTriangle");
MRT_SA(offsetof(mrt_light_triangle, index) == 96, "LightTriangle");

#ifdef __cplusplus
}
#endif
#endif /* MRT_ABI_H */
