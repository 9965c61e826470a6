// ORACLE — TEST INFRASTRUCTURE ONLY.
// ref_probe.cpp — compiles the REFERENCE's own renderer/Raytracing.h (+ Spectrum.h) where it lies under
// /root/reference (host half only; the device half is MSL behind __METAL_VERSION__) and prints, as JSON:
//   * sizeof / offsetof of every shared struct  -> tests/golden/abi_layout.json
//   * outputs of the header's host-callable functions on fixed inputs -> tests/golden/ref_host_kats.json
// Built by oracle/Makefile into oracle/_ref/ref_probe (container only; /root/reference is absent on the GPU box).
#include <cmath>
#include <cstddef>
#include <cstdio>
using std::pow;
#include "Raytracing.h"

// JSON-safe float: Python's json reads NaN / Infinity / -Infinity
static const char* J(double v) {
    static char buf[16][40]; static int k = 0; char* b = buf[k++ & 15];
    if (v != v) snprintf(b, 40, "NaN"); else if (v > 1e300) snprintf(b, 40, "Infinity"); else if (v < -1e300) snprintf(b, 40, "-Infinity");
    else snprintf(b, 40, "%.9g", v);
    return b;
}
#define SZ(T) printf("    \"%s\": {\"size\": %zu, \"align\": %zu", #T, sizeof(T), alignof(T))
#define OFF(T, m) printf(", \"%s\": %zu", #m, offsetof(T, m))
#define END() printf("}")

int main(int argc, char** argv) {
    bool kats = argc > 1 && argv[1][0] == 'k';
    if (!kats) {
        printf("{\n  \"source\": \"renderer/Raytracing.h + renderer/Spectrum.h compiled with g++ and oracle/ref_stub\",\n  \"structs\": {\n");
        SZ(Spectrum); OFF(Spectrum, values); END(); printf(",\n");
        SZ(SharedData); OFF(SharedData, frameIndex); OFF(SharedData, lightTrianglesCount); OFF(SharedData, time); END(); printf(",\n");
        SZ(Ray); OFF(Ray, origin); OFF(Ray, minDistance); OFF(Ray, direction); OFF(Ray, maxDistance); OFF(Ray, throughput);
        OFF(Ray, radiance); OFF(Ray, params); END(); printf(",\n");
        SZ(LightSamplingRay); OFF(LightSamplingRay, origin); OFF(LightSamplingRay, minDistance); OFF(LightSamplingRay, direction);
        OFF(LightSamplingRay, maxDistance); OFF(LightSamplingRay, throughput); OFF(LightSamplingRay, targetIndex); END(); printf(",\n");
        SZ(Intersection); OFF(Intersection, distance); OFF(Intersection, triangleIndex); OFF(Intersection, coordinates); END(); printf(",\n");
        SZ(Vertex); OFF(Vertex, v); OFF(Vertex, n); END(); printf(",\n");
        SZ(Material); OFF(Material, diffuse); OFF(Material, emissive); OFF(Material, ior); OFF(Material, materialType); END(); printf(",\n");
        SZ(TriangleReference); OFF(TriangleReference, tri); OFF(TriangleReference, materialIndex); OFF(TriangleReference, lightTriangleIndex); END(); printf(",\n");
        SZ(LightTriangle); OFF(LightTriangle, emissive); OFF(LightTriangle, v1); OFF(LightTriangle, v2); OFF(LightTriangle, v3);
        OFF(LightTriangle, area); OFF(LightTriangle, pdf); OFF(LightTriangle, cdf); OFF(LightTriangle, index); END(); printf("\n  },\n");
        printf("  \"macros\": {\"DISTANCE_EPSILON\": %.9g, \"ANGLE_EPSILON\": %.9g, \"PI\": %.9g, \"NOISE_DIMENSIONS\": %d, "
               "\"MAX_PATH_LENGTH\": %d, \"SPECTRUM_SAMPLES\": %d, \"ACCUMULATE_IMAGE\": %d, \"ANIMATE_NOISE\": %d, "
               "\"MATERIAL_DIFFUSE\": %u, \"MATERIAL_MIRROR\": %u, \"MATERIAL_SMOOTH_PLASTIC\": %u, \"MATERIAL_SMOOTH_DIELECTRIC\": %u}\n}\n",
               (double)(float)DISTANCE_EPSILON, (double)(float)ANGLE_EPSILON, (double)(float)PI, NOISE_DIMENSIONS, MAX_PATH_LENGTH,
               SPECTRUM_SAMPLES, (int)ACCUMULATE_IMAGE, ANIMATE_NOISE, MATERIAL_DIFFUSE, MATERIAL_MIRROR, MATERIAL_SMOOTH_PLASTIC,
               MATERIAL_SMOOTH_DIELECTRIC);
        return 0;
    }
    // host-callable functions of Raytracing.h:125-178, evaluated by the reference's own code
    const float xs[] = {0.0f, 1e-4f, 0.0031308f, 0.04045f, 0.25f, 0.5f, 0.999f, 1.0f, 2.5f, 17.0f};
    const int nx = sizeof(xs) / sizeof(xs[0]);
    printf("{\n  \"source\": \"renderer/Raytracing.h host functions run from the reference header\",\n");
    printf("  \"balanceHeuristic\": [");
    for (int i = 0; i < nx; ++i) for (int j = 0; j < nx; ++j)
        printf("%s[%.9g, %.9g, %s]", (i || j) ? ", " : "", xs[i], xs[j], J(balanceHeuristic(xs[i], xs[j])));
    printf("],\n  \"triangleSamplePDF\": [");
    for (int i = 1; i < nx; ++i) for (int j = 1; j < nx; ++j)
        printf("%s[%.9g, %.9g, %.9g, %s]", (i > 1 || j > 1) ? ", " : "", xs[i], xs[j], xs[(i + j) % nx],
               J(triangleSamplePDF(xs[i], xs[j], xs[(i + j) % nx])));
    printf("],\n  \"toSRGB\": [");
    for (int i = 0; i < nx; ++i) printf("%s[%.9g, %.9g]", i ? ", " : "", xs[i], (double)toSRGB(xs[i]));
    printf("],\n  \"toLinear\": [");
    for (int i = 0; i < nx; ++i) printf("%s[%.9g, %.9g]", i ? ", " : "", xs[i], (double)toLinear(xs[i]));
    printf("],\n  \"halton\": [");
    for (unsigned i = 0; i < 12; ++i) for (unsigned b = 2; b <= 5; ++b)
        printf("%s[%u, %u, %.9g]", (i || b > 2) ? ", " : "", i, b, (double)halton(i, b));
    printf("],\n  \"vanDerCorput\": [");
    for (unsigned i = 0; i < 12; ++i) for (unsigned b = 2; b <= 3; ++b)
        printf("%s[%u, %u, %.9g]", (i || b > 2) ? ", " : "", i, b, (double)vanDerCorput(i, b));
    printf("]\n}\n");
    return 0;
}
