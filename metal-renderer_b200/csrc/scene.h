// scene.h — host-side scene container in the reference's buffer ABI (include/mrt_abi.h).
// Stands in for what -initRaytracing assembles before upload (reference: renderer/Renderer.mm:257-448).
#pragma once
#include <cstdint>
#include <string>
#include <vector>
#include "mrt_abi.h"

struct mrt_scene {
    std::vector<mrt_vertex> vertices;
    std::vector<uint32_t> indices;                    // 3 per triangle (Renderer.mm:421-423)
    std::vector<mrt_triangle_reference> references;   // 1 per triangle (Renderer.mm:415-420)
    std::vector<mrt_material> materials;
    std::vector<mrt_light_triangle> lights;           // light_count + 1 sentinel
    uint32_t light_count = 0;
    std::string name;
};

namespace mrt {

// Adds one triangle with material `mat`; lights are derived later by finalize_lights().
void add_triangle(mrt_scene& s, uint32_t a, uint32_t b, uint32_t c, uint32_t mat);
uint32_t add_vertex(mrt_scene& s, float x, float y, float z, float nx, float ny, float nz);
uint32_t add_material(mrt_scene& s, float r, float g, float b, float er, float eg, float eb, float ior, uint32_t type);
// Light list, pdf/cdf and sentinel exactly as Renderer.mm:373-448 builds them.
void finalize_lights(mrt_scene& s);

bool make_procedural(const std::string& name, uint64_t seed, mrt_scene& out, std::string& err);
bool load_obj(const std::string& path, mrt_scene& out, std::string& err);
void noise_fill(uint64_t seed, uint32_t frame, float* tile);

}  // namespace mrt
