// obj_loader.cpp — Wavefront OBJ/MTL -> ABI buffers (host only).
//
// The reference loads Media/<scene>.obj through Apple SceneKit/ModelIO (renderer/Renderer.mm:265-268), which is
// closed source, so triangulation and vertex order are UNPINNED; this loader documents its own choices:
//   * a Vertex is a unique (position index, normal index) pair, in order of first use; texcoords are dropped
//     (Renderer.mm:365-369 drops them too);
//   * polygons are fan-triangulated (v0, v[i], v[i+1]) (the reference's water scenes need none: their 7,088 faces are
//     triangles already);
//   * faces without normals get the geometric face normal;
//   * triangles are emitted grouped by material in order of first `usemtl` (SceneKit makes one geometry element
//     per material, Renderer.mm:375-377), so triangle IDs are defined by THIS loader for oracle and GPU alike;
//   * MTL decode follows Renderer.mm:278-329: Kd = diffuse, Ka = emission, Ks = (roughness, metalness, +-ior);
//     no colour clamping.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <map>
#include <sstream>
#include "scene.h"

namespace mrt {

struct MtlEntry { float kd[3] = {1, 1, 1}, ka[3] = {0, 0, 0}, ks[3] = {1, 0, 0}; };

static std::string dirname_of(const std::string& p) {
    size_t k = p.find_last_of("/\\");
    return k == std::string::npos ? std::string() : p.substr(0, k + 1);
}

static void load_mtl(const std::string& path, std::map<std::string, MtlEntry>& out, std::vector<std::string>& order) {
    std::ifstream in(path);
    std::string line, cur;
    while (std::getline(in, line)) {
        std::istringstream ss(line);
        std::string key;
        if (!(ss >> key)) continue;
        if (key == "newmtl") { ss >> cur; if (!out.count(cur)) order.push_back(cur); out[cur] = MtlEntry(); }
        else if (cur.empty()) continue;
        else if (key == "Kd") ss >> out[cur].kd[0] >> out[cur].kd[1] >> out[cur].kd[2];
        else if (key == "Ka") ss >> out[cur].ka[0] >> out[cur].ka[1] >> out[cur].ka[2];
        else if (key == "Ks") ss >> out[cur].ks[0] >> out[cur].ks[1] >> out[cur].ks[2];
    }
}

// Renderer.mm:292-320
static mrt_material decode_material(const MtlEntry& e) {
    mrt_material m{};
    for (int i = 0; i < 3; ++i) { m.diffuse.values[i] = e.kd[i]; m.emissive.values[i] = e.ka[i]; }
    m.ior = e.ks[2];
    float roughness = e.ks[0], metallness = e.ks[1];
    m.materialType = MRT_MATERIAL_DIFFUSE;
    if (metallness > 0.0f) {
        if (roughness == 0.0f) m.materialType = MRT_MATERIAL_MIRROR;          // rough conductor: left diffuse
    } else if (roughness == 1.0f) {
        m.materialType = MRT_MATERIAL_DIFFUSE;
    } else if (m.ior <= 0.0f) {
        m.ior = std::fabs(m.ior);
        m.materialType = (roughness == 0.0f) ? MRT_MATERIAL_SMOOTH_PLASTIC : MRT_MATERIAL_DIFFUSE;
    } else {
        m.materialType = (roughness == 0.0f) ? MRT_MATERIAL_SMOOTH_DIELECTRIC : MRT_MATERIAL_DIFFUSE;
    }
    return m;
}

bool load_obj(const std::string& path, mrt_scene& s, std::string& err) {
    std::ifstream in(path);
    if (!in) { err = "cannot open '" + path + "'"; return false; }
    s = mrt_scene();
    s.name = path;
    std::vector<float> pos, nrm;
    std::map<std::string, MtlEntry> mtl;
    std::vector<std::string> mtlOrder;
    std::map<std::string, uint32_t> matIndex;                     // material name -> index (first use order)
    std::vector<std::vector<uint32_t>> groups;                    // per material: vertex ids, 3 per triangle
    std::map<std::pair<int, int>, uint32_t> unique;               // (pos, nrm) -> vertex id
    uint32_t curMat = 0xFFFFFFFFu;
    auto use_material = [&](const std::string& name) {
        auto it = matIndex.find(name);
        if (it == matIndex.end()) {
            it = matIndex.emplace(name, uint32_t(s.materials.size())).first;
            s.materials.push_back(decode_material(mtl.count(name) ? mtl[name] : MtlEntry()));
            groups.emplace_back();
        }
        curMat = it->second;
    };
    std::string line;
    while (std::getline(in, line)) {
        std::istringstream ss(line);
        std::string key;
        if (!(ss >> key)) continue;
        if (key == "v") { float x, y, z; ss >> x >> y >> z; pos.insert(pos.end(), {x, y, z}); }
        else if (key == "vn") { float x, y, z; ss >> x >> y >> z; nrm.insert(nrm.end(), {x, y, z}); }
        else if (key == "mtllib") { std::string f; ss >> f; load_mtl(dirname_of(path) + f, mtl, mtlOrder); }
        else if (key == "usemtl") { std::string n; ss >> n; use_material(n); }
        else if (key == "f") {
            if (curMat == 0xFFFFFFFFu) use_material("__default__");
            std::vector<std::pair<int, int>> corners;
            std::string tok;
            while (ss >> tok) {
                int vi = 0, ni = 0;
                const char* p = tok.c_str();
                vi = std::atoi(p);
                const char* s1 = std::strchr(p, '/');
                if (s1) { const char* s2 = std::strchr(s1 + 1, '/'); if (s2 && s2[1]) ni = std::atoi(s2 + 1); }
                if (vi < 0) vi = int(pos.size() / 3) + vi + 1;
                if (ni < 0) ni = int(nrm.size() / 3) + ni + 1;
                if (vi <= 0 || size_t(vi) > pos.size() / 3) { err = "bad vertex index in '" + line + "'"; return false; }
                if (ni < 0 || size_t(ni) > nrm.size() / 3) { err = "bad normal index in '" + line + "'"; return false; }
                corners.emplace_back(vi, ni);
            }
            if (corners.size() < 3) continue;
            float fn[3] = {0, 0, 0};
            bool needFace = false;
            for (auto& c : corners) needFace |= (c.second == 0);
            if (needFace) {
                const float* a = &pos[3 * (corners[0].first - 1)];
                const float* b = &pos[3 * (corners[1].first - 1)];
                const float* c = &pos[3 * (corners[2].first - 1)];
                float ux = b[0] - a[0], uy = b[1] - a[1], uz = b[2] - a[2], vx = c[0] - a[0], vy = c[1] - a[1], vz = c[2] - a[2];
                fn[0] = uy * vz - uz * vy; fn[1] = uz * vx - ux * vz; fn[2] = ux * vy - uy * vx;
                float l = std::sqrt(fn[0] * fn[0] + fn[1] * fn[1] + fn[2] * fn[2]);
                if (l > 0) { fn[0] /= l; fn[1] /= l; fn[2] /= l; }
            }
            std::vector<uint32_t> ids;
            for (auto& c : corners) {
                const float* p = &pos[3 * (c.first - 1)];
                if (c.second == 0) { ids.push_back(add_vertex(s, p[0], p[1], p[2], fn[0], fn[1], fn[2])); continue; }
                auto it = unique.find(c);
                if (it == unique.end()) {
                    const float* n = &nrm[3 * (c.second - 1)];
                    it = unique.emplace(c, add_vertex(s, p[0], p[1], p[2], n[0], n[1], n[2])).first;
                }
                ids.push_back(it->second);
            }
            for (size_t k = 1; k + 1 < ids.size(); ++k) { auto& g = groups[curMat]; g.push_back(ids[0]); g.push_back(ids[k]); g.push_back(ids[k + 1]); }
        }
    }
    for (uint32_t m = 0; m < groups.size(); ++m)
        for (size_t k = 0; k < groups[m].size(); k += 3) add_triangle(s, groups[m][k], groups[m][k + 1], groups[m][k + 2], m);
    if (s.references.empty()) { err = "no faces in '" + path + "'"; return false; }
    finalize_lights(s);
    return true;
}

}  // namespace mrt
