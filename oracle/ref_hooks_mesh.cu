// ref_hooks_mesh.cu — compiles the reference's src/mesh.cu as it lies (tinygltf + stb implementations included) and
// exposes what Scene::readFromGLTF (src/mesh.cu:80-307) parsed, so the oracle's loader restatement can be pinned
// against the reference's own loader in this container (host only: copySceneToDevice needs a GPU and is not called).
#include <fstream>
#include <iostream>
#include <map>
#include <memory>
#include <sstream>
#include <string>
#include <vector>
#include <thrust/device_vector.h>
#include "math/vec.h"
#include "material.h"
#include "camera.h"
#include "delta_light.h"
#define private public      // Scene keeps m_meshes / m_materials / m_lights private (include/mesh.cuh:106-112)
#include "mesh.cuh"
#undef private
#include "mesh.cu"

extern "C" {
void* ref_scene_load(const char* path) {
    try {
        return new Scene(path, "gltf");
    } catch (const std::exception& e) {
        std::cerr << "reference loader threw: " << e.what() << std::endl;
        return nullptr;
    }
}
void ref_scene_free(void* s) { delete static_cast<Scene*>(s); }
// counts: meshes, materials, lights
void ref_scene_counts(void* s_, int* out3) {
    Scene* s = static_cast<Scene*>(s_);
    out3[0] = (int)s->m_meshes.size(); out3[1] = (int)s->m_materials.size(); out3[2] = (int)s->m_lights.size();
}
void ref_scene_camera(void* s_, float* out19) {
    Scene* s = static_cast<Scene*>(s_);
    out19[0] = s->m_camera.getVFov(); out19[1] = s->m_camera.getAspRatio(); out19[2] = s->m_camera.getNearPlane();
    Mat4 m = s->m_camera.m_transform->localToWorld();
    memcpy(out19 + 3, &m, 64);
}
void ref_mesh_sizes(void* s_, int i, int* out2) {
    Mesh& m = static_cast<Scene*>(s_)->m_meshes[i];
    out2[0] = (int)m.m_polygons.size(); out2[1] = (int)m.m_vertexCoords.size();
}
// indices (u32), vertices, normals (xyz), local-to-world (16 floats), material name
void ref_mesh_data(void* s_, int i, uint32_t* idx, float* verts, float* norms, float* l2w16, char* name, int nameCap) {
    Mesh& m = static_cast<Scene*>(s_)->m_meshes[i];
    memcpy(idx, m.m_polygons.data(), m.m_polygons.size() * 4);
    memcpy(verts, m.m_vertexCoords.data(), m.m_vertexCoords.size() * 12);
    memcpy(norms, m.m_vertexNormals.data(), m.m_vertexNormals.size() * 12);
    Mat4 t = m.m_transform->localToWorld();
    memcpy(l2w16, &t, 64);
    snprintf(name, nameCap, "%s", m.m_material.c_str());
}
// materials / lights in std::map order (the order copySceneToDevice uploads them, src/mesh.cu:326-333,389-394)
void ref_material(void* s_, int i, void* out60, char* name, int nameCap) {
    static_assert(sizeof(Material) == 60, "Material layout");
    auto it = static_cast<Scene*>(s_)->m_materials.begin();
    std::advance(it, i);
    memcpy(out60, &it->second, 60);
    snprintf(name, nameCap, "%s", it->first.c_str());
}
void ref_light(void* s_, int i, void* out52, char* name, int nameCap) {
    auto it = static_cast<Scene*>(s_)->m_lights.begin();
    std::advance(it, i);
    memcpy(out52, &it->second, 52);
    snprintf(name, nameCap, "%s", it->first.c_str());
}
}   // extern "C"
