// host.cpp — host side of the drop-in boundary: the reference's classes (include/tpt/host.hpp) over the C ABI.
//
// Pure host C++ (no CUDA headers): glTF loading, scene flattening, image I/O and the frame driver; every device
// operation is a call into libtpt_b200.so.  Compile with -ffp-contract=off: the transforms below are evaluated one
// rounding per operation in the reference's order, so that the uploaded matrices equal the reference's bit for bit.
#include "tpt/host.hpp"

#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <sstream>

#include "tpt/json.hpp"

namespace tpt {

// ---- errors -------------------------------------------------------------------------------------------------
void throwLastError(const char* where) { throw std::runtime_error(std::string(where) + ": " + tpt_last_error()); }

// ---- math (include/math/mat.h:37-51,169-176,282-298; include/math/quat.h:52-69) ----------------------------------
Mat4 operator*(const Mat4& a, const Mat4& b) {
    Mat4 r;
    for (int i = 0; i < 16; ++i) r.m[i] = 0.0f;
    for (int row = 0; row < 4; ++row)
        for (int col = 0; col < 4; ++col)
            for (int k = 0; k < 4; ++k) r.m[col * 4 + row] += a.m[k * 4 + row] * b.m[col * 4 + k];
    return r;
}

Mat4 Transform::localToParent() const {
    if (m_hasMatrix) return m_matrix;
    const Real x = m_rotation.x, y = m_rotation.y, z = m_rotation.z, w = m_rotation.w;
    const Real x2 = x * x, y2 = y * y, z2 = z * z, xy = x * y, xz = x * z, yz = y * z, wx = w * x, wy = w * y, wz = w * z;
    Mat4 R, T, S;
    const Real r[16] = {1.0f - 2.0f * (y2 + z2), 2.0f * (xy + wz), 2.0f * (xz - wy), 0.0f,
                        2.0f * (xy - wz), 1.0f - 2.0f * (x2 + z2), 2.0f * (yz + wx), 0.0f,
                        2.0f * (xz + wy), 2.0f * (yz - wx), 1.0f - 2.0f * (x2 + y2), 0.0f,
                        0.0f, 0.0f, 0.0f, 1.0f};
    std::memcpy(R.m, r, sizeof r);
    T.at(3, 0) = m_location.x; T.at(3, 1) = m_location.y; T.at(3, 2) = m_location.z;
    S.at(0, 0) = m_scale.x; S.at(1, 1) = m_scale.y; S.at(2, 2) = m_scale.z;
    return (T * R) * S;
}

// Inverse transpose of the upper-left 3x3 (as a Mat4 with unit w), by cofactors: N = cof(A) / det(A).
// The reference expands a 4x4 cofactor inverse (include/math/mat.h:203-281); for this block-diagonal matrix both
// give the same matrix up to rounding (checked against the reference's values to 1e-6 relative in the tests).
Mat4 normalToWorld(const Mat4& l) {
    const Real a00 = l.at(0, 0), a01 = l.at(1, 0), a02 = l.at(2, 0);   // a[row][col]
    const Real a10 = l.at(0, 1), a11 = l.at(1, 1), a12 = l.at(2, 1);
    const Real a20 = l.at(0, 2), a21 = l.at(1, 2), a22 = l.at(2, 2);
    const Real c00 = a11 * a22 - a12 * a21, c01 = a12 * a20 - a10 * a22, c02 = a10 * a21 - a11 * a20;
    const Real c10 = a02 * a21 - a01 * a22, c11 = a00 * a22 - a02 * a20, c12 = a01 * a20 - a00 * a21;
    const Real c20 = a01 * a12 - a02 * a11, c21 = a02 * a10 - a00 * a12, c22 = a00 * a11 - a01 * a10;
    const Real inv = 1.0f / (a00 * c00 + a01 * c01 + a02 * c02);
    Mat4 n;   // (A^-1)^T = cof(A) / det: element [row][col] = c[row][col] * inv
    n.at(0, 0) = c00 * inv; n.at(1, 0) = c01 * inv; n.at(2, 0) = c02 * inv;
    n.at(0, 1) = c10 * inv; n.at(1, 1) = c11 * inv; n.at(2, 1) = c12 * inv;
    n.at(0, 2) = c20 * inv; n.at(1, 2) = c21 * inv; n.at(2, 2) = c22 * inv;
    return n;
}

static Vec3 mulPoint(const Mat4& m, Vec3 p, Real w) {   // MatrixVectorMultiply, include/math/mat.h:37-51
    Real v[4] = {p.x, p.y, p.z, w}, o[3];
    for (int row = 0; row < 3; ++row) {
        Real r = 0.0f;
        for (int k = 0; k < 4; ++k) r += m.m[k * 4 + row] * v[k];
        o[row] = r;
    }
    return Vec3(o[0], o[1], o[2]);
}

Material defaultMaterial() {   // include/material.h:88-103
    Material m;
    std::memset(&m, 0, sizeof m);
    m.base_color[0] = 0.82f; m.base_color[1] = 0.67f; m.base_color[2] = 0.16f;
    m.unused[1] = 0.5f;   // specular
    m.unused[2] = 0.5f;   // roughness
    m.unused[8] = 1.0f;
    return m;
}

tpt_camera Camera::toABI() const {
    tpt_camera c;
    c.vfov = m_verticalFOV;
    c.aspect = m_aspectRatio;
    c.near_plane = m_nearPlane;
    const Mat4 m = m_transform->localToWorld();
    std::memcpy(c.c2w, m.m, sizeof c.c2w);
    return c;
}

// ---- context -------------------------------------------------------------------------------------------------
Context::Context(int device) { check(tpt_context_create(device, &m_ctx), "tpt_context_create"); }
Context::~Context() { tpt_context_destroy(m_ctx); }
std::vector<std::unique_ptr<Context>> Context::initAll(int nGpus) {
    std::vector<tpt_context*> raw((size_t)std::max(nGpus, 1), nullptr);
    check(tpt_comm_init_all(nGpus, raw.data()), "tpt_comm_init_all");
    std::vector<std::unique_ptr<Context>> out;
    for (tpt_context* c : raw) out.emplace_back(new Context(c));
    return out;
}
void Context::synchronize() { check(tpt_context_synchronize(m_ctx), "tpt_context_synchronize"); }
tpt_counters Context::counters() {
    tpt_counters c;
    check(tpt_get_counters(m_ctx, &c), "tpt_get_counters");
    return c;
}

// ---- glTF loading (semantics of src/mesh.cu:80-307) -----------------------------------------------------------
static std::string readFile(const std::string& path) {
    std::ifstream f(path, std::ios::binary);
    if (!f) throw std::runtime_error("Failed to load .gltf : cannot open " + path);   // src/mesh.cu:92-94
    std::ostringstream ss;
    ss << f.rdbuf();
    return ss.str();
}

static std::vector<uint8_t> base64Decode(const std::string& s, size_t begin) {
    static int8_t lut[256];
    static bool init = false;
    if (!init) {
        std::memset(lut, -1, sizeof lut);
        const char* abc = "ABCDEFGHIJKLMNOPQRSTUVWXYZabcdefghijklmnopqrstuvwxyz0123456789+/";
        for (int i = 0; i < 64; ++i) lut[(uint8_t)abc[i]] = (int8_t)i;
        init = true;
    }
    std::vector<uint8_t> out;
    out.reserve((s.size() - begin) * 3 / 4);
    uint32_t acc = 0;
    int bits = 0;
    for (size_t i = begin; i < s.size(); ++i) {
        const int8_t v = lut[(uint8_t)s[i]];
        if (v < 0) continue;   // '=', whitespace
        acc = (acc << 6) | (uint32_t)v;
        bits += 6;
        if (bits >= 8) {
            bits -= 8;
            out.push_back((uint8_t)(acc >> bits));
        }
    }
    return out;
}

static Transform readTransform(const json::Value& node) {   // src/mesh.cu:103-139: missing rotation -> Quat(), scale -> 1, translation -> 0
    Transform t;
    const json::Value &r = node["rotation"], &s = node["scale"], &l = node["translation"];
    if (r.is_array() && r.size() == 4) t.m_rotation = Quat((Real)r[0].num, (Real)r[1].num, (Real)r[2].num, (Real)r[3].num);
    if (s.is_array() && s.size() == 3) t.m_scale = Vec3((Real)s[0].num, (Real)s[1].num, (Real)s[2].num);
    if (l.is_array() && l.size() == 3) t.m_location = Vec3((Real)l[0].num, (Real)l[1].num, (Real)l[2].num);
    return t;
}

namespace {
struct Gltf {
    json::Value root;
    std::vector<std::vector<uint8_t>> buffers;
    // Element `i` of an accessor lives at base + i * stride.  The reference assumes tightly packed data
    // (src/mesh.cu:165-170); `strided` (gltf-full) honours bufferView.byteStride.
    const uint8_t* data(int accessor, size_t elemBytes, size_t count, bool strided, size_t* stride) const {
        const json::Value& acc = root["accessors"][(size_t)accessor];
        if (!acc.is_object()) throw std::runtime_error("glTF: accessor index out of range");
        const json::Value& bv = root["bufferViews"][(size_t)acc["bufferView"].integer(-1)];
        if (!bv.is_object()) throw std::runtime_error("glTF: bufferView index out of range");
        const size_t b = (size_t)bv["buffer"].integer(0);
        if (b >= buffers.size()) throw std::runtime_error("glTF: buffer index out of range");
        const size_t off = (size_t)acc["byteOffset"].number(0) + (size_t)bv["byteOffset"].number(0);
        const size_t bs = strided ? (size_t)bv["byteStride"].number(0) : 0;
        *stride = bs ? bs : elemBytes;
        if (count && off + (count - 1) * *stride + elemBytes > buffers[b].size()) throw std::runtime_error("glTF: accessor reads past the end of its buffer");
        return buffers[b].data() + off;
    }
    size_t count(int accessor) const { return (size_t)root["accessors"][(size_t)accessor]["count"].number(0); }
};

uint32_t le32(const std::string& s, size_t off) {
    uint32_t v;
    std::memcpy(&v, s.data() + off, 4);
    return v;
}

// .gltf (JSON + external or base64 buffers) or .glb (12-byte header, JSON chunk, optional BIN chunk = buffer without uri)
void loadContainer(const std::string& filename, Gltf& g) {
    const std::string file = readFile(filename);
    std::string glbBin;
    bool hasBin = false;
    if (file.size() >= 12 && file.compare(0, 4, "glTF") == 0) {
        if (le32(file, 4) != 2) throw std::runtime_error("Failed to load .gltf : only GLB version 2 is supported");
        size_t off = 12;
        std::string jsonText;
        while (off + 8 <= file.size()) {
            const uint32_t len = le32(file, off), type = le32(file, off + 4);
            if (off + 8 + len > file.size()) throw std::runtime_error("Failed to load .gltf : truncated GLB chunk");
            if (type == 0x4E4F534Au) jsonText = file.substr(off + 8, len);                    // "JSON"
            else if (type == 0x004E4942u && !hasBin) { glbBin = file.substr(off + 8, len); hasBin = true; }   // "BIN\0"
            off += 8 + (size_t)len;
        }
        if (jsonText.empty()) throw std::runtime_error("Failed to load .gltf : GLB has no JSON chunk");
        g.root = json::parse(jsonText);
    } else {
        g.root = json::parse(file);
    }
    const std::string dir = filename.find_last_of("/\\") == std::string::npos ? "" : filename.substr(0, filename.find_last_of("/\\") + 1);
    for (size_t i = 0; i < g.root["buffers"].size(); ++i) {
        const json::Value& uriV = g.root["buffers"][i]["uri"];
        const std::string uri = uriV.string("");
        if (uri.empty()) {
            if (!hasBin) throw std::runtime_error("glTF: buffer without uri and no GLB BIN chunk");
            g.buffers.emplace_back(glbBin.begin(), glbBin.end());
        } else if (uri.compare(0, 5, "data:") == 0) {
            const size_t comma = uri.find(',');
            g.buffers.push_back(base64Decode(uri, comma == std::string::npos ? 5 : comma + 1));
        } else {
            const std::string raw = readFile(dir + uri);
            g.buffers.emplace_back(raw.begin(), raw.end());
        }
    }
}

Material readMaterial(const json::Value& gm) {   // src/mesh.cu:228-257
    Material m = defaultMaterial();
    const json::Value& pbr = gm["pbrMetallicRoughness"];
    m.unused[2] = (Real)pbr["roughnessFactor"].number(1.0);
    m.metallic = (Real)pbr["metallicFactor"].number(1.0);   // glTF default 1.0 -> a material that omits it is a mirror
    const json::Value& bc = pbr["baseColorFactor"];
    m.base_color[0] = bc.is_array() ? (Real)bc[0].num : 1.0f;
    m.base_color[1] = bc.is_array() ? (Real)bc[1].num : 1.0f;
    m.base_color[2] = bc.is_array() ? (Real)bc[2].num : 1.0f;
    const json::Value& ext = gm["extensions"];   // src/mesh.cu:243-257
    if (ext.has("KHR_materials_transmission")) m.unused[1] = 1.0f - (Real)ext["KHR_materials_transmission"]["transmissionFactor"].number(0) / 5.0f;
    if (ext.has("KHR_materials_emissive_strength")) m.emission = (Real)ext["KHR_materials_emissive_strength"]["emissiveStrength"].number(0);
    if (ext.has("KHR_materials_ior")) m.eta = (Real)ext["KHR_materials_ior"]["ior"].number(0);
    return m;
}

DeltaLight readLight(const json::Value& L, const Mat4& m) {   // src/mesh.cu:267-304
    const float wattsPerLumen = 1.0f / 683.0f;   // include/delta_light.h:7
    DeltaLight d;
    std::memset(&d, 0, sizeof d);
    const json::Value& col = L["color"];
    for (int k = 0; k < 3; ++k) d.color[k] = col.is_array() ? (Real)col[(size_t)k].num : 1.0f;
    const Real inten = (Real)L["intensity"].number(1.0);
    const std::string type = L["type"].string("");
    const Vec3 pos = mulPoint(m, Vec3(0, 0, 0), 1.0f), dirv = mulPoint(m, Vec3(0, 0, -1), 0.0f);
    if (type == "point") {
        d.type = 0;
        d.intensity = inten * wattsPerLumen;       // src/mesh.cu:276
        d.a[0] = pos.x; d.a[1] = pos.y; d.a[2] = pos.z;
    } else if (type == "directional") {
        d.type = 1;
        d.intensity = inten;                       // src/mesh.cu:283 (not scaled)
        d.a[0] = dirv.x; d.a[1] = dirv.y; d.a[2] = dirv.z;
    } else if (type == "spot") {
        d.type = 2;
        d.intensity = inten * wattsPerLumen;       // src/mesh.cu:290
        const Real inner = (Real)L["spot"]["innerConeAngle"].number(0.0);
        const Real outer = (Real)L["spot"]["outerConeAngle"].number(0.7853981633974483);
        const Real co = std::cos(outer), ci = std::cos(inner);
        d.cos_outer = co;
        d.inv_cone_diff = 1.0f / (ci - co);
        d.a[0] = pos.x; d.a[1] = pos.y; d.a[2] = pos.z;
        d.b[0] = dirv.x; d.b[1] = dirv.y; d.b[2] = dirv.z;
    } else {
        throw std::runtime_error("Unsupported light type");   // src/mesh.cu:303
    }
    return d;
}

// One primitive -> Mesh arrays.  `full` == false is the reference's reading (tight accessors, POSITION + NORMAL +
// indices required).  Returns false for a primitive that is not a triangle list (gltf-full skips those).
bool readPrimitive(const Gltf& g, const json::Value& prim, bool full, Mesh& mesh) {
    if (full && prim["mode"].integer(4) != 4) return false;
    const json::Value& attrs = prim["attributes"];
    const int aPos = attrs["POSITION"].integer(-1), aNrm = attrs["NORMAL"].integer(-1), aTex = attrs["TEXCOORD_0"].integer(-1);
    const int aIdx = prim["indices"].integer(-1);
    if (!full && (aPos < 0 || aNrm < 0 || aIdx < 0)) throw std::runtime_error("glTF: a mesh primitive needs POSITION, NORMAL and indices");
    if (aPos < 0) throw std::runtime_error("glTF: a mesh primitive needs POSITION");
    if (full && g.root["accessors"][(size_t)aPos]["componentType"].integer(5126) != 5126) throw std::runtime_error("glTF: POSITION must be float");
    const size_t nv = g.count(aPos);
    size_t sp = 12, sn = 12, st = 8;
    const uint8_t* pos = g.data(aPos, 12, nv, full, &sp);   // buffers are byte arrays: no alignment assumed
    mesh.m_vertexCoords.resize(nv);
    mesh.m_vertexNormals.assign(nv, Vec3(0, 0, 0));
    mesh.m_vertexTexCoords.resize(nv);
    for (size_t i = 0; i < nv; ++i) {
        Real p3[3];
        std::memcpy(p3, pos + sp * i, 12);
        mesh.m_vertexCoords[i] = Vec3(p3[0], p3[1], p3[2]);
    }
    if (aNrm >= 0) {
        const uint8_t* nrm = g.data(aNrm, 12, nv, full, &sn);
        for (size_t i = 0; i < nv; ++i) {
            Real n3[3];
            std::memcpy(n3, nrm + sn * i, 12);
            mesh.m_vertexNormals[i] = Vec3(n3[0], n3[1], n3[2]);
        }
    }
    if (aTex >= 0 && (!full || g.root["accessors"][(size_t)aTex]["componentType"].integer(5126) == 5126)) {
        const uint8_t* tex = g.data(aTex, 8, nv, full, &st);
        for (size_t i = 0; i < nv; ++i) std::memcpy(&mesh.m_vertexTexCoords[i], tex + st * i, 8);
    }
    if (aIdx >= 0) {
        const size_t ni = g.count(aIdx);
        const int ctype = g.root["accessors"][(size_t)aIdx]["componentType"].integer(5125);
        const size_t isz = ctype == 5121 ? 1 : (ctype == 5123 ? 2 : 4);
        size_t si = isz;
        const uint8_t* idx = g.data(aIdx, isz, ni, full, &si);
        mesh.m_polygons.resize(ni);
        for (size_t i = 0; i < ni; ++i) {
            uint32_t v = 0;
            std::memcpy(&v, idx + si * i, isz);
            if (full && v >= nv) throw std::runtime_error("glTF: index out of range");
            mesh.m_polygons[i] = v;
        }
    } else {   // non-indexed triangle list
        mesh.m_polygons.resize(nv - nv % 3);
        for (size_t i = 0; i < mesh.m_polygons.size(); ++i) mesh.m_polygons[i] = (uint32_t)i;
    }
    if (aNrm < 0) {   // area-weighted vertex normals from the faces
        for (size_t f = 0; f + 2 < mesh.m_polygons.size(); f += 3) {
            const Vec3 &a = mesh.m_vertexCoords[mesh.m_polygons[f]], &b = mesh.m_vertexCoords[mesh.m_polygons[f + 1]], &c = mesh.m_vertexCoords[mesh.m_polygons[f + 2]];
            const Vec3 e1(b.x - a.x, b.y - a.y, b.z - a.z), e2(c.x - a.x, c.y - a.y, c.z - a.z);
            const Vec3 n(e1.y * e2.z - e1.z * e2.y, e1.z * e2.x - e1.x * e2.z, e1.x * e2.y - e1.y * e2.x);
            for (int k = 0; k < 3; ++k) {
                Vec3& dst = mesh.m_vertexNormals[mesh.m_polygons[f + k]];
                dst.x += n.x; dst.y += n.y; dst.z += n.z;
            }
        }
        for (Vec3& n : mesh.m_vertexNormals) {
            const Real len = std::sqrt(n.x * n.x + n.y * n.y + n.z * n.z);
            if (len > 0) { n.x /= len; n.y /= len; n.z /= len; } else n = Vec3(0, 1, 0);
        }
    }
    return true;
}

std::string indexedKey(int idx, const std::string& name) {   // sorts by index, keeps the name readable
    char buf[16];
    std::snprintf(buf, sizeof buf, "%06d:", idx);
    return buf + name;
}
}   // namespace

Scene::Scene(const std::string& filename, const std::string& type) {
    if (type == "gltf") readFromGLTF(filename, false);
    else if (type == "gltf-full") readFromGLTF(filename, true);
    else throw std::runtime_error("Unsupported file type!");   // src/mesh.cu:76
}

void Scene::readFromGLTF(const std::string& filename, bool full) {
    Gltf g;
    loadContainer(filename, g);
    const json::Value& glMaterials = g.root["materials"];
    const json::Value& glLights = g.root["extensions"]["KHR_lights_punctual"]["lights"];
    const json::Value& nodes = g.root["nodes"];

    if (!full) {   // the reference's loop over the flat node list (src/mesh.cu:97-306)
        for (size_t ni = 0; ni < nodes.size(); ++ni) {
            const json::Value& node = nodes[ni];
            const int lightIdx = node["extensions"]["KHR_lights_punctual"]["light"].integer(-1);
            if (node["camera"].integer(-1) > -1) {
                const json::Value& cam = g.root["cameras"][(size_t)node["camera"].integer(0)];
                if (cam["type"].string("") == "perspective") {   // src/mesh.cu:146-156
                    const json::Value& p = cam["perspective"];
                    m_camera = Camera((Real)p["yfov"].number(0), (Real)p["aspectRatio"].number(0), (Real)p["znear"].number(0), readTransform(node));
                }
            } else if (node["mesh"].integer(-1) > -1) {
                const json::Value& prim = g.root["meshes"][(size_t)node["mesh"].integer(0)]["primitives"][0];   // primitives[0] only, src/mesh.cu:165
                Mesh mesh;
                readPrimitive(g, prim, false, mesh);
                if (glMaterials.size() > 0) {
                    const json::Value& gm = glMaterials[(size_t)prim["material"].integer(0)];
                    mesh.m_material = gm["name"].string("");
                    if (!m_materials.count(mesh.m_material)) m_materials[mesh.m_material] = readMaterial(gm);   // first definition wins, src/mesh.cu:228
                }
                mesh.m_transform = std::make_shared<Transform>(readTransform(node));
                m_meshes.push_back(std::move(mesh));
            } else if (lightIdx > -1) {
                m_lights[node["name"].string("")] = readLight(glLights[(size_t)lightIdx], readTransform(node).localToWorld());
            }
        }
    } else {   // scene graph: roots of scenes[scene], parents before children
        std::vector<int> roots;
        const json::Value& sceneNodes = g.root["scenes"][(size_t)g.root["scene"].integer(0)]["nodes"];
        if (sceneNodes.is_array()) {
            for (size_t i = 0; i < sceneNodes.size(); ++i) roots.push_back(sceneNodes[i].integer(-1));
        } else {   // no scenes: every node that is nobody's child
            std::vector<char> isChild(nodes.size(), 0);
            for (size_t i = 0; i < nodes.size(); ++i)
                for (size_t c = 0; c < nodes[i]["children"].size(); ++c) {
                    const int ch = nodes[i]["children"][c].integer(-1);
                    if (ch >= 0 && (size_t)ch < nodes.size()) isChild[(size_t)ch] = 1;
                }
            for (size_t i = 0; i < nodes.size(); ++i) if (!isChild[i]) roots.push_back((int)i);
        }
        bool haveCamera = false;
        struct Item { int node; std::shared_ptr<Transform> parent; int depth; };
        std::vector<Item> stack;
        for (auto it = roots.rbegin(); it != roots.rend(); ++it) stack.push_back({*it, nullptr, 0});
        while (!stack.empty()) {
            const Item item = stack.back();
            stack.pop_back();
            if (item.node < 0 || (size_t)item.node >= nodes.size()) throw std::runtime_error("glTF: node index out of range");
            if (item.depth > 256) throw std::runtime_error("glTF: node hierarchy deeper than 256 (cycle?)");
            const json::Value& node = nodes[(size_t)item.node];
            auto t = std::make_shared<Transform>(readTransform(node));
            t->m_parent = item.parent;
            const json::Value& mat = node["matrix"];
            if (mat.is_array() && mat.size() == 16) {
                t->m_hasMatrix = true;
                for (size_t k = 0; k < 16; ++k) t->m_matrix.m[k] = (Real)mat[k].num;
            }
            if (node["camera"].integer(-1) > -1 && !haveCamera) {
                const json::Value& cam = g.root["cameras"][(size_t)node["camera"].integer(0)];
                if (cam["type"].string("") == "perspective") {
                    const json::Value& p = cam["perspective"];
                    m_camera = Camera((Real)p["yfov"].number(0), (Real)p["aspectRatio"].number(1.77778), (Real)p["znear"].number(0.1), Transform());
                    m_camera.m_transform = t;
                    haveCamera = true;
                }
            }
            if (node["mesh"].integer(-1) > -1) {
                const json::Value& prims = g.root["meshes"][(size_t)node["mesh"].integer(0)]["primitives"];
                for (size_t pi = 0; pi < prims.size(); ++pi) {
                    Mesh mesh;
                    if (!readPrimitive(g, prims[pi], true, mesh) || mesh.m_polygons.empty()) continue;
                    const int mi = prims[pi]["material"].integer(-1);
                    if (mi >= 0 && (size_t)mi < glMaterials.size()) {
                        mesh.m_material = indexedKey(mi, glMaterials[(size_t)mi]["name"].string(""));
                        if (!m_materials.count(mesh.m_material)) m_materials[mesh.m_material] = readMaterial(glMaterials[(size_t)mi]);
                    } else {
                        mesh.m_material = "";   // sorts first: the default material
                        if (!m_materials.count("")) m_materials[""] = defaultMaterial();
                    }
                    mesh.m_transform = t;
                    m_meshes.push_back(std::move(mesh));
                }
            }
            const int lightIdx = node["extensions"]["KHR_lights_punctual"]["light"].integer(-1);
            if (lightIdx > -1) m_lights[indexedKey(item.node, node["name"].string(""))] = readLight(glLights[(size_t)lightIdx], t->localToWorld());
            const json::Value& ch = node["children"];
            for (size_t c = ch.size(); c-- > 0;) stack.push_back({ch[c].integer(-1), t, item.depth + 1});
        }
    }
    // declared deviation: a file without a `materials` array makes the reference read materials[0] of an empty vector
    // (src/mesh.cu:225,321,358); here it gets one default-constructed Material
    if (m_materials.empty()) m_materials[""] = defaultMaterial();
}

DeviceScene Scene::flatten() const {
    DeviceScene d;
    std::map<std::string, int> mtlIndex;
    for (const auto& kv : m_materials) {   // std::map order = byte-wise name order (src/mesh.cu:326-333)
        mtlIndex[kv.first] = (int)d.materials.size();
        d.materials.push_back(kv.second);
    }
    size_t nIdx = 0, nVtx = 0;
    for (const Mesh& mesh : m_meshes) {
        for (uint32_t i : mesh.m_polygons) d.indices.push_back(i + (uint32_t)nVtx);   // rebased, src/mesh.cu:340-346
        d.vertices.insert(d.vertices.end(), mesh.m_vertexCoords.begin(), mesh.m_vertexCoords.end());
        d.normals.insert(d.normals.end(), mesh.m_vertexNormals.begin(), mesh.m_vertexNormals.end());
        d.texCoords.insert(d.texCoords.end(), mesh.m_vertexTexCoords.begin(), mesh.m_vertexTexCoords.end());
        MtlInterval iv;
        iv.begin = (int32_t)(nIdx / 3);
        auto it = mtlIndex.find(mesh.m_material);
        iv.mtl_idx = it == mtlIndex.end() ? 0 : it->second;
        d.materialsLUT.push_back(iv);
        const Mat4 l2w = mesh.m_transform->localToWorld();
        d.vertTrans.push_back(l2w);
        d.normalTrans.push_back(normalToWorld(l2w));
        nIdx += mesh.m_polygons.size();
        nVtx += mesh.m_vertexCoords.size();
    }
    for (const auto& kv : m_lights) d.lights.push_back(kv.second);
    return d;
}

DeviceScene Scene::copySceneToDevice(Context& ctx) const {
    DeviceScene d = flatten();
    static_assert(sizeof(Vec3) == 12 && sizeof(Mat4) == 64 && sizeof(Material) == 60 && sizeof(DeltaLight) == 52, "upload layouts");
    tpt_scene_desc desc;
    std::memset(&desc, 0, sizeof desc);
    desc.indices = d.indices.data();
    desc.n_indices = d.indices.size();
    desc.vertices = reinterpret_cast<const float*>(d.vertices.data());
    desc.normals = reinterpret_cast<const float*>(d.normals.data());
    desc.n_vertices = d.vertices.size();
    desc.materials = d.materials.data();
    desc.n_materials = d.materials.size();
    desc.objects = d.materialsLUT.data();
    desc.n_objects = d.materialsLUT.size();
    desc.vert_trans = reinterpret_cast<const float*>(d.vertTrans.data());
    desc.normal_trans = reinterpret_cast<const float*>(d.normalTrans.data());
    desc.lights = d.lights.empty() ? nullptr : d.lights.data();
    desc.n_lights = d.lights.size();
    tpt_scene* s = nullptr;
    check(tpt_scene_create(ctx.handle(), &desc, &s), "tpt_scene_create");
    d.handle = std::shared_ptr<tpt_scene>(s, [](tpt_scene* p) { tpt_scene_destroy(p); });
    return d;
}

void DeviceScene::uploadTransforms() {
    if (!handle) throw std::runtime_error("DeviceScene::uploadTransforms: scene is not on a device");
    if (vertTrans.size() != materialsLUT.size() || normalTrans.size() != materialsLUT.size())
        throw std::runtime_error("DeviceScene::uploadTransforms: one matrix pair per object expected");
    check(tpt_scene_set_transforms(handle.get(), reinterpret_cast<const float*>(vertTrans.data()),
                                   reinterpret_cast<const float*>(normalTrans.data())),
          "tpt_scene_set_transforms");
}

// ---- BVH -------------------------------------------------------------------------------------------------------
void BVH::construct() { check(tpt_bvh_build(m_scene.get(), nullptr), "tpt_bvh_build"); }
void BVH::constructFromWorld(const float* dVerts, size_t nVertices, const uint32_t* dIndices, size_t nFaces, Context& ctx) {
    if (m_nFaces && nFaces != m_nFaces) throw std::runtime_error("BVH::construct: the index array does not hold the face count the BVH was made for");
    tpt_scene* s = m_scene.get();
    const bool fresh = s == nullptr;
    check(tpt_bvh_build_from_world(ctx.handle(), dVerts, nVertices, dIndices, nFaces, &s, nullptr), "tpt_bvh_build_from_world");
    if (fresh) m_scene = std::shared_ptr<tpt_scene>(s, [](tpt_scene* p) { tpt_scene_destroy(p); });
}
Context& Context::defaultContext() {
    static Context ctx(0);
    return ctx;
}
tpt_bvh_stats BVH::constructWithStats() {
    tpt_bvh_stats st;
    check(tpt_bvh_build(m_scene.get(), &st), "tpt_bvh_build");
    return st;
}
void BVH::refit() { check(tpt_bvh_refit(m_scene.get(), nullptr), "tpt_bvh_refit"); }
tpt_bvh_stats BVH::refitWithStats() {
    tpt_bvh_stats st;
    check(tpt_bvh_refit(m_scene.get(), &st), "tpt_bvh_refit");
    return st;
}
std::vector<BVHNode36> BVH::exportReferenceLayout(size_t nFaces) {
    static_assert(sizeof(BVHNode36) == 36, "reference BVHNode is 36 bytes");
    std::vector<BVHNode36> nodes(2 * nFaces - 1);
    check(tpt_bvh_export_reference_nodes(m_scene.get(), nodes.data(), nodes.size()), "tpt_bvh_export_reference_nodes");
    return nodes;
}
void BVH::traceRays(const tpt_ray* rays, size_t n, tpt_hit* hits, bool anyHit) {
    check(tpt_trace_rays(m_scene.get(), rays, n, hits, anyHit ? 1 : 0, 0), "tpt_trace_rays");
}

}   // namespace tpt
