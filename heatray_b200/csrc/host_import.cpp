// host_import.cpp — scene import for the host layer: Wavefront OBJ/MTL and glTF 2.0 (.gltf + buffers, .glb), own parsers.
//
// Reference: Source/HeatrayRenderer/Scene/AssimpMeshProvider.cpp:65-723 loads assets through the Assimp library with the
// post-processing flags of :680-690 and maps Assimp's material/light properties onto Heatray's (ProcessMaterial :362-514,
// ProcessGlassMaterial :264-360, ProcessLight :516-660, ProcessMesh :100-262, ProcessNode :73-98). Assimp is a large
// third-party dependency that is not linked here; this file reads the two formats Heatray's users feed it most directly and
// applies the SAME mapping rules, producing what MeshProvider exposes (MeshProvider.h:50-81): SoA vertex buffers, a
// triangle index buffer, a material index and a local transform per submesh, plus materials, lights and the scene AABB.
// What Assimp's post-processing contributes is restated: triangulation (fans), joined identical vertices, smooth normals
// when absent, tangent space when UVs exist (bitangent = cross(normal, tangent) as :217-222), global scale (cm -> m).
#include "host_api.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <new>
#include <stdexcept>
#include <string>
#include <unordered_map>
#include <vector>

namespace {

thread_local std::string g_importErr;

// ---------------------------------------------------------------------------------------
// Minimal JSON DOM (RFC 8259) for glTF
// ---------------------------------------------------------------------------------------
struct Json {
    enum Type { Null, Bool, Number, String, Array, Object } type = Null;
    double num = 0.0; bool b = false; std::string str;
    std::vector<Json> arr; std::vector<std::pair<std::string, Json>> obj;
    const Json* get(const char* key) const {
        if (type != Object) return nullptr;
        for (const auto& kv : obj) if (kv.first == key) return &kv.second;
        return nullptr;
    }
    const Json* at(size_t i) const { return (type == Array && i < arr.size()) ? &arr[i] : nullptr; }
    size_t size() const { return type == Array ? arr.size() : 0; }
    double number(double dflt) const { return type == Number ? num : dflt; }
};
double jnum(const Json* j, const char* key, double dflt) { const Json* v = j ? j->get(key) : nullptr; return v ? v->number(dflt) : dflt; }
int jint(const Json* j, const char* key, int dflt) { return (int)jnum(j, key, dflt); }
std::string jstr(const Json* j, const char* key, const char* dflt = "") { const Json* v = j ? j->get(key) : nullptr; return (v && v->type == Json::String) ? v->str : std::string(dflt); }

struct JsonParser {
    const char* p; const char* end; bool ok = true;
    void ws() { while (p < end && (*p == ' ' || *p == '\t' || *p == '\n' || *p == '\r')) ++p; }
    bool lit(const char* s) { const size_t n = std::strlen(s); if ((size_t)(end - p) >= n && !std::strncmp(p, s, n)) { p += n; return true; } return false; }
    void utf8(std::string& out, unsigned cp) {
        if (cp < 0x80) out.push_back((char)cp);
        else if (cp < 0x800) { out.push_back((char)(0xC0 | (cp >> 6))); out.push_back((char)(0x80 | (cp & 63))); }
        else if (cp < 0x10000) { out.push_back((char)(0xE0 | (cp >> 12))); out.push_back((char)(0x80 | ((cp >> 6) & 63))); out.push_back((char)(0x80 | (cp & 63))); }
        else { out.push_back((char)(0xF0 | (cp >> 18))); out.push_back((char)(0x80 | ((cp >> 12) & 63))); out.push_back((char)(0x80 | ((cp >> 6) & 63))); out.push_back((char)(0x80 | (cp & 63))); }
    }
    bool string(std::string& out) {
        if (p >= end || *p != '"') return false;
        ++p;
        while (p < end && *p != '"') {
            if (*p == '\\') {
                if (++p >= end) return false;
                switch (*p) {
                case 'n': out.push_back('\n'); break; case 't': out.push_back('\t'); break; case 'r': out.push_back('\r'); break;
                case 'b': out.push_back('\b'); break; case 'f': out.push_back('\f'); break;
                case 'u': {
                    if (end - p < 5) return false;
                    unsigned cp = (unsigned)std::strtoul(std::string(p + 1, p + 5).c_str(), nullptr, 16); p += 4;
                    if (cp >= 0xD800 && cp < 0xDC00 && end - p > 6 && p[1] == '\\' && p[2] == 'u') {
                        const unsigned lo = (unsigned)std::strtoul(std::string(p + 3, p + 7).c_str(), nullptr, 16);
                        cp = 0x10000 + ((cp - 0xD800) << 10) + (lo - 0xDC00); p += 6;
                    }
                    utf8(out, cp); break; }
                default: out.push_back(*p); break;
                }
                ++p;
            } else out.push_back(*p++);
        }
        if (p >= end) return false;
        ++p;
        return true;
    }
    Json value(int depth = 0) {
        Json j; ws();
        if (p >= end || depth > 200) { ok = false; return j; }
        if (*p == '{') {
            j.type = Json::Object; ++p; ws();
            if (p < end && *p == '}') { ++p; return j; }
            while (ok) {
                ws(); std::string k;
                if (!string(k)) { ok = false; break; }
                ws(); if (p >= end || *p != ':') { ok = false; break; } ++p;
                j.obj.emplace_back(std::move(k), value(depth + 1));
                ws(); if (p < end && *p == ',') { ++p; continue; }
                if (p < end && *p == '}') { ++p; break; }
                ok = false;
            }
        } else if (*p == '[') {
            j.type = Json::Array; ++p; ws();
            if (p < end && *p == ']') { ++p; return j; }
            while (ok) {
                j.arr.push_back(value(depth + 1));
                ws(); if (p < end && *p == ',') { ++p; continue; }
                if (p < end && *p == ']') { ++p; break; }
                ok = false;
            }
        } else if (*p == '"') { j.type = Json::String; if (!string(j.str)) ok = false; }
        else if (lit("true")) { j.type = Json::Bool; j.b = true; }
        else if (lit("false")) { j.type = Json::Bool; j.b = false; }
        else if (lit("null")) j.type = Json::Null;
        else {
            char* e = nullptr; j.num = std::strtod(p, &e);
            if (e == p || e > end) ok = false; else { j.type = Json::Number; p = e; }
        }
        return j;
    }
};

// ---------------------------------------------------------------------------------------
// Imported scene (what MeshProvider + Lighting receive from AssimpMeshProvider)
// ---------------------------------------------------------------------------------------
enum { kSlotBaseColor = 0, kSlotMetallicRoughness, kSlotEmissive, kSlotNormal, kSlotClearCoat, kSlotClearCoatRoughness, kSlotClearCoatNormal, kSlotCount };

struct ImpMaterial {
    std::string name; int type = HRC_MATERIAL_PBR;
    float baseColor[3] = {1, 1, 1}, emissive[3] = {0, 0, 0};
    float metallic = 0.0f, roughness = 1.0f, specularF0 = 0.5f, clearCoat = 0.0f, clearCoatRoughness = 0.0f;   // defaults of ProcessMaterial :379-383
    bool doubleSided = true, alphaMask = false, vertexColors = false;
    float ior = 1.33f, density = 0.05f;                                       // glass defaults, ProcessGlassMaterial :269-272
    std::string tex[kSlotCount];
};
struct ImpMesh {
    std::string name; int material = 0; float xf[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
    std::vector<float> pos, nrm, uv, tan, bit, col; std::vector<uint32_t> idx;
    float lo[3] = {0, 0, 0}, hi[3] = {0, 0, 0};
};
struct ImpLight { std::string name; int type = 0; float color[3] = {1, 1, 1}, intensity = 0, position[3] = {0, 0, 0}, theta = 0, phi = 0, inner = 0, outer = 0; };

} // namespace

struct HrhScene_t {
    std::vector<ImpMaterial> materials; std::vector<ImpMesh> meshes; std::vector<ImpLight> lights;
    float lo[3] = {0, 0, 0}, hi[3] = {0, 0, 0}; bool aabbValid = false;
    std::string dir;
    float sceneXf[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};      // Scene::applyTransform (Scene.cpp:38-49)
};

namespace {

// ---- column-major 4x4 helpers (glm layout) ----
void matIdentity(float* m) { for (int i = 0; i < 16; ++i) m[i] = (i % 5 == 0) ? 1.0f : 0.0f; }
void matMul(const float* a, const float* b, float* out) {       // out = a * b
    float r[16];
    for (int c = 0; c < 4; ++c) for (int row = 0; row < 4; ++row) {
        float s = 0.0f;
        for (int k = 0; k < 4; ++k) s += a[k * 4 + row] * b[c * 4 + k];
        r[c * 4 + row] = s;
    }
    std::memcpy(out, r, sizeof(r));
}
void matPoint(const float* m, const float* p, float* out) {
    for (int r = 0; r < 3; ++r) out[r] = m[r] * p[0] + m[4 + r] * p[1] + m[8 + r] * p[2] + m[12 + r];
}
void matFromTRS(const float* t, const float* q, const float* s, float* m) {      // q = (x, y, z, w)
    const float x = q[0], y = q[1], z = q[2], w = q[3];
    const float R[9] = {1 - 2 * (y * y + z * z), 2 * (x * y + z * w), 2 * (x * z - y * w),
                        2 * (x * y - z * w), 1 - 2 * (x * x + z * z), 2 * (y * z + x * w),
                        2 * (x * z + y * w), 2 * (y * z - x * w), 1 - 2 * (x * x + y * y)};   // columns
    for (int c = 0; c < 3; ++c) { for (int r = 0; r < 3; ++r) m[c * 4 + r] = R[c * 3 + r] * s[c]; m[c * 4 + 3] = 0.0f; }
    m[12] = t[0]; m[13] = t[1]; m[14] = t[2]; m[15] = 1.0f;
}

inline void cross3(const float* a, const float* b, float* o) { o[0] = a[1] * b[2] - a[2] * b[1]; o[1] = a[2] * b[0] - a[0] * b[2]; o[2] = a[0] * b[1] - a[1] * b[0]; }
inline float dot3(const float* a, const float* b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
inline bool normalize3(float* v) { const float l = std::sqrt(dot3(v, v)); if (!(l > 0.0f)) return false; v[0] /= l; v[1] /= l; v[2] /= l; return true; }

// aiProcess_GenSmoothNormals restated: the normal of a vertex is the normalised sum of the (normalised) face normals of every
// face that touches a vertex at the same POSITION, so split vertices (different uv) still shade smoothly.
void generateSmoothNormals(ImpMesh& m) {
    const size_t nv = m.pos.size() / 3;
    m.nrm.assign(nv * 3, 0.0f);
    struct Key { float x, y, z; bool operator<(const Key& o) const { return x != o.x ? x < o.x : (y != o.y ? y < o.y : z < o.z); } };
    std::map<Key, uint32_t> first; std::vector<uint32_t> rep(nv);
    for (size_t v = 0; v < nv; ++v) { const Key k{m.pos[3 * v], m.pos[3 * v + 1], m.pos[3 * v + 2]}; rep[v] = first.emplace(k, (uint32_t)v).first->second; }
    std::vector<float> acc(nv * 3, 0.0f);
    for (size_t t = 0; t + 2 < m.idx.size(); t += 3) {
        const float* a = &m.pos[3 * m.idx[t]]; const float* b = &m.pos[3 * m.idx[t + 1]]; const float* c = &m.pos[3 * m.idx[t + 2]];
        const float e1[3] = {b[0] - a[0], b[1] - a[1], b[2] - a[2]}, e2[3] = {c[0] - a[0], c[1] - a[1], c[2] - a[2]};
        float n[3]; cross3(e1, e2, n);
        if (!normalize3(n)) continue;
        for (int k = 0; k < 3; ++k) { float* d = &acc[3 * rep[m.idx[t + k]]]; d[0] += n[0]; d[1] += n[1]; d[2] += n[2]; }
    }
    for (size_t v = 0; v < nv; ++v) {
        float n[3] = {acc[3 * rep[v]], acc[3 * rep[v] + 1], acc[3 * rep[v] + 2]};
        if (!normalize3(n)) { n[0] = 0; n[1] = 1; n[2] = 0; }
        std::memcpy(&m.nrm[3 * v], n, 12);
    }
}

// aiProcess_CalcTangentSpace restated: per-face tangent from the uv derivatives, accumulated per vertex, made orthogonal to the
// vertex normal and normalised; the bitangent is then cross(normal, tangent) as AssimpMeshProvider.cpp:217-222 recomputes it.
void generateTangentSpace(ImpMesh& m) {
    const size_t nv = m.pos.size() / 3;
    if (m.uv.size() != nv * 2 || m.nrm.size() != nv * 3) return;
    std::vector<float> acc(nv * 3, 0.0f);
    for (size_t t = 0; t + 2 < m.idx.size(); t += 3) {
        const uint32_t i0 = m.idx[t], i1 = m.idx[t + 1], i2 = m.idx[t + 2];
        const float* p0 = &m.pos[3 * i0]; const float* p1 = &m.pos[3 * i1]; const float* p2 = &m.pos[3 * i2];
        const float e1[3] = {p1[0] - p0[0], p1[1] - p0[1], p1[2] - p0[2]}, e2[3] = {p2[0] - p0[0], p2[1] - p0[1], p2[2] - p0[2]};
        const float du1 = m.uv[2 * i1] - m.uv[2 * i0], dv1 = m.uv[2 * i1 + 1] - m.uv[2 * i0 + 1];
        const float du2 = m.uv[2 * i2] - m.uv[2 * i0], dv2 = m.uv[2 * i2 + 1] - m.uv[2 * i0 + 1];
        float det = du1 * dv2 - du2 * dv1;
        if (det == 0.0f) det = 1.0f;                                      // degenerate uv mapping: keep a usable direction (Assimp does the same)
        const float r = 1.0f / det;
        const float tg[3] = {(e1[0] * dv2 - e2[0] * dv1) * r, (e1[1] * dv2 - e2[1] * dv1) * r, (e1[2] * dv2 - e2[2] * dv1) * r};
        for (uint32_t v : {i0, i1, i2}) { acc[3 * v] += tg[0]; acc[3 * v + 1] += tg[1]; acc[3 * v + 2] += tg[2]; }
    }
    m.tan.assign(nv * 3, 0.0f); m.bit.assign(nv * 3, 0.0f);
    for (size_t v = 0; v < nv; ++v) {
        const float* n = &m.nrm[3 * v];
        float t[3] = {acc[3 * v], acc[3 * v + 1], acc[3 * v + 2]};
        const float d = dot3(t, n);
        t[0] -= n[0] * d; t[1] -= n[1] * d; t[2] -= n[2] * d;
        if (!normalize3(t)) {                                            // no usable uv gradient: any direction orthogonal to the normal
            const float ax[3] = {std::fabs(n[0]) < 0.9f ? 1.0f : 0.0f, std::fabs(n[0]) < 0.9f ? 0.0f : 1.0f, 0.0f};
            cross3(ax, n, t); normalize3(t);
        }
        std::memcpy(&m.tan[3 * v], t, 12);
        cross3(n, t, &m.bit[3 * v]);
    }
}

void finishMesh(ImpMesh& m, float globalScale) {
    if (globalScale != 1.0f) for (float& v : m.pos) v *= globalScale;     // aiProcess_GlobalScale on the vertices (ScaleProcess)
    if (m.nrm.size() != m.pos.size()) generateSmoothNormals(m);
    if (!m.uv.empty()) generateTangentSpace(m);
    for (int a = 0; a < 3; ++a) { m.lo[a] = 3.4e38f; m.hi[a] = -3.4e38f; }
    for (size_t v = 0; v + 2 < m.pos.size(); v += 3) for (int a = 0; a < 3; ++a) { m.lo[a] = std::min(m.lo[a], m.pos[v + a]); m.hi[a] = std::max(m.hi[a], m.pos[v + a]); }
}

// ProcessNode :88-95: the scene AABB grows by the transformed min and max corners of each submesh's AABB (two points, not eight).
void growSceneAABB(HrhScene_t& s, const ImpMesh& m) {
    if (m.pos.empty()) return;
    float a[3], b[3]; matPoint(m.xf, m.lo, a); matPoint(m.xf, m.hi, b);
    for (const float* p : {a, b}) for (int k = 0; k < 3; ++k) {
        if (!s.aabbValid) { s.lo[k] = s.hi[k] = p[k]; } else { s.lo[k] = std::min(s.lo[k], p[k]); s.hi[k] = std::max(s.hi[k], p[k]); }
        if (k == 2) s.aabbValid = true;
    }
}

std::string dirOf(const std::string& path) { const size_t k = path.find_last_of("/\\"); return k == std::string::npos ? std::string() : path.substr(0, k + 1); }
bool readTextFile(const std::string& path, std::string& out) {
    FILE* f = std::fopen(path.c_str(), "rb"); if (!f) return false;
    std::fseek(f, 0, SEEK_END); const long n = std::ftell(f); std::fseek(f, 0, SEEK_SET);
    out.resize(n > 0 ? (size_t)n : 0);
    const size_t got = n > 0 ? std::fread(&out[0], 1, (size_t)n, f) : 0;
    std::fclose(f);
    return got == out.size();
}

} // namespace

// ---------------------------------------------------------------------------------------
// Wavefront OBJ + MTL. Material keys follow what Assimp's OBJ importer exposes and ProcessMaterial then reads
// (AssimpMeshProvider.cpp:394-425): Kd -> base colour (default 0.6 grey), Ke -> emissive, Pr/Pm/Pc/Pcr -> roughness / metallic /
// clearcoat / clearcoat roughness when present, and Ni (default 1.0, ALWAYS present) -> specularF0 = ((1-ior)/(1+ior))^2.
// Textures: map_Kd -> base colour, map_Ke -> emissive, norm / map_Kn -> normal map. One submesh per (object/group, material).
// ---------------------------------------------------------------------------------------
namespace {

struct ObjMtl { ImpMaterial m; };

const char* skipWs(const char* p) { while (*p == ' ' || *p == '\t') ++p; return p; }
std::string restOfLine(const char* p) {
    p = skipWs(p); std::string s(p);
    while (!s.empty() && (s.back() == '\r' || s.back() == '\n' || s.back() == ' ' || s.back() == '\t')) s.pop_back();
    return s;
}
// "map_Kd -s 1 1 1 -o 0 0 0 file name.png": options start with '-', the file name is what remains
std::string mapFileName(const char* p) {
    std::string s = restOfLine(p);
    size_t pos = 0;
    while (pos < s.size() && s[pos] == '-') {
        size_t e = s.find(' ', pos); if (e == std::string::npos) return std::string();
        const std::string opt = s.substr(pos, e - pos);
        int args = (opt == "-s" || opt == "-o" || opt == "-t") ? 3 : (opt == "-mm") ? 2 : 1;
        pos = e;
        while (args-- > 0) { while (pos < s.size() && s[pos] == ' ') ++pos; while (pos < s.size() && s[pos] != ' ') ++pos; }
        while (pos < s.size() && s[pos] == ' ') ++pos;
    }
    std::string f = s.substr(pos);
    std::replace(f.begin(), f.end(), '\\', '/');
    return f;
}

void loadMtl(const std::string& path, std::vector<ImpMaterial>& mats, std::unordered_map<std::string, int>& byName) {
    std::string text;
    if (!readTextFile(path, text)) return;                              // a missing .mtl leaves the default material, like Assimp
    ImpMaterial* cur = nullptr; float ior = 1.0f; bool any = false;
    auto finish = [&]() { if (cur) cur->specularF0 = std::pow((1.0f - ior) / (1.0f + ior), 2.0f); };   // :419-422
    size_t pos = 0;
    while (pos < text.size()) {
        size_t e = text.find('\n', pos); if (e == std::string::npos) e = text.size();
        const std::string line = text.substr(pos, e - pos); pos = e + 1;
        const char* p = skipWs(line.c_str());
        if (*p == '#' || *p == 0 || *p == '\r') continue;
        float a, b, c;
        if (!std::strncmp(p, "newmtl", 6)) {
            finish();
            ImpMaterial m; m.name = restOfLine(p + 6); m.baseColor[0] = m.baseColor[1] = m.baseColor[2] = 0.6f; ior = 1.0f;
            byName[m.name] = (int)mats.size(); mats.push_back(m); cur = &mats.back(); any = true;
        } else if (!cur) continue;
        else if (!std::strncmp(p, "Kd ", 3)) { if (std::sscanf(p + 3, "%f %f %f", &a, &b, &c) == 3) { cur->baseColor[0] = a; cur->baseColor[1] = b; cur->baseColor[2] = c; } }
        else if (!std::strncmp(p, "Ke ", 3)) { if (std::sscanf(p + 3, "%f %f %f", &a, &b, &c) == 3) { cur->emissive[0] = a; cur->emissive[1] = b; cur->emissive[2] = c; } }
        else if (!std::strncmp(p, "Ni ", 3)) { std::sscanf(p + 3, "%f", &ior); }
        else if (!std::strncmp(p, "Pr ", 3)) { std::sscanf(p + 3, "%f", &cur->roughness); }
        else if (!std::strncmp(p, "Pm ", 3)) { std::sscanf(p + 3, "%f", &cur->metallic); }
        else if (!std::strncmp(p, "Pcr ", 4)) { std::sscanf(p + 4, "%f", &cur->clearCoatRoughness); }
        else if (!std::strncmp(p, "Pc ", 3)) { std::sscanf(p + 3, "%f", &cur->clearCoat); }
        else if (!std::strncmp(p, "map_Kd", 6)) cur->tex[kSlotBaseColor] = mapFileName(p + 6);
        else if (!std::strncmp(p, "map_Ke", 6)) cur->tex[kSlotEmissive] = mapFileName(p + 6);
        else if (!std::strncmp(p, "map_emissive", 12)) cur->tex[kSlotEmissive] = mapFileName(p + 12);
        else if (!std::strncmp(p, "map_Kn", 6)) cur->tex[kSlotNormal] = mapFileName(p + 6);
        else if (!std::strncmp(p, "norm", 4)) cur->tex[kSlotNormal] = mapFileName(p + 4);
    }
    finish();
    (void)any;
}

struct ObjKey { int v, t, n; bool operator==(const ObjKey& o) const { return v == o.v && t == o.t && n == o.n; } };
struct ObjKeyHash { size_t operator()(const ObjKey& k) const { return ((size_t)k.v * 73856093u) ^ ((size_t)k.t * 19349663u) ^ ((size_t)k.n * 83492791u); } };

bool importObj(const std::string& path, float globalScale, HrhScene_t& scene) {
    std::string text;
    if (!readTextFile(path, text)) { g_importErr = "cannot read " + path; return false; }
    std::vector<float> V, VT, VN, VC;                                   // positions (+ optional per-vertex colours "v x y z r g b")
    std::unordered_map<std::string, int> mtlByName;
    struct Group { std::string name; int material; std::vector<ObjKey> corners; std::vector<uint32_t> faceSizes; };
    std::vector<Group> groups;
    std::string curName = "defaultobject"; int curMat = -1;
    auto group = [&]() -> Group& {
        for (auto& g : groups) if (g.name == curName && g.material == curMat) return g;
        groups.push_back(Group{curName, curMat, {}, {}});
        return groups.back();
    };
    size_t pos = 0; bool anyColor = false;
    while (pos < text.size()) {
        size_t e = text.find('\n', pos); if (e == std::string::npos) e = text.size();
        std::string line = text.substr(pos, e - pos); pos = e + 1;
        while (!line.empty() && line.back() == '\\' && pos < text.size()) {          // line continuation
            line.pop_back(); size_t e2 = text.find('\n', pos); if (e2 == std::string::npos) e2 = text.size();
            line += text.substr(pos, e2 - pos); pos = e2 + 1;
        }
        const char* p = skipWs(line.c_str());
        if (p[0] == 'v' && (p[1] == ' ' || p[1] == '\t')) {
            float x[6] = {0, 0, 0, 1, 1, 1};
            const int n = std::sscanf(p + 2, "%f %f %f %f %f %f", x, x + 1, x + 2, x + 3, x + 4, x + 5);
            V.insert(V.end(), x, x + 3);
            if (n >= 6) anyColor = true;
            VC.insert(VC.end(), x + 3, x + 6);
        } else if (p[0] == 'v' && p[1] == 't') { float u = 0, v = 0; std::sscanf(p + 2, "%f %f", &u, &v); VT.push_back(u); VT.push_back(v); }
        else if (p[0] == 'v' && p[1] == 'n') { float x = 0, y = 0, z = 0; std::sscanf(p + 2, "%f %f %f", &x, &y, &z); VN.push_back(x); VN.push_back(y); VN.push_back(z); }
        else if (p[0] == 'f' && (p[1] == ' ' || p[1] == '\t')) {
            Group& g = group(); uint32_t count = 0; const char* q = p + 1;
            for (;;) {
                q = skipWs(q);
                if (*q == 0 || *q == '\r' || *q == '#') break;
                ObjKey k{0, 0, 0}; char* endp = nullptr;
                k.v = (int)std::strtol(q, &endp, 10); if (endp == q) break; q = endp;
                if (*q == '/') { ++q; if (*q != '/') { k.t = (int)std::strtol(q, &endp, 10); q = endp; } if (*q == '/') { ++q; k.n = (int)std::strtol(q, &endp, 10); q = endp; } }
                const int nv = (int)(V.size() / 3), nt = (int)(VT.size() / 2), nn = (int)(VN.size() / 3);
                k.v = k.v < 0 ? nv + k.v : k.v - 1; k.t = k.t < 0 ? nt + k.t : k.t - 1; k.n = k.n < 0 ? nn + k.n : k.n - 1;   // 1-based, negative = relative
                if (k.v < 0 || k.v >= nv) { g_importErr = "OBJ face references a missing vertex"; return false; }
                if (k.t >= nt) k.t = -1;
                if (k.n >= nn) k.n = -1;
                g.corners.push_back(k); ++count;
            }
            if (count >= 3) g.faceSizes.push_back(count); else g.corners.resize(g.corners.size() - count);   // points and lines are dropped (triangles only)
        } else if (!std::strncmp(p, "usemtl", 6)) {
            const std::string n = restOfLine(p + 6); auto it = mtlByName.find(n);
            if (it == mtlByName.end()) {                                                                     // unknown name: Assimp creates the material on the fly with defaults
                ImpMaterial m; m.name = n; m.baseColor[0] = m.baseColor[1] = m.baseColor[2] = 0.6f; m.specularF0 = 0.0f;
                mtlByName[n] = (int)scene.materials.size(); scene.materials.push_back(m); curMat = mtlByName[n];
            } else curMat = it->second;
        } else if (!std::strncmp(p, "mtllib", 6)) loadMtl(scene.dir + restOfLine(p + 6), scene.materials, mtlByName);
        else if ((p[0] == 'o' || p[0] == 'g') && (p[1] == ' ' || p[1] == '\t')) { curName = restOfLine(p + 1); if (curName.empty()) curName = "default"; }
    }
    int defaultMat = -1;
    for (auto& g : groups) {
        if (g.faceSizes.empty()) continue;
        if (g.material < 0) {
            if (defaultMat < 0) { ImpMaterial m; m.name = "DefaultMaterial"; m.baseColor[0] = m.baseColor[1] = m.baseColor[2] = 0.6f; m.specularF0 = 0.0f; defaultMat = (int)scene.materials.size(); scene.materials.push_back(m); }
            g.material = defaultMat;
        }
        ImpMesh mesh; mesh.name = g.name; mesh.material = g.material;
        bool hasT = true, hasN = true;
        for (const ObjKey& k : g.corners) { hasT = hasT && k.t >= 0; hasN = hasN && k.n >= 0; }
        std::unordered_map<ObjKey, uint32_t, ObjKeyHash> unique;                                             // aiProcess_JoinIdenticalVertices
        std::vector<uint32_t> cornerIndex; cornerIndex.reserve(g.corners.size());
        for (ObjKey k : g.corners) {
            if (!hasT) k.t = -1;
            if (!hasN) k.n = -1;
            auto it = unique.find(k);
            if (it == unique.end()) {
                const uint32_t id = (uint32_t)(mesh.pos.size() / 3);
                it = unique.emplace(k, id).first;
                mesh.pos.insert(mesh.pos.end(), &V[3 * k.v], &V[3 * k.v] + 3);
                if (hasT) { mesh.uv.push_back(VT[2 * k.t]); mesh.uv.push_back(VT[2 * k.t + 1]); }
                if (hasN) mesh.nrm.insert(mesh.nrm.end(), &VN[3 * k.n], &VN[3 * k.n] + 3);
                if (anyColor) mesh.col.insert(mesh.col.end(), &VC[3 * k.v], &VC[3 * k.v] + 3);
            }
            cornerIndex.push_back(it->second);
        }
        size_t c = 0;
        for (uint32_t fs : g.faceSizes) {                                                                    // fan, as ProcessMesh :242-250 does for n-gons
            for (uint32_t k = 2; k < fs; ++k) { mesh.idx.push_back(cornerIndex[c]); mesh.idx.push_back(cornerIndex[c + k - 1]); mesh.idx.push_back(cornerIndex[c + k]); }
            c += fs;
        }
        if (anyColor) scene.materials[mesh.material].vertexColors = true;                                    // :229 enableVertexColors()
        finishMesh(mesh, globalScale);
        growSceneAABB(scene, mesh);
        scene.meshes.push_back(std::move(mesh));
    }
    if (scene.meshes.empty()) { g_importErr = "no triangles in " + path; return false; }
    for (ImpMaterial& m : scene.materials) for (std::string& t : m.tex) if (!t.empty() && t[0] != '/') t = scene.dir + t;   // fileParent / texture path (:313, :449)
    return true;
}

} // namespace

// ---------------------------------------------------------------------------------------
// glTF 2.0 (.gltf with external / data-URI buffers, .glb)
// ---------------------------------------------------------------------------------------
namespace {

bool base64Decode(const std::string& s, size_t start, std::vector<uint8_t>& out) {
    uint32_t acc = 0; int bits = 0;
    for (size_t i = start; i < s.size(); ++i) {
        const char c = s[i]; int v;
        if (c >= 'A' && c <= 'Z') v = c - 'A'; else if (c >= 'a' && c <= 'z') v = c - 'a' + 26; else if (c >= '0' && c <= '9') v = c - '0' + 52;
        else if (c == '+' || c == '-') v = 62; else if (c == '/' || c == '_') v = 63; else if (c == '=') break; else continue;
        acc = (acc << 6) | (uint32_t)v; bits += 6;
        if (bits >= 8) { bits -= 8; out.push_back((uint8_t)(acc >> bits)); }
    }
    return true;
}

struct Gltf {
    Json doc; std::vector<std::vector<uint8_t>> buffers; std::string dir;

    // Reads accessor `index` as floats, `comps` per element (missing components 0, extra ones dropped); integers are normalised when flagged.
    // Byte range of an accessor inside its buffer, validated: offsets and stride come from the file and may be negative, huge or NaN.
    static bool span(const Json* bv, const Json* acc, int count, size_t elemBytes, size_t bufBytes, size_t& off, size_t& stride) {
        const double o1 = jnum(bv, "byteOffset", 0), o2 = jnum(acc, "byteOffset", 0), st = jnum(bv, "byteStride", 0);
        if (!(o1 >= 0.0) || !(o2 >= 0.0) || !(st >= 0.0) || o1 > (double)bufBytes || o2 > (double)bufBytes || st > 65536.0) return false;
        off = (size_t)o1 + (size_t)o2; stride = (size_t)st; if (!stride) stride = elemBytes;
        if (count <= 0) return count == 0;
        if ((size_t)count > bufBytes) return false;                        // more elements than bytes: also keeps the product below in range
        return off + stride * (size_t)(count - 1) + elemBytes <= bufBytes;
    }
    bool readFloats(int index, int comps, std::vector<float>& out, int* fileComps = nullptr) const {
        const Json* acc = doc.get("accessors") ? doc.get("accessors")->at((size_t)index) : nullptr;
        if (!acc) return false;
        const int count = jint(acc, "count", 0), ctype = jint(acc, "componentType", 0);
        const std::string type = jstr(acc, "type");
        const int nc = type == "SCALAR" ? 1 : type == "VEC2" ? 2 : type == "VEC3" ? 3 : type == "VEC4" ? 4 : 0;
        if (fileComps) *fileComps = nc;
        const Json* nrm = acc->get("normalized"); const bool normalized = nrm && nrm->type == Json::Bool && nrm->b;
        const int csize = (ctype == 5120 || ctype == 5121) ? 1 : (ctype == 5122 || ctype == 5123) ? 2 : (ctype == 5125 || ctype == 5126) ? 4 : 0;
        if (!nc || !csize || count < 0) return false;
        out.assign((size_t)count * comps, 0.0f);
        const Json* bvIdx = acc->get("bufferView");
        if (!bvIdx) return true;                                           // no bufferView: all zeros (sparse accessors are not applied)
        const Json* bv = doc.get("bufferViews") ? doc.get("bufferViews")->at((size_t)bvIdx->number(-1)) : nullptr;
        if (!bv) return false;
        const int buf = jint(bv, "buffer", -1);
        if (buf < 0 || buf >= (int)buffers.size()) return false;
        const std::vector<uint8_t>& b = buffers[buf];
        size_t off = 0, stride = 0;
        if (!span(bv, acc, count, (size_t)nc * csize, b.size(), off, stride)) return false;
        for (int i = 0; i < count; ++i) for (int c = 0; c < std::min(nc, comps); ++c) {
            const uint8_t* p = b.data() + off + stride * i + (size_t)c * csize; float v;
            switch (ctype) {
            case 5120: { int8_t x; std::memcpy(&x, p, 1); v = normalized ? std::max((float)x / 127.0f, -1.0f) : (float)x; break; }
            case 5121: { v = normalized ? (float)p[0] / 255.0f : (float)p[0]; break; }
            case 5122: { int16_t x; std::memcpy(&x, p, 2); v = normalized ? std::max((float)x / 32767.0f, -1.0f) : (float)x; break; }
            case 5123: { uint16_t x; std::memcpy(&x, p, 2); v = normalized ? (float)x / 65535.0f : (float)x; break; }
            case 5125: { uint32_t x; std::memcpy(&x, p, 4); v = (float)x; break; }
            default: std::memcpy(&v, p, 4); break;
            }
            out[(size_t)i * comps + c] = v;
        }
        return true;
    }
    bool readIndices(int index, std::vector<uint32_t>& out) const {
        std::vector<float> f;                                              // exact up to 2^24; larger u32 indices are read directly below
        const Json* acc = doc.get("accessors") ? doc.get("accessors")->at((size_t)index) : nullptr;
        if (!acc) return false;
        if (jint(acc, "componentType", 0) != 5125) { if (!readFloats(index, 1, f)) return false; out.resize(f.size()); for (size_t i = 0; i < f.size(); ++i) out[i] = (uint32_t)f[i]; return true; }
        const Json* bv = doc.get("bufferViews") && acc->get("bufferView") ? doc.get("bufferViews")->at((size_t)acc->get("bufferView")->number(-1)) : nullptr;
        if (!bv) return false;
        const int buf = jint(bv, "buffer", -1), count = jint(acc, "count", 0);
        if (buf < 0 || buf >= (int)buffers.size()) return false;
        size_t off = 0, stride = 0;
        if (count < 0 || !span(bv, acc, count, 4, buffers[buf].size(), off, stride)) return false;
        out.resize((size_t)count);
        for (int i = 0; i < count; ++i) std::memcpy(&out[i], buffers[buf].data() + off + stride * i, 4);
        return true;
    }
    std::string imagePath(const Json* texInfo) const {                      // textureInfo -> textures[i].source -> images[j].uri, relative to the asset
        if (!texInfo) return std::string();
        const Json* texs = doc.get("textures"); const Json* imgs = doc.get("images");
        const Json* tex = texs ? texs->at((size_t)jint(texInfo, "index", -1)) : nullptr;
        const Json* img = (tex && imgs) ? imgs->at((size_t)jint(tex, "source", -1)) : nullptr;
        const std::string uri = jstr(img, "uri");
        if (uri.empty() || uri.compare(0, 5, "data:") == 0) return std::string();   // embedded images: Assimp hands "*N" names the reference cannot open either
        std::string out;                                                    // percent-decoding
        for (size_t i = 0; i < uri.size(); ++i) {
            if (uri[i] == '%' && i + 2 < uri.size()) { out.push_back((char)std::strtol(uri.substr(i + 1, 2).c_str(), nullptr, 16)); i += 2; }
            else out.push_back(uri[i]);
        }
        return dir + out;
    }
};

void readVec(const Json* j, const char* key, float* out, int n) {
    const Json* a = j ? j->get(key) : nullptr;
    if (a && a->type == Json::Array) for (int i = 0; i < n && i < (int)a->arr.size(); ++i) out[i] = (float)a->arr[i].number(out[i]);
}

ImpMaterial gltfMaterial(const Gltf& g, const Json& m) {
    ImpMaterial out; out.name = jstr(&m, "name");
    const Json* pbr = m.get("pbrMetallicRoughness"); const Json* ext = m.get("extensions");
    float base[4] = {1, 1, 1, 1}; readVec(pbr, "baseColorFactor", base, 4);
    const std::string alphaMode = jstr(&m, "alphaMode", "OPAQUE");
    const Json* trans = ext ? ext->get("KHR_materials_transmission") : nullptr;
    const float transmission = (float)jnum(trans, "transmissionFactor", 0.0);
    const Json* iorExt = ext ? ext->get("KHR_materials_ior") : nullptr;
    std::memcpy(out.baseColor, base, 12);
    out.roughness = (float)jnum(pbr, "roughnessFactor", 1.0);
    out.tex[kSlotBaseColor] = g.imagePath(pbr ? pbr->get("baseColorTexture") : nullptr);
    out.tex[kSlotNormal] = g.imagePath(m.get("normalTexture"));
    out.tex[kSlotMetallicRoughness] = g.imagePath(pbr ? pbr->get("metallicRoughnessTexture") : nullptr);
    if (alphaMode == "BLEND" || transmission != 0.0f) {                     // ProcessMaterial :365-376 -> ProcessGlassMaterial :264-360
        out.type = HRC_MATERIAL_GLASS;
        out.ior = iorExt ? (float)jnum(iorExt, "ior", 1.5) : 1.33f;
        out.density = 0.05f;
        return out;
    }
    out.alphaMask = alphaMode == "MASK";                                     // :386-392
    out.metallic = (float)jnum(pbr, "metallicFactor", 1.0);
    const Json* cc = ext ? ext->get("KHR_materials_clearcoat") : nullptr;
    if (cc) {
        out.clearCoat = (float)jnum(cc, "clearcoatFactor", 0.0); out.clearCoatRoughness = (float)jnum(cc, "clearcoatRoughnessFactor", 0.0);
        out.tex[kSlotClearCoat] = g.imagePath(cc->get("clearcoatTexture")); out.tex[kSlotClearCoatRoughness] = g.imagePath(cc->get("clearcoatRoughnessTexture"));
        out.tex[kSlotClearCoatNormal] = g.imagePath(cc->get("clearcoatNormalTexture"));
    }
    readVec(&m, "emissiveFactor", out.emissive, 3);
    out.tex[kSlotEmissive] = g.imagePath(m.get("emissiveTexture"));
    const Json* ds = m.get("doubleSided"); out.doubleSided = ds && ds->type == Json::Bool && ds->b;   // glTF default false (AI_MATKEY_TWOSIDED is always written)
    if (iorExt) { const float ior = (float)jnum(iorExt, "ior", 1.5); out.specularF0 = std::pow((1.0f - ior) / (1.0f + ior), 2.0f); }   // :417-424
    return out;
}

// glm::extractEulerAngleXYZ on a column-major matrix (gtx/euler_angles.inl), as ProcessLight uses it (:573-586, :611-624)
void extractEulerXYZ(const float* M, float& t1, float& t2, float& t3) {
    auto m = [&](int c, int r) { return M[c * 4 + r]; };
    const float T1 = std::atan2(m(2, 1), m(2, 2));
    const float C2 = std::sqrt(m(0, 0) * m(0, 0) + m(1, 0) * m(1, 0));
    const float T2 = std::atan2(-m(2, 0), C2);
    const float S1 = std::sin(T1), C1 = std::cos(T1);
    const float T3 = std::atan2(S1 * m(0, 2) - C1 * m(0, 1), C1 * m(1, 1) - S1 * m(1, 2));
    t1 = -T1; t2 = -T2; t3 = -T3;
}

// Orientation of an imported directional / spot light, ProcessLight :564-586: local frame y = up (0,1,0), z = direction (0,0,-1),
// x = cross(y, z) — the reference builds the columns as vec4(axis, 1), so the node's translation leaks into the product — then
// glm::extractEulerAngleXYZ of lightTransform * local, yaw clamped to [0, 2 pi], pitch to [-pi/2, pi/2].
void lightAngles(const float* world, float& theta, float& phi) {
    const float up[3] = {0, 1, 0}, dir[3] = {0, 0, -1}; float x[3]; cross3(up, dir, x);
    const float localFrame[16] = {x[0], x[1], x[2], 1, up[0], up[1], up[2], 1, dir[0], dir[1], dir[2], 1, 0, 0, 0, 1};
    float fin[16]; matMul(world, localFrame, fin);
    float pitch, yaw, roll; extractEulerXYZ(fin, pitch, yaw, roll);
    phi = std::min(std::max(yaw, 0.0f), 6.283185307179586476925286766559f);
    theta = std::min(std::max(pitch, -1.5707963267948966192313216916398f), 1.5707963267948966192313216916398f);
}

} // namespace

namespace {

bool importGltf(const std::string& path, float globalScale, HrhScene_t& scene) {
    std::string file;
    if (!readTextFile(path, file)) { g_importErr = "cannot read " + path; return false; }
    Gltf g; g.dir = scene.dir;
    std::vector<uint8_t> glbBin; bool haveBin = false;
    std::string jsonText;
    if (file.size() >= 12 && !std::memcmp(file.data(), "glTF", 4)) {             // GLB container: 12-byte header, then chunks
        size_t pos = 12;
        while (pos + 8 <= file.size()) {
            uint32_t len, type; std::memcpy(&len, &file[pos], 4); std::memcpy(&type, &file[pos + 4], 4); pos += 8;
            if (pos + len > file.size()) { g_importErr = "truncated GLB chunk"; return false; }
            if (type == 0x4E4F534Au) jsonText.assign(&file[pos], len);
            else if (type == 0x004E4942u && !haveBin) { glbBin.assign(file.begin() + pos, file.begin() + pos + len); haveBin = true; }
            pos += (len + 3) & ~3u;
        }
    } else jsonText.swap(file);
    JsonParser jp{jsonText.data(), jsonText.data() + jsonText.size()};
    g.doc = jp.value();
    if (!jp.ok || g.doc.type != Json::Object) { g_importErr = "malformed glTF JSON"; return false; }

    if (const Json* bufs = g.doc.get("buffers")) for (size_t i = 0; i < bufs->size(); ++i) {
        const std::string uri = jstr(bufs->at(i), "uri");
        std::vector<uint8_t> data;
        if (uri.empty()) { if (i == 0 && haveBin) data = glbBin; }
        else if (uri.compare(0, 5, "data:") == 0) { const size_t c = uri.find(','); if (c != std::string::npos) base64Decode(uri, c + 1, data); }
        else { std::string raw; if (!readTextFile(g.dir + uri, raw)) { g_importErr = "cannot read buffer " + uri; return false; } data.assign(raw.begin(), raw.end()); }
        g.buffers.push_back(std::move(data));
    }
    if (const Json* mats = g.doc.get("materials")) for (size_t i = 0; i < mats->size(); ++i) scene.materials.push_back(gltfMaterial(g, mats->arr[i]));
    int defaultMat = -1;

    // one submesh per primitive, in mesh order (Assimp: one aiMesh per primitive); meshStart[m] = first submesh of glTF mesh m
    std::vector<int> meshStart, meshCount;
    if (const Json* meshes = g.doc.get("meshes")) for (size_t mi = 0; mi < meshes->size(); ++mi) {
        meshStart.push_back((int)scene.meshes.size()); int made = 0;
        const Json* prims = meshes->arr[mi].get("primitives");
        for (size_t pi = 0; prims && pi < prims->size(); ++pi) {
            const Json& pr = prims->arr[pi];
            const int mode = jint(&pr, "mode", 4);
            const Json* at = pr.get("attributes");
            if (!at || !at->get("POSITION") || (mode != 4 && mode != 5 && mode != 6)) continue;   // points / lines are not renderable here
            ImpMesh mesh; mesh.name = jstr(&meshes->arr[mi], "name");
            if (!g.readFloats((int)at->get("POSITION")->number(-1), 3, mesh.pos)) { g_importErr = "bad POSITION accessor"; return false; }
            const size_t nv = mesh.pos.size() / 3;
            if (const Json* a = at->get("NORMAL")) { if (!g.readFloats((int)a->number(-1), 3, mesh.nrm) || mesh.nrm.size() != nv * 3) mesh.nrm.clear(); }
            if (const Json* a = at->get("TEXCOORD_0")) {
                if (!g.readFloats((int)a->number(-1), 2, mesh.uv) || mesh.uv.size() != nv * 2) mesh.uv.clear();
                for (size_t v = 0; v < mesh.uv.size() / 2; ++v) mesh.uv[2 * v + 1] = 1.0f - mesh.uv[2 * v + 1];      // Assimp's glTF importer flips v; textures are loaded flipped too
            }
            bool fileTangents = false;
            if (const Json* a = at->get("TANGENT")) { fileTangents = g.readFloats((int)a->number(-1), 3, mesh.tan) && mesh.tan.size() == nv * 3 && !mesh.nrm.empty(); if (!fileTangents) mesh.tan.clear(); }
            if (const Json* a = at->get("COLOR_0")) { if (!g.readFloats((int)a->number(-1), 3, mesh.col) || mesh.col.size() != nv * 3) mesh.col.clear(); }
            std::vector<uint32_t> raw;
            if (const Json* ix = pr.get("indices")) { if (!g.readIndices((int)ix->number(-1), raw)) { g_importErr = "bad index accessor"; return false; } }
            else { raw.resize(nv); for (size_t v = 0; v < nv; ++v) raw[v] = (uint32_t)v; }
            for (uint32_t v : raw) if (v >= nv) { g_importErr = "index out of range"; return false; }
            if (mode == 4) { raw.resize(raw.size() / 3 * 3); mesh.idx = raw; }
            else if (mode == 5) for (size_t k = 2; k < raw.size(); ++k) {                                            // strip -> triangles, winding preserved
                if (k & 1) { mesh.idx.push_back(raw[k - 1]); mesh.idx.push_back(raw[k - 2]); mesh.idx.push_back(raw[k]); }
                else { mesh.idx.push_back(raw[k - 2]); mesh.idx.push_back(raw[k - 1]); mesh.idx.push_back(raw[k]); }
            } else for (size_t k = 2; k < raw.size(); ++k) { mesh.idx.push_back(raw[0]); mesh.idx.push_back(raw[k - 1]); mesh.idx.push_back(raw[k]); }
            const int mat = jint(&pr, "material", -1);
            if (mat >= 0 && mat < (int)scene.materials.size()) mesh.material = mat;
            else {
                if (defaultMat < 0) { ImpMaterial d = gltfMaterial(g, Json()); d.name = "DefaultMaterial"; defaultMat = (int)scene.materials.size(); scene.materials.push_back(d); }
                mesh.material = defaultMat;
            }
            if (!mesh.col.empty()) scene.materials[mesh.material].vertexColors = true;
            std::vector<float> keepTan; if (fileTangents) keepTan = mesh.tan;
            finishMesh(mesh, globalScale);
            if (fileTangents) {                                                                                      // tangents from the file win; bitangent = cross(n, t) (:217-222)
                mesh.tan = keepTan; mesh.bit.assign(nv * 3, 0.0f);
                for (size_t v = 0; v < nv; ++v) cross3(&mesh.nrm[3 * v], &mesh.tan[3 * v], &mesh.bit[3 * v]);
            }
            scene.meshes.push_back(std::move(mesh)); ++made;
        }
        meshCount.push_back(made);
    }

    // node hierarchy: world transform per node, depth first from the scene roots (ProcessNode :73-98)
    const Json* nodes = g.doc.get("nodes");
    const Json* lightDefs = nullptr;
    if (const Json* e = g.doc.get("extensions")) if (const Json* kl = e->get("KHR_lights_punctual")) lightDefs = kl->get("lights");
    struct Visit { int node; float parent[16]; };
    std::vector<Visit> stack;
    const Json* scenes = g.doc.get("scenes");
    const Json* sc = scenes ? scenes->at((size_t)jint(&g.doc, "scene", 0)) : nullptr;
    std::vector<int> roots;
    if (sc && sc->get("nodes")) for (const Json& r : sc->get("nodes")->arr) roots.push_back((int)r.number(-1));
    else if (nodes) for (size_t i = 0; i < nodes->size(); ++i) roots.push_back((int)i);
    for (auto it = roots.rbegin(); it != roots.rend(); ++it) { Visit v; v.node = *it; matIdentity(v.parent); stack.push_back(v); }
    int guard = 0;
    while (!stack.empty() && nodes && ++guard < 10000000) {
        const Visit cur = stack.back(); stack.pop_back();
        const Json* n = nodes->at((size_t)cur.node);
        if (!n) continue;
        float local[16]; matIdentity(local);
        if (const Json* mj = n->get("matrix")) { if (mj->size() == 16) for (int i = 0; i < 16; ++i) local[i] = (float)mj->arr[i].number(local[i]); }
        else {
            float t[3] = {0, 0, 0}, q[4] = {0, 0, 0, 1}, s[3] = {1, 1, 1};
            readVec(n, "translation", t, 3); readVec(n, "rotation", q, 4); readVec(n, "scale", s, 3);
            matFromTRS(t, q, s, local);
        }
        local[12] *= globalScale; local[13] *= globalScale; local[14] *= globalScale;      // aiProcess_GlobalScale on every node translation (ScaleProcess::applyScaling)
        float world[16]; matMul(cur.parent, local, world);
        const int mi = jint(n, "mesh", -1);
        if (mi >= 0 && mi < (int)meshStart.size())
            for (int k = 0; k < meshCount[mi]; ++k) std::memcpy(scene.meshes[meshStart[mi] + k].xf, world, 64);   // a mesh used by several nodes keeps the LAST node's transform (m_submeshes is indexed by mesh, :80-86)
        if (const Json* e = n->get("extensions")) if (const Json* kl = e->get("KHR_lights_punctual")) {
            const Json* ld = lightDefs ? lightDefs->at((size_t)jint(kl, "light", -1)) : nullptr;
            if (ld) {
                ImpLight L; L.name = jstr(n, "name");
                const std::string type = jstr(ld, "type");
                L.type = type == "directional" ? 0 : type == "point" ? 1 : 2;
                float col[3] = {1, 1, 1}; readVec(ld, "color", col, 3);
                const float inten = (float)jnum(ld, "intensity", 1.0);
                float ci[3] = {col[0] * inten, col[1] * inten, col[2] * inten};                                    // Assimp folds the intensity into the colour ...
                L.intensity = std::sqrt(dot3(ci, ci));                                                            // ... and ProcessLight splits it again by the vector length (:541-545)
                for (int k = 0; k < 3; ++k) L.color[k] = ci[k] / L.intensity;
                const float zero[3] = {0, 0, 0}; matPoint(world, zero, L.position);                               // mPosition = 0 in the light's own node
                if (L.type != 1) lightAngles(world, L.theta, L.phi);
                if (L.type == 2) { const Json* sp = ld->get("spot"); L.inner = (float)jnum(sp, "innerConeAngle", 0.0); L.outer = (float)jnum(sp, "outerConeAngle", 0.78539816339); }
                scene.lights.push_back(L);
            }
        }
        if (const Json* ch = n->get("children")) for (auto it = ch->arr.rbegin(); it != ch->arr.rend(); ++it) { Visit v; v.node = (int)it->number(-1); std::memcpy(v.parent, world, 64); stack.push_back(v); }
    }
    for (const ImpMesh& m : scene.meshes) growSceneAABB(scene, m);
    if (scene.meshes.empty()) { g_importErr = "no triangle primitives in " + path; return false; }
    return true;
}

bool endsWith(const std::string& s, const char* suf) {
    const size_t n = std::strlen(suf); if (s.size() < n) return false;
    for (size_t i = 0; i < n; ++i) if (std::tolower((unsigned char)s[s.size() - n + i]) != suf[i]) return false;
    return true;
}

} // namespace

// ---------------------------------------------------------------------------------------
// Public entry points (include/heatray_host.h)
// ---------------------------------------------------------------------------------------
extern "C" {

HRC_API const char* hrhImportLastError(void) { return g_importErr.c_str(); }
HRC_API void hrhLightAnglesFromTransform(const float worldFromLight[16], float* theta, float* phi) { lightAngles(worldFromLight, *theta, *phi); }

// ↔ HeatrayRenderer::SceneTransform::transform() (HeatrayRenderer.h:343-346): scale(scale) * yawPitchRoll(yaw, pitch, roll), column-major.
HRC_API void hrhSceneTransformMatrix(float yaw, float pitch, float roll, float scale, float out16[16]) {
    const float ch = std::cos(yaw), sh = std::sin(yaw), cp = std::cos(pitch), sp = std::sin(pitch), cb = std::cos(roll), sb = std::sin(roll);
    const float R[16] = {ch * cb + sh * sp * sb, sb * cp, -sh * cb + ch * sp * sb, 0.0f,
                         -ch * sb + sh * sp * cb, cb * cp, sb * sh + ch * sp * cb, 0.0f,
                         sh * cp, -sp, ch * cp, 0.0f,
                         0.0f, 0.0f, 0.0f, 1.0f};
    float S[16]; matIdentity(S); S[0] = S[5] = S[10] = scale;
    matMul(S, R, out16);
}
// ↔ util::AABB::center() / radius() / valid() (AABB.h:34-52). Quirks kept: only the two extreme corners go through the transform, radius() is the
// full diagonal, and valid() is true when ANY axis has min < max.
HRC_API int hrhAABBCamera(const float aabbMin[3], const float aabbMax[3], const float transform16[16], float center[3], float* radius) {
    float I[16]; matIdentity(I);
    const float* m = transform16 ? transform16 : I;
    float a[3], b[3];
    for (int r = 0; r < 3; ++r) {          // glm's mat4 * vec4: (m0 x + m1 y) + (m2 z + m3 w)
        a[r] = (m[r] * aabbMin[0] + m[4 + r] * aabbMin[1]) + (m[8 + r] * aabbMin[2] + m[12 + r] * 1.0f);
        b[r] = (m[r] * aabbMax[0] + m[4 + r] * aabbMax[1]) + (m[8 + r] * aabbMax[2] + m[12 + r] * 1.0f);
    }
    float d[3];
    for (int r = 0; r < 3; ++r) { if (center) center[r] = (a[r] + b[r]) * 0.5f; d[r] = b[r] - a[r]; }
    if (radius) *radius = std::sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
    return (aabbMin[0] < aabbMax[0] || aabbMin[1] < aabbMax[1] || aabbMin[2] < aabbMax[2]) ? 1 : 0;
}
// ↔ util::AABB::bottom() (AABB.h:46-51): center().y - |transformedMax.y - transformedMin.y| / 2.
HRC_API float hrhAABBBottom(const float aabbMin[3], const float aabbMax[3], const float transform16[16]) {
    float I[16]; matIdentity(I);
    const float* m = transform16 ? transform16 : I;
    const float ay = (m[1] * aabbMin[0] + m[5] * aabbMin[1]) + (m[9] * aabbMin[2] + m[13] * 1.0f);
    const float by = (m[1] * aabbMax[0] + m[5] * aabbMax[1]) + (m[9] * aabbMax[2] + m[13] * 1.0f);
    return (ay + by) * 0.5f - std::fabs(by - ay) * 0.5f;
}
// ↔ the viewer's "Add Ground Plane" (HeatrayRenderer.cpp:1167-1193): a square plane of side max(1, size_t(radius)) * 5 at the bottom of the
// scene AABB, PBR metallic 0 / roughness 0.9 / base colour 0.9 / specularF0 0.2. Returns the new mesh's index, or -1.
HRC_API int hrhSceneAddGroundPlane(HrhScene scene, const float aabbMin[3], const float aabbMax[3], const float transform16[16]) {
    if (!scene || !aabbMin || !aabbMax) return -1;
    float radius = 0.0f;
    hrhAABBCamera(aabbMin, aabbMax, transform16, nullptr, &radius);
    const size_t side = std::max<size_t>(1, (size_t)radius) * 5;               // the reference truncates the radius to an integer first
    ImpMaterial mat; mat.name = "Ground Plane"; mat.type = HRC_MATERIAL_PBR; mat.metallic = 0.0f; mat.roughness = 0.9f; mat.specularF0 = 0.2f;
    mat.baseColor[0] = mat.baseColor[1] = mat.baseColor[2] = 0.9f;
    scene->materials.push_back(mat);
    ImpMesh plane; plane.name = "Ground Plane"; plane.material = (int)scene->materials.size() - 1;
    plane.pos.resize(12); plane.nrm.resize(12); plane.uv.resize(8);
    uint32_t strip[4];
    hrhPlaneMesh((float)side, (float)side, plane.pos.data(), plane.nrm.data(), plane.uv.data(), strip);
    plane.idx = {strip[0], strip[1], strip[2], strip[2], strip[1], strip[3]};
    plane.xf[13] = hrhAABBBottom(aabbMin, aabbMax, transform16);
    scene->meshes.push_back(std::move(plane));
    return (int)scene->meshes.size() - 1;
}
// ↔ Scene::applyTransform (Scene.cpp:38-49): every mesh is uploaded with transform * its own worldFromEntity; winding stays the mesh's own (Mesh.cpp:86-91).
HRC_API void hrhSceneApplyTransform(HrhScene scene, const float transform16[16]) {
    if (!scene) return;
    if (transform16) std::memcpy(scene->sceneXf, transform16, 64); else matIdentity(scene->sceneXf);
}

// ↔ the built-in scenes of HeatrayRenderer::changeScene (HeatrayRenderer.cpp:123-286): "Editable PBR Material", "Editable Glass Material",
// "Multi-Material" (the default session's scene) and "Sphere Array", built from the same mesh providers (PlaneMeshProvider 15 x 15,
// SphereMeshProvider 50 x 50) with the same material parameters and transforms. forceEnableAllTextures of the two editable scenes only
// switches shader variants on for the material editor; with no texture bound the shading is the same, so it is not carried over.
// The reference leaves the camera alone for these scenes (no AABB framing): HrhSceneInfo reports an empty AABB (min == max == 0).
HRC_API int hrhBuildPresetScene(const char* name, HrhScene* out) {
    if (!name || !out) { g_importErr = "null argument"; return HRC_ERR_INVALID_VALUE; }
    *out = nullptr;
    std::unique_ptr<HrhScene_t> s(new HrhScene_t());
    auto sphere = [&](float radius, const char* meshName, int material, float tx, float ty, float tz) {
        ImpMesh m; m.name = meshName; m.material = material;
        uint32_t nv = 0, ni = 0;
        hrhSphereMesh(50, 50, radius, nullptr, nullptr, nullptr, nullptr, &nv, &ni);
        m.pos.resize(3 * (size_t)nv); m.nrm.resize(3 * (size_t)nv); m.uv.resize(2 * (size_t)nv); m.idx.resize(ni);
        hrhSphereMesh(50, 50, radius, m.pos.data(), m.nrm.data(), m.uv.data(), m.idx.data(), &nv, &ni);
        m.xf[12] = tx; m.xf[13] = ty; m.xf[14] = tz;                     // glm::translate(mat4(1), t)
        s->meshes.push_back(std::move(m));
    };
    auto pbr = [&](const char* matName, float metallic, float roughness, float base, float f0) {
        ImpMaterial m; m.name = matName; m.type = HRC_MATERIAL_PBR; m.metallic = metallic; m.roughness = roughness;
        m.baseColor[0] = m.baseColor[1] = m.baseColor[2] = base; m.specularF0 = f0;       // the other parameters keep PhysicallyBasedMaterial::Parameters' defaults
        s->materials.push_back(m); return (int)s->materials.size() - 1;
    };
    auto glass = [&](const char* matName, float r, float g, float b, float ior, float roughness, float density) {
        ImpMaterial m; m.name = matName; m.type = HRC_MATERIAL_GLASS; m.baseColor[0] = r; m.baseColor[1] = g; m.baseColor[2] = b;
        m.ior = ior; m.roughness = roughness; m.density = density;
        s->materials.push_back(m); return (int)s->materials.size() - 1;
    };
    const std::string n = name;
    if (n == "Editable PBR Material") sphere(1.0f, "PBR Sphere", pbr("PBR", 0.0f, 1.0f, 0.8f, 0.0f), 0, 0, 0);               // :123-138
    else if (n == "Editable Glass Material") sphere(1.0f, "Glass Sphere", glass("Glass", 0.8f, 0.8f, 0.8f, 1.33f, 0.0f, 0.8f), 0, 0, 0);   // :139-152
    else if (n == "Multi-Material") {                                                                                          // :153-240
        ImpMesh plane; plane.name = "Plane"; plane.material = pbr("Ground", 0.0f, 1.0f, 0.9f, 0.0f);
        plane.pos.resize(12); plane.nrm.resize(12); plane.uv.resize(8);
        uint32_t strip[4];
        hrhPlaneMesh(15.0f, 15.0f, plane.pos.data(), plane.nrm.data(), plane.uv.data(), strip);
        plane.idx = {strip[0], strip[1], strip[2], strip[2], strip[1], strip[3]};          // the provider's 4-index strip as a triangle list (odd triangle swapped)
        plane.xf[13] = -1.5f;
        s->meshes.push_back(std::move(plane));
        sphere(1.0f, "Sphere", pbr("PBR", 1.0f, 0.1f, 0.4f, 0.3f), -0.9f, -0.5f, -0.8f);
        sphere(1.0f, "Sphere", glass("Glass", 0.9f, 0.6f, 0.6f, 1.57f, 0.1f, 0.5f), 1.2f, -0.5f, 0.8f);
    } else if (n == "Sphere Array") {                                                                                          // :241-285
        const float radius = 0.5f, padding = radius * 0.2f;
        for (int row = 0; row < 2; ++row) {
            float roughness = 0.0f;                                                          // accumulated in fp32 exactly like the reference's loop
            float startX = (-5.0f * (radius * 2.0f + padding)) + ((radius * 2.0f + padding) * 0.5f);
            for (int i = 0; i < 10; ++i) {
                char label[96]; std::snprintf(label, sizeof(label), row ? "Sphere conductor roughness %f" : "Sphere dialectric roughness %f", roughness);
                sphere(radius, "Sphere", pbr(label, row ? 1.0f : 0.0f, roughness, 1.0f, 0.0f), startX, row ? 1.5f : 0.0f, 0.0f);
                roughness += 0.1f; startX += radius * 2.0f + padding;
            }
        }
    } else { g_importErr = "unknown built-in scene (Editable PBR Material, Editable Glass Material, Multi-Material, Sphere Array)"; return HRC_ERR_INVALID_VALUE; }
    *out = s.release();
    return HRC_OK;
}

// ↔ AssimpMeshProvider::AssimpMeshProvider / LoadScene (AssimpMeshProvider.cpp:65-75, 662-723); Scene::loadFromDisk (Scene.cpp:14-27)
HRC_API int hrhImportScene(const char* path, int32_t convertToMeters, HrhScene* out) {
    if (!path || !out) { g_importErr = "null argument"; return HRC_ERR_INVALID_VALUE; }
    *out = nullptr;
    std::unique_ptr<HrhScene_t> s(new HrhScene_t());
    const std::string p = path; s->dir = dirOf(p);
    const float scale = convertToMeters ? 0.01f : 1.0f;                          // :693-695: the asset is assumed to be in centimetres
    bool ok;
    try {                                                                        // no exception may cross the C boundary
        if (endsWith(p, ".obj")) ok = importObj(p, scale, *s);
        else if (endsWith(p, ".gltf") || endsWith(p, ".glb")) ok = importGltf(p, scale, *s);
        else { g_importErr = "unsupported scene format (OBJ and glTF 2.0 are read here)"; return HRC_ERR_UNSUPPORTED; }
    } catch (const std::bad_alloc&) { g_importErr = "out of memory while importing the scene"; return HRC_ERR_OUT_OF_MEMORY; }
      catch (const std::exception& e) { g_importErr = std::string("malformed scene file: ") + e.what(); return HRC_ERR_INVALID_VALUE; }
    if (!ok) return HRC_ERR_INVALID_VALUE;
    *out = s.release();
    return HRC_OK;
}
HRC_API void hrhFreeScene(HrhScene s) { delete s; }

HRC_API int hrhSceneGetInfo(HrhScene s, HrhSceneInfo* info) {
    if (!s || !info) return HRC_ERR_INVALID_VALUE;
    info->materialCount = (int32_t)s->materials.size(); info->meshCount = (int32_t)s->meshes.size(); info->lightCount = (int32_t)s->lights.size();
    std::memcpy(info->aabbMin, s->lo, 12); std::memcpy(info->aabbMax, s->hi, 12);
    return HRC_OK;
}

HRC_API int hrhSceneGetMaterial(HrhScene s, int32_t i, HrhSceneMaterial* out) {
    if (!s || !out || i < 0 || i >= (int)s->materials.size()) return HRC_ERR_INVALID_VALUE;
    const ImpMaterial& m = s->materials[i];
    std::memset(out, 0, sizeof(*out));
    out->type = m.type; out->name = m.name.c_str();
    for (int k = 0; k < kSlotCount; ++k) out->texturePath[k] = m.tex[k].empty() ? nullptr : m.tex[k].c_str();
    HrhPbrMaterial& p = out->pbr;
    std::memcpy(p.baseColor, m.baseColor, 12); std::memcpy(p.emissiveColor, m.emissive, 12);
    p.metallic = m.metallic; p.roughness = m.roughness; p.specularF0 = m.specularF0; p.clearCoat = m.clearCoat; p.clearCoatRoughness = m.clearCoatRoughness;
    p.doubleSided = m.doubleSided; p.alphaMask = m.alphaMask; p.enableVertexColors = m.vertexColors;
    p.baseColorTexture = p.metallicRoughnessTexture = p.emissiveTexture = p.normalmap = p.clearCoatTexture = p.clearCoatRoughnessTexture = p.clearCoatNormalmap = HRC_NULL_TEXTURE;
    HrhGlassMaterial& g = out->glass;
    std::memcpy(g.baseColor, m.baseColor, 12); g.roughness = m.type == HRC_MATERIAL_GLASS ? m.roughness : 0.0f; g.ior = m.ior; g.density = m.density;
    g.baseColorTexture = g.normalmap = g.metallicRoughnessTexture = HRC_NULL_TEXTURE;
    return HRC_OK;
}

HRC_API int hrhSceneGetMesh(HrhScene s, int32_t i, HrhSceneMesh* out) {
    if (!s || !out || i < 0 || i >= (int)s->meshes.size()) return HRC_ERR_INVALID_VALUE;
    const ImpMesh& m = s->meshes[i];
    std::memset(out, 0, sizeof(*out));
    out->name = m.name.c_str(); out->materialIndex = m.material;
    out->vertexCount = (uint32_t)(m.pos.size() / 3); out->indexCount = (uint32_t)m.idx.size();
    out->positions = m.pos.data(); out->normals = m.nrm.data();
    out->texCoords = m.uv.empty() ? nullptr : m.uv.data(); out->tangents = m.tan.empty() ? nullptr : m.tan.data();
    out->bitangents = m.bit.empty() ? nullptr : m.bit.data(); out->colors = m.col.empty() ? nullptr : m.col.data();
    out->indices = m.idx.data();
    std::memcpy(out->localTransform, m.xf, 64);
    return HRC_OK;
}

HRC_API int hrhSceneGetLight(HrhScene s, int32_t i, HrhSceneLight* out) {
    if (!s || !out || i < 0 || i >= (int)s->lights.size()) return HRC_ERR_INVALID_VALUE;
    const ImpLight& l = s->lights[i];
    std::memset(out, 0, sizeof(*out));
    out->name = l.name.c_str(); out->type = l.type; std::memcpy(out->color, l.color, 12); out->intensity = l.intensity;
    std::memcpy(out->position, l.position, 12); out->theta = l.theta; out->phi = l.phi; out->innerAngle = l.inner; out->outerAngle = l.outer;
    return HRC_OK;
}

} // extern "C"

// ---------------------------------------------------------------------------------------
// Upload of an imported scene into a backend context: what Scene::loadFromDisk (Scene.cpp:14-27) triggers in the reference —
// texture loads (AssimpMeshProvider.cpp:296-357, 433-511 through util::loadTexture), Material::build/modify, Mesh::Mesh
// (Mesh.cpp:55-155: one primitive per submesh, front face from the transform determinant, PBR + alpha mask = non-occluder),
// Lighting::clearAllButEnvironment + the imported punctual lights (Scene.cpp:16, ProcessLight).
// ---------------------------------------------------------------------------------------
namespace {

float det3(const float* m) {   // upper-left 3x3 of a column-major 4x4
    return m[0] * (m[5] * m[10] - m[9] * m[6]) - m[4] * (m[1] * m[10] - m[9] * m[2]) + m[8] * (m[1] * m[6] - m[5] * m[2]);
}

} // namespace

extern "C" HRC_API int hrhSceneUpload(HrhScene s, HrcContext ctx, int32_t clearOldScene, HrcTexture multiscatterLUT, HrhSceneUploadResult* result) {
    if (!s || !ctx) { g_importErr = "null argument"; return HRC_ERR_INVALID_VALUE; }
    HrhSceneUploadResult res{}; int rc;
    auto fail = [&](int code, const char* what) { g_importErr = std::string(what) + ": " + (hrcGetLastError(ctx) ? hrcGetLastError(ctx) : ""); if (result) *result = res; return code; };
    if (clearOldScene && (rc = hrcClearScene(ctx)) != HRC_OK) return fail(rc, "hrcClearScene");
    if (multiscatterLUT == HRC_NULL_TEXTURE) {                                    // loadMultiscatterTexture (MultiScatterUtil.cpp:141-150): LINEAR, CLAMP_TO_EDGE
        std::vector<float> lut(128 * 128);
        hr::generateMultiscatterLUT(lut.data(), 128, 4096, 0);
        HrcTextureDesc td{128, 128, HRC_TEX_R32F, HRC_WRAP_CLAMP, HRC_WRAP_CLAMP};
        if ((rc = hrcCreateTexture2D(ctx, &td, lut.data(), &multiscatterLUT)) != HRC_OK) return fail(rc, "multiscatter LUT");
    }
    res.multiscatterLUT = multiscatterLUT;

    std::map<std::pair<std::string, bool>, HrcTexture> cache;
    auto texture = [&](const std::string& path, bool linear) -> HrcTexture {
        if (path.empty()) return HRC_NULL_TEXTURE;
        const auto key = std::make_pair(path, linear);
        auto it = cache.find(key);
        if (it != cache.end()) return it->second;
        HrcTexture handle = HRC_NULL_TEXTURE;
        HrhLoadedTexture lt;
        if (hrhLoadTexture(path.c_str(), 1, linear ? 1 : 0, &lt) == HRC_OK) {     // generateMips = true at every call site (:314, :449 ...)
            HrcTextureDesc td{lt.image.width, lt.image.height, lt.format, lt.wrapS, lt.wrapT};
            int trc;
            if (lt.image.isFloat) trc = hrcCreateTexture2D(ctx, &td, lt.image.data, &handle);
            else {
                std::vector<uint8_t> rgba((size_t)lt.image.width * lt.image.height * 4);
                hrhExpandToRGBA8(&lt.image, rgba.data());
                trc = hrcCreateTexture2D(ctx, &td, rgba.data(), &handle);
            }
            hrhFreeImage(&lt.image);
            if (trc != HRC_OK) handle = HRC_NULL_TEXTURE; else ++res.textureCount;
        } else ++res.texturesMissing;                                              // the reference logs the error and renders without the map (TextureLoader.cpp:29-33)
        cache[key] = handle;
        return handle;
    };

    std::vector<HrcMaterial> handles;
    for (const ImpMaterial& m : s->materials) {
        HrcMaterial h = -1;
        if (m.type == HRC_MATERIAL_GLASS) {
            HrhGlassMaterial g{}; std::memcpy(g.baseColor, m.baseColor, 12); g.roughness = m.roughness; g.ior = m.ior; g.density = m.density;
            g.baseColorTexture = texture(m.tex[kSlotBaseColor], true); g.normalmap = texture(m.tex[kSlotNormal], false);
            g.metallicRoughnessTexture = texture(m.tex[kSlotMetallicRoughness], false);
            HrcGlassParams p{}; const uint32_t flags = hr::packGlassMaterial(g, p);
            rc = hrcCreateMaterial(ctx, HRC_MATERIAL_GLASS, flags, &p, sizeof(p), &h);
        } else {
            HrhPbrMaterial b{}; std::memcpy(b.baseColor, m.baseColor, 12); std::memcpy(b.emissiveColor, m.emissive, 12);
            b.metallic = m.metallic; b.roughness = m.roughness; b.specularF0 = m.specularF0; b.clearCoat = m.clearCoat; b.clearCoatRoughness = m.clearCoatRoughness;
            b.doubleSided = m.doubleSided; b.alphaMask = m.alphaMask; b.enableVertexColors = m.vertexColors;
            b.baseColorTexture = texture(m.tex[kSlotBaseColor], true); b.metallicRoughnessTexture = texture(m.tex[kSlotMetallicRoughness], false);
            b.emissiveTexture = texture(m.tex[kSlotEmissive], true); b.normalmap = texture(m.tex[kSlotNormal], false);
            b.clearCoatTexture = texture(m.tex[kSlotClearCoat], false); b.clearCoatRoughnessTexture = texture(m.tex[kSlotClearCoatRoughness], false);
            b.clearCoatNormalmap = texture(m.tex[kSlotClearCoatNormal], false);
            HrcPbrParams p{}; const uint32_t flags = hr::packPbrMaterial(b, multiscatterLUT, p);
            rc = hrcCreateMaterial(ctx, HRC_MATERIAL_PBR, flags, &p, sizeof(p), &h);
        }
        if (rc != HRC_OK) return fail(rc, "hrcCreateMaterial");
        handles.push_back(h); ++res.materialCount;
    }
    for (const ImpMesh& m : s->meshes) {
        if (m.idx.empty()) continue;
        HrcMeshDesc d{};
        d.positions = m.pos.data(); d.positionStride = 12; d.normals = m.nrm.data(); d.normalStride = 12;
        if (!m.uv.empty()) { d.texCoords = m.uv.data(); d.texCoordStride = 8; }
        if (!m.tan.empty()) { d.tangents = m.tan.data(); d.tangentStride = 12; d.bitangents = m.bit.data(); d.bitangentStride = 12; }
        if (!m.col.empty()) { d.colors = m.col.data(); d.colorStride = 12; }
        d.vertexCount = (uint32_t)(m.pos.size() / 3); d.indices = m.idx.data(); d.elementCount = (uint32_t)m.idx.size(); d.drawMode = HRC_DRAW_TRIANGLES;
        matMul(s->sceneXf, m.xf, d.worldFromEntity);                                                           // Scene.cpp:44: transform * submesh.transform
        d.frontFaceCW = det3(m.xf) < 0.0f;                                                                     // Mesh.cpp:86-91
        const ImpMaterial& mat = s->materials[m.material];
        d.isOccluder = !(mat.type == HRC_MATERIAL_PBR && mat.alphaMask);                                       // Mesh.cpp:95-100
        d.material = handles[m.material];
        HrcMesh mh;
        if ((rc = hrcCreateMesh(ctx, &d, &mh)) != HRC_OK) return fail(rc, "hrcCreateMesh");
        ++res.meshCount;
    }
    if ((rc = hrcCommitScene(ctx)) != HRC_OK) return fail(rc, "hrcCommitScene");
    std::vector<HrcDirLight> dl; std::vector<HrcPointLight> pl; std::vector<HrcSpotLight> sl;
    for (const ImpLight& l : s->lights) {                                                                      // at most 5 per type (ShaderLightingDefines.h:18-20)
        if (l.type == 0 && dl.size() < HRC_MAX_LIGHTS_PER_TYPE) { HrcDirLight o; hrhPackDirectionalLight(l.color, l.intensity, l.theta, l.phi, &o); dl.push_back(o); }
        else if (l.type == 1 && pl.size() < HRC_MAX_LIGHTS_PER_TYPE) { HrcPointLight o; hrhPackPointLight(l.color, l.intensity, l.position, &o); pl.push_back(o); }
        else if (l.type == 2 && sl.size() < HRC_MAX_LIGHTS_PER_TYPE) { HrcSpotLight o; hrhPackSpotLight(l.color, l.intensity, l.position, l.theta, l.phi, l.inner, l.outer, &o); sl.push_back(o); }
    }
    if ((rc = hrcSetDirectionalLights(ctx, (int32_t)dl.size(), dl.data())) != HRC_OK) return fail(rc, "hrcSetDirectionalLights");
    if ((rc = hrcSetPointLights(ctx, (int32_t)pl.size(), pl.data())) != HRC_OK) return fail(rc, "hrcSetPointLights");
    if ((rc = hrcSetSpotLights(ctx, (int32_t)sl.size(), sl.data())) != HRC_OK) return fail(rc, "hrcSetSpotLights");
    res.dirLights = (int32_t)dl.size(); res.pointLights = (int32_t)pl.size(); res.spotLights = (int32_t)sl.size();
    if (result) *result = res;
    return HRC_OK;
}
