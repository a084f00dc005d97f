// openrl_shim.cpp — libheatray_openrl.so: the OpenRL C entry points Heatray calls (3rdParty/OpenRL/rl.h:383-528, OpenRL.h:47-60),
// implemented as a state tracker over the CUDA backend's C-ABI (include/heatray_cuda.h). See include/heatray_openrl.h for the mapping.
// The reference's RLWrapper classes, PassGenerator, Scene, Mesh, Lighting, materials and lights compile unmodified against rl.h and
// link against this library (the test suite builds exactly that). No RLSL is compiled: a program is recognised by
// the RL_SHADER_NAME the reference gives its shaders and by the `#define` prefix string it passes as the first source string.
#include "../../include/heatray_openrl.h"
#include "../../include/heatray_cuda.h"

#include <dlfcn.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <map>
#include <memory>
#include <set>
#include <string>
#include <vector>

#define HRL_API extern "C" __attribute__((visibility("default")))

namespace {

// ---- backend binding -------------------------------------------------------------------------------------------------------------------
struct Backend {
    void* handle = nullptr; std::string prefix = "hrc";
    int (*CreateContext)(const HrcContextDesc*, HrcContext*) = nullptr;
    int (*SetOption)(HrcContext, int32_t, int32_t) = nullptr;          // optional: backends without options (the CPU oracle has few) still bind
    int (*DestroyContext)(HrcContext) = nullptr;
    const char* (*GetLastError)(HrcContext) = nullptr;
    int (*Resize)(HrcContext, int32_t, int32_t) = nullptr;
    int (*Clear)(HrcContext) = nullptr;
    int (*ReadPixels)(HrcContext, const float**) = nullptr;
    int (*ReleasePixels)(HrcContext) = nullptr;
    int (*CreateTexture2D)(HrcContext, const HrcTextureDesc*, const void*, HrcTexture*) = nullptr;
    int (*DestroyTexture)(HrcContext, HrcTexture) = nullptr;
    int (*CreateMaterial)(HrcContext, int32_t, uint32_t, const void*, size_t, HrcMaterial*) = nullptr;
    int (*UpdateMaterial)(HrcContext, HrcMaterial, const void*, size_t) = nullptr;
    int (*CreateMesh)(HrcContext, const HrcMeshDesc*, HrcMesh*) = nullptr;
    int (*DestroyMesh)(HrcContext, HrcMesh) = nullptr;
    int (*SetMeshTransform)(HrcContext, HrcMesh, const float*) = nullptr;
    int (*ClearScene)(HrcContext) = nullptr;
    int (*CommitScene)(HrcContext) = nullptr;
    int (*SetEnvironment)(HrcContext, const HrcEnvLight*) = nullptr;
    int (*SetDirectionalLights)(HrcContext, int32_t, const HrcDirLight*) = nullptr;
    int (*SetPointLights)(HrcContext, int32_t, const HrcPointLight*) = nullptr;
    int (*SetSpotLights)(HrcContext, int32_t, const HrcSpotLight*) = nullptr;
    int (*SetSequences)(HrcContext, int32_t, int32_t, const float*, const float*) = nullptr;
    int (*SetPixelTables)(HrcContext, const float*, size_t, const uint8_t*) = nullptr;
    int (*SetInteractiveBlocks)(HrcContext, int32_t, int32_t, const float*, int32_t, int32_t, const float*) = nullptr;
    int (*SetGlobals)(HrcContext, const HrcGlobals*) = nullptr;
    int (*RenderPass)(HrcContext, const HrcFrameParams*, HrcPassStats*) = nullptr;
    bool ok = false;
};
struct HostLib {
    void* handle = nullptr;
    size_t (*SequenceOffsetCount)(int32_t, int32_t) = nullptr;
    void (*GenerateSequenceOffsets)(int32_t, int32_t, float*, size_t) = nullptr;
    void (*GenerateSequenceIds)(int32_t, int32_t, int32_t, uint8_t*) = nullptr;
    void (*InteractiveBlockCount)(int32_t, int32_t, int32_t, int32_t, int32_t*, int32_t*) = nullptr;
    void (*GenerateInteractiveBlockRand)(int32_t, int32_t, float*) = nullptr;
    bool ok = false;
};
Backend g_backend; HostLib g_host;
std::string g_errorText;
std::map<std::string, uint64_t> g_calls;

template <class F> bool sym(void* h, const std::string& name, F& out) { out = reinterpret_cast<F>(dlsym(h, name.c_str())); return out != nullptr; }

bool resolveBackend(void* h, const std::string& p) {
    Backend b; b.handle = h; b.prefix = p;
    bool ok = sym(h, p + "CreateContext", b.CreateContext) & sym(h, p + "DestroyContext", b.DestroyContext) & sym(h, p + "GetLastError", b.GetLastError) &
              sym(h, p + "Resize", b.Resize) & sym(h, p + "Clear", b.Clear) & sym(h, p + "ReadPixels", b.ReadPixels) & sym(h, p + "ReleasePixels", b.ReleasePixels) &
              sym(h, p + "CreateTexture2D", b.CreateTexture2D) & sym(h, p + "DestroyTexture", b.DestroyTexture) & sym(h, p + "CreateMaterial", b.CreateMaterial) &
              sym(h, p + "UpdateMaterial", b.UpdateMaterial) & sym(h, p + "CreateMesh", b.CreateMesh) & sym(h, p + "DestroyMesh", b.DestroyMesh) &
              sym(h, p + "SetMeshTransform", b.SetMeshTransform) & sym(h, p + "ClearScene", b.ClearScene) & sym(h, p + "CommitScene", b.CommitScene) &
              sym(h, p + "SetEnvironment", b.SetEnvironment) & sym(h, p + "SetDirectionalLights", b.SetDirectionalLights) & sym(h, p + "SetPointLights", b.SetPointLights) &
              sym(h, p + "SetSpotLights", b.SetSpotLights) & sym(h, p + "SetSequences", b.SetSequences) & sym(h, p + "SetPixelTables", b.SetPixelTables) &
              sym(h, p + "SetInteractiveBlocks", b.SetInteractiveBlocks) & sym(h, p + "SetGlobals", b.SetGlobals) & sym(h, p + "RenderPass", b.RenderPass);
    if (!ok) return false;
    sym(h, p + "SetOption", b.SetOption);
    b.ok = true; g_backend = b;
    return true;
}
bool resolveHost(void* h) {
    HostLib l; l.handle = h;
    bool ok = sym(h, "hrhSequenceOffsetCount", l.SequenceOffsetCount) & sym(h, "hrhGenerateSequenceOffsets", l.GenerateSequenceOffsets) &
              sym(h, "hrhGenerateSequenceIds", l.GenerateSequenceIds) & sym(h, "hrhInteractiveBlockCount", l.InteractiveBlockCount) &
              sym(h, "hrhGenerateInteractiveBlockRand", l.GenerateInteractiveBlockRand);
    if (!ok) return false;
    l.ok = true; g_host = l;
    return true;
}
bool ensureBackend() {
    if (!g_backend.ok) {
        Dl_info info;
        std::string dir = ".";
        if (dladdr(reinterpret_cast<void*>(&ensureBackend), &info) && info.dli_fname) { dir = info.dli_fname; const size_t s = dir.rfind('/'); dir = s == std::string::npos ? "." : dir.substr(0, s); }
        const std::string path = dir + "/libheatray_cuda.so";
        void* h = dlopen(path.c_str(), RTLD_NOW | RTLD_GLOBAL);
        if (!h || !resolveBackend(h, "hrc")) { g_errorText = "cannot load the CUDA backend " + path + (h ? " (symbols missing)" : std::string(": ") + dlerror()); return false; }
    }
    if (!g_host.ok && !resolveHost(g_backend.handle)) { g_errorText = "the backend library does not export the hrh* host helpers (hrlBindHostLibrary)"; return false; }
    return true;
}

// ---- objects behind the handles --------------------------------------------------------------------------------------------------------
struct Buf { std::vector<uint8_t> data; std::string name; uint64_t version = 1; bool mapped = false; };
struct Tex {
    int w = 0, h = 0; RLenum format = RL_RGBA, type = RL_FLOAT; std::vector<uint8_t> data;
    RLenum wrapS = RL_REPEAT, wrapT = RL_REPEAT; uint64_t version = 1, uploaded = 0; HrcTexture handle = HRC_NULL_TEXTURE;
};
struct Fbo { Tex* color0 = nullptr; };
struct Shader { RLenum type = 0; std::string name; std::vector<std::string> sources; bool compiled = false; };
enum Kind { kUnknown, kFrame, kPbr, kGlass, kEnvLight, kDirLight, kPointLight, kSpotLight };
struct Program { std::vector<Shader*> shaders; bool linked = false; Kind kind = kUnknown; uint32_t flags = 0; std::string name; };
struct Uniform { float f[16] = {0}; int i[4] = {0}; Tex* tex = nullptr; int prim = 0; };
struct Attrib { Buf* buf = nullptr; int comps = 0; size_t stride = 0, offset = 0; };
struct Draw { bool valid = false; RLenum mode = RL_TRIANGLES; size_t count = 0, offset = 0; Buf* elements = nullptr; };
struct Prim {
    int id = 0; Program* program = nullptr;
    std::map<Program*, std::map<int, Uniform>> uniforms; std::map<Program*, std::map<int, Buf*>> blocks;
    std::map<int, Attrib> attribs; bool frontCW = false, isOccluder = true;
    Draw draw; HrcMesh mesh = -1; bool geometryDirty = false, transformDirty = false; float lastXf[16] = {0};
    HrcMaterial material = -1;
};
struct MatRec { HrcMaterial handle = -1; uint64_t version = 0; };

struct Ctx {
    HrcContext hrc = nullptr; RLenum error = RL_NO_ERROR;
    HrcGlobals lastGlobals{}; bool haveGlobals = false;      // the Globals block as last forwarded (sampleIndex zeroed)
    std::set<Buf*> bufs; std::set<Tex*> texs; std::set<Fbo*> fbos; std::set<Shader*> shaders; std::set<Program*> programs;
    std::map<int, std::unique_ptr<Prim>> prims; int nextPrim = 1;
    std::map<RLenum, Buf*> boundBuf; Tex* boundTex2D = nullptr; Fbo* boundFbo = nullptr; int boundPrim = 0; bool frontCW = false;
    std::vector<std::string> names;             // interned uniform / block / attribute names: location = index
    std::map<std::pair<Buf*, Program*>, MatRec> materials;
    int w = 0, h = 0; bool sceneDirty = false; int liveMeshes = 0;
    // what was last sent to the backend (buffer versions), so unchanged tables are not uploaded again
    uint64_t seqV = 0, apertureV = 0, offsetsV = 0, globalsV = 0; Buf* seqB = nullptr; Buf* apertureB = nullptr; Buf* offsetsB = nullptr; int tablesW = 0, tablesH = 0, tablesNs = 0;
    uint64_t blockTexV = 0; Tex* blockTex = nullptr;
    std::unique_ptr<Prim> framePrim;
    Ctx() { framePrim.reset(new Prim()); }
    Prim* prim(int id) { if (id == 0) return framePrim.get(); auto it = prims.find(id); return it == prims.end() ? nullptr : it->second.get(); }
    int intern(const char* n) { for (size_t k = 0; k < names.size(); ++k) if (names[k] == n) return (int)k; names.push_back(n); return (int)names.size() - 1; }
    int find(const char* n) const { for (size_t k = 0; k < names.size(); ++k) if (names[k] == n) return (int)k; return -1; }
};
Ctx* g_ctx = nullptr;

void setError(RLenum e, const std::string& text) { if (g_ctx && g_ctx->error == RL_NO_ERROR) g_ctx->error = e; g_errorText = text; }
#define CTX_OR_RETURN(...) Ctx* c = g_ctx; if (!c) { g_errorText = "no current OpenRL context"; return __VA_ARGS__; }
bool call(const char* name, int rc) {
    g_calls[name]++;
    if (rc == HRC_OK) return true;
    const char* t = g_backend.GetLastError && g_ctx ? g_backend.GetLastError(g_ctx->hrc) : "";
    setError(rc == HRC_ERR_OUT_OF_MEMORY ? RL_OUT_OF_MEMORY : (rc == HRC_ERR_INVALID_VALUE ? RL_INVALID_VALUE : (rc == HRC_ERR_INVALID_OPERATION ? RL_INVALID_OPERATION : RL_INTERNAL_ERROR)),
             std::string(name) + ": " + (t ? t : ""));
    return false;
}
#define BK(fn, ...) call("hrc" #fn, g_backend.fn(c->hrc, __VA_ARGS__))
#define BK0(fn) call("hrc" #fn, g_backend.fn(c->hrc))

// The C++ structs the reference maps its buffers with (x86-64 layout; RLtexture = pointer, RLprimitive = 32-bit enum, glm::vec3 = 3 packed floats).
struct EnvBlock { Tex* texture; float exposureCompensation, thetaRotation; int32_t primitive; };                                                      // ShaderLightingDefines.h:33-38
struct DirBlock { int32_t n; float directions[5][3], colors[5][3]; int32_t primitives[5]; };                                                          // :40-46
struct PointBlock { int32_t n; float positions[5][3], colors[5][3]; int32_t primitives[5]; };                                                         // :48-54
struct SpotBlock { int32_t n; float positions[5][3], directions[5][3], colors[5][3], angles[5][2]; int32_t primitives[5]; };                          // :56-64
struct PbrBlock { Tex *baseColor, *metallicRoughness, *emissive, *normalmap, *clearCoat, *clearCoatRoughness, *clearCoatNormalmap, *lut;               // PhysicallyBasedMaterial.cpp:16-36
                  float baseColorV[3], emissiveV[3], metallic, roughness, specularF0, roughnessAlpha, clearCoat_, clearCoatRoughness_, clearCoatRoughnessAlpha; };
struct GlassBlock { Tex *baseColor, *normalmap, *metallicRoughness; float baseColorV[3], roughness, roughnessAlpha, ior, density, specularF0; };      // GlassMaterial.cpp:15-28

Buf* blockOf(Ctx* c, Prim* p, const char* name) {
    if (!p || !p->program) return nullptr;
    const int idx = c->find(name); if (idx < 0) return nullptr;
    auto it = p->blocks.find(p->program); if (it == p->blocks.end()) return nullptr;
    auto jt = it->second.find(idx); return jt == it->second.end() ? nullptr : jt->second;
}
const Uniform* uniformOf(Ctx* c, Prim* p, const char* name) {
    if (!p || !p->program) return nullptr;
    const int idx = c->find(name); if (idx < 0) return nullptr;
    auto it = p->uniforms.find(p->program); if (it == p->uniforms.end()) return nullptr;
    auto jt = it->second.find(idx); return jt == it->second.end() ? nullptr : &jt->second;
}

// Textures reach the backend when something that is rendered references them (material slots, environment).
HrcTexture uploadTexture(Ctx* c, Tex* t) {
    if (!t || !c->texs.count(t) || t->w <= 0 || t->h <= 0 || t->data.empty()) return HRC_NULL_TEXTURE;
    if (t->handle != HRC_NULL_TEXTURE && t->uploaded == t->version) return t->handle;
    if (t->handle != HRC_NULL_TEXTURE) { BK(DestroyTexture, t->handle); t->handle = HRC_NULL_TEXTURE; }
    HrcTextureDesc d{}; d.width = t->w; d.height = t->h;
    d.wrapS = t->wrapS == RL_REPEAT ? HRC_WRAP_REPEAT : HRC_WRAP_CLAMP; d.wrapT = t->wrapT == RL_REPEAT ? HRC_WRAP_REPEAT : HRC_WRAP_CLAMP;
    if (t->type == RL_FLOAT) d.format = t->format == RL_RGBA ? HRC_TEX_RGBA32F : (t->format == RL_RGB ? HRC_TEX_RGB32F : HRC_TEX_R32F);
    else if (t->type == RL_UNSIGNED_BYTE && t->format == RL_RGBA) d.format = HRC_TEX_RGBA8;
    else { setError(RL_INVALID_ENUM, "texture format not supported by the backend (float R / RGB / RGBA, 8-bit RGBA)"); return HRC_NULL_TEXTURE; }
    HrcTexture h = HRC_NULL_TEXTURE;
    if (!BK(CreateTexture2D, &d, t->data.data(), &h)) return HRC_NULL_TEXTURE;
    t->handle = h; t->uploaded = t->version;
    return h;
}

bool packMaterial(Ctx* c, Program* prog, Buf* block, HrcPbrParams& pbr, HrcGlassParams& glass) {
    const uint32_t f = prog->flags;
    auto slot = [&](Tex* t, uint32_t bit) { return (f & bit) ? uploadTexture(c, t) : HRC_NULL_TEXTURE; };
    if (prog->kind == kPbr) {
        if (block->data.size() < sizeof(PbrBlock)) { setError(RL_INVALID_OPERATION, "\"Material\" block of a PhysicallyBased program is smaller than its ShaderParams"); return false; }
        PbrBlock b; std::memcpy(&b, block->data.data(), sizeof(b));
        pbr = HrcPbrParams{};
        pbr.baseColorTexture = slot(b.baseColor, HRC_MAT_HAS_BASE_COLOR_TEXTURE); pbr.metallicRoughnessTexture = slot(b.metallicRoughness, HRC_MAT_HAS_METALLIC_ROUGHNESS_TEXTURE);
        pbr.emissiveTexture = slot(b.emissive, HRC_MAT_HAS_EMISSIVE_TEXTURE); pbr.normalmap = slot(b.normalmap, HRC_MAT_HAS_NORMALMAP);
        pbr.clearCoatTexture = slot(b.clearCoat, HRC_MAT_HAS_CLEARCOAT_TEXTURE); pbr.clearCoatRoughnessTexture = slot(b.clearCoatRoughness, HRC_MAT_HAS_CLEARCOAT_ROUGHNESS_TEXTURE);
        pbr.clearCoatNormalmap = slot(b.clearCoatNormalmap, HRC_MAT_HAS_CLEARCOAT_NORMALMAP);
        pbr.multiscatterLUT = uploadTexture(c, b.lut);
        for (int k = 0; k < 3; ++k) { pbr.baseColor[k] = b.baseColorV[k]; pbr.emissiveColor[k] = b.emissiveV[k]; }
        pbr.metallic = b.metallic; pbr.roughness = b.roughness; pbr.specularF0 = b.specularF0; pbr.roughnessAlpha = b.roughnessAlpha;
        pbr.clearCoat = b.clearCoat_; pbr.clearCoatRoughness = b.clearCoatRoughness_; pbr.clearCoatRoughnessAlpha = b.clearCoatRoughnessAlpha;
        return true;
    }
    if (block->data.size() < sizeof(GlassBlock)) { setError(RL_INVALID_OPERATION, "\"Material\" block of a Glass program is smaller than its ShaderParams"); return false; }
    GlassBlock b; std::memcpy(&b, block->data.data(), sizeof(b));
    glass = HrcGlassParams{};
    glass.baseColorTexture = slot(b.baseColor, HRC_MAT_HAS_BASE_COLOR_TEXTURE); glass.normalmap = slot(b.normalmap, HRC_MAT_HAS_NORMALMAP);
    glass.metallicRoughnessTexture = slot(b.metallicRoughness, HRC_MAT_HAS_METALLIC_ROUGHNESS_TEXTURE);
    for (int k = 0; k < 3; ++k) glass.baseColor[k] = b.baseColorV[k];
    glass.roughness = b.roughness; glass.roughnessAlpha = b.roughnessAlpha; glass.ior = b.ior; glass.density = b.density; glass.specularF0 = b.specularF0;
    return true;
}

// One backend material per ("Material" block buffer, program); repacked when the buffer was rewritten (Material::modify -> Buffer::modify).
HrcMaterial materialFor(Ctx* c, Program* prog, Buf* block) {
    MatRec& m = c->materials[{block, prog}];
    if (m.handle >= 0 && m.version == block->version) return m.handle;
    HrcPbrParams pbr; HrcGlassParams glass;
    if (!packMaterial(c, prog, block, pbr, glass)) return -1;
    const void* params = prog->kind == kPbr ? (const void*)&pbr : (const void*)&glass;
    const size_t bytes = prog->kind == kPbr ? sizeof(pbr) : sizeof(glass);
    if (m.handle < 0) { if (!BK(CreateMaterial, prog->kind == kPbr ? HRC_MATERIAL_PBR : HRC_MATERIAL_GLASS, prog->flags, params, bytes, &m.handle)) { m.handle = -1; return -1; } }
    else if (!BK(UpdateMaterial, m.handle, params, bytes)) return -1;
    m.version = block->version;
    return m.handle;
}

const Attrib* attribOf(Ctx* c, Prim* p, const char* name) {
    const int idx = c->find(name); if (idx < 0) return nullptr;
    auto it = p->attribs.find(idx); return it == p->attribs.end() || !it->second.buf ? nullptr : &it->second;
}

// rlDrawElements on a primitive -> a backend mesh (Mesh.cpp:69-153).
bool syncPrimitive(Ctx* c, Prim* p) {
    if (!p->draw.valid || !p->program || (p->program->kind != kPbr && p->program->kind != kGlass)) return true;
    const Uniform* xf = uniformOf(c, p, "worldFromEntity");
    float M[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
    if (xf) std::memcpy(M, xf->f, sizeof(M));
    Buf* block = blockOf(c, p, "Material");
    if (!block) { setError(RL_INVALID_OPERATION, "primitive drawn without a \"Material\" uniform block"); return false; }
    if (p->geometryDirty || p->mesh < 0) {
        const HrcMaterial mat = materialFor(c, p->program, block);
        if (mat < 0) return false;
        const Attrib* pos = attribOf(c, p, "positionAttribute"); const Attrib* nrm = attribOf(c, p, "normalAttribute");
        if (!pos || !nrm || !p->draw.elements) { setError(RL_INVALID_OPERATION, "primitive drawn without position / normal attributes or an element buffer"); return false; }
        const uint32_t f = p->program->flags;
        const bool anyTex = (f & (HRC_MAT_HAS_BASE_COLOR_TEXTURE | HRC_MAT_HAS_METALLIC_ROUGHNESS_TEXTURE | HRC_MAT_HAS_EMISSIVE_TEXTURE | HRC_MAT_HAS_NORMALMAP | HRC_MAT_HAS_CLEARCOAT_TEXTURE |
                                  HRC_MAT_HAS_CLEARCOAT_ROUGHNESS_TEXTURE | HRC_MAT_HAS_CLEARCOAT_NORMALMAP)) != 0;
        const bool tangentSpace = anyTex && (f & (HRC_MAT_HAS_NORMALMAP | HRC_MAT_HAS_CLEARCOAT_NORMALMAP));
        // attributes exist in the vertex program only under their #define (vertex.rlsl:12-27): others are not consumed
        const Attrib* uv = anyTex ? attribOf(c, p, "texCoordAttribute") : nullptr;
        const Attrib* tan = tangentSpace ? attribOf(c, p, "tangentAttribute") : nullptr; const Attrib* bit = tangentSpace ? attribOf(c, p, "bitangentAttribute") : nullptr;
        const Attrib* col = (f & HRC_MAT_VERTEX_COLORS) ? attribOf(c, p, "colorAttribute") : nullptr;
        HrcMeshDesc d{};
        auto ptr = [](const Attrib* a) { return a ? reinterpret_cast<const float*>(a->buf->data.data() + a->offset) : nullptr; };
        auto stride = [](const Attrib* a) { return a ? (int32_t)(a->stride ? a->stride : (size_t)a->comps * 4) : 0; };
        d.positions = ptr(pos); d.positionStride = stride(pos); d.normals = ptr(nrm); d.normalStride = stride(nrm);
        d.texCoords = ptr(uv); d.texCoordStride = stride(uv); d.tangents = ptr(tan); d.tangentStride = stride(tan);
        d.bitangents = ptr(bit); d.bitangentStride = stride(bit); d.colors = ptr(col); d.colorStride = stride(col);
        const size_t ps = (size_t)d.positionStride;
        d.vertexCount = pos->buf->data.size() > pos->offset ? (uint32_t)((pos->buf->data.size() - pos->offset + ps - 12) / ps) : 0;
        for (const Attrib* a : {nrm, uv, tan, bit, col}) if (a) {          // the shortest attribute stream bounds the vertex count
            const size_t st = a->stride ? a->stride : (size_t)a->comps * 4, sz = (size_t)a->comps * 4;
            const size_t n = a->buf->data.size() > a->offset ? (a->buf->data.size() - a->offset + st - sz) / st : 0;
            d.vertexCount = std::min<uint32_t>(d.vertexCount, (uint32_t)n);
        }
        if (p->draw.offset + p->draw.count * sizeof(uint32_t) > p->draw.elements->data.size()) { setError(RL_INVALID_VALUE, "rlDrawElements range exceeds the element buffer"); return false; }
        d.indices = reinterpret_cast<const uint32_t*>(p->draw.elements->data.data() + p->draw.offset);      // Submesh::indexOffset is in bytes (rlDrawElements offset)
        d.elementCount = (uint32_t)p->draw.count;
        d.drawMode = p->draw.mode == RL_TRIANGLE_STRIP ? HRC_DRAW_TRIANGLE_STRIP : HRC_DRAW_TRIANGLES;
        std::memcpy(d.worldFromEntity, M, sizeof(M));
        d.frontFaceCW = p->frontCW ? 1 : 0; d.isOccluder = p->isOccluder ? 1 : 0; d.material = mat;
        if (p->mesh >= 0) { BK(DestroyMesh, p->mesh); p->mesh = -1; c->liveMeshes--; }
        if (!BK(CreateMesh, &d, &p->mesh)) { p->mesh = -1; return false; }
        c->liveMeshes++; p->material = mat;
        std::memcpy(p->lastXf, M, sizeof(M));
        p->geometryDirty = p->transformDirty = false; c->sceneDirty = true;
    } else {
        if (materialFor(c, p->program, block) < 0) return false;                      // parameter edits (Material::modify)
        if (p->transformDirty && std::memcmp(p->lastXf, M, sizeof(M)) != 0) {          // Scene::applyTransform rewrote the uniform
            if (!BK(SetMeshTransform, p->mesh, M)) return false;
            std::memcpy(p->lastXf, M, sizeof(M)); c->sceneDirty = true;
        }
        p->transformDirty = false;
    }
    return true;
}

bool syncLights(Ctx* c) {
    Prim* fp = c->framePrim.get();
    HrcEnvLight env{}; env.texture = HRC_NULL_TEXTURE;
    if (Buf* b = blockOf(c, fp, "EnvironmentLight")) if (b->data.size() >= sizeof(EnvBlock)) {
        EnvBlock e; std::memcpy(&e, b->data.data(), sizeof(e));
        env.enabled = e.primitive != 0 ? 1 : 0; env.exposureCompensation = e.exposureCompensation; env.thetaRotation = e.thetaRotation;
        env.texture = env.enabled ? uploadTexture(c, e.texture) : HRC_NULL_TEXTURE;
    }
    if (!BK(SetEnvironment, &env)) return false;
    // A light's miss shader reads colors[lightIndex] with lightIndex a uniform of ITS primitive (directionalLight.rlsl:23); slot i of the
    // block names that primitive. The backend keys everything by slot, which is the same thing as long as primitive[i].lightIndex == i —
    // the invariant Lighting::removeLight maintains through updateLightIndex (Lighting.cpp:239-271). It is checked, not assumed.
    auto indexOk = [&](int32_t prim, int slot) {
        Prim* lp = c->prim(prim); if (!lp) return false;
        const Uniform* u = uniformOf(c, lp, "lightIndex"); return u && u->i[0] == slot;
    };
    HrcDirLight dl[5]{}; HrcPointLight pl[5]{}; HrcSpotLight sl[5]{}; int nd = 0, np = 0, ns = 0;
    if (Buf* b = blockOf(c, fp, "DirectionalLights")) if (b->data.size() >= sizeof(DirBlock)) {
        DirBlock L; std::memcpy(&L, b->data.data(), sizeof(L)); nd = std::max(0, std::min(5, L.n));
        for (int i = 0; i < nd; ++i) { std::memcpy(dl[i].direction, L.directions[i], 12); std::memcpy(dl[i].color, L.colors[i], 12);
            if (!indexOk(L.primitives[i], i)) { setError(RL_INVALID_OPERATION, "directional light primitive's lightIndex does not match its slot"); return false; } }
    }
    if (Buf* b = blockOf(c, fp, "PointLights")) if (b->data.size() >= sizeof(PointBlock)) {
        PointBlock L; std::memcpy(&L, b->data.data(), sizeof(L)); np = std::max(0, std::min(5, L.n));
        for (int i = 0; i < np; ++i) { std::memcpy(pl[i].position, L.positions[i], 12); std::memcpy(pl[i].color, L.colors[i], 12);
            if (!indexOk(L.primitives[i], i)) { setError(RL_INVALID_OPERATION, "point light primitive's lightIndex does not match its slot"); return false; } }
    }
    if (Buf* b = blockOf(c, fp, "SpotLights")) if (b->data.size() >= sizeof(SpotBlock)) {
        SpotBlock L; std::memcpy(&L, b->data.data(), sizeof(L)); ns = std::max(0, std::min(5, L.n));
        for (int i = 0; i < ns; ++i) { std::memcpy(sl[i].position, L.positions[i], 12); std::memcpy(sl[i].direction, L.directions[i], 12); std::memcpy(sl[i].color, L.colors[i], 12);
            sl[i].angles[0] = L.angles[i][0]; sl[i].angles[1] = L.angles[i][1];
            if (!indexOk(L.primitives[i], i)) { setError(RL_INVALID_OPERATION, "spot light primitive's lightIndex does not match its slot"); return false; } }
    }
    return BK(SetDirectionalLights, nd, dl) && BK(SetPointLights, np, pl) && BK(SetSpotLights, ns, sl);
}

bool syncTables(Ctx* c, int& numSequences) {
    Prim* fp = c->framePrim.get();
    Buf* seq = blockOf(c, fp, "RandomSequences"); Buf* meta = blockOf(c, fp, "RandomSequenceMetadata"); Buf* ap = blockOf(c, fp, "ApertureSamples"); Buf* off = blockOf(c, fp, "SequenceOffsets");
    if (!seq || !meta || !ap || !off || meta->data.size() < 8) { setError(RL_INVALID_OPERATION, "frame program without its RandomSequences / RandomSequenceMetadata / ApertureSamples / SequenceOffsets blocks"); return false; }
    int32_t m[2]; std::memcpy(m, meta->data.data(), 8);
    const int ns = m[0], len = m[1]; numSequences = ns;
    const size_t need = (size_t)ns * len * 8;
    if (ns <= 0 || len <= 0 || seq->data.size() < need || ap->data.size() < need) { setError(RL_INVALID_OPERATION, "sequence tables smaller than their metadata says"); return false; }
    if (seq != c->seqB || ap != c->apertureB || seq->version != c->seqV || ap->version != c->apertureV) {
        if (!BK(SetSequences, ns, len, reinterpret_cast<const float*>(seq->data.data()), reinterpret_cast<const float*>(ap->data.data()))) return false;
        c->seqB = seq; c->apertureB = ap; c->seqV = seq->version; c->apertureV = ap->version;
    }
    if (off != c->offsetsB || off->version != c->offsetsV || c->tablesW != c->w || c->tablesH != c->h || c->tablesNs != ns) {
        // SequenceOffsets holds vec2 per pixel, only .x is read (perspective.rlsl:64). The shader's index formula uses the HEIGHT as row stride
        // and pixel centres (quirk B-Q1), which on square frames runs up to h/2 entries past the w*h table the reference builds; the backend
        // takes the table continued by h more points of the same Sobol sequence (hrhGenerateSequenceOffsets). The reference's own values are
        // used for every entry it supplies; the continuation is appended when its prefix is that same sequence, else the last value repeats.
        const size_t have = off->data.size() / 8, want = g_host.SequenceOffsetCount(c->w, c->h);
        std::vector<float> x(std::max(have, want));
        const float* src = reinterpret_cast<const float*>(off->data.data());
        for (size_t i = 0; i < have; ++i) x[i] = src[2 * i];
        if (want > have) {
            std::vector<float> ours(want); g_host.GenerateSequenceOffsets(c->w, c->h, ours.data(), want);
            const bool same = have > 0 && std::memcmp(ours.data(), x.data(), std::min(have, want) * sizeof(float)) == 0;
            for (size_t i = have; i < want; ++i) x[i] = same ? ours[i] : (have ? x[have - 1] : 0.0f);
        }
        std::vector<uint8_t> ids((size_t)c->w * c->h);
        g_host.GenerateSequenceIds(c->w, c->h, ns, ids.data());        // sequenceID = floor(random(FrameCoord / FrameSize) * numSequences), the host evaluates the sin hash (quirk B-Q2)
        if (!BK(SetPixelTables, x.data(), x.size(), ids.data())) return false;
        c->offsetsB = off; c->offsetsV = off->version; c->tablesW = c->w; c->tablesH = c->h; c->tablesNs = ns;
        c->blockTexV = 0;                                              // block hashes depend on the frame size too
    }
    return true;
}

bool syncFrame(Ctx* c, HrcFrameParams& fp) {
    Prim* p = c->framePrim.get();
    if (!p->program || p->program->kind != kFrame) { setError(RL_INVALID_OPERATION, "rlRenderFrame without a frame program on the null primitive"); return false; }
    auto f1 = [&](const char* n, float d) { const Uniform* u = uniformOf(c, p, n); return u ? u->f[0] : d; };
    auto i1 = [&](const char* n, int k, int d) { const Uniform* u = uniformOf(c, p, n); return u ? u->i[k] : d; };
    fp = HrcFrameParams{};
    fp.fovTan = f1("fovTan", 1.0f); fp.aspectRatio = f1("aspectRatio", 1.0f); fp.focusDistance = f1("focusDistance", 1.0f); fp.apertureRadius = f1("apertureRadius", 0.0f);
    if (const Uniform* u = uniformOf(c, p, "viewMatrix")) std::memcpy(fp.viewMatrix, u->f, 64); else { fp.viewMatrix[0] = fp.viewMatrix[5] = fp.viewMatrix[10] = fp.viewMatrix[15] = 1.0f; }
    fp.maxSampleIndex = f1("maxSampleIndex", 1.0f);
    fp.interactiveMode = i1("interactiveMode", 0, 0);
    fp.blockSize[0] = i1("blockSize", 0, 1); fp.blockSize[1] = i1("blockSize", 1, 1);
    fp.currentBlockPixelSample[0] = i1("currentBlockPixelSample", 0, 0); fp.currentBlockPixelSample[1] = i1("currentBlockPixelSample", 1, 0);
    fp.bandCount = 1; fp.bandHeight = 1;
    const Uniform* bt = uniformOf(c, p, "interactiveBlockSamplesTexture");
    if (bt && bt->tex && c->texs.count(bt->tex) && (bt->tex != c->blockTex || bt->tex->version != c->blockTexV) && bt->tex->type == RL_FLOAT && bt->tex->format == RL_RGB) {
        Tex* t = bt->tex; int bx = 0, by = 0;
        g_host.InteractiveBlockCount(c->w, c->h, t->w, t->h, &bx, &by);
        std::vector<float> rnd((size_t)bx * by * 2);
        g_host.GenerateInteractiveBlockRand(bx, by, rnd.data());
        if (!BK(SetInteractiveBlocks, t->w, t->h, reinterpret_cast<const float*>(t->data.data()), bx, by, rnd.data())) return false;
        c->blockTex = t; c->blockTexV = t->version;
    }
    return true;
}

Kind kindFromName(const std::string& n, RLenum type) {
    if (type == RL_FRAME_SHADER) return kFrame;
    if (n == "PhysicallyBased") return kPbr;
    if (n == "Glass") return kGlass;
    if (n == "Environment Light") return kEnvLight;
    if (n == "DirectionalLight") return kDirLight;
    if (n == "PointLight") return kPointLight;
    if (n == "SpotLight") return kSpotLight;
    return kUnknown;
}
uint32_t flagsFromSource(const Shader* s) {       // the #define prefix the material builders emit (PhysicallyBasedMaterial.cpp:49-113, GlassMaterial.cpp:40-74)
    static const struct { const char* name; uint32_t bit; } table[] = {
        {"DOUBLE_SIDED", HRC_MAT_DOUBLE_SIDED}, {"ALPHA_MASK", HRC_MAT_ALPHA_MASK}, {"HAS_BASE_COLOR_TEXTURE", HRC_MAT_HAS_BASE_COLOR_TEXTURE},
        {"HAS_METALLIC_ROUGHNESS_TEXTURE", HRC_MAT_HAS_METALLIC_ROUGHNESS_TEXTURE}, {"HAS_EMISSIVE_TEXTURE", HRC_MAT_HAS_EMISSIVE_TEXTURE}, {"HAS_NORMALMAP", HRC_MAT_HAS_NORMALMAP},
        {"HAS_CLEARCOAT_TEXTURE", HRC_MAT_HAS_CLEARCOAT_TEXTURE}, {"HAS_CLEARCOAT_ROUGHNESS_TEXTURE", HRC_MAT_HAS_CLEARCOAT_ROUGHNESS_TEXTURE},
        {"HAS_CLEARCOAT_NORMALMAP", HRC_MAT_HAS_CLEARCOAT_NORMALMAP}, {"VERTEX_COLORS", HRC_MAT_VERTEX_COLORS}};
    uint32_t f = 0;
    if (s->sources.empty()) return 0;
    const std::string& prefix = s->sources[0];
    size_t pos = 0;
    while ((pos = prefix.find("#define", pos)) != std::string::npos) {
        pos += 7;
        while (pos < prefix.size() && (prefix[pos] == ' ' || prefix[pos] == '\t')) ++pos;
        size_t end = pos; while (end < prefix.size() && (std::isalnum((unsigned char)prefix[end]) || prefix[end] == '_')) ++end;
        const std::string name = prefix.substr(pos, end - pos);
        for (auto& t : table) if (name == t.name) f |= t.bit;
        pos = end;
    }
    return f;
}

Buf* boundBuffer(Ctx* c, RLenum target) { auto it = c->boundBuf.find(target); return it == c->boundBuf.end() ? nullptr : it->second; }
Uniform* uniformSlot(Ctx* c, RLint location) {
    Prim* p = c->prim(c->boundPrim);
    if (!p || !p->program || location < 0 || location >= (RLint)c->names.size()) { setError(RL_INVALID_OPERATION, "uniform set without a program on the bound primitive, or bad location"); return nullptr; }
    if (c->names[location] == "worldFromEntity") p->transformDirty = true;
    return &p->uniforms[p->program][location];
}

} // namespace

// ---- exports ----------------------------------------------------------------------------------------------------------------------------
HRL_API int hrlBindBackend(void* handle, const char* prefix) { return handle && prefix && resolveBackend(handle, prefix) ? 0 : -1; }
HRL_API int hrlBindHostLibrary(void* handle) { return handle && resolveHost(handle) ? 0 : -1; }
HRL_API const char* hrlGetLastErrorString(void) { return g_errorText.c_str(); }
HRL_API uint64_t hrlBackendCallCount(const char* name) { auto it = g_calls.find(name ? name : ""); return it == g_calls.end() ? 0 : it->second; }

HRL_API OpenRLContext OpenRLCreateContext(const OpenRLContextAttribute*, OpenRLNotify, void*) {
    if (!ensureBackend()) { std::fprintf(stderr, "[heatray_openrl] %s\n", g_errorText.c_str()); return nullptr; }
    std::unique_ptr<Ctx> c(new Ctx());
    HrcContextDesc d{}; d.device = 0; d.collectStats = 0;
    if (const char* e = std::getenv("HEATRAY_CUDA_DEVICE")) d.device = std::atoi(e);
    g_calls["hrcCreateContext"]++;
    if (g_backend.CreateContext(&d, &c->hrc) != HRC_OK) {
        const char* t = g_backend.GetLastError(nullptr); g_errorText = std::string("hrcCreateContext: ") + (t ? t : "");
        std::fprintf(stderr, "[heatray_openrl] %s\n", g_errorText.c_str());
        return nullptr;
    }
    // The reference viewer renders one pass per rlRenderFrame and maps the pixel pack buffer after each (HeatrayRenderer.cpp:366-408). With
    // HEATRAY_OPENRL_SPECULATIVE_PASSES = 1..3 the backend keeps that many passes of the same progressive render running ahead (HRC_OPT_SPECULATIVE_PASSES,
    // include/heatray_cuda.h): the unmodified caller's readback then overlaps the next passes. Off by default: the picture then differs from the
    // one-pass-at-a-time sum by fp32 addition order.
    if (const char* e = std::getenv("HEATRAY_OPENRL_SPECULATIVE_PASSES")) {
        const int n = std::atoi(e);
        if (n > 0 && g_backend.SetOption) { g_calls["hrcSetOption"]++; g_backend.SetOption(c->hrc, HRC_OPT_SPECULATIVE_PASSES, n > HRC_MAX_DEFERRED_PASSES ? HRC_MAX_DEFERRED_PASSES : n); }
    }
    return c.release();
}
HRL_API int OpenRLSetCurrentContext(OpenRLContext ctx) { g_ctx = static_cast<Ctx*>(ctx); return 1; }
HRL_API int OpenRLDestroyContext(OpenRLContext ctx) {
    Ctx* c = static_cast<Ctx*>(ctx); if (!c) return 0;
    if (c->hrc) { g_calls["hrcDestroyContext"]++; g_backend.DestroyContext(c->hrc); }
    for (auto* b : c->bufs) delete b; for (auto* t : c->texs) delete t; for (auto* f : c->fbos) delete f; for (auto* s : c->shaders) delete s; for (auto* p : c->programs) delete p;
    if (g_ctx == c) g_ctx = nullptr;
    delete c; return 1;
}
HRL_API RLenum rlGetError(void) { if (!g_ctx) return RL_INVALID_OPERATION; const RLenum e = g_ctx->error; g_ctx->error = RL_NO_ERROR; return e; }

// buffers
HRL_API void rlGenBuffers(RLsize n, RLbuffer* out) { CTX_OR_RETURN(); for (RLsize i = 0; i < n; ++i) { Buf* b = new Buf(); c->bufs.insert(b); out[i] = reinterpret_cast<RLbuffer>(b); } }
HRL_API void rlDeleteBuffers(RLsize n, const RLbuffer* in) {
    CTX_OR_RETURN();
    for (RLsize i = 0; i < n; ++i) {
        Buf* b = reinterpret_cast<Buf*>(in[i]); if (!c->bufs.count(b)) continue;
        for (auto& kv : c->boundBuf) if (kv.second == b) kv.second = nullptr;
        auto scrub = [&](Prim* p) { for (auto& pr : p->blocks) for (auto& kv : pr.second) if (kv.second == b) kv.second = nullptr; for (auto& kv : p->attribs) if (kv.second.buf == b) kv.second.buf = nullptr; if (p->draw.elements == b) p->draw.elements = nullptr; };
        scrub(c->framePrim.get()); for (auto& kv : c->prims) scrub(kv.second.get());
        for (auto it = c->materials.begin(); it != c->materials.end();) if (it->first.first == b) it = c->materials.erase(it); else ++it;      // backend materials live until hrcClearScene
        if (c->seqB == b) c->seqB = nullptr; if (c->apertureB == b) c->apertureB = nullptr; if (c->offsetsB == b) c->offsetsB = nullptr;
        c->bufs.erase(b); delete b;
    }
}
HRL_API void rlBindBuffer(RLenum target, RLbuffer buffer) { CTX_OR_RETURN(); Buf* b = reinterpret_cast<Buf*>(buffer); if (b && !c->bufs.count(b)) { setError(RL_INVALID_VALUE, "rlBindBuffer: not a buffer"); return; } c->boundBuf[target] = b; }
HRL_API void rlBufferData(RLenum target, RLsize size, const void* data, RLenum) {
    CTX_OR_RETURN(); Buf* b = boundBuffer(c, target); if (!b) { setError(RL_INVALID_OPERATION, "rlBufferData without a bound buffer"); return; }
    b->data.assign(size, 0); if (data && size) std::memcpy(b->data.data(), data, size); b->version++;
}
HRL_API void rlBufferParameterString(RLenum target, RLenum pname, const char* param) { CTX_OR_RETURN(); Buf* b = boundBuffer(c, target); if (b && pname == RL_BUFFER_NAME && param) b->name = param; }
HRL_API void* rlMapBuffer(RLenum target, RLenum) {
    CTX_OR_RETURN(nullptr); Buf* b = boundBuffer(c, target); if (!b) { setError(RL_INVALID_OPERATION, "rlMapBuffer without a bound buffer"); return nullptr; }
    if (b->mapped) { setError(RL_INVALID_OPERATION, "buffer is already mapped"); return nullptr; }
    b->mapped = true; return b->data.data();
}
HRL_API RLboolean rlUnmapBuffer(RLenum target) {
    CTX_OR_RETURN(RL_FALSE); Buf* b = boundBuffer(c, target); if (!b || !b->mapped) { setError(RL_INVALID_OPERATION, "rlUnmapBuffer: buffer is not mapped"); return RL_FALSE; }
    b->mapped = false; b->version++;
    for (auto& kv : c->prims) for (auto& a : kv.second->attribs) if (a.second.buf == b) kv.second->geometryDirty = kv.second->draw.valid;   // vertex data rewritten after the draw
    return RL_TRUE;
}
HRL_API void rlVertexAttribBuffer(RLint indx, RLint size, RLenum type, RLboolean, RLsize stride, RLsize offset) {
    CTX_OR_RETURN(); Prim* p = c->prim(c->boundPrim); Buf* b = boundBuffer(c, RL_ARRAY_BUFFER);
    if (!p || c->boundPrim == 0 || !b || type != RL_FLOAT) { setError(RL_INVALID_OPERATION, "rlVertexAttribBuffer needs a bound primitive, a bound RL_ARRAY_BUFFER and RL_FLOAT data"); return; }
    p->attribs[indx] = Attrib{b, size, stride, offset};
}

// textures
HRL_API void rlGenTextures(RLsize n, RLtexture* out) { CTX_OR_RETURN(); for (RLsize i = 0; i < n; ++i) { Tex* t = new Tex(); c->texs.insert(t); out[i] = reinterpret_cast<RLtexture>(t); } }
HRL_API void rlDeleteTextures(RLsize n, const RLtexture* in) {
    CTX_OR_RETURN();
    for (RLsize i = 0; i < n; ++i) {
        Tex* t = reinterpret_cast<Tex*>(in[i]); if (!c->texs.count(t)) continue;
        if (t->handle != HRC_NULL_TEXTURE) BK(DestroyTexture, t->handle);
        if (c->boundTex2D == t) c->boundTex2D = nullptr; if (c->blockTex == t) c->blockTex = nullptr;
        for (auto* f : c->fbos) if (f->color0 == t) f->color0 = nullptr;
        c->texs.erase(t); delete t;
    }
}
HRL_API void rlBindTexture(RLenum target, RLtexture texture) { CTX_OR_RETURN(); if (target == RL_TEXTURE_2D) c->boundTex2D = reinterpret_cast<Tex*>(texture); }
HRL_API void rlTexParameteri(RLenum target, RLenum pname, RLint param) {
    CTX_OR_RETURN(); if (target != RL_TEXTURE_2D || !c->boundTex2D) return;
    if (pname == RL_TEXTURE_WRAP_S) c->boundTex2D->wrapS = param; else if (pname == RL_TEXTURE_WRAP_T) c->boundTex2D->wrapT = param;
    c->boundTex2D->version++;     // filters: level-0 bilinear always (policy A-12); NEAREST only matters for the block table, which the backend reads as texels
}
HRL_API void rlTexImage2D(RLenum target, RLint level, RLenum, RLint width, RLint height, RLint, RLenum format, RLenum type, const void* pixels) {
    CTX_OR_RETURN(); Tex* t = c->boundTex2D; if (target != RL_TEXTURE_2D || !t) { setError(RL_INVALID_OPERATION, "rlTexImage2D without a bound 2D texture"); return; }
    if (level != 0) return;
    const int comps = format == RL_RGBA ? 4 : (format == RL_RGB ? 3 : 1);
    const size_t bytes = (size_t)width * height * comps * (type == RL_FLOAT ? 4 : 1);
    t->w = width; t->h = height; t->format = format; t->type = type; t->data.assign(bytes, 0);
    if (pixels) std::memcpy(t->data.data(), pixels, bytes);
    t->version++;
}
HRL_API void rlTexImage3D(RLenum, RLint, RLenum, RLint, RLint, RLint, RLint, RLenum, RLenum, const void*) { setError(RL_UNIMPLEMENTED, "3D textures are not used by the path"); }
HRL_API void rlGenerateMipmap(RLenum) {}     // level-0 filtering (policy A-12)
HRL_API void rlGetTexImage(RLenum target, RLint, RLenum format, RLenum type, RLvoid* pixels) {
    CTX_OR_RETURN(); Tex* t = c->boundTex2D;
    if (target != RL_TEXTURE_2D || !t || format != RL_RGBA || type != RL_FLOAT) { setError(RL_INVALID_OPERATION, "rlGetTexImage: RGBA float read of the bound 2D texture only"); return; }
    Buf* pack = boundBuffer(c, RL_PIXEL_PACK_BUFFER);
    uint8_t* dst = pack ? pack->data.data() + reinterpret_cast<uintptr_t>(pixels) : static_cast<uint8_t*>(pixels);      // with a pack buffer bound, `pixels` is an offset into it
    const size_t bytes = (size_t)t->w * t->h * 16;
    if (pack && reinterpret_cast<uintptr_t>(pixels) + bytes > pack->data.size()) { setError(RL_INVALID_OPERATION, "pixel pack buffer is smaller than the texture"); return; }
    bool isTarget = false; for (auto* f : c->fbos) if (f->color0 == t) isTarget = true;
    if (isTarget && c->w == t->w && c->h == t->h) {              // the accumulation target: PixelPackBuffer::setPixelData (PixelPackBuffer.h:39-47)
        const float* src = nullptr;
        if (!BK(ReadPixels, &src)) return;
        std::memcpy(dst, src, bytes);
        BK0(ReleasePixels);
    } else if (t->data.size() == bytes) std::memcpy(dst, t->data.data(), bytes);
    else setError(RL_INVALID_OPERATION, "rlGetTexImage: texture is not RGBA float");
    if (pack) pack->version++;
}

// framebuffers
HRL_API void rlGenFramebuffers(RLsize n, RLframebuffer* out) { CTX_OR_RETURN(); for (RLsize i = 0; i < n; ++i) { Fbo* f = new Fbo(); c->fbos.insert(f); out[i] = reinterpret_cast<RLframebuffer>(f); } }
HRL_API void rlDeleteFramebuffers(RLsize n, const RLframebuffer* in) { CTX_OR_RETURN(); for (RLsize i = 0; i < n; ++i) { Fbo* f = reinterpret_cast<Fbo*>(in[i]); if (c->fbos.erase(f)) { if (c->boundFbo == f) c->boundFbo = nullptr; delete f; } } }
HRL_API void rlBindFramebuffer(RLenum, RLframebuffer fbo) { CTX_OR_RETURN(); c->boundFbo = reinterpret_cast<Fbo*>(fbo); }
HRL_API void rlFramebufferTexture2D(RLenum, RLenum attachment, RLenum, RLtexture texture, RLint) {
    CTX_OR_RETURN(); Fbo* f = c->boundFbo;
    // Framebuffer::addAttachment (Framebuffer.h:44-49) attaches without binding first; OpenRL evidently applies it to the framebuffer being built,
    // so with nothing bound the attachment goes to the most recently generated framebuffer.
    if (!f && !c->fbos.empty()) f = *c->fbos.rbegin();
    if (!f || attachment != RL_COLOR_ATTACHMENT0) { setError(RL_INVALID_OPERATION, "rlFramebufferTexture2D: colour attachment 0 of a framebuffer"); return; }
    f->color0 = reinterpret_cast<Tex*>(texture);
}
HRL_API RLenum rlCheckFramebufferStatus(RLenum) { CTX_OR_RETURN(0); Fbo* f = c->boundFbo ? c->boundFbo : (c->fbos.empty() ? nullptr : *c->fbos.rbegin()); return f && f->color0 ? RL_FRAMEBUFFER_COMPLETE : RL_FRAMEBUFFER_INCOMPLETE_MISSING_ATTACHMENT; }
HRL_API void rlViewport(RLint, RLint, RLint, RLint) {}     // the frame size follows the colour attachment (PassGenerator.cpp:301-323 resizes both together)
HRL_API void rlClear(RLbitfield mask) { CTX_OR_RETURN(); if ((mask & RL_COLOR_BUFFER_BIT) && c->w > 0) BK0(Clear); }

// primitives
HRL_API void rlGenPrimitives(RLsize n, RLprimitive* out) { CTX_OR_RETURN(); for (RLsize i = 0; i < n; ++i) { const int id = c->nextPrim++; c->prims[id].reset(new Prim()); c->prims[id]->id = id; out[i] = static_cast<RLprimitive>(id); } }
HRL_API void rlDeletePrimitives(RLsize n, const RLprimitive* in) {
    CTX_OR_RETURN();
    for (RLsize i = 0; i < n; ++i) {
        auto it = c->prims.find((int)in[i]); if (it == c->prims.end()) continue;
        if (it->second->mesh >= 0) { BK(DestroyMesh, it->second->mesh); c->liveMeshes--; c->sceneDirty = true; }
        if (c->boundPrim == (int)in[i]) c->boundPrim = 0;
        c->prims.erase(it);
    }
    if (c->liveMeshes == 0 && !c->materials.empty()) {          // Scene::clearMeshesAndMaterials: start the backend's material list over
        bool anyDrawn = false; for (auto& kv : c->prims) anyDrawn |= kv.second->draw.valid;
        if (!anyDrawn) { BK0(ClearScene); c->materials.clear(); c->sceneDirty = true; }
    }
}
HRL_API void rlBindPrimitive(RLenum, RLprimitive p) { CTX_OR_RETURN(); if ((int)p != 0 && !c->prims.count((int)p)) { setError(RL_INVALID_VALUE, "rlBindPrimitive: not a primitive"); return; } c->boundPrim = (int)p; }
HRL_API void rlPrimitiveParameter1i(RLenum, RLenum pname, RLint param) { CTX_OR_RETURN(); Prim* p = c->prim(c->boundPrim); if (p && pname == RL_PRIMITIVE_IS_OCCLUDER) { p->isOccluder = param != RL_FALSE; p->geometryDirty = p->draw.valid; } }
HRL_API void rlFrontFace(RLenum mode) { CTX_OR_RETURN(); c->frontCW = mode == RL_CW; }
HRL_API void rlDrawElements(RLenum mode, RLsize count, RLenum type, RLsize offset) {
    CTX_OR_RETURN(); Prim* p = c->prim(c->boundPrim); Buf* el = boundBuffer(c, RL_ELEMENT_ARRAY_BUFFER);
    if (!p || c->boundPrim == 0 || !el || type != RL_UNSIGNED_INT || (mode != RL_TRIANGLES && mode != RL_TRIANGLE_STRIP)) { setError(RL_INVALID_OPERATION, "rlDrawElements needs a bound primitive, a bound element buffer, RL_UNSIGNED_INT indices and triangles / strips"); return; }
    p->draw = Draw{true, mode, count, offset, el}; p->frontCW = c->frontCW; p->geometryDirty = true;
}

// shaders and programs
HRL_API RLshader rlCreateShader(RLenum type) { CTX_OR_RETURN(nullptr); Shader* s = new Shader(); s->type = type; c->shaders.insert(s); return reinterpret_cast<RLshader>(s); }
HRL_API void rlDeleteShader(RLshader sh) { CTX_OR_RETURN(); Shader* s = reinterpret_cast<Shader*>(sh); if (c->shaders.count(s)) { for (auto* p : c->programs) p->shaders.erase(std::remove(p->shaders.begin(), p->shaders.end(), s), p->shaders.end()); c->shaders.erase(s); delete s; } }
HRL_API void rlShaderString(RLshader sh, RLenum pname, const char* param) { CTX_OR_RETURN(); Shader* s = reinterpret_cast<Shader*>(sh); if (c->shaders.count(s) && pname == RL_SHADER_NAME && param) s->name = param; }
HRL_API void rlShaderSource(RLshader sh, RLsize count, const char* const* strings, const RLsize* lengths) {
    CTX_OR_RETURN(); Shader* s = reinterpret_cast<Shader*>(sh); if (!c->shaders.count(s)) { setError(RL_INVALID_VALUE, "rlShaderSource: not a shader"); return; }
    s->sources.clear(); for (RLsize i = 0; i < count; ++i) s->sources.emplace_back(lengths ? std::string(strings[i], lengths[i]) : std::string(strings[i]));
}
HRL_API void rlCompileShader(RLshader sh) { CTX_OR_RETURN(); Shader* s = reinterpret_cast<Shader*>(sh); if (c->shaders.count(s)) s->compiled = !s->sources.empty(); }
HRL_API void rlGetShaderiv(RLshader sh, RLenum pname, RLint* out) { CTX_OR_RETURN(); Shader* s = reinterpret_cast<Shader*>(sh); if (out) *out = (c->shaders.count(s) && pname == RL_COMPILE_STATUS && s->compiled) ? RL_TRUE : RL_FALSE; }
HRL_API void rlGetShaderString(RLshader, RLenum, const char** out) { if (out) *out = "RLSL is not compiled by this backend: a shader needs at least one non-empty source string"; }
HRL_API RLprogram rlCreateProgram(void) { CTX_OR_RETURN(nullptr); Program* p = new Program(); c->programs.insert(p); return reinterpret_cast<RLprogram>(p); }
HRL_API void rlDeleteProgram(RLprogram pr) {
    CTX_OR_RETURN(); Program* p = reinterpret_cast<Program*>(pr); if (!c->programs.count(p)) return;
    auto scrub = [&](Prim* q) { if (q->program == p) q->program = nullptr; q->uniforms.erase(p); q->blocks.erase(p); };
    scrub(c->framePrim.get()); for (auto& kv : c->prims) scrub(kv.second.get());
    for (auto it = c->materials.begin(); it != c->materials.end();) if (it->first.second == p) it = c->materials.erase(it); else ++it;
    c->programs.erase(p); delete p;
}
HRL_API void rlAttachShader(RLprogram pr, RLshader sh) { CTX_OR_RETURN(); Program* p = reinterpret_cast<Program*>(pr); Shader* s = reinterpret_cast<Shader*>(sh); if (c->programs.count(p) && c->shaders.count(s)) p->shaders.push_back(s); else setError(RL_INVALID_VALUE, "rlAttachShader: bad handle"); }
HRL_API void rlLinkProgram(RLprogram pr) {
    CTX_OR_RETURN(); Program* p = reinterpret_cast<Program*>(pr); if (!c->programs.count(p)) { setError(RL_INVALID_VALUE, "rlLinkProgram: not a program"); return; }
    p->linked = false; p->kind = kUnknown; p->flags = 0;
    for (Shader* s : p->shaders) {
        if (!s->compiled) return;
        if (s->type == RL_RAY_SHADER || s->type == RL_FRAME_SHADER) { p->kind = kindFromName(s->name, s->type); p->name = s->name; if (p->kind == kPbr || p->kind == kGlass) p->flags = flagsFromSource(s); }
    }
    // Unknown ray programs link (OpenRL would run them); the backend only shades the programs it knows and refuses to draw others.
    p->linked = !p->shaders.empty();
}
HRL_API void rlGetProgramiv(RLprogram pr, RLenum pname, RLint* out) { CTX_OR_RETURN(); Program* p = reinterpret_cast<Program*>(pr); if (out) *out = (c->programs.count(p) && pname == RL_LINK_STATUS && p->linked) ? 1 : 0; }
HRL_API void rlGetProgramString(RLprogram, RLenum, const char** out) { if (out) *out = "a program links when every attached shader compiled"; }
HRL_API void rlUseProgram(RLprogram pr) {
    CTX_OR_RETURN(); Program* p = reinterpret_cast<Program*>(pr); if (p && !c->programs.count(p)) { setError(RL_INVALID_VALUE, "rlUseProgram: not a program"); return; }
    Prim* q = c->prim(c->boundPrim); if (q && p) { if (q->program != p && q->draw.valid) q->geometryDirty = true; q->program = p; }      // unbinding (null) leaves the primitive's program in place
}
HRL_API RLint rlGetUniformLocation(RLprogram, const char* name) { CTX_OR_RETURN(-1); return name ? c->intern(name) : -1; }
HRL_API RLint rlGetUniformBlockIndex(RLprogram, const char* name) { CTX_OR_RETURN(-1); return name ? c->intern(name) : -1; }
HRL_API RLint rlGetAttribLocation(RLprogram, const char* name) { CTX_OR_RETURN(-1); return name ? c->intern(name) : -1; }
HRL_API void rlUniformBlockBuffer(RLint index, RLbuffer buffer) {
    CTX_OR_RETURN(); Prim* p = c->prim(c->boundPrim); Buf* b = reinterpret_cast<Buf*>(buffer);
    if (!p || !p->program || index < 0 || (b && !c->bufs.count(b))) { setError(RL_INVALID_OPERATION, "rlUniformBlockBuffer without a program on the bound primitive"); return; }
    p->blocks[p->program][index] = b;
}
HRL_API void rlUniform1f(RLint l, RLfloat x) { CTX_OR_RETURN(); if (Uniform* u = uniformSlot(c, l)) u->f[0] = x; }
HRL_API void rlUniform1i(RLint l, RLint x) { CTX_OR_RETURN(); if (Uniform* u = uniformSlot(c, l)) u->i[0] = x; }
HRL_API void rlUniform2fv(RLint l, RLsize, const RLfloat* v) { CTX_OR_RETURN(); if (Uniform* u = uniformSlot(c, l)) std::memcpy(u->f, v, 8); }
HRL_API void rlUniform2iv(RLint l, RLsize, const RLint* v) { CTX_OR_RETURN(); if (Uniform* u = uniformSlot(c, l)) std::memcpy(u->i, v, 8); }
HRL_API void rlUniform3fv(RLint l, RLsize, const RLfloat* v) { CTX_OR_RETURN(); if (Uniform* u = uniformSlot(c, l)) std::memcpy(u->f, v, 12); }
HRL_API void rlUniform4fv(RLint l, RLsize, const RLfloat* v) { CTX_OR_RETURN(); if (Uniform* u = uniformSlot(c, l)) std::memcpy(u->f, v, 16); }
HRL_API void rlUniform4iv(RLint l, RLsize, const RLint* v) { CTX_OR_RETURN(); if (Uniform* u = uniformSlot(c, l)) std::memcpy(u->i, v, 16); }
HRL_API void rlUniformMatrix4fv(RLint l, RLsize, RLboolean transpose, const RLfloat* v) {
    CTX_OR_RETURN(); Uniform* u = uniformSlot(c, l); if (!u) return;
    if (transpose) for (int r = 0; r < 4; ++r) for (int k = 0; k < 4; ++k) u->f[4 * k + r] = v[4 * r + k]; else std::memcpy(u->f, v, 64);
}
HRL_API void rlUniformt(RLint l, RLtexture t) { CTX_OR_RETURN(); if (Uniform* u = uniformSlot(c, l)) u->tex = reinterpret_cast<Tex*>(t); }
HRL_API void rlUniformp(RLint l, RLprimitive p) { CTX_OR_RETURN(); if (Uniform* u = uniformSlot(c, l)) u->prim = (int)p; }

// the pass
HRL_API void rlRenderFrame(void) {
    CTX_OR_RETURN();
    Fbo* f = c->boundFbo ? c->boundFbo : (c->fbos.empty() ? nullptr : *c->fbos.rbegin());
    if (!f || !f->color0 || f->color0->w <= 0) { setError(RL_INVALID_FRAMEBUFFER_OPERATION, "rlRenderFrame without a colour attachment"); return; }
    if (f->color0->w != c->w || f->color0->h != c->h) { if (!BK(Resize, f->color0->w, f->color0->h)) return; c->w = f->color0->w; c->h = f->color0->h; }
    for (auto& kv : c->prims) if (!syncPrimitive(c, kv.second.get())) return;
    if (c->sceneDirty) { if (!BK0(CommitScene)) return; c->sceneDirty = false; }
    if (!syncLights(c)) return;
    int ns = 0;
    if (!syncTables(c, ns)) return;
    HrcFrameParams fp;
    if (!syncFrame(c, fp)) return;
    Buf* gb = blockOf(c, c->framePrim.get(), "Globals");
    if (!gb || gb->data.size() < sizeof(HrcGlobals)) { setError(RL_INVALID_OPERATION, "frame program without its Globals block"); return; }
    HrcGlobals g; std::memcpy(&g, gb->data.data(), sizeof(g));
    fp.sampleIndex = g.sampleIndex;
    // PassGenerator rewrites the whole block every pass to bump Globals.sampleIndex (PassGenerator.cpp:349-360); the backend takes the sample index from
    // the frame parameters, so the block is forwarded only when something else in it changed (a forwarded block is a state change: it ends look-ahead)
    g.sampleIndex = 0;
    if (!c->haveGlobals || std::memcmp(&g, &c->lastGlobals, sizeof(g)) != 0) {
        if (!BK(SetGlobals, &g)) return;
        c->lastGlobals = g; c->haveGlobals = true;
    }
    BK(RenderPass, &fp, nullptr);
}
