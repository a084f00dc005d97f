// ref_scene.cpp — TEST INFRASTRUCTURE (build container only) -> oracle/_ref/libref_scene.so.
// The reference's OWN host code for the path — PassGenerator.cpp, Scene.cpp, Mesh.cpp, Lighting.cpp, PhysicallyBasedMaterial.cpp,
// GlassMaterial.cpp, MultiScatterUtil.cpp, the four Lights/*.cpp, ShaderCodeLoader.cpp, FileIO.cpp, Log.cpp and the header-only RLWrapper
// classes — compiled UNMODIFIED from /root/reference against the reference's rl.h and linked against libheatray_openrl.so, the product's
// implementation of the OpenRL entry points over the backend C-ABI. This file only drives it the way HeatrayRenderer.cpp does:
//   * the built-in scenes of HeatrayRenderer::changeScene (HeatrayRenderer.cpp:122-286), assembled with the reference's mesh providers,
//     materials and Scene::addMesh;
//   * Lighting::add*Light / updateLight / removeLight through PassGenerator::changeLighting;
//   * PassGenerator::renderPass with a PassCompleteCallback that maps the PixelPackBuffer (HeatrayRenderer.cpp:366-408).
// Stand-ins for what cannot be built here (listed, nothing else): util::loadTexture (TextureLoader.cpp needs FreeImage + stb: the stand-in
// serves the multiscatter LUT the caller supplies), the AssimpMeshProvider constructor (Assimp is not built; Scene::loadFromDisk is not
// used), FreeImage link stubs (ref_c_stubs.c).
#include <HeatrayRenderer/PassGenerator.h>
#include <HeatrayRenderer/OrbitCamera.h>
#include <HeatrayRenderer/Lights/DirectionalLight.h>
#include <HeatrayRenderer/Lights/EnvironmentLight.h>
#include <HeatrayRenderer/Lights/PointLight.h>
#include <HeatrayRenderer/Lights/SpotLight.h>
#include <HeatrayRenderer/Materials/GlassMaterial.h>
#include <HeatrayRenderer/Materials/PhysicallyBasedMaterial.h>
#include <HeatrayRenderer/Scene/AssimpMeshProvider.h>
#include <HeatrayRenderer/Scene/PlaneMeshProvider.h>
#include <HeatrayRenderer/Scene/Scene.h>
#include <HeatrayRenderer/Scene/SphereMeshProvider.h>
#include <RLWrapper/PixelPackBuffer.h>
#include <Utility/Log.h>
#include <Utility/TextureLoader.h>
#include <glm/glm/gtc/type_ptr.hpp>

#include <condition_variable>
#include <cstring>
#include <future>
#include <mutex>

// ---- stand-ins -------------------------------------------------------------------------------------------------------------------------
namespace {
std::vector<float> g_lut; int g_lutDim = 0;
class QuietLog final : public util::Log {
public:
    static void install() { setInstance(std::shared_ptr<QuietLog>(new QuietLog())); }
    void addNewItem(const std::string_view item, const Log::Type type) override { if (type == Log::Type::kError) std::fprintf(stderr, "[ref] %.*s\n", (int)item.size(), item.data()); }
};
}
namespace util {
// TextureLoader.cpp:37-42: without mips the sampler is LINEAR + CLAMP_TO_EDGE — the state the LUT is loaded with (MultiScatterUtil.cpp:145).
std::shared_ptr<openrl::Texture> loadTexture(const std::string_view path, bool generateMips, bool) {
    if (path.find("multiscatter_lut") == std::string_view::npos || g_lut.empty()) return nullptr;
    openrl::Texture::Descriptor desc; desc.dataType = RL_FLOAT; desc.format = RL_LUMINANCE; desc.internalFormat = RL_LUMINANCE; desc.width = g_lutDim; desc.height = g_lutDim;
    openrl::Texture::Sampler sampler;
    if (!generateMips) { sampler.magFilter = RL_LINEAR; sampler.minFilter = RL_LINEAR; sampler.wrapS = RL_CLAMP_TO_EDGE; sampler.wrapT = RL_CLAMP_TO_EDGE; }
    return openrl::Texture::create(g_lut.data(), desc, sampler, generateMips);
}
}
AssimpMeshProvider::AssimpMeshProvider(const std::string_view filename, bool, std::shared_ptr<Lighting>) : MeshProvider(filename), m_filename(filename) { std::abort(); }

// ---- the driver ------------------------------------------------------------------------------------------------------------------------
struct RefSceneOptions {          // plain-C view of the RenderOptions fields the probe sets
    int32_t maxRenderPasses, maxRayDepth, interactive, bokehShape;
    float envSolidColor[3], envExposure, envTheta; int32_t envEnabled;
    float cameraDistance, cameraPhi, cameraTheta, cameraTarget[3], focalLength, fstop /* <= 0: the reference's "no depth of field" entry */, focusDistance;
};
struct RefLight { int32_t type /* 0 directional, 1 point, 2 spot */; float color[3], power, position[3], theta, phi, inner, outer; };

struct Probe {
    PassGenerator gen;
    std::vector<std::shared_ptr<Light>> lights;      // in creation order, removed ones reset
    int w = 0, h = 0; bool everMapped = false;
    void wait() { std::promise<void> p; gen.runOpenRLTask([&p]() { p.set_value(); }); p.get_future().wait(); }
};

extern "C" {
#define REF_API __attribute__((visibility("default")))

REF_API void refscene_set_lut(const float* texels, int dim) { g_lut.assign(texels, texels + (size_t)dim * dim); g_lutDim = dim; }

REF_API void* refscene_create(int w, int h) {
    QuietLog::install();
    Probe* p = new Probe();
    p->w = w; p->h = h;
    p->gen.init(w, h);
    p->wait();
    return p;
}
REF_API int refscene_render(void* h, const struct RefSceneOptions* o, int passes, int resetFirst, float* rgbaOut);
REF_API void refscene_destroy(void* h) {
    Probe* p = static_cast<Probe*>(h);
    if (!p->everMapped) { RefSceneOptions o{}; o.maxRenderPasses = 1; o.maxRayDepth = 1; o.cameraDistance = 5.0f; o.focalLength = 50.0f; o.focusDistance = 5.0f; refscene_render(p, &o, 1, 1, nullptr); }   // runDestroyJob expects a mapped buffer
    p->lights.clear(); p->gen.destroy(); delete p;
}
REF_API void refscene_resize(void* h, int w, int hgt) { Probe* p = static_cast<Probe*>(h); p->w = w; p->h = hgt; p->gen.resize(w, hgt); p->wait(); }

// HeatrayRenderer::changeScene (HeatrayRenderer.cpp:122-286) for the four built-in scenes.
REF_API int refscene_load_preset(void* h, const char* name) {
    Probe* p = static_cast<Probe*>(h);
    const std::string sceneName = name;
    int ok = 1;
    p->gen.loadScene([&](std::shared_ptr<Scene> scene) {
        if (sceneName == "Editable PBR Material") {
            SphereMeshProvider sphereMeshProvider(50, 50, 1.0f, "PBR Sphere");
            std::shared_ptr<PhysicallyBasedMaterial> material = std::make_shared<PhysicallyBasedMaterial>("PBR");
            PhysicallyBasedMaterial::Parameters& params = material->parameters();
            params.metallic = 0.0f; params.roughness = 1.0f; params.baseColor = glm::vec3(0.8f); params.specularF0 = 0.0f; params.clearCoat = 0.0f; params.clearCoatRoughness = 0.0f;
            params.forceEnableAllTextures = true;
            scene->addMesh(&sphereMeshProvider, { material }, glm::mat4(1.0f));
        } else if (sceneName == "Editable Glass Material") {
            SphereMeshProvider sphereMeshProvider(50, 50, 1.0f, "Glass Sphere");
            std::shared_ptr<GlassMaterial> material = std::make_shared<GlassMaterial>("Glass");
            GlassMaterial::Parameters& params = material->parameters();
            params.baseColor = glm::vec3(0.8f); params.ior = 1.33f; params.roughness = 0.0f; params.density = 0.8f; params.forceEnableAllTextures = true;
            scene->addMesh(&sphereMeshProvider, { material }, glm::mat4(1.0f));
        } else if (sceneName == "Multi-Material") {
            PlaneMeshProvider planeMeshProvider(15, 15, "Plane");
            {
                std::shared_ptr<PhysicallyBasedMaterial> material = std::make_shared<PhysicallyBasedMaterial>("Ground");
                PhysicallyBasedMaterial::Parameters& params = material->parameters();
                params.metallic = 0.0f; params.roughness = 1.0f; params.baseColor = glm::vec3(0.9f); params.specularF0 = 0.0f;
                scene->addMesh(&planeMeshProvider, { material }, glm::translate(glm::mat4(1.0f), glm::vec3(0.0f, -1.5f, 0.0f)));
            }
            SphereMeshProvider sphereMeshProvider(50, 50, 1.0f, "Sphere");
            {
                std::shared_ptr<PhysicallyBasedMaterial> material = std::make_shared<PhysicallyBasedMaterial>("PBR");
                PhysicallyBasedMaterial::Parameters& params = material->parameters();
                params.metallic = 1.0f; params.roughness = 0.1f; params.baseColor = glm::vec3(0.4f); params.specularF0 = 0.3f;
                scene->addMesh(&sphereMeshProvider, { material }, glm::translate(glm::mat4(1.0f), glm::vec3(-0.9f, -0.5f, -0.8f)));
            }
            {
                std::shared_ptr<GlassMaterial> material = std::make_shared<GlassMaterial>("Glass");
                GlassMaterial::Parameters& params = material->parameters();
                params.roughness = 0.1f; params.baseColor = glm::vec3(0.9f, 0.6f, 0.6f); params.ior = 1.57f; params.density = 0.5f;
                scene->addMesh(&sphereMeshProvider, { material }, glm::translate(glm::mat4(1.0f), glm::vec3(1.2f, -0.5f, 0.8f)));
            }
        } else if (sceneName == "Sphere Array") {
            float radius = 0.5f;
            SphereMeshProvider sphereMeshProvider(50, 50, radius, "Sphere");
            float padding = radius * 0.2f;
            for (int metal = 0; metal < 2; ++metal) {
                float roughness = 0.0f;
                float startX = (-5.0f * (radius * 2.0f + padding)) + ((radius * 2.0f + padding) * 0.5f);
                for (int iSphere = 0; iSphere < 10; ++iSphere) {
                    std::shared_ptr<PhysicallyBasedMaterial> material = std::make_shared<PhysicallyBasedMaterial>("Sphere");
                    PhysicallyBasedMaterial::Parameters& params = material->parameters();
                    params.metallic = metal ? 1.0f : 0.0f; params.roughness = roughness; params.baseColor = glm::vec3(1.0f); params.specularF0 = 0.0f;
                    scene->addMesh(&sphereMeshProvider, { material }, glm::translate(glm::mat4(1.0f), glm::vec3(startX, metal ? 1.5f : 0.0f, 0.0f)));
                    roughness += 0.1f;
                    startX += radius * 2.0f + padding;
                }
            }
        } else ok = 0;
    });
    p->wait();
    return ok;
}

// Scene::applyTransform through PassGenerator::modifyScene (HeatrayRenderer.cpp: the scene transform sliders).
REF_API void refscene_apply_transform(void* h, const float* m16) {
    Probe* p = static_cast<Probe*>(h);
    const glm::mat4 t = glm::make_mat4(m16);
    p->gen.modifyScene([t](std::shared_ptr<Scene> scene) { scene->applyTransform(t); });
    p->wait();
}

// Lighting::add*Light + setParams + updateLight; returns the light's handle for refscene_remove_light (-1: the type is full, Lighting.cpp:201-217).
REF_API int refscene_add_light(void* h, const RefLight* L) {
    Probe* p = static_cast<Probe*>(h);
    int handle = -1;
    p->gen.changeLighting([&](std::shared_ptr<Lighting> lighting) {
        std::shared_ptr<Light> made;
        if (L->type == 0) {
            auto light = lighting->addDirectionalLight("dir");
            if (light) { DirectionalLight::Params q; q.color = glm::make_vec3(L->color); q.illuminance = L->power; q.orientation.theta = L->theta; q.orientation.phi = L->phi; light->setParams(q); lighting->updateLight(light); made = light; }
        } else if (L->type == 1) {
            auto light = lighting->addPointLight("point");
            if (light) { PointLight::Params q; q.color = glm::make_vec3(L->color); q.luminousIntensity = L->power; q.position = glm::make_vec3(L->position); light->setParams(q); lighting->updateLight(light); made = light; }
        } else {
            auto light = lighting->addSpotLight("spot");
            if (light) { SpotLight::Params q; q.color = glm::make_vec3(L->color); q.luminousIntensity = L->power; q.position = glm::make_vec3(L->position); q.orientation.theta = L->theta; q.orientation.phi = L->phi;
                         q.innerAngle = L->inner; q.outerAngle = L->outer; light->setParams(q); lighting->updateLight(light); made = light; }
        }
        if (made) { p->lights.push_back(made); handle = (int)p->lights.size() - 1; }
    });
    p->wait();
    return handle;
}
REF_API void refscene_remove_light(void* h, int handle) {
    Probe* p = static_cast<Probe*>(h);
    // Lighting::removeDirectionalLight (Lighting.cpp:239-271) moves the last light into the freed slot and renumbers its lightIndex uniform, but leaves that
    // slot's direction / colour / primitive entries as the removed light wrote them until the moved light's next updateLight — so, like any caller that
    // wants a consistent buffer, the probe refreshes the remaining lights right after the removal.
    p->gen.changeLighting([&](std::shared_ptr<Lighting> lighting) {
        lighting->removeLight(p->lights[handle]); p->lights[handle].reset();
        for (auto& l : p->lights) if (l) lighting->updateLight(l);
    });
    p->wait();
}

// `passes` progressive passes the way HeatrayRenderer::render kicks them (one renderPass job at a time, the callback maps the pixels);
// rgbaOut receives the last frame (w*h*4 floats, sums + sample count as the reference hands them to its display shader).
REF_API int refscene_render(void* h, const RefSceneOptions* o, int passes, int resetFirst, float* rgbaOut) {
    Probe* p = static_cast<Probe*>(h);
    PassGenerator::RenderOptions opt;
    opt.enableInteractiveMode = o->interactive != 0; opt.enableOfflineMode = false;
    opt.maxRenderPasses = (uint32_t)o->maxRenderPasses; opt.maxRayDepth = (uint32_t)o->maxRayDepth;
    opt.bokehShape = static_cast<PassGenerator::RenderOptions::BokehShape>(o->bokehShape);
    opt.environment.map = o->envEnabled ? std::string(EnvironmentLight::SOLID_COLOR) : std::string("<none>");
    opt.environment.solidColor = glm::make_vec3(o->envSolidColor); opt.environment.exposureCompensation = o->envExposure; opt.environment.thetaRotation = o->envTheta;
    OrbitCamera cam; cam.distance = o->cameraDistance; cam.phi = o->cameraPhi; cam.theta = o->cameraTheta; cam.target = glm::make_vec3(o->cameraTarget);
    opt.camera.viewMatrix = cam.createViewMatrix();
    opt.camera.focalLength = o->focalLength; opt.camera.focusDistance = o->focusDistance;
    opt.camera.fstop = o->fstop > 0.0f ? o->fstop : PassGenerator::RenderOptions::Camera::fstopOptions[0];
    opt.camera.setApertureRadius();
    opt.camera.aspectRatio = static_cast<float>(p->w) / static_cast<float>(p->h);      // HeatrayRenderer.cpp:71
    int width = 0, height = 0;
    for (int i = 0; i < passes; ++i) {
        opt.resetInternalState = (i == 0) && resetFirst;
        std::promise<void> done;
        p->gen.renderPass(opt, [&](bool frameDataAvailable, std::shared_ptr<openrl::PixelPackBuffer> pixels, float, size_t) {
            if (frameDataAvailable) {
                width = (int)pixels->width(); height = (int)pixels->height();
                // like the viewer (HeatrayRenderer.cpp:391-395) the callback maps the buffer and leaves it mapped: the next render job unmaps it first
                // (PassGenerator.cpp:329-333), and the destroy job unmaps it unconditionally (:424)
                const float* src = pixels->mapPixelData();
                p->everMapped = true;
                if (i + 1 == passes && rgbaOut) std::memcpy(rgbaOut, src, pixels->size());
            }
            done.set_value();
        });
        done.get_future().wait();
    }
    return width * 65536 + height;
}

} // extern "C"
