// heatray_headless — the viewer's render loop without the viewer: load an asset, kick progressive passes through the pass
// interface, resolve with the display colour pipeline, write the screenshot. Plain C over include/heatray_host.h only.
//
// Reference flow being mirrored: HeatrayRenderer::changeScene -> PassGenerator::loadScene -> Scene::loadFromDisk
// (HeatrayRenderer.cpp:288-310), camera framing from the scene AABB (:302-308), the pass-kicking loop of
// HeatrayRenderer::render (:366-408, one renderPass per completed pass until maxRenderPasses), the display pass
// (displayGL.frag via HeatrayRenderer.h:228-262) and saveScreenshot (:1619-1662). GLFW / ImGui / GL are absent by design.
#include "../../../include/heatray_host.h"

#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

namespace {

struct App {
    std::string scenePath, outPng = "heatray.png", outHdr, envMap, sessionIn, sessionOut;
    bool groundPlane = false;                                                     // --ground-plane: the viewer's "Add Ground Plane" button
    bool preset = false;                                                          // a built-in scene: the reference keeps the camera where it is
    bool haveSession = false; HrhSession session{};                               // --session: HeatrayRenderer::readSessionFile (HeatrayRenderer.cpp:581-688)
    int width = 800, height = 800, passes = 32, depth = 10, device = 0;          // reference defaults: 800x800 render area (main.cpp:42-43), 32 passes, depth 10
    bool meters = false, envOff = false; float envColor[3] = {0.5f, 0.5f, 0.5f}, envEV = 0.0f;
    float phi = 0.0f, theta = 0.0f, distanceScale = 1.0f, fstop = 0.0f, focalLength = 50.0f, exposureEV = 0.0f;
    HrcPostProcess post{0, 1.0f, 0.0f, 1.0f, 1.0f, 1.0f, 0.0f, 1.0f, 1.0f, 1.0f, 0.0f, 1.0f};
    // filled on the worker thread
    HrhSceneInfo info{}; HrhSceneUploadResult upload{}; HrcTexture envTex = HRC_NULL_TEXTURE; int rc = 0; std::string error;
    std::atomic<int> passesDone{0}; double passSeconds = 0.0;
};

void loadSceneTask(void* user, HrcContext ctx) {                                 // the LoadSceneCallback body (PassGenerator.cpp:124-136)
    App& a = *static_cast<App*>(user);
    HrhScene scene = nullptr;
    a.preset = false;
    for (const char* p : {"Editable PBR Material", "Editable Glass Material", "Multi-Material", "Sphere Array"}) if (a.scenePath == p) a.preset = true;
    a.rc = a.preset ? hrhBuildPresetScene(a.scenePath.c_str(), &scene)       // the built-in scenes of changeScene (HeatrayRenderer.cpp:123-286)
                    : hrhImportScene(a.scenePath.c_str(), a.meters ? 1 : 0, &scene);
    if (a.rc != HRC_OK) { a.error = hrhImportLastError(); return; }
    hrhSceneGetInfo(scene, &a.info);
    if (a.haveSession) {                                                         // readSessionFile :671-683: the session's yaw/pitch/roll/scale goes onto every mesh
        float xf[16]; hrhSceneTransformMatrix(a.session.rotationYaw, a.session.rotationPitch, a.session.rotationRoll, a.session.scale, xf);
        hrhSceneApplyTransform(scene, xf);
    }
    if (a.groundPlane) {
        float xf[16]; hrhSceneTransformMatrix(a.session.rotationYaw, a.session.rotationPitch, a.session.rotationRoll, a.session.scale, xf);
        const bool useSessionBox = a.haveSession && hrhAABBCamera(a.session.aabbMin, a.session.aabbMax, nullptr, nullptr, nullptr);
        hrhSceneAddGroundPlane(scene, useSessionBox ? a.session.aabbMin : a.info.aabbMin, useSessionBox ? a.session.aabbMax : a.info.aabbMax, a.haveSession ? xf : nullptr);
    }
    a.rc = hrhSceneUpload(scene, ctx, 1, HRC_NULL_TEXTURE, &a.upload);
    if (a.rc != HRC_OK) a.error = hrhImportLastError();
    hrhFreeScene(scene);
    if (a.rc == HRC_OK && !a.envMap.empty()) {                                   // EnvironmentLight::changeImageSource -> util::loadTexture(path, true, false) (EnvironmentLight.cpp:36-46)
        HrhLoadedTexture lt;
        a.rc = hrhLoadTexture(a.envMap.c_str(), 1, 0, &lt);
        if (a.rc != HRC_OK) { a.error = hrhImageLastError(); return; }
        HrcTextureDesc td{lt.image.width, lt.image.height, lt.format, HRC_WRAP_REPEAT, HRC_WRAP_CLAMP};
        if (lt.image.isFloat) a.rc = hrcCreateTexture2D(ctx, &td, lt.image.data, &a.envTex);
        else { std::vector<uint8_t> rgba((size_t)lt.image.width * lt.image.height * 4); hrhExpandToRGBA8(&lt.image, rgba.data()); a.rc = hrcCreateTexture2D(ctx, &td, rgba.data(), &a.envTex); }
        hrhFreeImage(&lt.image);
        if (a.rc != HRC_OK) a.error = hrcGetLastError(ctx);
    }
}

void passComplete(void* user, int32_t, const float*, int32_t, int32_t, float passTime, uint64_t) {
    App& a = *static_cast<App*>(user);
    a.passSeconds = passTime; a.passesDone++;                                    // passTime counts from the start of the job
}

void screenshotTask(void* user, HrcContext ctx) {                                // saveScreenshot, HeatrayRenderer.cpp:1619-1662
    App& a = *static_cast<App*>(user);
    const size_t n = (size_t)a.width * a.height;
    if (!a.outPng.empty()) {
        std::vector<uint8_t> rgb(n * 3);
        a.rc = hrcResolve(ctx, &a.post, HRC_RESOLVE_DISPLAY_RGB8, rgb.data());
        if (a.rc == HRC_OK) a.rc = hrhWritePNG(a.outPng.c_str(), rgb.data(), a.width, a.height, 3, 1);
        if (a.rc != HRC_OK) { a.error = "screenshot failed"; return; }
    }
    if (!a.outHdr.empty()) {
        std::vector<float> rgb(n * 3);
        a.rc = hrcResolve(ctx, nullptr, HRC_RESOLVE_HDR_RGB32F, rgb.data());
        const bool pfm = a.outHdr.size() > 4 && a.outHdr.compare(a.outHdr.size() - 4, 4, ".pfm") == 0;
        if (a.rc == HRC_OK) a.rc = pfm ? hrhWritePFM(a.outHdr.c_str(), rgb.data(), a.width, a.height, 1) : hrhWriteHDR(a.outHdr.c_str(), rgb.data(), a.width, a.height, 1);
        if (a.rc != HRC_OK) a.error = "HDR screenshot failed";
    }
}

int usage() {
    std::fprintf(stderr,
        "usage: heatray_headless <scene.obj|.gltf|.glb|built-in scene name> [--size W H] [--passes N] [--depth D] [--meters] [--device K]\n"
        "       [--env r g b | --env-map file.hdr | --no-env] [--ev EV] [--orbit PHI THETA] [--distance-scale S] [--focal MM] [--fstop F]\n"
        "       [--out image.png] [--hdr image.hdr|image.pfm] [--tonemap] [--exposure EV]\n"
        "       [--ground-plane]\n"
        "       [--session in.xml] [--save-session out.xml]     (a session supplies the scene, options, camera and post-processing;\n"
        "                                                        flags given after --session override it)\n");
    return 2;
}

} // namespace

int main(int argc, char** argv) {
    App a;
    for (int i = 1; i < argc; ++i) {
        const std::string k = argv[i];
        auto need = [&](int n) { if (i + n >= argc) { usage(); std::exit(2); } };
        if (k == "--size") { need(2); a.width = std::atoi(argv[++i]); a.height = std::atoi(argv[++i]); }
        else if (k == "--passes") { need(1); a.passes = std::atoi(argv[++i]); }
        else if (k == "--depth") { need(1); a.depth = std::atoi(argv[++i]); }
        else if (k == "--device") { need(1); a.device = std::atoi(argv[++i]); }
        else if (k == "--meters") a.meters = true;
        else if (k == "--no-env") a.envOff = true;
        else if (k == "--env") { need(3); for (int c = 0; c < 3; ++c) a.envColor[c] = (float)std::atof(argv[++i]); }
        else if (k == "--env-map") { need(1); a.envMap = argv[++i]; }
        else if (k == "--ev") { need(1); a.envEV = (float)std::atof(argv[++i]); }
        else if (k == "--orbit") { need(2); a.phi = (float)std::atof(argv[++i]); a.theta = (float)std::atof(argv[++i]); }
        else if (k == "--distance-scale") { need(1); a.distanceScale = (float)std::atof(argv[++i]); }
        else if (k == "--focal") { need(1); a.focalLength = (float)std::atof(argv[++i]); }
        else if (k == "--fstop") { need(1); a.fstop = (float)std::atof(argv[++i]); }
        else if (k == "--out") { need(1); a.outPng = argv[++i]; }
        else if (k == "--hdr") { need(1); a.outHdr = argv[++i]; }
        else if (k == "--ground-plane") a.groundPlane = true;
        else if (k == "--save-session") { need(1); a.sessionOut = argv[++i]; }
        else if (k == "--session") {
            need(1); a.sessionIn = argv[++i];
            hrhSessionDefaults(&a.session); int32_t missing = 0;
            if (hrhSessionRead(a.sessionIn.c_str(), &a.session, &missing) != HRC_OK) { std::fprintf(stderr, "session: %s\n", hrhSessionLastError()); return 1; }
            a.haveSession = true;
            HrhRenderOptions so; hrhDefaultRenderOptions(&so);
            hrhSessionApply(&a.session, &so, &a.post);
            a.exposureEV = a.session.exposure;
            a.passes = (int)so.maxRenderPasses; a.depth = (int)so.maxRayDepth; a.meters = a.session.units == 1;        // SceneUnits::kCentimeters (HeatrayRenderer.h:302-305)
            a.envOff = !so.envEnabled; a.envEV = so.envExposureCompensation; std::memcpy(a.envColor, so.envSolidColor, 12);
            a.focalLength = so.focalLength; a.fstop = so.fstop; a.phi = a.session.orbitPhi; a.theta = a.session.orbitTheta;
            if (a.scenePath.empty()) a.scenePath = a.session.scene;
            if (so.envEnabled && !so.envUseSolidColor)                           // EnvironmentLight::changeImageSource (EnvironmentLight.cpp:30-46)
                a.envMap = a.session.environmentBuiltIn ? std::string("Resources/Environments/") + a.session.environmentMap : std::string(a.session.environmentMap);
        }
        else if (k == "--tonemap") a.post.tonemappingEnabled = 1;
        else if (k == "--exposure") { need(1); a.exposureEV = (float)std::atof(argv[++i]); a.post.cameraExposure = std::pow(2.0f, a.exposureEV); }
        else if (k[0] == '-') return usage();
        else a.scenePath = k;
    }
    if (a.scenePath.empty() || a.width <= 0 || a.height <= 0 || a.passes <= 0) return usage();

    HrhPassGenerator pg = nullptr;
    if (hrhPassGeneratorCreate(a.device, a.width, a.height, &pg) != HRC_OK) {
        std::fprintf(stderr, "cannot create the CUDA backend (there is no CPU path): %s\n", pg ? hrhPassGeneratorLastError(pg) : "no device");
        return 1;
    }
    const auto t0 = std::chrono::steady_clock::now();
    hrhPassGeneratorRunTask(pg, loadSceneTask, &a);
    hrhPassGeneratorWaitIdle(pg);
    if (a.rc != HRC_OK) { std::fprintf(stderr, "scene load failed: %s\n", a.error.c_str()); hrhPassGeneratorDestroy(pg); return 1; }
    const double loadS = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();

    HrhRenderOptions o; hrhDefaultRenderOptions(&o);
    if (a.haveSession) hrhSessionApply(&a.session, &o, nullptr);                // sample mode, bokeh shape, channel clamp, env rotation, aperture ... (:588-622)
    o.enableInteractiveMode = 0; o.enableOfflineMode = 1;                        // no per-pass readback (PassGenerator.cpp:388-396); a headless run is always offline
    o.maxRenderPasses = (uint32_t)a.passes; o.maxRayDepth = (uint32_t)a.depth;
    o.envEnabled = a.envOff ? 0 : 1; o.envExposureCompensation = a.envEV;
    if (!a.envMap.empty()) { o.envUseSolidColor = 0; o.envTexture = a.envTex; } else { o.envUseSolidColor = 1; std::memcpy(o.envSolidColor, a.envColor, 12); }
    // camera framing, HeatrayRenderer.cpp:300-310 (AABB.h:34-44: radius() is the full diagonal). With a session the stored AABB under the
    // session's scene transform drives it when valid (:671-683); otherwise the stored orbit camera is used as is.
    float centre[3], radius = 0.0f, distance = 0.0f, xf[16];
    if (a.haveSession) {
        hrhSceneTransformMatrix(a.session.rotationYaw, a.session.rotationPitch, a.session.rotationRoll, a.session.scale, xf);
        if (hrhAABBCamera(a.session.aabbMin, a.session.aabbMax, xf, centre, &radius)) distance = radius * 3.0f;
        else { std::memcpy(centre, a.session.orbitTarget, 12); distance = a.session.orbitDistance; o.focusDistance = a.session.cameraFocusDistance; }
    } else if (a.preset) {
        centre[0] = centre[1] = centre[2] = 0.0f; distance = 19.0f; o.focusDistance = distance;       // OrbitCamera defaults (OrbitCamera.h:22-25), auto-focus on the centre
    } else {
        hrhAABBCamera(a.info.aabbMin, a.info.aabbMax, nullptr, centre, &radius);
        distance = radius * 3.0f;
    }
    distance *= a.distanceScale;
    hrhOrbitCameraViewMatrix(distance, a.phi, a.theta, centre, o.viewMatrix);
    o.aspectRatio = (float)a.width / (float)a.height; o.focalLength = a.focalLength;
    if (radius > 0.0f) o.focusDistance = distance;                               // auto-focus on the scene centre (:308)
    if (a.fstop > 0.0f) { o.fstop = a.fstop; o.apertureRadius = hrhApertureRadius(a.focalLength, a.fstop); }
    if (!a.sessionOut.empty()) {                                                 // writeSessionFile, HeatrayRenderer.cpp:503-579
        HrhSession w; if (a.haveSession) w = a.session; else hrhSessionDefaults(&w);
        w.interactiveMode = 0; w.maxRenderPasses = o.maxRenderPasses; w.maxRayDepth = o.maxRayDepth; w.maxChannelValue = o.maxChannelValue;
        std::snprintf(w.scene, sizeof(w.scene), "%s", a.scenePath.c_str());
        w.sampleMode = (uint32_t)o.sampleMode; w.bokehShape = (uint32_t)o.bokehShape;
        if (!a.envMap.empty()) { if (!a.haveSession || a.envMap != std::string("Resources/Environments/") + a.session.environmentMap) { std::snprintf(w.environmentMap, sizeof(w.environmentMap), "%s", a.envMap.c_str()); w.environmentBuiltIn = 0; } }
        else std::snprintf(w.environmentMap, sizeof(w.environmentMap), "%s", a.envOff ? "<none>" : "solid color");
        w.environmentExposureCompensation = o.envExposureCompensation; w.environmentThetaRotation = o.envThetaRotation; std::memcpy(w.environmentSolidColor, o.envSolidColor, 12);
        w.cameraAspectRatio = o.aspectRatio; w.cameraFocusDistance = o.focusDistance; w.cameraFocalLength = o.focalLength; w.cameraApertureRadius = o.apertureRadius; w.cameraFStop = o.fstop;
        w.orbitDistance = distance; w.orbitPhi = a.phi; w.orbitTheta = a.theta; std::memcpy(w.orbitTarget, centre, 12); w.orbitMaxDistance = radius * 10.0f;
        w.units = a.meters ? 1u : 0u;
        if (!a.haveSession) { std::memcpy(w.aabbMin, a.info.aabbMin, 12); std::memcpy(w.aabbMax, a.info.aabbMax, 12); }
        w.distanceScale = radius;
        w.tonemapEnable = a.post.tonemappingEnabled; w.exposure = a.exposureEV; w.brightness = a.post.brightness; w.contrast = a.post.contrast;
        w.hue = a.post.hue; w.saturation = a.post.saturation; w.vibrance = a.post.vibrance; w.red = a.post.red; w.green = a.post.green; w.blue = a.post.blue;
        w.vignetteIntensity = a.post.vignetteIntensity; w.vignetteFalloff = a.post.vignetteFalloff;
        if (hrhSessionWrite(a.sessionOut.c_str(), &w) != HRC_OK) std::fprintf(stderr, "warning: %s\n", hrhSessionLastError());
    }

    const auto t1 = std::chrono::steady_clock::now();
    o.resetInternalState = 1;
    hrhPassGeneratorRenderPass(pg, &o, passComplete, &a);                       // offline mode: ONE job runs every pass up to maxRenderPasses (PassGenerator.cpp:366-399)
    hrhPassGeneratorWaitIdle(pg);
    const double renderS = std::chrono::duration<double>(std::chrono::steady_clock::now() - t1).count();
    hrhPassGeneratorRunTask(pg, screenshotTask, &a);
    hrhPassGeneratorWaitIdle(pg);
    const char* err = hrhPassGeneratorLastError(pg);
    const bool ok = a.rc == HRC_OK && a.passesDone.load() == a.passes && (!err || !*err);
    std::printf("{\"scene\": \"%s\", \"meshes\": %d, \"materials\": %d, \"textures\": %d, \"textures_missing\": %d, \"lights\": [%d, %d, %d], "
                "\"size\": [%d, %d], \"passes\": %d, \"load_s\": %.3f, \"render_s\": %.3f, \"pass_s_sum\": %.3f, \"out\": \"%s\", \"ok\": %s}\n",
                a.scenePath.c_str(), a.upload.meshCount, a.upload.materialCount, a.upload.textureCount, a.upload.texturesMissing,
                a.upload.dirLights, a.upload.pointLights, a.upload.spotLights, a.width, a.height, a.passesDone.load(), loadS, renderS, a.passSeconds,
                a.outPng.c_str(), ok ? "true" : "false");
    if (!ok) std::fprintf(stderr, "error: %s %s\n", a.error.c_str(), err ? err : "");
    hrhPassGeneratorDestroy(pg);
    return ok ? 0 : 1;
}
