// host_pass_generator.cpp — hr::PassGenerator: the headless pass interface over the backend C-ABI.
//
// Mirrors the reference's PassGenerator (Source/HeatrayRenderer/PassGenerator.{h,cpp}) for the hot path:
//   * every public call is an asynchronous enqueue onto ONE worker thread that owns the backend context
//     (AsyncTaskQueue contract, Source/Utility/AsyncTaskQueue.h:129-164; PassGenerator.cpp:95-148);
//   * renderPass(options, callback): optional reset, per-pass frame uniforms, sample-index bookkeeping,
//     the pass itself, conditional readback, callback on the worker thread (PassGenerator.cpp:325-401);
//   * resetRenderingState diffs old/new options (PassGenerator.cpp:435-577);
//   * the result buffer stays mapped until the next pass unmaps it (PassGenerator.cpp:329-333).
// The OpenGL viewer, ImGui and session persistence are out of scope (no graphics stack).
#include "host_api.h"

#include <atomic>
#include <chrono>
#include <cmath>
#include <condition_variable>
#include <cstring>
#include <deque>
#include <functional>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

namespace hr {

constexpr int kNumRandomSequences = HRC_NUM_RANDOM_SEQUENCES;   // PassGenerator.h:193

// DebugVisualizationMode, PassGenerator.h:123-141
enum DebugVis { kNone, kGeometricNormals, kUVs, kTangents, kBitangents, kNormalmap, kFinalNormals, kBaseColor, kRoughness, kMetallic,
                kEmissive, kClearcoat, kClearcoatRoughness, kClearcoatNormalmap, kShader, kNANs, kInf };

class PassGenerator {
public:
    ~PassGenerator() { destroy(); }

    int init(int device, int width, int height) {                       // PassGenerator.cpp:30-42 + runInitJob :161-299
        m_worker = std::thread([this]() { workerLoop(); });
        int rc = HRC_OK;
        enqueue([this, device, width, height, &rc]() {
            HrcContextDesc desc{device, 0};
            rc = hrcCreateContext(&desc, &m_ctx);
            if (rc != HRC_OK) { m_error = hrcGetLastError(nullptr); return; }
            hrhDefaultRenderOptions(&m_options);
            m_options.maxRenderPasses = 0;                                // forces the first reset to build the QMC tables
            m_globals = HrcGlobals{}; m_globals.maxRayDepth = 5; m_globals.maxChannelValue = 3.14159265358979f;
            rc = applySize(width, height);
        });
        waitIdle();
        m_initialised = rc == HRC_OK;
        return rc;
    }

    void destroy() {                                                     // PassGenerator.cpp:100-106, runDestroyJob :414-433
        if (!m_worker.joinable()) return;
        enqueue([this]() {
            if (m_ctx) { if (m_mapped) hrcReleasePixels(m_ctx); hrcDestroyContext(m_ctx); m_ctx = nullptr; }
        });
        { std::lock_guard<std::mutex> g(m_mutex); m_quit = true; }
        m_cv.notify_all();
        m_worker.join();
    }

    void resize(int width, int height) { enqueue([this, width, height]() { check(applySize(width, height)); m_options.resetInternalState = 1; }); }   // :108-113, :301-323

    void renderPass(const HrhRenderOptions& opts, HrhPassCompleteCallback cb, void* user) {   // :115-122
        HrhRenderOptions copy = opts;
        enqueue([this, copy, cb, user]() { runRenderFrameJob(copy, cb, user); }, &copy);
    }

    void runTask(HrhContextTask task, void* user) { enqueue([this, task, user]() { task(user, m_ctx); }); }   // loadScene / changeLighting / modifyScene / runOpenRLTask

    void waitIdle() {
        std::unique_lock<std::mutex> g(m_mutex);
        m_idle.wait(g, [this]() { return m_jobs.empty() && !m_busy; });
    }

    const char* lastError() const { return m_error.c_str(); }

private:
    struct Job { std::function<void()> fn; bool isRender = false; HrhRenderOptions options{}; };
    void enqueue(std::function<void()> job, const HrhRenderOptions* renderOptions = nullptr) {
        Job j; j.fn = std::move(job);
        if (renderOptions) { j.isRender = true; j.options = *renderOptions; }
        { std::lock_guard<std::mutex> g(m_mutex); m_jobs.push_back(std::move(j)); }
        m_cv.notify_one();
    }
    // Look-ahead for runRenderFrameJob: is the job right behind the running one another pass of the same progressive render (no reset)?
    // How many of the jobs queued right behind the running one (up to HRC_MAX_DEFERRED_PASSES) continue the same progressive render.
    int queuedContinuations() {
        std::lock_guard<std::mutex> g(m_mutex);
        int k = 0;
        for (const auto& j : m_jobs) {
            if (k >= HRC_MAX_DEFERRED_PASSES || !j.isRender) break;
            const HrhRenderOptions& n = j.options;
            if (n.resetInternalState || n.enableInteractiveMode != m_options.enableInteractiveMode || n.enableOfflineMode != m_options.enableOfflineMode) break;
            ++k;
        }
        return k;
    }
    void workerLoop() {
        for (;;) {
            std::function<void()> job; bool isRender = false;
            {
                std::unique_lock<std::mutex> g(m_mutex);
                m_cv.wait(g, [this]() { return m_quit || !m_jobs.empty(); });
                if (m_jobs.empty()) { if (m_quit) return; continue; }
                job = std::move(m_jobs.front().fn); isRender = m_jobs.front().isRender; m_jobs.pop_front(); m_busy = true;
            }
            if (!isRender) dropPrefetchedPass();                          // anything else (scene, size, lights ...) invalidates a pass rendered ahead
            job();
            { std::lock_guard<std::mutex> g(m_mutex); m_busy = false; }
            m_idle.notify_all();
        }
    }
    void check(int rc) { if (rc != HRC_OK && m_ctx) m_error = hrcGetLastError(m_ctx); }

    int applySize(int w, int h) {
        if (m_mapped) { hrcReleasePixels(m_ctx); m_mapped = false; }
        int rc = hrcResize(m_ctx, w, h); if (rc) { check(rc); return rc; }
        m_width = w; m_height = h;
        // generateSequenceOffsets (PassGenerator.cpp:150-159) + the per-pixel sequence ids (quirk B-Q2)
        const size_t n = hrhSequenceOffsetCount(w, h);
        std::vector<float> off(n); hrhGenerateSequenceOffsets(w, h, off.data(), n);
        std::vector<uint8_t> ids((size_t)w * h); hrhGenerateSequenceIds(w, h, kNumRandomSequences, ids.data());
        rc = hrcSetPixelTables(m_ctx, off.data(), n, ids.data()); check(rc);
        if (rc) return rc;
        // interactive block-coordinate texture + per-block hashes (PassGenerator.cpp:266-294, perspective.rlsl:42-57)
        const int bs = 3;                                                 // RenderOptions::kInteractiveBlockSize, PassGenerator.h:59
        std::vector<float> coords((size_t)bs * bs * 3);
        hrhGenerateInteractiveBlockCoords(bs, bs, m_blockShuffleSeed, coords.data());
        const int bx = (w + bs - 1) / bs, by = (h + bs - 1) / bs;
        std::vector<float> brand((size_t)bx * by * 2);
        hrhGenerateInteractiveBlockRand(bx, by, brand.data());
        rc = hrcSetInteractiveBlocks(m_ctx, bs, bs, coords.data(), bx, by, brand.data()); check(rc);
        return rc;
    }

    void setDebugGlobals(int mode) {                                     // PassGenerator.cpp:473-568
        HrcGlobals& g = m_globals;
        g.enableDebugVisualizer = 0; g.showGeometricNormals = g.showUVs = g.showTangents = g.showBitangents = g.showNormalmap = 0;
        g.showFinalNormals = g.showBaseColor = g.showRoughness = g.showMetallic = g.showEmissive = 0;
        g.showClearcoat = g.showClearcoatRoughness = g.showClearcoatNormalmap = g.showShader = 0;
        g.enableAccumulatorVisualizer = g.showNANs = g.showInf = 0;
        if (mode != kNone && mode != kNANs && mode != kInf) g.enableDebugVisualizer = 1;
        switch (mode) {
            case kGeometricNormals: g.showGeometricNormals = 1; break;
            case kUVs: g.showUVs = 1; break;
            case kTangents: g.showTangents = 1; break;
            case kBitangents: g.showBitangents = 1; break;
            case kNormalmap: g.showNormalmap = 1; break;
            case kFinalNormals: g.showFinalNormals = 1; break;
            case kBaseColor: g.showBaseColor = 1; break;
            case kRoughness: g.showRoughness = 1; break;
            case kMetallic: g.showMetallic = 1; break;
            case kEmissive: g.showEmissive = 1; break;
            case kClearcoat: g.showClearcoat = 1; break;
            case kClearcoatRoughness: g.showClearcoatRoughness = 1; break;
            case kClearcoatNormalmap: g.showClearcoatNormalmap = 1; break;
            case kShader: g.showShader = 1; break;
            case kNANs: g.enableAccumulatorVisualizer = 1; g.showNANs = 1; break;
            case kInf: g.enableAccumulatorVisualizer = 1; g.showInf = 1; break;
            default: break;
        }
    }

    void changeEnvironment(const HrhRenderOptions& o) {                  // PassGenerator.cpp:580-603 + EnvironmentLight.cpp:48-98
        HrcEnvLight env{};
        env.enabled = o.envEnabled;
        env.exposureCompensation = std::pow(2.0f, o.envExposureCompensation);
        env.thetaRotation = o.envThetaRotation;
        env.texture = HRC_NULL_TEXTURE;
        if (o.envEnabled) {
            if (o.envUseSolidColor) {
                if (m_solidTex >= 0) hrcDestroyTexture(m_ctx, m_solidTex);
                HrcTextureDesc d{1, 1, HRC_TEX_RGB32F, HRC_WRAP_CLAMP, HRC_WRAP_CLAMP};
                check(hrcCreateTexture2D(m_ctx, &d, o.envSolidColor, &m_solidTex));
                env.texture = m_solidTex;
            } else env.texture = o.envTexture;
        }
        check(hrcSetEnvironment(m_ctx, &env));
    }

    void dropPrefetchedPass() { if (!m_prefetched.empty() && m_ctx) hrcDiscardDeferred(m_ctx); m_prefetched.clear(); }

    void resetRenderingState(const HrhRenderOptions& n) {                // PassGenerator.cpp:435-577
        dropPrefetchedPass();
        m_currentSampleIndex = 0;
        m_currentBlockPixelSample[0] = m_currentBlockPixelSample[1] = 0;
        check(hrcClear(m_ctx));
        const HrhRenderOptions& o = m_options;
        if (!m_envApplied || o.envEnabled != n.envEnabled || o.envUseSolidColor != n.envUseSolidColor || o.envTexture != n.envTexture ||
            o.envExposureCompensation != n.envExposureCompensation || o.envThetaRotation != n.envThetaRotation ||
            std::memcmp(o.envSolidColor, n.envSolidColor, sizeof(o.envSolidColor)) != 0) {
            changeEnvironment(n); m_envApplied = true;
        }
        if (o.sampleMode != n.sampleMode || o.maxRenderPasses != n.maxRenderPasses || o.bokehShape != n.bokehShape) {
            const uint32_t count = n.maxRenderPasses;
            std::vector<float> xy(2 * (size_t)kNumRandomSequences * count), ap(xy.size());
            hrhGenerateRandomSequences(n.sampleMode, n.bokehShape, kNumRandomSequences, count, xy.data(), ap.data());
            check(hrcSetSequences(m_ctx, kNumRandomSequences, (int)count, xy.data(), ap.data()));
        }
        m_globals.maxRayDepth = (int)n.maxRayDepth;
        m_globals.maxChannelValue = n.maxChannelValue;
        if (o.debugVisMode != n.debugVisMode || !m_globalsApplied) setDebugGlobals(n.debugVisMode);
        check(hrcSetGlobals(m_ctx, &m_globals)); m_globalsApplied = true;
        if (n.debugPassRendering) m_currentSampleIndex = (uint32_t)n.debugPassIndex;   // :570-572
        m_options = n;
        m_options.resetInternalState = 0;
    }

    void runRenderFrameJob(const HrhRenderOptions& n, HrhPassCompleteCallback cb, void* user) {   // PassGenerator.cpp:325-401
        auto t0 = std::chrono::steady_clock::now();
        if (!m_ctx) return;
        if (m_mapped) { hrcReleasePixels(m_ctx); m_mapped = false; }       // :329-333
        if (n.enableInteractiveMode != m_options.enableInteractiveMode || n.enableOfflineMode != m_options.enableOfflineMode || n.resetInternalState)
            resetRenderingState(n);
        HrcFrameParams f{};
        f.fovTan = hrhFovTan(m_options.focalLength);                       // :341-343, :357
        f.aspectRatio = m_options.aspectRatio; f.focusDistance = m_options.focusDistance; f.apertureRadius = m_options.apertureRadius;
        std::memcpy(f.viewMatrix, m_options.viewMatrix, sizeof(f.viewMatrix));
        f.blockSize[0] = f.blockSize[1] = 3;                               // RenderOptions::kInteractiveBlockSize, PassGenerator.h:59
        f.maxSampleIndex = float(m_options.maxRenderPasses);
        f.bandCount = 1; f.bandHeight = 1;
        bool jobCompleted = false;
        do {
            f.sampleIndex = (int)m_currentSampleIndex;
            f.interactiveMode = m_options.enableInteractiveMode ? 1 : 0;
            f.currentBlockPixelSample[0] = m_currentBlockPixelSample[0]; f.currentBlockPixelSample[1] = m_currentBlockPixelSample[1];
            if (m_options.enableInteractiveMode) {                         // :372-381
                if (++m_currentBlockPixelSample[0] == 3) {
                    m_currentBlockPixelSample[0] = 0;
                    if (++m_currentBlockPixelSample[1] == 3) { m_currentBlockPixelSample[0] = m_currentBlockPixelSample[1] = 0; ++m_currentSampleIndex; }
                }
            } else ++m_currentSampleIndex;
            int rc;
            if (!m_prefetched.empty() && std::memcmp(&m_prefetched.front(), &f, sizeof(f)) == 0) { rc = hrcCommitDeferred(m_ctx); m_prefetched.pop_front(); }   // this pass was rendered ahead
            else { dropPrefetchedPass(); rc = hrcEnqueuePass(m_ctx, &f); }  // rlRenderFrame(), :386
            check(rc);
            jobCompleted = rc != HRC_OK || !(m_options.enableOfflineMode && m_currentSampleIndex < m_options.maxRenderPasses);   // :390
            // Look-ahead: when the caller has already queued the next passes of this progressive render (up to three), start them now,
            // into deferred buffers, so that they render while this pass is copied to the host and handed to the callback. The frame
            // constants of those passes are fully determined here (no reset: the queued jobs' own options are not consulted, :334-339).
            if (jobCompleted && rc == HRC_OK && !m_options.enableOfflineMode && !m_options.enableInteractiveMode && !m_options.debugPassRendering) {
                const int want = queuedContinuations();
                while ((int)m_prefetched.size() < want && m_currentSampleIndex + m_prefetched.size() < (size_t)m_options.maxRenderPasses) {
                    HrcFrameParams ahead = f;
                    ahead.sampleIndex = (int)(m_currentSampleIndex + m_prefetched.size());
                    if (hrcEnqueuePassDeferred(m_ctx, &ahead) != HRC_OK) break;
                    m_prefetched.push_back(ahead);
                }
            }
            const float* pixels = nullptr;
            if (jobCompleted && rc == HRC_OK) {                            // setPixelData + map, :392-394 and HeatrayRenderer.cpp:391-395
                rc = hrcReadPixels(m_ctx, &pixels); check(rc);
                m_mapped = rc == HRC_OK;
            }
            const float passTime = std::chrono::duration<float>(std::chrono::steady_clock::now() - t0).count();
            if (cb) cb(user, jobCompleted && pixels ? 1 : 0, pixels, m_width, m_height, passTime, m_currentSampleIndex);   // :397-398
        } while (!jobCompleted);
    }

    std::thread m_worker;
    std::mutex m_mutex;
    std::condition_variable m_cv, m_idle;
    std::deque<Job> m_jobs;
    bool m_quit = false, m_busy = false, m_initialised = false;

    HrcContext m_ctx = nullptr;
    std::string m_error;
    HrhRenderOptions m_options{};
    HrcGlobals m_globals{};
    bool m_envApplied = false, m_globalsApplied = false, m_mapped = false;
    std::deque<HrcFrameParams> m_prefetched;                           // passes rendered ahead into the backend's deferred buffers, oldest first
    HrcTexture m_solidTex = HRC_NULL_TEXTURE;
    uint32_t m_currentSampleIndex = 0;
    uint32_t m_blockShuffleSeed = 0x5eed;   // the reference seeds this shuffle from std::random_device (quirk B-Q6)
    int m_currentBlockPixelSample[2] = {0, 0};
    int m_width = 0, m_height = 0;
};

} // namespace hr

extern "C" {

HRC_API void hrhDefaultRenderOptions(HrhRenderOptions* o) {             // defaults of PassGenerator.h:49-150
    std::memset(o, 0, sizeof(*o));
    o->enableInteractiveMode = 1; o->enableOfflineMode = 0; o->resetInternalState = 1;
    o->maxRenderPasses = 32; o->maxRayDepth = 10; o->maxChannelValue = 3.14159265358979f;
    o->envEnabled = 1; o->envUseSolidColor = 1; o->envSolidColor[0] = o->envSolidColor[1] = o->envSolidColor[2] = 0.5f; o->envTexture = HRC_NULL_TEXTURE;
    o->aspectRatio = -1.0f; o->focusDistance = 1.0f; o->focalLength = 50.0f; o->apertureRadius = 0.0f; o->fstop = 32.0f;
    for (int i = 0; i < 4; ++i) o->viewMatrix[5 * i] = 1.0f;
    o->sampleMode = HRH_SAMPLE_SOBOL; o->bokehShape = HRH_BOKEH_CIRCULAR;
}
HRC_API int hrhPassGeneratorCreate(int32_t device, int32_t width, int32_t height, HrhPassGenerator* out) {
    if (!out) return HRC_ERR_INVALID_VALUE;
    auto* pg = new hr::PassGenerator();
    int rc = pg->init(device, width, height);
    *out = reinterpret_cast<HrhPassGenerator>(pg);     // returned even on failure so the caller can read the error, then destroy
    return rc;
}
HRC_API int hrhPassGeneratorDestroy(HrhPassGenerator pg) { delete reinterpret_cast<hr::PassGenerator*>(pg); return HRC_OK; }
HRC_API int hrhPassGeneratorResize(HrhPassGenerator pg, int32_t w, int32_t h) { reinterpret_cast<hr::PassGenerator*>(pg)->resize(w, h); return HRC_OK; }
HRC_API int hrhPassGeneratorRenderPass(HrhPassGenerator pg, const HrhRenderOptions* o, HrhPassCompleteCallback cb, void* user) {
    if (!pg || !o) return HRC_ERR_INVALID_VALUE;
    reinterpret_cast<hr::PassGenerator*>(pg)->renderPass(*o, cb, user); return HRC_OK;
}
HRC_API int hrhPassGeneratorRunTask(HrhPassGenerator pg, HrhContextTask task, void* user) {
    if (!pg || !task) return HRC_ERR_INVALID_VALUE;
    reinterpret_cast<hr::PassGenerator*>(pg)->runTask(task, user); return HRC_OK;
}
HRC_API int hrhPassGeneratorWaitIdle(HrhPassGenerator pg) { reinterpret_cast<hr::PassGenerator*>(pg)->waitIdle(); return HRC_OK; }
HRC_API const char* hrhPassGeneratorLastError(HrhPassGenerator pg) { return reinterpret_cast<hr::PassGenerator*>(pg)->lastError(); }

}
