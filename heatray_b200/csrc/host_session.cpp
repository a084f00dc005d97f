// host_session.cpp — Heatray session files (XML): read, write, and the mapping onto render options / post-processing.
//
// Reference: Source/HeatrayRenderer/Session/Session.h:21-78 (the X-macro table of groups, names, types and defaults),
// Session.cpp:88-214 (parse / write through tinyxml2: <HeatraySession><Group><Variable value="..."/>...), and
// HeatrayRenderer::writeSessionFile / readSessionFile (HeatrayRenderer.cpp:503-690) for what the variables mean. tinyxml2 is not
// used here: the format is a two-level element tree with one attribute per leaf, read by the small scanner below and written
// with tinyxml2's own number formatting ("%.8g" floats, "%u", true/false; tinyxml2.cpp:580-597).
#include "host_api.h"

#include <cmath>
#include <cstddef>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

namespace {

enum VarType { kUInt, kFloat, kBool, kString };
struct VarDesc { const char* group; const char* name; VarType type; size_t offset; double dflt; const char* sdflt; };

#define OFF(f) offsetof(HrhSession, f)
// Same order, names, types and defaults as HEATRAY_SESSION_VARIABLES (Session.h:21-78).
const VarDesc kVars[] = {
    {"RenderOptions", "InteractiveMode", kBool, OFF(interactiveMode), 0, nullptr},
    {"RenderOptions", "MaxRenderPasses", kUInt, OFF(maxRenderPasses), 32, nullptr},
    {"RenderOptions", "MaxChannelValue", kFloat, OFF(maxChannelValue), 1.0, nullptr},
    {"RenderOptions", "Scene", kString, OFF(scene), 0, "Multi-Material"},
    {"RenderOptions", "MaxRayDepth", kUInt, OFF(maxRayDepth), 10, nullptr},
    {"RenderOptions", "SampleMode", kUInt, OFF(sampleMode), 0, nullptr},
    {"RenderOptions", "BokehShape", kUInt, OFF(bokehShape), 0, nullptr},
    {"RenderOptions", "EnvironmentMap", kString, OFF(environmentMap), 0, "studio.hdr"},
    {"RenderOptions", "EnvironmentBuiltIn", kBool, OFF(environmentBuiltIn), 1, nullptr},
    {"RenderOptions", "EnvironmentExposureCompensation", kFloat, OFF(environmentExposureCompensation), 0, nullptr},
    {"RenderOptions", "EnvironmentThetaRotation", kFloat, OFF(environmentThetaRotation), 0, nullptr},
    {"RenderOptions", "EnvironmentMapSolidColorX", kFloat, OFF(environmentSolidColor[0]), 0, nullptr},
    {"RenderOptions", "EnvironmentMapSolidColorY", kFloat, OFF(environmentSolidColor[1]), 0, nullptr},
    {"RenderOptions", "EnvironmentMapSolidColorZ", kFloat, OFF(environmentSolidColor[2]), 0, nullptr},
    {"RenderOptions", "CameraAspectRatio", kFloat, OFF(cameraAspectRatio), -1.0, nullptr},
    {"RenderOptions", "CameraFocusDistance", kFloat, OFF(cameraFocusDistance), 1.0, nullptr},
    {"RenderOptions", "CameraFocalLength", kFloat, OFF(cameraFocalLength), 50.0, nullptr},
    {"RenderOptions", "CameraApertureRadius", kFloat, OFF(cameraApertureRadius), 0, nullptr},
    {"RenderOptions", "CameraFStop", kFloat, OFF(cameraFStop), 0, nullptr},
    {"Camera", "OrbitDistance", kFloat, OFF(orbitDistance), 10.0, nullptr},
    {"Camera", "OrbitPhi", kFloat, OFF(orbitPhi), 10.0, nullptr},
    {"Camera", "OrbitTheta", kFloat, OFF(orbitTheta), 10.0, nullptr},
    {"Camera", "OrbitTargetX", kFloat, OFF(orbitTarget[0]), 10.0, nullptr},
    {"Camera", "OrbitTargetY", kFloat, OFF(orbitTarget[1]), 10.0, nullptr},
    {"Camera", "OrbitTargetZ", kFloat, OFF(orbitTarget[2]), 10.0, nullptr},
    {"Camera", "OrbitMaxDistance", kFloat, OFF(orbitMaxDistance), 10.0, nullptr},
    {"Scene", "Units", kUInt, OFF(units), 0, nullptr},
    {"Scene", "AABB_MinX", kFloat, OFF(aabbMin[0]), 0, nullptr},
    {"Scene", "AABB_MinY", kFloat, OFF(aabbMin[1]), 0, nullptr},
    {"Scene", "AABB_MinZ", kFloat, OFF(aabbMin[2]), 0, nullptr},
    {"Scene", "AABB_MaxX", kFloat, OFF(aabbMax[0]), 0, nullptr},
    {"Scene", "AABB_MaxY", kFloat, OFF(aabbMax[1]), 0, nullptr},
    {"Scene", "AABB_MaxZ", kFloat, OFF(aabbMax[2]), 0, nullptr},
    {"Scene", "DistanceScale", kFloat, OFF(distanceScale), 1.0, nullptr},
    {"Scene", "RotationYaw", kFloat, OFF(rotationYaw), 0, nullptr},
    {"Scene", "RotationPitch", kFloat, OFF(rotationPitch), 0, nullptr},
    {"Scene", "RotationRoll", kFloat, OFF(rotationRoll), 0, nullptr},
    {"Scene", "Scale", kFloat, OFF(scale), 1.0, nullptr},
    {"PostProcessing", "TonemapEnable", kBool, OFF(tonemapEnable), 0, nullptr},
    {"PostProcessing", "Exposure", kFloat, OFF(exposure), 0, nullptr},
    {"PostProcessing", "Brightness", kFloat, OFF(brightness), 0, nullptr},
    {"PostProcessing", "Contrast", kFloat, OFF(contrast), 0, nullptr},
    {"PostProcessing", "Hue", kFloat, OFF(hue), 0, nullptr},
    {"PostProcessing", "Saturation", kFloat, OFF(saturation), 0, nullptr},
    {"PostProcessing", "Vibrance", kFloat, OFF(vibrance), 0, nullptr},
    {"PostProcessing", "Red", kFloat, OFF(red), 0, nullptr},
    {"PostProcessing", "Green", kFloat, OFF(green), 0, nullptr},
    {"PostProcessing", "Blue", kFloat, OFF(blue), 0, nullptr},
    {"PostProcessing", "VignetteIntensity", kFloat, OFF(vignetteIntensity), 0, nullptr},
    {"PostProcessing", "VignetteFalloff", kFloat, OFF(vignetteFalloff), 0, nullptr},
};
#undef OFF
constexpr size_t kNumVars = sizeof(kVars) / sizeof(kVars[0]);
// Group order on disk: the reference iterates an unordered_map keyed by group name (Session.cpp:166), so the order is the standard
// library's; this is libstdc++'s (what the compiled reference writes here). Readers look groups up by name, so any order parses.
const char* const kGroups[] = {"Scene", "Camera", "PostProcessing", "RenderOptions"};

thread_local std::string g_sessionErr;

void setDefault(HrhSession* s, const VarDesc& v) {
    char* base = reinterpret_cast<char*>(s) + v.offset;
    switch (v.type) {
    case kUInt: *reinterpret_cast<uint32_t*>(base) = (uint32_t)v.dflt; break;
    case kFloat: *reinterpret_cast<float*>(base) = (float)v.dflt; break;
    case kBool: *reinterpret_cast<int32_t*>(base) = v.dflt != 0; break;
    case kString: std::snprintf(base, HRH_SESSION_STRING_SIZE, "%s", v.sdflt); break;
    }
}

// ---- a scanner for the subset of XML a session file uses: declaration, comments, nested elements, attributes ----
struct Element { std::string name; std::vector<std::pair<std::string, std::string>> attrs; std::vector<Element> children; };

std::string decodeEntities(const std::string& s) {
    std::string out; out.reserve(s.size());
    for (size_t i = 0; i < s.size(); ++i) {
        if (s[i] != '&') { out.push_back(s[i]); continue; }
        const size_t e = s.find(';', i);
        if (e == std::string::npos) { out.push_back('&'); continue; }
        const std::string ent = s.substr(i + 1, e - i - 1);
        if (ent == "amp") out.push_back('&'); else if (ent == "lt") out.push_back('<'); else if (ent == "gt") out.push_back('>');
        else if (ent == "quot") out.push_back('"'); else if (ent == "apos") out.push_back('\'');
        else if (!ent.empty() && ent[0] == '#') { const long cp = ent.size() > 1 && ent[1] == 'x' ? std::strtol(ent.c_str() + 2, nullptr, 16) : std::strtol(ent.c_str() + 1, nullptr, 10); if (cp > 0 && cp < 128) out.push_back((char)cp); }
        else { out.append(s, i, e - i + 1); }
        i = e;
    }
    return out;
}

struct XmlScanner {
    const std::string& t; size_t p = 0; bool ok = true;
    explicit XmlScanner(const std::string& text) : t(text) {}
    void skipWs() { while (p < t.size() && (t[p] == ' ' || t[p] == '\t' || t[p] == '\n' || t[p] == '\r')) ++p; }
    bool skipMisc() {          // whitespace, text, <?...?>, <!--...-->, <!DOCTYPE ...>
        for (;;) {
            while (p < t.size() && t[p] != '<') ++p;
            if (p >= t.size()) return false;
            if (t.compare(p, 4, "<!--") == 0) { const size_t e = t.find("-->", p + 4); if (e == std::string::npos) { ok = false; return false; } p = e + 3; }
            else if (t.compare(p, 2, "<?") == 0) { const size_t e = t.find("?>", p + 2); if (e == std::string::npos) { ok = false; return false; } p = e + 2; }
            else if (t.compare(p, 2, "<!") == 0) { const size_t e = t.find('>', p + 2); if (e == std::string::npos) { ok = false; return false; } p = e + 1; }
            else return true;
        }
    }
    std::string name() { const size_t s = p; while (p < t.size() && !std::strchr(" \t\r\n/>=", t[p])) ++p; return t.substr(s, p - s); }
    bool element(Element& out, int depth = 0) {      // at '<' of an opening tag
        if (depth > 64) { ok = false; return false; }
        ++p; out.name = name();
        if (out.name.empty()) { ok = false; return false; }
        for (;;) {
            skipWs();
            if (p >= t.size()) { ok = false; return false; }
            if (t[p] == '/') { if (p + 1 < t.size() && t[p + 1] == '>') { p += 2; return true; } ok = false; return false; }
            if (t[p] == '>') { ++p; break; }
            const std::string an = name(); skipWs();
            if (an.empty() || p >= t.size() || t[p] != '=') { ok = false; return false; }
            ++p; skipWs();
            if (p >= t.size() || (t[p] != '"' && t[p] != '\'')) { ok = false; return false; }
            const char q = t[p++]; const size_t e = t.find(q, p);
            if (e == std::string::npos) { ok = false; return false; }
            out.attrs.emplace_back(an, decodeEntities(t.substr(p, e - p))); p = e + 1;
        }
        for (;;) {                                    // children until the matching close tag
            if (!skipMisc()) { ok = false; return false; }
            if (t.compare(p, 2, "</") == 0) { const size_t e = t.find('>', p); if (e == std::string::npos) { ok = false; return false; } p = e + 1; return true; }
            Element c; if (!element(c, depth + 1)) return false;
            out.children.push_back(std::move(c));
        }
    }
};

const Element* child(const Element& e, const char* name) { for (const Element& c : e.children) if (c.name == name) return &c; return nullptr; }
const std::string* attr(const Element& e, const char* name) { for (const auto& a : e.attrs) if (a.first == name) return &a.second; return nullptr; }

bool toBool(const std::string& s, bool& out) {      // XMLUtil::ToBool (tinyxml2.cpp:644-666): integers, then true/True/TRUE, false/False/FALSE
    char* e = nullptr; const long v = std::strtol(s.c_str(), &e, 10);
    if (e != s.c_str() && *e == 0) { out = v != 0; return true; }
    if (s == "true" || s == "True" || s == "TRUE") { out = true; return true; }
    if (s == "false" || s == "False" || s == "FALSE") { out = false; return true; }
    return false;
}

std::string escapeAttr(const char* s) {
    std::string o;
    for (; *s; ++s) { switch (*s) { case '&': o += "&amp;"; break; case '<': o += "&lt;"; break; case '>': o += "&gt;"; break; case '"': o += "&quot;"; break; case '\'': o += "&apos;"; break; default: o.push_back(*s); } }
    return o;
}

} // namespace

extern "C" {

HRC_API const char* hrhSessionLastError(void) { return g_sessionErr.c_str(); }

HRC_API void hrhSessionDefaults(HrhSession* s) {
    if (!s) return;
    std::memset(s, 0, sizeof(*s));
    for (size_t i = 0; i < kNumVars; ++i) setDefault(s, kVars[i]);
}

// Session::parseSessionFile (Session.cpp:88-160). `session` keeps its current values for anything the file does not hold (the reference's
// variable table is a static that keeps earlier values too); *missing counts such variables. A group that is absent is skipped like :112-116.
HRC_API int hrhSessionRead(const char* path, HrhSession* s, int32_t* missing) {
    if (!path || !s) { g_sessionErr = "null argument"; return HRC_ERR_INVALID_VALUE; }
    if (missing) *missing = 0;
    FILE* f = std::fopen(path, "rb");
    if (!f) { g_sessionErr = std::string("unable to load session file ") + path; return HRC_ERR_INVALID_VALUE; }
    std::string text; char buf[4096]; size_t n;
    while ((n = std::fread(buf, 1, sizeof(buf), f)) > 0) text.append(buf, n);
    std::fclose(f);
    XmlScanner sc(text);
    Element root;
    bool found = false;
    while (sc.skipMisc()) {
        Element e; if (!sc.element(e)) break;
        if (e.name == "HeatraySession") { root = std::move(e); found = true; break; }
    }
    if (!sc.ok) { g_sessionErr = "malformed XML in session file"; return HRC_ERR_INVALID_VALUE; }
    if (!found) { g_sessionErr = "unable to load root session node HeatraySession"; return HRC_ERR_INVALID_VALUE; }
    int miss = 0;
    for (size_t i = 0; i < kNumVars; ++i) {
        const VarDesc& v = kVars[i];
        const Element* g = child(root, v.group);
        const Element* e = g ? child(*g, v.name) : nullptr;
        const std::string* val = e ? attr(*e, "value") : nullptr;
        if (!val) { ++miss; continue; }
        char* base = reinterpret_cast<char*>(s) + v.offset;
        switch (v.type) {
        case kUInt: { char* end = nullptr; const unsigned long u = std::strtoul(val->c_str(), &end, 10); if (end != val->c_str()) *reinterpret_cast<uint32_t*>(base) = (uint32_t)u; else ++miss; break; }
        case kFloat: { char* end = nullptr; const float x = std::strtof(val->c_str(), &end); if (end != val->c_str()) *reinterpret_cast<float*>(base) = x; else ++miss; break; }
        case kBool: { bool b; if (toBool(*val, b)) *reinterpret_cast<int32_t*>(base) = b; else ++miss; break; }
        case kString: std::snprintf(base, HRH_SESSION_STRING_SIZE, "%s", val->c_str()); break;
        }
    }
    if (missing) *missing = miss;
    return HRC_OK;
}

// Session::writeSessionFile (Session.cpp:162-214); groups in the table's order (the reference iterates an unordered_map), tinyxml2's layout.
HRC_API int hrhSessionWrite(const char* path, const HrhSession* s) {
    if (!path || !s) { g_sessionErr = "null argument"; return HRC_ERR_INVALID_VALUE; }
    std::string out = "<HeatraySession>\n";
    for (const char* group : kGroups) {
        out += std::string("    <") + group + ">\n";
        for (size_t i = 0; i < kNumVars; ++i) {
            const VarDesc& v = kVars[i];
            if (std::strcmp(v.group, group) != 0) continue;
            const char* base = reinterpret_cast<const char*>(s) + v.offset;
            char num[64]; std::string val;
            switch (v.type) {
            case kUInt: std::snprintf(num, sizeof(num), "%u", *reinterpret_cast<const uint32_t*>(base)); val = num; break;
            case kFloat: std::snprintf(num, sizeof(num), "%.8g", *reinterpret_cast<const float*>(base)); val = num; break;
            case kBool: val = *reinterpret_cast<const int32_t*>(base) ? "true" : "false"; break;
            case kString: val = escapeAttr(base); break;
            }
            out += std::string("        <") + v.name + " value=\"" + val + "\"/>\n";
        }
        out += std::string("    </") + group + ">\n";
    }
    out += "</HeatraySession>\n";
    FILE* f = std::fopen(path, "wb");
    if (!f) { g_sessionErr = "unable to save Heatray session file"; return HRC_ERR_INVALID_VALUE; }
    const bool ok = std::fwrite(out.data(), 1, out.size(), f) == out.size();
    return (std::fclose(f) == 0 && ok) ? HRC_OK : HRC_ERR_INVALID_VALUE;
}

// HeatrayRenderer::readSessionFile (HeatrayRenderer.cpp:587-667): session variables -> RenderOptions, post-processing and the orbit camera.
// Quirk kept: InteractiveMode feeds BOTH enableInteractiveMode and enableOfflineMode (:594-595).
HRC_API void hrhSessionApply(const HrhSession* s, HrhRenderOptions* o, HrcPostProcess* post) {
    if (!s) return;
    if (o) {
        o->enableInteractiveMode = s->interactiveMode; o->enableOfflineMode = s->interactiveMode;
        o->maxRenderPasses = s->maxRenderPasses; o->maxRayDepth = s->maxRayDepth; o->maxChannelValue = s->maxChannelValue;
        o->sampleMode = (int32_t)s->sampleMode; o->bokehShape = (int32_t)s->bokehShape;
        o->envExposureCompensation = s->environmentExposureCompensation; o->envThetaRotation = s->environmentThetaRotation;
        std::memcpy(o->envSolidColor, s->environmentSolidColor, sizeof(o->envSolidColor));
        // PassGenerator::changeEnvironment (PassGenerator.cpp:580-601): "<none>" removes the light, "solid color" (EnvironmentLight::SOLID_COLOR) uses the colour, anything else is a map the caller loads
        o->envEnabled = std::strcmp(s->environmentMap, "<none>") != 0;
        o->envUseSolidColor = std::strcmp(s->environmentMap, "solid color") == 0;
        o->aspectRatio = s->cameraAspectRatio; o->focusDistance = s->cameraFocusDistance; o->focalLength = s->cameraFocalLength;
        o->apertureRadius = s->cameraApertureRadius; o->fstop = s->cameraFStop;
        hrhOrbitCameraViewMatrix(s->orbitDistance, s->orbitPhi, s->orbitTheta, s->orbitTarget, o->viewMatrix);   // resetRenderer(), HeatrayRenderer.cpp:1603
        o->resetInternalState = 1;
    }
    if (post) {
        post->tonemappingEnabled = s->tonemapEnable; post->cameraExposure = std::pow(2.0f, s->exposure);        // HeatrayRenderer.h:238
        post->brightness = s->brightness; post->contrast = s->contrast; post->hue = s->hue; post->saturation = s->saturation; post->vibrance = s->vibrance;
        post->red = s->red; post->green = s->green; post->blue = s->blue; post->vignetteIntensity = s->vignetteIntensity; post->vignetteFalloff = s->vignetteFalloff;
    }
}

} // extern "C"
