// The reference's Session class (Source/HeatrayRenderer/Session/Session.cpp, unmodified) + its vendored tinyxml2, driven through a
// flat C view so tests can (a) have the reference write a session file that host_session.cpp must read and (b) have the reference
// read a file host_session.cpp wrote. Output: oracle/_ref/libref_session.so. Test infrastructure.
#include <HeatrayRenderer/Session/Session.h>
#include <Utility/Log.h>
#include <cstring>
#include <memory>
#include <string>
namespace util { std::shared_ptr<Log> Log::m_instance = nullptr; }
namespace { struct NullLog : util::Log { NullLog() = default; void addNewItem(const std::string_view, const Type) override {} static void install() { setInstance(std::make_shared<NullLog>()); } };
struct Installer { Installer() { NullLog::install(); } } g_installer; }   // the reference logs through a singleton the viewer installs
// values: one double per session variable in table order (bools as 0/1); strings passed separately (Scene, EnvironmentMap).
static const int kTypes[] = {
#define X(group, name, type, dflt) Session::SessionVariable::k##name == Session::SessionVariable::k##name ? (#type[1] == 'B' ? 0 : #type[1] == 'U' ? 1 : #type[1] == 'F' ? 2 : 3) : 0,
    HEATRAY_SESSION_VARIABLES
#undef X
};
extern "C" int ref_session_kind(int i) { return kTypes[i]; }   // 0 bool, 1 uint, 2 float, 3 string
extern "C" int ref_session_count() { return (int)Session::SessionVariable::kNumSessionVariables; }
extern "C" int ref_session_write(const char* path, const double* values, const char* scene, const char* envMap) {
    Session s; int i = 0;
#define X(group, name, type, dflt) { const auto v = Session::SessionVariable::k##name; \
        if (kTypes[i] == 0) s.setVariableValue(v, values[i] != 0.0); else if (kTypes[i] == 1) s.setVariableValue(v, (uint32_t)values[i]); \
        else if (kTypes[i] == 2) s.setVariableValue(v, (float)values[i]); \
        else s.setVariableValue(v, std::string_view(std::strcmp(#name, "Scene") == 0 ? scene : envMap)); ++i; }
    HEATRAY_SESSION_VARIABLES
#undef X
    return s.writeSessionFile(path) ? 1 : 0;
}
extern "C" int ref_session_read(const char* path, double* values, char* scene, char* envMap) {
    Session s;
    if (!s.parseSessionFile(path)) return 0;
    int i = 0;
#define X(group, name, type, dflt) { const auto v = Session::SessionVariable::k##name; \
        if (kTypes[i] == 0) { bool b; s.getVariableValue(v, b); values[i] = b; } else if (kTypes[i] == 1) { uint32_t u; s.getVariableValue(v, u); values[i] = u; } \
        else if (kTypes[i] == 2) { float f; s.getVariableValue(v, f); values[i] = f; } \
        else { std::string str; s.getVariableValue(v, str); std::strcpy(std::strcmp(#name, "Scene") == 0 ? scene : envMap, str.c_str()); values[i] = 0; } ++i; }
    HEATRAY_SESSION_VARIABLES
#undef X
    return 1;
}
