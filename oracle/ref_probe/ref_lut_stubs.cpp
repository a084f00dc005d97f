// Link stubs only (never called by the probes): the LUT generator's translation unit references
// OpenRL, FreeImage and the logger for its save/load paths, which the probes do not execute.
#include <Utility/Log.h>
#include <Utility/TextureLoader.h>
namespace util {
std::shared_ptr<Log> Log::m_instance = nullptr;
std::shared_ptr<openrl::Texture> loadTexture(const std::string_view, bool, bool) { return nullptr; }
}
