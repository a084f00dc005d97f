// Force-included before reference headers: libstdc++ lacks std::sqrtf/cosf/sinf..., which the
// reference uses (Random.h:278-281,302). Adds names only; no reference code is altered.
#include <cmath>
#include <cstring>
#include <cstdint>
#include <limits>
namespace std { using ::sqrtf; using ::cosf; using ::sinf; using ::tanf; using ::powf; using ::fabsf; }
