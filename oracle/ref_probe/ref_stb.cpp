// stb_image.h as vendored by the reference (3rdParty/stb/stb_image.h, the decoder behind util::loadTexture,
// TextureLoader.cpp:7-8,89-107), compiled unmodified. Output: oracle/_ref/libref_stb.so. Used by tests/test_images.py and
// oracle/ref_probe/make_golden.py to pin heatray_b200/csrc/host_image.cpp's own decoders.
#define STB_IMAGE_IMPLEMENTATION
#include "stb/stb_image.h"
#include <cstring>
extern "C" int ref_stb_is_hdr(const unsigned char* bytes, int n) { return stbi_is_hdr_from_memory(bytes, n); }
// Returns the number of elements written (w*h*comp), 0 on failure. `out` may be null to query the size.
extern "C" long ref_stb_load(const unsigned char* bytes, int n, int flip, int* w, int* h, int* comp, void* out, long capacityElems) {
    stbi_set_flip_vertically_on_load(flip);
    const int hdr = stbi_is_hdr_from_memory(bytes, n);
    void* px = hdr ? (void*)stbi_loadf_from_memory(bytes, n, w, h, comp, 0) : (void*)stbi_load_from_memory(bytes, n, w, h, comp, 0);
    if (!px) return 0;
    const long elems = (long)*w * *h * *comp;
    if (out && elems <= capacityElems) std::memcpy(out, px, (size_t)elems * (hdr ? 4 : 1));
    stbi_image_free(px);
    return elems;
}
