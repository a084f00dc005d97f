#!/bin/sh
# Builds oracle/_ref/*.so from the reference sources where they lie (/root/reference; read-only, nothing copied
# into the repo). Only runs where the reference is mounted (the build container); the GPU box uses the prebuilt files.
set -e
REF=${REF:-/root/reference}
HERE=$(cd "$(dirname "$0")" && pwd)
OUT="$HERE/../_ref"
[ -d "$REF/Source" ] || { echo "reference not mounted; skipping oracle/_ref"; exit 0; }
mkdir -p "$OUT"
cd "$HERE"
FL="-std=c++20 -O2 -fPIC -include std_shim.h -I$REF/Source -I$REF/3rdParty -I$REF/3rdParty/glm -w"
g++ $FL -shared ref_qmc.cpp -o "$OUT/libref_qmc.so"
gcc -fPIC -c ref_c_stubs.c -o /tmp/ref_c_stubs.o
g++ $FL -shared ref_lut.cpp ref_lut_stubs.cpp /tmp/ref_c_stubs.o -o "$OUT/libref_lut.so" -Wl,-z,defs
grep -v '^#include' "$REF/Resources/shaders/utility.rlsl" | sed 's/dot(st\.xy,/dot(st,/' > /tmp/ref_rlsl_utility.inc
grep -v '^#include' "$REF/Resources/shaders/brdfs.rlsl" > /tmp/ref_rlsl_brdfs.inc
g++ -std=c++20 -O2 -fPIC -shared -fsingle-precision-constant -I$REF/3rdParty -I$REF/3rdParty/glm -w ref_rlsl.cpp -o "$OUT/libref_rlsl.so"
g++ $FL -DGLM_ENABLE_EXPERIMENTAL -shared ref_host.cpp -o "$OUT/libref_host.so"
# displayGL.frag as C++: drop #version, turn uniform/in/out globals into statics, rename main; two swizzle proxies get an explicit
# vec3() (glm proxies do not convert inside operator/ and mix<>): nothing else is touched. -D_MSC_EXTENSIONS only selects glm's
# member-swizzle syntax (GLM_LANG_CXXMS_FLAG) without enabling SIMD types.
sed -e '/^#version/d' -e 's/^uniform sampler2D raytracedTexture;/struct sampler2D {}; static sampler2D raytracedTexture;/' \
    -e 's/^uniform /static /' -e 's/^in vec2 /static vec2 /' -e 's/^out vec4 /static vec4 /' -e 's/^void main()/static void display_main()/' \
    -e 's|result\.xyz / result\.w|vec3(result.xyz) / result.w|' -e 's|mix(k\.xxx,|mix(vec3(k.xxx),|' \
    "$REF/Resources/shaders/displayGL.frag" > /tmp/ref_display.inc
g++ -std=gnu++20 -D_MSC_EXTENSIONS -O2 -fPIC -shared -fsingle-precision-constant -I$REF/3rdParty -I$REF/3rdParty/glm -w ref_display.cpp -o "$OUT/libref_display.so"
g++ -std=c++17 -O2 -fPIC -shared -I$REF/3rdParty -w ref_stb.cpp -o "$OUT/libref_stb.so"
# Session.h relies on MSVC pulling in <cstdint>: -include cstdint supplies it, nothing else changes
g++ -std=c++20 -O1 -fPIC -shared -include cstdint -I$REF/Source -I$REF/3rdParty -w ref_session.cpp $REF/Source/HeatrayRenderer/Session/Session.cpp $REF/3rdParty/tinyxml2/tinyxml2.cpp -o "$OUT/libref_session.so"
# The shader programs themselves (frame shader, PBR / glass per #define set, the four light shaders) as C++: rlsl_to_cpp.py resolves the
# #include lines like ShaderCodeLoader.cpp and rewrites the RLSL-only declarations (listed in its header) into /tmp; rlsl_shim.h supplies the built-ins.
python3 rlsl_to_cpp.py /tmp/hr_ref_shaders
g++ -std=gnu++20 -D_MSC_EXTENSIONS -O2 -ffp-contract=off -fPIC -shared -fsingle-precision-constant -I/tmp/hr_ref_shaders -I. -I$REF/3rdParty -I$REF/3rdParty/glm -w ref_shaders.cpp -o "$OUT/libref_shaders.so"
# The reference's own host code for the path, unmodified, against the product's OpenRL entry points (libheatray_openrl.so): PassGenerator, Scene, Mesh,
# Lighting, the materials, the lights, the shader loader and the header-only RLWrapper classes. assimp/config.h is the one generated header the Assimp
# headers need (cmake's configure_file of config.h.in with nothing defined); TextureLoader.cpp (FreeImage + stb) and Assimp itself are not built: ref_scene.cpp
# holds the stand-ins.
SHIM="$HERE/../../heatray_b200"
if [ -f "$SHIM/libheatray_openrl.so" ]; then
  mkdir -p /tmp/hr_ref_scene/inc/assimp /tmp/hr_ref_scene/obj
  sed 's/#cmakedefine.*//' "$REF/3rdParty/assimp/include/assimp/config.h.in" > /tmp/hr_ref_scene/inc/assimp/config.h
  SFL="$FL -O1 -I$REF/3rdParty/assimp/include -I/tmp/hr_ref_scene/inc -DGLM_ENABLE_EXPERIMENTAL"
  OBJS=""
  for f in HeatrayRenderer/PassGenerator HeatrayRenderer/Scene/Scene HeatrayRenderer/Scene/Mesh HeatrayRenderer/Scene/Lighting HeatrayRenderer/Materials/PhysicallyBasedMaterial \
           HeatrayRenderer/Materials/GlassMaterial HeatrayRenderer/Materials/MultiScatterUtil HeatrayRenderer/Lights/DirectionalLight HeatrayRenderer/Lights/PointLight \
           HeatrayRenderer/Lights/SpotLight HeatrayRenderer/Lights/EnvironmentLight Utility/ShaderCodeLoader Utility/FileIO Utility/Log; do
    o=/tmp/hr_ref_scene/obj/$(basename $f).o
    g++ $SFL -c "$REF/Source/$f.cpp" -o $o &
    OBJS="$OBJS $o"
  done
  g++ $SFL -c ref_scene.cpp -o /tmp/hr_ref_scene/obj/ref_scene.o &
  gcc -fPIC -DREF_SCENE_STUBS -c ref_c_stubs.c -o /tmp/hr_ref_scene/obj/ref_c_stubs.o
  wait
  g++ -shared -o "$OUT/libref_scene.so" $OBJS /tmp/hr_ref_scene/obj/ref_scene.o /tmp/hr_ref_scene/obj/ref_c_stubs.o -L"$SHIM" -lheatray_openrl -Wl,-rpath,'$ORIGIN/../../heatray_b200' -pthread -Wl,-z,defs
fi
echo "oracle/_ref built"
