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
echo "oracle/_ref built"
