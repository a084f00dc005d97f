#!/usr/bin/env python
"""RLSL -> C++ text rewriter for oracle/_ref/libref_shaders.so (TEST INFRASTRUCTURE, build container only).

Reads the reference's shader programs WHERE THEY LIE (/root/reference/Resources/shaders/*.rlsl, read-only), resolves their
`#include` lines exactly as the reference's loader does (Source/Utility/ShaderCodeLoader.cpp:28-64: depth first, every file once,
an included file's text goes BEFORE the text of the file that names it) and rewrites the handful of RLSL-only constructs into
C++ that compiles against glm + oracle/ref_probe/rlsl_shim.h. Output goes to a scratch directory (default /tmp/hr_ref_shaders),
never into the repository. Shader arithmetic is not touched; every rewrite is listed here:

  uniformblock X { ... };         ->  static struct X_block { ... } X;
  T name[1];  (inside a block)    ->  const T* name;          (OpenRL blocks end in an unsized array; the host binds a buffer)
  uniform T name;                 ->  static T name;
  varying T name;                 ->  static T name;
  attribute T name;               ->  static T name;          (vertex.rlsl)
  rayattribute T name;            ->  (dropped: the four attributes are members of the shim's Ray, rayAttributes.rlsl:8-11)
  out T name   (parameters)       ->  T& name
  void main() / void setup()      ->  static void rlsl_main() / static void rlsl_setup()
  free functions / globals        ->  left as they are (each program is compiled inside its own C++ namespace)
and three glm-only token patches (glm swizzle proxies do not take part in template deduction or in constructors from proxies
of another vector type; the value is the same vector):
  dot(st.xy,                      ->  dot(st,                 (utility.rlsl:17, st is already a vec2)
  vec2(blockIndex.yx)             ->  vec2(ivec2(blockIndex.yx))
  ivec2(texture2D(...).xy)        ->  ivec2(vec2(texture2D(...).xy))
  v *= <expr>.rgb;                ->  v *= vec3(<expr>.rgb);   (regex; physicallyBased.rlsl:62, glass.rlsl:155)
The literal ones are PATCHES below; a patch that stops matching the reference text aborts the build.
"""
from __future__ import annotations

import re
import sys
from pathlib import Path

REF = Path("/root/reference/Resources/shaders")

# (old, new) literal replacements applied to the concatenated program text; each must match at least once somewhere
# in the set of programs (checked at the end) so a stale entry is noticed.
PATCHES = [
    ("dot(st.xy,", "dot(st,"),
    ("vec2(blockIndex.yx)", "vec2(ivec2(blockIndex.yx))"),
    ("ivec2(texture2D(interactiveBlockSamplesTexture, blockSampleUV).xy)", "ivec2(vec2(texture2D(interactiveBlockSamplesTexture, blockSampleUV).xy))"),
]


def load_program(name: str) -> str:
    """ShaderCodeLoader.cpp:28-64."""
    seen: set[str] = set()
    chunks: list[str] = []

    def rec(fn: str) -> None:
        src = (REF / fn).read_text()
        while True:
            off = src.find("#include")
            if off < 0:
                break
            q0 = src.find('"', off)
            q1 = src.find('"', q0 + 1)
            inc = src[q0 + 1:q1]
            src = src[:off] + src[q1 + 1:]
            if inc in seen:
                continue
            seen.add(inc)
            rec(inc)
        chunks.append(f"// ---- {fn} ----\n" + src + "\n")

    rec(name)
    return "".join(chunks)


def rewrite(text: str, used: set[int]) -> str:
    # uniformblock X { ... };  (blocks never nest)
    def block(m: re.Match) -> str:
        name, body = m.group(1), m.group(2)
        body = re.sub(r"(\w+)\s+(\w+)\[1\];", r"const \1* \2;", body)
        return f"static struct {name}_block {{{body}}} {name};"
    text = re.sub(r"uniformblock\s+(\w+)\s*\{(.*?)\};", block, text, flags=re.S)
    text = re.sub(r"^uniform\s+", "static ", text, flags=re.M)
    text = re.sub(r"^varying\s+", "static ", text, flags=re.M)
    text = re.sub(r"^attribute\s+", "static ", text, flags=re.M)
    text = re.sub(r"^rayattribute\s+[^;]*;", "", text, flags=re.M)
    text = re.sub(r"\bout\s+(vec[234]|float|int|primitive)\s+(\w+)", r"\1& \2", text)
    text = re.sub(r"^void\s+main\s*\(\s*\)", "static void rlsl_main()", text, flags=re.M)
    text = re.sub(r"^void\s+setup\s*\(\s*\)", "static void rlsl_setup()", text, flags=re.M)
    text = re.sub(r"(\w+) \*= ([^;=]+)\.(rgb|xyz);", r"\1 *= vec3(\2.\3);", text)      # vec3 *= <swizzle proxy>
    for i, (old, new) in enumerate(PATCHES):
        if old in text:
            used.add(i)
            text = text.replace(old, new)
    return text


# Feature-flag sets (HRC_MAT_* bits, include/heatray_cuda.h) each material program is compiled for: every #define of
# PhysicallyBasedMaterial.cpp:49-113 / GlassMaterial.cpp:40-74 alone and in the combinations the scenes use.
DS, AM, BCT, MRT, EMT, NM, CCT, CCRT, CCNM, VC = (1 << i for i in range(10))
PBR_VARIANTS = [0, DS, AM | BCT, DS | AM | BCT, DS | BCT, DS | MRT, DS | EMT, DS | NM, DS | CCT, DS | CCRT, DS | CCNM, DS | VC, VC,
                DS | BCT | NM, DS | BCT | MRT | CCT | CCRT, DS | BCT | MRT | NM, BCT | MRT | NM | VC,
                DS | BCT | MRT | EMT | NM | CCT | CCRT | CCNM | VC, DS | AM | BCT | MRT | EMT | NM | CCT | CCRT | CCNM | VC]
GLASS_VARIANTS = [0, BCT, MRT, NM, BCT | MRT, BCT | MRT | NM]
DEFINES = [(DS, "DOUBLE_SIDED"), (AM, "ALPHA_MASK"), (BCT, "HAS_BASE_COLOR_TEXTURE"), (MRT, "HAS_METALLIC_ROUGHNESS_TEXTURE"), (EMT, "HAS_EMISSIVE_TEXTURE"),
           (NM, "HAS_NORMALMAP"), (CCT, "HAS_CLEARCOAT_TEXTURE"), (CCRT, "HAS_CLEARCOAT_ROUGHNESS_TEXTURE"), (CCNM, "HAS_CLEARCOAT_NORMALMAP"), (VC, "VERTEX_COLORS")]
ANY_TEXTURE = BCT | MRT | EMT | NM | CCT | CCRT | CCNM


def variants_text() -> str:
    out = []
    table = {1: [], 2: []}
    for kind, prog, flags_list in ((1, "physicallyBased", PBR_VARIANTS), (2, "glass", GLASS_VARIANTS)):
        for f in flags_list:
            defs = [name for bit, name in DEFINES if f & bit]
            if f & ANY_TEXTURE:                       # PhysicallyBasedMaterial.cpp:104-109 / GlassMaterial.cpp:66-71
                if f & (NM | CCNM):
                    defs.append("USE_TANGENT_SPACE")
                defs.append("HAS_TEXTURES")
            ns = f"{prog}_{f:03x}"
            out.append(f"namespace {ns} {{")
            out += [f"#define {d}" for d in defs]
            out.append(f"#define PROGRAM_KIND {kind}")
            out.append(f'#include "{prog}.inc"')
            out.append('#include "ref_shaders_program.inc"')
            out.append("#undef PROGRAM_KIND")
            out += [f"#undef {d}" for d in defs]
            out.append("}")
            table[kind].append((f, ns))
    # vertex.rlsl under the #define sets Material::build hands the vertex shader (bit 1 HAS_TEXTURES, 2 USE_TANGENT_SPACE, 4 VERTEX_COLORS)
    vtab = []
    for f in (0, 1, 3, 4, 5, 7):
        defs = [name for bit, name in ((1, "HAS_TEXTURES"), (2, "USE_TANGENT_SPACE"), (4, "VERTEX_COLORS")) if f & bit]
        ns = f"vertex_{f}"
        out.append(f"namespace {ns} {{")
        out += [f"#define {d}" for d in defs]
        out.append('#include "vertex.inc"')
        out.append('#include "ref_vertex_program.inc"')
        out += [f"#undef {d}" for d in defs]
        out.append("}")
        vtab.append((f, ns))
    out.append("static const VertexVariant kVertexVariants[] = {" + ", ".join(f"{{{f}u, {ns}::run}}" for f, ns in vtab) + "};")
    for kind, name in ((1, "kPbrVariants"), (2, "kGlassVariants")):
        out.append(f"static const MaterialVariant {name}[] = {{" + ", ".join(f"{{0x{f:x}u, {ns}::run}}" for f, ns in table[kind]) + "};")
    return "\n".join(out) + "\n"


def main() -> None:
    out = Path(sys.argv[1] if len(sys.argv) > 1 else "/tmp/hr_ref_shaders")
    out.mkdir(parents=True, exist_ok=True)
    used: set[int] = set()
    for prog in ("perspective", "physicallyBased", "glass", "environmentLight", "directionalLight", "pointLight", "spotLight", "vertex"):
        (out / f"{prog}.inc").write_text(rewrite(load_program(prog + ".rlsl"), used))
    (out / "variants.inc").write_text(variants_text())
    stale = [PATCHES[i][0] for i in range(len(PATCHES)) if i not in used]
    if stale:
        raise SystemExit(f"rlsl_to_cpp: patches that no longer match the reference text: {stale}")


if __name__ == "__main__":
    main()
