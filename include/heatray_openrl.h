/* heatray_openrl.h — the OpenRL entry points Heatray calls, implemented over the CUDA backend (libheatray_openrl.so).
 *
 * The reference talks to its ray tracer through the OpenRL 1.4 C API (3rdParty/OpenRL/rl.h:383-528, OpenRL.h:47-60) wrapped by the
 * header-only openrl:: classes (Source/RLWrapper/*.h). libheatray_openrl.so exports those same symbols — the 62 `rl*` / `OpenRL*`
 * functions that Source/ actually calls (the rest of rl.h is not exported and fails at link time, loudly) — with the same
 * signatures, handle types and enum values, so the reference's own RLWrapper, PassGenerator, Scene, Mesh, Lighting, Material and
 * Light code compiles UNMODIFIED against the reference's rl.h and links against this library instead of OpenRL.lib / OpenRL.framework.
 * This header repeats the declarations for callers that do not have rl.h; a translation unit that includes rl.h must not include it.
 *
 * What the GL-style state machine turns into (heatray_b200/csrc/openrl_shim.cpp):
 *   buffers / textures / framebuffers / primitives / shaders / programs   host-side objects behind the opaque handles
 *   rlShaderString(RL_SHADER_NAME) + the `#define` prefix string           program kind (frame, PBR, glass, 4 light programs) + HRC_MAT_* flags;
 *                                                                          RLSL text is never compiled (no CUDA meaning)
 *   rlDrawElements on a primitive                                          hrcCreateMesh (attributes, indices, worldFromEntity uniform, front face,
 *                                                                          RL_PRIMITIVE_IS_OCCLUDER) + hrcCreateMaterial from the "Material" block
 *   the four light blocks, Globals, RandomSequences, ApertureSamples,      hrcSet*Lights / hrcSetEnvironment / hrcSetGlobals / hrcSetSequences /
 *   SequenceOffsets, frame-program uniforms                                hrcSetPixelTables / hrcSetInteractiveBlocks + HrcFrameParams, at rlRenderFrame
 *   rlRenderFrame / rlClear / rlViewport + colour attachment size          hrcRenderPass / hrcClear / hrcResize
 *   rlGetTexImage of the colour attachment into a RL_PIXEL_PACK_BUFFER     hrcReadPixels
 * Errors: like OpenRL, calls return nothing; rlGetError() reports and clears the first error since the last call (RL_INVALID_* for
 * misuse, RL_OUT_OF_MEMORY / RL_INTERNAL_ERROR for backend failures; hrlGetLastErrorString() has the backend's text).
 * Threading: one thread per context, as OpenRL (PassGenerator.h:229-231). */
#ifndef HEATRAY_OPENRL_H
#define HEATRAY_OPENRL_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Which backend the shim drives: a dlopen handle exporting the heatray_cuda.h C-ABI under `prefix` ("hrc"). Without this call the shim
 * loads libheatray_cuda.so from its own directory on first use and fails (OpenRLCreateContext returns NULL) when it is missing — there
 * is no CPU path. Tests bind other implementations of the same ABI here. */
int hrlBindBackend(void* dlopenHandle, const char* prefix);
/* Host-layer helpers the shim needs (per-pixel sequence ids, block hashes, offset-table continuation): a dlopen handle exporting the
 * heatray_host.h `hrh*` functions. Defaults to the backend library. */
int hrlBindHostLibrary(void* dlopenHandle);
const char* hrlGetLastErrorString(void);
/* Counters for tests: backend calls the shim has made so far (hrcCreateMesh, hrcCommitScene, hrcRenderPass, ...), by entry-point name. */
uint64_t hrlBackendCallCount(const char* entryPointName);

#ifndef RL_VERSION_1_4   /* rl.h not included: repeat its declarations (rl.h:11-49, 57-382) */
typedef int RLint; typedef unsigned int RLuint; typedef int RLboolean; typedef int RLenum; typedef unsigned int RLbitfield;
typedef float RLfloat; typedef size_t RLsize; typedef void RLvoid; typedef intptr_t RLintptr;
enum RLprimitive { RL_NULL_PRIMITIVE = 0, RL_MAX_PRIMITIVE_HANDLE = ~(RLint)0 };
typedef struct _RLbuffer* RLbuffer; typedef enum RLprimitive RLprimitive; typedef struct _RLtexture* RLtexture;
typedef struct _RLframebuffer* RLframebuffer; typedef struct _RLshader* RLshader; typedef struct _RLprogram* RLprogram;
typedef void* OpenRLContext; typedef RLintptr OpenRLContextAttribute;
typedef void (*OpenRLNotify)(RLenum error, const void* private_data, size_t private_size, const char* message, void* user_data);
#define RL_VERSION_1_4 1
#define RL_COLOR_BUFFER_BIT 0x00004000
#define RL_FALSE 0
#define RL_TRUE 1
#define RL_TRIANGLES 0x0004
#define RL_TRIANGLE_STRIP 0x0005
#define RL_ARRAY_BUFFER 0x8892
#define RL_ELEMENT_ARRAY_BUFFER 0x8893
#define RL_STATIC_DRAW 0x88E4
#define RL_TEXTURE_2D 0x0DE1
#define RL_TEXTURE_3D 0x806F
#define RL_NO_ERROR 0
#define RL_INVALID_ENUM 0x0500
#define RL_INVALID_VALUE 0x0501
#define RL_INVALID_OPERATION 0x0502
#define RL_OUT_OF_MEMORY 0x0505
#define RL_INVALID_FRAMEBUFFER_OPERATION 0x0506
#define RL_UNIMPLEMENTED 0x0507
#define RL_INTERNAL_ERROR 0x0508
#define RL_CW 0x0900
#define RL_CCW 0x0901
#define RL_UNSIGNED_BYTE 0x1401
#define RL_UNSIGNED_INT 0x1405
#define RL_FLOAT 0x1406
#define RL_RGB 0x1907
#define RL_RGBA 0x1908
#define RL_LUMINANCE 0x1909
#define RL_VERTEX_SHADER 0x8B31
#define RL_RAY_SHADER 0x10200
#define RL_FRAME_SHADER 0x10201
#define RL_PREFIX_RAY_SHADER 0x10203
#define RL_LINK_STATUS 0x8B82
#define RL_COMPILE_STATUS 0x8B81
#define RL_NEAREST 0x2600
#define RL_LINEAR 0x2601
#define RL_LINEAR_MIPMAP_LINEAR 0x2703
#define RL_TEXTURE_MAG_FILTER 0x2800
#define RL_TEXTURE_MIN_FILTER 0x2801
#define RL_TEXTURE_WRAP_S 0x2802
#define RL_TEXTURE_WRAP_T 0x2803
#define RL_TEXTURE_WRAP_R 0x8072
#define RL_REPEAT 0x2901
#define RL_CLAMP_TO_EDGE 0x812F
#define RL_FRAMEBUFFER 0x8D40
#define RL_COLOR_ATTACHMENT0 0x8CE0
#define RL_FRAMEBUFFER_COMPLETE 0x8CD5
#define RL_FRAMEBUFFER_INCOMPLETE_MISSING_ATTACHMENT 0x8CD7
#define RL_READ_ONLY 0x88B8
#define RL_WRITE_ONLY 0x88B9
#define RL_READ_WRITE 0x88BA
#define RL_PRIMITIVE 0x10000
#define RL_PRIMITIVE_IS_OCCLUDER 0x10102
#define RL_BUFFER_NAME 0x10113
#define RL_SHADER_NAME 0x10115
#define RL_COMPILE_LOG 0x10117
#define RL_LINK_LOG 0x10118
#define RL_UNIFORM_BLOCK_BUFFER 0x10122
#define RL_PIXEL_PACK_BUFFER 0x88EB

OpenRLContext OpenRLCreateContext(const OpenRLContextAttribute* attributes, OpenRLNotify notify, void* user_data);   /* OpenRL.h:47-50 */
int OpenRLDestroyContext(OpenRLContext ctx);
int OpenRLSetCurrentContext(OpenRLContext ctx);
void rlAttachShader(RLprogram program, RLshader shader);
void rlBindBuffer(RLenum target, RLbuffer buffer);
void rlBindFramebuffer(RLenum target, RLframebuffer framebuffer);
void rlBindPrimitive(RLenum target, RLprimitive primitive);
void rlBindTexture(RLenum target, RLtexture texture);
void rlBufferData(RLenum target, RLsize size, const void* data, RLenum usage);
void rlBufferParameterString(RLenum target, RLenum pname, const char* param);
RLenum rlCheckFramebufferStatus(RLenum target);
void rlClear(RLbitfield mask);
void rlCompileShader(RLshader shader);
RLprogram rlCreateProgram(void);
RLshader rlCreateShader(RLenum type);
void rlDeleteBuffers(RLsize n, const RLbuffer* buffers);
void rlDeleteFramebuffers(RLsize n, const RLframebuffer* framebuffers);
void rlDeletePrimitives(RLsize n, const RLprimitive* primitives);
void rlDeleteProgram(RLprogram program);
void rlDeleteShader(RLshader shader);
void rlDeleteTextures(RLsize n, const RLtexture* textures);
void rlDrawElements(RLenum mode, RLsize count, RLenum type, RLsize offset);
void rlFramebufferTexture2D(RLenum target, RLenum attachment, RLenum textarget, RLtexture texture, RLint level);
void rlFrontFace(RLenum mode);
void rlGenBuffers(RLsize n, RLbuffer* buffers);
void rlGenerateMipmap(RLenum target);
void rlGenFramebuffers(RLsize n, RLframebuffer* framebuffers);
void rlGenPrimitives(RLsize n, RLprimitive* primitives);
void rlGenTextures(RLsize n, RLtexture* textures);
RLint rlGetAttribLocation(RLprogram program, const char* name);
RLenum rlGetError(void);
void rlGetProgramiv(RLprogram program, RLenum pname, RLint* params);
void rlGetProgramString(RLprogram program, RLenum pname, const char** param);
void rlGetShaderiv(RLshader shader, RLenum pname, RLint* params);
void rlGetShaderString(RLshader shader, RLenum pname, const char** param);
void rlGetTexImage(RLenum target, RLint level, RLenum format, RLenum type, RLvoid* pixels);
RLint rlGetUniformBlockIndex(RLprogram program, const char* name);
RLint rlGetUniformLocation(RLprogram program, const char* name);
void rlLinkProgram(RLprogram program);
void* rlMapBuffer(RLenum target, RLenum access);
void rlPrimitiveParameter1i(RLenum target, RLenum pname, RLint param);
void rlRenderFrame(void);
void rlShaderSource(RLshader shader, RLsize count, const char* const* string, const RLsize* length);
void rlShaderString(RLshader shader, RLenum pname, const char* param);
void rlTexImage2D(RLenum target, RLint level, RLenum internalformat, RLint width, RLint height, RLint border, RLenum format, RLenum type, const void* pixels);
void rlTexImage3D(RLenum target, RLint level, RLenum internalformat, RLint width, RLint height, RLint depth, RLint border, RLenum format, RLenum type, const void* pixels);
void rlTexParameteri(RLenum target, RLenum pname, RLint param);
void rlUniform1f(RLint location, RLfloat x);
void rlUniform1i(RLint location, RLint x);
void rlUniform2fv(RLint location, RLsize count, const RLfloat* v);
void rlUniform2iv(RLint location, RLsize count, const RLint* v);
void rlUniform3fv(RLint location, RLsize count, const RLfloat* v);
void rlUniform4fv(RLint location, RLsize count, const RLfloat* v);
void rlUniform4iv(RLint location, RLsize count, const RLint* v);
void rlUniformp(RLint location, RLprimitive p);
void rlUniformt(RLint location, RLtexture t);
void rlUniformMatrix4fv(RLint location, RLsize count, RLboolean transpose, const RLfloat* value);
void rlUniformBlockBuffer(RLint index, RLbuffer buffer);
RLboolean rlUnmapBuffer(RLenum target);
void rlUseProgram(RLprogram program);
void rlVertexAttribBuffer(RLint indx, RLint size, RLenum type, RLboolean normalized, RLsize stride, RLsize offset);
void rlViewport(RLint x, RLint y, RLint width, RLint height);
#endif /* RL_VERSION_1_4 */

#ifdef __cplusplus
}
#endif
#endif /* HEATRAY_OPENRL_H */
