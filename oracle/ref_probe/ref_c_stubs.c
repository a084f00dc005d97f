/* C link stubs (never called): FreeImage + the two OpenRL entry points pulled in by headers. */
void FreeImage_AllocateT(void) {} void FreeImage_DeInitialise(void) {} void FreeImage_GetBits(void) {}
void FreeImage_GetFIFFromFilename(void) {} void FreeImage_GetFormatFromFIF(void) {} void FreeImage_Initialise(void) {}
void FreeImage_Save(void) {} void FreeImage_SetOutputMessage(void) {}
#ifndef REF_SCENE_STUBS   /* libref_scene.so links the product's OpenRL entry points instead */
int rlGetError(void) { return 0; } void rlDeleteTextures(void) {}
#endif
