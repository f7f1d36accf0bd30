// libmitsuba-hw is reduced to basicshader.cpp (the Constant*Texture classes); the two Renderer members it references
// belong to the OpenGL layer ("unavailable on this box") and are never called on the software path.
#include <mitsuba/hw/renderer.h>
MTS_NAMESPACE_BEGIN
Shader *Renderer::registerShaderForResource(const HWResource *) { SLog(EError, "no hardware renderer in the oracle build"); return NULL; }
void Renderer::unregisterShaderForResource(const HWResource *) { SLog(EError, "no hardware renderer in the oracle build"); }
MTS_NAMESPACE_END
