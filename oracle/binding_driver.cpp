// oracle/binding_driver.cpp — TEST INFRASTRUCTURE. Compiles integration/Renderer_B200.h (the binding a qaray maintainer adds)
// against the UNMODIFIED reference headers and drives it exactly as src/main.cpp:55-59 drives a renderer:
//     renderer->Init(); LoadScene(file); renderer->ComputeScene(renderImage, scene); renderer->Render(); renderer->Terminate();
// Built by `make -C oracle ref` into oracle/_ref/libqaref_binding.so (the reference's objects + this file, linked against
// qaray_b200/libqaray_b200.so). The access specifiers of the reference's classes are opened for this translation unit only
// (standard headers first, so they are not affected); a maintainer would add `friend struct qaray::Flattened;` instead.
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <iostream>
#include <map>
#include <memory>
#include <string>
#include <vector>
#include <unistd.h>

#define private public
#define protected public
#include "math/math.h"
#include "core/core.h"
#include "tasking/parallel_for.h"
#include "samplers/sampler_selection.h"
#include "parser/xmlload.h"
#include "Renderer_B200.h"
#undef private
#undef protected

using namespace qaray;

namespace {
RendererParam g_bparam;
Renderer_B200 *g_brenderer = nullptr;
Flattened *g_flat = nullptr;
bool g_inited = false;
void need_init()
{
  if (!g_inited) g_brenderer->Init();
  g_inited = true;
}
}  // namespace

extern "C" {

// LoadScene + ComputeScene (Init comes first in main.cpp; here it is deferred so the CPU tests can inspect the tables). Returns 1 on success (LoadScene's convention).
int qabind_load(const char *xml_path, const char *chdir_to, int spp, int bounce, int srgb)
{
  if (chdir_to && chdir_to[0] && chdir(chdir_to) != 0) return 0;
  g_bparam.SetSPPMin(spp);
  g_bparam.SetSPPMax(spp);
  g_bparam.SetSRGBFlag(srgb != 0);
  Material::maxBounce = bounce;
  if (!g_brenderer) g_brenderer = new Renderer_B200(g_bparam);  // Init() (the CUDA device) is taken when something renders
  LoadSceneInSilentMode(true);
  if (!LoadScene(xml_path)) return 0;
  g_brenderer->ComputeScene(renderImage, scene);
  delete g_flat;
  g_flat = new Flattened(scene);
  return 1;
}

// the description Flattened makes of the loaded qaray::Scene (host memory owned by this library)
const qgSceneDesc *qabind_desc(void) { return g_flat ? &g_flat->desc : nullptr; }

// Render() of the subclass into the reference's own FrameBuffer; copies the planes PixelRender fills. Returns 1 on success.
int qabind_render(uint8_t *rgb8, float *depth, uint8_t *sample_count, uint8_t *mask, uint64_t *rays)
{
  if (!g_brenderer) return 0;
  need_init();
  g_brenderer->Render();
  if (!g_brenderer->GpuScene()) return 0;
  const size_t n = (size_t) renderImage.GetWidth() * renderImage.GetHeight();
  if (rgb8) memcpy(rgb8, renderImage.GetPixels(), 3 * n);
  if (depth) memcpy(depth, renderImage.GetZBuffer(), n * sizeof(float));
  if (sample_count) memcpy(sample_count, renderImage.GetSampleCount(), n);
  if (mask) memcpy(mask, renderImage.GetMasks(), n);
  if (rays) *rays = g_brenderer->stats.closest_rays + g_brenderer->stats.shadow_rays;
  return 1;
}

// the qgScene the subclass created from the flattened description (for the ray-level parity checks)
qgScene *qabind_scene(void)
{
  if (!g_brenderer) return nullptr;
  need_init();
  if (!g_brenderer->GpuScene()) g_brenderer->Upload();
  return g_brenderer->GpuScene();
}

void qabind_terminate(void)
{
  if (g_brenderer) g_brenderer->Terminate();
}

}  // extern "C"
