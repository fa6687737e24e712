// oracle/ref_driver.cpp — TEST INFRASTRUCTURE, not product code.
//
// Headless C-ABI driver around the UNMODIFIED reference renderer (wilsonCernWq/qaray).
// It is compiled by oracle/Makefile together with the reference's own sources where they lie
// under /root/reference (nothing of the reference is copied into this repository); the only
// outputs are oracle/_ref/libqaref_parity.so and oracle/_ref/qaref_perf.
//
// Everything here goes through the reference's public classes:
//   LoadScene                      src/parser/xmlload.cpp:71
//   Renderer::ComputeScene         src/renderers/renderer.cpp:71
//   Renderer::ThreadRender         src/renderers/renderer.cpp:370
//   Renderer::PixelRender          src/renderers/renderer.cpp:302
//   Scene::TraceNodeNormal/Shadow  src/scene/scene.cpp:35-74
//   Material::Shade                src/core/material.h:49
//
// The per-thread RNG of the reference (Sampler_Marsaglia, seeded from time()) keeps its seed
// private; to make single-threaded runs reproducible this translation unit opens the class
// (access specifiers do not change layout) and sets the xorshift32 state directly.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cstdint>
#include <map>
#include <vector>
#include <string>
#include <chrono>
#include <unistd.h>

#include <iostream>
#include <atomic>
#include <algorithm>
#include <functional>
#include <cmath>
#include "math/math.h"
#include "core/core.h"
#include "tasking/parallel_for.h"

#define private public
#include "samplers/Sampler_Marsaglia.h"
#undef private
// TriObj keeps its cyBVH private; qaref_bvh_dump reads it through cyBVH's public node accessors
#define private public
#include "objects/objects.h"
#undef private

#include "scene/scene.h"
#include "renderers/renderer.h"
#include "parser/xmlload.h"
#include "objects/objects.h"
#include "lights/lights.h"
#include "materials/materials.h"
#include "samplers/sampler_selection.h"
#include "tiny_obj_loader.h"  // the reference's vendored tinyobjloader v1.1.0 (external/tinyobjloader), implementation in mesh/TriMesh.cpp

#ifdef QAREF_COUNTERS
// defined by the patched copy of scene/scene.cpp (see oracle/Makefile)
extern "C" unsigned long long qaref_count_closest;
extern "C" unsigned long long qaref_count_shadow;
#endif

using namespace qaray;

namespace {

class HeadlessRenderer : public Renderer {
 public:
  explicit HeadlessRenderer(RendererParam &p) : Renderer(p) {}
  void Render() override { ThreadRender(); }
  // quiet timers: the caller times ThreadRender itself
  void StartTimer() override {}
  void StopTimer() override {}
  void KillTimer() override {}
  // read-only views of the protected frame ComputeScene derived
  const Point3 &U() const { return screenU; }
  const Point3 &V() const { return screenV; }
  const Point3 &A() const { return screenA; }
  const Point3 &SX() const { return screenX; }
  const Point3 &SY() const { return screenY; }
  float Dof() const { return dof; }
  size_t W() const { return pixelW; }
  size_t H() const { return pixelH; }
};

RendererParam g_param;
HeadlessRenderer *g_renderer = nullptr;
std::vector<Node *> g_nodes;          // pre-order (the order TraceNodeNormal visits)
std::vector<int> g_parent;
std::map<const Node *, int> g_nodeIndex;
bool g_loaded = false;

void CollectNodes(Node *n, int parent)
{
  int me = (int) g_nodes.size();
  g_nodes.push_back(n);
  g_parent.push_back(parent);
  g_nodeIndex[n] = me;
  for (int c = 0; c < n->GetNumChild(); ++c) CollectNodes(n->GetChild(c), me);
}

}  // namespace

extern "C" {

// record layout shared with the python side (oracle/refapi.py)
struct QarefHit {
  int32_t hit;          // 0/1
  int32_t node;         // pre-order node index of hInfo.c.node (-1 on miss)
  int32_t face;         // TriObj face id of the accepted triangle (-1 for sphere/plane); needs the QAREF_FACEID patch
  int32_t mtlID;
  int32_t front;
  int32_t hasTexture;
  float z;
  float p[3];
  float N[3];
  float uvw[3];
  float duvw[6];
};

int qaref_load(const char *xml_path, const char *chdir_to, int silent)
{
  if (chdir_to && chdir_to[0]) {
    if (chdir(chdir_to) != 0) { return 0; }
  }
  LoadSceneInSilentMode(silent != 0);
  if (!LoadScene(xml_path)) { return 0; }
  if (!g_renderer) g_renderer = new HeadlessRenderer(g_param);
  g_renderer->ComputeScene(renderImage, scene);
  g_nodes.clear();
  g_parent.clear();
  g_nodeIndex.clear();
  CollectNodes(&scene.rootNode, -1);
  g_loaded = true;
  return 1;
}

void qaref_set_params(int sppMin, int sppMax, int maxBounce, int srgb, int threads)
{
  g_param.SetSPPMin(sppMin);
  g_param.SetSPPMax(sppMax);
  g_param.SetSRGBFlag(srgb != 0);
  Material::maxBounce = maxBounce;
  if (threads > 0) tasking::set_num_of_threads((size_t) threads);
}

// xorshift32 state of the calling thread's sampler (serial runs only)
void qaref_seed(uint32_t s)
{
  Sampler_Marsaglia &smp = rng->local();
  smp.initialized = true;
  smp.seed[0] = s;
}

uint32_t qaref_rng_state(void) { return rng->local().seed[0]; }

int qaref_num_nodes(void) { return (int) g_nodes.size(); }
int qaref_num_lights(void) { return (int) scene.lights.size(); }

// per node: meta[4] = {parent, objType (0 none, 1 sphere, 2 plane, 3 obj), numChild, numFaces}
//           xf[21]  = {tm column-major 9, itm column-major 9, pos 3}
void qaref_dump_nodes(int32_t *meta, float *xf)
{
  for (size_t i = 0; i < g_nodes.size(); ++i) {
    Node *n = g_nodes[i];
    Object *o = n->GetNodeObj();
    int type = 0, nf = 0;
    if (o == &theSphere) type = 1;
    else if (o == &thePlane) type = 2;
    else if (o != nullptr) { type = 3; nf = (int) static_cast<TriObj *>(o)->NF(); }
    meta[4 * i + 0] = g_parent[i];
    meta[4 * i + 1] = type;
    meta[4 * i + 2] = n->GetNumChild();
    meta[4 * i + 3] = nf;
    const Matrix3 &tm = n->GetTransform();
    const Matrix3 &itm = n->GetInverseTransform();
    const Point3 &pos = n->GetPosition();
    for (int c = 0; c < 3; ++c)
      for (int r = 0; r < 3; ++r) {
        xf[21 * i + 3 * c + r] = tm[c][r];
        xf[21 * i + 9 + 3 * c + r] = itm[c][r];
      }
    xf[21 * i + 18] = pos.x;
    xf[21 * i + 19] = pos.y;
    xf[21 * i + 20] = pos.z;
  }
}

// The reference's own BVH of node `node_index`'s TriObj (cyBVH, built by TriObj::Load with 4 elements per leaf), read through
// BVH's public accessors (src/ext/cyBVH.h:75-128) into cyBVH's array form: slot 0 unused, root 1; data = leaf: bit 31 |
// (count - 1) << 28 | element offset, internal: first child. Returns the number of node slots (largest id + 1), 0 when the
// node holds no mesh; *num_elems = NF.
int64_t qaref_bvh_dump(int node_index, int64_t cap_nodes, float *boxes6, uint32_t *data, int64_t cap_elems, uint32_t *elements, int64_t *num_elems)
{
  if (node_index < 0 || node_index >= (int) g_nodes.size()) return 0;
  Object *o = g_nodes[node_index]->GetNodeObj();
  if (o == nullptr || o == &theSphere || o == &thePlane) return 0;
  TriObj *t = static_cast<TriObj *>(o);
  if (num_elems) *num_elems = (int64_t) t->NF();
  if (t->NF() == 0) return 0;
  const auto &bvh = t->bvh;
  // element offsets are relative to the first element: the leftmost leaf holds offset 0
  unsigned first = bvh.GetRootNodeID();
  while (!bvh.IsLeafNode(first)) first = bvh.GetFirstChildNode(first);
  const unsigned int *base = bvh.GetNodeElements(first);
  std::vector<unsigned> stack(1, bvh.GetRootNodeID());
  int64_t slots = 0;
  while (!stack.empty()) {
    const unsigned id = stack.back();
    stack.pop_back();
    slots = std::max<int64_t>(slots, (int64_t) id + 1);
    if ((int64_t) id < cap_nodes) {
      if (boxes6) memcpy(boxes6 + 6 * id, bvh.GetNodeBounds(id), 6 * sizeof(float));
      if (bvh.IsLeafNode(id)) {
        const unsigned cnt = bvh.GetNodeElementCount(id);
        const unsigned off = (unsigned) (bvh.GetNodeElements(id) - base);
        if (data) data[id] = 0x80000000u | ((cnt - 1) << 28) | off;
        for (unsigned k = 0; k < cnt; ++k)
          if (elements && (int64_t) off + k < cap_elems) elements[off + k] = bvh.GetNodeElements(id)[k];
      } else if (data) {
        data[id] = bvh.GetFirstChildNode(id);
      }
    }
    if (!bvh.IsLeafNode(id)) {
      stack.push_back(bvh.GetFirstChildNode(id));
      stack.push_back(bvh.GetSecondChildNode(id));
    }
  }
  return slots;
}

// out[14] = pos3, dir3, up3, fovy, focal, dof, W, H
void qaref_camera(float *out)
{
  const Camera &c = scene.camera;
  out[0] = c.pos.x; out[1] = c.pos.y; out[2] = c.pos.z;
  out[3] = c.dir.x; out[4] = c.dir.y; out[5] = c.dir.z;
  out[6] = c.up.x;  out[7] = c.up.y;  out[8] = c.up.z;
  out[9] = c.fovy; out[10] = c.focalDistance; out[11] = c.depthOfField;
  out[12] = (float) c.imgWidth; out[13] = (float) c.imgHeight;
}

// the image-plane frame Renderer::ComputeScene derived (renderer.cpp:77-91): U3 V3 A3
void qaref_screen(float *out)
{
  const Point3 &U = g_renderer->U(), &V = g_renderer->V(), &A = g_renderer->A();
  out[0] = U.x; out[1] = U.y; out[2] = U.z;
  out[3] = V.x; out[4] = V.y; out[5] = V.z;
  out[6] = A.x; out[7] = A.y; out[8] = A.z;
}

// primary DiffRay of pixel (i,j), sample s, exactly as PixelRender builds it (renderer.cpp:312-326)
static DiffRay PrimaryRay(int i, int j, int s, Point3 *texposOut)
{
  const Point3 texpos = Point3(Halton(s, 11), Halton(s, 13), 0.f) + Point3(i, j, 0.f);
  const Point3 &A = g_renderer->A(), &U = g_renderer->U(), &V = g_renderer->V();
  const Point3 cpt = A + texpos.x * U + texpos.y * V;
  const Point3 xpt = A + (texpos.x + DiffRay::dx) * U + texpos.y * V;
  const Point3 ypt = A + texpos.x * U + (texpos.y + DiffRay::dy) * V;
  Point3 campos = scene.camera.pos;
  if (g_renderer->Dof() > 0.1f) {  // lens sample (renderer.cpp:319-322): the sample's first two random numbers
    SuperSamplerHalton lens(Color3f(0.f), 1, 1);
    const Point3 dofSample = lens.NewDofSample(g_renderer->Dof());
    campos += dofSample.x * g_renderer->SX() + dofSample.y * g_renderer->SY();
  }
  DiffRay ray(campos, cpt - campos, campos, xpt - campos, campos, ypt - campos);
  ray.Normalize();
  if (texposOut) *texposOut = texpos;
  return ray;
}

void qaref_primary_ray(int i, int j, int s, float *out18)
{
  DiffRay r = PrimaryRay(i, j, s, nullptr);
  const Ray *rr[3] = {&r.c, &r.x, &r.y};
  for (int k = 0; k < 3; ++k) {
    out18[6 * k + 0] = rr[k]->p.x; out18[6 * k + 1] = rr[k]->p.y; out18[6 * k + 2] = rr[k]->p.z;
    out18[6 * k + 3] = rr[k]->dir.x; out18[6 * k + 4] = rr[k]->dir.y; out18[6 * k + 5] = rr[k]->dir.z;
  }
}

static void FillHit(bool hit, const DiffHitInfo &h, QarefHit *o)
{
  memset(o, 0, sizeof(*o));
  o->hit = hit ? 1 : 0;
  o->node = -1;
  o->face = -1;
  o->z = h.c.z;
  if (!hit) return;
  auto it = g_nodeIndex.find(h.c.node);
  o->node = it == g_nodeIndex.end() ? -1 : it->second;
#ifdef QAREF_FACEID
  o->face = h.c.faceID;
#endif
  o->mtlID = h.c.mtlID;
  o->front = h.c.hasFrontHit ? 1 : 0;
  o->hasTexture = h.c.hasTexture ? 1 : 0;
  o->p[0] = h.c.p.x; o->p[1] = h.c.p.y; o->p[2] = h.c.p.z;
  o->N[0] = h.c.N.x; o->N[1] = h.c.N.y; o->N[2] = h.c.N.z;
  o->uvw[0] = h.c.uvw.x; o->uvw[1] = h.c.uvw.y; o->uvw[2] = h.c.uvw.z;
  for (int k = 0; k < 2; ++k) {
    o->duvw[3 * k + 0] = h.c.duvw[k].x;
    o->duvw[3 * k + 1] = h.c.duvw[k].y;
    o->duvw[3 * k + 2] = h.c.duvw[k].z;
  }
}

// rays: [n][18] = c.p c.dir x.p x.dir y.p y.dir ; hasDiff[n] ; directions are used as given (not renormalised)
void qaref_trace_closest(int64_t n, const float *rays, const uint8_t *hasDiff, QarefHit *out)
{
  for (int64_t k = 0; k < n; ++k) {
    const float *r = rays + 18 * k;
    DiffRay ray(Point3(r[0], r[1], r[2]), Point3(r[3], r[4], r[5]),
                Point3(r[6], r[7], r[8]), Point3(r[9], r[10], r[11]),
                Point3(r[12], r[13], r[14]), Point3(r[15], r[16], r[17]));
    ray.hasDiffRay = hasDiff ? (hasDiff[k] != 0) : true;
    DiffHitInfo h;
    h.Init();
#ifdef QAREF_FACEID
    h.c.faceID = -1;
#endif
    bool hit = scene.TraceNodeNormal(scene.rootNode, ray, h);
    FillHit(hit, h, out + k);
  }
}

// rays: [n][6] = p dir ; tmax[n] ; occluded[n]  — GenLight::Shadow's call (lights.cpp:39-48)
void qaref_trace_shadow(int64_t n, const float *rays, const float *tmax, uint8_t *occluded)
{
  for (int64_t k = 0; k < n; ++k) {
    const float *r = rays + 6 * k;
    Ray ray(Point3(r[0], r[1], r[2]), Point3(r[3], r[4], r[5]));
    HitInfo h;
    h.z = tmax[k];
    occluded[k] = scene.TraceNodeShadow(scene.rootNode, ray, h) ? 1 : 0;
  }
}

// One radiance sample of pixel (i,j), sample index s — the body of PixelRender's loop
// (renderer.cpp:312-339) through public APIs. Returns hasHit; rgb = localColor; z = depth.
int qaref_sample(int i, int j, int s, float *rgb, float *z)
{
  Point3 texpos;
  DiffRay ray = PrimaryRay(i, j, s, &texpos);
  DiffHitInfo hInfo;
  hInfo.c.z = BIGFLOAT;
  bool hasHit = scene.TraceNodeNormal(scene.rootNode, ray, hInfo);
  Color3f c;
  if (hasHit) {
    c = hInfo.c.node->GetMaterial()->Shade(ray, hInfo, scene.lights, Material::maxBounce);
  } else {
    const float u = texpos.x / g_renderer->W();
    const float v = texpos.y / g_renderer->H();
    c = scene.background.Sample(Point3(u, v, 0.f));
  }
  rgb[0] = c.r; rgb[1] = c.g; rgb[2] = c.b;
  if (z) *z = hasHit ? hInfo.c.z : BIGFLOAT;
  return hasHit ? 1 : 0;
}

// Counter-based stream (needs the P4 patch of oracle/Makefile): the calling thread's sampler returns draw n of
// the stream `key` from now on. The key derivation restates qaray_b200/csrc/qmath.h rng_key.
static uint32_t qaref_mix(uint32_t h)
{
  h ^= h >> 16; h *= 0x85EBCA6Bu; h ^= h >> 13; h *= 0xC2B2AE35u; h ^= h >> 16;
  return h;
}
uint32_t qaref_path_key(uint32_t pixel, uint32_t sample, uint32_t seed)
{
  return qaref_mix(qaref_mix(pixel * 0x9E3779B1u + 0x7F4A7C15u) ^ (sample * 0x85EBCA77u) ^ (seed * 0xC2B2AE3Du + 0x27D4EB2Fu));
}
void qaref_seed_path(uint32_t key)
{
  Sampler_Marsaglia &smp = rng->local();
  smp.initialized = true;
  smp.seed[0] = key;
  smp.seed[1] = 0;
  smp.seed[2] = 0xC0FFEE42u;
}
// n samples (i, j, s): the reference's own Shade recursion fed with the counter-based stream of (pixel, s, seed)
void qaref_samples(int64_t n, const int32_t *pix, uint32_t seed, float *rgb, float *z, int32_t *hit)
{
  const int W = (int) g_renderer->W();
  for (int64_t k = 0; k < n; ++k) {
    const int i = pix[3 * k], j = pix[3 * k + 1], s = pix[3 * k + 2];
    qaref_seed_path(qaref_path_key((uint32_t) (j * W + i), (uint32_t) s, seed));
    float zz = 0.f;
    const int h = qaref_sample(i, j, s, rgb + 3 * k, &zz);
    if (z) z[k] = zz;
    if (hit) hit[k] = h;
  }
}


// ---- photon / caustics maps (the reference's -use-photon-map path) ------------------------------------------------------------
// The body of the emission loops of Renderer::ComputeScene (renderer.cpp:143-196 photon map, :214-266 caustics map) for
// emissions [first, first + n), each fed with the counter-based stream of (emission, map, seed) (P4 patch) and without the
// map-full stop: counts[e] photons, rec[(e * bounceMax + k) * 9] = position, incoming direction, power before normalisation.
int qaref_emit_photons(int which, uint32_t first, int64_t n, uint32_t seed, int bounceMax, int32_t *counts, float *rec)
{
  std::vector<Light *> photonLights;
  for (auto light : scene.lights) {
    if (light->IsPhotonSource()) { photonLights.push_back(light); }
  }
  if (photonLights.empty()) return 0;
  const qaFLOAT lightScale = 1.f / static_cast<qaFLOAT>(photonLights.size());
  for (int64_t e = 0; e < n; ++e) {
    qaref_seed_path(qaref_path_key(first + (uint32_t) e, 0x50484F54u + (uint32_t) which, seed));
    Light *light;
    if (photonLights.size() == 1) { light = photonLights[0]; }
    else {
      qaFLOAT r;
      rng->local().Get1f(r);
      size_t id = which == 0 ? (size_t) MIN(FLOOR(r * photonLights.size()), photonLights.size() - 1)
                             : MIN(static_cast<size_t>(CEIL(r * photonLights.size())), photonLights.size() - 1);
      light = photonLights[id];
    }
    DiffRay ray = light->RandomPhoton();
    ray.Normalize();
    DiffHitInfo hInfo;
    hInfo.Init();
    Color3f intensity = light->GetPhotonIntensity(ray.c.dir) * lightScale;
    size_t bounce = 0;
    int cnt = 0;
    while (bounce < (size_t) bounceMax) {
      if (scene.TraceNodeNormal(scene.rootNode, ray, hInfo)) {
        const Material *mtl = hInfo.c.node->GetMaterial();
        if (!mtl) break;
        if (mtl->IsPhotonSurface(0) && bounce != 0 && (which == 0 || !hInfo.c.hasDiffuseHit) && cnt < bounceMax) {
          float *o = rec + ((size_t) e * bounceMax + cnt) * 9;
          o[0] = hInfo.c.p.x; o[1] = hInfo.c.p.y; o[2] = hInfo.c.p.z;
          o[3] = ray.c.dir.x; o[4] = ray.c.dir.y; o[5] = ray.c.dir.z;
          o[6] = intensity.r; o[7] = intensity.g; o[8] = intensity.b;
          ++cnt;
        }
        if (mtl->RandomPhotonBounce(ray, intensity, hInfo)) {
          bool diffuseHit = hInfo.c.hasDiffuseHit;
          ++bounce;
          ray.Normalize();
          hInfo.Init();
          if (which == 1) hInfo.c.hasDiffuseHit = (diffuseHit || mtl->IsPhotonSurface(0));
        } else { break; }
      } else { break; }
    }
    counts[e] = cnt;
  }
  return 1;
}

// cyPhotonMap itself: a map built from the given photons (AddPhoton quantises them, ScalePhotonPowers, the balanced kd-tree of
// PrepareForIrradianceEstimation) queried with EstimateIrradiance<100>(..., FILTER_TYPE_QUADRATIC) as Shade does
// (MtlBlinn_PhotonMap.cpp:430-433). Also returns the photons as the map hands them back (GetPower / GetDirection).
void qaref_photon_estimate(int64_t np, const float *pos, const float *dir, const float *power, float scale, float radius, int64_t nq,
                           const float *qpos, const float *qnrm, float *irrad, float *dout, float *back_dir, float *back_power)
{
  cyPhotonMap map;
  map.AllocatePhotons((unsigned) np);
  for (int64_t i = 0; i < np; ++i)
    map.AddPhoton(Point3(pos[3 * i], pos[3 * i + 1], pos[3 * i + 2]), Point3(dir[3 * i], dir[3 * i + 1], dir[3 * i + 2]),
                  Color3f(power[3 * i], power[3 * i + 1], power[3 * i + 2]));
  map.ScalePhotonPowers(scale);
  for (int64_t i = 0; i < np && back_dir && back_power; ++i) {  // before the kd-tree reorders them
    Point3 d;
    Color3f c;
    map[(unsigned) i].GetDirection(d);
    map[(unsigned) i].GetPower(c);
    back_dir[3 * i] = d.x; back_dir[3 * i + 1] = d.y; back_dir[3 * i + 2] = d.z;
    back_power[3 * i] = c.r; back_power[3 * i + 1] = c.g; back_power[3 * i + 2] = c.b;
  }
  map.PrepareForIrradianceEstimation();
  for (int64_t q = 0; q < nq; ++q) {
    Color3f I;
    Point3 D;
    const Point3 N(qnrm[3 * q], qnrm[3 * q + 1], qnrm[3 * q + 2]);
    map.EstimateIrradiance<100>(I, D, radius, Point3(qpos[3 * q], qpos[3 * q + 1], qpos[3 * q + 2]), &N, 1.f, cyPhotonMap::FILTER_TYPE_QUADRATIC);
    irrad[3 * q] = I.r; irrad[3 * q + 1] = I.g; irrad[3 * q + 2] = I.b;
    dout[3 * q] = D.x; dout[3 * q + 1] = D.y; dout[3 * q + 2] = D.z;
  }
}

// RendererParam's photon switches (src/main.cpp:35-40); takes effect at the next qaref_load (ComputeScene builds the maps)
void qaref_set_photon(int use, int photonMapSize, int causticsMapSize)
{
  g_param.SetPhotonMappingFlag(use != 0);
  if (photonMapSize > 0) g_param.SetPhotonMapSize((size_t) photonMapSize);
  if (causticsMapSize > 0) g_param.SetCausticsMapSize((size_t) causticsMapSize);
}
// the maps ComputeScene built, as the estimate reads them: n photons -> pos, dir, power (each 3n floats); returns the count
int64_t qaref_photon_map(int which, int64_t cap, float *pos, float *dir, float *power)
{
  cyPhotonMap &m = which == 0 ? scene.photonmap.map : scene.causticsmap.map;
  const int64_t n = (int64_t) m.NumPhotons();
  for (int64_t i = 0; i < n && i < cap; ++i) {
    Point3 d;
    Color3f c;
    m[(unsigned) i].GetDirection(d);
    m[(unsigned) i].GetPower(c);
    pos[3 * i] = m[(unsigned) i].position.x; pos[3 * i + 1] = m[(unsigned) i].position.y; pos[3 * i + 2] = m[(unsigned) i].position.z;
    dir[3 * i] = d.x; dir[3 * i + 1] = d.y; dir[3 * i + 2] = d.z;
    power[3 * i] = c.r; power[3 * i + 1] = c.g; power[3 * i + 2] = c.b;
  }
  return n;
}

// tinyobjloader itself, as TriMesh::LoadFromFileObj calls it (src/mesh/TriMesh.cpp:63-100: triangulate = true): the attribute
// arrays and, per triangle, 9 indices (v n t per corner) + the material id, in file order. For the parser fuzz test.
// counts[5] = floats in vertices / normals / texcoords, triangles, materials. Returns 0 when LoadObj fails.
int qaref_tinyobj_load(const char *obj_path, const char *mtl_dir, int64_t *counts, float *vertices, float *normals, float *texcoords,
                       int32_t *tri10, int64_t cap_floats, int64_t cap_tris)
{
  tinyobj::attrib_t attrib;
  std::vector<tinyobj::shape_t> shapes;
  std::vector<tinyobj::material_t> materials;
  std::string err;
  if (!tinyobj::LoadObj(&attrib, &shapes, &materials, &err, obj_path, mtl_dir, true)) return 0;
  counts[0] = (int64_t) attrib.vertices.size(); counts[1] = (int64_t) attrib.normals.size(); counts[2] = (int64_t) attrib.texcoords.size();
  for (size_t i = 0; i < attrib.vertices.size() && (int64_t) i < cap_floats; ++i) vertices[i] = attrib.vertices[i];
  for (size_t i = 0; i < attrib.normals.size() && (int64_t) i < cap_floats; ++i) normals[i] = attrib.normals[i];
  for (size_t i = 0; i < attrib.texcoords.size() && (int64_t) i < cap_floats; ++i) texcoords[i] = attrib.texcoords[i];
  int64_t nt = 0;
  for (auto &sh : shapes) {
    size_t off = 0;
    for (size_t f = 0; f < sh.mesh.num_face_vertices.size(); ++f) {
      const int fv = sh.mesh.num_face_vertices[f];
      if (fv == 3 && nt < cap_tris) {
        for (int k = 0; k < 3; ++k) {
          const tinyobj::index_t &ix = sh.mesh.indices[off + k];
          tri10[10 * nt + 3 * k + 0] = ix.vertex_index; tri10[10 * nt + 3 * k + 1] = ix.normal_index; tri10[10 * nt + 3 * k + 2] = ix.texcoord_index;
        }
        tri10[10 * nt + 9] = sh.mesh.material_ids[f];
      }
      if (fv == 3) ++nt;
      off += fv;
    }
  }
  counts[3] = nt;
  counts[4] = (int64_t) materials.size();
  return 1;
}

// Serial fp32 render in the reference's own pixel order (tiles of 4x4 in k order, pixels in idx order,
// renderer.cpp:387-403) with the reference's running mean (scene.cpp:113-121); fixed spp.
// rgb: W*H*3 linear means. Used to pin the restatement bit-for-bit (same seed, same draw order).
void qaref_render_adaptive_serial(int sppMin, int sppMax, float *rgb, float *depth, int32_t *count);
void qaref_render_linear_serial(int spp, float *rgb, float *depth) { qaref_render_adaptive_serial(spp, spp, rgb, depth, nullptr); }

// PixelRender's loop (renderer.cpp:304-346) with the reference's SuperSamplerHalton, serial, linear output:
// mean colour, depth of sample 0 and the number of samples each pixel took (sppMin < sppMax = the adaptive stop).
void qaref_render_adaptive_serial(int sppMin, int sppMax, float *rgb, float *depth, int32_t *count)
{
  const int W = (int) g_renderer->W(), H = (int) g_renderer->H();
  const int ts = 4;
  const int tdx = (W + ts - 1) / ts, tdy = (H + ts - 1) / ts;
  for (int k = 0; k < tdx * tdy; ++k) {
    const int tx = k % tdx, ty = k / tdx;
    const int i0 = std::min(W, tx * ts), j0 = std::min(H, ty * ts);
    const int i1 = std::min(W, (tx + 1) * ts), j1 = std::min(H, (ty + 1) * ts);
    const int np = (i1 - i0) * (j1 - j0);
    for (int idx = 0; idx < np; ++idx) {
      const int j = j0 + idx / (i1 - i0);
      const int i = i0 + idx % (i1 - i0);
      SuperSamplerHalton sampler(Color3f(0.005f, 0.001f, 0.005f), sppMin, sppMax);
      float d = 0.f;
      while (sampler.Loop()) {
        float c[3], z;
        qaref_sample(i, j, sampler.GetSampleID(), c, &z);
        if (sampler.GetSampleID() == 0) d = z;
        sampler.Accumulate(Color3f(c[0], c[1], c[2]));
        sampler.Increment();
      }
      const Color3f &m = sampler.GetColor();
      rgb[3 * (j * W + i) + 0] = m.r;
      rgb[3 * (j * W + i) + 1] = m.g;
      rgb[3 * (j * W + i) + 2] = m.b;
      if (depth) depth[j * W + i] = d;
      if (count) count[j * W + i] = sampler.GetSampleID();
    }
  }
}

// The same loop over the pixel rows [j0, j1) only (fixed spp): lets tools/make_image_golden.py spread a 1024-spp frame
// over several processes (the reference's scene and RNG are process-global). rgb / depth are full-frame buffers.
void qaref_render_linear_rows(int spp, int j0, int j1, float *rgb, float *depth)
{
  const int W = (int) g_renderer->W(), H = (int) g_renderer->H();
  for (int j = std::max(0, j0); j < std::min(H, j1); ++j)
    for (int i = 0; i < W; ++i) {
      SuperSamplerHalton sampler(Color3f(0.005f, 0.001f, 0.005f), spp, spp);
      float d = 0.f;
      while (sampler.Loop()) {
        float c[3], z;
        qaref_sample(i, j, sampler.GetSampleID(), c, &z);
        if (sampler.GetSampleID() == 0) d = z;
        sampler.Accumulate(Color3f(c[0], c[1], c[2]));
        sampler.Increment();
      }
      const Color3f &m = sampler.GetColor();
      rgb[3 * (j * W + i) + 0] = m.r;
      rgb[3 * (j * W + i) + 1] = m.g;
      rgb[3 * (j * W + i) + 2] = m.b;
      if (depth) depth[j * W + i] = d;
    }
}

// The reference's real render (ThreadRender with its own scheduler): fills rgb8 / depth / sampleCount.
// Returns elapsed seconds of ThreadRender.
double qaref_render(uint8_t *rgb8, float *depth, uint8_t *sampleCount)
{
  auto t0 = std::chrono::steady_clock::now();
  g_renderer->Render();
  auto t1 = std::chrono::steady_clock::now();
  const size_t n = g_renderer->W() * g_renderer->H();
  if (rgb8) memcpy(rgb8, renderImage.GetPixels(), n * 3);
  if (depth) memcpy(depth, renderImage.GetZBuffer(), n * sizeof(float));
  if (sampleCount) memcpy(sampleCount, renderImage.GetSampleCount(), n);
  return std::chrono::duration<double>(t1 - t0).count();
}

void qaref_counters(uint64_t *closest, uint64_t *shadow, int reset)
{
#ifdef QAREF_COUNTERS
  *closest = qaref_count_closest;
  *shadow = qaref_count_shadow;
  if (reset) { qaref_count_closest = 0; qaref_count_shadow = 0; }
#else
  *closest = 0; *shadow = 0; (void) reset;
#endif
}

int qaref_threads(void) { return (int) tasking::get_num_of_threads(); }

}  // extern "C"

#ifdef QAREF_MAIN
// qaref_perf <scene.xml> <asset_dir> <spp> <bounce> [threads] [repeat] [rgb8_out_prefix]
// Times the reference's own ThreadRender (its TBB/OpenMP tile loop) and prints one JSON line.
int main(int argc, char **argv)
{
  if (argc < 5) {
    fprintf(stderr, "usage: %s scene.xml asset_dir spp bounce [threads] [repeat] [rgb8_out_prefix]\n", argv[0]);
    return 2;
  }
  int spp = atoi(argv[3]), bounce = atoi(argv[4]);
  int threads = argc > 5 ? atoi(argv[5]) : 0;
  int repeat = argc > 6 ? atoi(argv[6]) : 1;
  qaref_set_params(spp, spp, bounce, 1, threads);
  if (!qaref_load(argv[1], argv[2], 1)) { fprintf(stderr, "load failed\n"); return 1; }
  const size_t W = g_renderer->W(), H = g_renderer->H();
  // rgb8_out_prefix: the colour plane PixelRender wrote (renderer.cpp:348-361) of repeat r goes to <prefix><r>.rgb8
  // (raw W*H*3 bytes), so bench.py can state the image RMSE of the GPU frame against the reference's frame
  const char *dump = argc > 7 ? argv[7] : nullptr;
  std::vector<uint8_t> img(dump ? W * H * 3 : 0);
  for (int r = 0; r < repeat; ++r) {
    double dt = qaref_render(dump ? img.data() : nullptr, nullptr, nullptr);
    if (dump) {
      const std::string path = std::string(dump) + std::to_string(r) + ".rgb8";
      if (FILE *fp = fopen(path.c_str(), "wb")) { fwrite(img.data(), 1, img.size(), fp); fclose(fp); }
    }
    uint64_t c = 0, s = 0;
    qaref_counters(&c, &s, 1);
    printf("{\"scene\": \"%s\", \"width\": %zu, \"height\": %zu, \"spp\": %d, \"bounce\": %d, "
           "\"threads\": %d, \"seconds\": %.6f, \"samples\": %zu, \"closest_rays\": %llu, \"shadow_rays\": %llu}\n",
           argv[1], W, H, spp, bounce, qaref_threads(), dt, W * H * (size_t) spp,
           (unsigned long long) c, (unsigned long long) s);
    fflush(stdout);
  }
  return 0;
}
#endif
