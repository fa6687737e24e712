// api.cu — the C ABI of include/qaray_b200.h: scene upload (host tables -> HBM), the wavefront render driver and
// the ray-buffer entry points. Host code here is compiled with -ffp-contract=off (triangle normals and the camera
// frame must round as the reference's glm expressions do). There is no CPU fallback: without a CUDA device every
// compute entry point returns QG_ENODEVICE.
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <string>
#include <cstdlib>
#include <utility>
#include <vector>

#include "../../include/qaray_b200.h"
#include "png.h"
#include "scene_host.h"
#include "wavefront.h"

using namespace qb;

namespace {

thread_local std::string g_err;
int g_device = -1;  // the device qg_init chose: where scenes are created from then on (each scene remembers its own)

int fail(int code, const std::string &msg)
{
  g_err = msg;
  return code;
}
#define CU(call)                                                                                         \
  do {                                                                                                   \
    cudaError_t e_ = (call);                                                                             \
    if (e_ != cudaSuccess) return fail(QG_ECUDA, std::string(#call) + ": " + cudaGetErrorString(e_));    \
  } while (0)

int ensure_device()
{
  if (g_device >= 0) {
    cudaSetDevice(g_device);
    return QG_OK;
  }
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) return fail(QG_ENODEVICE, "no CUDA device: qaray_b200 has no CPU path");
  g_device = 0;
  cudaSetDevice(0);
  return QG_OK;
}

template <class T>
struct DevBuf {
  T *p = nullptr;
  size_t n = 0;
  ~DevBuf() { release(); }
  void release()
  {
    if (p) cudaFree(p);
    p = nullptr;
    n = 0;
  }
  cudaError_t alloc(size_t count)
  {
    release();
    n = count;
    return cudaMalloc((void **) &p, std::max<size_t>(count, 1) * sizeof(T));
  }
  cudaError_t ensure(size_t count)  // keep the allocation when it is already large enough (no cudaMalloc / cudaFree on the hot path)
  {
    if (p && n >= count) return cudaSuccess;
    return alloc(count);
  }
  cudaError_t upload(const std::vector<T> &v)
  {
    cudaError_t e = alloc(v.size());
    if (e != cudaSuccess || v.empty()) return e;
    return cudaMemcpy(p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice);
  }
  size_t bytes() const { return n * sizeof(T); }
};

}  // namespace

// one photon map of a scene: the photons in emission order as cyPhotonMap would return them (GetPower / GetDirection), and
// their grid-sorted copy on the device (api_photon.inc)
struct PhotonMapStore {
  std::vector<float> pos, dir, power;  // 3, 3, 1 per photon; power = Photon::power (the largest channel)
  std::vector<uint32_t> color;         // Color24: r | g << 8 | b << 16
  float radius = 0.f;
  DevBuf<float4> pa, pb;
  DevBuf<uint32_t> cells;
  DPhotonMap dev{};
};

struct qgScene {
  int device = 0;  // every entry point that takes the scene makes this the calling thread's current device
  FlatScene flat;
  DScene dev;
  DevBuf<DNode> nodes;
  DevBuf<int32_t> obj_node, chain_off, chain, multi;
  DevBuf<DMesh> meshes;
  DevBuf<DBvhNode> bvh;
  DevBuf<DTri> tris;
  DevBuf<uint32_t> tri_face;
  DevBuf<DFace> faces;
  DevBuf<DMaterial> materials;
  DevBuf<DLight> lights;
  DevBuf<DTexture> textures;
  DevBuf<uchar4> texels;
  int64_t num_faces = 0, num_bvh = 0, device_bytes = 0;
  int num_obj = 0, max_bvh_depth = 0;
  // wavefront buffers (allocated on first render, sized by capacity)
  WfBuffers wf;
  std::vector<void *> wf_allocs;
  uint32_t wf_capacity = 0;
  // frame-sized device planes of qg_render / qg_resolve, kept between calls
  DevBuf<float> fr_acc, fr_dep, rs_mean, rs_dep;
  DevBuf<uint8_t> rs_rgb8, rs_sc, rs_mask;
  PhotonMapStore pm[2];  // [0] Scene::photonmap, [1] Scene::causticsmap (src/scene/scene.h:65-66)
  bool pm_built = false;
  bool want_gather = false;  // the next ensure_wf also allocates the gather-request queue
  int dev_num_area = 0;  // PointLight / SpotLight with size > 0.01 (lights.cpp:52,87)
  DevBuf<unsigned long long> stats;
  cudaStream_t stream = nullptr;
  std::vector<cudaEvent_t> ev_pool;  // timestamp events of KernelClock
  cudaStream_t shadow_stream = nullptr;  // shadow(d) runs here, beside trace(d + 1) (run_wave)
  cudaEvent_t ev_shade = nullptr, ev_shadow = nullptr;
  ~qgScene()
  {
    for (cudaEvent_t e : ev_pool) cudaEventDestroy(e);
    if (ev_shade) cudaEventDestroy(ev_shade);
    if (ev_shadow) cudaEventDestroy(ev_shadow);
    if (shadow_stream) cudaStreamDestroy(shadow_stream);
    for (void *p : wf_allocs) cudaFree(p);
    if (stream) cudaStreamDestroy(stream);
  }
};

namespace {

int use_scene(const qgScene *sc)
{
  if (cudaSetDevice(sc->device) != cudaSuccess) return fail(QG_ENODEVICE, "no CUDA device: qaray_b200 has no CPU path");
  return QG_OK;
}

void put_texcolor_dev(const qgTexturedColor &q, DTexColor *o)
{
  memset(o, 0, sizeof *o);
  for (int k = 0; k < 3; ++k) { o->color[k] = q.color[k]; o->pos[k] = q.xf.pos[k]; }
  o->texture = q.texture;
  for (int k = 0; k < 9; ++k) o->itm[k] = q.xf.itm[k];
  bool id = q.xf.pos[0] == 0.f && q.xf.pos[1] == 0.f && q.xf.pos[2] == 0.f;
  for (int c = 0; c < 3; ++c)
    for (int r = 0; r < 3; ++r) id = id && q.xf.itm[3 * c + r] == (c == r ? 1.f : 0.f);
  o->identity = id ? 1 : 0;
}

// what the BVH build reads of a qgMesh: vertex positions and the faces' vertex indices
bool mesh_positions(const qgMesh &m, ObjMesh *om)
{
  om->vertices.assign(m.vertices, m.vertices + 3 * (size_t) m.num_vertices);
  om->faces.resize(m.num_faces);
  for (int f = 0; f < m.num_faces; ++f)
    for (int k = 0; k < 3; ++k) {
      const int vi = m.face_indices[9 * f + 3 * k];
      if (vi < 0 || vi >= m.num_vertices) return false;
      om->faces[f].idx[k].v = vi;
    }
  return true;
}

int upload_scene(qgScene *sc)
{
  const FlatScene &F = sc->flat;
  const int nn = (int) F.nodes.size();
  if (nn < 1 || F.nodes[0].parent != -1) return fail(QG_EINVAL, "node 0 must be the root (parent -1)");
  // nodes
  std::vector<DNode> nodes(nn);
  for (int i = 0; i < nn; ++i) {
    const qgNode &q = F.nodes[i];
    if (i > 0 && (q.parent < 0 || q.parent >= i)) return fail(QG_EINVAL, "nodes must be listed in pre-order (parent before child)");
    DNode &d = nodes[i];
    d.i0 = make_float4(q.xf.itm[0], q.xf.itm[1], q.xf.itm[2], q.xf.pos[0]);
    d.i1 = make_float4(q.xf.itm[3], q.xf.itm[4], q.xf.itm[5], q.xf.pos[1]);
    d.i2 = make_float4(q.xf.itm[6], q.xf.itm[7], q.xf.itm[8], q.xf.pos[2]);
    d.t0 = make_float4(q.xf.tm[0], q.xf.tm[1], q.xf.tm[2], 0.f);
    d.t1 = make_float4(q.xf.tm[3], q.xf.tm[4], q.xf.tm[5], 0.f);
    d.t2 = make_float4(q.xf.tm[6], q.xf.tm[7], q.xf.tm[8], 0.f);
    d.type = q.object_type;
    d.mesh = q.mesh;
    d.material = q.material;
    d.parent = q.parent;
    if (q.object_type == QG_OBJ_MESH && (q.mesh < 0 || q.mesh >= (int) F.meshes.size())) return fail(QG_EINVAL, "node references a missing mesh");
    if (q.material >= (int) F.materials.size()) return fail(QG_EINVAL, "node references a missing material");
  }
  // object list + ancestor chains (root excluded, object node last)
  std::vector<int32_t> obj_node, chain_off{0}, chain;
  for (int i = 0; i < nn; ++i) {
    if (F.nodes[i].object_type == QG_OBJ_NONE) continue;
    std::vector<int> up;
    for (int c = i; c > 0; c = F.nodes[c].parent) up.push_back(c);
    if ((int) up.size() > QB_MAX_DEPTH) return fail(QG_EINVAL, "scene tree deeper than supported");
    obj_node.push_back(i);
    for (auto it = up.rbegin(); it != up.rend(); ++it) chain.push_back(*it);
    chain_off.push_back((int32_t) chain.size());
  }
  sc->num_obj = (int) obj_node.size();
  // meshes: BVH in the reference's topology, triangles in leaf order
  std::vector<DMesh> meshes(F.meshes.size());
  std::vector<DBvhNode> bvh;
  std::vector<DTri> tris;
  std::vector<uint32_t> tri_face;
  std::vector<DFace> faces;
  for (size_t mi = 0; mi < F.meshes.size(); ++mi) {
    const qgMesh &m = F.meshes[mi];
    ObjMesh om;
    if (!mesh_positions(m, &om)) return fail(QG_EINVAL, "mesh face references a missing vertex");
    CyBvh cy;
    build_cy_bvh(om, 4, &cy);  // TriObj::Load: bvh.SetMesh(this, 4) (objects.h:71)
    DMesh &dm = meshes[mi];
    for (int k = 0; k < 3; ++k) { dm.bmin[k] = m.bound_min[k]; dm.bmax[k] = m.bound_max[k]; }
    dm.face_base = (uint32_t) faces.size();
    dm.num_faces = (uint32_t) m.num_faces;
    dm.pad = 0;
    dm.root_ref = 0;
    const uint32_t tri_base = (uint32_t) tris.size();
    const uint32_t bvh_base = (uint32_t) bvh.size();
    if ((size_t) tri_base + m.num_faces >= (1u << 28)) return fail(QG_EINVAL, "more than 2^28 triangles");
    // triangles in leaf (elements[]) order; N = normalize(cross(B-A, C-A)) exactly as IntersectTriangle computes it
    for (uint32_t s = 0; s < (uint32_t) m.num_faces; ++s) {
      const uint32_t f = cy.elements[s];
      const float *a = &m.vertices[3 * m.face_indices[9 * f + 0]], *b = &m.vertices[3 * m.face_indices[9 * f + 3]],
                  *c = &m.vertices[3 * m.face_indices[9 * f + 6]];
      const V3 A = v3(a[0], a[1], a[2]), B = v3(b[0], b[1], b[2]), C = v3(c[0], c[1], c[2]);
      const V3 N = normalize(cross(B - A, C - A));
      DTri t;
      t.a = make_float4(A.x, A.y, A.z, N.x);
      t.b = make_float4(B.x, B.y, B.z, N.y);
      t.c = make_float4(C.x, C.y, C.z, N.z);
      const bool has_uv = m.face_indices[9 * f + 2] >= 0 && m.face_indices[9 * f + 5] >= 0 && m.face_indices[9 * f + 8] >= 0;
      const uint32_t face_word = (dm.face_base + f) | (has_uv ? 0x80000000u : 0u);
      // per-triangle constants of IntersectTriangle (objects.cpp:235-247), same expressions, no contraction
      const int axis = tri_axis(N);
      const float rcp_area = 1.f / tri_area(axis, A, B, C);
      uint32_t axis_bits = (uint32_t) axis;
      float axis_f, face_f;
      memcpy(&axis_f, &axis_bits, 4);
      memcpy(&face_f, &face_word, 4);
      t.d = make_float4(rcp_area, axis_f, face_f, 0.f);
      tris.push_back(t);
      tri_face.push_back(face_word);
    }
    for (int f = 0; f < m.num_faces; ++f) {
      DFace df;
      memset(&df, 0, sizeof df);
      bool has_uv = true;
      for (int k = 0; k < 3; ++k) {
        const int ni = m.face_indices[9 * f + 3 * k + 1], ti = m.face_indices[9 * f + 3 * k + 2];
        if (ni >= 0 && ni < m.num_normals)
          for (int c = 0; c < 3; ++c) df.n[3 * k + c] = m.normals[3 * ni + c];
        if (ti >= 0 && ti < m.num_texcoords) { df.uv[2 * k] = m.texcoords[2 * ti]; df.uv[2 * k + 1] = m.texcoords[2 * ti + 1]; }
        if (ti < 0) has_uv = false;
      }
      df.mtl_uv = ((m.face_material[f] + 1) << 1) | (has_uv ? 1 : 0);
      faces.push_back(df);
    }
    if (m.num_faces > 0) {
      // cy node id -> child ref
      std::vector<uint32_t> internal_index(cy.nodes.size(), 0);
      uint32_t ni = 0;
      for (size_t id = 1; id < cy.nodes.size(); ++id)
        if (!cy.nodes[id].leaf()) internal_index[id] = bvh_base + ni++;
      auto ref_of = [&](uint32_t id) -> uint32_t {
        const CyBvh::Node &n = cy.nodes[id];
        if (n.leaf()) return QB_LEAF_BIT | ((n.count() - 1) << 28) | (tri_base + n.offset());
        return internal_index[id];
      };
      // depth of the tree = most far children a ray can have deferred at once (the shared-memory stack of traverse.cuh)
      {
        std::vector<std::pair<uint32_t, int>> todo{{1u, 0}};
        int max_depth = 0;
        while (!todo.empty()) {
          const auto [id, dep] = todo.back();
          todo.pop_back();
          max_depth = std::max(max_depth, dep);
          if (!cy.nodes[id].leaf()) { todo.push_back({cy.nodes[id].child(), dep + 1}); todo.push_back({cy.nodes[id].child() + 1, dep + 1}); }
        }
        if (max_depth >= QB_STACK) return fail(QG_EINVAL, "mesh BVH deeper than the traversal stack (QB_STACK)");
        sc->max_bvh_depth = std::max(sc->max_bvh_depth, max_depth);
      }
      bvh.resize(bvh_base + ni);
      for (size_t id = 1; id < cy.nodes.size(); ++id) {
        const CyBvh::Node &n = cy.nodes[id];
        if (n.leaf()) continue;
        const uint32_t c0 = n.child(), c1 = c0 + 1;
        const float *b0 = cy.nodes[c0].box, *b1 = cy.nodes[c1].box;
        DBvhNode &o = bvh[internal_index[id]];
        o.a = make_float4(b0[0], b0[1], b0[2], b0[3]);
        o.b = make_float4(b0[4], b0[5], b1[0], b1[1]);
        o.c = make_float4(b1[2], b1[3], b1[4], b1[5]);
        o.d = make_uint4(ref_of(c0), ref_of(c1), 0u, 0u);
      }
      dm.root_ref = ref_of(1);
    }
  }
  sc->num_faces = (int64_t) faces.size();
  sc->num_bvh = (int64_t) bvh.size();
  // materials, lights, textures
  std::vector<DMaterial> mats(F.materials.size());
  for (size_t i = 0; i < mats.size(); ++i) {
    const qgMaterial &q = F.materials[i];
    DMaterial &d = mats[i];
    memset(&d, 0, sizeof d);
    d.kind = q.kind;
    d.sub_first = q.sub_first;
    d.sub_count = q.sub_count;
    put_texcolor_dev(q.diffuse, &d.diffuse);
    put_texcolor_dev(q.specular, &d.specular);
    put_texcolor_dev(q.emission, &d.emission);
    put_texcolor_dev(q.reflection, &d.reflection);
    put_texcolor_dev(q.refraction, &d.refraction);
    d.has_maps = (q.diffuse.texture >= 0 || q.specular.texture >= 0 || q.emission.texture >= 0 || q.reflection.texture >= 0 ||
                  q.refraction.texture >= 0);
    for (int k = 0; k < 3; ++k) d.absorption[k] = q.absorption[k];
    d.ior = q.ior;
    d.gloss = q.glossiness;
    d.refl_gloss = q.reflection_glossiness;
    d.refr_gloss = q.refraction_glossiness;
    d.kill = q.kill;
    if (q.kind == QG_MTL_MULTI && (q.sub_first < 0 || q.sub_first + q.sub_count > (int) F.multi.size()))
      return fail(QG_EINVAL, "multi-material list out of range");
  }
  for (int32_t s : F.multi)
    if (s < 0 || s >= (int) mats.size()) return fail(QG_EINVAL, "multi-material entry out of range");
  std::vector<DLight> lights(F.lights.size());
  int nonambient = 0;
  for (size_t i = 0; i < lights.size(); ++i) {
    const qgLight &q = F.lights[i];
    DLight &d = lights[i];
    d.type = q.type;
    for (int k = 0; k < 3; ++k) { d.intensity[k] = q.intensity[k]; d.position[k] = q.position[k]; d.direction[k] = q.direction[k]; }
    d.size = q.size;
    d.inner = q.inner;
    d.outer = q.outer;
    if (q.type != QG_LIGHT_AMBIENT) ++nonambient;
  }
  std::vector<DTexture> texs(F.textures.size());
  std::vector<uchar4> texels;
  for (size_t i = 0; i < texs.size(); ++i) {
    const qgTexture &q = F.textures[i];
    DTexture &d = texs[i];
    memset(&d, 0, sizeof d);
    d.type = q.type;
    d.width = q.width;
    d.height = q.height;
    for (int k = 0; k < 3; ++k) { d.c1[k] = q.color1[k]; d.c2[k] = q.color2[k]; }
    d.texel_base = (uint32_t) texels.size();
    if (q.type == QG_TEX_IMAGE) {
      if (!q.rgb8 || q.width <= 0 || q.height <= 0) return fail(QG_EINVAL, "image texture without pixels");
      const size_t np = (size_t) q.width * q.height;
      for (size_t p = 0; p < np; ++p) texels.push_back(make_uchar4(q.rgb8[3 * p], q.rgb8[3 * p + 1], q.rgb8[3 * p + 2], 255));
    }
  }
  auto tex_ok = [&](const qgTexturedColor &t) { return t.texture < (int) texs.size(); };
  for (const qgMaterial &q : F.materials)
    if (!tex_ok(q.diffuse) || !tex_ok(q.specular) || !tex_ok(q.emission) || !tex_ok(q.reflection) || !tex_ok(q.refraction))
      return fail(QG_EINVAL, "material references a missing texture");
  if (!tex_ok(F.background) || !tex_ok(F.environment)) return fail(QG_EINVAL, "background/environment references a missing texture");

  CU(sc->nodes.upload(nodes));
  CU(sc->obj_node.upload(obj_node));
  CU(sc->chain_off.upload(chain_off));
  CU(sc->chain.upload(chain));
  CU(sc->meshes.upload(meshes));
  CU(sc->bvh.upload(bvh));
  CU(sc->tris.upload(tris));
  CU(sc->tri_face.upload(tri_face));
  CU(sc->faces.upload(faces));
  CU(sc->materials.upload(mats));
  CU(sc->multi.upload(F.multi));
  CU(sc->lights.upload(lights));
  CU(sc->textures.upload(texs));
  CU(sc->texels.upload(texels));
  CU(sc->stats.alloc(8));
  CU(cudaMemset(sc->stats.p, 0, sc->stats.bytes()));
  sc->device_bytes = (int64_t) (sc->nodes.bytes() + sc->obj_node.bytes() + sc->chain_off.bytes() + sc->chain.bytes() + sc->meshes.bytes() +
                                sc->bvh.bytes() + sc->tris.bytes() + sc->tri_face.bytes() + sc->faces.bytes() + sc->materials.bytes() +
                                sc->multi.bytes() + sc->lights.bytes() + sc->textures.bytes() + sc->texels.bytes());
  DScene &D = sc->dev;
  memset(&D, 0, sizeof D);
  D.nodes = sc->nodes.p; D.obj_node = sc->obj_node.p; D.chain_off = sc->chain_off.p; D.chain = sc->chain.p;
  D.meshes = sc->meshes.p; D.bvh = sc->bvh.p; D.tris = sc->tris.p; D.tri_face = sc->tri_face.p; D.faces = sc->faces.p;
  D.materials = sc->materials.p; D.multi = sc->multi.p; D.lights = sc->lights.p; D.textures = sc->textures.p; D.texels = sc->texels.p;
  D.num_nodes = nn; D.num_obj = sc->num_obj; D.num_meshes = (int) meshes.size(); D.num_materials = (int) mats.size();
  D.num_lights = (int) lights.size(); D.num_textures = (int) texs.size(); D.num_nonambient = nonambient;
  D.bvh_stack = std::max(1, sc->max_bvh_depth + 1);
  // ancestor-ray cache of the traversal (traverse.cuh, QB_ANC_CACHE): worth its 7 stack entries per lane when consecutive
  // objects share their first-level group node and the larger stacks keep the same number of blocks resident
  {
    int shared = 0;
    for (int k = 1; k < sc->num_obj; ++k) {
      const int a0 = chain_off[k - 1], a1 = chain_off[k], b1 = chain_off[k + 1];
      if (a1 - a0 >= 2 && b1 - a1 >= 2 && chain[a0] == chain[a1]) ++shared;
    }
    if (const char *e = getenv("QB_ANC_CACHE_OFF")) shared = atoi(e) ? 0 : shared;  // A/B
    if (shared > 0 && trace_stack_fits(D.bvh_stack, QB_ANC_WORDS)) {
      D.anc_cache = 1;
      D.bvh_stack += QB_ANC_WORDS;
    }
  }
  sc->dev_num_area = 0;
  for (const qgLight &l : F.lights)
    if ((l.type == QG_LIGHT_POINT || l.type == QG_LIGHT_SPOT) && l.size > 0.01f) sc->dev_num_area++;
  put_texcolor_dev(F.background, &D.background);
  put_texcolor_dev(F.environment, &D.environment);
  // Texture::Sample's tap table (core/texture.cpp:38-49): Halton(2,3) points mapped onto the unit-diameter disc
  float taps[64];
  taps[0] = taps[1] = 0.f;
  for (int i = 1; i < 32; ++i) {
    float x = halton(i, 2), y = halton(i, 3);
    const float r = sqrtf(x) * 0.5f;
    x = r * sinf(y * (float) M_PI * 2);
    y = r * cosf(y * (float) M_PI * 2);
    taps[2 * i] = x;
    taps[2 * i + 1] = y;
  }
  upload_texture_taps(taps);
  CU(cudaStreamCreateWithFlags(&sc->stream, cudaStreamNonBlocking));
  CU(cudaStreamCreateWithFlags(&sc->shadow_stream, cudaStreamNonBlocking));
  CU(cudaEventCreateWithFlags(&sc->ev_shade, cudaEventDisableTiming));
  CU(cudaEventCreateWithFlags(&sc->ev_shadow, cudaEventDisableTiming));
  CU(cudaDeviceSynchronize());
  return QG_OK;
}

}  // namespace

// ---------------------------------------------------------------------------------------------- render driver
namespace {

int ensure_wf(qgScene *sc, uint32_t capacity, bool carry_diff = false)
{
  if (sc->wf_capacity == capacity && (!carry_diff || sc->wf.h4 != nullptr) && (!sc->want_gather || sc->wf.g0 != nullptr)) return QG_OK;
  carry_diff = carry_diff || sc->wf.h4 != nullptr;  // keep what an earlier render asked for
  for (void *p : sc->wf_allocs) cudaFree(p);
  sc->wf_allocs.clear();
  sc->wf_capacity = 0;
  // room for the dead entries of the shade kernels' output reservations: at most one QB_OUT_CHUNK tail per shade warp and
  // queue, for the largest grid any material class's shade kernel may be launched with
  const uint32_t requested = capacity;
  capacity += (uint32_t) wf_shade_grid_max() * (QB_BLOCK / 32) * QB_OUT_CHUNK;
  WfBuffers &B = sc->wf;
  memset(&B, 0, sizeof B);
  auto alloc4 = [&](float4 **p, size_t n) -> cudaError_t {
    cudaError_t e = cudaMalloc((void **) p, std::max<size_t>(n, 1) * sizeof(float4));
    if (e == cudaSuccess) sc->wf_allocs.push_back(*p);
    return e;
  };
  const size_t shadow_cap = (size_t) capacity * (size_t) std::max(1, sc->dev.num_nonambient);
  for (int k = 0; k < 2; ++k) {
    CU(alloc4(&B.ps0[k], capacity));
    CU(alloc4(&B.ps1[k], capacity));
    CU(alloc4(&B.ps2[k], capacity));
    CU(alloc4(&B.ps3[k], capacity));
  }
  CU(alloc4(&B.h0, capacity));
  CU(alloc4(&B.h1, capacity));
  CU(alloc4(&B.h2, capacity));
  CU(alloc4(&B.h3, capacity));
  if (carry_diff) {  // differential rays of secondary paths (Whitted-style material classes)
    for (int k = 0; k < 2; ++k) {
      CU(alloc4(&B.ps4[k], capacity));
      CU(alloc4(&B.ps5[k], capacity));
      CU(alloc4(&B.ps6[k], capacity));
    }
    CU(alloc4(&B.h4, capacity));
  }
  if (sc->want_gather) {  // photon-map gather requests: at most one per shaded hit
    CU(alloc4(&B.g0, capacity));
    CU(alloc4(&B.g1, capacity));
    CU(alloc4(&B.g2, capacity));
    CU(alloc4(&B.g3, capacity));
    CU(alloc4(&B.g4, capacity));
  }
  CU(alloc4(&B.s0, shadow_cap));
  CU(alloc4(&B.s1, shadow_cap));
  CU(alloc4(&B.s2, shadow_cap));
  if (sc->dev_num_area > 0) {  // area-light groups: one per (path, area light)
    const size_t area_cap = std::min<size_t>((size_t) capacity * (size_t) sc->dev_num_area, 0xFFFFFFFFull / QB_AREA_SPP_MAX);
    CU(alloc4(&B.a0, area_cap));
    CU(alloc4(&B.a1, area_cap));
    CU(alloc4(&B.a2, area_cap));
    CU(cudaMalloc((void **) &B.amask, area_cap * 2 * sizeof(uint32_t)));
    sc->wf_allocs.push_back(B.amask);
    CU(cudaMemset(B.amask, 0, area_cap * 2 * sizeof(uint32_t)));
    B.area_capacity = (uint32_t) area_cap;
  }
  B.anyq_capacity = 1u << 22;  // beyond this the lanes repeat the mesh themselves (exact, slower)
  CU(cudaMalloc((void **) &B.anyq, (size_t) B.anyq_capacity * sizeof(uint32_t)));
  sc->wf_allocs.push_back(B.anyq);
  B.redo_capacity = 1u << 18;
  CU(cudaMalloc((void **) &B.redo, (size_t) B.redo_capacity * sizeof(uint4)));
  sc->wf_allocs.push_back(B.redo);
  CU(cudaMalloc((void **) &B.perm, std::max<size_t>(capacity, 1) * sizeof(uint32_t)));
  sc->wf_allocs.push_back(B.perm);
  CU(cudaMalloc((void **) &B.mat_bins, 2 * QB_MAT_BINS * sizeof(uint32_t)));
  sc->wf_allocs.push_back(B.mat_bins);
  CU(cudaMemset(B.mat_bins, 0, 2 * QB_MAT_BINS * sizeof(uint32_t)));
  CU(cudaMalloc((void **) &B.counters, QB_CNT_WORDS * sizeof(uint32_t)));
  sc->wf_allocs.push_back(B.counters);
  B.stats = sc->stats.p;
  B.capacity = capacity;
  B.shadow_capacity = (uint32_t) std::min<size_t>(shadow_cap, 0xFFFFFFFFu);
  sc->wf_capacity = requested;
  return QG_OK;
}

// Renderer::ComputeScene's camera frame (renderer.cpp:77-91)
void camera_frame(const qgCamera &cam, int W, int H, V3 *U, V3 *Vv, V3 *A, V3 *Xo = nullptr, V3 *Yo = nullptr)
{
  const float PI = (float) M_PI;
  const V3 pos = v3(cam.pos[0], cam.pos[1], cam.pos[2]), dir = v3(cam.dir[0], cam.dir[1], cam.dir[2]), up = v3(cam.up[0], cam.up[1], cam.up[2]);
  const float focal = cam.focal;
  const float aspect = W / (float) H;
  const float screenH = 2.f * focal * std::tan(cam.fovy * PI / 2.f / 180.f);
  const float screenW = aspect * screenH;
  const V3 X = normalize(cross(dir, up));
  const V3 Y = normalize(cross(X, dir));
  const V3 Z = normalize(-dir);
  *U = X * (screenW / W);
  *Vv = -Y * (screenH / H);
  *A = pos - Z * focal + Y * screenH / 2.f - X * screenW / 2.f;
  if (Xo) *Xo = X;  // screenX / screenY: the lens plane of the depth-of-field jitter (renderer.cpp:88-89,319-322)
  if (Yo) *Yo = Y;
}

bool class_carries_diff(int cls) { return cls == QG_CLASS_MONTECARLOGI || cls == QG_CLASS_BLINN_BASIC || cls == QG_CLASS_PHONG_BASIC; }

struct Shard {
  int W, H, tiles_x, tiles_y, tiles_local, s0, s1;
  unsigned long long total_paths, valid_samples, owned_pixels;  // 64-bit: 4K x 1024 spp is 8.5e9 samples
};

int make_params(const qgScene *sc, const qgRenderParams *rp, WfParams *P, Shard *sh)
{
  if (!rp || rp->spp <= 0) return fail(QG_EINVAL, "spp must be positive");
  if (rp->max_bounce < 0 || rp->max_bounce >= QB_MAX_BOUNCE) return fail(QG_EINVAL, "max_bounce out of range [0, 14]");
  const qgCamera &cam = sc->flat.camera;
  const int W = rp->width > 0 ? rp->width : cam.width, H = rp->height > 0 ? rp->height : cam.height;
  if (W <= 0 || H <= 0) return fail(QG_EINVAL, "image size must be positive");
  const int ranks = rp->tile_ranks > 0 ? rp->tile_ranks : 1;
  if (rp->tile_rank < 0 || rp->tile_rank >= ranks) return fail(QG_EINVAL, "tile_rank out of range");
  int s0 = rp->sample_begin, s1 = rp->sample_end;
  if (s0 == 0 && s1 == 0) s1 = rp->spp;
  if (s0 < 0 || s1 > rp->spp || s0 > s1) return fail(QG_EINVAL, "sample shard out of range");
  memset(P, 0, sizeof *P);
  camera_frame(cam, W, H, &P->scr_u, &P->scr_v, &P->scr_a, &P->scr_x, &P->scr_y);
  P->dof = cam.dof;
  P->cam_pos = v3(cam.pos[0], cam.pos[1], cam.pos[2]);
  P->width = W;
  P->height = H;
  P->mode = 0;
  sh->W = W;
  sh->H = H;
  sh->tiles_x = (W + 3) / 4;  // tileDimX = ceil(W / tileSize), tileSize = 4 (renderer.h:93, renderer.cpp:104-108)
  sh->tiles_y = (H + 3) / 4;
  const int tile_count = sh->tiles_x * sh->tiles_y;
  sh->tiles_local = tile_count > rp->tile_rank ? (tile_count - rp->tile_rank + ranks - 1) / ranks : 0;
  sh->s0 = s0;
  sh->s1 = s1;
  sh->total_paths = (unsigned long long) sh->tiles_local * 16ull * (unsigned long long) (s1 - s0);
  // pixels this shard really owns (edge tiles are partly outside the image)
  unsigned long long px = 0;
  for (int lt = 0; lt < sh->tiles_local; ++lt) {
    const int k = rp->tile_rank + lt * ranks, tx = k % sh->tiles_x, ty = k / sh->tiles_x;
    px += (unsigned long long) (std::min(W, tx * 4 + 4) - tx * 4) * (unsigned long long) (std::min(H, ty * 4 + 4) - ty * 4);
  }
  sh->owned_pixels = px;
  sh->valid_samples = px * (unsigned long long) (s1 - s0);
  P->tiles_x = sh->tiles_x;
  P->tile_rank = rp->tile_rank;
  P->tile_ranks = ranks;
  P->tiles_local = sh->tiles_local;
  P->sample_begin = s0;
  P->seed = rp->seed;
  P->max_bounce = rp->max_bounce;
  P->stale_tex = 1;
  P->sort_mat = rp->sort_by_material ? 1 : 0;
  if (rp->material_class < QG_CLASS_PHOTONMAP || rp->material_class > QG_CLASS_PHONG_BASIC) return fail(QG_EINVAL, "unknown material_class");
  P->mtl_class = rp->material_class;
  P->carry_diff = class_carries_diff(rp->material_class) ? 1 : 0;
  // MtlBlinn_PhotonMap in queue order: the shade kernel computes the hit attributes itself (no k_wf_finish pass, no h0..h3
  // round trip). The material sort needs mtlID before shade and the other classes read h0..h4 in their own kernels: those
  // keep the separate pass. QB_FUSED_FINISH=0 restores it everywhere (A/B).
  P->fused_finish = rp->material_class == QG_CLASS_PHOTONMAP && !rp->sort_by_material ? 1 : 0;
  if (const char *e = getenv("QB_FUSED_FINISH")) P->fused_finish = P->fused_finish && atoi(e) != 0;
  // the constructors of MtlBlinn_Basic / MtlPhong_Basic switch the lights' inverse-square falloff off for the whole
  // process (MtlBlinn_Basic.cpp:18, lights.cpp:18-30); it happens as soon as the scene file declares one material
  if (rp->use_photon_map) {
    if (rp->material_class != QG_CLASS_PHOTONMAP) return fail(QG_EINVAL, "use_photon_map needs material_class 0 (MtlBlinn_PhotonMap)");
    if (!sc->pm_built) return fail(QG_EINVAL, "use_photon_map without maps: call qg_photon_maps_build first (Renderer::ComputeScene)");
    if ((unsigned long long) W * (unsigned long long) H >= (1ull << 30)) return fail(QG_EINVAL, "frame too large for the gather queue's slot field");
    P->use_photon_map = 1;
    P->pm[0] = sc->pm[0].dev;
    P->pm[1] = sc->pm[1].dev;
  }
  P->no_falloff = (rp->material_class == QG_CLASS_BLINN_BASIC || rp->material_class == QG_CLASS_PHONG_BASIC) && !sc->flat.materials.empty();
  return QG_OK;
}

// Worst-case number of paths at any depth per primary path: the Whitted-style classes spawn a refraction and a
// reflection ray per hit, MtlBlinn_MonteCarloGI adds 10 hemisphere samples at depth 0 and 1 below (:176-178). The wave
// size is the buffer capacity divided by this, so no queue can overflow whatever the scene does.
unsigned long long class_growth(int cls, int max_bounce)
{
  unsigned long long g = 1, worst = 1;
  for (int d = 0; d < max_bounce; ++d) {
    if (cls == QG_CLASS_BLINN_BASIC || cls == QG_CLASS_PHONG_BASIC) g *= 2ull;
    else if (cls == QG_CLASS_MONTECARLOGI) g *= (d == 0 ? 12ull : 3ull);
    worst = std::max(worst, g);
  }
  return worst;
}

// Per-kernel-class device time: one timestamp event after every launch (class = 0 trace, 1 shade, 2 shadow,
// 3 other); a launch's duration is the gap to the previous timestamp on the same stream. Events come from a pool
// owned by the scene so a render creates none after its first call.
struct KernelClock {
  qgScene *sc;
  cudaStream_t st;
  bool on;
  size_t used = 0;
  std::vector<int> cls;   // class of the interval that ENDS at this event; -1 = the event only opens an interval
  std::vector<int> prev;  // the event that opens it: the previous one recorded on the same stream
  int last_main = -1, last_side = -1;
  void stamp(int c, cudaStream_t s = nullptr)
  {
    if (!on) return;
    if (used == sc->ev_pool.size()) {
      cudaEvent_t e;
      if (cudaEventCreate(&e) != cudaSuccess) { on = false; return; }
      sc->ev_pool.push_back(e);
    }
    const bool side = s != nullptr && s == sc->shadow_stream;
    cudaEventRecord(sc->ev_pool[used], s ? s : st);
    int &last = side ? last_side : last_main;
    cls.push_back(c);
    prev.push_back(last);
    last = (int) used++;
  }
  // out[0] trace + finish, out[1] shade, out[2] any-hit time NOT already covered by a trace interval (the any-hit kernels
  // run beside trace(d + 1) on a second stream), out[3] other: out[0] + out[2] is the time some traversal kernel was running
  void sums(double out[4]) const
  {
    for (int k = 0; k < 4; ++k) out[k] = 0.0;
    std::vector<std::pair<float, float>> tr, sh;
    for (size_t i = 0; i < used; ++i) {
      float t0 = 0.f, t1 = 0.f;
      if (cls[i] < 0 || prev[i] < 0) continue;
      if (cudaEventElapsedTime(&t0, sc->ev_pool[0], sc->ev_pool[prev[i]]) != cudaSuccess) continue;
      if (cudaEventElapsedTime(&t1, sc->ev_pool[0], sc->ev_pool[i]) != cudaSuccess) continue;
      if (cls[i] == 0) tr.push_back({t0, t1});
      if (cls[i] == 2) sh.push_back({t0, t1});
      else out[cls[i]] += (double) (t1 - t0);
    }
    std::sort(sh.begin(), sh.end());
    float covered_to = -1.f;
    for (auto iv : sh) {  // any-hit intervals are serial on their stream; subtract what the trace intervals cover
      float lo = std::max(iv.first, covered_to), hi = iv.second;
      if (hi <= lo) continue;
      double len = (double) (hi - lo);
      for (auto t : tr) {
        const float a = std::max(lo, t.first), b = std::min(hi, t.second);
        if (b > a) len -= (double) (b - a);
      }
      out[2] += std::max(0.0, len);
      covered_to = hi;
    }
  }
};

// One wave: the depth loop trace -> finish -> shade -> shadow with no host synchronisation. The shadow rays of depth d
// only add into the frame, so they run on a second stream beside trace(d + 1) / finish(d + 1): every traversal launch
// ends in a latency-bound tail (its longest rays), and the other kernel's blocks fill the SMs that tail leaves idle.
// trace(d + 1) is launched first so that it owns the SMs and shadow(d) fills in; shade(d + 1) rewrites the shadow queue
// and therefore waits for shadow(d). (Running the main chain on a higher-priority stream changed nothing: measured.)
#ifndef QB_FUSED_DEPTHS_DEFAULT
#define QB_FUSED_DEPTHS_DEFAULT 0xFFFFFFFEu
#endif
void run_wave(qgScene *sc, WfParams P, cudaStream_t st, uint64_t *launches, KernelClock *clk = nullptr)
{
  const WfBuffers &B = sc->wf;
  static const bool serial = getenv("QB_SERIAL_SHADOW") != nullptr;  // A/B: everything on one stream
  cudaStream_t st2 = serial ? st : sc->shadow_stream;
  cudaMemsetAsync(B.counters, 0, QB_CNT_WORDS * sizeof(uint32_t), st);
  if (clk) clk->stamp(3, st);
  const int gt = wf_trace_grid(), gs = wf_shade_grid(), gsh = wf_shadow_grid();
  const bool any_shadow = sc->dev.num_nonambient > 0 || sc->dev_num_area > 0;
  // which depths run the fused shade kernel (bit d; QB_FUSED_DEPTHS overrides for A/B). Depth 0 keeps the separate pass: its
  // attributes include the two differential rays (three chain transforms, texture footprints) and run better at k_wf_finish's
  // 8 blocks / SM. Measured, ms per step, none / d0 only / d>=1 / all: tower 71.88 / 71.92 / 71.48 / 71.57, xmas 426.4 / 431.0 / 425.8 / 430.3
  static const unsigned fused_depths = getenv("QB_FUSED_DEPTHS") ? (unsigned) strtoul(getenv("QB_FUSED_DEPTHS"), nullptr, 0) : QB_FUSED_DEPTHS_DEFAULT;
  const bool fused_ok = P.fused_finish != 0;
  auto shade = [&](int d) {
    P.depth = d;
    P.fused_finish = fused_ok && ((fused_depths >> d) & 1u) ? 1 : 0;
    if (P.sort_mat) {
      launch_mat_sort(sc->dev, B, P, st);
      *launches += 3;
    }
    launch_wf_shade(sc->dev, B, P, gs, st);
    *launches += 1;
    if (P.use_photon_map) {
      launch_photon_gather(sc->dev, B, P, st);
      *launches += 1;
    }
    if (clk) clk->stamp(1, st);
    if (any_shadow && st2 != st) cudaEventRecord(sc->ev_shade, st);
  };
  auto trace = [&](int d) {
    P.depth = d;
    P.fused_finish = fused_ok && ((fused_depths >> d) & 1u) ? 1 : 0;
    launch_wf_trace(sc->dev, B, P, gt, st);
    if (clk) clk->stamp(0, st);
    *launches += P.fused_finish ? 3 : 4;  // traversal, the two cull-check kernels, hit attributes (unless the shade kernel computes them)
  };
  trace(0);
  shade(0);
  for (int d = 0; d <= P.max_bounce; ++d) {
    if (d < P.max_bounce) trace(d + 1);
    if (any_shadow) {
      P.depth = d;
      if (st2 != st) {
        cudaStreamWaitEvent(st2, sc->ev_shade, 0);
        if (clk) clk->stamp(-1, st2);
      }
      if (sc->dev.num_nonambient > 0) {
        launch_wf_shadow(sc->dev, B, P, gsh, st2);
        if (clk) clk->stamp(2, st2);
        *launches += 2;  // any-hit traversal + the warp-per-ray exact walk of what it handed over
      }
      if (sc->dev_num_area > 0) {
        launch_wf_area(sc->dev, B, P, gsh, st2);
        if (clk) clk->stamp(2, st2);
        *launches += 3;
      }
      if (st2 != st) {
        cudaEventRecord(sc->ev_shadow, st2);
        cudaStreamWaitEvent(st, sc->ev_shadow, 0);  // before shade(d + 1), or the end of the wave
        if (clk) clk->stamp(-1, st);
      }
    }
    if (d < P.max_bounce) shade(d + 1);
  }
  launch_wf_stats(B, P.max_bounce, st);
  if (clk) clk->stamp(3, st);
  *launches += 1;
}

int render_adaptive(qgScene *sc, const qgRenderParams *rp, float *d_accum, float *d_depth, cudaStream_t st, qgRenderStats *stats);

int render_accumulate(qgScene *sc, const qgRenderParams *rp, float *d_accum, float *d_depth, cudaStream_t st, qgRenderStats *stats)
{
  if (rp && rp->spp_min > 0 && rp->spp_min < rp->spp) return render_adaptive(sc, rp, d_accum, d_depth, st, stats);
  WfParams P;
  Shard sh;
  int rc = make_params(sc, rp, &P, &sh);
  if (rc) return rc;
  // Paths per wave. Every launch ends with a latency-bound tail (its longest ray walks its BVH nodes one dependent
  // load after another), so waves are made as large as a fixed slice of HBM allows: 16 GiB of path, hit and
  // shadow-queue state (12 + 3 * lights float4 per path), at most 2^26 paths.
  uint32_t cap = (uint32_t) rp->batch_paths;
  if (rp->batch_paths <= 0) {
    const unsigned long long per_path = 16ull * ((P.carry_diff ? 19ull : 12ull) + 3ull * (unsigned long long) std::max(1, sc->dev.num_nonambient)) +
                                        56ull * (unsigned long long) sc->dev_num_area;
    cap = (uint32_t) std::min<unsigned long long>(1ull << 26, (16ull << 30) / per_path);
  }
  cap = (cap + 31u) & ~31u;
  // paths per wave: the capacity over the class's worst-case growth (1 for the two path-tracing classes)
  const unsigned long long growth = class_growth(P.mtl_class, P.max_bounce);
  uint32_t wave = (uint32_t) std::max<unsigned long long>(32ull, ((unsigned long long) cap / growth) & ~31ull);
  if ((unsigned long long) wave > sh.total_paths) wave = (uint32_t) std::max<unsigned long long>(32ull, (sh.total_paths + 31ull) & ~31ull);
  if ((unsigned long long) wave * growth > (1ull << 31)) return fail(QG_EINVAL, "max_bounce too deep for this material class");
  cap = (uint32_t) ((unsigned long long) wave * growth);
  sc->want_gather = sc->want_gather || P.use_photon_map != 0;
  rc = ensure_wf(sc, cap, P.carry_diff != 0);
  if (rc) return rc;
  P.accum = d_accum;
  P.depth_plane = d_depth;
  cudaEvent_t e0, e1;
  CU(cudaEventCreate(&e0));
  CU(cudaEventCreate(&e1));
  CU(cudaMemsetAsync(sc->stats.p, 0, sc->stats.bytes(), st));
  CU(cudaEventRecord(e0, st));
  uint64_t launches = 0;
  KernelClock clk{sc, st, stats != nullptr};
  for (unsigned long long base = 0; base < sh.total_paths; base += wave) {
    P.wave_base = base;
    P.n_paths = (uint32_t) std::min<unsigned long long>(wave, sh.total_paths - base);
    run_wave(sc, P, st, &launches, &clk);
  }
  CU(cudaEventRecord(e1, st));
  CU(cudaStreamSynchronize(st));
  CU(cudaGetLastError());
  if (stats) {
    unsigned long long hs[8];
    CU(cudaMemcpy(hs, sc->stats.p, sizeof hs, cudaMemcpyDeviceToHost));
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    memset(stats, 0, sizeof *stats);
    stats->samples = sh.valid_samples;
    stats->closest_rays = hs[0] + sh.valid_samples;
    stats->shadow_rays = hs[1];
    stats->nodes_visited = hs[2];
    stats->tris_tested = hs[3];
    stats->objects_tested = hs[4];
    stats->shadow_nodes_visited = hs[5];
    stats->shadow_tris_tested = hs[6];
    stats->shadow_objects_tested = hs[7];
    stats->kernel_launches = launches;
    stats->device_ms = ms;
    double cms[4];
    clk.sums(cms);
    stats->trace_ms = cms[0];
    stats->shade_ms = cms[1];
    stats->shadow_ms = cms[2];
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  return QG_OK;
}

// Renderer::PixelRender with sppMin < sppMax: SuperSamplerHalton's variance-driven stop (src/scene/scene.cpp:92-97).
// Per round: the pixels still running get their next `chunk` samples as mode-1 paths (radiance kept per path), then
// k_adaptive_update replays Accumulate / Loop over them in sample order. One host read of the active count per round.
int render_adaptive(qgScene *sc, const qgRenderParams *rp, float *d_accum, float *d_depth, cudaStream_t st, qgRenderStats *stats)
{
  WfParams P;
  Shard sh;
  qgRenderParams q = *rp;
  if (q.sample_begin != 0 || q.sample_end != 0) return fail(QG_EINVAL, "adaptive sampling does not shard by sample index (its stop depends on all earlier samples)");
  int rc = make_params(sc, &q, &P, &sh);
  if (rc) return rc;
  const int W = sh.W, H = sh.H, npix = W * H, spp_min = rp->spp_min, spp_max = rp->spp;
  const unsigned long long growth = class_growth(P.mtl_class, P.max_bounce);
  // paths per round: bounded by 2^24 (radiance + triples + wave state stay around a GiB)
  const unsigned long long round_cap = std::max<unsigned long long>(1ull << 16, (1ull << 24) / growth);
  DevBuf<int32_t> d_count, d_list, d_trip;
  DevBuf<uint8_t> d_active;
  DevBuf<float> d_mean, d_sdev, d_rad, d_rdepth;
  DevBuf<uint32_t> d_nact;
  CU(d_count.alloc(npix));
  CU(d_active.alloc(npix));
  CU(d_list.alloc(npix));
  CU(d_mean.alloc(4 * (size_t) npix));
  CU(d_sdev.alloc(4 * (size_t) npix));
  CU(d_nact.alloc(1));
  CU(cudaMemsetAsync(d_count.p, 0, d_count.bytes(), st));
  CU(cudaMemsetAsync(d_active.p, 1, d_active.bytes(), st));
  CU(cudaMemsetAsync(d_mean.p, 0, d_mean.bytes(), st));
  CU(cudaMemsetAsync(d_sdev.p, 0, d_sdev.bytes(), st));
  CU(cudaMemsetAsync(sc->stats.p, 0, sc->stats.bytes(), st));
  cudaEvent_t e0, e1;
  CU(cudaEventCreate(&e0));
  CU(cudaEventCreate(&e1));
  CU(cudaEventRecord(e0, st));
  uint64_t launches = 0, traced = 0;
  KernelClock clk{sc, st, stats != nullptr};
  P.mode = 1;
  size_t trip_cap = 0;
  // pixels this shard owns (valid_samples = pixels * spp is 64-bit: 4K x 1024 spp = 8.5e9 must not be narrowed first)
  uint32_t n_active = (uint32_t) sh.owned_pixels;
  for (int round = 0; n_active > 0; ++round) {
    // first round: everyone needs sppMin samples; later rounds look 4 samples ahead
    int chunk = round == 0 ? spp_min : std::min(4, spp_max - spp_min);
    chunk = (int) std::max<unsigned long long>(1ull, std::min<unsigned long long>((unsigned long long) chunk, round_cap / std::max<uint32_t>(1u, n_active)));
    const size_t need = (size_t) n_active * (size_t) chunk;
    if (need > trip_cap) {
      CU(d_trip.alloc(3 * need));
      CU(d_rad.alloc(4 * need));
      CU(d_rdepth.alloc(need));
      trip_cap = need;
    }
    CU(cudaMemsetAsync(d_nact.p, 0, sizeof(uint32_t), st));
    launch_adaptive_build(W, H, P.tile_rank, P.tile_ranks, d_count.p, d_active.p, chunk, d_nact.p, d_list.p, d_trip.p, st);
    uint32_t na = 0;
    CU(cudaMemcpyAsync(&na, d_nact.p, sizeof na, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    launches += 1;
    if ((size_t) na * (size_t) chunk > trip_cap) return fail(QG_EINVAL, "adaptive round larger than its buffers");
    n_active = na;
    if (na == 0) break;
    const size_t np = (size_t) na * (size_t) chunk;
    CU(cudaMemsetAsync(d_rad.p, 0, 4 * np * sizeof(float), st));
    uint32_t wave = (uint32_t) std::min<unsigned long long>(np, round_cap);
    wave = (wave + 31u) & ~31u;
    sc->want_gather = sc->want_gather || P.use_photon_map != 0;
    rc = ensure_wf(sc, (uint32_t) ((unsigned long long) wave * growth), P.carry_diff != 0);
    if (rc) return rc;
    for (size_t base = 0; base < np; base += wave) {
      P.pix_list = d_trip.p + 3 * base;
      P.accum = d_rad.p + 4 * base;
      P.depth_plane = d_rdepth.p + base;
      P.wave_base = 0;
      P.n_paths = (uint32_t) std::min<size_t>(wave, np - base);
      run_wave(sc, P, st, &launches, &clk);
    }
    traced += np;
    launch_adaptive_update(na, d_list.p, chunk, d_rad.p, d_rdepth.p, spp_min, spp_max, d_count.p, d_active.p, d_mean.p, d_sdev.p, d_depth, st);
    launches += 1;
  }
  launch_adaptive_finish(npix, d_count.p, d_mean.p, d_accum, st);
  launches += 1;
  CU(cudaEventRecord(e1, st));
  CU(cudaStreamSynchronize(st));
  CU(cudaGetLastError());
  if (stats) {
    unsigned long long hs[8];
    CU(cudaMemcpy(hs, sc->stats.p, sizeof hs, cudaMemcpyDeviceToHost));
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    memset(stats, 0, sizeof *stats);
    stats->samples = traced;  // samples traced, including look-ahead samples the stop rule then discarded
    stats->closest_rays = hs[0] + traced;
    stats->shadow_rays = hs[1];
    stats->nodes_visited = hs[2]; stats->tris_tested = hs[3]; stats->objects_tested = hs[4];
    stats->shadow_nodes_visited = hs[5]; stats->shadow_tris_tested = hs[6]; stats->shadow_objects_tested = hs[7];
    stats->kernel_launches = launches;
    stats->device_ms = ms;
    double cms[4];
    clk.sums(cms);
    stats->trace_ms = cms[0]; stats->shade_ms = cms[1]; stats->shadow_ms = cms[2];
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  return QG_OK;
}

int resolve_to_host(qgScene *sc, int W, int H, int spp, int srgb, const float *d_accum, const float *d_depth, float *rgb_mean, uint8_t *rgb8,
                    float *depth, uint8_t *sample_count, uint8_t *mask, cudaStream_t st)
{
  const size_t n = (size_t) W * H;
  DevBuf<float> &d_mean = sc->rs_mean, &d_dep = sc->rs_dep;
  DevBuf<uint8_t> &d_rgb8 = sc->rs_rgb8, &d_sc = sc->rs_sc, &d_mask = sc->rs_mask;
  if (rgb_mean) CU(d_mean.ensure(3 * n));
  if (rgb8) CU(d_rgb8.ensure(3 * n));
  if (depth) CU(d_dep.ensure(n));
  if (sample_count) CU(d_sc.ensure(n));
  if (mask) CU(d_mask.ensure(n));
  launch_resolve(d_accum, d_depth, (int) n, spp, srgb, rgb_mean ? d_mean.p : nullptr, rgb8 ? d_rgb8.p : nullptr, depth ? d_dep.p : nullptr,
                 sample_count ? d_sc.p : nullptr, mask ? d_mask.p : nullptr, st);
  if (rgb_mean) CU(cudaMemcpyAsync(rgb_mean, d_mean.p, 3 * n * sizeof(float), cudaMemcpyDeviceToHost, st));
  if (rgb8) CU(cudaMemcpyAsync(rgb8, d_rgb8.p, 3 * n, cudaMemcpyDeviceToHost, st));
  if (depth) CU(cudaMemcpyAsync(depth, d_dep.p, n * sizeof(float), cudaMemcpyDeviceToHost, st));
  if (sample_count) CU(cudaMemcpyAsync(sample_count, d_sc.p, n, cudaMemcpyDeviceToHost, st));
  if (mask) CU(cudaMemcpyAsync(mask, d_mask.p, n, cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  CU(cudaGetLastError());
  return QG_OK;
}

}  // namespace

#include "api_entry.inc"  // the extern "C" entry points
#include "api_multi.inc"  // ... and the multi-GPU ones (NCCL)
#include "api_photon.inc"  // ... and the photon / caustics maps
