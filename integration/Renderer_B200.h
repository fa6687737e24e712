// integration/Renderer_B200.h — the binding a qaray maintainer adds as src/renderers/Renderer_B200.h: a third subclass of
// qaray::Renderer (src/renderers/renderer.h:70-111) next to Renderer_GUI / Renderer_MPI, driven by the same five calls
// (src/main.cpp:55-59), that renders on a B200 through include/qaray_b200.h.
//
//   Flattened       qaray::Scene (src/scene/scene.h:55-67) -> qgSceneDesc: node tree in pre-order, TriObj meshes, MtlBlinn /
//                   MultiMtl materials, textures, lights, camera. Nothing in qaray's classes changes; the flattening reads
//                   a handful of private members (material colours, TextureMap::texture, light parameters), so either add
//                   `friend struct qaray::Flattened;` to MtlBlinn_PhotonMap, TextureMap, TextureFile, TextureChecker and the
//                   four light classes, or build this one translation unit with their access opened (what oracle/Makefile
//                   does to compile it against the unmodified sources: `make -C oracle binding`).
//   Renderer_B200   Init / ComputeScene (inherited) / Render / Terminate over qg_init / qg_scene_create / qg_render.
//
// This file is compiled and run by tests/test_binding.py (tables equal to qg_scene_load_xml's on every fixture scene, hits
// equal to the fixtures, frames equal to the library's own loader path); it is not part of libqaray_b200.so.
#pragma once
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <string>
#include <vector>

#include "renderers/renderer.h"
#include "scene/scene.h"
#include "objects/objects.h"
#include "materials/materials.h"
#include "lights/lights.h"
#include "textures/texture.h"

#include "qaray_b200.h"

namespace qaray {

struct Flattened {
  qgSceneDesc desc;
  std::vector<qgNode> nodes;
  std::vector<qgMesh> meshes;
  std::vector<qgMaterial> materials;
  std::vector<int32_t> multi;
  std::vector<qgLight> lights;
  std::vector<qgTexture> textures;
  // storage behind the descriptors' pointers
  std::vector<std::vector<float>> fstore;
  std::vector<std::vector<int32_t>> istore;
  std::vector<std::vector<uint8_t>> bstore;
  std::map<const Object *, int> meshIndex;
  std::map<const Material *, int> materialIndex;
  std::map<const Texture *, int> textureIndex;
  int nullTexture = -1;

  static void PutXform(const Transformation &t, qgXform *o)
  {
    const Matrix3 &tm = t.GetTransform(), &itm = t.GetInverseTransform();
    const Point3 &pos = t.GetPosition();
    for (int c = 0; c < 3; ++c)
      for (int r = 0; r < 3; ++r) {  // column-major like glm::mat3
        o->tm[3 * c + r] = tm[c][r];
        o->itm[3 * c + r] = itm[c][r];
      }
    o->pos[0] = pos.x; o->pos[1] = pos.y; o->pos[2] = pos.z;
  }

  int AddTexture(const Texture *tex)
  {
    if (!tex) {  // a TextureMap whose file failed to load keeps a null texture and samples black (core/texture.cpp:60-66)
      if (nullTexture < 0) {
        qgTexture q;
        memset(&q, 0, sizeof q);
        q.type = QG_TEX_NULL;
        q.color2[0] = q.color2[1] = q.color2[2] = 1.f;
        textures.push_back(q);
        bstore.emplace_back();
        nullTexture = (int) textures.size() - 1;
      }
      return nullTexture;
    }
    auto it = textureIndex.find(tex);
    if (it != textureIndex.end()) return it->second;
    qgTexture q;
    memset(&q, 0, sizeof q);
    q.color2[0] = q.color2[1] = q.color2[2] = 1.f;
    bstore.emplace_back();
    if (const TextureChecker *ck = dynamic_cast<const TextureChecker *>(tex)) {
      q.type = QG_TEX_CHECKER;
      q.color1[0] = ck->color1.r; q.color1[1] = ck->color1.g; q.color1[2] = ck->color1.b;
      q.color2[0] = ck->color2.r; q.color2[1] = ck->color2.g; q.color2[2] = ck->color2.b;
    } else if (const TextureFile *tf = dynamic_cast<const TextureFile *>(tex)) {
      q.type = QG_TEX_IMAGE;
      q.width = tf->width;
      q.height = tf->height;
      std::vector<uint8_t> &px = bstore.back();
      px.resize(tf->data.size() * 3);
      for (size_t i = 0; i < tf->data.size(); ++i) { px[3 * i] = tf->data[i].r; px[3 * i + 1] = tf->data[i].g; px[3 * i + 2] = tf->data[i].b; }
    }
    textures.push_back(q);
    textureIndex[tex] = (int) textures.size() - 1;
    return (int) textures.size() - 1;
  }

  void PutTexturedColor(const TexturedColor &tc, qgTexturedColor *o)
  {
    memset(o, 0, sizeof *o);
    const Color3f c = tc.GetColor();
    o->color[0] = c.r; o->color[1] = c.g; o->color[2] = c.b;
    const TextureMap *map = tc.GetTexture();
    if (map) {
      o->texture = AddTexture(map->texture);
      PutXform(*map, &o->xf);
    } else {
      o->texture = -1;
      PutXform(Transformation(), &o->xf);
    }
  }

  void PutBlinn(const MtlBlinn &m, qgMaterial *q)
  {
    memset(q, 0, sizeof *q);
    q->kind = QG_MTL_BLINN;
    PutTexturedColor(m.diffuse, &q->diffuse);
    PutTexturedColor(m.specular, &q->specular);
    PutTexturedColor(m.emission, &q->emission);
    PutTexturedColor(m.reflection, &q->reflection);
    PutTexturedColor(m.refraction, &q->refraction);
    q->absorption[0] = m.absorption.r; q->absorption[1] = m.absorption.g; q->absorption[2] = m.absorption.b;
    q->ior = m.ior;
    q->glossiness = m.specularGlossiness;
    q->reflection_glossiness = m.reflectionGlossiness;
    q->refraction_glossiness = m.refractionGlossiness;
    q->kill = m.kill;
  }

  // scene.materials in list order; the sub-materials of a MultiMtl follow it as unnamed entries
  void AddMaterials(const MaterialList &list)
  {
    for (const Material *mtl : list) {
      if (const MultiMtl *mm = dynamic_cast<const MultiMtl *>(mtl)) {
        qgMaterial q;
        PutBlinn(MtlBlinn(), &q);  // a MultiMtl carries no parameters of its own
        q.kind = QG_MTL_MULTI;
        const size_t at = materials.size();
        materials.push_back(q);
        materialIndex[mtl] = (int) at;
        std::vector<int32_t> subs;
        for (const Material *s : mm->mtls) {
          qgMaterial qs;
          PutBlinn(*static_cast<const MtlBlinn *>(s), &qs);
          materials.push_back(qs);
          subs.push_back((int32_t) materials.size() - 1);
        }
        materials[at].sub_first = (int32_t) multi.size();
        materials[at].sub_count = (int32_t) subs.size();
        multi.insert(multi.end(), subs.begin(), subs.end());
      } else {
        qgMaterial q;
        PutBlinn(*static_cast<const MtlBlinn *>(mtl), &q);
        materialIndex[mtl] = (int) materials.size();
        materials.push_back(q);
      }
    }
    int32_t off = 0;
    for (qgMaterial &q : materials) {  // blinn entries: an empty range at the running offset
      if (q.kind == QG_MTL_MULTI) off = q.sub_first + q.sub_count;
      else { q.sub_first = off; q.sub_count = 0; }
    }
  }

  int AddMesh(const TriObj *obj)
  {
    auto it = meshIndex.find(obj);
    if (it != meshIndex.end()) return it->second;
    const TriMesh &m = *obj;
    qgMesh q;
    memset(&q, 0, sizeof q);
    std::vector<float> v(3 * m.NV()), n(3 * m.NVN()), t(2 * m.NVT());
    for (size_t i = 0; i < m.NV(); ++i) { v[3 * i] = m.V((int) i).x; v[3 * i + 1] = m.V((int) i).y; v[3 * i + 2] = m.V((int) i).z; }
    for (size_t i = 0; i < m.NVN(); ++i) { n[3 * i] = m.VN((int) i).x; n[3 * i + 1] = m.VN((int) i).y; n[3 * i + 2] = m.VN((int) i).z; }
    for (size_t i = 0; i < m.NVT(); ++i) { t[2 * i] = m.VT((int) i).x; t[2 * i + 1] = m.VT((int) i).y; }
    std::vector<int32_t> fi(9 * m.NF()), fm(m.NF());
    for (size_t f = 0; f < m.NF(); ++f) {  // faces as TriMesh holds them after its material sort (TriMesh.cpp:107-114)
      for (int k = 0; k < 3; ++k) {
        fi[9 * f + 3 * k + 0] = m.F(f).v[k]->vertex_index;
        fi[9 * f + 3 * k + 1] = m.F(f).v[k]->normal_index;
        fi[9 * f + 3 * k + 2] = m.F(f).v[k]->texcoord_index;
      }
      fm[f] = m.F(f).mtl;
    }
    fstore.push_back(std::move(v)); q.vertices = fstore.back().data();
    fstore.push_back(std::move(n)); q.normals = fstore.back().data();
    fstore.push_back(std::move(t)); q.texcoords = fstore.back().data();
    istore.push_back(std::move(fi)); q.face_indices = istore.back().data();
    istore.push_back(std::move(fm)); q.face_material = istore.back().data();
    q.num_vertices = (int32_t) m.NV(); q.num_normals = (int32_t) m.NVN(); q.num_texcoords = (int32_t) m.NVT(); q.num_faces = (int32_t) m.NF();
    const vec3f lo = m.GetBoundMin(), hi = m.GetBoundMax();
    q.bound_min[0] = lo.x; q.bound_min[1] = lo.y; q.bound_min[2] = lo.z;
    q.bound_max[0] = hi.x; q.bound_max[1] = hi.y; q.bound_max[2] = hi.z;
    meshes.push_back(q);
    meshIndex[obj] = (int) meshes.size() - 1;
    return (int) meshes.size() - 1;
  }

  // pre-order = the order Scene::TraceNodeNormal visits (src/scene/scene.cpp:50-74)
  void AddNode(Node *n, int parent)
  {
    qgNode q;
    memset(&q, 0, sizeof q);
    q.parent = parent;
    q.mesh = -1;
    Object *o = n->GetNodeObj();
    if (o == &theSphere) q.object_type = QG_OBJ_SPHERE;
    else if (o == &thePlane) q.object_type = QG_OBJ_PLANE;
    else if (o != nullptr) { q.object_type = QG_OBJ_MESH; q.mesh = AddMesh(static_cast<TriObj *>(o)); }
    const Material *mtl = n->GetMaterial();
    q.material = -1;
    if (mtl) {
      auto it = materialIndex.find(mtl);
      if (it != materialIndex.end()) q.material = it->second;
    }
    PutXform(*n, &q.xf);
    const int me = (int) nodes.size();
    nodes.push_back(q);
    for (int c = 0; c < n->GetNumChild(); ++c) AddNode(n->GetChild(c), me);
  }

  void AddLights(const LightList &list)
  {
    for (const Light *l : list) {
      qgLight q;
      memset(&q, 0, sizeof q);
      q.direction[2] = 1.f;
      if (const AmbientLight *a = dynamic_cast<const AmbientLight *>(l)) {
        q.type = QG_LIGHT_AMBIENT;
        q.intensity[0] = a->intensity.r; q.intensity[1] = a->intensity.g; q.intensity[2] = a->intensity.b;
      } else if (const DirectLight *d = dynamic_cast<const DirectLight *>(l)) {
        q.type = QG_LIGHT_DIRECT;
        q.intensity[0] = d->intensity.r; q.intensity[1] = d->intensity.g; q.intensity[2] = d->intensity.b;
        q.direction[0] = d->direction.x; q.direction[1] = d->direction.y; q.direction[2] = d->direction.z;
      } else if (const SpotLight *s = dynamic_cast<const SpotLight *>(l)) {
        q.type = QG_LIGHT_SPOT;
        q.intensity[0] = s->intensity.r; q.intensity[1] = s->intensity.g; q.intensity[2] = s->intensity.b;
        q.position[0] = s->position.x; q.position[1] = s->position.y; q.position[2] = s->position.z;
        q.direction[0] = s->direction.x; q.direction[1] = s->direction.y; q.direction[2] = s->direction.z;
        q.size = s->size; q.inner = s->inner; q.outer = s->outer;
      } else if (const PointLight *p = dynamic_cast<const PointLight *>(l)) {
        q.type = QG_LIGHT_POINT;
        q.intensity[0] = p->intensity.r; q.intensity[1] = p->intensity.g; q.intensity[2] = p->intensity.b;
        q.position[0] = p->position.x; q.position[1] = p->position.y; q.position[2] = p->position.z;
        q.size = p->size;
      }
      lights.push_back(q);
    }
  }

  explicit Flattened(Scene &s)
  {
    memset(&desc, 0, sizeof desc);
    AddMaterials(s.materials);
    AddNode(&s.rootNode, -1);
    AddLights(s.lights);
    PutTexturedColor(s.background, &desc.background);
    PutTexturedColor(s.environment, &desc.environment);
    const Camera &c = s.camera;
    qgCamera &q = desc.camera;
    q.pos[0] = c.pos.x; q.pos[1] = c.pos.y; q.pos[2] = c.pos.z;
    q.dir[0] = c.dir.x; q.dir[1] = c.dir.y; q.dir[2] = c.dir.z;
    q.up[0] = c.up.x; q.up[1] = c.up.y; q.up[2] = c.up.z;
    q.fovy = c.fovy; q.focal = c.focalDistance; q.dof = c.depthOfField;
    q.width = c.imgWidth; q.height = c.imgHeight;
    desc.nodes = nodes.data(); desc.num_nodes = (int32_t) nodes.size();
    desc.meshes = meshes.data(); desc.num_meshes = (int32_t) meshes.size();
    desc.materials = materials.data(); desc.num_materials = (int32_t) materials.size();
    desc.multi = multi.data(); desc.num_multi = (int32_t) multi.size();
    desc.lights = lights.data(); desc.num_lights = (int32_t) lights.size();
    for (size_t i = 0; i < textures.size(); ++i) textures[i].rgb8 = bstore[i].empty() ? nullptr : bstore[i].data();
    desc.textures = textures.data(); desc.num_textures = (int32_t) textures.size();
  }
};


// The renderer: main.cpp selects it next to Renderer_GUI / Renderer_MPI (`-b200`), then runs the usual
//   renderer->Init(); LoadScene(file); renderer->ComputeScene(renderImage, scene); renderer->Render(); renderer->Terminate();
class Renderer_B200 : public Renderer {
 protected:
  qgScene *gpu = nullptr;
  int device = 0;
 public:
  uint32_t seed = 0x5EED;  // the library's counter-based streams need a seed; the reference seeds from the clock
  qgRenderStats stats;
  qgPhotonStats photonStats;
  explicit Renderer_B200(RendererParam &p, int cudaDevice = 0) : Renderer(p), device(cudaDevice)
  {
    memset(&stats, 0, sizeof stats);
    memset(&photonStats, 0, sizeof photonStats);
  }
  qgScene *GpuScene() { return gpu; }
  void StartTimer() override {}
  void StopTimer() override {}
  void KillTimer() override {}
  void Init() override
  {
    if (qg_init(device) != QG_OK) {
      fprintf(stderr, "Renderer_B200::Init: %s\n", qg_last_error());
      exit(EXIT_FAILURE);
    }
  }
  // ComputeScene (renderer.cpp:71-291) has run: scene, image, colorBuffer .. maskBuffer, pixelW / pixelH, mpiRank / mpiSize
  bool Upload()
  {
    if (gpu) return true;
    Flattened f(*scene);
    if (qg_scene_create(&f.desc, &gpu) != QG_OK) {
      fprintf(stderr, "Renderer_B200: %s\n", qg_last_error());
      return false;
    }
    if (param.usePhotonMap && param.photonMapSize > 0 && param.causticsMapSize > 0) {
      // the base ComputeScene has already run the reference's own emission pass on the host; the device maps are built by
      // the library from the same parameters (renderer.cpp:123-129)
      qgPhotonParams pp;
      pp.photon_map_size = (int32_t) param.photonMapSize; pp.photon_map_bounce = (int32_t) param.photonMapBounce;
      pp.photon_map_radius = param.photonMapRadius;
      pp.caustics_map_size = (int32_t) param.causticsMapSize; pp.caustics_map_bounce = (int32_t) param.causticsMapBounce;
      pp.caustics_map_radius = param.causticsMapRadius;
      pp.seed = seed;
      if (qg_photon_maps_build(gpu, &pp, &photonStats) != QG_OK) {
        fprintf(stderr, "Renderer_B200: %s\n", qg_last_error());
        return false;
      }
    }
    return true;
  }
  void Render() override
  {
    if (!Upload()) return;
    qgRenderParams rp;
    memset(&rp, 0, sizeof rp);
    rp.spp = (int32_t) param.sppMax;  // `-spp N` sets min = max (main.cpp:21-24)
    rp.spp_min = param.sppMin < param.sppMax ? (int32_t) param.sppMin : 0;  // SuperSamplerHalton's adaptive stop (scene.cpp:92-97)
    rp.material_class = QG_CLASS_PHOTONMAP;  // what `using MtlBlinn = ...` names in this build (materials.h:57-61)
    rp.max_bounce = Material::maxBounce;     // src/core/material.cpp:31
    rp.srgb = param.useSRGB ? 1 : 0;
    rp.seed = seed;
    rp.width = (int32_t) pixelW;
    rp.height = (int32_t) pixelH;
    rp.tile_rank = (int32_t) mpiRank;        // renderer.cpp:383-387
    rp.tile_ranks = (int32_t) mpiSize;
    rp.use_photon_map = (param.usePhotonMap && param.photonMapSize > 0 && param.causticsMapSize > 0) ? 1 : 0;
    StartTimer();
    // the planes PixelRender fills (renderer.cpp:348-365): Color3c RGB8, depth of sample 0, 255 * s / sppMax, mask
    if (qg_render(gpu, &rp, nullptr, reinterpret_cast<uint8_t *>(colorBuffer), depthBuffer, sampleCountBuffer, maskBuffer, &stats) != QG_OK)
      fprintf(stderr, "Renderer_B200::Render: %s\n", qg_last_error());
    StopTimer();
  }
  void Terminate() override
  {
    if (gpu) qg_scene_destroy(gpu);
    gpu = nullptr;
    Renderer::Terminate();
  }
};

}  // namespace qaray
