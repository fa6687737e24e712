// scene_host.h — host-side scene: what the reference holds in qaray::Scene after LoadScene
// (src/scene/scene.h:55-67, src/parser/xmlload.cpp), in flat vectors ready for the C ABI's qgSceneDesc.
#pragma once
#include <cstdint>
#include <memory>
#include <string>
#include <vector>

#include "../../include/qaray_b200.h"
#include "objload.h"
#include "qmath.h"

namespace qb {

// Transformation (src/core/transform.h:36-79) with the reference's update rules
struct HostXform {
  Xform t;
  HostXform();
  void translate(V3 p);
  void rotate(V3 axis, float degree);
  void scale(float sx, float sy, float sz);
  void transform(const M3 &m);
};

struct HostTexture {
  std::string name;
  int type = QG_TEX_NULL;
  V3 color1 = {0, 0, 0}, color2 = {1, 1, 1};
  int width = 0, height = 0;
  std::vector<uint8_t> rgb;
};

struct HostTexColor {
  V3 color = {0, 0, 0};
  int texture = -1;  // -1 = no TextureMap
  HostXform xf;
};

struct HostMaterial {
  std::string name;
  int kind = QG_MTL_BLINN;
  HostTexColor diffuse, specular, emission, reflection, refraction;
  V3 absorption = {0, 0, 0};
  float ior = 1.f, glossiness = 20.f, reflection_glossiness = 0.f, refraction_glossiness = 0.f, kill = 0.1f;
  std::vector<int> subs;  // MULTI
  HostMaterial();
};

struct HostLight {
  std::string name;
  int type = QG_LIGHT_AMBIENT;
  V3 intensity = {0, 0, 0}, position = {0, 0, 0}, direction = {0, 0, 1};
  float size = 0.f, inner = 0.f, outer = 0.f, blend = 0.f;
};

struct HostNode {
  std::string name, mtl_name;
  int parent = -1;
  int object_type = QG_OBJ_NONE;
  int mesh = -1;
  int material = -1;
  HostXform xf;
};

struct HostScene {
  std::vector<HostNode> nodes;  // pre-order, node 0 = root
  std::vector<std::unique_ptr<ObjMesh>> meshes;
  std::vector<std::string> mesh_names;
  std::vector<HostMaterial> materials;   // named scene materials first come first found; MULTI subs appended unnamed
  std::vector<HostLight> lights;
  std::vector<HostTexture> textures;
  HostTexColor background, environment;
  qgCamera camera;
  std::string log;  // what the reference prints while loading, kept instead of written to stdout
};

// LoadScene (src/parser/xmlload.cpp:71-149). Relative asset names resolve against asset_dir ("" = cwd).
bool load_scene_xml(const char *xml_path, const char *asset_dir, HostScene *scene, std::string *err);

// TriMesh::ComputeNormals / ComputeBoundingBox (src/mesh/TriMesh.cpp:117-158)
void mesh_compute_normals(ObjMesh *m);
void mesh_compute_bounds(ObjMesh *m);

// BVH in the reference's topology (src/ext/cyBVH.h:144-166,322-421 via src/mesh/TriBVH.cpp:32-56, leaf size 4)
struct CyBvh {
  struct Node {
    float box[6];
    uint32_t data;  // leaf: bit31 | (count-1) << 28 | element offset ; internal: first child index
    bool leaf() const { return (data & 0x80000000u) != 0; }
    uint32_t child() const { return data & 0x7FFFFFFFu; }
    uint32_t count() const { return ((data >> 28) & 7u) + 1u; }
    uint32_t offset() const { return data & 0x0FFFFFFFu; }
  };
  std::vector<Node> nodes;         // nodes[0] unused, root = 1
  std::vector<uint32_t> elements;  // face ids in leaf order
};
void build_cy_bvh(const ObjMesh &mesh, unsigned max_leaf, CyBvh *out);

// Owning copy of a scene in C-ABI form.
struct FlatScene {
  std::vector<qgNode> nodes;
  std::vector<qgMesh> meshes;
  std::vector<qgMaterial> materials;
  std::vector<int32_t> multi;
  std::vector<qgLight> lights;
  std::vector<qgTexture> textures;
  qgTexturedColor background, environment;
  qgCamera camera;
  // storage the descriptors point into
  std::vector<std::vector<float>> f_store;
  std::vector<std::vector<int32_t>> i_store;
  std::vector<std::vector<uint8_t>> b_store;
  qgSceneDesc desc() const;
};
void flatten(const HostScene &hs, FlatScene *out);
void copy_desc(const qgSceneDesc &d, FlatScene *out);  // deep copy of caller-owned buffers

}  // namespace qb
