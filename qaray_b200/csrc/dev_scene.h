// dev_scene.h — device-resident scene tables (HBM layout) shared by the traversal and shading kernels.
//
//   nodes[]      scene-tree nodes in pre-order (src/scene/scene.cpp:50-74 visits them in this order)
//   obj_node[]   the nodes that hold an object, pre-order; chain[] = for each, its ancestors root-first (root excluded)
//   bvh[]        64-byte internal nodes: both children's boxes + two child refs (4 x LDG.128 per visit)
//   tris[]       64-byte triangles in leaf order: A,B,C in xyz, the unit face normal in the three w lanes, and the
//                per-triangle constants of IntersectTriangle (1/area, dropped axis) + the face word
//   tri_face[]   global face id of each triangle slot; face_*[] per-face shading attributes (read once per hit)
#pragma once
#define QB_ANC_WORDS 7  // stack entries per lane the ancestor-ray cache of the traversal takes (traverse.cuh, QB_ANC_CACHE)
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/qaray_b200.h"
#include "qmath.h"

namespace qb {

#define QB_LEAF_BIT 0x80000000u
#define QB_BIGFLOAT 1.0e30f
#define QB_BIAS 0.005f
#define QB_MAX_DEPTH 8      // scene-tree depth supported by the chain replay
#define QB_BLOCK 128        // threads per block of every persistent kernel
#ifndef QB_MIN_BLOCKS
#define QB_MIN_BLOCKS 8      // resident blocks per SM of the closest-hit kernels: 64 registers, possible since the lanes' cold state lives in shared memory (traverse.cuh)
#endif
#ifndef QB_MIN_BLOCKS_ANY
#define QB_MIN_BLOCKS_ANY 8  // any-hit kernels keep less state: 64 registers without spills, 7 % faster than at 6 (measured)
#endif
#ifndef QB_STACK
#define QB_STACK 40         // deferred far children per ray = max BVH depth supported (the reference keeps 40 entries too, objects.cpp:331)
#endif

struct DNode {
  float4 i0, i1, i2;  // itm column k in xyz, pos[k] in w      (ToNodeCoords)
  float4 t0, t1, t2;  // tm column k in xyz                     (FromNodeCoords)
  int32_t type, mesh, material, parent;
};

struct DMesh {
  float bmin[3], bmax[3];
  uint32_t root_ref;    // child-ref encoding, see DBvhNode
  uint32_t face_base;   // first global face id of this mesh
  uint32_t num_faces;
  uint32_t pad;
};

// child ref: bit31 = leaf; leaf: bits 28..30 = count-1, bits 0..27 = first triangle slot; internal: index into bvh[]
struct DBvhNode {
  float4 a;  // c0.min.xyz, c0.max.x
  float4 b;  // c0.max.yz, c1.min.xy
  float4 c;  // c1.min.z, c1.max.xyz
  uint4 d;   // child0 ref, child1 ref, 0, 0
};

struct DTri {       // 64 B
  float4 a, b, c;  // A.xyz|N.x  B.xyz|N.y  C.xyz|N.z
  float4 d;        // 1 / TriangleArea(axis, A, B, C), axis (int bits), tri_face word (uint bits), 0
};

struct DFace {      // 64 B
  float n[9];       // vertex normals of the three corners
  float uv[6];      // texture coordinates (valid when has_uv)
  int32_t mtl_uv;   // (mtl + 1) << 1 | has_uv   (mtl -1 -> 0)
};

struct DTexColor {
  float color[3];
  int32_t texture;     // -1 none
  float itm[9], pos[3];
  int32_t identity;    // transform is the identity (skip the matrix)
  float pad[3];
};

struct DTexture {
  int32_t type, width, height;
  uint32_t texel_base;  // into texels[] (uchar4 RGBX)
  float c1[3], c2[3];
};

struct DMaterial {
  int32_t kind, sub_first, sub_count, has_maps;
  DTexColor diffuse, specular, emission, reflection, refraction;
  float absorption[3];
  float ior, gloss, refl_gloss, refr_gloss, kill;
};

struct DLight {
  int32_t type;
  float intensity[3], position[3], direction[3];
  float size, inner, outer;
};

struct DScene {
  const DNode *nodes;
  const int32_t *obj_node;
  const int32_t *chain_off;   // num_obj + 1
  const int32_t *chain;       // ancestor node ids, root-first, root excluded, the object node last
  const DMesh *meshes;
  const DBvhNode *bvh;
  const DTri *tris;
  const uint32_t *tri_face;
  const DFace *faces;
  const DMaterial *materials;
  const int32_t *multi;
  const DLight *lights;
  const DTexture *textures;
  const uchar4 *texels;
  int32_t num_nodes, num_obj, num_meshes, num_materials, num_lights, num_textures;
  int32_t num_nonambient;
  int32_t bvh_stack;  // deferred far children a ray can hold at once = deepest mesh BVH of the scene + 1 (sizes the pool's stacks),
                      // plus QB_ANC_WORDS entries when anc_cache is set
  int32_t anc_cache;  // 1: the top QB_ANC_WORDS stack entries of a lane cache its ray in a shared ancestor's space (traverse.cuh)
  DTexColor background, environment;
};

}  // namespace qb
