// objload.h — Wavefront OBJ/MTL reader producing what the reference's TriMesh holds after
// TriMesh::LoadFromFileObj (src/mesh/TriMesh.cpp:63-116), which goes through tinyobjloader v1.1.0
// (a vendored third-party library of the reference, external/tinyobjloader/tiny_obj_loader.h) with
// triangulate=true. The semantics restated here, because they decide vertex bits, face order and face ids:
//   * numbers are parsed by tinyobj's own decimal parser (tiny_obj_loader.h:498-611), not strtod;
//   * polygons become triangle fans (i0, i[k-1], i[k]) (tiny_obj_loader.h:974-1030);
//   * indices are 1-based, negative = relative to the count so far (tiny_obj_loader.h:432-452);
//   * a face takes the material active at its `f` line; unknown usemtl names give -1;
//   * faces are then std::sort'ed by material id with the reference's comparator (TriMesh.cpp:107-114) —
//     std::sort is not stable and the comparator is not a strict weak order when ids are negative, so the
//     resulting order is whatever libstdc++'s introsort does; the same call is made here.
#pragma once
#include <string>
#include <vector>

namespace qb {

struct ObjIndex {
  int v, n, t;  // -1 when absent
};

struct ObjMaterial {
  std::string name;
  float diffuse[3], specular[3], transmittance[3];
  float shininess, ior;
  int illum;
  std::string diffuse_texname, specular_texname;
};

struct ObjFace {
  ObjIndex idx[3];
  int mtl;       // material id or -1
  int file_idx;  // position in file order (the reference's TriFace::idx)
};

struct ObjMesh {
  std::string dir, name, file;          // directory (with trailing '/'), base name, full path
  std::vector<float> vertices;          // xyz
  std::vector<float> normals;           // xyz
  std::vector<float> texcoords;         // uv
  std::vector<ObjMaterial> materials;
  std::vector<ObjFace> faces;           // sorted as the reference sorts them
  float bound_min[3], bound_max[3];

  size_t NV() const { return vertices.size() / 3; }
  size_t NVN() const { return normals.size() / 3; }
  size_t NF() const { return faces.size(); }
};

// Loads file (path relative to the current directory, like the reference). load_mtl mirrors the reference's
// flag (ignored there as well: tinyobj always follows mtllib). Returns false and fills err on failure.
bool load_obj(const char *filename, ObjMesh *mesh, std::string *err);

// tinyobj's decimal parser; exposed for the unit test.
bool tinyobj_parse_double(const char *s, const char *s_end, double *result);

}  // namespace qb
