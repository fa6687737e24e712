// scene_host.cpp — scene XML -> HostScene, following src/parser/xmlload.cpp of the reference tag by tag so the
// same file gives the same node tree, transforms (bit for bit), materials, lights and textures.
// Compiled with -ffp-contract=off: the matrix arithmetic must round exactly as the reference's glm calls do.
#include "scene_host.h"

#include <strings.h>

#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>

#include "png.h"
#include "xml.h"

namespace qb {

static const float kPI = (float) M_PI;  // src/math/math.cpp:13

// ------------------------------------------------------------------------------------------ Transformation
HostXform::HostXform()
{
  t.tm = m3_identity();
  t.itm = m3_identity();
  t.pos = v3(0, 0, 0);
}
void HostXform::transform(const M3 &m)  // transform.cpp:36-41
{
  t.tm = mul(m, t.tm);
  t.pos = mul(m, t.pos);
  t.itm = inverse(t.tm);
}
void HostXform::translate(V3 p) { t.pos += p; }

// upper-left 3x3 of glm::rotate(mat4(1), angle, axis) (glm/gtc/matrix_transform.inl), evaluated as glm does:
// Result[c] = m[0]*R[c][0] + m[1]*R[c][1] + m[2]*R[c][2] with m = identity
static M3 glm_rotate_identity(float angle, V3 v)
{
  const float a = angle;
  const float c = cosf(a);
  const float s = sinf(a);
  const V3 axis = normalize(v);
  const V3 temp = (1.0f - c) * axis;
  float R[3][3];
  R[0][0] = c + temp.x * axis.x;
  R[0][1] = temp.x * axis.y + s * axis.z;
  R[0][2] = temp.x * axis.z - s * axis.y;
  R[1][0] = temp.y * axis.x - s * axis.z;
  R[1][1] = c + temp.y * axis.y;
  R[1][2] = temp.y * axis.z + s * axis.x;
  R[2][0] = temp.z * axis.x + s * axis.y;
  R[2][1] = temp.z * axis.y - s * axis.x;
  R[2][2] = c + temp.z * axis.z;
  M3 out;
  const float I[3][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
  for (int col = 0; col < 3; ++col)
    for (int r = 0; r < 3; ++r) out.c[col][r] = I[0][r] * R[col][0] + I[1][r] * R[col][1] + I[2][r] * R[col][2];
  return out;
}
void HostXform::rotate(V3 axis, float degree)  // transform.h:62-66
{
  transform(glm_rotate_identity(degree * kPI / 180.0f, axis));
}
void HostXform::scale(float sx, float sy, float sz)  // transform.h:67-71; glm::scale: Result[i] = m[i] * v[i]
{
  M3 m;
  const float I[3][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
  const float s[3] = {sx, sy, sz};
  for (int col = 0; col < 3; ++col)
    for (int r = 0; r < 3; ++r) m.c[col][r] = I[col][r] * s[col];
  transform(m);
}

HostMaterial::HostMaterial()  // MtlBlinn_PhotonMap ctor, MtlBlinn_PhotonMap.cpp:43-55
{
  diffuse.color = v3(0.5f, 0.5f, 0.5f);
  specular.color = v3(0.7f, 0.7f, 0.7f);
}

// ------------------------------------------------------------------------------------------ XML helpers
namespace {

struct Loader {
  HostScene *sc;
  std::string asset_dir;
  struct NodeMtl { int node; std::string name; };
  std::vector<NodeMtl> node_mtl;  // xmlload.cpp:62-67

  void logf(const char *fmt, ...)
  {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    sc->log += buf;
  }
  std::string resolve(const std::string &name) const
  {
    if (name.empty() || name[0] == '/' || asset_dir.empty()) return name;
    return asset_dir + "/" + name;
  }
};

bool tag_is(const XmlElement &e, const char *tag) { return strcasecmp(e.name.c_str(), tag) == 0; }

void read_float(const XmlElement &e, float &f, const char *name = "value")  // xmlload.cpp:563-568
{
  double d = (double) f;
  e.query_double(name, &d);
  f = (float) d;
}
void read_vector(const XmlElement &e, V3 &v)  // xmlload.cpp:527-543
{
  double x = v.x, y = v.y, z = v.z;
  e.query_double("x", &x);
  e.query_double("y", &y);
  e.query_double("z", &z);
  v = v3((float) x, (float) y, (float) z);
  float f = 1;
  read_float(e, f);
  v *= f;
}
void read_color(const XmlElement &e, V3 &c)  // xmlload.cpp:545-561
{
  double r = c.x, g = c.y, b = c.z;
  e.query_double("r", &r);
  e.query_double("g", &g);
  e.query_double("b", &b);
  c = v3((float) r, (float) g, (float) b);
  float f = 1;
  read_float(e, f);
  c *= f;
}

void load_transform(HostXform *xf, const XmlElement &e)  // xmlload.cpp:293-320
{
  for (auto &ch : e.children) {
    if (tag_is(*ch, "scale")) {
      V3 s = v3(1, 1, 1);
      read_vector(*ch, s);
      xf->scale(s.x, s.y, s.z);
    } else if (tag_is(*ch, "rotate")) {
      V3 s = v3(0, 0, 0);
      read_vector(*ch, s);
      s = normalize(s);
      float a = 0.f;  // the reference leaves this uninitialised when `angle` is missing
      read_float(*ch, a, "angle");
      xf->rotate(s, a);
    } else if (tag_is(*ch, "translate")) {
      V3 t = v3(0, 0, 0);
      read_vector(*ch, t);
      xf->translate(t);
    }
  }
}

// ReadTexture(const char*) (xmlload.cpp:612-628): file textures are shared by name; a file that fails to load
// leaves a TextureMap with a null texture, which samples black (core/texture.cpp:67-70)
int read_texture_file(Loader &L, const std::string &name)
{
  for (size_t i = 0; i < L.sc->textures.size(); ++i)
    if (L.sc->textures[i].type == QG_TEX_IMAGE && L.sc->textures[i].name == name) return (int) i;
  HostTexture t;
  t.name = name;
  bool ok = false;
  const size_t len = name.size();
  if (len >= 3) {
    const std::string full = L.resolve(name);
    if (strcasecmp(name.c_str() + len - 3, "png") == 0) {
      std::string err;
      ok = png_decode_rgb8(full.c_str(), &t.rgb, &t.width, &t.height, &err);
    } else if (strcasecmp(name.c_str() + len - 3, "ppm") == 0) {
      FILE *fp = fopen(full.c_str(), "rb");
      if (fp) {  // binary P6, maxval 255 (textures/texture.cpp:28-53)
        int w = 0, h = 0, mx = 0;
        char magic[3] = {0, 0, 0};
        if (fscanf(fp, "%2s %d %d %d", magic, &w, &h, &mx) == 4 && !strcmp(magic, "P6") && mx == 255 && w > 0 && h > 0) {
          fgetc(fp);
          t.rgb.resize((size_t) w * h * 3);
          ok = fread(t.rgb.data(), 1, t.rgb.size(), fp) == t.rgb.size();
          t.width = w;
          t.height = h;
        }
        fclose(fp);
      }
    }
  }
  if (ok) {
    t.type = QG_TEX_IMAGE;
    L.sc->textures.push_back(std::move(t));
    return (int) L.sc->textures.size() - 1;
  }
  L.logf("texture \"%s\" could not be loaded (samples black)\n", name.c_str());
  for (size_t i = 0; i < L.sc->textures.size(); ++i)
    if (L.sc->textures[i].type == QG_TEX_NULL) return (int) i;
  HostTexture nul;
  nul.name = "#null";
  L.sc->textures.push_back(nul);
  return (int) L.sc->textures.size() - 1;
}

// ReadTexture(TiXmlElement*) (xmlload.cpp:575-608): returns false when the element has no texture attribute
bool read_texture(Loader &L, const XmlElement &e, HostTexColor *tc)
{
  tc->texture = -1;
  tc->xf = HostXform();
  const char *tex_name = e.attribute("texture");
  if (!tex_name) return false;
  if (strcasecmp(tex_name, "checkerboard") == 0) {
    HostTexture t;
    t.name = tex_name;
    t.type = QG_TEX_CHECKER;
    for (auto &ch : e.children) {
      if (tag_is(*ch, "color1")) { V3 c = v3(0, 0, 0); read_color(*ch, c); t.color1 = c; }
      else if (tag_is(*ch, "color2")) { V3 c = v3(0, 0, 0); read_color(*ch, c); t.color2 = c; }
    }
    L.sc->textures.push_back(t);
    tc->texture = (int) L.sc->textures.size() - 1;
  } else {
    tc->texture = read_texture_file(L, tex_name);
  }
  load_transform(&tc->xf, e);
  return true;
}

}  // namespace

namespace {

void set_gloss(float *dst, float gloss) { *dst = gloss > 0.00001f ? gloss : -1.f; }  // MtlBlinn_PhotonMap.cpp:56-63

void load_material(Loader &L, const XmlElement &e)  // xmlload.cpp:324-399
{
  const char *name = e.attribute("name");
  const char *type = e.attribute("type");
  if (!type || strcasecmp(type, "blinn") != 0) {
    L.logf("material [%s] - UNKNOWN\n", name ? name : "");
    return;  // only `blinn` creates a material in the reference
  }
  HostMaterial m;
  m.name = name ? name : "";
  for (auto &chp : e.children) {
    const XmlElement &ch = *chp;
    V3 c = v3(1, 1, 1);
    float f = 1;
    if (tag_is(ch, "diffuse")) {
      read_color(ch, c);
      m.diffuse.color = c;
      read_texture(L, ch, &m.diffuse);
    } else if (tag_is(ch, "specular")) {
      read_color(ch, c);
      m.specular.color = c;
      read_texture(L, ch, &m.specular);
    } else if (tag_is(ch, "glossiness")) {
      read_float(ch, f);
      m.glossiness = f;
    } else if (tag_is(ch, "emission")) {
      read_color(ch, c);
      m.emission.color = c;
      read_texture(L, ch, &m.emission);
    } else if (tag_is(ch, "reflection")) {
      read_color(ch, c);
      m.reflection.color = c;
      read_texture(L, ch, &m.reflection);
      f = 0;
      read_float(ch, f, "glossiness");
      set_gloss(&m.reflection_glossiness, f);
    } else if (tag_is(ch, "refraction")) {
      read_color(ch, c);
      m.refraction.color = c;
      read_float(ch, f, "index");
      m.ior = f;
      read_texture(L, ch, &m.refraction);
      f = 0;
      read_float(ch, f, "glossiness");
      set_gloss(&m.refraction_glossiness, f);
    } else if (tag_is(ch, "absorption")) {
      read_color(ch, c);
      m.absorption = c;
    }
  }
  L.sc->materials.push_back(m);
}

// 4-vector helpers for SpotLight::SetRotation (lights.cpp:115-119): normalize(rotate(mat4(1), radians(deg), axis) * (0,0,-1,0))
V3 spot_direction(float degree, V3 axis)
{
  const M3 R = glm_rotate_identity(degree * 0.01745329251994329576923690768489f, axis);
  // mat4 * vec4 as glm evaluates it: (m0*v0 + m1*v1) + (m2*v2 + m3*v3), 4th row/column of the rotation = (0,0,0,1)
  const float v[4] = {0.f, 0.f, -1.f, 0.f};
  float r[4];
  for (int k = 0; k < 4; ++k) {
    const float m0 = k < 3 ? R.c[0][k] : 0.f, m1 = k < 3 ? R.c[1][k] : 0.f, m2 = k < 3 ? R.c[2][k] : 0.f, m3 = k < 3 ? 0.f : 1.f;
    r[k] = (m0 * v[0] + m1 * v[1]) + (m2 * v[2] + m3 * v[3]);
  }
  const float d = (r[0] * r[0] + r[1] * r[1]) + (r[2] * r[2] + r[3] * r[3]);  // glm dot(vec4)
  const float inv = 1.0f / sqrtf(d);
  return v3(r[0] * inv, r[1] * inv, r[2] * inv);
}
void spot_set_angle(HostLight *l, float s)  // lights.cpp:120-123
{
  s = qmin(qmax(s / 2.f, 1.f), 89.f) / 180.f * kPI;
  l->outer = tanf(s);
}
void spot_set_blend(HostLight *l, float s)  // lights.cpp:124-127
{
  l->blend = qmax(qmin(s, 1.f), 0.f);
  l->inner = sqrtf(l->outer * l->outer * (1.f - l->blend));
}

void load_light(Loader &L, const XmlElement &e)  // xmlload.cpp:403-523
{
  const char *name = e.attribute("name");
  const char *type = e.attribute("type");
  if (!type) return;
  HostLight l;
  l.name = name ? name : "";
  if (strcasecmp(type, "ambient") == 0) {
    l.type = QG_LIGHT_AMBIENT;
    for (auto &ch : e.children)
      if (tag_is(*ch, "intensity")) { V3 c = v3(1, 1, 1); read_color(*ch, c); l.intensity = c; }
  } else if (strcasecmp(type, "direct") == 0) {
    l.type = QG_LIGHT_DIRECT;
    l.direction = v3(0, 0, 1);
    for (auto &ch : e.children) {
      if (tag_is(*ch, "intensity")) { V3 c = v3(1, 1, 1); read_color(*ch, c); l.intensity = c; }
      else if (tag_is(*ch, "direction")) { V3 v = v3(1, 1, 1); read_vector(*ch, v); l.direction = normalize(v); }
    }
  } else if (strcasecmp(type, "point") == 0) {
    l.type = QG_LIGHT_POINT;
    for (auto &ch : e.children) {
      if (tag_is(*ch, "intensity")) { V3 c = v3(1, 1, 1); read_color(*ch, c); l.intensity = c; }
      else if (tag_is(*ch, "position")) { V3 v = v3(0, 0, 0); read_vector(*ch, v); l.position = v; }
      else if (tag_is(*ch, "size")) { float f = 0; read_float(*ch, f); l.size = f; }
    }
  } else if (strcasecmp(type, "spot") == 0) {
    l.type = QG_LIGHT_SPOT;
    l.direction = v3(1, 0, 0);  // lights.h:126
    spot_set_angle(&l, 45);
    spot_set_blend(&l, 1.f);
    for (auto &ch : e.children) {
      if (tag_is(*ch, "intensity")) { V3 c = v3(1, 1, 1); read_color(*ch, c); l.intensity = c; }
      else if (tag_is(*ch, "position")) { V3 v = v3(0, 0, 0); read_vector(*ch, v); l.position = v; }
      else if (tag_is(*ch, "size")) { float f = 0; read_float(*ch, f); l.size = f; }
      else if (tag_is(*ch, "rotation")) {
        V3 v = v3(0, 0, 0);
        float f = 0;
        read_vector(*ch, v);
        read_float(*ch, f, "angle");
        l.direction = spot_direction(f, v);
      }
      else if (tag_is(*ch, "angle")) { float f = 0; read_float(*ch, f); spot_set_angle(&l, f); }
      else if (tag_is(*ch, "blend")) { float f = 0; read_float(*ch, f); spot_set_blend(&l, f); }
    }
  } else {
    return;
  }
  L.sc->lights.push_back(l);
}

int find_material(const HostScene &sc, const std::string &name)  // MaterialList::Find: first match by name
{
  for (size_t i = 0; i < sc.materials.size(); ++i)
    if (sc.materials[i].name == name && !sc.materials[i].name.empty()) return (int) i;
  return -1;
}

// TriObj::Load (objects.h:65-73) + the MultiMtl generation of LoadNode (xmlload.cpp:222-272)
int load_mesh(Loader &L, int node_index, const std::string &name, bool has_mtl_attr)
{
  for (size_t i = 0; i < L.sc->mesh_names.size(); ++i)
    if (L.sc->mesh_names[i] == name) return (int) i;  // objList.Find(name): instances share the mesh
  std::unique_ptr<ObjMesh> mesh(new ObjMesh);
  std::string err;
  if (!load_obj(L.resolve(name).c_str(), mesh.get(), &err)) {
    L.logf(" -- ERROR: Cannot load file \"%s.\" %s\n", name.c_str(), err.c_str());
    return -1;
  }
  if (mesh->NVN() == 0) mesh_compute_normals(mesh.get());
  mesh_compute_bounds(mesh.get());
  const int mesh_index = (int) L.sc->meshes.size();
  // TriMesh::GetDirectoryName() as the reference sees it: the directory part of the name given in the XML
  std::string xml_dir = name;
  {
    const size_t p = xml_dir.find_last_of("/\\");
    xml_dir = p == std::string::npos ? std::string() : xml_dir.substr(0, p + 1);
  }
  if (!has_mtl_attr && !mesh->materials.empty() && find_material(*L.sc, name) < 0) {
    HostMaterial mm;
    mm.kind = QG_MTL_MULTI;
    mm.name = name;
    std::vector<HostMaterial> subs;
    for (const ObjMaterial &om : mesh->materials) {
      HostMaterial m;
      m.diffuse.color = v3(om.diffuse[0], om.diffuse[1], om.diffuse[2]);
      m.specular.color = v3(om.specular[0], om.specular[1], om.specular[2]);
      m.glossiness = om.shininess;
      m.ior = om.ior;
      if (!om.diffuse_texname.empty()) m.diffuse.texture = read_texture_file(L, xml_dir + om.diffuse_texname);
      // the reference assigns the specular map to the DIFFUSE slot (xmlload.cpp:249-252); kept
      if (!om.specular_texname.empty()) m.diffuse.texture = read_texture_file(L, xml_dir + om.specular_texname);
      if (om.illum > 2 && om.illum <= 7) {
        m.reflection.color = v3(om.specular[0], om.specular[1], om.specular[2]);
        if (!om.specular_texname.empty()) m.reflection.texture = read_texture_file(L, xml_dir + om.specular_texname);
        if (om.illum >= 6)
          m.refraction.color = v3(1.f - om.transmittance[0], 1.f - om.transmittance[1], 1.f - om.transmittance[2]);
      }
      subs.push_back(m);
    }
    // the MultiMtl takes its place in the scene material list now; its sub-materials are unnamed entries
    L.sc->materials.push_back(mm);
    const int mm_index = (int) L.sc->materials.size() - 1;
    for (auto &s : subs) {
      L.sc->materials.push_back(s);
      L.sc->materials[mm_index].subs.push_back((int) L.sc->materials.size() - 1);
    }
    Loader::NodeMtl nm;
    nm.node = node_index;
    nm.name = name;
    L.node_mtl.push_back(nm);
  }
  L.sc->meshes.push_back(std::move(mesh));
  L.sc->mesh_names.push_back(name);
  return mesh_index;
}

void load_node(Loader &L, int parent, const XmlElement &e, int level)  // xmlload.cpp:188-289
{
  const int me = (int) L.sc->nodes.size();
  L.sc->nodes.emplace_back();
  {
    HostNode &n = L.sc->nodes[me];
    n.parent = parent;
    const char *name = e.attribute("name");
    n.name = name ? name : "";
  }
  const char *mtl_name = e.attribute("material");
  if (mtl_name) {
    Loader::NodeMtl nm;
    nm.node = me;
    nm.name = mtl_name;
    L.node_mtl.push_back(nm);
    L.sc->nodes[me].mtl_name = mtl_name;
  }
  const char *type = e.attribute("type");
  if (type) {
    if (strcasecmp(type, "sphere") == 0) L.sc->nodes[me].object_type = QG_OBJ_SPHERE;
    else if (strcasecmp(type, "plane") == 0) L.sc->nodes[me].object_type = QG_OBJ_PLANE;
    else if (strcasecmp(type, "obj") == 0) {
      const std::string nm = L.sc->nodes[me].name;
      const int mesh = load_mesh(L, me, nm, mtl_name != nullptr);
      if (mesh >= 0) {
        L.sc->nodes[me].object_type = QG_OBJ_MESH;
        L.sc->nodes[me].mesh = mesh;
      }
    } else {
      L.logf("object [%s] - UNKNOWN TYPE\n", L.sc->nodes[me].name.c_str());
    }
  }
  for (auto &ch : e.children)
    if (tag_is(*ch, "object")) load_node(L, me, *ch, level + 1);
  load_transform(&L.sc->nodes[me].xf, e);
}

}  // namespace

// ------------------------------------------------------------------------------------------ LoadScene
bool load_scene_xml(const char *xml_path, const char *asset_dir, HostScene *scene, std::string *err)
{
  XmlParser parser;
  std::unique_ptr<XmlElement> doc = parser.parse_file(xml_path);
  if (!doc) {
    if (err) *err = std::string("Failed to load the file \"") + xml_path + "\": " + parser.error();
    return false;
  }
  const XmlElement *xml = doc->first_child("xml");
  if (!xml) { if (err) *err = "No \"xml\" tag found."; return false; }
  const XmlElement *sc = xml->first_child("scene");
  if (!sc) { if (err) *err = "No \"scene\" tag found."; return false; }
  const XmlElement *cam = xml->first_child("camera");
  if (!cam) { if (err) *err = "No \"camera\" tag found."; return false; }

  *scene = HostScene();
  Loader L;
  L.sc = scene;
  L.asset_dir = asset_dir ? asset_dir : "";
  while (!L.asset_dir.empty() && L.asset_dir.back() == '/' && L.asset_dir.size() > 1) L.asset_dir.pop_back();
  scene->nodes.emplace_back();  // root node: no object, identity transform (Node::Init)

  for (auto &chp : sc->children) {  // LoadScene(TiXmlElement*) xmlload.cpp:160-185
    const XmlElement &ch = *chp;
    if (tag_is(ch, "background")) {
      V3 c = v3(1, 1, 1);
      read_color(ch, c);
      scene->background.color = c;
      read_texture(L, ch, &scene->background);
    } else if (tag_is(ch, "environment")) {
      V3 c = v3(1, 1, 1);
      read_color(ch, c);
      scene->environment.color = c;
      read_texture(L, ch, &scene->environment);
    } else if (tag_is(ch, "object")) {
      load_node(L, 0, ch, 0);
    } else if (tag_is(ch, "material")) {
      load_material(L, ch);
    } else if (tag_is(ch, "light")) {
      load_light(L, ch);
    }
  }
  // assign materials (xmlload.cpp:106-112); later entries for the same node overwrite earlier ones
  for (auto &nm : L.node_mtl) {
    const int m = find_material(*scene, nm.name);
    if (m >= 0) scene->nodes[nm.node].material = m;
  }
  // instances of a multi-material OBJ that carry no material attribute: the reference leaves them without a
  // material (null dereference when shaded); they get the OBJ's MultiMtl here.
  for (auto &n : scene->nodes)
    if (n.object_type == QG_OBJ_MESH && n.material < 0 && n.mtl_name.empty()) n.material = find_material(*scene, n.name);

  // camera (xmlload.cpp:114-144, camera.cpp:31-41)
  V3 pos = v3(0, 0, 0), dir = v3(0, 0, -1), up = v3(0, 1, 0);
  float fovy = 40, focal = 1, dof = 0;
  int W = 200, H = 150;
  dir += pos;
  for (auto &chp : cam->children) {
    const XmlElement &ch = *chp;
    if (tag_is(ch, "position")) read_vector(ch, pos);
    else if (tag_is(ch, "target")) read_vector(ch, dir);
    else if (tag_is(ch, "up")) read_vector(ch, up);
    else if (tag_is(ch, "fov")) read_float(ch, fovy);
    else if (tag_is(ch, "focaldist")) read_float(ch, focal);
    else if (tag_is(ch, "dof")) read_float(ch, dof);
    else if (tag_is(ch, "width")) ch.query_int("value", &W);
    else if (tag_is(ch, "height")) ch.query_int("value", &H);
  }
  dir = dir - pos;
  dir = normalize(dir);
  const V3 x = cross(dir, up);
  up = normalize(cross(x, dir));
  qgCamera &c = scene->camera;
  c.pos[0] = pos.x; c.pos[1] = pos.y; c.pos[2] = pos.z;
  c.dir[0] = dir.x; c.dir[1] = dir.y; c.dir[2] = dir.z;
  c.up[0] = up.x; c.up[1] = up.y; c.up[2] = up.z;
  c.fovy = fovy; c.focal = focal; c.dof = dof;
  c.width = W; c.height = H;
  return true;
}

// ------------------------------------------------------------------------------------------ mesh post-processing
void mesh_compute_bounds(ObjMesh *m)  // TriMesh.cpp:117-133
{
  if (m->NV() > 0) {
    for (int k = 0; k < 3; ++k) m->bound_min[k] = m->bound_max[k] = m->vertices[k];
    for (size_t i = 1; i < m->NV(); ++i)
      for (int k = 0; k < 3; ++k) {
        const float v = m->vertices[3 * i + k];
        if (m->bound_min[k] > v) m->bound_min[k] = v;
        if (m->bound_max[k] < v) m->bound_max[k] = v;
      }
  } else {
    for (int k = 0; k < 3; ++k) { m->bound_min[k] = 1; m->bound_max[k] = 0; }
  }
}

// TriMesh::ComputeNormals (TriMesh.cpp:134-158): area-weighted vertex normals, normal index := vertex index.
// The reference's zeroing and normalising loops run over NF() where NV() is meant: with NF > NV it writes past
// the array (undefined behaviour, not reproducible); with NF <= NV only the first NF normals get normalised.
// Here the loops are clamped to the array (min(NF, NV)), which is the reference's defined behaviour.
void mesh_compute_normals(ObjMesh *m)
{
  m->normals.assign(3 * m->NV(), 0.f);
  auto V = [&](int i) { return v3(m->vertices[3 * i], m->vertices[3 * i + 1], m->vertices[3 * i + 2]); };
  for (ObjFace &f : m->faces) {
    const V3 N = cross(V(f.idx[1].v) - V(f.idx[0].v), V(f.idx[2].v) - V(f.idx[0].v));
    for (int k = 0; k < 3; ++k) {
      float *n = &m->normals[3 * f.idx[k].v];
      n[0] += N.x; n[1] += N.y; n[2] += N.z;
      f.idx[k].n = f.idx[k].v;
    }
  }
  const size_t lim = m->NF() < m->NV() ? m->NF() : m->NV();
  for (size_t i = 0; i < lim; ++i) {
    float *n = &m->normals[3 * i];
    const V3 u = normalize(v3(n[0], n[1], n[2]));
    n[0] = u.x; n[1] = u.y; n[2] = u.z;
  }
}

// ------------------------------------------------------------------------------------------ cyBVH topology
namespace {
struct BvhBuilder {
  const ObjMesh &mesh;
  unsigned max_leaf;
  std::vector<uint32_t> &elements;
  struct Temp { int c1 = -1, c2 = -1; float box[6]; unsigned count, offset; };
  std::vector<Temp> temps;

  BvhBuilder(const ObjMesh &m, unsigned ml, std::vector<uint32_t> &el) : mesh(m), max_leaf(ml), elements(el) {}

  float vtx(unsigned face, int corner, int dim) const { return mesh.vertices[3 * mesh.faces[face].idx[corner].v + dim]; }
  void elem_bounds(unsigned i, float box[6]) const  // TriBVH.cpp:33-46
  {
    for (int k = 0; k < 3; ++k) box[k] = box[k + 3] = vtx(i, 0, k);
    for (int j = 1; j < 3; ++j)
      for (int k = 0; k < 3; ++k) {
        const float p = vtx(i, j, k);
        if (box[k] > p) box[k] = p;
        if (box[k + 3] < p) box[k + 3] = p;
      }
  }
  float elem_center(unsigned i, int dim) const  // TriBVH.cpp:49-56
  {
    return (vtx(i, 0, dim) + vtx(i, 1, dim) + vtx(i, 2, dim)) / 3.0f;
  }
  static void box_init(float b[6]) { b[0] = b[1] = b[2] = 1e30f; b[3] = b[4] = b[5] = -1e30f; }
  static void box_add(float b[6], const float o[6])
  {
    for (int i = 0; i < 3; ++i) {
      if (b[i] > o[i]) b[i] = o[i];
      if (b[i + 3] < o[i + 3]) b[i + 3] = o[i + 3];
    }
  }
  unsigned mean_split(unsigned count, uint32_t *el, const float *box) const  // cyBVH.h:381-421
  {
    if (count <= max_leaf) return 0;
    const float d[3] = {box[3] - box[0], box[4] - box[1], box[5] - box[2]};
    unsigned sd[3];
    sd[0] = d[0] >= d[1] ? (d[0] >= d[2] ? 0 : 2) : (d[1] >= d[2] ? 1 : 2);
    sd[1] = (sd[0] + 1) % 3;
    sd[2] = (sd[0] + 2) % 3;
    if (d[sd[1]] < d[sd[2]]) { const unsigned t = sd[1]; sd[1] = sd[2]; sd[2] = t; }
    unsigned child1 = 0;
    for (int s = 0; s < 3; ++s) {
      const unsigned dim = sd[s];
      const float split = 0.5f * (box[dim] + box[dim + 3]);
      unsigned i = 0, j = count;
      while (i < j) {
        const float center = elem_center(el[i], (int) dim);
        if (center <= split) {
          i++;
        } else {
          j--;
          const uint32_t t = el[i]; el[i] = el[j]; el[j] = t;
        }
      }
      if (i < count && i > 0) { child1 = i; break; }
    }
    return child1;
  }
  void split(int ti)  // cyBVH.h:322-360 (explicit stack instead of recursion; child 1 subtree first, as there)
  {
    std::vector<int> todo{ti};
    while (!todo.empty()) {
      const int t = todo.back();
      todo.pop_back();
      const unsigned count = temps[t].count, offset = temps[t].offset;
      uint32_t *el = &elements[offset];
      unsigned c1 = mean_split(count, el, temps[t].box);
      if (c1 == 0 || c1 >= count) {
        if (count > 8) c1 = count / 2;  // CY_BVH_MAX_ELEMENT_COUNT
        else continue;
      }
      Temp a, b;
      box_init(a.box);
      box_init(b.box);
      float eb[6];
      for (unsigned i = 0; i < c1; ++i) { elem_bounds(el[i], eb); box_add(a.box, eb); }
      for (unsigned i = c1; i < count; ++i) { elem_bounds(el[i], eb); box_add(b.box, eb); }
      a.count = c1; a.offset = offset;
      b.count = count - c1; b.offset = offset + c1;
      temps[t].c1 = (int) temps.size(); temps.push_back(a);
      temps[t].c2 = (int) temps.size(); temps.push_back(b);
      todo.push_back(temps[t].c2);
      todo.push_back(temps[t].c1);
    }
  }
};
}  // namespace

void build_cy_bvh(const ObjMesh &mesh, unsigned max_leaf, CyBvh *out)  // cyBVH.h:144-166
{
  out->nodes.clear();
  out->elements.clear();
  const unsigned n = (unsigned) mesh.NF();
  if (n == 0) return;
  if (max_leaf > 8) max_leaf = 8;
  out->elements.resize(n);
  for (unsigned i = 0; i < n; ++i) out->elements[i] = i;
  BvhBuilder B(mesh, max_leaf, out->elements);
  BvhBuilder::Temp root;
  BvhBuilder::box_init(root.box);
  float eb[6];
  for (unsigned i = 0; i < n; ++i) { B.elem_bounds(i, eb); BvhBuilder::box_add(root.box, eb); }
  root.count = n;
  root.offset = 0;
  B.temps.push_back(root);
  B.split(0);
  // ConvertTempData (cyBVH.h:362-377): siblings adjacent, child-1 subtree numbered before child 2's children
  out->nodes.resize(B.temps.size() + 1);
  struct Item { unsigned node_id; int temp; };
  // emulate the recursion: convert(nodeID, t, childIndex) returns the next free index
  struct Frame { unsigned node_id; int temp; unsigned child_index; int stage; };
  std::vector<Frame> st;
  st.push_back({1u, 0, 2u, 0});
  unsigned ret = 0;
  while (!st.empty()) {
    Frame &f = st.back();
    const BvhBuilder::Temp &t = B.temps[f.temp];
    CyBvh::Node &nd = out->nodes[f.node_id];
    if (t.c1 < 0) {
      memcpy(nd.box, t.box, sizeof nd.box);
      nd.data = (t.offset & 0x0FFFFFFFu) | ((t.count - 1) << 28) | 0x80000000u;
      ret = f.child_index;
      st.pop_back();
      continue;
    }
    if (f.stage == 0) {
      memcpy(nd.box, t.box, sizeof nd.box);
      nd.data = f.child_index & 0x7FFFFFFFu;
      f.stage = 1;
      st.push_back({f.child_index, t.c1, f.child_index + 2, 0});
    } else if (f.stage == 1) {
      f.stage = 2;
      st.push_back({f.child_index + 1, t.c2, ret, 0});
    } else {
      st.pop_back();  // ret already holds child 2's return value
    }
  }
}

// ------------------------------------------------------------------------------------------ C-ABI form
static void put_xform(const Xform &t, qgXform *o)
{
  for (int c = 0; c < 3; ++c)
    for (int r = 0; r < 3; ++r) {
      o->tm[3 * c + r] = t.tm.c[c][r];
      o->itm[3 * c + r] = t.itm.c[c][r];
    }
  o->pos[0] = t.pos.x; o->pos[1] = t.pos.y; o->pos[2] = t.pos.z;
}
static void put_texcolor(const HostTexColor &h, qgTexturedColor *o)
{
  o->color[0] = h.color.x; o->color[1] = h.color.y; o->color[2] = h.color.z;
  o->texture = h.texture;
  put_xform(h.xf.t, &o->xf);
}

qgSceneDesc FlatScene::desc() const
{
  qgSceneDesc d;
  memset(&d, 0, sizeof d);
  d.nodes = nodes.data(); d.num_nodes = (int32_t) nodes.size();
  d.meshes = meshes.data(); d.num_meshes = (int32_t) meshes.size();
  d.materials = materials.data(); d.num_materials = (int32_t) materials.size();
  d.multi = multi.data(); d.num_multi = (int32_t) multi.size();
  d.lights = lights.data(); d.num_lights = (int32_t) lights.size();
  d.textures = textures.data(); d.num_textures = (int32_t) textures.size();
  d.background = background;
  d.environment = environment;
  d.camera = camera;
  return d;
}

void flatten(const HostScene &hs, FlatScene *out)
{
  *out = FlatScene();
  for (const HostNode &n : hs.nodes) {
    qgNode q;
    q.parent = n.parent;
    q.object_type = n.object_type;
    q.mesh = n.mesh;
    q.material = n.material;
    put_xform(n.xf.t, &q.xf);
    out->nodes.push_back(q);
  }
  for (const auto &mp : hs.meshes) {
    const ObjMesh &m = *mp;
    qgMesh q;
    memset(&q, 0, sizeof q);
    out->f_store.push_back(m.vertices);   q.vertices = out->f_store.back().data();
    out->f_store.push_back(m.normals);    q.normals = out->f_store.back().data();
    out->f_store.push_back(m.texcoords);  q.texcoords = out->f_store.back().data();
    std::vector<int32_t> fi(9 * m.NF()), fm(m.NF());
    for (size_t f = 0; f < m.NF(); ++f) {
      for (int k = 0; k < 3; ++k) {
        fi[9 * f + 3 * k + 0] = m.faces[f].idx[k].v;
        fi[9 * f + 3 * k + 1] = m.faces[f].idx[k].n;
        fi[9 * f + 3 * k + 2] = m.faces[f].idx[k].t;
      }
      fm[f] = m.faces[f].mtl;
    }
    out->i_store.push_back(std::move(fi)); q.face_indices = out->i_store.back().data();
    out->i_store.push_back(std::move(fm)); q.face_material = out->i_store.back().data();
    q.num_vertices = (int32_t) m.NV();
    q.num_normals = (int32_t) m.NVN();
    q.num_texcoords = (int32_t) (m.texcoords.size() / 2);
    q.num_faces = (int32_t) m.NF();
    for (int k = 0; k < 3; ++k) { q.bound_min[k] = m.bound_min[k]; q.bound_max[k] = m.bound_max[k]; }
    out->meshes.push_back(q);
  }
  for (const HostMaterial &m : hs.materials) {
    qgMaterial q;
    memset(&q, 0, sizeof q);
    q.kind = m.kind;
    put_texcolor(m.diffuse, &q.diffuse);
    put_texcolor(m.specular, &q.specular);
    put_texcolor(m.emission, &q.emission);
    put_texcolor(m.reflection, &q.reflection);
    put_texcolor(m.refraction, &q.refraction);
    q.absorption[0] = m.absorption.x; q.absorption[1] = m.absorption.y; q.absorption[2] = m.absorption.z;
    q.ior = m.ior;
    q.glossiness = m.glossiness;
    q.reflection_glossiness = m.reflection_glossiness;
    q.refraction_glossiness = m.refraction_glossiness;
    q.kill = m.kill;
    q.sub_first = (int32_t) out->multi.size();
    q.sub_count = (int32_t) m.subs.size();
    for (int s : m.subs) out->multi.push_back(s);
    out->materials.push_back(q);
  }
  for (const HostLight &l : hs.lights) {
    qgLight q;
    memset(&q, 0, sizeof q);
    q.type = l.type;
    q.intensity[0] = l.intensity.x; q.intensity[1] = l.intensity.y; q.intensity[2] = l.intensity.z;
    q.position[0] = l.position.x; q.position[1] = l.position.y; q.position[2] = l.position.z;
    q.direction[0] = l.direction.x; q.direction[1] = l.direction.y; q.direction[2] = l.direction.z;
    q.size = l.size;
    q.inner = l.inner;
    q.outer = l.outer;
    out->lights.push_back(q);
  }
  for (const HostTexture &t : hs.textures) {
    qgTexture q;
    memset(&q, 0, sizeof q);
    q.type = t.type;
    q.color1[0] = t.color1.x; q.color1[1] = t.color1.y; q.color1[2] = t.color1.z;
    q.color2[0] = t.color2.x; q.color2[1] = t.color2.y; q.color2[2] = t.color2.z;
    q.width = t.width;
    q.height = t.height;
    out->b_store.push_back(t.rgb);
    q.rgb8 = out->b_store.back().data();
    out->textures.push_back(q);
  }
  put_texcolor(hs.background, &out->background);
  put_texcolor(hs.environment, &out->environment);
  out->camera = hs.camera;
}

void copy_desc(const qgSceneDesc &d, FlatScene *out)
{
  *out = FlatScene();
  out->nodes.assign(d.nodes, d.nodes + d.num_nodes);
  for (int i = 0; i < d.num_meshes; ++i) {
    qgMesh q = d.meshes[i];
    out->f_store.emplace_back(q.vertices, q.vertices + 3 * (size_t) q.num_vertices); q.vertices = out->f_store.back().data();
    out->f_store.emplace_back(q.normals, q.normals + 3 * (size_t) q.num_normals);    q.normals = out->f_store.back().data();
    if (q.num_texcoords > 0 && q.texcoords) out->f_store.emplace_back(q.texcoords, q.texcoords + 2 * (size_t) q.num_texcoords);
    else out->f_store.emplace_back();
    q.texcoords = out->f_store.back().data();
    out->i_store.emplace_back(q.face_indices, q.face_indices + 9 * (size_t) q.num_faces); q.face_indices = out->i_store.back().data();
    out->i_store.emplace_back(q.face_material, q.face_material + (size_t) q.num_faces);   q.face_material = out->i_store.back().data();
    out->meshes.push_back(q);
  }
  out->materials.assign(d.materials, d.materials + d.num_materials);
  if (d.num_multi > 0) out->multi.assign(d.multi, d.multi + d.num_multi);
  out->lights.assign(d.lights, d.lights + d.num_lights);
  for (int i = 0; i < d.num_textures; ++i) {
    qgTexture q = d.textures[i];
    if (q.type == QG_TEX_IMAGE && q.rgb8) out->b_store.emplace_back(q.rgb8, q.rgb8 + (size_t) q.width * q.height * 3);
    else out->b_store.emplace_back();
    q.rgb8 = out->b_store.back().data();
    out->textures.push_back(q);
  }
  out->background = d.background;
  out->environment = d.environment;
  out->camera = d.camera;
}

}  // namespace qb
