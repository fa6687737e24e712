// objload.cpp — see objload.h. Compiled with -ffp-contract=off.
#include "objload.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>

namespace qb {

static inline bool is_digit(char c) { return c >= '0' && c <= '9'; }
static inline bool is_space(char c) { return c == ' ' || c == '\t'; }
static inline bool is_newline(char c) { return c == '\r' || c == '\n' || c == '\0'; }

// Restatement of tinyobjloader v1.1.0 tryParseDouble (tiny_obj_loader.h:498-611): sign, integer digits
// accumulated as mantissa*10+d, fraction digits added as d * 10^-k (table for k < 8, pow beyond), optional
// exponent applied as ldexp(mantissa * 5^e, e).
bool tinyobj_parse_double(const char *s, const char *s_end, double *result)
{
  if (s >= s_end) return false;
  double mantissa = 0.0;
  int exponent = 0;
  char sign = '+', exp_sign = '+';
  const char *curr = s;
  int read = 0;
  bool end_not_reached = false;

  if (*curr == '+' || *curr == '-') {
    sign = *curr;
    curr++;
  } else if (!is_digit(*curr)) {
    return false;
  }
  end_not_reached = (curr != s_end);
  while (end_not_reached && is_digit(*curr)) {
    mantissa *= 10;
    mantissa += (int) (*curr - '0');
    curr++;
    read++;
    end_not_reached = (curr != s_end);
  }
  if (read == 0) return false;
  if (end_not_reached) {
    bool has_exp = false;
    if (*curr == '.') {
      curr++;
      read = 1;
      end_not_reached = (curr != s_end);
      while (end_not_reached && is_digit(*curr)) {
        static const double pow_lut[] = {1.0, 0.1, 0.01, 0.001, 0.0001, 0.00001, 0.000001, 0.0000001};
        const int lut_entries = (int) (sizeof pow_lut / sizeof pow_lut[0]);
        mantissa += (int) (*curr - '0') * (read < lut_entries ? pow_lut[read] : std::pow(10.0, -read));
        read++;
        curr++;
        end_not_reached = (curr != s_end);
      }
      has_exp = end_not_reached && (*curr == 'e' || *curr == 'E');
    } else if (*curr == 'e' || *curr == 'E') {
      has_exp = true;
    }
    if (has_exp) {
      curr++;
      end_not_reached = (curr != s_end);
      if (end_not_reached && (*curr == '+' || *curr == '-')) {
        exp_sign = *curr;
        curr++;
      } else if (!is_digit(*curr)) {
        return false;
      }
      read = 0;
      end_not_reached = (curr != s_end);
      while (end_not_reached && is_digit(*curr)) {
        exponent *= 10;
        exponent += (int) (*curr - '0');
        curr++;
        read++;
        end_not_reached = (curr != s_end);
      }
      exponent *= (exp_sign == '+' ? 1 : -1);
      if (read == 0) return false;
    }
  }
  *result = (sign == '+' ? 1 : -1) * (exponent ? std::ldexp(mantissa * std::pow(5.0, exponent), exponent) : mantissa);
  return true;
}

static float parse_real(const char **token, double default_value = 0.0)
{
  (*token) += strspn(*token, " \t");
  const char *end = (*token) + strcspn(*token, " \t\r");
  double val = default_value;
  tinyobj_parse_double(*token, end, &val);
  *token = end;
  return (float) val;
}

static bool fix_index(int idx, int n, int *ret)
{
  if (idx > 0) { *ret = idx - 1; return true; }
  if (idx == 0) return false;
  *ret = n + idx;
  return true;
}

// i, i/j, i//k, i/j/k (tiny_obj_loader.h:747-798)
static bool parse_triple(const char **token, int vsize, int vnsize, int vtsize, ObjIndex *out)
{
  ObjIndex vi = {-1, -1, -1};
  if (!fix_index(atoi(*token), vsize, &vi.v)) return false;
  (*token) += strcspn(*token, "/ \t\r");
  if ((*token)[0] != '/') { *out = vi; return true; }
  (*token)++;
  if ((*token)[0] == '/') {
    (*token)++;
    if (!fix_index(atoi(*token), vnsize, &vi.n)) return false;
    (*token) += strcspn(*token, "/ \t\r");
    *out = vi;
    return true;
  }
  if (!fix_index(atoi(*token), vtsize, &vi.t)) return false;
  (*token) += strcspn(*token, "/ \t\r");
  if ((*token)[0] != '/') { *out = vi; return true; }
  (*token)++;
  if (!fix_index(atoi(*token), vnsize, &vi.n)) return false;
  (*token) += strcspn(*token, "/ \t\r");
  *out = vi;
  return true;
}

static bool read_lines(const std::string &path, std::vector<std::string> *lines)
{
  FILE *fp = fopen(path.c_str(), "rb");
  if (!fp) return false;
  std::string buf;
  char tmp[1 << 16];
  size_t n;
  while ((n = fread(tmp, 1, sizeof tmp, fp)) > 0) buf.append(tmp, n);
  fclose(fp);
  size_t b = 0;
  while (b < buf.size()) {
    size_t e = b;
    while (e < buf.size() && buf[e] != '\n' && buf[e] != '\r') ++e;
    lines->emplace_back(buf, b, e - b);
    if (e < buf.size() && buf[e] == '\r' && e + 1 < buf.size() && buf[e + 1] == '\n') ++e;
    b = e + 1;
  }
  return true;
}

static void init_material(ObjMaterial *m)
{
  m->name.clear();
  m->diffuse_texname.clear();
  m->specular_texname.clear();
  for (int i = 0; i < 3; ++i) m->diffuse[i] = m->specular[i] = m->transmittance[i] = 0.f;
  m->illum = 0;
  m->shininess = 1.f;
  m->ior = 1.f;
}

// tiny_obj_loader.h:1049-1432 for the fields the reference consumes (xmlload.cpp:232-272)
static bool load_mtl(const std::string &path, std::vector<ObjMaterial> *materials, std::map<std::string, int> *map)
{
  std::vector<std::string> lines;
  if (!read_lines(path, &lines)) return false;
  ObjMaterial material;
  init_material(&material);
  for (std::string &linebuf : lines) {
    if (!linebuf.empty()) linebuf = linebuf.substr(0, linebuf.find_last_not_of(" \t") + 1);
    if (linebuf.empty()) continue;
    const char *token = linebuf.c_str();
    token += strspn(token, " \t");
    if (token[0] == '\0' || token[0] == '#') continue;
    if (0 == strncmp(token, "newmtl", 6) && is_space(token[6])) {
      if (!material.name.empty()) {
        map->insert(std::make_pair(material.name, (int) materials->size()));
        materials->push_back(material);
      }
      init_material(&material);
      material.name = token + 7;
      continue;
    }
    if (token[0] == 'K' && token[1] == 'd' && is_space(token[2])) {
      token += 2;
      for (int i = 0; i < 3; ++i) material.diffuse[i] = parse_real(&token);
      continue;
    }
    if (token[0] == 'K' && token[1] == 's' && is_space(token[2])) {
      token += 2;
      for (int i = 0; i < 3; ++i) material.specular[i] = parse_real(&token);
      continue;
    }
    if ((token[0] == 'K' && token[1] == 't' && is_space(token[2])) || (token[0] == 'T' && token[1] == 'f' && is_space(token[2]))) {
      token += 2;
      for (int i = 0; i < 3; ++i) material.transmittance[i] = parse_real(&token);
      continue;
    }
    if (token[0] == 'N' && token[1] == 'i' && is_space(token[2])) { token += 2; material.ior = parse_real(&token); continue; }
    if (token[0] == 'N' && token[1] == 's' && is_space(token[2])) { token += 2; material.shininess = parse_real(&token); continue; }
    if (0 == strncmp(token, "illum", 5) && is_space(token[5])) {
      token += 6;
      token += strspn(token, " \t");
      material.illum = atoi(token);
      continue;
    }
    if (0 == strncmp(token, "map_Kd", 6) && is_space(token[6])) {
      token += 7;
      token += strspn(token, " \t");
      material.diffuse_texname = token;
      continue;
    }
    if (0 == strncmp(token, "map_Ks", 6) && is_space(token[6])) {
      token += 7;
      token += strspn(token, " \t");
      material.specular_texname = token;
      continue;
    }
  }
  map->insert(std::make_pair(material.name, (int) materials->size()));
  materials->push_back(material);
  return true;
}

bool load_obj(const char *filename, ObjMesh *mesh, std::string *err)
{
  // ComputePath (src/mesh/TriMesh.cpp:47-60)
  std::string fpath = filename;
  std::replace(fpath.begin(), fpath.end(), '\\', '/');
  size_t p = fpath.find_last_of('/');
  if (p != std::string::npos) {
    mesh->dir = fpath.substr(0, p + 1);
    mesh->name = fpath.substr(p + 1);
  } else {
    mesh->dir.clear();
    mesh->name = fpath;
  }
  mesh->file = fpath;
  mesh->vertices.clear();
  mesh->normals.clear();
  mesh->texcoords.clear();
  mesh->materials.clear();
  mesh->faces.clear();

  std::vector<std::string> lines;
  if (!read_lines(fpath, &lines)) {
    if (err) *err = "Cannot open file [" + fpath + "]";
    return false;
  }
  std::map<std::string, int> material_map;
  int material = -1;
  std::vector<ObjIndex> face;
  for (const std::string &linebuf : lines) {
    if (linebuf.empty()) continue;
    const char *token = linebuf.c_str();
    token += strspn(token, " \t");
    if (token[0] == '\0' || token[0] == '#') continue;
    if (token[0] == 'v' && is_space(token[1])) {
      token += 2;
      for (int i = 0; i < 3; ++i) mesh->vertices.push_back(parse_real(&token));
      continue;
    }
    if (token[0] == 'v' && token[1] == 'n' && is_space(token[2])) {
      token += 3;
      for (int i = 0; i < 3; ++i) mesh->normals.push_back(parse_real(&token));
      continue;
    }
    if (token[0] == 'v' && token[1] == 't' && is_space(token[2])) {
      token += 3;
      for (int i = 0; i < 2; ++i) mesh->texcoords.push_back(parse_real(&token));
      continue;
    }
    if (token[0] == 'f' && is_space(token[1])) {
      token += 2;
      token += strspn(token, " \t");
      face.clear();
      while (!is_newline(token[0])) {
        ObjIndex vi;
        if (!parse_triple(&token, (int) (mesh->vertices.size() / 3), (int) (mesh->normals.size() / 3),
                          (int) (mesh->texcoords.size() / 2), &vi)) {
          if (err) *err = "Failed parse `f' line(e.g. zero value for face index).";
          return false;
        }
        face.push_back(vi);
        token += strspn(token, " \t\r");
      }
      for (size_t k = 2; k < face.size(); ++k) {  // triangle fan
        ObjFace f;
        f.idx[0] = face[0];
        f.idx[1] = face[k - 1];
        f.idx[2] = face[k];
        f.mtl = material;
        f.file_idx = (int) mesh->faces.size();
        mesh->faces.push_back(f);
      }
      continue;
    }
    if (0 == strncmp(token, "usemtl", 6) && is_space(token[6])) {
      std::string namebuf = token + 7;
      auto it = material_map.find(namebuf);
      material = it != material_map.end() ? it->second : -1;
      continue;
    }
    if (0 == strncmp(token, "mtllib", 6) && is_space(token[6])) {
      std::string rest = token + 7;
      size_t b = 0;
      while (b <= rest.size()) {  // SplitString(' ')
        size_t e = rest.find(' ', b);
        if (e == std::string::npos) e = rest.size();
        std::string fn = rest.substr(b, e - b);
        if (!fn.empty() && load_mtl(mesh->dir + fn, &mesh->materials, &material_map)) break;
        b = e + 1;
      }
      continue;
    }
  }
  // the reference's face sort (TriMesh.cpp:107-114); same libstdc++ call, same comparator
  std::sort(mesh->faces.begin(), mesh->faces.end(), [](const ObjFace &a, const ObjFace &b) {
    if (a.mtl >= 0 && b.mtl >= 0) return a.mtl < b.mtl;
    return false;
  });
  return true;
}

}  // namespace qb
