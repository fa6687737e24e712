// objload.cpp — see objload.h. Compiled with -ffp-contract=off.
//
// An OBJ / MTL reader of this repository's own (cursor + keyword table), written to honour a FORMAT CONTRACT rather than a
// grammar: the reference reads its meshes through tinyobjloader v1.1.0, and vertex bits, face ids and face order decide
// hit distances, primitive ids and BVH topology. What the contract says is listed at each function; that the result is
// bit-identical to tinyobjloader's is checked by fuzzing against the reference's own copy (tests/test_objload_fuzz.py,
// oracle/ref_driver.cpp qaref_tinyobj_*) and, for the fixture meshes, by tests/test_binding.py.
#include "objload.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>

namespace qb {

namespace {

inline bool digit(char c) { return c >= '0' && c <= '9'; }
inline bool blank(char c) { return c == ' ' || c == '\t'; }

// A decimal literal split into its parts; `ok` false = tinyobjloader rejects the token and keeps the default value.
struct Decimal {
  bool ok = false, negative = false, exp_negative = false;
  const char *int_b = nullptr, *int_e = nullptr;    // integer digits (at least one)
  const char *frac_b = nullptr, *frac_e = nullptr;  // fraction digits (may be empty)
  const char *exp_b = nullptr, *exp_e = nullptr;    // exponent digits (empty = no exponent)
};

// Lexing rules (observed behaviour of tinyobjloader 1.1.0's number parser, which is NOT strtod):
//   * optional sign, then at least one digit: ".5" and "-.5" are rejected;
//   * an optional '.' with any number of digits, then an optional e/E exponent; an exponent may follow the integer part directly;
//   * anything else ends the literal where it stands ("1abc" is 1);
//   * 'e' must be followed by an optional sign and at least one digit, else the WHOLE token is rejected ("1e", "1e+" give the default).
Decimal lex_decimal(const char *b, const char *e)
{
  Decimal d;
  const char *p = b;
  if (p >= e) return d;
  if (*p == '+' || *p == '-') d.negative = *p++ == '-';
  else if (!digit(*p)) return d;
  d.int_b = p;
  while (p != e && digit(*p)) ++p;
  d.int_e = p;
  if (d.int_e == d.int_b) return d;
  d.frac_b = d.frac_e = d.exp_b = d.exp_e = p;
  bool exponent = false;
  if (p != e && *p == '.') {
    d.frac_b = ++p;
    while (p != e && digit(*p)) ++p;
    d.frac_e = p;
    exponent = p != e && (*p == 'e' || *p == 'E');
  } else if (p != e && (*p == 'e' || *p == 'E')) {
    exponent = true;
  }
  if (exponent) {
    ++p;
    if (p != e && (*p == '+' || *p == '-')) d.exp_negative = *p++ == '-';
    else if (p == e || !digit(*p)) return d;  // at the token's end the library looks at the separator, which is no digit
    d.exp_b = p;
    while (p != e && digit(*p)) ++p;
    d.exp_e = p;
    if (d.exp_e == d.exp_b) return d;
  }
  d.ok = true;
  return d;
}

// Arithmetic contract (every step in double, in this order — the sums are not correctly rounded, and the float it is cast
// to inherits that):
//   mantissa = ((d1 * 10 + d2) * 10 + ...) over the integer digits,
//            + digit_k * 10^-k for fraction digit k = 1, 2, ...: a literal table up to 10^-7, pow(10, -k) beyond;
//   value    = mantissa, or ldexp(mantissa * pow(5, e), e) when an exponent e != 0 is given; then the sign.
double eval_decimal(const Decimal &d)
{
  static const double tenth[8] = {1.0, 0.1, 0.01, 0.001, 0.0001, 0.00001, 0.000001, 0.0000001};
  double m = 0.0;
  for (const char *p = d.int_b; p != d.int_e; ++p) {
    m *= 10;
    m += (int) (*p - '0');
  }
  int k = 1;
  for (const char *p = d.frac_b; p != d.frac_e; ++p, ++k) m += (int) (*p - '0') * (k < 8 ? tenth[k] : std::pow(10.0, -k));
  int ex = 0;
  for (const char *p = d.exp_b; p != d.exp_e; ++p) {
    ex *= 10;
    ex += (int) (*p - '0');
  }
  if (d.exp_negative) ex = -ex;
  const double v = ex ? std::ldexp(m * std::pow(5.0, ex), ex) : m;
  return d.negative ? -v : v;
}

// One line of an OBJ / MTL file.
struct Cursor {
  const char *p;
  explicit Cursor(const std::string &line) : p(line.c_str()) {}
  void skip_blank() { p += strspn(p, " \t"); }
  bool at_end() const { return *p == '\0' || *p == '\r' || *p == '\n'; }
  // the keyword followed by a blank; consumes the keyword and that one blank
  bool keyword(const char *kw)
  {
    const size_t n = strlen(kw);
    if (strncmp(p, kw, n) != 0 || !blank(p[n])) return false;
    p += n + 1;
    return true;
  }
  // next real: the token runs to the next blank or '\r'; a rejected token gives `fallback`
  float real(double fallback = 0.0)
  {
    skip_blank();
    const char *e = p + strcspn(p, " \t\r");
    const Decimal d = lex_decimal(p, e);
    p = e;
    return (float) (d.ok ? eval_decimal(d) : fallback);
  }
  int integer()
  {
    skip_blank();
    return atoi(p);
  }
  std::string rest()
  {
    skip_blank();
    return std::string(p);
  }
};

// Index contract: 1-based; negative = counted back from the elements read SO FAR; 0 is an error.
bool resolve_index(int raw, int count, int *out)
{
  if (raw == 0) return false;
  *out = raw > 0 ? raw - 1 : count + raw;
  return true;
}

// A face corner: v, v/t, v//n or v/t/n. Absent parts stay -1. The cursor ends after the corner.
bool read_corner(Cursor &c, int nv, int nn, int nt, ObjIndex *out)
{
  ObjIndex k = {-1, -1, -1};
  auto field = [&](int count, int *dst) {
    const bool ok = resolve_index(atoi(c.p), count, dst);
    c.p += strcspn(c.p, "/ \t\r");
    return ok;
  };
  if (!field(nv, &k.v)) return false;
  if (*c.p == '/') {
    ++c.p;
    if (*c.p == '/') {  // v//n
      ++c.p;
      if (!field(nn, &k.n)) return false;
    } else {            // v/t[/n]
      if (!field(nt, &k.t)) return false;
      if (*c.p == '/') {
        ++c.p;
        if (!field(nn, &k.n)) return false;
      }
    }
  }
  *out = k;
  return true;
}

// The file as lines: '\n', '\r\n' and a lone '\r' all end a line.
bool slurp_lines(const std::string &path, std::vector<std::string> *lines)
{
  FILE *fp = fopen(path.c_str(), "rb");
  if (!fp) return false;
  std::string text;
  char chunk[1 << 16];
  for (size_t n; (n = fread(chunk, 1, sizeof chunk, fp)) > 0;) text.append(chunk, n);
  fclose(fp);
  for (size_t b = 0; b < text.size();) {
    const size_t e = text.find_first_of("\r\n", b);
    if (e == std::string::npos) {
      lines->emplace_back(text, b);
      break;
    }
    lines->emplace_back(text, b, e - b);
    b = e + ((text[e] == '\r' && e + 1 < text.size() && text[e + 1] == '\n') ? 2 : 1);
  }
  return true;
}

ObjMaterial blank_material()
{
  ObjMaterial m;
  for (int i = 0; i < 3; ++i) m.diffuse[i] = m.specular[i] = m.transmittance[i] = 0.f;
  m.illum = 0;
  m.shininess = 1.f;
  m.ior = 1.f;
  return m;
}

// MTL contract, for the fields the reference consumes (src/parser/xmlload.cpp:232-272): Kd, Ks, Kt | Tf, Ni, Ns, illum,
// map_Kd, map_Ks. `newmtl` closes the material before it IF that one has a name; the last material is always appended (so a
// file without `newmtl` yields one unnamed material). Trailing blanks of a line do not belong to names.
bool read_mtl(const std::string &path, std::vector<ObjMaterial> *materials, std::map<std::string, int> *by_name)
{
  std::vector<std::string> lines;
  if (!slurp_lines(path, &lines)) return false;
  ObjMaterial cur = blank_material();
  auto close = [&]() {
    by_name->insert(std::make_pair(cur.name, (int) materials->size()));
    materials->push_back(cur);
  };
  for (std::string &line : lines) {
    line.erase(line.find_last_not_of(" \t") + 1);
    Cursor c(line);
    c.skip_blank();
    if (c.at_end() || *c.p == '#') continue;
    if (c.keyword("newmtl")) {
      if (!cur.name.empty()) close();
      cur = blank_material();
      cur.name = c.p;
    } else if (c.keyword("Kd")) {
      for (float &x : cur.diffuse) x = c.real();
    } else if (c.keyword("Ks")) {
      for (float &x : cur.specular) x = c.real();
    } else if (c.keyword("Kt") || c.keyword("Tf")) {
      for (float &x : cur.transmittance) x = c.real();
    } else if (c.keyword("Ni")) {
      cur.ior = c.real();
    } else if (c.keyword("Ns")) {
      cur.shininess = c.real();
    } else if (c.keyword("illum")) {
      cur.illum = c.integer();
    } else if (c.keyword("map_Kd")) {
      cur.diffuse_texname = c.rest();
    } else if (c.keyword("map_Ks")) {
      cur.specular_texname = c.rest();
    }
  }
  close();
  return true;
}

}  // namespace

bool tinyobj_parse_double(const char *s, const char *s_end, double *result)
{
  const Decimal d = lex_decimal(s, s_end);
  if (!d.ok) return false;
  *result = eval_decimal(d);
  return true;
}

// OBJ contract: `v` 3 reals, `vn` 3 reals, `vt` 2 reals; `f` any number of corners, triangulated as the fan
// (c0, c[k-1], c[k]); a face takes the material of the last `usemtl` (-1 for an unknown name); `mtllib` tries its
// space-separated names in turn and stops at the first file that opens. Everything else is ignored.
bool load_obj(const char *filename, ObjMesh *mesh, std::string *err)
{
  // TriMesh's ComputePath (src/mesh/TriMesh.cpp:47-60): directory with trailing '/', base name
  std::string full = filename;
  std::replace(full.begin(), full.end(), '\\', '/');
  const size_t slash = full.find_last_of('/');
  mesh->dir = slash == std::string::npos ? std::string() : full.substr(0, slash + 1);
  mesh->name = slash == std::string::npos ? full : full.substr(slash + 1);
  mesh->file = full;
  mesh->vertices.clear();
  mesh->normals.clear();
  mesh->texcoords.clear();
  mesh->materials.clear();
  mesh->faces.clear();

  std::vector<std::string> lines;
  if (!slurp_lines(full, &lines)) {
    if (err) *err = "Cannot open file [" + full + "]";
    return false;
  }
  std::map<std::string, int> by_name;
  int active_material = -1;
  std::vector<ObjIndex> corners;
  for (const std::string &line : lines) {
    Cursor c(line);
    c.skip_blank();
    if (c.at_end() || *c.p == '#') continue;
    if (c.keyword("v")) {
      for (int i = 0; i < 3; ++i) mesh->vertices.push_back(c.real());
    } else if (c.keyword("vn")) {
      for (int i = 0; i < 3; ++i) mesh->normals.push_back(c.real());
    } else if (c.keyword("vt")) {
      for (int i = 0; i < 2; ++i) mesh->texcoords.push_back(c.real());
    } else if (c.keyword("f")) {
      corners.clear();
      for (c.skip_blank(); !c.at_end(); c.p += strspn(c.p, " \t\r")) {
        ObjIndex k;
        if (!read_corner(c, (int) mesh->NV(), (int) mesh->NVN(), (int) (mesh->texcoords.size() / 2), &k)) {
          if (err) *err = "Failed parse `f' line(e.g. zero value for face index).";
          return false;
        }
        corners.push_back(k);
      }
      for (size_t k = 2; k < corners.size(); ++k) {
        ObjFace f;
        f.idx[0] = corners[0];
        f.idx[1] = corners[k - 1];
        f.idx[2] = corners[k];
        f.mtl = active_material;
        f.file_idx = (int) mesh->faces.size();
        mesh->faces.push_back(f);
      }
    } else if (c.keyword("usemtl")) {
      const auto it = by_name.find(std::string(c.p));
      active_material = it == by_name.end() ? -1 : it->second;
    } else if (c.keyword("mtllib")) {
      const std::string names(c.p);
      for (size_t b = 0; b <= names.size();) {
        size_t e = names.find(' ', b);
        if (e == std::string::npos) e = names.size();
        if (e > b && read_mtl(mesh->dir + names.substr(b, e - b), &mesh->materials, &by_name)) break;
        b = e + 1;
      }
    }
  }
  // the reference's face sort (src/mesh/TriMesh.cpp:107-114): the same libstdc++ std::sort with the same comparator — it
  // is neither stable nor, with negative ids, a strict weak order, so only the identical call gives the identical order
  std::sort(mesh->faces.begin(), mesh->faces.end(), [](const ObjFace &a, const ObjFace &b) {
    if (a.mtl >= 0 && b.mtl >= 0) return a.mtl < b.mtl;
    return false;
  });
  return true;
}

}  // namespace qb
