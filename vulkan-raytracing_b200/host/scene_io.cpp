// scene_io.cpp — host-side asset loading and synthetic scene generation (librt_host.so).
// Replaces the CPU prologue of the reference's main(): OBJ parse + buffer packing
// (src/main.cpp:1603-1729) and skybox decode (src/main.cpp:2061-2113).  No ray tracing.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../include/rt_host.h"
#include "camera.hpp"
#include "vec3.hpp"

#ifdef RTH_HAVE_STB
// The reference's own JPEG decoder (include/stb_image.h, v2.27), used in place from the
// reference checkout through the include path; never copied into this repository.
#define STB_IMAGE_IMPLEMENTATION
#define STBI_ONLY_JPEG
#define STBI_NO_STDIO_FAILURE_STRINGS
#pragma GCC diagnostic push
#pragma GCC diagnostic ignored "-Wunused-function"
#pragma GCC diagnostic ignored "-Wmissing-field-initializers"
#include "stb_image.h"
#pragma GCC diagnostic pop
#endif

namespace {

thread_local std::string g_err;
int fail(int code, const std::string& msg) {
  g_err = msg;
  return code;
}

template <class T>
T* dup_vec(const std::vector<T>& v) {
  T* p = (T*)malloc(std::max<size_t>(1, v.size()) * sizeof(T));
  if (p && !v.empty()) memcpy(p, v.data(), v.size() * sizeof(T));
  return p;
}

struct Corner {
  int v, n;
};

// ---- OBJ --------------------------------------------------------------------------
// Only what the path consumes: v, vn and f records (v, v/vt, v//vn, v/vt/vn, negative
// indices).  Polygons are fan-triangulated; every face in the reference's assets is
// already a triangle.
bool parse_obj(const char* path, std::vector<float>& pos, std::vector<float>& nrm, std::vector<Corner>& corners) {
  FILE* f = fopen(path, "rb");
  if (!f) return false;
  fseek(f, 0, SEEK_END);
  long sz = ftell(f);
  fseek(f, 0, SEEK_SET);
  std::vector<char> buf((size_t)sz + 2);
  size_t rd = fread(buf.data(), 1, (size_t)sz, f);
  fclose(f);
  buf[rd] = '\n';
  buf[rd + 1] = 0;
  char* p = buf.data();
  char* end = p + rd + 1;
  std::vector<Corner> poly;
  while (p < end) {
    char* eol = (char*)memchr(p, '\n', (size_t)(end - p));
    if (!eol) eol = end - 1;
    *eol = 0;
    while (*p == ' ' || *p == '\t') p++;
    if (p[0] == 'v' && (p[1] == ' ' || p[1] == '\t')) {
      char* q = p + 1;
      for (int k = 0; k < 3; k++) pos.push_back(strtof(q, &q));
    } else if (p[0] == 'v' && p[1] == 'n' && (p[2] == ' ' || p[2] == '\t')) {
      char* q = p + 2;
      for (int k = 0; k < 3; k++) nrm.push_back(strtof(q, &q));
    } else if (p[0] == 'f' && (p[1] == ' ' || p[1] == '\t')) {
      poly.clear();
      char* q = p + 1;
      int nv = (int)(pos.size() / 3), nn = (int)(nrm.size() / 3);
      while (true) {
        while (*q == ' ' || *q == '\t' || *q == '\r') q++;
        if (!*q) break;
        Corner c{-1, -1};
        long vi = strtol(q, &q, 10);
        c.v = vi > 0 ? (int)vi - 1 : nv + (int)vi;
        if (*q == '/') {
          q++;
          if (*q != '/') strtol(q, &q, 10);  // vt, unused
          if (*q == '/') {
            q++;
            long ni = strtol(q, &q, 10);
            c.n = ni > 0 ? (int)ni - 1 : nn + (int)ni;
          }
        }
        poly.push_back(c);
      }
      for (size_t k = 2; k < poly.size(); k++) {
        corners.push_back(poly[0]);
        corners.push_back(poly[k - 1]);
        corners.push_back(poly[k]);
      }
    }
    p = eol + 1;
  }
  return true;
}

void area_weighted_normals(const std::vector<float>& pos, const std::vector<uint32_t>& idx, std::vector<float>& nrm) {
  size_t nv = pos.size() / 3;
  std::vector<double> acc(3 * nv, 0.0);
  for (size_t t = 0; t + 2 < idx.size(); t += 3) {
    const float* a = &pos[3 * (size_t)idx[t]];
    const float* b = &pos[3 * (size_t)idx[t + 1]];
    const float* c = &pos[3 * (size_t)idx[t + 2]];
    double e1[3] = {(double)b[0] - a[0], (double)b[1] - a[1], (double)b[2] - a[2]};
    double e2[3] = {(double)c[0] - a[0], (double)c[1] - a[1], (double)c[2] - a[2]};
    double n[3] = {e1[1] * e2[2] - e1[2] * e2[1], e1[2] * e2[0] - e1[0] * e2[2], e1[0] * e2[1] - e1[1] * e2[0]};
    for (int k = 0; k < 3; k++)
      for (int j = 0; j < 3; j++) acc[3 * (size_t)idx[t + k] + j] += n[j];
  }
  nrm.resize(3 * nv);
  for (size_t v = 0; v < nv; v++) {
    double l = std::sqrt(acc[3 * v] * acc[3 * v] + acc[3 * v + 1] * acc[3 * v + 1] + acc[3 * v + 2] * acc[3 * v + 2]);
    if (l == 0.0) { nrm[3 * v] = 0.f; nrm[3 * v + 1] = 1.f; nrm[3 * v + 2] = 0.f; continue; }
    for (int j = 0; j < 3; j++) nrm[3 * v + j] = (float)(acc[3 * v + j] / l);
  }
}

// ---- value noise --------------------------------------------------------------------
inline uint32_t hash3(int x, int y, int z, uint32_t seed) {
  uint32_t h = seed;
  h ^= (uint32_t)x * 0x9E3779B1u; h = (h << 13) | (h >> 19); h *= 0x85EBCA6Bu;
  h ^= (uint32_t)y * 0xC2B2AE35u; h = (h << 15) | (h >> 17); h *= 0x27D4EB2Fu;
  h ^= (uint32_t)z * 0x165667B1u; h = (h << 11) | (h >> 21); h *= 0x9E3779B1u;
  h ^= h >> 15; h *= 0x2C1B3C6Du; h ^= h >> 12; h *= 0x297A2D39u; h ^= h >> 15;
  return h;
}
inline double lattice(int x, int y, int z, uint32_t seed) {  // [-1,1]
  return (double)hash3(x, y, z, seed) * (2.0 / 4294967295.0) - 1.0;
}
inline double smooth(double t) { return t * t * (3.0 - 2.0 * t); }
double value_noise(double x, double y, double z, uint32_t seed) {
  double fx = std::floor(x), fy = std::floor(y), fz = std::floor(z);
  int ix = (int)fx, iy = (int)fy, iz = (int)fz;
  double tx = smooth(x - fx), ty = smooth(y - fy), tz = smooth(z - fz);
  double c[2][2][2];
  for (int a = 0; a < 2; a++)
    for (int b = 0; b < 2; b++)
      for (int d = 0; d < 2; d++) c[a][b][d] = lattice(ix + a, iy + b, iz + d, seed);
  auto lerp = [](double a, double b, double t) { return a + (b - a) * t; };
  double x00 = lerp(c[0][0][0], c[1][0][0], tx), x10 = lerp(c[0][1][0], c[1][1][0], tx);
  double x01 = lerp(c[0][0][1], c[1][0][1], tx), x11 = lerp(c[0][1][1], c[1][1][1], tx);
  return lerp(lerp(x00, x10, ty), lerp(x01, x11, ty), tz);
}
double fbm4(double x, double y, double z, uint32_t seed) {  // 4 octaves, in about [-1,1]
  double s = 0.0, amp = 1.0, f = 2.0, norm = 0.0;
  for (int o = 0; o < 4; o++) {
    s += amp * value_noise(x * f, y * f, z * f, seed + 101u * (uint32_t)o);
    norm += amp;
    amp *= 0.5;
    f *= 2.0;
  }
  return s / norm;
}

}  // namespace

extern "C" {

const char* rth_last_error(void) { return g_err.c_str(); }
void rth_free(void* p) { free(p); }

int rth_load_obj(const char* path, float** vertices, uint32_t* n_vertices, uint32_t** indices,
                 uint32_t* n_triangles, int* policy) {
  if (!path || !vertices || !n_vertices || !indices || !n_triangles) return fail(-1, "null argument");
  std::vector<float> pos, nrm;
  std::vector<Corner> corners;
  if (!parse_obj(path, pos, nrm, corners)) return fail(-1, std::string("cannot open ") + path);
  size_t nv = pos.size() / 3, nn = nrm.size() / 3;
  for (const Corner& c : corners)
    if (c.v < 0 || (size_t)c.v >= nv || (c.n >= 0 && (size_t)c.n >= nn))
      return fail(-1, std::string("index out of range in ") + path);
  std::vector<float> vb;
  std::vector<uint32_t> ib;
  int pol;
  bool has_n = nn > 0 && std::all_of(corners.begin(), corners.end(), [](const Corner& c) { return c.n >= 0; });
  bool one_to_one = has_n && nn == nv && std::all_of(corners.begin(), corners.end(), [](const Corner& c) { return c.n == c.v; });
  if (one_to_one) {
    // exactly the reference: vertex i carries normals[i]; index = vertex_index
    pol = RTH_PACK_REFERENCE;
    vb.resize(6 * nv);
    for (size_t i = 0; i < nv; i++) {
      memcpy(&vb[6 * i], &pos[3 * i], 12);
      memcpy(&vb[6 * i + 3], &nrm[3 * i], 12);
    }
    ib.reserve(corners.size());
    for (const Corner& c : corners) ib.push_back((uint32_t)c.v);
  } else if (has_n) {
    pol = RTH_PACK_DEINDEXED;
    std::map<std::pair<int, int>, uint32_t> seen;
    for (const Corner& c : corners) {
      auto key = std::make_pair(c.v, c.n);
      auto it = seen.find(key);
      if (it == seen.end()) {
        uint32_t id = (uint32_t)(vb.size() / 6);
        seen.emplace(key, id);
        for (int k = 0; k < 3; k++) vb.push_back(pos[3 * (size_t)c.v + k]);
        for (int k = 0; k < 3; k++) vb.push_back(nrm[3 * (size_t)c.n + k]);
        ib.push_back(id);
      } else {
        ib.push_back(it->second);
      }
    }
  } else {
    pol = RTH_PACK_GENERATED;
    for (const Corner& c : corners) ib.push_back((uint32_t)c.v);
    std::vector<float> gen;
    area_weighted_normals(pos, ib, gen);
    vb.resize(6 * nv);
    for (size_t i = 0; i < nv; i++) {
      memcpy(&vb[6 * i], &pos[3 * i], 12);
      memcpy(&vb[6 * i + 3], &gen[3 * i], 12);
    }
  }
  *vertices = dup_vec(vb);
  *indices = dup_vec(ib);
  *n_vertices = (uint32_t)(vb.size() / 6);
  *n_triangles = (uint32_t)(ib.size() / 3);
  if (policy) *policy = pol;
  return 0;
}

int rth_generate_normals(float* vertices, uint32_t nv, const uint32_t* indices, uint32_t nt) {
  if (!vertices || (nt && !indices)) { g_err = "rth_generate_normals: null mesh"; return 1; }
  for (uint64_t i = 0; i < 3ull * nt; i++)
    if (indices[i] >= nv) { g_err = "rth_generate_normals: vertex index out of range"; return 1; }
  std::vector<float> pos(3 * (size_t)nv), nrm;
  for (size_t v = 0; v < nv; v++) std::memcpy(&pos[3 * v], &vertices[6 * v], 3 * sizeof(float));
  std::vector<uint32_t> idx(indices, indices + 3 * (size_t)nt);
  area_weighted_normals(pos, idx, nrm);
  for (size_t v = 0; v < nv; v++) std::memcpy(&vertices[6 * v + 3], &nrm[3 * v], 3 * sizeof(float));
  return 0;
}

int rth_load_mtl(const char* path, rth_material* out, uint32_t max_materials, uint32_t* n_found) {
  if (!path || !n_found || (max_materials && !out)) { g_err = "rth_load_mtl: null argument"; return 1; }
  FILE* f = std::fopen(path, "rb");
  if (!f) { g_err = std::string("cannot open ") + path; return 1; }
  uint32_t n = 0;
  rth_material cur;
  bool open_block = false;
  auto flush = [&]() {
    if (!open_block) return;
    if (n < max_materials) out[n] = cur;
    n++;
  };
  char line[1024];
  while (std::fgets(line, sizeof line, f)) {
    char key[32] = "", name[64] = "";
    float v[3] = {0.f, 0.f, 0.f};
    if (std::sscanf(line, " %31s", key) != 1 || key[0] == '#') continue;
    if (!std::strcmp(key, "newmtl")) {
      flush();
      std::memset(&cur, 0, sizeof cur);
      const float ka[3] = {0.1f, 0.3f, 0.1f}, kd[3] = {0.2f, 1.0f, 0.2f}, ks[3] = {0.8f, 0.8f, 0.8f};  // shader.rgen:52-55
      std::memcpy(cur.ka, ka, sizeof ka); std::memcpy(cur.kd, kd, sizeof kd); std::memcpy(cur.ks, ks, sizeof ks);
      cur.ns = 100.0f; cur.ni = 1.52f;                                                                 // :125, :51
      if (std::sscanf(line, " %*s %63s", name) == 1) std::snprintf(cur.name, sizeof cur.name, "%s", name);
      open_block = true;
    } else if (open_block) {
      const int got = std::sscanf(line, " %*s %f %f %f", &v[0], &v[1], &v[2]);
      if (got == 3 && !std::strcmp(key, "Ka")) std::memcpy(cur.ka, v, sizeof v);
      else if (got == 3 && !std::strcmp(key, "Kd")) std::memcpy(cur.kd, v, sizeof v);
      else if (got == 3 && !std::strcmp(key, "Ks")) std::memcpy(cur.ks, v, sizeof v);
      else if (got >= 1 && !std::strcmp(key, "Ns") && v[0] > 0.0f) cur.ns = v[0];
      else if (got >= 1 && !std::strcmp(key, "Ni") && v[0] > 0.0f) cur.ni = v[0];
    }
  }
  flush();
  std::fclose(f);
  *n_found = n;
  return 0;
}

int rth_make_standin_armadillo(uint32_t level, uint32_t seed, float radius, float** vertices,
                               uint32_t* n_vertices, uint32_t** indices, uint32_t* n_triangles) {
  if (level > 10) return fail(-1, "level too large");
  // icosahedron
  const double t = (1.0 + std::sqrt(5.0)) / 2.0;
  std::vector<double> P = {-1, t, 0, 1, t, 0, -1, -t, 0, 1, -t, 0, 0, -1, t, 0, 1, t,
                           0, -1, -t, 0, 1, -t, t, 0, -1, t, 0, 1, -t, 0, -1, -t, 0, 1};
  std::vector<uint32_t> F = {0, 11, 5, 0, 5, 1, 0, 1, 7, 0, 7, 10, 0, 10, 11, 1, 5, 9, 5, 11, 4, 11, 10, 2, 10, 7, 6, 7, 1, 8,
                             3, 9, 4, 3, 4, 2, 3, 2, 6, 3, 6, 8, 3, 8, 9, 4, 9, 5, 2, 4, 11, 6, 2, 10, 8, 6, 7, 9, 8, 1};
  auto norm3 = [&](size_t i) {
    double l = std::sqrt(P[3 * i] * P[3 * i] + P[3 * i + 1] * P[3 * i + 1] + P[3 * i + 2] * P[3 * i + 2]);
    P[3 * i] /= l; P[3 * i + 1] /= l; P[3 * i + 2] /= l;
  };
  for (size_t i = 0; i < 12; i++) norm3(i);
  for (uint32_t l = 0; l < level; l++) {
    std::unordered_map<uint64_t, uint32_t> mid;
    mid.reserve(F.size() * 2);
    auto midpoint = [&](uint32_t a, uint32_t b) {
      uint64_t key = a < b ? ((uint64_t)a << 32) | b : ((uint64_t)b << 32) | a;
      auto it = mid.find(key);
      if (it != mid.end()) return it->second;
      uint32_t id = (uint32_t)(P.size() / 3);
      for (int k = 0; k < 3; k++) P.push_back(0.5 * (P[3 * (size_t)a + k] + P[3 * (size_t)b + k]));
      norm3(id);
      mid.emplace(key, id);
      return id;
    };
    std::vector<uint32_t> G;
    G.reserve(F.size() * 4);
    for (size_t f = 0; f < F.size(); f += 3) {
      uint32_t a = F[f], b = F[f + 1], c = F[f + 2];
      uint32_t ab = midpoint(a, b), bc = midpoint(b, c), ca = midpoint(c, a);
      uint32_t tri[12] = {a, ab, ca, b, bc, ab, c, ca, bc, ab, bc, ca};
      G.insert(G.end(), tri, tri + 12);
    }
    F.swap(G);
  }
  size_t nv = P.size() / 3;
  std::vector<float> pos(3 * nv);
  for (size_t i = 0; i < nv; i++) {
    double d = 1.0 + 0.15 * fbm4(P[3 * i], P[3 * i + 1], P[3 * i + 2], seed);
    for (int k = 0; k < 3; k++) pos[3 * i + k] = (float)(radius * d * P[3 * i + k]);
  }
  std::vector<float> nrm;
  area_weighted_normals(pos, F, nrm);
  std::vector<float> vb(6 * nv);
  for (size_t i = 0; i < nv; i++) {
    memcpy(&vb[6 * i], &pos[3 * i], 12);
    memcpy(&vb[6 * i + 3], &nrm[3 * i], 12);
  }
  *vertices = dup_vec(vb);
  *indices = dup_vec(F);
  *n_vertices = (uint32_t)nv;
  *n_triangles = (uint32_t)(F.size() / 3);
  return 0;
}

int rth_make_grid(const float* vertices, uint32_t n_vertices, const uint32_t* indices, uint32_t n_triangles,
                  uint32_t nx, uint32_t nz, float pitch, float** out_vertices, uint32_t* out_n_vertices,
                  uint32_t** out_indices, uint32_t* out_n_triangles) {
  uint64_t copies = (uint64_t)nx * nz;
  if (copies == 0 || copies * n_vertices > 0xFFFFFFFFull || copies * n_triangles > 0xFFFFFFFFull / 3)
    return fail(-1, "grid too large for 32-bit indices");
  float* vb = (float*)malloc((size_t)copies * n_vertices * 24);
  uint32_t* ib = (uint32_t*)malloc((size_t)copies * n_triangles * 12);
  if (!vb || !ib) { free(vb); free(ib); return fail(-4, "out of memory"); }
  for (uint32_t iz = 0; iz < nz; iz++)
    for (uint32_t ix = 0; ix < nx; ix++) {
      size_t c = (size_t)iz * nx + ix;
      float dx = ((float)ix - 0.5f * (float)(nx - 1)) * pitch, dz = ((float)iz - 0.5f * (float)(nz - 1)) * pitch;
      float* v = vb + c * n_vertices * 6;
      for (uint32_t i = 0; i < n_vertices; i++) {
        v[6 * (size_t)i + 0] = vertices[6 * (size_t)i + 0] + dx;
        v[6 * (size_t)i + 1] = vertices[6 * (size_t)i + 1];
        v[6 * (size_t)i + 2] = vertices[6 * (size_t)i + 2] + dz;
        memcpy(&v[6 * (size_t)i + 3], &vertices[6 * (size_t)i + 3], 12);
      }
      uint32_t* id = ib + c * n_triangles * 3;
      uint32_t base = (uint32_t)(c * n_vertices);
      for (size_t i = 0; i < (size_t)n_triangles * 3; i++) id[i] = indices[i] + base;
    }
  *out_vertices = vb;
  *out_indices = ib;
  *out_n_vertices = (uint32_t)(copies * n_vertices);
  *out_n_triangles = (uint32_t)(copies * n_triangles);
  return 0;
}

int rth_has_jpeg_decoder(void) {
#ifdef RTH_HAVE_STB
  return 1;
#else
  return 0;
#endif
}

int rth_load_skybox_dir(const char* dir, uint8_t** rgba8, uint32_t* w, uint32_t* h) {
#ifdef RTH_HAVE_STB
  static const char* names[6] = {"right", "left", "top", "bottom", "front", "back"};  // main.cpp:2064-2071
  uint8_t* all = nullptr;
  int W = 0, H = 0;
  for (int i = 0; i < 6; i++) {
    std::string p = std::string(dir) + "/" + names[i] + ".jpg";
    int fw, fh, ch;
    unsigned char* d = stbi_load(p.c_str(), &fw, &fh, &ch, 4);  // STBI_rgb_alpha, main.cpp:2078
    if (!d) { free(all); return fail(-1, "cannot decode " + p); }
    if (i == 0) {
      W = fw; H = fh;
      all = (uint8_t*)malloc((size_t)6 * W * H * 4);
      if (!all) { stbi_image_free(d); return fail(-4, "out of memory"); }
    } else if (fw != W || fh != H) {  // the reference silently assumes this (main.cpp:2074-2080)
      stbi_image_free(d); free(all);
      return fail(-1, "skybox faces differ in size: " + p);
    }
    memcpy(all + (size_t)i * W * H * 4, d, (size_t)W * H * 4);
    stbi_image_free(d);
  }
  *rgba8 = all; *w = (uint32_t)W; *h = (uint32_t)H;
  return 0;
#else
  (void)dir; (void)rgba8; (void)w; (void)h;
  return fail(-5, "built without a JPEG decoder");
#endif
}

int rth_make_procedural_skybox(uint32_t size, uint8_t* out) {
  if (!size || !out) return fail(-1, "bad argument");
  const double sun[3] = {0.45, 0.55, 0.70};
  double sl = std::sqrt(sun[0] * sun[0] + sun[1] * sun[1] + sun[2] * sun[2]);
#pragma omp parallel for schedule(static)
  for (int64_t row = 0; row < (int64_t)6 * size; row++) {
    int face = (int)(row / size);
    uint32_t j = (uint32_t)(row % size);
    for (uint32_t i = 0; i < size; i++) {
      double sc = 2.0 * ((i + 0.5) / size) - 1.0, tc = 2.0 * ((j + 0.5) / size) - 1.0;
      double r[3];
      switch (face) {  // inverse of the Vulkan cube face table
        case 0: r[0] = 1; r[1] = -tc; r[2] = -sc; break;
        case 1: r[0] = -1; r[1] = -tc; r[2] = sc; break;
        case 2: r[0] = sc; r[1] = 1; r[2] = tc; break;
        case 3: r[0] = sc; r[1] = -1; r[2] = -tc; break;
        case 4: r[0] = sc; r[1] = -tc; r[2] = 1; break;
        default: r[0] = -sc; r[1] = -tc; r[2] = -1; break;
      }
      double l = std::sqrt(r[0] * r[0] + r[1] * r[1] + r[2] * r[2]);
      for (double& x : r) x /= l;
      double col[3];
      if (r[1] >= 0.0) {
        double h = std::pow(1.0 - r[1], 3.0);
        double base[3] = {0.25 + 0.55 * h, 0.45 + 0.45 * h, 0.85 + 0.12 * h};
        double cl = fbm4(r[0] / (r[1] + 0.25) * 1.5, 3.7, r[2] / (r[1] + 0.25) * 1.5, 7u);
        double cov = std::min(1.0, std::max(0.0, (cl - 0.05) * 3.0)) * std::min(1.0, r[1] * 6.0);
        double sd = (r[0] * sun[0] + r[1] * sun[1] + r[2] * sun[2]) / sl;
        double glow = std::pow(std::max(0.0, sd), 64.0) * 0.6 + (sd > 0.9985 ? 1.0 : 0.0);
        for (int k = 0; k < 3; k++) col[k] = base[k] * (1.0 - cov) + 0.95 * cov + glow;
      } else {
        double px = r[0] / -r[1], pz = r[2] / -r[1];
        int cx = (int)std::floor(px * 2.0), cz = (int)std::floor(pz * 2.0);
        double chk = ((cx + cz) & 1) ? 0.65 : 0.25;
        double fade = std::exp(-0.15 * std::sqrt(px * px + pz * pz));
        double g[3] = {0.55 * chk + 0.1, 0.5 * chk + 0.12, 0.4 * chk + 0.1};
        double hz[3] = {0.80, 0.88, 0.95};
        for (int k = 0; k < 3; k++) col[k] = g[k] * fade + hz[k] * (1.0 - fade);
      }
      uint8_t* px8 = out + (((size_t)face * size + j) * size + i) * 4;
      for (int k = 0; k < 3; k++) px8[k] = (uint8_t)std::lrint(255.0 * std::min(1.0, std::max(0.0, col[k])));
      px8[3] = 255;
    }
  }
  return 0;
}

void* rth_camera_create(float px, float py, float pz) { return new vkrt::Camera(vkrt::vec3(px, py, pz)); }
void rth_camera_destroy(void* c) { delete (vkrt::Camera*)c; }
void rth_camera_move(void* c, int dir, float d) { ((vkrt::Camera*)c)->move((vkrt::CameraMovementDirection)dir, d); }
void rth_camera_mouse(void* c, float x, float y) { ((vkrt::Camera*)c)->processMouseMovement(x, y); }
void rth_camera_look(void* c, int dir) { ((vkrt::Camera*)c)->look((vkrt::CameraMovementDirection)dir); }
void rth_camera_get(void* c, float o[12]) {
  vkrt::Camera* cam = (vkrt::Camera*)c;
  vkrt::vec3 v[4] = {cam->getPosition(), cam->getRightVector(), cam->getUpVector(), cam->getFrontVector()};
  for (int i = 0; i < 4; i++) { o[3 * i] = v[i].x; o[3 * i + 1] = v[i].y; o[3 * i + 2] = v[i].z; }
}
void rth_camera_view_matrix(void* c, int without_translation, float o[16]) {
  vkrt::Camera* cam = (vkrt::Camera*)c;
  vkrt::mat4 m = without_translation ? cam->getViewingMatrixWithoutTranslation() : cam->getViewingMatrix();
  memcpy(o, m.m, sizeof(float) * 16);
}

void rth_orbit_transforms(float t, float out24[24]) {
  using namespace vkrt;
  toRowMajor3x4(mat4::identity(), out24);
  mat4 m = translate(rotate(translate(mat4::identity(), vec3(0, 0, -5)), t * 3.14159265358979323846f, vec3(0, 1, 0)),
                     vec3(0, 0, 10));
  toRowMajor3x4(m, out24 + 12);
}

int rth_write_ppm(const char* path, uint32_t w, uint32_t h, const uint8_t* rgba8) {
  FILE* f = fopen(path, "wb");
  if (!f) return fail(-1, std::string("cannot write ") + path);
  fprintf(f, "P6\n%u %u\n255\n", w, h);
  std::vector<uint8_t> row(3 * (size_t)w);
  for (uint32_t y = 0; y < h; y++) {
    for (uint32_t x = 0; x < w; x++) memcpy(&row[3 * (size_t)x], &rgba8[((size_t)y * w + x) * 4], 3);
    fwrite(row.data(), 1, row.size(), f);
  }
  fclose(f);
  return 0;
}

}  // extern "C"
