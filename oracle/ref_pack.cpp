// ref_pack — parses an OBJ with the reference's own parser (include/tiny_obj_loader.h, used in
// place from the reference checkout) and packs it by the rules of src/main.cpp:1638-1700:
//   index list   = vertex_index of every face corner over all shapes      (main.cpp:1643-1649)
//   vertex buffer = per OBJ vertex i: [vertices[3i..], normals[3i..]]      (main.cpp:1673-1683)
// The reference reads normals[] out of bounds when a file has fewer normals than vertices;
// this tool refuses such files instead (exit 3).  Output: uint32 n_vertices, uint32
// n_triangles, floats[6*n_vertices], uint32[3*n_triangles].  Test infrastructure.
#define TINYOBJLOADER_IMPLEMENTATION
#include "tiny_obj_loader.h"

#include <cstdint>
#include <cstdio>
#include <vector>

int main(int argc, char** argv) {
  if (argc != 3) { fprintf(stderr, "usage: ref_pack in.obj out.bin\n"); return 2; }
  tinyobj::ObjReaderConfig cfg;
  tinyobj::ObjReader reader;
  if (!reader.ParseFromFile(argv[1], cfg)) { fprintf(stderr, "%s\n", reader.Error().c_str()); return 1; }
  const tinyobj::attrib_t& at = reader.GetAttrib();
  if (at.normals.size() < at.vertices.size()) return 3;
  std::vector<uint32_t> idx;
  uint32_t prims = 0;
  for (const tinyobj::shape_t& sh : reader.GetShapes()) {
    prims += (uint32_t)sh.mesh.num_face_vertices.size();
    for (const tinyobj::index_t& i : sh.mesh.indices) idx.push_back((uint32_t)i.vertex_index);
  }
  const size_t nfl = at.vertices.size();
  std::vector<float> vb(2 * nfl);
  for (size_t i = 0; i < nfl; i += 3)
    for (int k = 0; k < 3; k++) { vb[2 * i + k] = at.vertices[i + k]; vb[2 * i + 3 + k] = at.normals[i + k]; }
  FILE* f = fopen(argv[2], "wb");
  if (!f) return 1;
  const uint32_t hdr[2] = {(uint32_t)(nfl / 3), prims};
  fwrite(hdr, 4, 2, f);
  fwrite(vb.data(), 4, vb.size(), f);
  fwrite(idx.data(), 4, idx.size(), f);
  fclose(f);
  return idx.size() == 3 * (size_t)prims ? 0 : 4;
}
