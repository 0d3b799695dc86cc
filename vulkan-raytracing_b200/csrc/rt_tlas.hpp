// rt_tlas.hpp — top-level structure over the instances (host side, plain C++).
//
// Replaces createTLAS / the per-frame TLAS UPDATE of the reference (main.cpp:555-793,
// 2836-2861).  The reference rebuilds its two-instance TLAS on the GPU every frame behind a
// blocking fence; here the instance transforms are a small host array, so the top level is a
// binary tree over the instances' padded world boxes rebuilt on the host in microseconds
// (median split on the widest centroid axis) whenever rt_scene_set_transforms() is called, and
// uploaded with the instance array.  Node layout = the binary BLAS node (4 x float4: two child
// boxes + two child references); a negative reference ~i names instance i.  Scenes with at
// most RT_TLAS_MIN_INSTANCES instances skip it: the traversal loops over them directly.
#pragma once
#include <algorithm>
#include <cstring>
#include <vector>

#include "rt_trav.cuh"

namespace rt {

static const uint32_t RT_TLAS_MIN_INSTANCES = 4;

struct TlasBuilder {
  const std::vector<InstDev>& inst;
  std::vector<float4>& nodes;
  std::vector<uint32_t> ids;

  TlasBuilder(const std::vector<InstDev>& i, std::vector<float4>& n) : inst(i), nodes(n) {}

  void box_of(uint32_t begin, uint32_t end, float* lo, float* hi) const {
    for (int k = 0; k < 3; k++) { lo[k] = 3.0e38f; hi[k] = -3.0e38f; }
    for (uint32_t i = begin; i < end; i++)
      for (int k = 0; k < 3; k++) {
        lo[k] = std::min(lo[k], inst[ids[i]].wlo[k]);
        hi[k] = std::max(hi[k], inst[ids[i]].whi[k]);
      }
  }

  // returns the child reference of the subtree over ids[begin, end)
  int build(uint32_t begin, uint32_t end) {
    if (end - begin == 1) return ~(int)ids[begin];
    float clo[3] = {3.0e38f, 3.0e38f, 3.0e38f}, chi[3] = {-3.0e38f, -3.0e38f, -3.0e38f};
    for (uint32_t i = begin; i < end; i++)
      for (int k = 0; k < 3; k++) {
        const float c = 0.5f * (inst[ids[i]].wlo[k] + inst[ids[i]].whi[k]);
        clo[k] = std::min(clo[k], c);
        chi[k] = std::max(chi[k], c);
      }
    int axis = 0;
    if (chi[1] - clo[1] > chi[axis] - clo[axis]) axis = 1;
    if (chi[2] - clo[2] > chi[axis] - clo[axis]) axis = 2;
    const uint32_t mid = begin + (end - begin) / 2;
    std::nth_element(ids.begin() + begin, ids.begin() + mid, ids.begin() + end, [&](uint32_t a, uint32_t b) {
      const float ca = inst[a].wlo[axis] + inst[a].whi[axis], cb = inst[b].wlo[axis] + inst[b].whi[axis];
      return ca < cb || (ca == cb && a < b);
    });
    const size_t me = nodes.size() / 4;
    nodes.resize(nodes.size() + 4);
    const int c0 = build(begin, mid), c1 = build(mid, end);
    float l0[3], h0[3], l1[3], h1[3];
    box_of(begin, mid, l0, h0);
    box_of(mid, end, l1, h1);
    uint32_t u0, u1;
    memcpy(&u0, &c0, 4);
    memcpy(&u1, &c1, 4);
    float f0, f1;
    memcpy(&f0, &u0, 4);
    memcpy(&f1, &u1, 4);
    float4* o = &nodes[4 * me];
    o[0] = make_float4(l0[0], l0[1], l0[2], h0[0]);
    o[1] = make_float4(h0[1], h0[2], l1[0], l1[1]);
    o[2] = make_float4(l1[2], h1[0], h1[1], h1[2]);
    o[3] = make_float4(f0, f1, 0.0f, 0.0f);
    return (int)me;
  }
};

// nodes is left empty when the scene is small enough for the direct loop
inline void build_tlas(const std::vector<InstDev>& inst, std::vector<float4>& nodes) {
  nodes.clear();
  if (inst.size() <= RT_TLAS_MIN_INSTANCES) return;
  TlasBuilder b(inst, nodes);
  b.ids.resize(inst.size());
  for (uint32_t i = 0; i < inst.size(); i++) b.ids[i] = i;
  b.build(0, (uint32_t)inst.size());
}

}  // namespace rt
