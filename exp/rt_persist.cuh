// rt_persist.cuh — warp-persistent traversal of the compressed 8-wide BVH (device only).
//
// The run-to-completion loop of rt_trav.cuh leaves most lanes idle once rays diverge
// (ncu r01: 17.9 of 32 threads in the node test, 5.3 in the triangle test).  Here every
// lane is a small state machine and the warp iterates it in lock step:
//   * a lane whose ray has finished is refilled from the ray queue (one warp-aggregated
//     atomic for all idle lanes) instead of waiting for the slowest ray of a batch of 32;
//   * triangle groups are postponed (pushed on the stack) while fewer than `tri_thresh`
//     lanes have triangles to test and the lane still has nodes to visit (Ylitie et al.
//     2017, §4.3), so the triangle test runs with more lanes switched on.
// Results do not depend on the visiting order: closest hit = smallest t with the
// (instance, primitive) tie rule; any-hit is a boolean.
#pragma once
#include "rt_trav.cuh"

namespace rt {

#ifndef RT_FETCH
#define RT_FETCH 128  // rays a warp takes from the queue per atomic
#endif

// Source: bool load(uint32_t idx, f3& o, f3& d, float& tmin, float& tmax)  (per lane)
// Sink:   void done(bool finished, uint32_t idx, bool hit, const HitRec& best)
//         called by ALL lanes of the warp whenever at least one lane finished a ray.
template <bool ANY, bool COUNT, class Stack, class Source, class Sink>
__device__ __forceinline__ void trace_wide_persistent(const SceneDev& scene, uint32_t n, uint32_t* work_counter,
                                                      Stack& st, Source& src, Sink& sink, TravCounters& tc,
                                                      int refill_thresh, int tri_thresh) {
  const uint32_t FULL = 0xFFFFFFFFu;
  const uint32_t lane = threadIdx.x & 31u;
  const uint32_t lt_mask = (1u << lane) - 1u;
  bool has_ray = false, finished = false, exhausted = false;
  uint32_t idx = 0, inst = 0;
  f3 wo = mk3(0, 0, 0), wd = mk3(0, 0, 0);
  float tmin = 0.0f;
  RayObj r;
  HitRec best;
  best.inst = RT_MISS; best.prim = RT_MISS; best.t = 0.0f; best.u = 0.0f; best.v = 0.0f;
  BlasDev blas = {};
  uint2 ngroup = make_uint2(0u, 0u), tgroup = make_uint2(0u, 0u);
  st.reset();

  uint32_t batch_next = 0, batch_end = 0;  // warp-private slice of the ray queue (warp-uniform)
  while (true) {
    // ---- idle lanes: flush finished rays to the sink and refill, once enough lanes wait -----
    // (both need atomics on shared counters, so they are batched: one queue fetch per
    // RT_FETCH rays, one sink call per >= refill_thresh finished rays)
    const uint32_t idle = __ballot_sync(FULL, !has_ray || finished);
    if (idle != 0u && (__popc(idle) >= refill_thresh || idle == FULL)) {
      if (__any_sync(FULL, finished)) {
        sink.done(finished, idx, best.inst != RT_MISS, best);
        if (finished) { has_ray = false; finished = false; }
      }
      if (batch_next == batch_end && !exhausted) {
        uint32_t base = 0;
        if (lane == 0) base = atomicAdd(work_counter, (uint32_t)RT_FETCH);
        base = __shfl_sync(FULL, base, 0);
        if (base >= n) { exhausted = true; }
        else { batch_next = base; batch_end = base + RT_FETCH < n ? base + RT_FETCH : n; }
      }
      const uint32_t avail = batch_end - batch_next;
      if (avail != 0u) {
        const uint32_t rank = (uint32_t)__popc(idle & lt_mask);
        if (!has_ray && rank < avail) {
          float tmax;
          idx = batch_next + rank;
          if (src.load(idx, wo, wd, tmin, tmax)) {
            has_ray = true;
            inst = 0xFFFFFFFFu;
            best.inst = RT_MISS; best.prim = RT_MISS; best.u = 0.0f; best.v = 0.0f;
            best.t = __uint_as_float(__float_as_uint(tmax) - 1u);  // "t <= best.t" == "t < tmax"
            ngroup = make_uint2(0u, 0u);
            tgroup = make_uint2(0u, 0u);
            st.reset();
          }
        }
        const uint32_t want = (uint32_t)__popc(idle);
        batch_next += want < avail ? want : avail;
      }
    }
    if (!__any_sync(FULL, has_ray)) {
      if (exhausted && batch_next == batch_end) break;
      continue;
    }
    // ---- one traversal step ---------------------------------------------------------------
    if (has_ray && !finished) {
      if (ngroup.y <= 0x00FFFFFFu && tgroup.y == 0u) {
        if (!st.empty()) {
          ngroup = st.pop();
        } else {  // this instance is done: move to the next one the ray can reach
          bool found = false;
          while (++inst < scene.n_inst) {
            const InstDev& in = scene.inst[inst];
            if (scene.n_inst > 1) {
              RayObj rw;
              rw.o = wo;
              rw.idir = mk3(safe_rcp_dir(wd.x), safe_rcp_dir(wd.y), safe_rcp_dir(wd.z));
              if (!slab_test(rw, in.wlo[0], in.wlo[1], in.wlo[2], in.whi[0], in.whi[1], in.whi[2], tmin, best.t)) continue;
            }
            blas = scene.blas[in.mesh];
            if (blas.ntri == 0u) continue;
            r = ray_setup(xform_point(in.w2o, wo), xform_vector(in.w2o, wd));
            ngroup = make_uint2(0u, 0x80000000u);
            found = true;
            break;
          }
          if (!found) finished = true;
        }
      }
      if (!finished) {
        if (ngroup.y > 0x00FFFFFFu) {
          const uint32_t hits = ngroup.y;
          const int child_bit = bfind32(hits);
          const uint32_t child_base = ngroup.x;
          ngroup.y &= ~(1u << child_bit);
          if (ngroup.y > 0x00FFFFFFu) st.push(ngroup);
          const uint32_t slot = (uint32_t)(child_bit - 24) ^ (r.octinv4 & 7u);
          const uint32_t rel = (uint32_t)__popc(hits & ~(0xFFFFFFFFu << slot) & 0xFFu);
          const size_t ni = 5 * (size_t)(child_base + rel);
          const uint4* nodes = (const uint4*)blas.nodes;
          const float4 n0 = __ldg(&blas.nodes[ni]);
          const uint4 n1 = __ldg(&nodes[ni + 1]);
          const uint4 n2 = __ldg(&nodes[ni + 2]);
          const uint4 n3 = __ldg(&nodes[ni + 3]);
          const uint4 n4 = __ldg(&nodes[ni + 4]);
          if (COUNT) tc.nodes++;
          const uint32_t hm = wide_node_test<COUNT>(n0, n1, n2, n3, n4, r, tmin, best.t);
          ngroup.x = n1.x;
          tgroup.x = n1.y;
          ngroup.y = (hm & 0xFF000000u) | (__float_as_uint(n0.w) >> 24);
          tgroup.y = hm & 0x00FFFFFFu;
        } else {
          tgroup = ngroup;
          ngroup = make_uint2(0u, 0u);
        }
      }
    }
    // ---- triangle phase: test now if enough lanes take part, otherwise postpone ------------
    const bool has_tri = has_ray && !finished && tgroup.y != 0u;
    const uint32_t tmask = __ballot_sync(FULL, has_tri);
    if (has_tri) {
      if (__popc(tmask) >= tri_thresh || ngroup.y <= 0x00FFFFFFu) {
        while (tgroup.y != 0u) {
          const int ti = bfind32(tgroup.y);
          tgroup.y &= ~(1u << ti);
          if (test_leaf_tri<ANY, COUNT>(blas, tgroup.x + (uint32_t)ti, r, tmin, inst, best, &tc)) {
            finished = true;  // any-hit: first intersection ends the ray
            tgroup.y = 0u;
          }
        }
      } else {
        st.push(tgroup);
        tgroup.y = 0u;
      }
    }
  }
}

}  // namespace rt
