// rt_wave.cuh — warp-synchronous traversal of the compressed 8-wide BVH (device only).
//
// What traceRayEXT (shader.rgen:86-87 closest hit, :107-112 terminate on first hit) costs on a
// GPU without RT cores is dominated by SIMD efficiency: ncu on the run-to-completion loop shows
// 14.6 / 11.4 / 6.7 of 32 threads active per instruction in bounces 1 / 2 / 3+ (rays inside the
// refractive mesh).  Two losses: lanes whose ray finished wait for the slowest ray of the
// batch, and the triangle test runs with the few lanes that happen to hold a leaf hit.
//
// Here every lane is a small state machine and the whole warp advances in lock step:
//   1. refill  - when at least `refill_min` lanes are idle, finished rays are handed to the sink
//                (hit record + material-queue append) and idle lanes take new rays from a
//                warp-private chunk of the ray queue (one atomic per RT_WAVE_CHUNK rays);
//   2. node    - every lane that has an inner child pending pops it and tests one wide node;
//   3. leaf    - triangle groups are tested when at least `tri_min` lanes hold one (or when a
//                lane has nothing else to do), otherwise they are pushed and picked up later
//                (Ylitie, Karras, Laine 2017, section 4.3).
// Results do not depend on the visiting order: closest hit = smallest t with the
// (instance, primitive) tie rule of consider_hit(); any-hit is a boolean.
#pragma once
#include "rt_trav.cuh"

namespace rt {

#ifndef RT_WAVE_CHUNK
#define RT_WAVE_CHUNK 256
#endif

enum { LANE_IDLE = 0, LANE_RUN = 1, LANE_WAIT = 2 };

// Source: void load(uint32_t idx, f3& o, f3& d, float& tmax)          per lane
// Sink:   void done(bool finished, uint32_t idx, bool hit, const HitRec&)  all lanes of the warp
template <bool ANY, bool COUNT, class Stack, class Source, class Sink>
__device__ __forceinline__ void wave_trace(const SceneDev& scene, uint32_t n, uint32_t* work_counter, Stack& st,
                                           const Source& src, const Sink& sink, TravCounters& tc, int refill_min,
                                           int tri_min) {
  const uint32_t FULL = 0xFFFFFFFFu;
  const uint32_t lane = threadIdx.x & 31u;
  const uint32_t lt_mask = (1u << lane) - 1u;
  const float tmin = 0.001f;  // shader.rgen:87 and :112
  int state = LANE_IDLE;
  uint32_t idx = 0, inst = 0, cand = 0;
  RayObj r;
  r.o = mk3(0, 0, 0); r.d = mk3(0, 0, 1); r.idir = mk3(1, 1, 1);
  r.kx = 0; r.ky = 1; r.kz = 2; r.Sx = r.Sy = 0.0f; r.Sz = 1.0f; r.octinv4 = 0;
  HitRec best;
  best.inst = RT_MISS; best.prim = RT_MISS; best.t = 0.0f; best.u = 0.0f; best.v = 0.0f;
  const uint4* nodes = nullptr;
  const float4* tris = nullptr;
  uint2 ngroup = make_uint2(0u, 0u), tgroup = make_uint2(0u, 0u);
  st.reset();
  uint32_t chunk_next = 0, chunk_end = 0;  // warp-uniform
  bool exhausted = false;                  // warp-uniform
  // rays a warp takes per atomic: large enough to keep the shared counter cold, small enough
  // that every warp of the grid gets several chunks
  const uint32_t total_warps = gridDim.x * (blockDim.x >> 5);
  const uint32_t chunk = min((uint32_t)RT_WAVE_CHUNK, max(32u, (n / (total_warps * 8u)) & ~31u));

  while (true) {
    // ---- 1. refill: all the rare, expensive work (sink, fetch, instance entry with its ray
    // transform and divisions) happens here, for many lanes at once ---------------------------
    const uint32_t run_mask = __ballot_sync(FULL, state == LANE_RUN);
    const int idle = 32 - __popc(run_mask);
    if (idle >= refill_min || run_mask == 0u) {
      // a waiting lane with no instance left has finished its ray
      const bool fin = state == LANE_WAIT && cand == 0u;
      if (__any_sync(FULL, fin)) {
        sink.done(fin, idx, best.inst != RT_MISS, best);
        if (fin) state = LANE_IDLE;
      }
      const uint32_t idle_mask = __ballot_sync(FULL, state == LANE_IDLE);
      if (chunk_next == chunk_end && !exhausted && idle_mask != 0u) {
        uint32_t base = 0;
        if (lane == 0) base = atomicAdd(work_counter, chunk);
        base = __shfl_sync(FULL, base, 0);
        if (base >= n) exhausted = true;
        else { chunk_next = base; chunk_end = min(base + chunk, n); }
      }
      const uint32_t avail = chunk_end - chunk_next;
      f3 o = mk3(0, 0, 0), d = mk3(0, 0, 1);
      float tmax = 0.0f;
      if (state == LANE_IDLE) {
        const uint32_t rank = (uint32_t)__popc(idle_mask & lt_mask);
        if (rank < avail) {
          idx = chunk_next + rank;
          src.load(idx, o, d, tmax);
          best.inst = RT_MISS; best.prim = RT_MISS; best.u = 0.0f; best.v = 0.0f;
          best.t = __uint_as_float(__float_as_uint(tmax) - 1u);  // "t <= best.t" == "t < tmax"
          cand = scene.n_inst >= 32u ? 0xFFFFFFFFu : (1u << scene.n_inst) - 1u;
          st.reset();
          state = LANE_WAIT;
        }
      } else if (state == LANE_WAIT) {
        src.load(idx, o, d, tmax);
      }
      {
        const uint32_t want = (uint32_t)__popc(idle_mask);
        chunk_next += want < avail ? want : avail;
      }
      if (state == LANE_WAIT) {
        // enter the next instance whose padded world box the ray can still reach
        RayObj rw;
        rw.o = o;
        rw.idir = mk3(safe_rcp_dir(d.x), safe_rcp_dir(d.y), safe_rcp_dir(d.z));
        while (cand != 0u) {
          inst = (uint32_t)__ffs((int)cand) - 1u;
          cand &= cand - 1u;
          const InstDev& in = scene.inst[inst];
          if (scene.n_inst > 1u &&
              !slab_test(rw, in.wlo[0], in.wlo[1], in.wlo[2], in.whi[0], in.whi[1], in.whi[2], tmin, best.t))
            continue;
          const BlasDev& b = scene.blas[in.mesh];
          if (b.ntri == 0u) continue;
          nodes = (const uint4*)b.nodes;
          tris = b.tris;
          r = ray_setup(xform_point(in.w2o, o), xform_vector(in.w2o, d));
          ngroup = make_uint2(0u, 0x80000000u);
          tgroup = make_uint2(0u, 0u);
          state = LANE_RUN;
          break;
        }
      }
      if (!__any_sync(FULL, state != LANE_IDLE)) {
        if (exhausted && chunk_next == chunk_end) break;  // nothing in flight, nothing left
        continue;
      }
      if (!__any_sync(FULL, state == LANE_RUN)) continue;   // only finished rays: sink them first
    }
    // ---- 2. node step -----------------------------------------------------------------------
    if (state == LANE_RUN) {
      if (ngroup.y <= 0x00FFFFFFu) {  // (tgroup is always empty here: tested or pushed in step 3)
        if (!st.empty()) ngroup = st.pop();
        else state = LANE_WAIT;       // this instance is finished
      }
      if (state == LANE_RUN) {
        if (ngroup.y > 0x00FFFFFFu) {
          const uint32_t hits = ngroup.y;
          const int child_bit = bfind32(hits);
          const uint32_t child_base = ngroup.x;
          ngroup.y &= ~(1u << child_bit);
          if (ngroup.y > 0x00FFFFFFu) st.push(ngroup);
          const uint32_t slot = (uint32_t)(child_bit - 24) ^ (r.octinv4 & 7u);
          const uint32_t rel = (uint32_t)__popc(hits & ~(0xFFFFFFFFu << slot) & 0xFFu);
          const uint4* np = nodes + 5 * (size_t)(child_base + rel);
          const uint4 m0 = __ldg(np), n1 = __ldg(np + 1), n2 = __ldg(np + 2), n3 = __ldg(np + 3), n4 = __ldg(np + 4);
          const float4 n0 = make_float4(__uint_as_float(m0.x), __uint_as_float(m0.y), __uint_as_float(m0.z), __uint_as_float(m0.w));
          if (COUNT) tc.nodes++;
          const uint32_t hm = wide_node_test<COUNT>(n0, n1, n2, n3, n4, r, tmin, best.t);
          ngroup.x = n1.x;
          tgroup.x = n1.y;
          ngroup.y = (hm & 0xFF000000u) | (m0.w >> 24);
          tgroup.y = hm & 0x00FFFFFFu;
        } else {  // a popped leaf group (or nothing)
          tgroup = ngroup;
          ngroup = make_uint2(0u, 0u);
        }
      }
    }
    // ---- 3. leaf step: test now if enough lanes take part, otherwise keep it for later --------
    const bool has_tri = state == LANE_RUN && tgroup.y != 0u;
    const uint32_t tmask = __ballot_sync(FULL, has_tri);
    if (has_tri) {
      if (__popc(tmask) >= tri_min || ngroup.y <= 0x00FFFFFFu) {
        do {
          const int ti = bfind32(tgroup.y);
          tgroup.y &= ~(1u << ti);
          const float4* tp = tris + 3 * (size_t)(tgroup.x + (uint32_t)ti);
          const float4 a = __ldg(tp), bb = __ldg(tp + 1), c = __ldg(tp + 2);
          if (COUNT) tc.tris++;
          float t, u, v;
          if (tri_test(r, mk3(a.x, a.y, a.z), mk3(bb.x, bb.y, bb.z), mk3(c.x, c.y, c.z), tmin, best.t, t, u, v)) {
            if (ANY) { best.inst = inst; state = LANE_WAIT; cand = 0u; tgroup.y = 0u; }
            else consider_hit(best, inst, __float_as_uint(a.w), t, u, v);
          }
        } while (tgroup.y != 0u);
      } else {  // too few lanes: keep the group on the stack; it is popped right after this subtree
        st.push(tgroup);
        tgroup.y = 0u;
      }
    }
  }
}

}  // namespace rt
