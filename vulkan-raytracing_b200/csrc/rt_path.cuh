// rt_path.cuh — the bounce-fused, path-persistent frame kernel for the binary BVH (device only;
// included by rt_render.cu after the wavefront kernels, whose helpers it reuses).
//
// What it replaces: the per-bounce launch chain k_trace -> k_shade_miss -> k_shade_diffuse ->
// k_shadow -> k_shade_spec of the wavefront pipeline, i.e. the whole bounce loop of
// shader.rgen:84-178 including both traceRayEXT calls (:86-87 closest hit, :107-112 shadow).
//
// Why: measured on C3 (profiles/r02_baseline_bounces.txt) bounces 3-8 carry 0.15-0.56 M rays per
// launch against 151 552 resident lanes, so a launch is one 32-ray batch per warp and lasts as long
// as its slowest ray while the finished lanes idle (8-10 of 32 lanes active, 42 % of the traversal
// time for 17 % of the rays).  Here a lane owns a *path*: when its ray is finished the lane shades
// the hit and carries on with the reflected / refracted / shadow ray of the same path, and a lane
// whose path has ended takes the next primary ray of the queue k_raygen filled.  No ray state goes
// through HBM between bounces and there is one launch (one tail) per sub-frame.
//
// SIMD discipline: every lane is a small state machine and the warp advances phase by phase, each
// phase run only when enough lanes wait for it:
//   node  - lanes with an inner node pop/visit one binary node per step; a lane that reaches a
//           leaf parks it and walks on (Aila & Laine 2009, speculative traversal);
//   leaf  - run when >= t_leaf lanes hold a parked triangle (or nobody can walk);
//   serve - run when >= t_serve lanes have finished a mesh / a ray / a path: instance entry (world
//           box cull, ray into object space, watertight shear constants), shading (shader.rchit,
//           mirror / refraction, sky lookup, shadow-ray emission, Phong), refill from the queue.
// Results do not depend on any of this scheduling: closest hit = smallest t with the (instance,
// primitive) tie rule, any-hit is a boolean, and every sample slot is written by its own path in
// shader order (ambient at ray generation, then at most one overwrite).
#pragma once

namespace rt {

constexpr int PATH_STACK_SM = 12;                          // stack entries per lane in shared memory
constexpr int PATH_STACK_CAP = RT_MAX_TREE_DEPTH;          // build_blas rejects deeper trees
constexpr int PATH_STACK_LOC = PATH_STACK_CAP - PATH_STACK_SM;
constexpr int PATH_NONE = (int)0x80000000;                 // "no node" / "no parked leaf" (~slot is never this)
enum { PS_IDLE = 0, PS_WAIT = 1, PS_RUN = 2, PS_DONE = 3 };
// per-lane words parked in shared memory (column layout): world origin, world direction, and for a
// shadow ray the view direction, normal and instance of the diffuse hit it belongs to
enum { PK_OX = 0, PK_OY, PK_OZ, PK_DX, PK_DY, PK_DZ, PK_VX, PK_VY, PK_VZ, PK_NX, PK_NY, PK_NZ, PK_INST, PK_WORDS };

struct PathArgs {
  uint32_t* cnt;                 // counter block of the sub-frame: [C_NRAY] queue length, [C_WTRACE] queue cursor
  const float4* rayO;            // primary rays that reach the scene (k_raygen): origin | sample slot
  const float4* rayD;
  float4* sample;                // one radiance slot per sample
  rt_hit* primary;               // optional export of the primary hit records
  unsigned long long* stats64;   // COUNT: nodes visited / triangles tested (closest, shadow)
  int t_serve, t_leaf, t_left, chunk;
  uint32_t bounce0;              // bounce index of the rays in the queue (0 = primary rays)
};

#ifndef RT_PATH_MINBLK
#define RT_PATH_MINBLK 32
#endif

template <bool COUNT>
__global__ void __launch_bounds__(32, RT_PATH_MINBLK) k_path(SceneDev scene, FrameUniforms U, PathArgs a) {
  __shared__ uint32_t s_stack[PATH_STACK_SM * 32];
  __shared__ float s_keep[PK_WORDS * 32];
  uint32_t loc[PATH_STACK_LOC];
  const uint32_t FULL = 0xFFFFFFFFu;
  const uint32_t lane = threadIdx.x & 31u;
  const uint32_t lt_mask = (1u << lane) - 1u;
  uint32_t* stk = &s_stack[lane];
  float* keep = &s_keep[lane];

  int st = PS_IDLE;
  uint32_t slot = 0, bounce = 0, cand = 0, inst = 0;
  bool shadow = false;
  RayObj r;
  r.o = mk3(0, 0, 0); r.d = mk3(0, 0, 1); r.idir = mk3(1, 1, 1);
  r.kx = 0; r.ky = 1; r.kz = 2; r.Sx = r.Sy = 0.0f; r.Sz = 1.0f; r.octinv4 = 0;
  HitRec best;
  best.inst = RT_MISS; best.prim = RT_MISS; best.t = 0.0f; best.u = 0.0f; best.v = 0.0f;
  int node = PATH_NONE, parked = PATH_NONE, sp = 0;
  const float4* nodes = nullptr;
  const float4* tris = nullptr;
  uint32_t chunk_next = 0, chunk_end = 0;  // warp-uniform: this warp's slice of the ray queue
  bool exhausted = false;                  // warp-uniform
  const uint32_t n = a.cnt[C_NRAY];
  const uint32_t chunk = (uint32_t)a.chunk;
  const int t_serve = a.t_serve, t_leaf = a.t_leaf;
  uint32_t n_sec = 0, n_shadow = 0, n_miss = 0, n_diff = 0, n_spec = 0;
  uint32_t c_nodes[2] = {0, 0}, c_tris[2] = {0, 0};
  uint32_t c_iter[2] = {0, 0}, c_walk[2] = {0, 0}, c_cat[5] = {0, 0, 0, 0, 0};   // COUNT: node-loop iterations / walking lanes, [1] = after the queue ran dry

  auto pop_or_none = [&]() -> int {
    if (sp == 0) return PATH_NONE;
    --sp;
    return (int)(sp < PATH_STACK_SM ? stk[sp * 32] : loc[sp - PATH_STACK_SM]);
  };
  auto begin_ray = [&](f3 o, f3 d, float tmax) {
    keep[PK_OX * 32] = o.x; keep[PK_OY * 32] = o.y; keep[PK_OZ * 32] = o.z;
    keep[PK_DX * 32] = d.x; keep[PK_DY * 32] = d.y; keep[PK_DZ * 32] = d.z;
    best.inst = RT_MISS; best.prim = RT_MISS; best.u = 0.0f; best.v = 0.0f;
    best.t = __uint_as_float(__float_as_uint(tmax) - 1u);  // nextafter(tmax, -inf): "t <= best.t" == "t < tmax"
    cand = scene.n_inst >= 32u ? 0xFFFFFFFFu : (1u << scene.n_inst) - 1u;
    st = PS_WAIT;
  };

  while (true) {
    // ======================= serve: instance entry, shading, refill ==========================
    while (true) {
      // ---- (1) a lane with a ray and no mesh: enter the next instance whose world box it reaches
      if (__any_sync(FULL, st == PS_WAIT)) {
        if (st == PS_WAIT) {
          const f3 o = mk3(keep[PK_OX * 32], keep[PK_OY * 32], keep[PK_OZ * 32]);
          const f3 d = mk3(keep[PK_DX * 32], keep[PK_DY * 32], keep[PK_DZ * 32]);
          RayObj rw;
          rw.o = o;
          rw.idir = mk3(safe_rcp_dir(d.x), safe_rcp_dir(d.y), safe_rcp_dir(d.z));
          st = PS_DONE;  // unless an instance is left
          while (cand != 0u) {
            inst = (uint32_t)__ffs((int)cand) - 1u;
            cand &= cand - 1u;
            const InstDev& in = scene.inst[inst];
            if (scene.n_inst > 1u &&
                !slab_test(rw, in.wlo[0], in.wlo[1], in.wlo[2], in.whi[0], in.whi[1], in.whi[2], 0.001f, best.t))
              continue;
            const BlasDev& b = scene.blas[in.mesh];
            const uint32_t ntri = b.ntri;
            if (ntri == 0u) continue;
            r = ray_setup(xform_point(in.w2o, o), xform_vector(in.w2o, d));
            nodes = b.nodes;
            tris = b.tris;
            sp = 0;
            if (ntri == 1u) { node = PATH_NONE; parked = ~0; }  // a one-triangle mesh has no node
            else { node = 0; parked = PATH_NONE; }
            st = PS_RUN;
            break;
          }
        }
      }
      // ---- (2) a lane whose ray is finished: shade it (shader.rgen:90-178) -----------------------
      if (__any_sync(FULL, st == PS_DONE)) {
        if (st == PS_DONE) {
          const f3 d = mk3(keep[PK_DX * 32], keep[PK_DY * 32], keep[PK_DZ * 32]);
          st = PS_IDLE;  // the path ends here unless a new ray is begun below
          if (shadow) {
            shadow = false;
            if (best.inst == RT_MISS) {  // shader_shadow.rmiss ran: light the sample (shader.rgen:114-129)
              const f3 V = mk3(keep[PK_VX * 32], keep[PK_VY * 32], keep[PK_VZ * 32]);
              const f3 N = mk3(keep[PK_NX * 32], keep[PK_NY * 32], keep[PK_NZ * 32]);
              const f3 c = lit_color(scene, U, keep[PK_INST * 32], V, N, d, slot % U.spp);
              a.sample[slot] = make_float4(c.x, c.y, c.z, 1.0f);
            }
          } else {
            const bool hit = best.inst != RT_MISS;
            if (a.primary && bounce == 0u) {   // (bounce0 > 0: the wavefront launches exported them)
              uint32_t x, y;
              slot_to_pixel(U, slot / U.spp, x, y);
              rt_hit out;
              out.instance = best.inst; out.primitive = best.prim; out.t = hit ? best.t : 0.0f; out.u = best.u; out.v = best.v;
              a.primary[((size_t)y * U.width + x) * U.spp + (slot % U.spp)] = out;
            }
            if (!hit) {  // shader.rmiss + shader.rgen:90-94
              const f3 c = sky_sample(scene, mk3(d.x, d.y, -d.z));
              a.sample[slot] = make_float4(c.x, c.y, c.z, 1.0f);
              n_miss++;
            } else {
              f3 P, N;
              rchit(scene, best.inst, best.prim, best.u, best.v, P, N);
              const uint32_t type = scene.inst[best.inst].type;
              if (type == 0u) {  // diffuse, shader.rgen:97-131
                n_diff++;
                if (!(dot3(d, N) >= 0.0f)) {  // :104-105: a back face keeps the ambient colour
                  if (scene.mat) {
                    const f3 amb = material_ambient(scene.mat[best.inst]);
                    a.sample[slot] = make_float4(amb.x, amb.y, amb.z, 1.0f);
                  }
                  f3 so, L;
                  float dist;
                  shadow_ray(U, P, N, so, L, dist);
                  keep[PK_VX * 32] = d.x; keep[PK_VY * 32] = d.y; keep[PK_VZ * 32] = d.z;
                  keep[PK_NX * 32] = N.x; keep[PK_NY * 32] = N.y; keep[PK_NZ * 32] = N.z;
                  keep[PK_INST * 32] = __uint_as_float(best.inst);
                  shadow = true;
                  n_shadow++;
                  begin_ray(so, L, dist);  // shader.rgen:107-112
                }
              } else {  // mirror / refraction / any other type value, shader.rgen:132-165
                n_spec++;
                if (bounce < U.max_bounce) {  // the loop of :84 runs j = 0..maxBounceCount
                  f3 o = mk3(keep[PK_OX * 32], keep[PK_OY * 32], keep[PK_OZ * 32]), nd = d;
                  scatter(type, P, N, o, nd, scene.mat ? scene.mat[best.inst].ni : 1.52f);
                  bounce++;
                  n_sec++;
                  begin_ray(o, nd, 10000.0f);
                }
              }
            }
          }
        }
      }
      // ---- (3) a lane without a path: next primary ray of the queue ---------------------------------
      {
        const uint32_t idle = __ballot_sync(FULL, st == PS_IDLE);
        if (idle != 0u && !(exhausted && chunk_next == chunk_end)) {
          if (chunk_next == chunk_end) {
            uint32_t base = 0;
            if (lane == 0) base = atomicAdd(&a.cnt[C_WTRACE], chunk);
            base = __shfl_sync(FULL, base, 0);
            if (base >= n) exhausted = true;
            else { chunk_next = base; chunk_end = min(base + chunk, n); }
          }
          const uint32_t avail = chunk_end - chunk_next;
          if (st == PS_IDLE) {
            const uint32_t rank = (uint32_t)__popc(idle & lt_mask);
            if (rank < avail) {
              const float4 ro = a.rayO[chunk_next + rank], rd = a.rayD[chunk_next + rank];
              slot = __float_as_uint(ro.w);
              bounce = a.bounce0;
              shadow = false;
              begin_ray(mk3(ro.x, ro.y, ro.z), mk3(rd.x, rd.y, rd.z), 10000.0f);  // shader.rgen:86-87
            }
          }
          chunk_next += min((uint32_t)__popc(idle), avail);
        }
      }
      const uint32_t mw = __ballot_sync(FULL, st == PS_WAIT), md = __ballot_sync(FULL, st == PS_DONE);
      const uint32_t mr = __ballot_sync(FULL, st == PS_RUN);
      // once here, serve everybody: a lane left in DONE would wait through the whole next traverse phase
      // (it also counts toward t_serve there, so leftovers make the next serve come early and small).
      // Only a last straggler or two (rays that missed every instance box in this very pass) may stay.
      if (mw == 0u && __popc(md) < a.t_left && (mr != 0u || md == 0u)) break;
    }
    if (!__any_sync(FULL, st != PS_IDLE)) break;  // nothing in flight and the queue is empty

    // ======================= traverse: node steps, parked leaves ==============================
    {
      const bool queue_empty = exhausted && chunk_next == chunk_end;
      // lanes that already wait for the next serve phase
      while (true) {
        const uint32_t runm = __ballot_sync(FULL, st == PS_RUN);
        const int waiting0 = __popc(__ballot_sync(FULL, st == PS_DONE || st == PS_WAIT || (st == PS_IDLE && !queue_empty)));
        while (true) {
          if (node >= 0) {
            const float8 lo8 = ldg256(&nodes[4 * (size_t)node]), hi8 = ldg256(&nodes[4 * (size_t)node + 2]);
            const float4 n0 = lo8.a, n1 = lo8.b, n2 = hi8.a, n3 = hi8.b;
            if (COUNT) c_nodes[shadow ? 1 : 0]++;
            float tn0, tf0, tn1, tf1;
#ifdef RT_PATH_DEBUG
#define PDBG(...) do { if (slot == RT_PATH_DEBUG && st == PS_RUN) printf(__VA_ARGS__); } while (0)
#else
#define PDBG(...) ((void)0)
#endif
            slab_interval(r, n0.x, n0.y, n0.z, n0.w, n1.x, n1.y, 0.001f, best.t, tn0, tf0);
            slab_interval(r, n1.z, n1.w, n2.x, n2.y, n2.z, n2.w, 0.001f, best.t, tn1, tf1);
            const bool h0 = tn0 <= tf0, h1 = tn1 <= tf1;
            const int c0 = (int)__float_as_uint(n3.x), c1 = (int)__float_as_uint(n3.y);
            const bool both = h0 && h1;
            const bool swap = both ? (tn1 < tn0) : !h0;   // the child to visit first is c1
            const int cnear = swap ? c1 : c0, cfar = swap ? c0 : c1;
            bool need_pop = !(h0 || h1);
            node = cnear;
            if (both) {
              if (sp < PATH_STACK_SM) stk[sp * 32] = (uint32_t)cfar;
              else if (sp < PATH_STACK_CAP) loc[sp - PATH_STACK_SM] = (uint32_t)cfar;
              if (sp < PATH_STACK_CAP) sp++;
              else RT_FLAG_OVERFLOW();   // cannot happen for a tree build_blas accepted
            }
            if (!need_pop && node < 0 && parked == PATH_NONE) {  // park the first leaf and walk on
              parked = node;
              need_pop = true;
            }
            if (need_pop) node = pop_or_none();
            PDBG("node step: h0 %d h1 %d tn %g %g tf %g %g c0 %d c1 %d -> node %d parked %d sp %d best.t %g\n", (int)h0, (int)h1, tn0, tn1, tf0, tf1, c0, c1, node, parked, sp, best.t);
          }
          const uint32_t walk = __ballot_sync(FULL, node >= 0);
          const uint32_t pk = __ballot_sync(FULL, parked != PATH_NONE);
          if (COUNT) {
            c_iter[queue_empty ? 1 : 0]++;
            c_walk[queue_empty ? 1 : 0] += (uint32_t)__popc(walk);
            if (!queue_empty) {   // what the lanes that do not walk are waiting for
              c_cat[0] += (uint32_t)__popc(__ballot_sync(FULL, st == PS_IDLE));
              c_cat[1] += (uint32_t)__popc(__ballot_sync(FULL, st == PS_DONE || st == PS_WAIT));
              c_cat[2] += (uint32_t)__popc(__ballot_sync(FULL, st == PS_RUN && node == PATH_NONE && parked == PATH_NONE));
              c_cat[3] += (uint32_t)__popc(__ballot_sync(FULL, st == PS_RUN && node == PATH_NONE && parked != PATH_NONE));
              c_cat[4] += (uint32_t)__popc(__ballot_sync(FULL, st == PS_RUN && node < 0 && node != PATH_NONE));
            }
          }
          if (walk == 0u || __popc(pk) >= t_leaf) break;
          if (waiting0 + __popc(runm & ~walk & ~pk) >= t_serve) break;
        }
        // ---- parked triangles, together -------------------------------------------------------
        if (parked != PATH_NONE) {
          const float4* tp = tris + 3 * (size_t)(uint32_t)~parked;
          const float4 ta = __ldg(tp), tb = __ldg(tp + 1), tc = __ldg(tp + 2);
          if (COUNT) c_tris[shadow ? 1 : 0]++;
          PDBG("tri test: leaf %d prim %u best.t %g node %d sp %d\n", ~parked, __float_as_uint(ta.w), best.t, node, sp);
          parked = PATH_NONE;
          float t, u, v;
          if (tri_test(r, mk3(ta.x, ta.y, ta.z), mk3(tb.x, tb.y, tb.z), mk3(tc.x, tc.y, tc.z), 0.001f, best.t, t, u, v)) {
            if (shadow) {  // terminate on first hit (shader.rgen:111): the ray is finished
              best.inst = inst;
              node = PATH_NONE;
              sp = 0;
              cand = 0u;
            } else {
              consider_hit(best, inst, __float_as_uint(ta.w), t, u, v);
            }
          }
        }
        if (node < 0 && node != PATH_NONE) {  // a second leaf was reached while the first was parked
          parked = node;
          node = pop_or_none();
        }
        const bool fin = st == PS_RUN && node == PATH_NONE && parked == PATH_NONE;
        if (fin) st = cand != 0u ? PS_WAIT : PS_DONE;
        const uint32_t still = __ballot_sync(FULL, st == PS_RUN);
        if (still == 0u || waiting0 + __popc(runm & ~still) >= t_serve) break;
      }
    }
  }

  // ray counts by kind (exactly the oracle's) and, in the counting build, traversal work
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    n_sec += __shfl_xor_sync(FULL, n_sec, o);
    n_shadow += __shfl_xor_sync(FULL, n_shadow, o);
    n_miss += __shfl_xor_sync(FULL, n_miss, o);
    n_diff += __shfl_xor_sync(FULL, n_diff, o);
    n_spec += __shfl_xor_sync(FULL, n_spec, o);
  }
  if (lane == 0) {
    if (n_sec) atomicAdd(&a.cnt[C_NSEC], n_sec);
    if (n_shadow) atomicAdd(&a.cnt[C_NSHADOW], n_shadow);
    if (n_miss) atomicAdd(&a.cnt[C_NMISS], n_miss);
    if (n_diff) atomicAdd(&a.cnt[C_NDIFF], n_diff);
    if (n_spec) atomicAdd(&a.cnt[C_NSPEC], n_spec);
  }
  if (COUNT) {
    atomicAdd(&a.stats64[0], (unsigned long long)c_nodes[0]);
    atomicAdd(&a.stats64[1], (unsigned long long)c_tris[0]);
    atomicAdd(&a.stats64[2], (unsigned long long)c_nodes[1]);
    atomicAdd(&a.stats64[3], (unsigned long long)c_tris[1]);
    if (lane == 0) {   // SIMD occupancy of the node loop (RT_DEBUG_PATH prints it)
      atomicAdd(&a.stats64[4], (unsigned long long)c_iter[0]);
      atomicAdd(&a.stats64[5], (unsigned long long)c_walk[0]);
      atomicAdd(&a.stats64[6], (unsigned long long)c_iter[1]);
      atomicAdd(&a.stats64[7], (unsigned long long)c_walk[1]);
      for (int k = 0; k < 5; k++) atomicAdd(&a.stats64[8 + k], (unsigned long long)c_cat[k]);
    }
  }
}

}  // namespace rt
