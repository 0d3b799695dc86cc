// rt_render.cu — the wavefront frame pipeline for sm_100a: what one
// vkCmdTraceRaysKHR(width,height,1) (main.cpp:2620-2624) does in the reference.
//
//   k_raygen                 shader.rgen:61-81  primary rays -> ray queue
//   per bounce j = 0..maxBounceCount (shader.rgen:84):
//     k_trace                traceRayEXT closest hit (shader.rgen:86-87): persistent warps
//                            pull 32 rays at a time, traverse with a short stack in shared
//                            memory, then ballot-compact the ray ids into the miss /
//                            diffuse / specular queues (material-sorted)
//     k_shade_miss           shader.rmiss + skybox lookup (shader.rgen:90-94)
//     k_shade_diffuse        shader.rchit + back-face test, emits shadow rays (:97-110)
//     k_shadow               traceRayEXT terminate-on-first-hit + Phong (:111-130)
//     k_shade_spec           shader.rchit + mirror / refraction (:132-165) -> next queue
//   k_resolve                color /= samples; imageStore (:183-185)
//
// Ray state in HBM between stages: 2 x float4 per ray (origin|sample slot, direction),
// one uint4 hit record, 4-byte queue entries.  Each sample owns one float4 radiance slot
// written exactly once, so the resolve sums samples in shader order and the image is
// bit-reproducible whatever order warps retire in.
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>

// Traversal stacks are sized for the deepest tree build_blas accepts (RT_MAX_TREE_DEPTH); should an
// entry ever not fit it is dropped *safely* (the stack pointer does not move) and this device word
// records it, so a frame with missing geometry cannot pass silently: rt_get_stats reports it.
#include <cuda_runtime.h>
#include <cstdint>
namespace rt { __device__ uint32_t g_trav_flags; }
#if defined(__CUDA_ARCH__)
// a plain store of a constant: an atomic here drags warp-aggregation code (votes, uniform-datapath
// reduction) into the push path of the traversal loop and cost the closest-hit kernel 10 % (measured)
#define RT_FLAG_OVERFLOW() (*(volatile uint32_t*)&::rt::g_trav_flags = 1u)
#else
#define RT_FLAG_OVERFLOW() ((void)0)   // host pass of the RT_HD functions: never executed
#endif

#include "rt_internal.h"
#include "rt_wave.cuh"

namespace rt {

enum { C_NRAY = 0, C_NMISS, C_NDIFF, C_NSPEC, C_NSHADOW, C_WTRACE, C_WSHADOW, C_NCULL, C_NSEC, C_SHCUR, C_STRIDE = 16 };
// Block size of the persistent traversal kernels.  One warp per CTA: the warps share nothing but the
// stack array, and a one-warp CTA hands its slot to the next launch the moment its last batch is done
// (C3 frame 9.08 ms at 128 threads, 9.01 at 64, 8.93 at 32, 9.24 at 256: r01_sweep_occupancy.txt).
#ifndef RT_TRACE_THREADS
#define RT_TRACE_THREADS 32
#endif
constexpr int TRACE_THREADS = RT_TRACE_THREADS;
#ifndef RT_STACK_SM
#define RT_STACK_SM 8
#endif
constexpr int STACK_SM = RT_STACK_SM;      // stack entries per thread kept in shared memory
constexpr int STACK_LOCAL = RT_MAX_TREE_DEPTH - STACK_SM;  // overflow entries in local memory
// 8-byte words of shared memory per thread: STACK_SM 64-bit entries (8-wide layout), or STACK_SM 32-bit entries plus the
// 32 bytes of the parked ray (binary / 4-wide layouts, SlimStack)
constexpr int STACK_U2 = STACK_SM > STACK_SM / 2 + 4 ? STACK_SM : STACK_SM / 2 + 4;

// Traversal stack: the first STACK_SM entries live in shared memory (column per thread,
// bank-conflict free), the rarely used rest in local memory.
// The overflow array is a separate local array the struct only points to: with the array inside
// the struct the compiler kept the whole struct, stack pointer included, in local memory.
struct HybridStack {
  typedef uint2 Entry;
  uint2* sm;      // &smem[threadIdx.x], stride blockDim.x
  uint2* loc;     // STACK_LOCAL entries of thread-local memory
  int sp;
  __device__ __forceinline__ void reset() { sp = 0; }
  __device__ __forceinline__ bool empty() const { return sp == 0; }
  __device__ __forceinline__ void push(uint2 v) {
    if (sp < STACK_SM) sm[sp * TRACE_THREADS] = v;
    else if (sp < STACK_SM + STACK_LOCAL) loc[sp - STACK_SM] = v;
    else { RT_FLAG_OVERFLOW(); return; }   // dropped, flagged; pop() stays in bounds
    sp++;
  }
  __device__ __forceinline__ uint2 pop() {
    --sp;
    return sp < STACK_SM ? sm[sp * TRACE_THREADS] : loc[sp - STACK_SM];
  }
  // 32-bit entries (binary traversal); a packed 2 x STACK_SM word view of the same storage
  // measured 4 % slower than simply using the low half of an 8-byte slot
  __device__ __forceinline__ void push1(uint32_t v) { push(make_uint2(v, 0u)); }
  // n = 1..3 entries, e1 nearest (popped first).  Usual case: everything lands in the shared-memory part, three
  // predicated stores and one add instead of three pushes with their shared / local branches
  __device__ __forceinline__ void push_far_first(int n, uint32_t e1, uint32_t e2, uint32_t e3) {
    if (sp + 3 <= STACK_SM) {
      sm[(sp + n - 1) * TRACE_THREADS] = make_uint2(e1, 0u);
      if (n >= 2) sm[(sp + n - 2) * TRACE_THREADS] = make_uint2(e2, 0u);
      if (n >= 3) sm[sp * TRACE_THREADS] = make_uint2(e3, 0u);
      sp += n;
    } else {
      if (n >= 3) push1(e3);
      if (n >= 2) push1(e2);
      push1(e1);
    }
  }
  __device__ __forceinline__ uint32_t pop1() { return pop().x; }
  __device__ __forceinline__ void init(uint2* smem, Entry* local) { sm = smem + threadIdx.x; loc = local; sp = 0; }
  __device__ __forceinline__ void park_ray(const RayObj&) {}
  __device__ __forceinline__ const RayObj& tri_ray(const RayObj& r) const { return r; }
};

// Stack of the binary and 4-wide layouts (RT_SLIM_STACK, default): 32-bit entries, so the 8 shared-memory entries of a
// lane take half of the CTA's 2 KB; the other half keeps what only the triangle test reads of the object-space ray
// (origin, shear constants, axis permutation: 7 words), written once per mesh entry (park_ray) and read back with two
// LDS.128 at each leaf (tri_ray).  The node loop - where the kernel spends its issue slots and its registers - is left
// with 1/d and -o/d.  (`volatile` asm: the compiler must not keep the values in registers instead.)
#ifndef RT_SLIM_STACK
#define RT_SLIM_STACK 1
#endif
struct SlimStack {
  typedef uint32_t Entry;
  uint32_t* sm;    // &words[threadIdx.x], stride TRACE_THREADS
  float4* park;    // &park4[threadIdx.x], two float4 per lane, stride TRACE_THREADS
  uint32_t* loc;
  int sp;
  __device__ __forceinline__ void init(uint2* smem, Entry* local) {
    sm = (uint32_t*)smem + threadIdx.x;
    park = (float4*)((uint32_t*)smem + STACK_SM * TRACE_THREADS) + threadIdx.x;
    loc = local;
    sp = 0;
  }
  __device__ __forceinline__ void reset() { sp = 0; }
  __device__ __forceinline__ bool empty() const { return sp == 0; }
  __device__ __forceinline__ void push1(uint32_t v) {
    if (sp < STACK_SM) sm[sp * TRACE_THREADS] = v;
    else if (sp < STACK_SM + STACK_LOCAL) loc[sp - STACK_SM] = v;
    else { RT_FLAG_OVERFLOW(); return; }
    sp++;
  }
  __device__ __forceinline__ uint32_t pop1() {
    --sp;
    return sp < STACK_SM ? sm[sp * TRACE_THREADS] : loc[sp - STACK_SM];
  }
  __device__ __forceinline__ void push_far_first(int n, uint32_t e1, uint32_t e2, uint32_t e3) {
    if (sp + 3 <= STACK_SM) {
      sm[(sp + n - 1) * TRACE_THREADS] = e1;
      if (n >= 2) sm[(sp + n - 2) * TRACE_THREADS] = e2;
      if (n >= 3) sm[sp * TRACE_THREADS] = e3;
      sp += n;
    } else {
      if (n >= 3) push1(e3);
      if (n >= 2) push1(e2);
      push1(e1);
    }
  }
  // 64-bit entries belong to the 8-wide layout, which never runs on this stack (the calls only have to compile)
  __device__ __forceinline__ void push(uint2 v) { push1(v.x); }
  __device__ __forceinline__ uint2 pop() { return make_uint2(pop1(), 0u); }
  __device__ __forceinline__ void park_ray(const RayObj& r) {
    park[0] = make_float4(r.o.x, r.o.y, r.o.z, r.Sx);
    park[TRACE_THREADS] = make_float4(r.Sy, r.Sz, __uint_as_float((uint32_t)(r.kx | (r.ky << 2) | (r.kz << 4))), 0.0f);
  }
  __device__ __forceinline__ RayObj tri_ray(const RayObj& r) const {
    float4 a, b;
    const unsigned p0 = (unsigned)__cvta_generic_to_shared(park);
    asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w) : "r"(p0));
    asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w) : "r"(p0 + 16u * TRACE_THREADS));
    RayObj t;
    t.o = mk3(a.x, a.y, a.z);
    t.Sx = a.w; t.Sy = b.x; t.Sz = b.y;
    const uint32_t k = __float_as_uint(b.z);
    t.kx = (int)(k & 3u); t.ky = (int)((k >> 2) & 3u); t.kz = (int)(k >> 4);
    t.d = r.d; t.idir = r.idir; t.noid = r.noid; t.octinv4 = r.octinv4;   // not used by the triangle test
    return t;
  }
};

// the stack type a layout traverses with
template <int ACCEL> struct StackFor { typedef HybridStack type; };
#if RT_SLIM_STACK
template <> struct StackFor<ACCEL_BVH2> { typedef SlimStack type; };
template <> struct StackFor<ACCEL_BVH4> { typedef SlimStack type; };
#endif

template <int ACCEL, bool ANY, bool COUNT>
__device__ __forceinline__ bool trace_scene_t(const SceneDev& s, f3 o, f3 d, float tmin, float tmax, HitRec& best,
                                              typename StackFor<ACCEL>::type& st, TravCounters* cnt) {
  SceneDev sc = s;
  sc.accel = ACCEL;  // compile-time constant: the untaken traversal variants fold away
  return trace_scene<typename StackFor<ACCEL>::type, ANY, COUNT>(sc, o, d, tmin, tmax, best, st, cnt);
}

__device__ __forceinline__ uint32_t lane_id() { return threadIdx.x & 31u; }

// warp-aggregated append: returns the destination index for lanes with pred, ~0u otherwise
__device__ __forceinline__ uint32_t warp_append(bool pred, uint32_t* counter) {
  const uint32_t mask = __ballot_sync(0xFFFFFFFFu, pred);
  if (mask == 0u) return 0xFFFFFFFFu;
  const int leader = __ffs(mask) - 1;
  uint32_t base = 0;
  if ((int)lane_id() == leader) base = atomicAdd(counter, (uint32_t)__popc(mask));
  base = __shfl_sync(0xFFFFFFFFu, base, leader);
  return pred ? base + __popc(mask & ((1u << lane_id()) - 1u)) : 0xFFFFFFFFu;
}

// ---- ray generation ---------------------------------------------------------------------
// A primary ray that reaches no instance's padded world box is answered here, at the top level
// of the scene (what the TLAS does for traceRayEXT): the miss shader's sky lookup runs at once and
// the ray never enters the queue.  Only rays with something to traverse are written to HBM.
// Works on the slots [slot_base, slot_base + nslots) of the frame (one sub-frame).
__global__ void __launch_bounds__(256) k_raygen(SceneDev scene, FrameUniforms U, uint32_t slot_base, uint32_t nslots,
                                                float4* __restrict__ rayO, float4* __restrict__ rayD,
                                                float4* __restrict__ sample, uint32_t* __restrict__ cnt,
                                                rt_hit* __restrict__ primary, uint32_t* __restrict__ tile_live) {
  const uint32_t warps = (gridDim.x * blockDim.x) >> 5;
  const uint32_t w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const f3 amb = ambient_color();
  uint32_t culled = 0;
  // slot -> (pixel, sample) is seven integer divisions per slot in the general case (slot_to_pixel); with samples per pixel
  // and tile sides that are powers of two (the defaults: 4, 64 x 64) six of them are shifts and masks.  ncu: this kernel is
  // issue-bound (76 % of the slots busy) and a fifth of its instructions were those divisions.
  const uint32_t tile_px = U.tile_w * U.tile_h;
  const bool p2 = (U.spp & (U.spp - 1)) == 0 && (tile_px & (tile_px - 1)) == 0 && (U.tile_w & (U.tile_w - 1)) == 0 && U.tile_w >= 8;
  const int sh_spp = __ffs((int)U.spp) - 1, sh_tpx = __ffs((int)tile_px) - 1, sh_bx = __ffs((int)(U.tile_w >> 3)) - 1;
  for (uint64_t base = (uint64_t)w * 32; base < nslots; base += (uint64_t)warps * 32) {
    const uint32_t s = slot_base + (uint32_t)base + lane_id();
    bool valid = false, reach = false;
    f3 o = mk3(0, 0, 0), d = mk3(0, 0, 0);
    uint32_t smp = 0;
    if ((uint32_t)base + lane_id() < nslots) {
      uint32_t x, y;
      if (p2) {
        const uint32_t pix = s >> sh_spp, lt = pix >> sh_tpx, in = pix & (tile_px - 1u);
        smp = s & (U.spp - 1u);
        const uint32_t tile = lt * U.shard_count + U.shard_index;
        const uint32_t blk = in >> 5, l = in & 31u;
        const uint32_t tx = ((blk & ((U.tile_w >> 3) - 1u)) << 3) + (l & 7u), ty = ((blk >> sh_bx) << 2) + (l >> 3);
        const uint32_t trow = tile / U.tiles_x;
        x = (tile - trow * U.tiles_x) * U.tile_w + tx;
        y = trow * U.tile_h + ty;
        valid = x < U.width && y < U.height;
      } else {
        valid = slot_to_pixel(U, s / U.spp, x, y);
        smp = s % U.spp;
      }
      if (valid) {
        raygen(U, x, y, smp, o, d);
        RayObj rw;
        rw.o = o;
        rw.idir = mk3(safe_rcp_dir(d.x), safe_rcp_dir(d.y), safe_rcp_dir(d.z));
        if (scene.tlas) {  // many instances: the box around all of them decides here
          reach = slab_test(rw, scene.slo[0], scene.slo[1], scene.slo[2], scene.shi[0], scene.shi[1], scene.shi[2], 0.001f, 10000.0f);
        } else {
          for (uint32_t i = 0; i < scene.n_inst && !reach; i++) {
            const InstDev& in = scene.inst[i];
            reach = slab_test(rw, in.wlo[0], in.wlo[1], in.wlo[2], in.whi[0], in.whi[1], in.whi[2], 0.001f, 10000.0f);
          }
        }
        if (reach) {
          // every sample starts as tmpColor = Iamb*ka (shader.rgen:81)
          sample[s] = make_float4(amb.x, amb.y, amb.z, 1.0f);
        } else {
          const f3 c = sky_sample(scene, mk3(d.x, d.y, -d.z));  // shader.rmiss + shader.rgen:90-94
          sample[s] = make_float4(c.x, c.y, c.z, 1.0f);
          culled++;
          if (primary) {
            rt_hit out;
            out.instance = RT_MISS; out.primitive = RT_MISS; out.t = 0.0f; out.u = 0.0f; out.v = 0.0f;
            primary[((size_t)y * U.width + x) * U.spp + smp] = out;
          }
        }
      } else {
        sample[s] = make_float4(0.f, 0.f, 0.f, 0.f);  // padding outside the frame
      }
    }
    // a tile none of whose rays reaches the scene is final here: k_resolve stores it while the rest is traced
    // (a warp's 32 slots lie in one tile)
    if (tile_live && __any_sync(0xFFFFFFFFu, reach) && lane_id() == 0)
      tile_live[p2 ? (slot_base + (uint32_t)base) >> (sh_tpx + sh_spp) : (slot_base + (uint32_t)base) / (tile_px * U.spp)] = 1u;
    const uint32_t dst = warp_append(reach, &cnt[C_NRAY]);
    if (reach) {
      rayO[dst] = make_float4(o.x, o.y, o.z, __uint_as_float(s));
      rayD[dst] = make_float4(d.x, d.y, d.z, 0.0f);
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) culled += __shfl_xor_sync(0xFFFFFFFFu, culled, o);
  if (lane_id() == 0 && culled) atomicAdd(&cnt[C_NCULL], culled);
}

// ---- closest-hit traversal ----------------------------------------------------------------
struct TraceSource {
  const float4* rayO;
  const float4* rayD;
  __device__ __forceinline__ bool load(uint32_t idx, f3& o, f3& d, float& tmin, float& tmax) const {
    const float4 ro = rayO[idx], rd = rayD[idx];
    o = mk3(ro.x, ro.y, ro.z);
    d = mk3(rd.x, rd.y, rd.z);
    tmin = 0.001f;     // shader.rgen:87
    tmax = 10000.0f;
    return true;
  }
};

// hit record + material-sorted queue append for the rays a warp has just finished
struct TraceSink {
  const SceneDev& scene;
  const FrameUniforms& U;
  uint32_t* cnt;
  const float4* rayO;
  uint4* hits;
  uint32_t *qMiss, *qDiff, *qSpec;
  rt_hit* primary;
  __device__ __forceinline__ void done(bool finished, uint32_t idx, bool hit, const HitRec& h) const {
    int kind = -1;
    if (finished) {
      hits[idx] = make_uint4(h.inst, h.prim, __float_as_uint(h.u), __float_as_uint(h.v));
      kind = !hit ? 0 : (scene.inst[h.inst].type == 0u ? 1 : 2);
      if (primary) {
        const uint32_t s = __float_as_uint(rayO[idx].w);
        uint32_t x, y;
        slot_to_pixel(U, s / U.spp, x, y);
        rt_hit out;
        out.instance = h.inst; out.primitive = h.prim; out.t = hit ? h.t : 0.0f; out.u = h.u; out.v = h.v;
        primary[((size_t)y * U.width + x) * U.spp + (s % U.spp)] = out;
      }
    }
    uint32_t dst = warp_append(kind == 0, &cnt[C_NMISS]);
    if (kind == 0) qMiss[dst] = idx;
    dst = warp_append(kind == 1, &cnt[C_NDIFF]);
    if (kind == 1) qDiff[dst] = idx;
    dst = warp_append(kind == 2, &cnt[C_NSPEC]);
    if (kind == 2) qSpec[dst] = idx;
  }
};

// 32 resident warps per SM (64 registers): the traversal is latency- and issue-bound, and 32
// resident warps beat the 20 a 96-register build allows (profiles/r01_sweep_occupancy.txt)
#ifndef RT_TRACE_MINBLK
#define RT_TRACE_MINBLK (1024 / RT_TRACE_THREADS)
#endif
// the 4-wide layout holds a 128-byte node in registers: RT_BVH4_MINBLK trades resident warps for registers
#ifndef RT_BVH4_MINBLK
#define RT_BVH4_MINBLK RT_TRACE_MINBLK
#endif
#define RT_TRACE_BOUNDS __launch_bounds__(TRACE_THREADS, (ACCEL == ACCEL_BVH4 ? RT_BVH4_MINBLK : RT_TRACE_MINBLK))
template <int ACCEL, bool COUNT>
__global__ void RT_TRACE_BOUNDS
    k_trace(SceneDev scene, FrameUniforms U, uint32_t* __restrict__ cnt, const float4* __restrict__ rayO,
            const float4* __restrict__ rayD, uint4* __restrict__ hits, uint32_t* __restrict__ qMiss,
            uint32_t* __restrict__ qDiff, uint32_t* __restrict__ qSpec, rt_hit* __restrict__ primary,
            unsigned long long* __restrict__ stats64, const uint32_t* __restrict__ perm) {
  __shared__ uint2 stack_sm[STACK_U2 * TRACE_THREADS];
  typedef typename StackFor<ACCEL>::type Stack;
  typename Stack::Entry stack_loc[STACK_LOCAL];
  Stack st;
  st.init(stack_sm, stack_loc);
  TravCounters tc{0ull, 0ull};
  const uint32_t n = cnt[C_NRAY];
  while (true) {
    // (32 rays per pull also at the end of the queue: shrinking the last batches to even out the
    // finishing times of the warps costs more in SIMD width than the shorter tail returns,
    // profiles/r01_sweep_persistent.txt)
    uint32_t base = 0;
    if (lane_id() == 0) base = atomicAdd(&cnt[C_WTRACE], 32u);
    base = __shfl_sync(0xFFFFFFFFu, base, 0);
    if (base >= n) break;
    const bool active = base + lane_id() < n;
    const uint32_t idx = (perm && active) ? perm[base + lane_id()] : base + lane_id();
    int kind = -1;  // 0 miss, 1 diffuse, 2 specular/other
    if (active) {
      const float4 ro = rayO[idx], rd = rayD[idx];
      HitRec h;
      const bool hit = trace_scene_t<ACCEL, false, COUNT>(scene, mk3(ro.x, ro.y, ro.z), mk3(rd.x, rd.y, rd.z), 0.001f,
                                                          10000.0f, h, st, &tc);
      hits[idx] = make_uint4(h.inst, h.prim, __float_as_uint(h.u), __float_as_uint(h.v));
      kind = !hit ? 0 : (scene.inst[h.inst].type == 0u ? 1 : 2);
      if (primary) {
        const uint32_t s = __float_as_uint(ro.w);
        uint32_t x, y;
        slot_to_pixel(U, s / U.spp, x, y);
        rt_hit out;
        out.instance = h.inst; out.primitive = h.prim; out.t = hit ? h.t : 0.0f; out.u = h.u; out.v = h.v;
        primary[((size_t)y * U.width + x) * U.spp + (s % U.spp)] = out;
      }
    }
    uint32_t dst = warp_append(kind == 0, &cnt[C_NMISS]);
    if (kind == 0) qMiss[dst] = idx;
    dst = warp_append(kind == 1, &cnt[C_NDIFF]);
    if (kind == 1) qDiff[dst] = idx;
    dst = warp_append(kind == 2, &cnt[C_NSPEC]);
    if (kind == 2) qSpec[dst] = idx;
  }
  if (COUNT) {
    atomicAdd(&stats64[0], tc.nodes);
    atomicAdd(&stats64[1], tc.tris);
  }
}

// ---- miss: skybox -------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_shade_miss(SceneDev scene, const uint32_t* __restrict__ cnt,
                                                    const uint32_t* __restrict__ qMiss, const float4* __restrict__ rayO,
                                                    const float4* __restrict__ rayD, float4* __restrict__ sample) {
  const uint32_t n = cnt[C_NMISS];
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const uint32_t idx = qMiss[i];
    const float4 rd = rayD[idx];
    const uint32_t s = __float_as_uint(rayO[idx].w);
    const f3 c = sky_sample(scene, mk3(rd.x, rd.y, -rd.z));  // shader.rgen:92
    sample[s] = make_float4(c.x, c.y, c.z, 1.0f);
  }
}

// ---- diffuse hits: back-face test, shadow-ray emission --------------------------------------
__global__ void __launch_bounds__(256)
    k_shade_diffuse(SceneDev scene, uint32_t* __restrict__ cnt, const uint32_t* __restrict__ qDiff,
                    const float4* __restrict__ rayO, const float4* __restrict__ rayD, const uint4* __restrict__ hits,
                    float4* __restrict__ shadowQ, float4* __restrict__ sample) {
  const uint32_t n = cnt[C_NDIFF];
  const uint32_t warps = (gridDim.x * blockDim.x) >> 5;
  const uint32_t w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  for (uint64_t base = (uint64_t)w * 32; base < n; base += (uint64_t)warps * 32) {
    const uint32_t i = (uint32_t)base + lane_id();
    bool emit = false;
    f3 P = mk3(0, 0, 0), N = mk3(0, 0, 0), d = mk3(0, 0, 0);
    uint32_t s = 0, inst = 0;
    if (i < n) {
      const uint32_t idx = qDiff[i];
      const uint4 h = hits[idx];
      const float4 rd = rayD[idx];
      s = __float_as_uint(rayO[idx].w);
      d = mk3(rd.x, rd.y, rd.z);
      inst = h.x;
      rchit(scene, h.x, h.y, __uint_as_float(h.z), __uint_as_float(h.w), P, N);
      emit = !(dot3(d, N) >= 0.0f);  // shader.rgen:104-105: back face keeps the ambient colour
    }
    const uint32_t dst = warp_append(emit, &cnt[C_NSHADOW]);
    if (emit) {
      if (scene.mat) {  // rt_scene_set_materials: the surface's own ambient term, lit or not
        const f3 amb = material_ambient(scene.mat[inst]);
        sample[s] = make_float4(amb.x, amb.y, amb.z, 1.0f);
      }
      shadowQ[3 * (size_t)dst + 0] = make_float4(P.x, P.y, P.z, __uint_as_float(s));
      shadowQ[3 * (size_t)dst + 1] = make_float4(N.x, N.y, N.z, __uint_as_float(inst));
      shadowQ[3 * (size_t)dst + 2] = make_float4(d.x, d.y, d.z, 0.0f);
    }
  }
}

// ---- shadow rays + Phong ----------------------------------------------------------------------
// colour of an unshadowed diffuse hit; inst_bits = the instance id k_shade_diffuse left in the queue
__device__ __forceinline__ f3 lit_color(const SceneDev& scene, const FrameUniforms& U, float inst_bits, f3 d, f3 N, f3 L,
                                        uint32_t i) {
  return scene.mat ? phong_material(U, scene.mat[__float_as_uint(inst_bits)], d, N, L, i) : phong(U, d, N, L, i);
}

struct ShadowSource {
  const FrameUniforms& U;
  const float4* shadowQ;
  __device__ __forceinline__ bool load(uint32_t i, f3& o, f3& d, float& tmin, float& tmax) const {
    const float4 q0 = shadowQ[3 * (size_t)i], q1 = shadowQ[3 * (size_t)i + 1];
    shadow_ray(U, mk3(q0.x, q0.y, q0.z), mk3(q1.x, q1.y, q1.z), o, d, tmax);
    tmin = 0.001f;     // shader.rgen:112
    return true;
  }
};

struct ShadowSink {
  const SceneDev& scene;
  const FrameUniforms& U;
  const float4* shadowQ;
  float4* sample;
  __device__ __forceinline__ void done(bool finished, uint32_t i, bool occluded, const HitRec&) const {
    if (finished && !occluded) {  // shader_shadow.rmiss ran: light the sample (shader.rgen:114-129)
      const float4 q0 = shadowQ[3 * (size_t)i], q1 = shadowQ[3 * (size_t)i + 1], q2 = shadowQ[3 * (size_t)i + 2];
      const f3 P = mk3(q0.x, q0.y, q0.z), N = mk3(q1.x, q1.y, q1.z);
      f3 so, L;
      float dist;
      shadow_ray(U, P, N, so, L, dist);
      const uint32_t s = __float_as_uint(q0.w);
      const f3 c = lit_color(scene, U, q1.w, mk3(q2.x, q2.y, q2.z), N, L, s % U.spp);
      sample[s] = make_float4(c.x, c.y, c.z, 1.0f);
    }
  }
};

// trace_scene for looped instances (no top-level tree, <= 4 instances), called by the WHOLE warp: the instances are
// taken one after the other with the warp in step - an instance no lane's ray reaches is skipped by one vote, the lanes
// meet again after every instance - and nearest first for the warp as a whole (mean distance at which its rays enter the
// instance's world box), so that a hit in the nearer instance culls the box of the farther one.  Per-lane loops over the
// instances let the lanes of a warp drift into different instances, which then run the same traversal code one after the
// other (profiles/r02_migrate.txt).  The order only steers the search: closest hit = smallest t with the (instance,
// primitive) tie rule, any-hit = a boolean.
template <int ACCEL, bool ANY, bool COUNT>
__device__ __forceinline__ bool trace_scene_stepped(const SceneDev& s, bool have, f3 o, f3 d, float tmin, float tmax, HitRec& best,
                                                    typename StackFor<ACCEL>::type& st, TravCounters* cnt) {
  typedef typename StackFor<ACCEL>::type Stack;
  SceneDev sc = s;
  sc.accel = ACCEL;
  best.inst = RT_MISS; best.prim = RT_MISS; best.t = u2f(f2u(tmax) - 1u); best.u = 0.0f; best.v = 0.0f;
  RayObj rw;
  rw.o = o;
  rw.idir = mk3(safe_rcp_dir(d.x), safe_rcp_dir(d.y), safe_rcp_dir(d.z));
  uint32_t order = 0x3210u;
  if (s.n_inst > 1) {
    float key[4] = {3.0e38f, 3.0e38f, 3.0e38f, 3.0e38f};
#pragma unroll
    for (uint32_t i = 0; i < 4; i++) {
      if (i >= s.n_inst) break;
      const InstDev& in = s.inst[i];
      float tn = 0.0f, tf = -1.0f;
      if (have) slab_interval(rw, in.wlo[0], in.wlo[1], in.wlo[2], in.whi[0], in.whi[1], in.whi[2], tmin, tmax, tn, tf);
      const bool h = have && tn <= tf;
      float sum = h ? tn : 0.0f;
      uint32_t c = h ? 1u : 0u;
#pragma unroll
      for (int k = 16; k > 0; k >>= 1) { sum += __shfl_xor_sync(0xFFFFFFFFu, sum, k); c += __shfl_xor_sync(0xFFFFFFFFu, c, k); }
      key[i] = c ? sum / (float)c : 3.0e38f;
    }
    uint32_t id[4] = {0u, 1u, 2u, 3u};
#pragma unroll
    for (int a = 1; a < 4; a++)
#pragma unroll
      for (int bq = a; bq > 0; bq--)
        if (key[bq] < key[bq - 1]) { const float tk = key[bq]; key[bq] = key[bq - 1]; key[bq - 1] = tk; const uint32_t ti = id[bq]; id[bq] = id[bq - 1]; id[bq - 1] = ti; }
    order = id[0] | (id[1] << 4) | (id[2] << 8) | (id[3] << 12);
  }
  bool found = false;   // ANY: an occluder was hit
  for (uint32_t oi = 0; oi < s.n_inst; oi++) {
    const uint32_t i = (order >> (4 * oi)) & 0xFu;
    const InstDev& in = s.inst[i];
    bool go = have && !found;
    if (go && s.n_inst > 1) go = slab_test(rw, in.wlo[0], in.wlo[1], in.wlo[2], in.whi[0], in.whi[1], in.whi[2], tmin, best.t);
    if (!__any_sync(0xFFFFFFFFu, go)) continue;
    if (go) {
      const bool hit = trace_instance<Stack, ANY, COUNT>(sc, i, o, d, rw, false, tmin, best, st, cnt);
      if (ANY && hit) found = true;
    }
    __syncwarp();
  }
  return ANY ? found : best.inst != RT_MISS;
}

// The same for a scene with a top-level tree, called by the WHOLE warp.  A lane walks the top-level tree (nearest child
// first, as trace_scene does) only until it reaches an instance; then the warp meets, and all lanes that hold an instance
// traverse theirs together - different instances, the same traversal code in step.  In trace_scene a lane enters its
// instance the moment it reaches it, so at any time the lanes of a warp are spread over the top-level walk and the
// instance traversals at different depths of the call, and those run one after the other.
template <int ACCEL, bool ANY, bool COUNT>
__device__ __forceinline__ bool trace_scene_tlas_stepped(const SceneDev& s, bool have, f3 o, f3 d, float tmin, float tmax, HitRec& best,
                                                         typename StackFor<ACCEL>::type& st, TravCounters* cnt) {
  typedef typename StackFor<ACCEL>::type Stack;
  SceneDev sc = s;
  sc.accel = ACCEL;
  best.inst = RT_MISS; best.prim = RT_MISS; best.t = u2f(f2u(tmax) - 1u); best.u = 0.0f; best.v = 0.0f;
  RayObj rw;
  rw.o = o;
  rw.idir = mk3(safe_rcp_dir(d.x), safe_rcp_dir(d.y), safe_rcp_dir(d.z));
  int tstack[RT_MAX_TLAS_DEPTH];
  int tsp = 0, node = 0;
  bool walking = have;   // the lane's top-level walk is not finished
  while (true) {
    // ---- top level: until this lane stands on an instance (node < 0) or has nothing left
    while (walking && node >= 0) {
      const float8 lo8 = ldg256(&s.tlas[4 * (size_t)node]), hi8 = ldg256(&s.tlas[4 * (size_t)node + 2]);
      const float4 n0 = lo8.a, n1 = lo8.b, n2 = hi8.a, n3 = hi8.b;
      float tn0, tf0, tn1, tf1;
      slab_interval(rw, n0.x, n0.y, n0.z, n0.w, n1.x, n1.y, tmin, best.t, tn0, tf0);
      slab_interval(rw, n1.z, n1.w, n2.x, n2.y, n2.z, n2.w, tmin, best.t, tn1, tf1);
      const bool h0 = tn0 <= tf0, h1 = tn1 <= tf1;
      int c0 = (int)f2u(n3.x), c1 = (int)f2u(n3.y);
      if (h0 && h1) {
        if (tn1 < tn0) { const int t = c0; c0 = c1; c1 = t; }
        if (tsp < RT_MAX_TLAS_DEPTH) tstack[tsp++] = c1;
        else RT_FLAG_OVERFLOW();
        node = c0;
      } else if (h0) {
        node = c0;
      } else if (h1) {
        node = c1;
      } else if (tsp > 0) {
        node = tstack[--tsp];
      } else {
        walking = false;
      }
    }
    if (!__any_sync(0xFFFFFFFFu, walking)) break;
    // ---- the instances the lanes stand on, together
    if (walking) {
      const bool hit = trace_instance<Stack, ANY, COUNT>(sc, (uint32_t)~node, o, d, rw, false, tmin, best, st, cnt);
      if (ANY && hit) { walking = false; }
      else if (tsp > 0) node = tstack[--tsp];
      else walking = false;
    }
    __syncwarp();
  }
  return best.inst != RT_MISS;
}

template <int ACCEL, bool COUNT, bool STEP>
__global__ void RT_TRACE_BOUNDS
    k_shadow(SceneDev scene, FrameUniforms U, uint32_t* __restrict__ cnt, const float4* __restrict__ shadowQ,
             float4* __restrict__ sample, unsigned long long* __restrict__ stats64, const uint32_t* __restrict__ cnt0,
             uint32_t bounce) {
  __shared__ uint2 stack_sm[STACK_U2 * TRACE_THREADS];
  typedef typename StackFor<ACCEL>::type Stack;
  typename Stack::Entry stack_loc[STACK_LOCAL];
  Stack st;
  st.init(stack_sm, stack_loc);
  TravCounters tc{0ull, 0ull};
  const uint32_t n = cnt[C_NSHADOW];
  if (n == 0u) return;
  const bool stepped = STEP;
  if (cnt0) {   // append-only queue (k_trace_shade): this bounce's rays start behind those of the earlier bounces
    uint32_t start = 0;
    for (uint32_t j = 0; j < bounce; j++) start += cnt0[(size_t)j * C_STRIDE + C_NSHADOW];
    shadowQ += 3 * (size_t)start;
  }
  while (true) {
    uint32_t base = 0;
    if (lane_id() == 0) base = atomicAdd(&cnt[C_WSHADOW], 32u);
    base = __shfl_sync(0xFFFFFFFFu, base, 0);
    if (base >= n) break;
    const uint32_t i = base + lane_id();
    if (stepped) {   // looped instances: the warp takes them in step (trace_scene_stepped)
      const bool have = i < n;
      float4 q0 = make_float4(0, 0, 0, 0), q1 = make_float4(0, 0, 1, 0), q2 = make_float4(0, 0, 1, 0);
      if (have) { q0 = shadowQ[3 * (size_t)i]; q1 = shadowQ[3 * (size_t)i + 1]; q2 = shadowQ[3 * (size_t)i + 2]; }
      const f3 P = mk3(q0.x, q0.y, q0.z), N = mk3(q1.x, q1.y, q1.z), d = mk3(q2.x, q2.y, q2.z);
      const uint32_t s = __float_as_uint(q0.w);
      f3 so, L;
      float dist;
      shadow_ray(U, P, N, so, L, dist);
      HitRec h;
      const bool occluded = scene.tlas ? trace_scene_tlas_stepped<ACCEL, true, COUNT>(scene, have, so, L, 0.001f, dist, h, st, &tc)
                                       : trace_scene_stepped<ACCEL, true, COUNT>(scene, have, so, L, 0.001f, dist, h, st, &tc);
      if (have && !occluded) {
        const f3 c = lit_color(scene, U, q1.w, d, N, L, s % U.spp);
        sample[s] = make_float4(c.x, c.y, c.z, 1.0f);
      }
    } else if (i < n) {
      const float4 q0 = shadowQ[3 * (size_t)i], q1 = shadowQ[3 * (size_t)i + 1], q2 = shadowQ[3 * (size_t)i + 2];
      const f3 P = mk3(q0.x, q0.y, q0.z), N = mk3(q1.x, q1.y, q1.z), d = mk3(q2.x, q2.y, q2.z);
      const uint32_t s = __float_as_uint(q0.w);
      f3 so, L;
      float dist;
      shadow_ray(U, P, N, so, L, dist);
      HitRec h;
      const bool occluded = trace_scene_t<ACCEL, true, COUNT>(scene, so, L, 0.001f, dist, h, st, &tc);
      if (!occluded) {
        const f3 c = lit_color(scene, U, q1.w, d, N, L, s % U.spp);
        sample[s] = make_float4(c.x, c.y, c.z, 1.0f);
      }
    }
  }
  if (COUNT) {
    atomicAdd(&stats64[2], tc.nodes);
    atomicAdd(&stats64[3], tc.tris);
  }
}

// ---- closest hit + the shading of what was hit, one launch per bounce ---------------------------------------
// traceRayEXT (shader.rgen:86-87) and the rest of the loop body (:90-165) for the rays of one bounce: a lane that has
// finished its ray runs shader.rmiss / shader.rchit right here instead of handing the hit to k_shade_miss /
// k_shade_diffuse / k_shade_spec through a hit record and an index queue.  Per bounce and sub-frame that leaves this
// launch and the shadow launch (4 launches fewer); what it buys is not the launch overhead but the *wait*: the three
// small kernels could only start when blocks of the other sub-frame's persistent traversal launch retired (all 32 CTA
// slots and the whole register file of every SM are taken), which left 0.1-0.5 ms between consecutive closest-hit
// launches of a sub-frame in bounces >= 3, where the frame is a chain of short launches (profiles/r02_fused_shade.txt).
// Shadow rays are appended to a queue that only grows during the frame (every path emits at most one): the shadow
// launch of bounce j reads its part while bounce j + 1 already appends behind it, so nothing waits for a shadow launch.
template <int ACCEL, bool COUNT, bool TLAS>
__global__ void RT_TRACE_BOUNDS
    k_trace_shade(SceneDev scene, FrameUniforms U, uint32_t* __restrict__ cnt0, uint32_t bounce, const float4* __restrict__ rayO,
                  const float4* __restrict__ rayD, float4* __restrict__ nextO, float4* __restrict__ nextD,
                  float4* __restrict__ shadowQ, float4* __restrict__ sample, rt_hit* __restrict__ primary,
                  unsigned long long* __restrict__ stats64) {
  __shared__ uint2 stack_sm[STACK_U2 * TRACE_THREADS];
  typedef typename StackFor<ACCEL>::type Stack;
  typename Stack::Entry stack_loc[STACK_LOCAL];
  Stack st;
  st.init(stack_sm, stack_loc);
  TravCounters tc{0ull, 0ull};
  uint32_t* cnt = cnt0 + (size_t)bounce * C_STRIDE;
  uint32_t* cntNext = cnt + C_STRIDE;
  const uint32_t n = cnt[C_NRAY];
  const bool last = bounce >= U.max_bounce;   // the loop of shader.rgen:84 runs j = 0..maxBounceCount
  uint32_t n_miss = 0, n_diff = 0, n_spec = 0, n_shadow = 0;
  while (true) {
    uint32_t base = 0;
    if (lane_id() == 0) base = atomicAdd(&cnt[C_WTRACE], 32u);
    base = __shfl_sync(0xFFFFFFFFu, base, 0);
    if (base >= n) break;
    const uint32_t idx = base + lane_id();
    bool emit = false, cont = false;
    f3 P = mk3(0, 0, 0), N = mk3(0, 0, 0), o = mk3(0, 0, 0), d = mk3(0, 0, 0), dn = mk3(0, 0, 0);
    uint32_t slot = 0, inst = 0;
    const bool have = idx < n;
    HitRec h;
    bool hit = false;
    if (have) {
      const float4 ro = rayO[idx], rd = rayD[idx];
      slot = __float_as_uint(ro.w);
      o = mk3(ro.x, ro.y, ro.z);
      d = mk3(rd.x, rd.y, rd.z);
    }
    // shader.rgen:86-87.  (TLAS: scene with a top-level tree, traversed with the warp in step)
    if (TLAS) hit = trace_scene_tlas_stepped<ACCEL, false, COUNT>(scene, have, o, d, 0.001f, 10000.0f, h, st, &tc);
    else if (have) hit = trace_scene_t<ACCEL, false, COUNT>(scene, o, d, 0.001f, 10000.0f, h, st, &tc);
    if (have) {
      if (primary) {
        uint32_t x, y;
        slot_to_pixel(U, slot / U.spp, x, y);
        rt_hit out;
        out.instance = h.inst; out.primitive = h.prim; out.t = hit ? h.t : 0.0f; out.u = h.u; out.v = h.v;
        primary[((size_t)y * U.width + x) * U.spp + (slot % U.spp)] = out;
      }
      if (!hit) {   // shader.rmiss + shader.rgen:90-94
        const f3 c = sky_sample(scene, mk3(d.x, d.y, -d.z));
        sample[slot] = make_float4(c.x, c.y, c.z, 1.0f);
        n_miss++;
      } else {
        inst = h.inst;
        rchit(scene, h.inst, h.prim, h.u, h.v, P, N);
        const uint32_t type = scene.inst[h.inst].type;
        if (type == 0u) {   // diffuse, shader.rgen:97-110: a back face keeps the ambient colour, a front face asks the light
          n_diff++;
          emit = !(dot3(d, N) >= 0.0f);
          if (emit && scene.mat) {   // rt_scene_set_materials: the surface's own ambient term, lit or not
            const f3 amb = material_ambient(scene.mat[inst]);
            sample[slot] = make_float4(amb.x, amb.y, amb.z, 1.0f);
          }
        } else {            // mirror / refraction, shader.rgen:132-165
          n_spec++;
          if (!last) {
            dn = d;
            scatter(type, P, N, o, dn, scene.mat ? scene.mat[inst].ni : 1.52f);
            cont = true;
          }
        }
      }
    }
    uint32_t dst = warp_append(emit, &cnt0[C_SHCUR]);
    if (emit) {
      n_shadow++;
      shadowQ[3 * (size_t)dst + 0] = make_float4(P.x, P.y, P.z, __uint_as_float(slot));
      shadowQ[3 * (size_t)dst + 1] = make_float4(N.x, N.y, N.z, __uint_as_float(inst));
      shadowQ[3 * (size_t)dst + 2] = make_float4(d.x, d.y, d.z, 0.0f);
    }
    dst = warp_append(cont, &cntNext[C_NRAY]);
    if (cont) {
      nextO[dst] = make_float4(o.x, o.y, o.z, __uint_as_float(slot));
      nextD[dst] = make_float4(dn.x, dn.y, dn.z, 0.0f);
    }
  }
#pragma unroll
  for (int k = 16; k > 0; k >>= 1) {
    n_miss += __shfl_xor_sync(0xFFFFFFFFu, n_miss, k);
    n_diff += __shfl_xor_sync(0xFFFFFFFFu, n_diff, k);
    n_spec += __shfl_xor_sync(0xFFFFFFFFu, n_spec, k);
    n_shadow += __shfl_xor_sync(0xFFFFFFFFu, n_shadow, k);
  }
  if (lane_id() == 0) {
    if (n_miss) atomicAdd(&cnt[C_NMISS], n_miss);
    if (n_diff) atomicAdd(&cnt[C_NDIFF], n_diff);
    if (n_spec) atomicAdd(&cnt[C_NSPEC], n_spec);
    if (n_shadow) atomicAdd(&cnt[C_NSHADOW], n_shadow);
  }
  if (COUNT) {
    atomicAdd(&stats64[0], tc.nodes);
    atomicAdd(&stats64[1], tc.tris);
  }
}

// ---- closest hit + shading for looped instances, the warp in step (k_trace_shade_step; default for the binary layout) ----
// Same work as k_trace_shade; two differences in how the warp goes through a batch of 32 rays.
//  * The instances are taken one after the other by the whole warp: a vote skips an instance no lane's ray reaches, the
//    lanes meet again after every instance, and the order is nearest first for the warp as a whole (mean distance at
//    which its rays enter the instance's world box), so the hit found in the nearer instance culls the farther one's box.
//    trace_scene's per-lane loop lets lanes drift into different instances, which then run the same traversal code one
//    after the other, and visits the instances in index order: C3 8.48 -> 8.31 ms by the stepping, -> 7.74 ms by the
//    order (798 M instead of 884 M node visits per frame); profiles/r02_migrate.txt.
//  * CUT = true (RT_MIGRATE = T > 0, opt-in, measured slower): a batch ends when no more than T of the lanes traversing
//    an instance are left; those lanes write their traversal state (node, stack, closest hit so far, instance) to a
//    buffer private to the warp and the warp takes the next batch; when the buffer holds a batch's worth (or the queue
//    is empty) the warp resumes those rays together, from where they stopped - long rays share warps with long rays.
//    The check (one LDS, one compare, one more loop exit) costs the loop 20-25 % however rarely it fires.
// Binary layout, instances looped (no top-level tree), one-warp CTAs.
constexpr int MIG_STACK = 40;   // deepest stack a ray may have when it is set aside (deeper ones keep running)
struct __align__(16) MigEntry {
  uint32_t idx, where, node;     // ray, (instance | stack depth << 8), node to visit next
  uint32_t b_inst, b_prim;
  float b_t, b_u, b_v;
  uint32_t stack[MIG_STACK];
};

// trace_bvh2_direct with an exit (CUT): returns true when the lane was asked to stop (node and stack say where).  The
// live count is read beside the node and looked at when the node's boxes have been tested.  Measured: the check costs
// the loop 20-25 % wherever it is placed (end of the node visit, every 8th visit, only where the stack is popped) and
// whatever T is - more than setting the stragglers aside returns (profiles/r02_migrate.txt) - so CUT is opt-in.
template <bool COUNT, bool CUT>
__device__ __forceinline__ bool trace_bvh2_mig(const BlasDev& b, const RayObj& r, float tmin, uint32_t inst, HitRec& best,
                                               SlimStack& st, int& node, uint32_t live_addr, int T, TravCounters* cnt) {
  st.park_ray(r);
  while (true) {
    if (node < 0) {
      test_leaf_tri<false, COUNT>(b, (uint32_t)~node, st.tri_ray(r), tmin, inst, best, cnt);
      if (st.empty()) break;
      node = (int)st.pop1();
      continue;
    }
    const float8 lo8 = ldg256(&b.nodes[4 * (size_t)node]), hi8 = ldg256(&b.nodes[4 * (size_t)node + 2]);
    int lv = 0x7FFFFFFF;
    if (CUT) asm volatile("ld.volatile.shared.u32 %0, [%1];" : "=r"(lv) : "r"(live_addr));
    const float4 n0 = lo8.a, n1 = lo8.b, n2 = hi8.a, n3 = hi8.b;
    if (COUNT) cnt->nodes++;
    float tn0, tf0, tn1, tf1;
    RT_SLAB(r, n0.x, n0.y, n0.z, n0.w, n1.x, n1.y, tmin, best.t, tn0, tf0);
    RT_SLAB(r, n1.z, n1.w, n2.x, n2.y, n2.z, n2.w, tmin, best.t, tn1, tf1);
    const bool h0 = tn0 <= tf0, h1 = tn1 <= tf1;
    int c0 = (int)f2u(n3.x), c1 = (int)f2u(n3.y);
    if (h0 && h1) {
      if (tn1 < tn0) { const int t = c0; c0 = c1; c1 = t; }
      st.push1((uint32_t)c1);
      node = c0;
    } else if (h0) {
      node = c0;
    } else if (h1) {
      node = c1;
    } else {
      if (st.empty()) break;
      node = (int)st.pop1();
    }
    if (CUT && lv <= T && st.sp <= MIG_STACK) return true;   // `node` (an inner node or a leaf) is where the ray goes on
  }
  return false;
}

template <bool COUNT, bool CUT>
__global__ void __launch_bounds__(32, 32)
    k_trace_shade_step(SceneDev scene, FrameUniforms U, uint32_t* __restrict__ cnt0, uint32_t bounce, const float4* __restrict__ rayO,
                      const float4* __restrict__ rayD, float4* __restrict__ nextO, float4* __restrict__ nextD,
                      float4* __restrict__ shadowQ, float4* __restrict__ sample, rt_hit* __restrict__ primary,
                      unsigned long long* __restrict__ stats64, MigEntry* __restrict__ mig, int T, int F) {
  __shared__ uint2 stack_sm[STACK_U2 * 32];
  __shared__ int s_live, s_nbuf;
  uint32_t stack_loc[STACK_LOCAL];
  SlimStack st;
  st.init(stack_sm, stack_loc);
  TravCounters tc{0ull, 0ull};
  MigEntry* my = mig + (size_t)blockIdx.x * 32;
  // shared-memory address of s_live, made opaque so that it stays in a register (rebuilt from SR_CgaCtaId with an
  // S2R per node visit otherwise)
  uint32_t live_addr = (uint32_t)__cvta_generic_to_shared(&s_live);
  asm volatile("mov.u32 %0, %0;" : "+r"(live_addr));
  uint32_t* cnt = cnt0 + (size_t)bounce * C_STRIDE;
  uint32_t* cntNext = cnt + C_STRIDE;
  const uint32_t n = cnt[C_NRAY];
  const bool last = bounce >= U.max_bounce;
  uint32_t n_miss = 0, n_diff = 0, n_spec = 0, n_shadow = 0;
  bool main_done = false;   // warp-uniform
  if (lane_id() == 0) s_nbuf = 0;
  __syncwarp();
  while (true) {
    const int nbuf = s_nbuf;   // rays this warp has set aside (uniform: read behind a warp barrier)
    bool have = false, resumed = false;
    uint32_t idx = 0;
    const int room = T * (int)min(scene.n_inst, 8u);   // a batch can set aside up to T rays per instance
    if (!main_done && nbuf + room <= 32 && nbuf < F) {
      uint32_t base = 0;
      if (lane_id() == 0) base = atomicAdd(&cnt[C_WTRACE], 32u);
      base = __shfl_sync(0xFFFFFFFFu, base, 0);
      if (base >= n) main_done = true;
      else { idx = base + lane_id(); have = idx < n; }
    }
    int inst_i = 0, node = 0;
    HitRec best;
    best.inst = RT_MISS; best.prim = RT_MISS; best.t = u2f(f2u(10000.0f) - 1u); best.u = 0.0f; best.v = 0.0f;
    st.sp = 0;
    if (main_done || nbuf + room > 32 || nbuf >= F) {   // resume what was set aside
      if (nbuf == 0) break;             // (only reached with the queue empty)
      resumed = true;
      have = (int)lane_id() < nbuf;
      if (have) {
        const MigEntry& e = my[lane_id()];   // (L2 loads: the entry was stored by another lane of this warp)
        const uint4 h0 = __ldcg((const uint4*)&e), h1 = __ldcg((const uint4*)&e + 1);
        idx = h0.x;
        inst_i = (int)(h0.y & 0xFFu);
        const int sp = (int)(h0.y >> 8);
        node = (int)h0.z;
        best.inst = h0.w; best.prim = h1.x; best.t = __uint_as_float(h1.y); best.u = __uint_as_float(h1.z); best.v = __uint_as_float(h1.w);
        for (int k = 0; k < sp; k++) st.push1(__ldcg(&e.stack[k]));
      }
      __syncwarp();
      if (lane_id() == 0) s_nbuf = 0;
    }
    // a batch that nothing can follow (queue empty) runs to completion
    const int Tnow = main_done ? -1 : T;
    bool emit = false, cont = false, aside = false;
    f3 P = mk3(0, 0, 0), N = mk3(0, 0, 0), o = mk3(0, 0, 0), d = mk3(0, 0, 0), dn = mk3(0, 0, 0);
    uint32_t slot = 0, inst = 0;
    RayObj rw;
    rw.o = o; rw.idir = mk3(1, 1, 1);
    if (have) {
      const float4 ro = rayO[idx], rd = rayD[idx];
      slot = __float_as_uint(ro.w);
      o = mk3(ro.x, ro.y, ro.z);
      d = mk3(rd.x, rd.y, rd.z);
      rw.o = o;
      rw.idir = mk3(safe_rcp_dir(d.x), safe_rcp_dir(d.y), safe_rcp_dir(d.z));
    }
    // The instances one after the other, the whole warp in step (a lane that has nothing to do in instance i waits for
    // the others at the end of the iteration, as it does in trace_scene): s_live counts the lanes traversing instance i.
    // Order of the instances for this batch: nearest first for the warp as a whole (mean distance at which its rays enter
    // the instance's world box), so that the hit found in the nearer instance culls the box of the farther one.  The order
    // only steers the search: the closest hit with the (instance, primitive) tie rule does not depend on it.
    uint32_t order = 0x3210u;   // 4 bits per position
    if (!CUT && scene.n_inst > 1) {
      float key[4] = {3.0e38f, 3.0e38f, 3.0e38f, 3.0e38f};
#pragma unroll
      for (uint32_t i = 0; i < 4; i++) {
        if (i >= scene.n_inst) break;
        const InstDev& in = scene.inst[i];
        float tn = 0.0f, tf = -1.0f;
        if (have) slab_interval(rw, in.wlo[0], in.wlo[1], in.wlo[2], in.whi[0], in.whi[1], in.whi[2], 0.001f, 10000.0f, tn, tf);
        const bool h = have && tn <= tf;
        float sum = h ? tn : 0.0f;
        uint32_t c = h ? 1u : 0u;
#pragma unroll
        for (int k = 16; k > 0; k >>= 1) { sum += __shfl_xor_sync(0xFFFFFFFFu, sum, k); c += __shfl_xor_sync(0xFFFFFFFFu, c, k); }
        key[i] = c ? sum / (float)c : 3.0e38f;
      }
      // (xor-butterfly sums are the same in every lane: the order is warp-uniform)
      uint32_t id[4] = {0u, 1u, 2u, 3u};
#pragma unroll
      for (int a = 1; a < 4; a++)
#pragma unroll
        for (int bq = a; bq > 0; bq--)
          if (key[bq] < key[bq - 1]) { const float tk = key[bq]; key[bq] = key[bq - 1]; key[bq - 1] = tk; const uint32_t ti = id[bq]; id[bq] = id[bq - 1]; id[bq - 1] = ti; }
      order = id[0] | (id[1] << 4) | (id[2] << 8) | (id[3] << 12);
    }
    for (uint32_t oi = 0; oi < scene.n_inst; oi++) {
      const uint32_t i = CUT ? oi : ((order >> (4 * oi)) & 0xFu);
      const InstDev& in = scene.inst[i];
      const bool mid = resumed && (int)i == inst_i;   // this lane enters instance i in the middle of its traversal
      bool go = have && !aside && (int)i >= inst_i;
      if (go && !mid && scene.n_inst > 1)
        go = slab_test(rw, in.wlo[0], in.wlo[1], in.wlo[2], in.whi[0], in.whi[1], in.whi[2], 0.001f, best.t);
      const BlasDev b = scene.blas[in.mesh];
      if (go && b.ntri == 0) go = false;
      const uint32_t gm = __ballot_sync(0xFFFFFFFFu, go);
      if (gm == 0u) continue;
      if (lane_id() == 0) s_live = __popc(gm);
      __syncwarp();
      // rays are only set aside where many started together: the last T of them, at most a quarter (a phase that begins
      // with three lanes has no stragglers, only three rays)
      const int thr = Tnow < 0 ? -1 : min(Tnow, __popc(gm) >> 2);
      if (go) {
        const RayObj r = ray_setup(xform_point(in.w2o, o), xform_vector(in.w2o, d));
        if (b.ntri == 1) {
          test_leaf_tri<false, COUNT>(b, 0, r, 0.001f, i, best, &tc);
        } else {
          if (!mid) { st.sp = 0; node = 0; }
          if (trace_bvh2_mig<COUNT, CUT>(b, r, 0.001f, i, best, st, node, live_addr, thr, &tc)) {
            const int pos = atomicAdd(&s_nbuf, 1);
            MigEntry& e = my[pos];
            e.idx = idx;
            e.where = i | ((uint32_t)st.sp << 8);
            e.node = (uint32_t)node;
            e.b_inst = best.inst; e.b_prim = best.prim; e.b_t = best.t; e.b_u = best.u; e.b_v = best.v;
            const int sp = st.sp;
            for (int k = 0; k < sp; k++) e.stack[k] = k < STACK_SM ? st.sm[k * 32] : st.loc[k - STACK_SM];
            aside = true;
          }
        }
        atomicSub(&s_live, 1);
      }
      __syncwarp();
    }
    if (have) {
      if (!aside) {
        const bool hit = best.inst != RT_MISS;
        if (primary) {
          uint32_t x, y;
          slot_to_pixel(U, slot / U.spp, x, y);
          rt_hit out;
          out.instance = best.inst; out.primitive = best.prim; out.t = hit ? best.t : 0.0f; out.u = best.u; out.v = best.v;
          primary[((size_t)y * U.width + x) * U.spp + (slot % U.spp)] = out;
        }
        if (!hit) {
          const f3 c = sky_sample(scene, mk3(d.x, d.y, -d.z));
          sample[slot] = make_float4(c.x, c.y, c.z, 1.0f);
          n_miss++;
        } else {
          inst = best.inst;
          rchit(scene, best.inst, best.prim, best.u, best.v, P, N);
          const uint32_t type = scene.inst[best.inst].type;
          if (type == 0u) {
            n_diff++;
            emit = !(dot3(d, N) >= 0.0f);
            if (emit && scene.mat) {
              const f3 amb = material_ambient(scene.mat[inst]);
              sample[slot] = make_float4(amb.x, amb.y, amb.z, 1.0f);
            }
          } else {
            n_spec++;
            if (!last) {
              dn = d;
              scatter(type, P, N, o, dn, scene.mat ? scene.mat[inst].ni : 1.52f);
              cont = true;
            }
          }
        }
      }
    }
    uint32_t dst = warp_append(emit, &cnt0[C_SHCUR]);
    if (emit) {
      n_shadow++;
      shadowQ[3 * (size_t)dst + 0] = make_float4(P.x, P.y, P.z, __uint_as_float(slot));
      shadowQ[3 * (size_t)dst + 1] = make_float4(N.x, N.y, N.z, __uint_as_float(inst));
      shadowQ[3 * (size_t)dst + 2] = make_float4(d.x, d.y, d.z, 0.0f);
    }
    dst = warp_append(cont, &cntNext[C_NRAY]);
    if (cont) {
      nextO[dst] = make_float4(o.x, o.y, o.z, __uint_as_float(slot));
      nextD[dst] = make_float4(dn.x, dn.y, dn.z, 0.0f);
    }
    __syncwarp();   // s_nbuf and the entries written by this batch are visible to the whole warp
  }
#pragma unroll
  for (int k = 16; k > 0; k >>= 1) {
    n_miss += __shfl_xor_sync(0xFFFFFFFFu, n_miss, k);
    n_diff += __shfl_xor_sync(0xFFFFFFFFu, n_diff, k);
    n_spec += __shfl_xor_sync(0xFFFFFFFFu, n_spec, k);
    n_shadow += __shfl_xor_sync(0xFFFFFFFFu, n_shadow, k);
  }
  if (lane_id() == 0) {
    if (n_miss) atomicAdd(&cnt[C_NMISS], n_miss);
    if (n_diff) atomicAdd(&cnt[C_NDIFF], n_diff);
    if (n_spec) atomicAdd(&cnt[C_NSPEC], n_spec);
    if (n_shadow) atomicAdd(&cnt[C_NSHADOW], n_shadow);
  }
  if (COUNT) {
    atomicAdd(&stats64[0], tc.nodes);
    atomicAdd(&stats64[1], tc.tris);
  }
}

}  // namespace rt
#include "rt_path.cuh"   // the bounce-fused path kernel (uses lit_color, warp helpers, the counter slots)
namespace rt {

// ---- warp-synchronous variants (rt_wave.cuh) ---------------------------------------------------
struct WaveTraceSource {
  const float4* rayO;
  const float4* rayD;
  const uint32_t* perm;
  __device__ __forceinline__ void load(uint32_t i, f3& o, f3& d, float& tmax) const {
    const uint32_t idx = perm ? perm[i] : i;
    const float4 ro = rayO[idx], rd = rayD[idx];
    o = mk3(ro.x, ro.y, ro.z);
    d = mk3(rd.x, rd.y, rd.z);
    tmax = 10000.0f;   // shader.rgen:87
  }
};

struct WaveTraceSink {
  const SceneDev& scene;
  const FrameUniforms& U;
  uint32_t* cnt;
  const float4* rayO;
  const uint32_t* perm;
  uint4* hits;
  uint32_t *qMiss, *qDiff, *qSpec;
  rt_hit* primary;
  __device__ __forceinline__ void done(bool finished, uint32_t i, bool hit, const HitRec& h) const {
    int kind = -1;
    uint32_t idx = 0;
    if (finished) {
      idx = perm ? perm[i] : i;
      hits[idx] = make_uint4(h.inst, h.prim, __float_as_uint(h.u), __float_as_uint(h.v));
      kind = !hit ? 0 : (scene.inst[h.inst].type == 0u ? 1 : 2);
      if (primary) {
        const uint32_t s = __float_as_uint(rayO[idx].w);
        uint32_t x, y;
        slot_to_pixel(U, s / U.spp, x, y);
        rt_hit out;
        out.instance = h.inst; out.primitive = h.prim; out.t = hit ? h.t : 0.0f; out.u = h.u; out.v = h.v;
        primary[((size_t)y * U.width + x) * U.spp + (s % U.spp)] = out;
      }
    }
    uint32_t dst = warp_append(kind == 0, &cnt[C_NMISS]);
    if (kind == 0) qMiss[dst] = idx;
    dst = warp_append(kind == 1, &cnt[C_NDIFF]);
    if (kind == 1) qDiff[dst] = idx;
    dst = warp_append(kind == 2, &cnt[C_NSPEC]);
    if (kind == 2) qSpec[dst] = idx;
  }
};

template <bool COUNT>
__global__ void __launch_bounds__(TRACE_THREADS, RT_TRACE_MINBLK)
    k_trace_wave(SceneDev scene, FrameUniforms U, uint32_t* __restrict__ cnt, const float4* __restrict__ rayO,
                 const float4* __restrict__ rayD, uint4* __restrict__ hits, uint32_t* __restrict__ qMiss,
                 uint32_t* __restrict__ qDiff, uint32_t* __restrict__ qSpec, rt_hit* __restrict__ primary,
                 unsigned long long* __restrict__ stats64, int refill_min, int tri_min, const uint32_t* __restrict__ perm) {
  __shared__ uint2 stack_sm[STACK_U2 * TRACE_THREADS];
  uint2 stack_loc[STACK_LOCAL];
  HybridStack st;
  st.sm = &stack_sm[threadIdx.x];
  st.loc = stack_loc;
  TravCounters tc{0ull, 0ull};
  const WaveTraceSource src{rayO, rayD, perm};
  const WaveTraceSink sink{scene, U, cnt, rayO, perm, hits, qMiss, qDiff, qSpec, primary};
  wave_trace<false, COUNT>(scene, cnt[C_NRAY], &cnt[C_WTRACE], st, src, sink, tc, refill_min, tri_min);
  if (COUNT) {
    atomicAdd(&stats64[0], tc.nodes);
    atomicAdd(&stats64[1], tc.tris);
  }
}

struct WaveShadowSource {
  const FrameUniforms& U;
  const float4* shadowQ;
  __device__ __forceinline__ void load(uint32_t i, f3& o, f3& d, float& tmax) const {
    const float4 q0 = shadowQ[3 * (size_t)i], q1 = shadowQ[3 * (size_t)i + 1];
    shadow_ray(U, mk3(q0.x, q0.y, q0.z), mk3(q1.x, q1.y, q1.z), o, d, tmax);
  }
};

template <bool COUNT>
__global__ void __launch_bounds__(TRACE_THREADS, RT_TRACE_MINBLK)
    k_shadow_wave(SceneDev scene, FrameUniforms U, uint32_t* __restrict__ cnt, const float4* __restrict__ shadowQ,
                  float4* __restrict__ sample, unsigned long long* __restrict__ stats64, int refill_min, int tri_min) {
  __shared__ uint2 stack_sm[STACK_U2 * TRACE_THREADS];
  uint2 stack_loc[STACK_LOCAL];
  HybridStack st;
  st.sm = &stack_sm[threadIdx.x];
  st.loc = stack_loc;
  TravCounters tc{0ull, 0ull};
  const WaveShadowSource src{U, shadowQ};
  const ShadowSink sink{scene, U, shadowQ, sample};
  wave_trace<true, COUNT>(scene, cnt[C_NSHADOW], &cnt[C_WSHADOW], st, src, sink, tc, refill_min, tri_min);
  if (COUNT) {
    atomicAdd(&stats64[2], tc.nodes);
    atomicAdd(&stats64[3], tc.tris);
  }
}

// Sort key of a secondary ray: 21-bit Morton cell of the origin inside the scene bounds, then
// a 9-bit Morton cell of the direction.  Rays of one warp then start close together and head
// the same way, so they visit the same nodes and finish after a similar number of steps.
__device__ __forceinline__ uint32_t spread3_10(uint32_t v) {
  v &= 0x3FFu;
  v = (v | (v << 16)) & 0x030000FFu;
  v = (v | (v << 8)) & 0x0300F00Fu;
  v = (v | (v << 4)) & 0x030C30C3u;
  v = (v | (v << 2)) & 0x09249249u;
  return v;
}
__device__ __forceinline__ uint32_t ray_sort_key(const SceneDev& s, f3 o, f3 d, int bits) {
  const float cx = fminf(fmaxf((o.x - s.slo[0]) * s.sinv[0], 0.0f), 0.9999f);
  const float cy = fminf(fmaxf((o.y - s.slo[1]) * s.sinv[1], 0.0f), 0.9999f);
  const float cz = fminf(fmaxf((o.z - s.slo[2]) * s.sinv[2], 0.0f), 0.9999f);
  const uint32_t om = (spread3_10((uint32_t)(cx * 128.0f)) << 2) | (spread3_10((uint32_t)(cy * 128.0f)) << 1) |
                      spread3_10((uint32_t)(cz * 128.0f));
  const float il = 1.0f / (fabsf(d.x) + fabsf(d.y) + fabsf(d.z) + 1e-30f);
  const uint32_t qx = (uint32_t)fminf(fmaf(d.x * il, 4.0f, 4.0f), 7.0f), qy = (uint32_t)fminf(fmaf(d.y * il, 4.0f, 4.0f), 7.0f),
                 qz = (uint32_t)fminf(fmaf(d.z * il, 4.0f, 4.0f), 7.0f);
  const uint32_t dm = (spread3_10(qx) << 2) | (spread3_10(qy) << 1) | spread3_10(qz);
  return ((om << 9) | (dm & 0x1FFu)) >> (30 - bits);  // keep the most significant `bits`
}

// ---- mirror / refraction: next bounce ----------------------------------------------------------
__global__ void __launch_bounds__(256)
    k_shade_spec(SceneDev scene, const uint32_t* __restrict__ cnt, uint32_t* __restrict__ cntNext,
                 const uint32_t* __restrict__ qSpec, const float4* __restrict__ rayO, const float4* __restrict__ rayD,
                 const uint4* __restrict__ hits, float4* __restrict__ nextO, float4* __restrict__ nextD,
                 uint32_t* __restrict__ skeys, uint32_t* __restrict__ svals, int sort_bits) {
  const uint32_t n = cnt[C_NSPEC];
  const uint32_t warps = (gridDim.x * blockDim.x) >> 5;
  const uint32_t w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  for (uint64_t base = (uint64_t)w * 32; base < n; base += (uint64_t)warps * 32) {
    const uint32_t i = (uint32_t)base + lane_id();
    const bool active = i < n;
    f3 o = mk3(0, 0, 0), d = mk3(0, 0, 0);
    float slot = 0.0f;
    if (active) {
      const uint32_t idx = qSpec[i];
      const uint4 h = hits[idx];
      const float4 ro = rayO[idx], rd = rayD[idx];
      slot = ro.w;
      o = mk3(ro.x, ro.y, ro.z);
      d = mk3(rd.x, rd.y, rd.z);
      f3 P, N;
      rchit(scene, h.x, h.y, __uint_as_float(h.z), __uint_as_float(h.w), P, N);
      scatter(scene.inst[h.x].type, P, N, o, d, scene.mat ? scene.mat[h.x].ni : 1.52f);
    }
    const uint32_t dst = warp_append(active, &cntNext[C_NRAY]);
    if (active) {
      nextO[dst] = make_float4(o.x, o.y, o.z, slot);
      nextD[dst] = make_float4(d.x, d.y, d.z, 0.0f);
      if (skeys) {
        skeys[dst] = ray_sort_key(scene, o, d, sort_bits);
        svals[dst] = dst;
      }
    }
  }
}

// ---- resolve: color /= samples; imageStore ---------------------------------------------------
// Three output layouts: full frame (row-major), compact shard (tile after tile, what the NCCL
// gather ships) and shard-in-frame, where a shard stores its tiles at their final position in a
// full frame that may sit in another GPU's memory.  For that one the threads walk the tiles row
// by row, so a warp's store is one 128-byte run of a frame row (NVLink moves it as one packet
// instead of four 32-byte ones); the samples are read through the 8x4 micro-block order they
// were accumulated in.
// `part`: 0 = every tile, 1 = only the tiles k_raygen finished (tile_live == 0: stored early, beside the
// traversal - for a frame in host memory that hides their trip over PCIe), 2 = only the others.
__global__ void __launch_bounds__(256) k_resolve(FrameUniforms U, uint32_t pix_begin, uint32_t npix, const float4* __restrict__ sample,
                                                 void* __restrict__ out, uint32_t format,
                                                 const uint32_t* __restrict__ tile_live, int part) {
  const uint32_t tile_px = U.tile_w * U.tile_h;
  for (uint32_t p = pix_begin + blockIdx.x * blockDim.x + threadIdx.x; p < npix; p += gridDim.x * blockDim.x) {
    if (part != 0 && (tile_live[p / tile_px] != 0u) == (part == 1)) continue;
    uint32_t x, y, pslot = p;
    bool valid;
    size_t dst;
    if (U.in_frame) {
      const uint32_t lt = p / tile_px, in = p % tile_px;
      const uint32_t tx = in % U.tile_w, ty = in / U.tile_w;
      pslot = lt * tile_px + ((ty / 4) * (U.tile_w / 8) + tx / 8) * 32 + (ty % 4) * 8 + (tx % 8);
      const uint32_t tile = lt * U.shard_count + U.shard_index;
      x = (tile % U.tiles_x) * U.tile_w + tx;
      y = (tile / U.tiles_x) * U.tile_h + ty;
      valid = x < U.width && y < U.height;
      if (!valid) continue;
      dst = (size_t)y * U.width + x;
    } else {
      valid = slot_to_pixel(U, p, x, y);
      if (U.shard_count <= 1) {
        if (!valid) continue;
        dst = (size_t)y * U.width + x;
      } else {  // compact shard layout: tile after tile, each tile row-major
        const uint32_t lt = p / tile_px;
        const uint32_t tile = lt * U.shard_count + U.shard_index;
        const uint32_t tx = x - (tile % U.tiles_x) * U.tile_w, ty = y - (tile / U.tiles_x) * U.tile_h;
        dst = (size_t)lt * tile_px + (size_t)ty * U.tile_w + tx;
      }
    }
    float c[4] = {0.f, 0.f, 0.f, 0.f};
    if (valid) {
      for (uint32_t i = 0; i < U.spp; i++) {
        const float4 s = sample[(size_t)pslot * U.spp + i];
        c[0] += s.x; c[1] += s.y; c[2] += s.z; c[3] += 1.0f;   // color += vec4(tmpColor, 1)
      }
      const float fs = (float)U.spp;
      c[0] = c[0] / fs; c[1] = c[1] / fs; c[2] = c[2] / fs; c[3] = c[3] / fs;
    }
    if (format == RT_FORMAT_RGBA8) {
      ((uint32_t*)out)[dst] = to_unorm8(c[0]) | (to_unorm8(c[1]) << 8) | (to_unorm8(c[2]) << 16) | (to_unorm8(c[3]) << 24);
    } else {
      ((float4*)out)[dst] = make_float4(c[0], c[1], c[2], c[3]);
    }
  }
}

// gathered shard buffers -> row-major frame
__global__ void __launch_bounds__(256) k_assemble(FrameUniforms U, uint32_t format, const void* __restrict__ gathered,
                                                  void* __restrict__ frame, uint32_t shard_pixels) {
  const uint32_t npix = U.width * U.height;
  const uint32_t tile_px = U.tile_w * U.tile_h;
  for (uint32_t p = blockIdx.x * blockDim.x + threadIdx.x; p < npix; p += gridDim.x * blockDim.x) {
    const uint32_t x = p % U.width, y = p / U.width;
    const uint32_t tile = (y / U.tile_h) * U.tiles_x + (x / U.tile_w);
    const uint32_t shard = tile % U.shard_count, lt = tile / U.shard_count;
    const size_t src = (size_t)shard * shard_pixels + (size_t)lt * tile_px + (size_t)(y % U.tile_h) * U.tile_w + (x % U.tile_w);
    if (format == RT_FORMAT_RGBA8) ((uint32_t*)frame)[p] = ((const uint32_t*)gathered)[src];
    else ((float4*)frame)[p] = ((const float4*)gathered)[src];
  }
}

// ---- arbitrary ray batches (rt_trace_rays) ------------------------------------------------------
template <int ACCEL, bool ANY, bool COUNT>
__global__ void __launch_bounds__(TRACE_THREADS)
    k_trace_batch(SceneDev scene, uint64_t n, const float* __restrict__ o, const float* __restrict__ d, float tmin,
                  float tmax, rt_hit* __restrict__ out, unsigned long long* __restrict__ stats64) {
  __shared__ uint2 stack_sm[STACK_U2 * TRACE_THREADS];
  typedef typename StackFor<ACCEL>::type Stack;
  typename Stack::Entry stack_loc[STACK_LOCAL];
  Stack st;
  st.init(stack_sm, stack_loc);
  TravCounters tc{0ull, 0ull};
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
    HitRec h;
    const bool hit = trace_scene_t<ACCEL, ANY, COUNT>(scene, mk3(o[3 * i], o[3 * i + 1], o[3 * i + 2]),
                                                      mk3(d[3 * i], d[3 * i + 1], d[3 * i + 2]), tmin, tmax, h, st, &tc);
    rt_hit r;
    if (ANY) { r.instance = hit ? 0u : RT_MISS; r.primitive = RT_MISS; r.t = 0.f; r.u = 0.f; r.v = 0.f; }
    else { r.instance = h.inst; r.primitive = h.prim; r.t = hit ? h.t : 0.0f; r.u = h.u; r.v = h.v; }
    out[i] = r;
  }
  if (COUNT) {
    atomicAdd(&stats64[0], tc.nodes);
    atomicAdd(&stats64[1], tc.tris);
  }
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
static const int RT_MAX_SUB = 8;
struct MigEntry;
struct FramePipeline {
  size_t cap = 0;  // ray slots the buffers hold
  float4 *rayO[2] = {nullptr, nullptr}, *rayD[2] = {nullptr, nullptr};
  uint4* hits = nullptr;
  uint32_t *qMiss = nullptr, *qDiff = nullptr, *qSpec = nullptr;
  float4* shadowQ = nullptr;
  float4* sample = nullptr;
  uint32_t* sortK[2] = {nullptr, nullptr};
  uint32_t* sortV[2] = {nullptr, nullptr};
  uint32_t* sortH = nullptr;
  uint32_t* cnt = nullptr;
  size_t cnt_blocks = 0;
  unsigned long long* stats64 = nullptr;
  std::vector<uint32_t> h_cnt;
  cudaEvent_t ev[2] = {nullptr, nullptr};
  // A frame is rendered as up to two independent sub-frames (halves of the tile list), each with its
  // own slice of the buffers, counters, streams and events: see pipeline_render.
  struct Sub {
    cudaStream_t own = nullptr;    // main stream of sub-frame 1 (sub-frame 0 runs on the caller's stream)
    cudaStream_t aux = nullptr;    // shadow rays of bounce j run here, beside bounce j+1's closest-hit rays
    std::vector<cudaEvent_t> kev;  // per bounce: trace begin/end, shadow begin/end
    std::vector<cudaEvent_t> sev;  // sev[2j]: shadow queue of bounce j complete, sev[2j+1]: its shadow rays done
    std::vector<cudaEvent_t> aev;  // aev[j]: h_alive[j] has landed
    uint32_t* h_alive = nullptr;   // pinned: rays entering bounce j, copied back without stalling the stream
    size_t alive_cap = 0;
    cudaEvent_t done = nullptr;
  } sub[RT_MAX_SUB];
  MigEntry* mig = nullptr;       // k_trace_shade_step: 32 entries per resident warp and sub-frame
  size_t mig_blocks = 0;
  // early resolve of the tiles k_raygen finished
  uint32_t* tile_live = nullptr;
  size_t tile_cap = 0;
  cudaStream_t early_st[RT_MAX_SUB] = {nullptr, nullptr, nullptr, nullptr};   // highest priority
  cudaEvent_t ray_done[RT_MAX_SUB] = {nullptr, nullptr, nullptr, nullptr};
  cudaEvent_t early_done[RT_MAX_SUB] = {nullptr, nullptr, nullptr, nullptr};
  cudaEvent_t start = nullptr;
  // what pipeline_frame_stats needs of the last frame (the event arithmetic is done on demand)
  int last_nsub = 0;
  uint32_t last_bounce[RT_MAX_SUB] = {0, 0, 0, 0};
  uint64_t last_launches = 0;
  std::vector<uint32_t> h_all;
  unsigned long long h64[16] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
  uint32_t h_flags = 0;          // g_trav_flags of the last frame / ray batch
  bool stats_pending = false;
};

FramePipeline* pipeline_create() { return new FramePipeline(); }

static void pipeline_free_buffers(FramePipeline* p) {
  for (int i = 0; i < 2; i++) { cudaFree(p->rayO[i]); cudaFree(p->rayD[i]); p->rayO[i] = p->rayD[i] = nullptr; }
  cudaFree(p->hits); cudaFree(p->qMiss); cudaFree(p->qDiff); cudaFree(p->qSpec); cudaFree(p->shadowQ); cudaFree(p->sample);
  for (int i = 0; i < 2; i++) { cudaFree(p->sortK[i]); cudaFree(p->sortV[i]); p->sortK[i] = nullptr; p->sortV[i] = nullptr; }
  cudaFree(p->sortH); p->sortH = nullptr;
  p->hits = nullptr; p->qMiss = p->qDiff = p->qSpec = nullptr; p->shadowQ = nullptr; p->sample = nullptr;
  p->cap = 0;
}

void pipeline_destroy(FramePipeline* p) {
  if (!p) return;
  pipeline_free_buffers(p);
  cudaFree(p->cnt); cudaFree(p->stats64);
  for (auto& e : p->ev) if (e) cudaEventDestroy(e);
  for (auto& sb : p->sub) {
    for (auto& e : sb.kev) cudaEventDestroy(e);
    for (auto& e : sb.aev) cudaEventDestroy(e);
    for (auto& e : sb.sev) cudaEventDestroy(e);
    if (sb.done) cudaEventDestroy(sb.done);
    if (sb.aux) cudaStreamDestroy(sb.aux);
    if (sb.own) cudaStreamDestroy(sb.own);
    if (sb.h_alive) cudaFreeHost(sb.h_alive);
  }
  if (p->start) cudaEventDestroy(p->start);
  cudaFree(p->tile_live);
  cudaFree(p->mig);
  for (auto& e : p->early_st) if (e) cudaStreamDestroy(e);
  for (auto& e : p->ray_done) if (e) cudaEventDestroy(e);
  for (auto& e : p->early_done) if (e) cudaEventDestroy(e);
  delete p;
}

static int pipeline_reserve(FramePipeline* p, size_t slots, size_t bounces, std::string& err) {
  if (slots > p->cap) {
    pipeline_free_buffers(p);
    for (int i = 0; i < 2; i++) {
      RT_CUDA_TRY(cudaMalloc(&p->rayO[i], slots * sizeof(float4)), err);
      RT_CUDA_TRY(cudaMalloc(&p->rayD[i], slots * sizeof(float4)), err);
    }
    RT_CUDA_TRY(cudaMalloc(&p->hits, slots * sizeof(uint4)), err);
    RT_CUDA_TRY(cudaMalloc(&p->qMiss, slots * sizeof(uint32_t)), err);
    RT_CUDA_TRY(cudaMalloc(&p->qDiff, slots * sizeof(uint32_t)), err);
    RT_CUDA_TRY(cudaMalloc(&p->qSpec, slots * sizeof(uint32_t)), err);
    RT_CUDA_TRY(cudaMalloc(&p->shadowQ, slots * 3 * sizeof(float4)), err);
    RT_CUDA_TRY(cudaMalloc(&p->sample, slots * sizeof(float4)), err);
    for (int i = 0; i < 2; i++) {
      RT_CUDA_TRY(cudaMalloc(&p->sortK[i], slots * sizeof(uint32_t)), err);
      RT_CUDA_TRY(cudaMalloc(&p->sortV[i], slots * sizeof(uint32_t)), err);
    }
    RT_CUDA_TRY(cudaMalloc(&p->sortH, sort_hist_words(slots) * sizeof(uint32_t)), err);
    p->cap = slots;
  }
  if (bounces + 2 > p->cnt_blocks) {   // one set of per-bounce counter blocks per sub-frame
    cudaFree(p->cnt);
    p->cnt = nullptr;
    p->cnt_blocks = bounces + 2;
    RT_CUDA_TRY(cudaMalloc(&p->cnt, RT_MAX_SUB * p->cnt_blocks * C_STRIDE * sizeof(uint32_t)), err);
  }
  if (!p->stats64) RT_CUDA_TRY(cudaMalloc(&p->stats64, 16 * sizeof(unsigned long long)), err);
  if (!p->ev[0]) {
    RT_CUDA_TRY(cudaEventCreate(&p->ev[0]), err);
    RT_CUDA_TRY(cudaEventCreate(&p->ev[1]), err);
    RT_CUDA_TRY(cudaEventCreateWithFlags(&p->start, cudaEventDisableTiming), err);
  }
  if (!p->early_st[0]) {
    int least = 0, greatest = 0;
    RT_CUDA_TRY(cudaDeviceGetStreamPriorityRange(&least, &greatest), err);
    for (auto& e : p->early_st) RT_CUDA_TRY(cudaStreamCreateWithPriority(&e, cudaStreamNonBlocking, greatest), err);
    for (auto& e : p->ray_done) RT_CUDA_TRY(cudaEventCreateWithFlags(&e, cudaEventDisableTiming), err);
    for (auto& e : p->early_done) RT_CUDA_TRY(cudaEventCreateWithFlags(&e, cudaEventDisableTiming), err);
  }
  for (auto& sb : p->sub) {
    if (!sb.aux) {
      RT_CUDA_TRY(cudaStreamCreateWithFlags(&sb.aux, cudaStreamNonBlocking), err);
      RT_CUDA_TRY(cudaStreamCreateWithFlags(&sb.own, cudaStreamNonBlocking), err);
      RT_CUDA_TRY(cudaEventCreateWithFlags(&sb.done, cudaEventDisableTiming), err);
    }
    while (sb.kev.size() < 4 * (bounces + 1)) {
      cudaEvent_t e;
      RT_CUDA_TRY(cudaEventCreate(&e), err);
      sb.kev.push_back(e);
    }
    if (bounces + 2 > sb.alive_cap) {
      if (sb.h_alive) cudaFreeHost(sb.h_alive);
      sb.h_alive = nullptr;
      sb.alive_cap = bounces + 2;
      RT_CUDA_TRY(cudaHostAlloc(&sb.h_alive, sb.alive_cap * sizeof(uint32_t), cudaHostAllocDefault), err);
    }
    while (sb.sev.size() < 2 * (bounces + 1)) {
      cudaEvent_t e;
      RT_CUDA_TRY(cudaEventCreateWithFlags(&e, cudaEventDisableTiming), err);
      sb.sev.push_back(e);
    }
    while (sb.aev.size() < bounces + 2) {
      cudaEvent_t e;
      RT_CUDA_TRY(cudaEventCreateWithFlags(&e, cudaEventDisableTiming), err);
      sb.aev.push_back(e);
    }
  }
  return RT_OK;
}

template <class K>
static int persistent_blocks(K kernel, int threads) {
  // RT_CARVEOUT = percent of the SM's 228 KB the traversal kernels ask for as shared memory (a hint; the rest is L1)
  static const int carve = getenv("RT_CARVEOUT") ? atoi(getenv("RT_CARVEOUT")) : -1;
  if (carve >= 0) cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, carve);
  int per_sm = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, threads, 0) != cudaSuccess || per_sm < 1) per_sm = 1;
  return per_sm * sm_count();
}

static int stride_blocks(uint64_t n, int threads) {
  const uint64_t want = (n + threads - 1) / threads;
  return (int)std::max<uint64_t>(1, std::min<uint64_t>(want, (uint64_t)sm_count() * 8));
}

#define RT_DISPATCH_ACCEL(accel, counting, CALL)                          \
  do {                                                                    \
    if ((accel) == ACCEL_WIDE) { if (counting) { CALL(ACCEL_WIDE, true); } else { CALL(ACCEL_WIDE, false); } }        \
    else if ((accel) == ACCEL_BVH2) { if (counting) { CALL(ACCEL_BVH2, true); } else { CALL(ACCEL_BVH2, false); } }   \
    else if ((accel) == ACCEL_BVH4) { if (counting) { CALL(ACCEL_BVH4, true); } else { CALL(ACCEL_BVH4, false); } }   \
    else { if (counting) { CALL(ACCEL_BRUTE, true); } else { CALL(ACCEL_BRUTE, false); } }                            \
  } while (0)

// One frame.  The tile list is cut into RT_SUBFRAMES consecutive parts that are rendered as independent
// sub-frames, each on its own stream (plus one stream each for their shadow rays).  A persistent traversal launch ends
// with a few warps still working on their last 32-ray batch while the rest of the GPU idles
// (≈ 0.1 ms per launch, 18 launches per frame); with several sub-frames in flight the blocks of another
// part's launch move in as soon as blocks retire, so those tails are filled.  Each sample
// owns its radiance slot, so the split does not change a single bit of the image.
// RT_SUBFRAMES=1 renders the frame in one piece.
namespace {
struct SubCtx {
  FramePipeline::Sub* sb;
  cudaStream_t st, sst;
  uint32_t slot_base, nslots;
  float4 *rayO[2], *rayD[2];
  uint4* hits;
  uint32_t *qMiss, *qDiff, *qSpec;
  float4* shadowQ;
  uint32_t* cnt;
  int cur = 0;
  bool dead = false;
  uint32_t last_bounce = 0;
  const uint32_t* perm = nullptr;
};
}  // namespace

int pipeline_render(FramePipeline* p, const FrameRequest& rq, cudaStream_t st, rt_stats* stats, std::string& err) {
  const FrameUniforms& U = rq.U;
  const uint64_t npix = (uint64_t)U.n_local_tiles * U.tile_w * U.tile_h;
  const uint64_t nslots = npix * U.spp;
  if (nslots == 0) return RT_OK;
  if (nslots > 0xFFFFFF00ull) { err = "frame too large: ray slots exceed 32 bits"; return RT_ERR_UNSUPPORTED; }
  int rc = pipeline_reserve(p, nslots, U.max_bounce, err);
  if (rc != RT_OK) return rc;
  const uint32_t accel = rq.scene.accel;
  const bool counting = rq.counting;
  // tuning knobs of the persistent traversal (defaults from the sweeps under profiles/)
  static const int sort_from = getenv("RT_SORT_FROM") ? atoi(getenv("RT_SORT_FROM")) : 1000;
  static const int sort_to = getenv("RT_SORT_TO") ? atoi(getenv("RT_SORT_TO")) : 1000;
  static const int sort_bits = std::min(30, std::max(8, getenv("RT_SORT_BITS") ? atoi(getenv("RT_SORT_BITS")) : 16));
  static const bool wave_on = getenv("RT_WAVE") ? atoi(getenv("RT_WAVE")) != 0 : false;
  static const int wave_refill = getenv("RT_WAVE_REFILL") ? atoi(getenv("RT_WAVE_REFILL")) : 12;
  static const int wave_tri = getenv("RT_WAVE_TRI") ? atoi(getenv("RT_WAVE_TRI")) : 6;
  // Shadow rays of bounce j only feed the radiance slots of paths that have ended, so they need
  // not finish before bounce j+1 starts: they go to a second stream and their last batches (and
  // the launches that carry a handful of rays) overlap the next closest-hit launch.  The shadow
  // queue is single-buffered: k_shade_diffuse of bounce j+1 waits for them.  RT_SHADOW_STREAM=0
  // puts everything back on one stream.
  static const bool shadow_async = getenv("RT_SHADOW_STREAM") ? atoi(getenv("RT_SHADOW_STREAM")) != 0 : true;
  static const int sub_env = getenv("RT_SUBFRAMES") ? atoi(getenv("RT_SUBFRAMES")) : 2;
  static const uint64_t sub_min_slots = getenv("RT_SUBFRAMES_MIN") ? strtoull(getenv("RT_SUBFRAMES_MIN"), nullptr, 10) : (6ull << 20);
  const bool may_sort = accel != ACCEL_BRUTE && sort_from <= (int)U.max_bounce && sort_from <= sort_to;
  const bool wave = accel == ACCEL_WIDE && wave_on && rq.scene.tlas == nullptr;
  // The bounce-fused path kernel (rt_path.cuh): binary BVH, instances looped (no top-level tree).
  // Opt-in (RT_FUSED=1): parity-green, but measured slower than the per-bounce wavefront launches on
  // every workload (C3 13.1 vs 8.9 ms, a 1/8 shard of C4 7.9 vs 5.1 ms): the lock-step phases cost 10.2
  // warp instructions per node visit against 6.6 (profiles/r02_path_kernel.txt).
  static const bool fused_env = getenv("RT_FUSED") ? atoi(getenv("RT_FUSED")) != 0 : false;
  static const int path_serve = getenv("RT_PATH_SERVE") ? atoi(getenv("RT_PATH_SERVE")) : 16;
  static const int path_leaf = getenv("RT_PATH_LEAF") ? atoi(getenv("RT_PATH_LEAF")) : 8;
  static const int path_left = getenv("RT_PATH_LEFT") ? atoi(getenv("RT_PATH_LEFT")) : 3;
  static const int path_chunk = getenv("RT_PATH_CHUNK") ? atoi(getenv("RT_PATH_CHUNK")) : 64;
  // Shading inside the closest-hit launch (k_trace_shade), the default; RT_FUSE_SHADE=0 puts the three shading kernels
  // back between the traversal launches.  Ray reordering and the warp-synchronous 8-wide variant keep the separate kernels.
  static const bool fuse_shade_env = getenv("RT_FUSE_SHADE") ? atoi(getenv("RT_FUSE_SHADE")) != 0 : true;
  const bool fuse_shade = fuse_shade_env && !may_sort && !(accel == ACCEL_WIDE && wave_on && rq.scene.tlas == nullptr);
  // RT_MIGRATE = T: a batch of the closest-hit launch ends when no more than T of its rays are still traversing; the
  // stragglers are set aside with their traversal state and resumed together (k_trace_shade_step)
  static const int migrate_F = getenv("RT_MIGRATE_FLUSH") ? std::max(1, atoi(getenv("RT_MIGRATE_FLUSH"))) : 32;   // resume when this many are set aside
  static const int migrate_env = getenv("RT_MIGRATE") ? std::max(0, std::min(16, atoi(getenv("RT_MIGRATE")))) : 0;
  // Binary layout, instances looped: the closest-hit launch takes the instances one after the other with the whole warp in
  // step (k_trace_shade_step<., false>: an instance no lane reaches is skipped by one vote, the lanes meet again after
  // each instance): C3 8.48 -> 8.31 ms, a 1/8 shard of C4 4.38 -> 4.28.  RT_INSTANCE_STEP=0: the generic k_trace_shade.
  static const bool step_env = getenv("RT_INSTANCE_STEP") ? atoi(getenv("RT_INSTANCE_STEP")) != 0 : true;
  const bool steppable = fuse_shade && accel == ACCEL_BVH2 && rq.scene.tlas == nullptr && rq.scene.n_inst <= 4u && RT_TRACE_THREADS == 32;
  const bool stepped = step_env && steppable;
  // RT_TLAS_STEP: scenes with a top-level tree - the warp meets between the top-level walk and the instance traversals
  static const bool tlas_step_env = getenv("RT_TLAS_STEP") ? atoi(getenv("RT_TLAS_STEP")) != 0 : true;
  const bool tlas_step = tlas_step_env && rq.scene.tlas != nullptr;
  // RT_SHADOW_STEP=1: the shadow launch steps through looped instances too (any layout).  Off: measured slower
  // (C3 7.78 -> 7.95 ms; the any-hit traversal parks leaves by votes of its own and gains nothing from the order)
  static const bool shadow_step_env = getenv("RT_SHADOW_STEP") ? atoi(getenv("RT_SHADOW_STEP")) != 0 : false;
  static const bool shadow_tlas_env = getenv("RT_SHADOW_TLAS_STEP") ? atoi(getenv("RT_SHADOW_TLAS_STEP")) != 0 : true;
  const bool shadow_step = !wave && ((shadow_step_env && rq.scene.tlas == nullptr && rq.scene.n_inst <= 4u && rq.scene.n_inst > 1u) ||
                                     (shadow_tlas_env && rq.scene.tlas != nullptr));
  const bool migrate = migrate_env > 0 && steppable;
  const int migrate_T = migrate ? std::min(migrate_env, 16 / (int)std::max(1u, rq.scene.n_inst)) : 0;   // the buffer holds 32
  if (migrate && !p->mig) {
    p->mig_blocks = (size_t)std::min(persistent_blocks(k_trace_shade_step<false, true>, 32), persistent_blocks(k_trace_shade_step<true, true>, 32));
    RT_CUDA_TRY(cudaMalloc(&p->mig, (size_t)RT_MAX_SUB * p->mig_blocks * 32 * sizeof(MigEntry)), err);
  }
  const bool fused = fused_env && accel == ACCEL_BVH2 && rq.scene.tlas == nullptr && !may_sort && rq.scene.n_inst <= 32u;
  // Deep bounce limits (the reference's default MAX_BOUNCE_COUNT is 63): from bounce RT_TAIL_FROM (9) on, the rest of every
  // path runs in ONE launch of the path kernel per sub-frame instead of two launches and a host look per bounce - by then
  // a few thousand rays are left, and what a bounce costs is its launches and the round trip of its ray count.
  static const uint32_t tail_from = getenv("RT_TAIL_FROM") ? (uint32_t)atoi(getenv("RT_TAIL_FROM")) : 9u;
  const bool tail = !fused && fuse_shade && tail_from > 0u && U.max_bounce >= tail_from && accel == ACCEL_BVH2 &&
                    rq.scene.tlas == nullptr && rq.scene.n_inst <= 32u;
  const int nsub = (sub_env >= 2 && nslots >= sub_min_slots && !may_sort && !wave)
                       ? (int)std::min<uint32_t>((uint32_t)std::min(sub_env, RT_MAX_SUB), U.n_local_tiles) : 1;

  SubCtx ctx[RT_MAX_SUB];
  {
    const uint64_t tile_slots = (uint64_t)U.tile_w * U.tile_h * U.spp;
    for (int k = 0; k < nsub; k++) {
      SubCtx& c = ctx[k];
      c.sb = &p->sub[k];
      c.st = k == 0 ? st : c.sb->own;
      c.sst = shadow_async ? c.sb->aux : c.st;
      const uint64_t off = (uint64_t)U.n_local_tiles * k / nsub * tile_slots;
      const uint64_t end = (uint64_t)U.n_local_tiles * (k + 1) / nsub * tile_slots;
      c.slot_base = (uint32_t)off;
      c.nslots = (uint32_t)(end - off);
      for (int i = 0; i < 2; i++) { c.rayO[i] = p->rayO[i] + off; c.rayD[i] = p->rayD[i] + off; }
      c.hits = p->hits + off; c.qMiss = p->qMiss + off; c.qDiff = p->qDiff + off; c.qSpec = p->qSpec + off;
      c.shadowQ = p->shadowQ + 3 * off;
      c.cnt = p->cnt + (size_t)k * p->cnt_blocks * C_STRIDE;
    }
  }
  uint64_t launches = 0;
  RT_CUDA_TRY(cudaEventRecord(p->ev[0], st), err);
  RT_CUDA_TRY(cudaMemsetAsync(p->cnt, 0, (size_t)nsub * p->cnt_blocks * C_STRIDE * sizeof(uint32_t), st), err);
  RT_CUDA_TRY(cudaMemsetAsync(p->stats64, 0, 16 * sizeof(unsigned long long), st), err);
  {
    void* fl = nullptr;
    RT_CUDA_TRY(cudaGetSymbolAddress(&fl, g_trav_flags), err);
    RT_CUDA_TRY(cudaMemsetAsync(fl, 0, sizeof(uint32_t), st), err);
  }
  // Tiles whose primary rays all miss the scene (4/5 of C3's) are final after k_raygen: every sub-frame resolves
  // its own on a high-priority side stream, one block per SM, launched before its first traversal launch so the
  // blocks are placed first; the stores (PCIe for a frame in host memory) then run beside the traversal.  The
  // other tiles are resolved at the end.  RT_EARLY_RESOLVE=0: one resolve at the end.
  static const bool early_env = getenv("RT_EARLY_RESOLVE") ? atoi(getenv("RT_EARLY_RESOLVE")) != 0 : true;
  const bool early = early_env && rq.d_out != nullptr;
  if (early) {
    if (U.n_local_tiles > p->tile_cap) {
      cudaFree(p->tile_live);
      p->tile_live = nullptr;
      p->tile_cap = 0;
      RT_CUDA_TRY(cudaMalloc(&p->tile_live, (size_t)U.n_local_tiles * sizeof(uint32_t)), err);
      p->tile_cap = U.n_local_tiles;
    }
    RT_CUDA_TRY(cudaMemsetAsync(p->tile_live, 0, (size_t)U.n_local_tiles * sizeof(uint32_t), st), err);
  }
  if (nsub > 1) {
    RT_CUDA_TRY(cudaEventRecord(p->start, st), err);
    for (int k = 1; k < nsub; k++) RT_CUDA_TRY(cudaStreamWaitEvent(ctx[k].st, p->start, 0), err);
  }
  for (int k = 0; k < nsub; k++) {
    SubCtx& c = ctx[k];
    k_raygen<<<stride_blocks(c.nslots, 256), 256, 0, c.st>>>(rq.scene, U, c.slot_base, c.nslots, c.rayO[0], c.rayD[0], p->sample,
                                                             c.cnt, rq.want_primary ? rq.d_primary : nullptr,
                                                             early ? p->tile_live : nullptr);
    launches++;
    if (early) {
      RT_CUDA_TRY(cudaEventRecord(p->ray_done[k], c.st), err);
      RT_CUDA_TRY(cudaStreamWaitEvent(p->early_st[k], p->ray_done[k], 0), err);
      const uint32_t pix0 = c.slot_base / U.spp, pix1 = (c.slot_base + c.nslots) / U.spp;
      k_resolve<<<sm_count(), 256, 0, p->early_st[k]>>>(U, pix0, pix1, p->sample, rq.d_out, rq.output_format, p->tile_live, 1);
      launches++;
      RT_CUDA_TRY(cudaEventRecord(p->early_done[k], p->early_st[k]), err);
    }
  }
  // one bounce of one sub-frame, enqueued without waiting for anything
  auto enqueue_bounce = [&](SubCtx& c, uint32_t j) -> int {
    FramePipeline::Sub& sb = *c.sb;
    cudaStream_t cs = c.st, sst = c.sst;
    const int cur = c.cur;
    uint32_t* cj = c.cnt + (size_t)j * C_STRIDE;
    uint32_t* cn = c.cnt + (size_t)(j + 1) * C_STRIDE;
    rt_hit* prim = (j == 0 && rq.want_primary) ? rq.d_primary : nullptr;
    const uint32_t* perm = c.perm;
    const uint32_t* sh_cnt0 = nullptr;   // the shadow queue is refilled from its start by every bounce
#define CALL_TRACE(A, C)                                                                                          \
  {                                                                                                               \
    static int blocks = 0;                                                                                        \
    if (!blocks) blocks = persistent_blocks(k_trace<A, C>, TRACE_THREADS);                                        \
    k_trace<A, C><<<blocks, TRACE_THREADS, 0, cs>>>(rq.scene, U, cj, c.rayO[cur], c.rayD[cur], c.hits, c.qMiss,   \
                                                    c.qDiff, c.qSpec, prim, p->stats64, perm);   \
  }
    RT_CUDA_TRY(cudaEventRecord(sb.kev[4 * j + 0], cs), err);
    if (wave) {
      static int wb[2] = {0, 0};
      if (!wb[0]) { wb[0] = persistent_blocks(k_trace_wave<false>, TRACE_THREADS); wb[1] = persistent_blocks(k_trace_wave<true>, TRACE_THREADS); }
      if (counting) k_trace_wave<true><<<wb[1], TRACE_THREADS, 0, cs>>>(rq.scene, U, cj, c.rayO[cur], c.rayD[cur], c.hits, c.qMiss, c.qDiff, c.qSpec, prim, p->stats64, wave_refill, wave_tri, perm);
      else k_trace_wave<false><<<wb[0], TRACE_THREADS, 0, cs>>>(rq.scene, U, cj, c.rayO[cur], c.rayD[cur], c.hits, c.qMiss, c.qDiff, c.qSpec, prim, p->stats64, wave_refill, wave_tri, perm);
    } else
    RT_DISPATCH_ACCEL(accel, counting, CALL_TRACE);
    RT_CUDA_TRY(cudaEventRecord(sb.kev[4 * j + 1], cs), err);
#undef CALL_TRACE
    k_shade_miss<<<sm_count() * 8, 256, 0, cs>>>(rq.scene, cj, c.qMiss, c.rayO[cur], c.rayD[cur], p->sample);
    if (shadow_async && j > 0) RT_CUDA_TRY(cudaStreamWaitEvent(cs, sb.sev[2 * (j - 1) + 1], 0), err);
    k_shade_diffuse<<<sm_count() * 8, 256, 0, cs>>>(rq.scene, cj, c.qDiff, c.rayO[cur], c.rayD[cur], c.hits, c.shadowQ, p->sample);
    if (shadow_async) {
      RT_CUDA_TRY(cudaEventRecord(sb.sev[2 * j], cs), err);
      RT_CUDA_TRY(cudaStreamWaitEvent(sst, sb.sev[2 * j], 0), err);
    }
#define CALL_SHADOW(A, C)                                                                                 \
  {                                                                                                       \
    static int blocks = 0;                                                                                \
    if (!blocks) blocks = std::min(persistent_blocks(k_shadow<A, C, false>, TRACE_THREADS), persistent_blocks(k_shadow<A, C, true>, TRACE_THREADS)); \
    if (shadow_step) k_shadow<A, C, true><<<blocks, TRACE_THREADS, 0, sst>>>(rq.scene, U, cj, c.shadowQ, p->sample, p->stats64, sh_cnt0, j); \
    else k_shadow<A, C, false><<<blocks, TRACE_THREADS, 0, sst>>>(rq.scene, U, cj, c.shadowQ, p->sample, p->stats64, sh_cnt0, j); \
  }
    RT_CUDA_TRY(cudaEventRecord(sb.kev[4 * j + 2], sst), err);
    if (wave) {
      static int wb[2] = {0, 0};
      if (!wb[0]) { wb[0] = persistent_blocks(k_shadow_wave<false>, TRACE_THREADS); wb[1] = persistent_blocks(k_shadow_wave<true>, TRACE_THREADS); }
      if (counting) k_shadow_wave<true><<<wb[1], TRACE_THREADS, 0, sst>>>(rq.scene, U, cj, c.shadowQ, p->sample, p->stats64, wave_refill, wave_tri);
      else k_shadow_wave<false><<<wb[0], TRACE_THREADS, 0, sst>>>(rq.scene, U, cj, c.shadowQ, p->sample, p->stats64, wave_refill, wave_tri);
    } else
    RT_DISPATCH_ACCEL(accel, counting, CALL_SHADOW);
    RT_CUDA_TRY(cudaEventRecord(sb.kev[4 * j + 3], sst), err);
    if (shadow_async) RT_CUDA_TRY(cudaEventRecord(sb.sev[2 * j + 1], sst), err);
#undef CALL_SHADOW
    launches += 4;
    c.last_bounce = j;
    if (j < U.max_bounce) {
      const bool sorting = may_sort && (int)(j + 1) >= sort_from && (int)(j + 1) <= sort_to;
      k_shade_spec<<<sm_count() * 8, 256, 0, cs>>>(rq.scene, cj, cn, c.qSpec, c.rayO[cur], c.rayD[cur], c.hits, c.rayO[cur ^ 1],
                                                   c.rayD[cur ^ 1], sorting ? p->sortK[0] : nullptr,
                                                   sorting ? p->sortV[0] : nullptr, sort_bits);
      launches++;
      c.cur ^= 1;
      c.perm = nullptr;
      if (sorting) {  // reorder the next bounce's rays; the count stays on the device (single sub-frame only)
        int rc2 = sort_pairs_u32(p->sortK[0], p->sortK[1], p->sortV[0], p->sortV[1], p->cap, cn + C_NRAY, sort_bits, p->sortH, cs, err);
        if (rc2 != RT_OK) return rc2;
        launches += sort_launches(sort_bits);
        c.perm = p->sortV[0];
      }
      // Deep bounce limits (MAX_BOUNCE_COUNT 63): stop as soon as no ray survives.  The count
      // travels to pinned memory behind the bounce; bounces 1-8 only look at counts that have
      // already landed (no stall), deeper ones wait for theirs.
      RT_CUDA_TRY(cudaMemcpyAsync(&sb.h_alive[j + 1], cn + C_NRAY, sizeof(uint32_t), cudaMemcpyDeviceToHost, cs), err);
      RT_CUDA_TRY(cudaEventRecord(sb.aev[j + 1], cs), err);
      if (j + 1 > 8) { RT_CUDA_TRY(cudaEventSynchronize(sb.aev[j + 1]), err); }
      for (uint32_t k = 1; k <= j + 1 && !c.dead; k++)
        c.dead = cudaEventQuery(sb.aev[k]) == cudaSuccess && sb.h_alive[k] == 0u;
    }
    return RT_OK;
  };
  // one bounce of one sub-frame with the shading done by the closest-hit launch itself (k_trace_shade): two launches,
  // and nothing on the sub-frame's main stream ever waits for a shadow launch
  auto enqueue_bounce_fused = [&](SubCtx& c, uint32_t j) -> int {
    FramePipeline::Sub& sb = *c.sb;
    cudaStream_t cs = c.st, sst = c.sst;
    const int cur = c.cur;
    uint32_t* cj = c.cnt + (size_t)j * C_STRIDE;
    uint32_t* cn = c.cnt + (size_t)(j + 1) * C_STRIDE;
    rt_hit* prim = (j == 0 && rq.want_primary) ? rq.d_primary : nullptr;
    const uint32_t* sh_cnt0 = c.cnt;
#define CALL_TRACE_SHADE(A, C)                                                                                        \
  {                                                                                                                   \
    static int blocks[2] = {0, 0};                                                                                    \
    if (!blocks[0]) { blocks[0] = persistent_blocks(k_trace_shade<A, C, false>, TRACE_THREADS);                       \
                      blocks[1] = persistent_blocks(k_trace_shade<A, C, true>, TRACE_THREADS); }                      \
    if (tlas_step) k_trace_shade<A, C, true><<<blocks[1], TRACE_THREADS, 0, cs>>>(rq.scene, U, c.cnt, j, c.rayO[cur], c.rayD[cur], c.rayO[cur ^ 1], \
                                                          c.rayD[cur ^ 1], c.shadowQ, p->sample, prim, p->stats64);    \
    else k_trace_shade<A, C, false><<<blocks[0], TRACE_THREADS, 0, cs>>>(rq.scene, U, c.cnt, j, c.rayO[cur], c.rayD[cur], c.rayO[cur ^ 1], \
                                                          c.rayD[cur ^ 1], c.shadowQ, p->sample, prim, p->stats64);    \
  }
#define CALL_SHADOW_F(A, C)                                                                               \
  {                                                                                                       \
    static int blocks = 0;                                                                                \
    if (!blocks) blocks = std::min(persistent_blocks(k_shadow<A, C, false>, TRACE_THREADS), persistent_blocks(k_shadow<A, C, true>, TRACE_THREADS)); \
    if (shadow_step) k_shadow<A, C, true><<<blocks, TRACE_THREADS, 0, sst>>>(rq.scene, U, cj, c.shadowQ, p->sample, p->stats64, sh_cnt0, j); \
    else k_shadow<A, C, false><<<blocks, TRACE_THREADS, 0, sst>>>(rq.scene, U, cj, c.shadowQ, p->sample, p->stats64, sh_cnt0, j); \
  }
    RT_CUDA_TRY(cudaEventRecord(sb.kev[4 * j + 0], cs), err);
#define CALL_STEPPED(C, CUT, BLOCKS, MG, T)                                                                                   \
  k_trace_shade_step<C, CUT><<<BLOCKS, 32, 0, cs>>>(rq.scene, U, c.cnt, j, c.rayO[cur], c.rayD[cur], c.rayO[cur ^ 1], c.rayD[cur ^ 1], \
                                                   c.shadowQ, p->sample, prim, p->stats64, MG, T, migrate_F)
    if (migrate) {
      MigEntry* mg = p->mig + (size_t)(&c - ctx) * p->mig_blocks * 32;
      if (counting) CALL_STEPPED(true, true, (int)p->mig_blocks, mg, migrate_T);
      else CALL_STEPPED(false, true, (int)p->mig_blocks, mg, migrate_T);
    } else if (stepped) {
      static int sb_blocks[2] = {0, 0};
      if (!sb_blocks[0]) { sb_blocks[0] = persistent_blocks(k_trace_shade_step<false, false>, 32); sb_blocks[1] = persistent_blocks(k_trace_shade_step<true, false>, 32); }
      if (counting) CALL_STEPPED(true, false, sb_blocks[1], nullptr, 0);
      else CALL_STEPPED(false, false, sb_blocks[0], nullptr, 0);
    } else
#undef CALL_STEPPED
    RT_DISPATCH_ACCEL(accel, counting, CALL_TRACE_SHADE);
    RT_CUDA_TRY(cudaEventRecord(sb.kev[4 * j + 1], cs), err);
    if (shadow_async) {
      RT_CUDA_TRY(cudaEventRecord(sb.sev[2 * j], cs), err);
      RT_CUDA_TRY(cudaStreamWaitEvent(sst, sb.sev[2 * j], 0), err);
    }
    RT_CUDA_TRY(cudaEventRecord(sb.kev[4 * j + 2], sst), err);
    RT_DISPATCH_ACCEL(accel, counting, CALL_SHADOW_F);
    RT_CUDA_TRY(cudaEventRecord(sb.kev[4 * j + 3], sst), err);
    if (shadow_async) RT_CUDA_TRY(cudaEventRecord(sb.sev[2 * j + 1], sst), err);
#undef CALL_TRACE_SHADE
#undef CALL_SHADOW_F
    launches += 2;
    c.last_bounce = j;
    if (j < U.max_bounce) {
      c.cur ^= 1;
      RT_CUDA_TRY(cudaMemcpyAsync(&sb.h_alive[j + 1], cn + C_NRAY, sizeof(uint32_t), cudaMemcpyDeviceToHost, cs), err);
      RT_CUDA_TRY(cudaEventRecord(sb.aev[j + 1], cs), err);
      if (j + 1 > 8 && !tail) { RT_CUDA_TRY(cudaEventSynchronize(sb.aev[j + 1]), err); }
      for (uint32_t k = 1; k <= j + 1 && !c.dead; k++)
        c.dead = cudaEventQuery(sb.aev[k]) == cudaSuccess && sb.h_alive[k] == 0u;
    }
    return RT_OK;
  };
  // the whole bounce loop of one sub-frame in one persistent launch
  auto enqueue_path = [&](SubCtx& c, uint32_t j0) -> int {
    FramePipeline::Sub& sb = *c.sb;
    PathArgs pa;
    pa.cnt = c.cnt + (size_t)j0 * C_STRIDE;
    pa.rayO = c.rayO[c.cur];
    pa.rayD = c.rayD[c.cur];
    pa.bounce0 = j0;
    pa.sample = p->sample;
    pa.primary = rq.want_primary ? rq.d_primary : nullptr;
    pa.stats64 = p->stats64;
    pa.t_serve = std::max(1, std::min(32, path_serve));
    pa.t_leaf = std::max(1, std::min(32, path_leaf));
    pa.t_left = std::max(1, std::min(32, path_left));
    pa.chunk = std::max(32, path_chunk) & ~31;
    static int blocks[2] = {0, 0};
    if (!blocks[0]) { blocks[0] = persistent_blocks(k_path<false>, 32); blocks[1] = persistent_blocks(k_path<true>, 32); }
    RT_CUDA_TRY(cudaEventRecord(sb.kev[4 * j0 + 0], c.st), err);
    if (counting) k_path<true><<<blocks[1], 32, 0, c.st>>>(rq.scene, U, pa);
    else k_path<false><<<blocks[0], 32, 0, c.st>>>(rq.scene, U, pa);
    RT_CUDA_TRY(cudaEventRecord(sb.kev[4 * j0 + 1], c.st), err);
    RT_CUDA_TRY(cudaEventRecord(sb.kev[4 * j0 + 2], c.st), err);   // no separate shadow launches: empty interval
    RT_CUDA_TRY(cudaEventRecord(sb.kev[4 * j0 + 3], c.st), err);
    // the shadow rays of the last wavefront bounce run beside this launch; the join below waits for sev[2 j0 + 1]
    if (shadow_async && j0 > 0) RT_CUDA_TRY(cudaStreamWaitEvent(c.st, sb.sev[2 * (j0 - 1) + 1], 0), err);
    if (shadow_async) RT_CUDA_TRY(cudaEventRecord(sb.sev[2 * j0 + 1], c.st), err);
    launches++;
    c.last_bounce = j0;
    return RT_OK;
  };
  // RT_FUSE_FROM = j: bounces < j run as wavefront launches (coherent rays, full batches), the rest of
  // every path in the fused kernel
  static const uint32_t fuse_from = getenv("RT_FUSE_FROM") ? (uint32_t)atoi(getenv("RT_FUSE_FROM")) : 0u;
  // RT_STAGGER = b: the sub-frames are started one after the other - sub-frame k traces its primary rays when sub-frame
  // k - 1 has finished bounce b - so that the thin late bounces of one sub-frame run beside the full early bounces of the next
  static const int stagger = getenv("RT_STAGGER") ? atoi(getenv("RT_STAGGER")) : -1;
  if (stagger >= 0 && fuse_shade && !fused && nsub > 1) {
    for (int k = 0; k < nsub; k++) {
      if (k > 0) {
        const uint32_t b = std::min<uint32_t>((uint32_t)stagger, ctx[k - 1].last_bounce);
        RT_CUDA_TRY(cudaStreamWaitEvent(ctx[k].st, ctx[k - 1].sb->kev[4 * b + 1], 0), err);
      }
      for (uint32_t j = 0; j <= U.max_bounce && !ctx[k].dead; j++) {
        rc = enqueue_bounce_fused(ctx[k], j);
        if (rc != RT_OK) return rc;
      }
    }
  } else
  for (uint32_t j = 0; j <= U.max_bounce; j++) {
    if ((fused && j >= fuse_from) || (tail && j >= tail_from)) {
      for (int k = 0; k < nsub; k++) {
        if (ctx[k].dead) continue;
        rc = enqueue_path(ctx[k], j);
        if (rc != RT_OK) return rc;
      }
      break;
    }
    bool any = false;
    for (int k = 0; k < nsub; k++) {
      if (ctx[k].dead) continue;
      rc = fuse_shade ? enqueue_bounce_fused(ctx[k], j) : enqueue_bounce(ctx[k], j);
      if (rc != RT_OK) return rc;
      any = any || !ctx[k].dead;
    }
    if (!any) break;
  }
  // join: every sub-frame's shadow rays, then sub-frame 1 into the caller's stream
  for (int k = nsub - 1; k >= 0; k--) {
    SubCtx& c = ctx[k];
    if (shadow_async) RT_CUDA_TRY(cudaStreamWaitEvent(c.st, c.sb->sev[2 * c.last_bounce + 1], 0), err);
    if (k > 0) {
      RT_CUDA_TRY(cudaEventRecord(c.sb->done, c.st), err);
      RT_CUDA_TRY(cudaStreamWaitEvent(st, c.sb->done, 0), err);
    }
  }
  if (rq.d_out) {
    k_resolve<<<stride_blocks(npix, 256), 256, 0, st>>>(U, 0u, (uint32_t)npix, p->sample, rq.d_out, rq.output_format,
                                                        p->tile_live, early ? 2 : 0);
    launches++;
    if (early)
      for (int k = 0; k < nsub; k++) RT_CUDA_TRY(cudaStreamWaitEvent(st, p->early_done[k], 0), err);
  }
  RT_CUDA_TRY(cudaEventRecord(p->ev[1], st), err);
  RT_CUDA_TRY(cudaGetLastError(), err);
  // counters to the host (also the synchronisation point of the frame); the stage times are
  // worked out from the events when somebody asks (pipeline_frame_stats)
  const size_t cnt_words = p->cnt_blocks * C_STRIDE;
  p->h_all.resize((size_t)nsub * cnt_words);
  RT_CUDA_TRY(cudaMemcpyAsync(p->h_all.data(), p->cnt, p->h_all.size() * sizeof(uint32_t), cudaMemcpyDeviceToHost, st), err);
  RT_CUDA_TRY(cudaMemcpyAsync(p->h64, p->stats64, sizeof p->h64, cudaMemcpyDeviceToHost, st), err);
  RT_CUDA_TRY(cudaMemcpyFromSymbolAsync(&p->h_flags, g_trav_flags, sizeof(uint32_t), 0, cudaMemcpyDeviceToHost, st), err);
  RT_CUDA_TRY(cudaStreamSynchronize(st), err);
  if (stats) stats->stack_overflow = p->h_flags & 1u;
  p->last_nsub = nsub;
  for (int k = 0; k < nsub; k++) p->last_bounce[k] = ctx[k].last_bounce;
  p->last_launches = launches;
  p->stats_pending = true;
  (void)stats;
  return RT_OK;
}

// Frame fields of rt_stats for the last pipeline_render (no-op if they are already up to date).
void pipeline_frame_stats(FramePipeline* p, rt_stats* stats) {
  if (!p || !stats || !p->stats_pending) return;
  p->stats_pending = false;
  const int nsub = p->last_nsub;
  const size_t cnt_words = p->cnt_blocks * C_STRIDE;
  const std::vector<uint32_t>& h_all = p->h_all;
  const unsigned long long* h64 = p->h64;
  p->h_cnt.assign(cnt_words, 0u);
  for (int k = 0; k < nsub; k++)
    for (size_t i = 0; i < cnt_words; i++) p->h_cnt[i] += h_all[(size_t)k * cnt_words + i];
  {
    float ms = 0.f;
    cudaEventElapsedTime(&ms, p->ev[0], p->ev[1]);
    stats->frame_ms = ms;
    uint32_t last_bounce = 0;
    for (int k = 0; k < nsub; k++) last_bounce = std::max(last_bounce, p->last_bounce[k]);
    // trace_ms = time during which a closest-hit launch was running, shadow_ms = the part of the
    // shadow launches not hidden behind one (they run on other streams), so trace_ms + shadow_ms is
    // the time a traversal kernel was running: unions of the launch intervals of all sub-frames
    auto union_ms = [](std::vector<std::pair<float, float>>& iv) {
      std::sort(iv.begin(), iv.end());
      double uni = 0.0;
      float cur_end = -1.0f;
      for (auto& x : iv) {
        const float s0 = std::max(x.first, cur_end);
        if (x.second > s0) { uni += x.second - s0; cur_end = x.second; }
      }
      return uni;
    };
    std::vector<std::pair<float, float>> iv_trace, iv_all;
    static const bool dbg = getenv("RT_DEBUG_BOUNCES") != nullptr;
    for (int k = 0; k < nsub; k++)
      for (uint32_t j = 0; j <= p->last_bounce[k]; j++) {
        const auto& kev = p->sub[k].kev;
        float a0 = 0.f, a1 = 0.f, b0 = 0.f, b1 = 0.f;
        cudaEventElapsedTime(&a0, p->ev[0], kev[4 * j + 0]);
        cudaEventElapsedTime(&a1, p->ev[0], kev[4 * j + 1]);
        cudaEventElapsedTime(&b0, p->ev[0], kev[4 * j + 2]);
        cudaEventElapsedTime(&b1, p->ev[0], kev[4 * j + 3]);
        iv_trace.push_back({a0, a1});
        iv_all.push_back({a0, a1});
        iv_all.push_back({b0, b1});
        if (dbg) {
          const uint32_t* c = &h_all[(size_t)k * cnt_words + (size_t)j * C_STRIDE];
          fprintf(stderr, "sub %d bounce %u: trace %.3f..%.3f ms shadow %.3f..%.3f ms  rays %u miss %u diff %u spec %u shadow %u\n", k,
                  j, a0, a1, b0, b1, c[C_NRAY], c[C_NMISS], c[C_NDIFF], c[C_NSPEC], c[C_NSHADOW]);
        }
      }
    const double tr = union_ms(iv_trace), uni = union_ms(iv_all);
    const double sh = std::max(0.0, uni - tr);
    stats->trace_ms = tr;
    stats->shadow_ms = sh;
    stats->shade_ms = std::max(0.0, (double)ms - tr - sh);
    // primary traces = rays traversed + rays the top level answered in k_raygen
    stats->rays_primary = (uint64_t)p->h_cnt[C_NRAY] + p->h_cnt[C_NCULL];
    stats->rays_secondary = stats->rays_shadow = stats->n_hit = 0;
    stats->n_miss = p->h_cnt[C_NCULL];
    stats->rays_top_level = p->h_cnt[C_NCULL];
    for (uint32_t j = 0; j <= last_bounce; j++) {
      const uint32_t* c = &p->h_cnt[(size_t)j * C_STRIDE];
      if (j > 0) stats->rays_secondary += c[C_NRAY];
      stats->rays_secondary += c[C_NSEC];   // the fused path kernel counts all bounces >= 1 here
      stats->rays_shadow += c[C_NSHADOW];
      stats->n_miss += c[C_NMISS];
      stats->n_hit += (uint64_t)c[C_NDIFF] + c[C_NSPEC];
    }
    stats->nodes_visited = h64[0] + h64[2];
    stats->tris_tested = h64[1] + h64[3];
    static const bool dbg_path = getenv("RT_DEBUG_PATH") != nullptr;
    if (dbg_path && h64[4] + h64[6])
      fprintf(stderr, "k_path node loop: %.1f M warp iterations at %.2f walking lanes while the queue had rays, %.1f M at %.2f after\n",
              h64[4] / 1e6, h64[4] ? (double)h64[5] / h64[4] : 0.0, h64[6] / 1e6, h64[6] ? (double)h64[7] / h64[6] : 0.0);
    if (dbg_path && h64[4])
      fprintf(stderr, "  lanes per iteration not walking (queue has rays): no path %.2f, ray done / between meshes %.2f, mesh done %.2f, "
              "only a parked leaf %.2f, second leaf (blocked) %.2f\n", (double)h64[8] / h64[4], (double)h64[9] / h64[4],
              (double)h64[10] / h64[4], (double)h64[11] / h64[4], (double)h64[12] / h64[4]);
    stats->kernel_launches = p->last_launches;
    stats->ray_state_bytes = 2 * sizeof(float4) + sizeof(uint4) + sizeof(uint32_t);
  }
}

int pipeline_trace_batch(FramePipeline* p, const SceneDev& scene, uint64_t n, const float* d_o, const float* d_d,
                         float tmin, float tmax, bool any_hit, bool counting, rt_hit* d_out, cudaStream_t st,
                         rt_stats* stats, std::string& err) {
  if (n == 0) return RT_OK;
  p->stats_pending = false;   // the stats below replace those of the last frame
  if (!p->stats64) RT_CUDA_TRY(cudaMalloc(&p->stats64, 16 * sizeof(unsigned long long)), err);
  RT_CUDA_TRY(cudaMemsetAsync(p->stats64, 0, 16 * sizeof(unsigned long long), st), err);
  {
    void* fl = nullptr;
    RT_CUDA_TRY(cudaGetSymbolAddress(&fl, g_trav_flags), err);
    RT_CUDA_TRY(cudaMemsetAsync(fl, 0, sizeof(uint32_t), st), err);
  }
  const int blocks = (int)std::max<uint64_t>(1, std::min<uint64_t>((n + TRACE_THREADS - 1) / TRACE_THREADS, (uint64_t)sm_count() * 8));
#define CALL_BATCH(A, C)                                                                                             \
  {                                                                                                                  \
    if (any_hit) k_trace_batch<A, true, C><<<blocks, TRACE_THREADS, 0, st>>>(scene, n, d_o, d_d, tmin, tmax, d_out, p->stats64); \
    else k_trace_batch<A, false, C><<<blocks, TRACE_THREADS, 0, st>>>(scene, n, d_o, d_d, tmin, tmax, d_out, p->stats64);        \
  }
  RT_DISPATCH_ACCEL(scene.accel, counting, CALL_BATCH);
#undef CALL_BATCH
  RT_CUDA_TRY(cudaGetLastError(), err);
  unsigned long long h64[16];
  RT_CUDA_TRY(cudaMemcpyAsync(h64, p->stats64, sizeof h64, cudaMemcpyDeviceToHost, st), err);
  RT_CUDA_TRY(cudaMemcpyFromSymbolAsync(&p->h_flags, g_trav_flags, sizeof(uint32_t), 0, cudaMemcpyDeviceToHost, st), err);
  RT_CUDA_TRY(cudaStreamSynchronize(st), err);
  if (stats) {
    stats->stack_overflow = p->h_flags & 1u;
    stats->nodes_visited = h64[0];
    stats->tris_tested = h64[1];
    stats->kernel_launches = 1;
  }
  return RT_OK;
}

// FP32 FMA peak of the CUDA cores (denominator of the FLOP term of the traversal roofline):
// 8 independent FMA chains per thread, enough resident warps to saturate the FMA pipes.
__global__ void __launch_bounds__(256) k_fma_peak(float* out, int iters, float a, float b) {
  float x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
  for (int i = 0; i < iters; i++) {
#pragma unroll
    for (int k = 0; k < 8; k++) {
      x0 = fmaf(x0, a, b); x1 = fmaf(x1, a, b); x2 = fmaf(x2, a, b); x3 = fmaf(x3, a, b);
      x4 = fmaf(x4, a, b); x5 = fmaf(x5, a, b); x6 = fmaf(x6, a, b); x7 = fmaf(x7, a, b);
    }
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = ((x0 + x1) + (x2 + x3)) + ((x4 + x5) + (x6 + x7));
}

int measure_fp32_tflops(double* tflops, std::string& err) {
  const int blocks = sm_count() * 8, threads = 256, iters = 4096;
  float* d = nullptr;
  RT_CUDA_TRY(cudaMalloc(&d, (size_t)blocks * threads * sizeof(float)), err);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  double best = 0.0;
  for (int rep = 0; rep < 5; rep++) {
    cudaEventRecord(e0);
    k_fma_peak<<<blocks, threads>>>(d, iters, 0.999f, 0.001f);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    const double fl = 2.0 * 64.0 * iters * (double)blocks * threads;
    if (rep > 0 && ms > 0.f) best = std::max(best, fl / (ms * 1e-3) / 1e12);
  }
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  cudaFree(d);
  RT_CUDA_TRY(cudaGetLastError(), err);
  *tflops = best;
  return RT_OK;
}

int assemble_shards(const FrameUniforms& U, uint32_t format, const void* gathered, void* frame, cudaStream_t st,
                    std::string& err) {
  const uint32_t tiles = U.tiles_x * U.tiles_y;
  const uint32_t per_shard_tiles = (tiles + U.shard_count - 1) / U.shard_count;
  const uint32_t shard_pixels = per_shard_tiles * U.tile_w * U.tile_h;
  k_assemble<<<stride_blocks((uint64_t)U.width * U.height, 256), 256, 0, st>>>(U, format, gathered, frame, shard_pixels);
  RT_CUDA_TRY(cudaGetLastError(), err);
  return RT_OK;
}

}  // namespace rt
