// rt_sort.cu — one-sweep LSD radix sort of (key, value) pairs for sm_100a (Adinets & Merrill,
// "Onesweep: a faster least significant digit radix sort for GPUs", 2022).
//
// Used by the LBVH build (63-bit Morton keys -> triangle ids) and by the ray reordering between
// bounces (30-bit origin/direction cells -> ray ids).  HBM-bound integer work, no tensor cores:
//   k_os_hist   one read of the keys builds the digit histograms of ALL passes (8 B / key)
//   k_os_scan   256-bin exclusive scans -> first output position of every digit, per pass
//   k_os_pass   per 8-bit digit: each 4096-key tile ranks its keys (stable, warp match_any),
//               publishes its per-digit counts, finds the counts of all earlier tiles with a
//               decoupled look-back over the tile status words, reorders the tile in shared memory
//               and writes it out in output order (coalesced runs per digit): keys and values
//               are read once and written once per pass
//               (24 B / pair / pass for 64-bit keys; the three-kernel sort it replaces read the
//               keys twice and needed 3 launches per pass).
// Tiles are handed out by an atomic counter, so a tile only ever waits for tiles that are
// already running: the look-back cannot deadlock, whatever the grid size.
// The element count may live in device memory (count_dev): the ray sort runs without a host
// round trip, with a grid sized for the upper bound.
#include <algorithm>
#include <cstdio>

#include "rt_internal.h"

namespace rt {

constexpr int OS_THREADS = 256;
constexpr int OS_ITEMS = 16;
constexpr int OS_TILE = OS_THREADS * OS_ITEMS;
constexpr int OS_WARPS = OS_THREADS / 32;
constexpr int OS_MAX_PASSES = 8;
constexpr uint32_t OS_FLAG_AGG = 1u << 30, OS_FLAG_PREFIX = 2u << 30, OS_VALUE_MASK = (1u << 30) - 1u;

// scratch layout (uint32 words): [0, 8*256) histograms / digit starts, then 16 control words
// (tile counters per pass), then the tile status words of all passes
static size_t os_tiles(uint64_t n) { return (size_t)((n + OS_TILE - 1) / OS_TILE); }
size_t sort_hist_words(uint64_t n) { return (size_t)OS_MAX_PASSES * 256 + 16 + (size_t)OS_MAX_PASSES * 256 * std::max<size_t>(os_tiles(n), 1); }

template <class K>
__global__ void __launch_bounds__(OS_THREADS) k_os_hist(const K* __restrict__ keys, uint64_t n_host, const uint32_t* __restrict__ count_dev,
                                                        int passes, uint32_t* __restrict__ hist) {
  __shared__ uint32_t sh[OS_MAX_PASSES * 256];
  const uint64_t n = count_dev ? (uint64_t)*count_dev : n_host;
  for (int i = threadIdx.x; i < passes * 256; i += OS_THREADS) sh[i] = 0;
  __syncthreads();
  for (uint64_t i = (uint64_t)blockIdx.x * OS_THREADS + threadIdx.x; i < n; i += (uint64_t)gridDim.x * OS_THREADS) {
    const K k = keys[i];
#pragma unroll
    for (int p = 0; p < OS_MAX_PASSES; p++)
      if (p < passes) atomicAdd(&sh[p * 256 + ((uint32_t)(k >> (8 * p)) & 0xFFu)], 1u);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < passes * 256; i += OS_THREADS)
    if (sh[i]) atomicAdd(&hist[i], sh[i]);
}

// one block per pass: counts -> exclusive prefix (first output position of each digit)
__global__ void __launch_bounds__(256) k_os_scan(uint32_t* __restrict__ hist) {
  __shared__ uint32_t wsum[8];
  uint32_t* h = hist + blockIdx.x * 256;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const uint32_t v = h[threadIdx.x];
  uint32_t x = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t y = __shfl_up_sync(0xFFFFFFFFu, x, o);
    if (lane >= o) x += y;
  }
  if (lane == 31) wsum[wid] = x;
  __syncthreads();
  uint32_t wb = 0;
  for (int w = 0; w < wid; w++) wb += wsum[w];
  h[threadIdx.x] = wb + x - v;
}

__device__ __forceinline__ uint32_t ld_volatile_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(p));
  return v;
}

// One digit pass.  status: [tile][256] words of this pass, zero on entry; ctrl[0]: tile counter.
template <class K>
__global__ void __launch_bounds__(OS_THREADS, 2)
    k_os_pass(const K* __restrict__ kin, const uint32_t* __restrict__ vin, K* __restrict__ kout, uint32_t* __restrict__ vout,
              uint64_t n_host, const uint32_t* __restrict__ count_dev, int shift, const uint32_t* __restrict__ digit_start,
              uint32_t* __restrict__ status, uint32_t* __restrict__ ctrl) {
  __shared__ uint32_t warp_cnt[OS_WARPS][256];
  __shared__ uint32_t digit_base[256];     // global position of the tile's first key of digit d, minus its position in the tile
  __shared__ uint32_t wsum[OS_WARPS];
  __shared__ K stage[OS_TILE];             // the tile in output order: keys first, then (same storage) the values
  __shared__ uint32_t s_tile;
  const uint64_t n = count_dev ? (uint64_t)*count_dev : n_host;
  const uint32_t num_tiles = (uint32_t)((n + OS_TILE - 1) / OS_TILE);
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const uint32_t lt_mask = (1u << lane) - 1u;
  while (true) {
    if (threadIdx.x == 0) s_tile = atomicAdd(ctrl, 1u);
#pragma unroll
    for (int w = 0; w < OS_WARPS; w++) warp_cnt[w][threadIdx.x] = 0;
    __syncthreads();
    const uint32_t tile = s_tile;
    if (tile >= num_tiles) break;
    K key[OS_ITEMS];
    uint32_t val[OS_ITEMS];
    uint32_t rank[OS_ITEMS];
    const uint64_t tile_base = (uint64_t)tile * OS_TILE;
    const uint32_t nvalid = (uint32_t)std::min<uint64_t>((uint64_t)OS_TILE, n - tile_base);
    const uint64_t base = tile_base + (uint64_t)wid * (OS_ITEMS * 32) + lane;
#pragma unroll
    for (int r = 0; r < OS_ITEMS; r++) {
      const uint64_t p = base + (uint64_t)r * 32;
      const bool valid = p < n;
      key[r] = valid ? kin[p] : (K)0;
      val[r] = valid ? vin[p] : 0u;
    }
    // stable rank of every key among the keys of its warp with the same digit
#pragma unroll
    for (int r = 0; r < OS_ITEMS; r++) {
      const bool valid = base + (uint64_t)r * 32 < n;
      const uint32_t d = valid ? ((uint32_t)(key[r] >> shift) & 0xFFu) : 0x100u;
      const uint32_t peers = __match_any_sync(0xFFFFFFFFu, d);
      const uint32_t prior = valid ? warp_cnt[wid][d] : 0u;
      __syncwarp();
      const uint32_t before = __popc(peers & lt_mask);
      if (valid && before == 0) warp_cnt[wid][d] = prior + __popc(peers);
      __syncwarp();
      rank[r] = prior + before;
    }
    __syncthreads();
    {  // thread d owns digit d: offsets of the warps inside the tile, position of the digit inside the
       // tile (block scan), then the look-back
      const uint32_t d = threadIdx.x;
      uint32_t run = 0;
#pragma unroll
      for (int w = 0; w < OS_WARPS; w++) {
        const uint32_t c = warp_cnt[w][d];
        warp_cnt[w][d] = run;
        run += c;
      }
      uint32_t x = run;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t y = __shfl_up_sync(0xFFFFFFFFu, x, o);
        if (lane >= o) x += y;
      }
      if (lane == 31) wsum[wid] = x;
      uint32_t* my = status + (size_t)tile * 256 + d;
      uint32_t excl = 0;
      if (tile == 0) {
        atomicExch(my, OS_FLAG_PREFIX | run);
      } else {
        atomicExch(my, OS_FLAG_AGG | run);
        int t = (int)tile - 1;
        while (true) {
          uint32_t s;
          do { s = ld_volatile_u32(status + (size_t)t * 256 + d); } while ((s >> 30) == 0u);
          excl += s & OS_VALUE_MASK;
          if (s & OS_FLAG_PREFIX) break;
          t--;
        }
        atomicExch(my, OS_FLAG_PREFIX | (excl + run));
      }
      __syncthreads();
      uint32_t in_tile = x - run;   // keys of smaller digits in this tile
      for (int w = 0; w < wid; w++) in_tile += wsum[w];
#pragma unroll
      for (int w = 0; w < OS_WARPS; w++) warp_cnt[w][d] += in_tile;   // now: position in the tile of warp w's first key of digit d
      digit_base[d] = digit_start[d] + excl - in_tile;
    }
    __syncthreads();
    // Reorder the tile in shared memory, then write it out in output order: consecutive threads write
    // consecutive addresses of a digit's run (the direct scatter wrote 32 different runs per warp store).
    uint32_t pos[OS_ITEMS];
#pragma unroll
    for (int r = 0; r < OS_ITEMS; r++) {
      const bool valid = base + (uint64_t)r * 32 < n;
      const uint32_t d = (uint32_t)(key[r] >> shift) & 0xFFu;
      pos[r] = valid ? warp_cnt[wid][d] + rank[r] : 0xFFFFFFFFu;
      if (valid) stage[pos[r]] = key[r];
    }
    __syncthreads();
    uint32_t dst[OS_ITEMS];
#pragma unroll
    for (int r = 0; r < OS_ITEMS; r++) {
      const uint32_t i = threadIdx.x + (uint32_t)r * OS_THREADS;
      dst[r] = 0xFFFFFFFFu;
      if (i < nvalid) {
        const K k = stage[i];
        dst[r] = digit_base[(uint32_t)(k >> shift) & 0xFFu] + i;
        kout[dst[r]] = k;
      }
    }
    __syncthreads();
    uint32_t* stage_v = reinterpret_cast<uint32_t*>(stage);
#pragma unroll
    for (int r = 0; r < OS_ITEMS; r++)
      if (pos[r] != 0xFFFFFFFFu) stage_v[pos[r]] = val[r];
    __syncthreads();
#pragma unroll
    for (int r = 0; r < OS_ITEMS; r++)
      if (dst[r] != 0xFFFFFFFFu) vout[dst[r]] = stage_v[threadIdx.x + (uint32_t)r * OS_THREADS];
    __syncthreads();
  }
}

// zero the tile status words actually needed when the element count lives on the device
__global__ void __launch_bounds__(256) k_os_clear(uint32_t* __restrict__ status, const uint32_t* __restrict__ count_dev, int passes,
                                                  size_t tiles_max) {
  const size_t tiles = ((size_t)*count_dev + OS_TILE - 1) / OS_TILE;
  for (int p = 0; p < passes; p++) {
    uint32_t* s = status + (size_t)p * 256 * tiles_max;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < tiles * 256; i += (size_t)gridDim.x * blockDim.x) s[i] = 0u;
  }
}

template <class K>
static int sort_pairs_t(K* keys, K* keys_alt, uint32_t* vals, uint32_t* vals_alt, uint64_t n_max, const uint32_t* count_dev,
                        int key_bits, uint32_t* scratch, cudaStream_t stream, std::string& err) {
  if (n_max == 0) return RT_OK;
  if (n_max > OS_VALUE_MASK) { err = "sort_pairs: more than 2^30 elements"; return RT_ERR_UNSUPPORTED; }
  int passes = (key_bits + 7) / 8;
  if (passes & 1) passes++;  // even number of passes: the result lands in keys/vals
  if (passes > OS_MAX_PASSES) { err = "sort_pairs: key too wide"; return RT_ERR_UNSUPPORTED; }
  const size_t tiles = os_tiles(n_max);
  uint32_t* hist = scratch;
  uint32_t* ctrl = scratch + OS_MAX_PASSES * 256;
  uint32_t* status = ctrl + 16;
  const int grid = (int)std::min<size_t>(tiles, (size_t)sm_count() * 4);
  if (count_dev) {
    RT_CUDA_TRY(cudaMemsetAsync(scratch, 0, ((size_t)OS_MAX_PASSES * 256 + 16) * sizeof(uint32_t), stream), err);
    k_os_clear<<<grid, 256, 0, stream>>>(status, count_dev, passes, tiles);
  } else {
    RT_CUDA_TRY(cudaMemsetAsync(scratch, 0, ((size_t)OS_MAX_PASSES * 256 + 16 + (size_t)passes * 256 * tiles) * sizeof(uint32_t), stream), err);
  }
  k_os_hist<K><<<grid, OS_THREADS, 0, stream>>>(keys, n_max, count_dev, passes, hist);
  k_os_scan<<<passes, 256, 0, stream>>>(hist);
  K* kin = keys; K* kout = keys_alt;
  uint32_t* vin = vals; uint32_t* vout = vals_alt;
  for (int p = 0; p < passes; p++) {
    k_os_pass<K><<<grid, OS_THREADS, 0, stream>>>(kin, vin, kout, vout, n_max, count_dev, 8 * p, hist + p * 256,
                                                 status + (size_t)p * 256 * tiles, ctrl + p);
    std::swap(kin, kout);
    std::swap(vin, vout);
  }
  RT_CUDA_TRY(cudaGetLastError(), err);
  return RT_OK;
}

int sort_pairs(uint64_t* keys, uint64_t* keys_alt, uint32_t* vals, uint32_t* vals_alt, uint64_t n, int key_bits,
               uint32_t* scratch, cudaStream_t stream, std::string& err) {
  return sort_pairs_t<uint64_t>(keys, keys_alt, vals, vals_alt, n, nullptr, key_bits, scratch, stream, err);
}

int sort_pairs_u32(uint32_t* keys, uint32_t* keys_alt, uint32_t* vals, uint32_t* vals_alt, uint64_t n_max,
                   const uint32_t* count_dev, int key_bits, uint32_t* scratch, cudaStream_t stream, std::string& err) {
  return sort_pairs_t<uint32_t>(keys, keys_alt, vals, vals_alt, n_max, count_dev, key_bits, scratch, stream, err);
}

int sort_launches(int key_bits) { return 3 + ((((key_bits + 7) / 8) + 1) & ~1); }

}  // namespace rt

// ---- rt_sort_pairs_noalloc: the same sort for callers that must not allocate ---------------------------
// (no counterpart in the reference, see include/rt_b200.h)
// The sort's scratch - the ping-pong copies of the keys and values and the histogram / tile-status words -
// lives in ONE file-scope device array sized for RT_SORT_NOALLOC_MAX_PAIRS pairs instead of cudaMalloc'd
// buffers.  Every GPU that loads the library holds the array until the process exits; every call shares it.
namespace rt {

constexpr uint64_t NOALLOC_MAX = RT_SORT_NOALLOC_MAX_PAIRS;
constexpr size_t NOALLOC_KEYS_BYTES = (size_t)NOALLOC_MAX * sizeof(uint64_t);          // keys_alt
constexpr size_t NOALLOC_VALS_BYTES = (size_t)NOALLOC_MAX * sizeof(uint32_t);          // vals_alt
constexpr size_t NOALLOC_HIST_WORDS =
    (size_t)OS_MAX_PASSES * 256 + 16 + (size_t)OS_MAX_PASSES * 256 * ((NOALLOC_MAX + OS_TILE - 1) / OS_TILE);   // = sort_hist_words(max)
constexpr size_t NOALLOC_BYTES = NOALLOC_KEYS_BYTES + NOALLOC_VALS_BYTES + NOALLOC_HIST_WORDS * sizeof(uint32_t);
static_assert(NOALLOC_KEYS_BYTES % 256 == 0 && NOALLOC_VALS_BYTES % 256 == 0, "pieces of the scratch array stay 256-byte aligned");

static __device__ __align__(256) unsigned char g_sort_scratch[NOALLOC_BYTES];   // 14.0 MiB for 2^20 pairs

// In host code `g_sort_scratch` is only a stand-in for the symbol, not a device address: the address is asked
// for once per device, with that device current, and kept here.
static const int NOALLOC_MAX_DEVICES = 64;
static unsigned char* g_sort_scratch_addr[NOALLOC_MAX_DEVICES] = {nullptr};

extern thread_local std::string g_err;   // rt_api.cu: message behind rt_last_error()

}  // namespace rt

extern "C" int rt_sort_pairs_noalloc(int device_id, uint64_t n, uint64_t* d_keys, uint32_t* d_vals, void* stream) {
  using namespace rt;
  if (n > NOALLOC_MAX) {
    g_err = "rt_sort_pairs_noalloc: " + std::to_string(n) + " pairs exceed the fixed scratch of RT_SORT_NOALLOC_MAX_PAIRS = " +
            std::to_string(NOALLOC_MAX) + " pairs (use the allocating sort for larger inputs)";
    return RT_ERR_UNSUPPORTED;
  }
  if (n && (!d_keys || !d_vals)) { g_err = "rt_sort_pairs_noalloc: null device pointer"; return RT_ERR_INVALID_ARGUMENT; }
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1) {
    cudaGetLastError();
    g_err = "no CUDA device: librt_b200 has no CPU fallback";
    return RT_ERR_NO_DEVICE;
  }
  if (device_id < 0) device_id = 0;
  if (device_id >= ndev || device_id >= NOALLOC_MAX_DEVICES) { g_err = "rt_sort_pairs_noalloc: bad device_id"; return RT_ERR_INVALID_ARGUMENT; }
  if (n == 0) return RT_OK;
  RT_CUDA_TRY(cudaSetDevice(device_id), g_err);
  if (!g_sort_scratch_addr[device_id]) {   // set-up, once per device, that device current
    unsigned char* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress((void**)&p, g_sort_scratch);
    if (e != cudaSuccess || !p) {
      cudaGetLastError();
      g_err = std::string("rt_sort_pairs_noalloc: cudaGetSymbolAddress(g_sort_scratch): ") + cudaGetErrorString(e);
      return RT_ERR_CUDA;
    }
    g_sort_scratch_addr[device_id] = p;
  }
  unsigned char* base = g_sort_scratch_addr[device_id];
  uint64_t* keys_alt = (uint64_t*)base;
  uint32_t* vals_alt = (uint32_t*)(base + NOALLOC_KEYS_BYTES);
  uint32_t* hist = (uint32_t*)(base + NOALLOC_KEYS_BYTES + NOALLOC_VALS_BYTES);
  // sort_pairs clears the histogram, control and tile-status words it uses at the start of every call (the array
  // keeps the last call's contents); keys_alt / vals_alt are fully written before they are read.  64 key bits =
  // 8 passes, an even number: the result lands in d_keys / d_vals.  Nothing here allocates or synchronises.
  return sort_pairs(d_keys, keys_alt, d_vals, vals_alt, n, 64, hist, (cudaStream_t)stream, g_err);
}
