// rt_internal.h — host-side declarations shared by the translation units of librt_b200.so.
#pragma once
#include <cuda_runtime.h>

#include <string>
#include <vector>

#include "../../include/rt_b200.h"
#include "rt_shade.cuh"

namespace rt {

#define RT_CUDA_TRY(expr, err)                                                              \
  do {                                                                                      \
    cudaError_t e_ = (expr);                                                                \
    if (e_ != cudaSuccess) {                                                                \
      (err) = std::string(#expr) + ": " + cudaGetErrorString(e_);                           \
      return e_ == cudaErrorMemoryAllocation ? RT_ERR_OUT_OF_MEMORY : RT_ERR_CUDA;          \
    }                                                                                       \
  } while (0)

// One built bottom-level structure, owning its device memory.
struct BlasHost {
  BlasDev dev{};
  float lo[3] = {0, 0, 0}, hi[3] = {0, 0, 0};   // padded object-space bounds
  float* d_verts = nullptr;
  uint32_t* d_idx = nullptr;
  void* d_nodes = nullptr;
  void* d_tris = nullptr;
  uint64_t node_count = 0, bytes = 0;
  uint32_t depth = 0;                          // levels of binary nodes above the deepest leaf
};

// Builds the acceleration structure of one mesh already resident on the device.
int build_blas(BlasHost& b, uint32_t nvert, uint32_t ntri, int accel, bool refine, cudaStream_t stream, float* kernel_ms,
               std::string& err);
void free_blas(BlasHost& b);

// One-sweep LSD radix sort of (key,value) pairs (rt_sort.cu), 8 bits per pass over `key_bits`
// bits, stable; the result lands in keys/vals (the alt buffers are scratch).  scratch must hold
// sort_hist_words(n) uint32.  The u32 variant can take its element count from device memory
// (count_dev != nullptr; n_max is then the upper bound the grid and scratch are sized for).
int sort_pairs(uint64_t* keys, uint64_t* keys_alt, uint32_t* vals, uint32_t* vals_alt, uint64_t n, int key_bits,
               uint32_t* scratch, cudaStream_t stream, std::string& err);
int sort_pairs_u32(uint32_t* keys, uint32_t* keys_alt, uint32_t* vals, uint32_t* vals_alt, uint64_t n_max,
                   const uint32_t* count_dev, int key_bits, uint32_t* scratch, cudaStream_t stream, std::string& err);
size_t sort_hist_words(uint64_t n);
int sort_launches(int key_bits);

// Area-weighted vertex normals of a packed mesh on the device (rt_normals.cu).
int generate_normals(float* d_verts, uint32_t nv, const uint32_t* d_idx, uint32_t nt, cudaStream_t st, float* ms_out,
                     std::string& err);

// Per-scene wavefront buffers and the frame pipeline.
struct FramePipeline;
FramePipeline* pipeline_create();
void pipeline_destroy(FramePipeline*);

struct FrameRequest {
  SceneDev scene;
  FrameUniforms U;
  uint32_t output_format;
  bool counting;
  bool want_primary;     // also record t of primary hits
  void* d_out;           // device output (format per output_format), may be null
  rt_hit* d_primary;     // device, n_slots*spp records in (y*w+x)*spp+s order, may be null
};
int pipeline_render(FramePipeline*, const FrameRequest&, cudaStream_t, rt_stats* stats, std::string& err);
// fills the frame fields of rt_stats for the last pipeline_render (event arithmetic on demand)
void pipeline_frame_stats(FramePipeline*, rt_stats* stats);
int pipeline_trace_batch(FramePipeline*, const SceneDev&, uint64_t n, const float* d_o, const float* d_d, float tmin,
                         float tmax, bool any_hit, bool counting, rt_hit* d_out, cudaStream_t, rt_stats* stats,
                         std::string& err);
int assemble_shards(const FrameUniforms& U, uint32_t format, const void* gathered, void* frame, cudaStream_t,
                    std::string& err);
int sm_count();
int measure_fp32_tflops(double* tflops, std::string& err);

}  // namespace rt
