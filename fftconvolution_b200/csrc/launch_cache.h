// Host-side caches for the two CUDA runtime calls that used to precede every launch: setting a kernel's dynamic
// shared-memory limit and asking for its occupancy.  Both cost microseconds of CPU time while the GPU waits -- a
// third of a single convolve() call -- and their answers never change for a (device, kernel,
// block size, shared memory) tuple.
#pragma once
#include <cuda_runtime.h>

#include <map>
#include <mutex>
#include <tuple>

namespace b2c {

inline cudaError_t ensure_dyn_smem(const void* fn, size_t bytes) {
  static std::mutex mu;
  static std::map<std::pair<int, const void*>, size_t> seen;
  int dev = 0;
  cudaGetDevice(&dev);
  std::lock_guard<std::mutex> g(mu);
  auto it = seen.find({dev, fn});
  if (it != seen.end() && it->second >= bytes) return cudaSuccess;
  const cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
  if (e == cudaSuccess) seen[{dev, fn}] = bytes;
  return e;
}

// resident CTAs per SM of `fn` at this block size and dynamic shared memory (0 when it cannot be resident)
inline int cached_occupancy(const void* fn, int threads, size_t smem) {
  static std::mutex mu;
  static std::map<std::tuple<int, const void*, int, size_t>, int> seen;
  int dev = 0;
  cudaGetDevice(&dev);
  {
    std::lock_guard<std::mutex> g(mu);
    auto it = seen.find(std::make_tuple(dev, fn, threads, smem));
    if (it != seen.end()) return it->second;
  }
  if (ensure_dyn_smem(fn, smem) != cudaSuccess) { cudaGetLastError(); return 0; }
  int n = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, fn, threads, smem) != cudaSuccess) { cudaGetLastError(); n = 0; }
  std::lock_guard<std::mutex> g(mu);
  seen[std::make_tuple(dev, fn, threads, smem)] = n;
  return n;
}

}  // namespace b2c
