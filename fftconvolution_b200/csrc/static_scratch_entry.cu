// b2c_convolve_batch_device_static -- a batched shared-kernel convolution that never asks the
// CUDA allocator for anything.
//
// Who needs this: callers that are not allowed, or cannot afford, to allocate device memory at the
// moment they convolve.  Examples: a plug-in running inside a host application that has already
// handed every byte of the GPU to its own memory pool (a driver allocation would fail, or fragment
// that pool); a latency-critical loop in which cudaMalloc's device-wide synchronisation would stall
// the other streams; code that runs under a fixed, audited memory budget.  The regular entry
// points create their scratch lazily with cudaMalloc (kernel spectrum, twiddle tables), so they
// do not fit those callers.
//
// What it does instead: the scratch of the period-108 kernel (the (12, 9) member of conv_tuned2.cu,
// the same kernel the regular entry point runs at that period) -- the kernel's spectrum and the
// table of 108th roots of unity -- lives in ONE file-scope device array
// of 98 496 bytes that every GPU loading this library carries until the process exits.  The caller
// passes only its own input and output buffers.  Any problem whose alias-safe periods fit in 108 on
// both axes is run at the fixed period 108 x 108; anything larger is refused (B2C_ESIZE).
//
// Rules that follow from sharing one array:
//   * the first call on each device performs a one-off set-up (looks up the array's device address
//     and uploads the root table with a blocking copy).  Make that first call during start-up,
//     before any stream capture is begun on that device.
//   * the array keeps whatever the previous call left in it.  Nothing in it has to start at zero:
//     the root table is written once and never changes, and every call recomputes the whole kernel
//     spectrum before the convolution kernel reads it.
//   * two calls whose GPU work overlaps (different streams, or different host threads) would write
//     the spectrum while the other reads it: they race.  Serialise calls on one device -- same
//     stream, or an event between them.
//   * no cudaMalloc / cudaFree / allocating library routine is called here, on any path.
#include <cuda_runtime.h>

#include <cmath>
#include <cstdio>
#include <mutex>

#include "../../include/b2conv.h"
#include "fft_device.cuh"

namespace b2c {
// conv_tuned2.cu: kernel spectrum + convolution of a square menu period; allocates nothing
cudaError_t launch_tuned2(int P, int hs, int ws, int hk, int wk, int hd, int wd, int off_h, int off_w,
                          int hs_in, int ws_in, int sh_h, int sh_w, const double* d_src, const double* d_ker,
                          double* d_dst, long n, cplx* kspec, const cplx* tw, int sms, cudaStream_t st, int* launches, int kmod_h, int kmod_w);
// capi.cu
int set_last_error(int code, const char* text);
}  // namespace b2c

namespace {

constexpr int kPeriod = 108;                                   // the only period this entry point runs
constexpr size_t kSpecBytes = sizeof(double2) * 108 * 56;      // kernel spectrum: 108 positions x 55 frequencies, row pitch 56
constexpr size_t kRootsOffset = 96768;                         // = kSpecBytes, a multiple of 256
constexpr size_t kRootsBytes = sizeof(double2) * kPeriod;
constexpr size_t kScratchBytes = kRootsOffset + kRootsBytes;   // 98 496 bytes per GPU, for the life of the process
static_assert(kSpecBytes <= kRootsOffset, "layout");

// The scratch itself.  In host code this name is only a handle for cudaGetSymbolAddress -- it is
// NOT a device pointer and must never be handed to a kernel directly.
static __device__ __align__(256) unsigned char g_static_scratch[kScratchBytes];

constexpr int kMaxDevices = 64;
struct PerDevice {
  bool ready = false;
  unsigned char* base = nullptr;   // device address of g_static_scratch on this device
  int sms = 0;
};
PerDevice g_dev[kMaxDevices];
std::mutex g_dev_mu;

int cuda_fail(const char* what, cudaError_t e) {
  char buf[256];
  snprintf(buf, sizeof buf, "b2c_convolve_batch_device_static: %s failed: %s", what, cudaGetErrorString(e));
  return b2c::set_last_error(B2C_ECUDA, buf);
}

// One-off, per device, with that device current.
int set_up_device(int dev, PerDevice** out) {
  std::lock_guard<std::mutex> lk(g_dev_mu);
  PerDevice& d = g_dev[dev];
  if (!d.ready) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
    if (e != cudaSuccess) return cuda_fail("cudaGetSymbolAddress", e);
    d.base = static_cast<unsigned char*>(p);
    double2 roots[kPeriod];   // exp(-2*pi*i*k/108), on the host stack
    const long double two_pi = 6.283185307179586476925286766559005768L;
    for (int k = 0; k < kPeriod; ++k) {
      const long double a = two_pi * (long double)k / (long double)kPeriod;
      roots[k] = make_double2((double)cosl(a), (double)(-sinl(a)));
    }
    e = cudaMemcpy(d.base + kRootsOffset, roots, kRootsBytes, cudaMemcpyHostToDevice);
    if (e != cudaSuccess) return cuda_fail("cudaMemcpy(root table)", e);
    // the copy comes from pageable memory and may return before its DMA has landed; the kernels below run on
    // the caller's stream, which need not be ordered after the legacy stream: wait here, once per device
    e = cudaStreamSynchronize(0);
    if (e != cudaSuccess) return cuda_fail("cudaStreamSynchronize(root table)", e);
    e = cudaDeviceGetAttribute(&d.sms, cudaDevAttrMultiProcessorCount, dev);
    if (e != cudaSuccess) return cuda_fail("cudaDeviceGetAttribute", e);
    d.ready = true;
  }
  *out = &d;
  return 0;
}

// Output extent, crop offset and smallest alias-safe period of one axis (same table as DESIGN.md 1).
struct Axis { int n_dst, off, min_period; };
bool plan_axis(int mode, int ns, int nk, Axis* a) {
  const int full = ns + nk - 1;
  switch (mode) {
    case B2C_LINEAR_FULL: case B2C_CIRCULAR_FULL_UNPADDED: case B2C_CIRCULAR_FULL:
      *a = {full, 0, full}; return true;
    case B2C_LINEAR_SAME_UNPADDED: case B2C_LINEAR_SAME:
      *a = {ns, nk / 2, ns + nk / 2}; return true;
    case B2C_LINEAR_VALID:
      *a = {ns - nk + 1, nk - 1, ns}; return true;
    case B2C_CIRCULAR_SAME_PADDED:
      *a = {ns, 0, full}; return true;
    case B2C_CIRCULAR_SAME:   // needs the period to equal the source size exactly
      *a = {ns, 0, ns == kPeriod ? kPeriod : kPeriod + 1}; return true;
  }
  return false;
}

}  // namespace

extern "C" int b2c_convolve_batch_device_static(int mode, int h_src, int w_src, int h_kernel, int w_kernel,
                                                const double* d_src, const double* d_kernel, double* d_dst,
                                                long n_images, void* cuda_stream) {
  if (mode < 0 || mode > 7) return b2c::set_last_error(B2C_EINVAL, "b2c_convolve_batch_device_static: unrecognized convolution mode");
  if (h_src < 1 || w_src < 1 || h_kernel < 1 || w_kernel < 1 || n_images < 0)
    return b2c::set_last_error(B2C_EINVAL, "b2c_convolve_batch_device_static: sizes must be positive");
  if (mode == B2C_LINEAR_VALID && (h_kernel > h_src || w_kernel > w_src)) return 0;   // empty result, like convolution_fftw.h:218-219
  if (n_images == 0) return 0;
  if (!d_src || !d_kernel || !d_dst) return b2c::set_last_error(B2C_EINVAL, "b2c_convolve_batch_device_static: null buffer");
  Axis ah, aw;
  plan_axis(mode, h_src, h_kernel, &ah);
  plan_axis(mode, w_src, w_kernel, &aw);
  if (ah.min_period > kPeriod || aw.min_period > kPeriod) {
    char buf[320];
    if (mode == B2C_CIRCULAR_SAME)
      snprintf(buf, sizeof buf, "b2c_convolve_batch_device_static: CIRCULAR_SAME needs the period to equal the source size; "
               "the static scratch only serves 108 x 108 (got %d x %d)", h_src, w_src);
    else
      snprintf(buf, sizeof buf, "b2c_convolve_batch_device_static: this problem needs periods >= %d x %d but the static "
               "scratch (%zu bytes) serves at most 108 x 108; use b2c_convolve_batch_device", ah.min_period, aw.min_period,
               kScratchBytes);
    return b2c::set_last_error(B2C_ESIZE, buf);
  }
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return cuda_fail("cudaGetDevice", e);
  if (dev < 0 || dev >= kMaxDevices) return b2c::set_last_error(B2C_ESIZE, "b2c_convolve_batch_device_static: device index beyond 63");
  PerDevice* d = nullptr;
  if (int rc = set_up_device(dev, &d)) return rc;

  b2c::cplx* spectrum = reinterpret_cast<b2c::cplx*>(d->base);
  const b2c::cplx* roots = reinterpret_cast<const b2c::cplx*>(d->base + kRootsOffset);
  cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
  // every call rewrites the whole spectrum before it is read (see the rules at the top of the file)
  int launches = 0;
  e = b2c::launch_tuned2(kPeriod, h_src, w_src, h_kernel, w_kernel, ah.n_dst, aw.n_dst, ah.off, aw.off, h_src, w_src, 0, 0,
                         d_src, d_kernel, d_dst, n_images, spectrum, roots, d->sms, st, &launches, kPeriod, kPeriod);
  if (e != cudaSuccess) return cuda_fail("kernel launch", e);
  return 0;
}
