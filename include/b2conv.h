/* b2conv.h -- C-ABI of libb2conv.so, the B200 (sm_100a) batched FP64 convolution engine that
 * replaces the hot path of jeremyfix/FFTConvolution:
 *   FFTW_Convolution::{init_workspace, convolve, clear_workspace}  (Convolution/src/convolution_fftw.h)
 *   + the size planner                                             (Convolution/src/factorize.h)
 *
 * Plain C: ints, longs, pointers.  No C++ types, no exceptions, no torch types cross this line.
 * Every function returns 0 on success and a non-zero B2C_E* code on failure; b2c_last_error()
 * then describes the failure (thread-local string).  There is NO CPU fallback: a compute call on
 * a machine without a usable GPU fails with B2C_ECUDA.
 *
 * The reference-side binding (the drop-in C++ header) is include/convolution_fftw.h; the ctypes
 * binding used by tests/bench is fftconvolution_b200/capi.py.  See INTEGRATION.md.
 */
#ifndef B2CONV_H
#define B2CONV_H

#ifdef __cplusplus
extern "C" {
#endif

/* Convolution_Mode ordinals -- identical to convolution_fftw.h:17-27 (all eight, same order). */
enum {
  B2C_LINEAR_FULL = 0,
  B2C_LINEAR_SAME_UNPADDED = 1,
  B2C_LINEAR_SAME = 2,
  B2C_LINEAR_VALID = 3,
  B2C_CIRCULAR_SAME = 4,
  B2C_CIRCULAR_SAME_PADDED = 5,
  B2C_CIRCULAR_FULL_UNPADDED = 6,
  B2C_CIRCULAR_FULL = 7
};

enum { B2C_OK = 0, B2C_EINVAL = 1, B2C_ECUDA = 2, B2C_ENOMEM = 3, B2C_ESIZE = 4 };

/* Compute path selection (b2c_plan_set_path). AUTO = planner's measured crossover. */
/* FFT_GENERIC = the any-size FFT kernel even where a tuned kernel exists (tests, comparisons). */
enum { B2C_PATH_AUTO = 0, B2C_PATH_FFT = 1, B2C_PATH_DIRECT = 2, B2C_PATH_FFT_GENERIC = 3 };

typedef struct b2c_plan b2c_plan;

/* Replaces init_workspace (convolution_fftw.h:44-148).  Pure host planning: computes h_dst/w_dst
 * exactly as the reference does (bit-for-bit, including 0x0 for an impossible LINEAR_VALID,
 * :75-92) and picks the 2/3/5/7-smooth FFT periods for the shared-memory radix schedule
 * (replaces find_closest_factor, factorize.h:92-104).  Allocates no device memory; device tables
 * are created lazily by the first compute call.  Unknown mode -> B2C_EINVAL. */
int b2c_plan_create(b2c_plan** plan, int mode, int h_src, int w_src, int h_kernel, int w_kernel);

/* Replaces reading ws.h_fftw/w_fftw/h_dst/w_dst (convolution_fftw.h:29-42).  h_fft/w_fft are the
 * NEW planner's periods (no reference caller reads them).  Any out-pointer may be NULL. */
int b2c_plan_query(const b2c_plan* plan, int* h_fft, int* w_fft, int* h_dst, int* w_dst);

/* Extra planning facts: path actually used (B2C_PATH_FFT/DIRECT), crop offsets, fold periods
 * (CIRCULAR_SAME computed as fold of the linear result; 0 = no fold), work-buffer bytes per CTA
 * and whether it lives in shared memory (1) or in a global scratch slab (0). */
int b2c_plan_info(const b2c_plan* plan, int* path, int* off_h, int* off_w, int* fold_h, int* fold_w,
                  long* work_bytes, int* work_in_smem);

/* Force the FFT or the direct kernel (crossover measurements, tests). */
int b2c_plan_set_path(b2c_plan* plan, int path);

/* Replaces clear_workspace (convolution_fftw.h:150-164).  NULL is accepted. */
int b2c_plan_destroy(b2c_plan* plan);

/* Replaces convolve (convolution_fftw.h:216-284) for one image with HOST pointers:
 * H2D, kernel spectrum, fused convolution kernel, D2H, synchronous.  dst is h_dst x w_dst
 * row-major.  An empty plan (LINEAR_VALID with kernel > source) is a no-op, like :218-219. */
int b2c_convolve_host(b2c_plan* plan, const double* src, const double* kernel, double* dst);

/* Batched form with HOST buffers (pinned or pageable): src [n][h_src][w_src], kernel
 * [h_kernel][w_kernel] if kernel_is_shared else [n][h_kernel][w_kernel], dst [n][h_dst][w_dst].
 * Chunks the batch and overlaps H2D, compute and D2H on three streams.  Synchronous. */
int b2c_convolve_batch_host(b2c_plan* plan, const double* src, const double* kernel,
                            int kernel_is_shared, double* dst, long n_images);

/* Batched form with DEVICE buffers on the current device, same layouts, 16-byte aligned.
 * Asynchronous on `cuda_stream` (a cudaStream_t; NULL = default stream).  With a shared kernel
 * the kernel spectrum is recomputed by every call (it is inside the timed region of bench.py). */
int b2c_convolve_batch_device(b2c_plan* plan, const double* d_src, const double* d_kernel,
                              int kernel_is_shared, double* d_dst, long n_images, void* cuda_stream);

/* No counterpart in the reference (it is double-real only, convolution_fftw.h:216): the same
 * batch with SINGLE-PRECISION device buffers (float in, float out; the arithmetic stays FP64: inputs are widened and
 * the result narrowed through a staging buffer the plan owns, in chunks of at most 256 MB), and with COMPLEX
 * double-precision buffers (interleaved re, im; kernel complex too; computed as four real convolutions).  Layouts
 * otherwise as b2c_convolve_batch_device; asynchronous on `cuda_stream`; the first call may allocate the staging. */
int b2c_convolve_batch_device_f32(b2c_plan* plan, const float* d_src, const float* d_kernel, int kernel_is_shared,
                                  float* d_dst, long n_images, void* cuda_stream);
int b2c_convolve_batch_device_c64(b2c_plan* plan, const double* d_src, const double* d_kernel, int kernel_is_shared,
                                  double* d_dst, long n_images, void* cuda_stream);

/* No counterpart in the reference.  Batched shared-kernel convolution with DEVICE buffers that
 * calls no cudaMalloc / cudaFree / allocating routine at all -- for callers that cannot allocate
 * GPU memory when they convolve: plug-ins inside an application whose memory pool already owns the
 * whole GPU, latency-critical loops that cannot take cudaMalloc's device-wide synchronisation, code
 * under a fixed memory budget.  Its scratch (kernel spectrum + root table of the tuned 108 x 108
 * kernel) is one file-scope device array of 98 496 bytes that every GPU using the library holds
 * until exit; the caller passes only inputs and outputs.
 * Limit: both alias-safe periods must be <= 108 (e.g. LINEAR_SAME needs h_src + h_kernel/2 <= 108,
 * LINEAR_FULL h_src + h_kernel - 1 <= 108, same for widths; CIRCULAR_SAME only for 108 x 108
 * sources); larger problems return B2C_ESIZE with a message naming the required periods.
 * The first call on a device does a one-off set-up (address look-up + blocking upload of the root
 * table): make it before any stream capture on that device.  All calls share the array, which
 * keeps the previous call's spectrum; each call rewrites it before use, so nothing needs clearing,
 * but calls whose GPU work overlaps race -- serialise them (one stream, or events).
 * Layouts as b2c_convolve_batch_device with kernel_is_shared = 1; asynchronous on `cuda_stream`. */
int b2c_convolve_batch_device_static(int mode, int h_src, int w_src, int h_kernel, int w_kernel,
                                     const double* d_src, const double* d_kernel, double* d_dst,
                                     long n_images, void* cuda_stream);

/* The reference planner, restated for callers that want its sizes (factorize.h:92-104 with
 * FFTW_FACTORS of convolution_fftw.h:15, including the "no multiple of 4" quirk of :74). */
int b2c_find_closest_factor(int n);

/* The new planner's period for a minimum alias-safe length n: the cheapest 2/3/5/7-smooth
 * size >= n under the radix-schedule cost model. */
int b2c_smooth_size(int n);

/* Batch partitioner: contiguous slice [begin,end) of n images owned by `rank` of
 * `world` (one process per GPU; no data-path collective). */
int b2c_partition(long n, int world, int rank, long* begin, long* end);

/* The same partitioner inside ONE process (no reference counterpart; the reference is single-threaded,
 * convolution_benchmark.cc:178-186 loops over convolve): the batch in host memory is cut into the contiguous
 * shards of b2c_partition and device g of `n_devices` (0 = all visible devices) convolves shard g on its own
 * host thread, plan and copy/compute pipeline.  Arguments as b2c_plan_create + b2c_convolve_batch_host.
 * Pinned host buffers (b2c_host_alloc) are needed for full PCIe rate. */
int b2c_multi_convolve_batch_host(int mode, int h_src, int w_src, int h_kernel, int w_kernel, const double* src,
                                  const double* kernel, int kernel_is_shared, double* dst, long n_images,
                                  int n_devices);

/* Order-independent verification checksum of n doubles on the current device: wrapping 64-bit
 * sum and xor of the IEEE-754 bit patterns, plus the plain double sum.  out = 3 x 64 bit on the
 * HOST: {sum_bits, xor_bits, double_sum_bits}.  Synchronous on `cuda_stream`.  Ranks combine
 * the first with an integer-sum all-reduce, the second with xor (or sum of per-rank values). */
int b2c_checksum_device(const double* d, long n, unsigned long long* out3, void* cuda_stream);

/* Counter-based synthetic data: d[i] = (splitmix64(seed + i) >> 11) * 2^-53
 * for i in [first, first+n).  Asynchronous on `cuda_stream`. */
int b2c_fill_uniform_device(double* d, long n, unsigned long long seed, long first, void* cuda_stream);

/* Pinned host memory for the host-buffer entry points (cudaHostAlloc / cudaFreeHost). */
int b2c_host_alloc(void** p, long bytes);
int b2c_host_free(void* p);

/* Make `device` current for this thread (cudaSetDevice) -- one process per GPU sets it once. */
int b2c_set_device(int device);

/* Number of engine kernels launched by this process so far (bench.py's gpu_launches). */
long b2c_launch_count(void);
/* Average device time in ms of the dominant convolution kernel over the launches recorded since
 * b2c_kernel_timer(1) (CUDA events on the launching stream); b2c_kernel_timer(0) stops. */
int b2c_kernel_timer(int enable);
int b2c_kernel_timer_read(double* avg_ms, long* launches);

const char* b2c_last_error(void);
const char* b2c_version(void);

#ifdef __cplusplus
}
#endif
#endif /* B2CONV_H */
