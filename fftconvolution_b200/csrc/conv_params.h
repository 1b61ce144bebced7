// Kernel parameter block shared by host (capi.cu) and device (conv_generic.cu, conv_direct.cu).
#pragma once
#include <cuda_runtime.h>

#include "fft_device.cuh"

namespace b2c {

struct ConvParams {
  // geometry (reference names: h_src,w_src,h_kernel,w_kernel,h_dst,w_dst -- convolution_fftw.h:29-42)
  int hs, ws, hk, wk, hd, wd;
  int off_h, off_w;      // crop origin inside the linear result (convolution_fftw.h:227-272)
  int fold_h, fold_w;    // >0: fold the linear result modulo this on store (CIRCULAR_SAME)
  int lin_h, lin_w;      // extent of the linear result that has to be produced
  int kwrap_h, kwrap_w;  // kernel taps wrap-accumulate modulo these (convolution_fftw.h:185-187)
  int ld;                // leading dimension (complex elements) of the work buffer
  AxisDev ah, aw;
  double scale;          // 1 / (P_h * P_w)            (convolution_fftw.h:210-211)
  const double* src;     // [n][hs][ws]
  const double* ker;     // [hk][wk] or [n][hk][wk]
  double* dst;           // [n][hd][wd]
  long n_images;
  cplx* kspec;           // shared-kernel spectrum, P_h x P_w, scrambled order, pre-scaled
  cplx* work;            // global work slabs [gridDim.x][P_h*ld] when the image does not fit in smem
  int work_in_smem;
};

struct DirectParams {
  int hs, ws, hk, wk, hd, wd;
  int off_h, off_w;
  int wrap_h, wrap_w;    // >0: source indices wrap modulo this (CIRCULAR_SAME), else zero outside
  int tile_h;            // output rows per CTA
  int tiles_per_image;
  int group;             // > 1: a work unit is `group` whole consecutive images (tiles_per_image == 1, shared kernel)
  int sp_w;              // padded staged-source row length (doubles)
  int kernel_is_shared;
  // bulk-staged variant (conv_direct.cu: direct_bulk_kernel): whole images, rows copied by the TMA engine
  int bulk;              // 0: element-wise staging; 1 / 2: bulk-staged with that many tile buffers per CTA
  int ti;                // output rows per thread of the bulk-staged kernel (direct_bulk_ti)
  int odd;               // that kernel's odd-width form: per-row column shift, edge element by cp.async
  int col_shift;         // staged column of source column c is c - (off_w - (wk - 1)) + col_shift (keeps copies 16-byte aligned)
  const double* src;
  const double* ker;
  double* dst;
  long n_images;
};

// strip-mode kernel (conv_big2d.cu)
struct BigParams {
  int hs, ws, hk, wk, hd, wd;
  int off_h, off_w;
  int kwrap_h, kwrap_w;
  int half_w;        // P_w / 2 + 1 half-spectrum columns
  int ldw;           // row pitch (complex) of the row-phase buffer, odd
  int rows_chunk;    // rows per row-phase chunk
  int cols_strip;    // half-spectrum columns per column-phase strip
  int ldc;           // row pitch (complex) of the column-phase buffer = 2*cols_strip (+1)
  int rows_y;        // rows of the linear result that are produced: off_h + hd
  int bulk_rows;     // conv_big2ds.cu: source / kernel rows and result-spectrum rows enter the row buffer by bulk copies
  int src_h, src_w;  // conv_big2ds.cu: extents of the stored source; (hs, ws) are the input extents the transforms see --
  int sh_h, sh_w;    //   larger by the shift when the source is extended periodically: input (r, c) = source ((r - sh_h) mod src_h, (c - sh_w) mod src_w)
  AxisDev ah, aw;
  double scale;
  const double* src;
  const double* ker;
  long ker_stride;   // 0: one shared kernel
  double* dst;
  long n_images;
  cplx* slab;        // [gridDim.x][max(hs + hke, rows_y) * half_w]: Y replaces S|K strip by strip
  const double* alpha;   // [n_images]: power-of-two kernel scale of the packed row transforms (balance_kernel)
};

}  // namespace b2c
