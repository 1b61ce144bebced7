/* TEST INFRASTRUCTURE ONLY -- stand-in for <fftw3.h>.
 *
 * The build container (and, unless probed otherwise, the GPU box) has no libfftw3, so the
 * UNMODIFIED reference header Convolution/src/convolution_fftw.h cannot be compiled as shipped.
 * It uses exactly nine FFTW names (convolution_fftw.h:5,36-38,135,137,143-147,153-163,190,191,208):
 * fftw_complex, fftw_plan, FFTW_ESTIMATE, fftw_malloc, fftw_free, fftw_plan_dft_r2c_2d,
 * fftw_plan_dft_c2r_2d, fftw_execute, fftw_destroy_plan.  This header declares those and nothing
 * else; oracle/shim/fftw3_shim.cpp implements them on oracle/mixed_fft.h with FFTW's documented
 * conventions (unnormalised transforms, r2c output h x (w/2+1) interleaved complex).
 * Numbers timed through this shim are labelled "reference header over shim FFT -- not FFTW".
 */
#ifndef ORACLE_SHIM_FFTW3_H
#define ORACLE_SHIM_FFTW3_H
#include <stddef.h>
#ifdef __cplusplus
extern "C" {
#endif
typedef double fftw_complex[2];
typedef struct orc_fftw_plan_s* fftw_plan;
#define FFTW_ESTIMATE (1U << 6)
void* fftw_malloc(size_t n);
void fftw_free(void* p);
fftw_plan fftw_plan_dft_r2c_2d(int n0, int n1, double* in, fftw_complex* out, unsigned flags);
fftw_plan fftw_plan_dft_c2r_2d(int n0, int n1, fftw_complex* in, double* out, unsigned flags);
void fftw_execute(const fftw_plan p);
void fftw_destroy_plan(fftw_plan p);
#ifdef __cplusplus
}
#endif
#endif
