/* TEST INFRASTRUCTURE ONLY - stand-in for <fftw3.h>.
 *
 * FFTW3 is an un-vendored, unpinned system dependency of the reference (Makefile.rules:22-23, `pkg-config fftw3`) and
 * is not installed in this image.  The reference's libgadget/petapm.cpp touches FFTW directly only to switch its
 * threading on (petapm.cpp:187-189); the transforms themselves go through heFFTe, whose stand-in is ref_shim/heffte.h.
 * Nothing in the product (shenqi_b200/) includes this file.
 */
#ifndef SHENQI_ORACLE_FFTW3_SHIM_H
#define SHENQI_ORACLE_FFTW3_SHIM_H
#ifdef __cplusplus
extern "C" {
#endif
typedef double fftw_complex[2];
static inline int fftw_init_threads(void) { return 1; }
static inline void fftw_plan_with_nthreads(int nthreads) { (void) nthreads; }
static inline void fftw_cleanup_threads(void) { }
#ifdef __cplusplus
}
#endif
#endif
