// host/shim/gsl/gsl_randist.h -- build shim (not part of the GPU core).
// The reference's src/cameras/realisticEye.h:45 includes GSL, which is not
// vendored and not installed.  The camera itself is out of the hot path and is
// not compiled (host/Makefile); this opaque type only lets src/core/api.cpp
// parse the class declaration.
#ifndef SPTX_GSL_SHIM_H
#define SPTX_GSL_SHIM_H
typedef struct gsl_rng_opaque gsl_rng;
#endif
