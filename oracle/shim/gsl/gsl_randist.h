// oracle/shim/gsl/gsl_randist.h -- TEST INFRASTRUCTURE ONLY.
// The reference's src/cameras/realisticEye.h:45 includes GSL, which is not
// vendored and not installed.  The camera itself is out of the hot path and is
// not compiled (oracle/Makefile); this opaque type only lets src/core/api.cpp
// parse the class declaration.
#ifndef SPTX_ORACLE_GSL_SHIM_H
#define SPTX_ORACLE_GSL_SHIM_H
typedef struct gsl_rng_opaque gsl_rng;
#endif
