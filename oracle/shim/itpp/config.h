/* Hand-written stand-in for the config.h that IT++'s own build system would
 * generate (test infrastructure only; see oracle/README.md).  It declares what
 * glibc/libstdc++ on this image provide and switches every optional external
 * library (BLAS, LAPACK, FFT) off: none of them is on the LUT decode path. */
#ifndef CONFIG_H
#define CONFIG_H
#define HAVE_ACOSH 1
#define HAVE_ASINH 1
#define HAVE_ATANH 1
#define HAVE_CBRT 1
#define HAVE_CMATH 1
#define HAVE_COMPLEX 1
#define HAVE_DECL_ISFINITE 1
#define HAVE_DECL_ISINF 1
#define HAVE_DECL_ISNAN 1
#define HAVE_DECL_SIGNGAM 1
#define HAVE_DLFCN_H 1
#define HAVE_ERF 1
#define HAVE_ERFC 1
#define HAVE_EXPM1 1
#define HAVE_FINITE 1
#define HAVE_INTTYPES_H 1
#define HAVE_ISFINITE 1
#define HAVE_ISINF 1
#define HAVE_ISNAN 1
#define HAVE_LGAMMA 1
#define HAVE_LOG1P 1
#define HAVE_LOG2 1
#define HAVE_MEMORY_H 1
#define HAVE_RINT 1
#define HAVE_STDINT_H 1
#define HAVE_STDLIB_H 1
#define HAVE_STD_ISFINITE 1
#define HAVE_STD_ISINF 1
#define HAVE_STD_ISNAN 1
#define HAVE_STRINGS_H 1
#define HAVE_STRING_H 1
#define HAVE_SYS_STAT_H 1
#define HAVE_SYS_TYPES_H 1
#define HAVE_TGAMMA 1
#define HAVE_UNISTD_H 1
#define STDC_HEADERS 1
#define TIME_WITH_SYS_TIME 1
#define PACKAGE "itpp"
#define PACKAGE_NAME "IT++"
#define PACKAGE_VERSION "4.3.1"
#define PACKAGE_STRING "IT++ 4.3.1"
#define VERSION "4.3.1"
#include <cmath>
#endif
