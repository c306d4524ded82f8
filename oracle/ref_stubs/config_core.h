/* Hand-written stand-in for the file the reference's CMake generates from
 * configure/config_core.in (CMakeLists.txt:389).  Default option values of the reference
 * build: MATHBACKEND 4, NATIVE_SIZE 64, int128 available, BE2/BE4 compiled in, no NTL/TCM. */
#ifndef ORACLE_REF_CONFIG_CORE_H
#define ORACLE_REF_CONFIG_CORE_H
#define WITH_BE2
#define WITH_BE4
#define CKKS_M_FACTOR 1
#define HAVE_INT128 TRUE
#define HAVE_INT64 TRUE
#define MATHBACKEND 4
#define NATIVEINT 64
#endif
