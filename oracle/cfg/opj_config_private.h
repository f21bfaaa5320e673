/* Hand-written stand-in for grok's cmake-generated opj_config_private.h
 * (template: ext/grok/src/lib/openjp2/opj_config_private.h.cmake.in). Test infrastructure only. */
#define OPJ_HAVE_INTTYPES_H 1
#define OPJ_PACKAGE_VERSION "2.1.1"
#define OPJ_HAVE_FSEEKO 1
#define OPJ_HAVE_MALLOC_H
#define OPJ_HAVE_POSIX_MEMALIGN
#if !defined(_POSIX_C_SOURCE)
#define _POSIX_C_SOURCE 200112L
#endif
