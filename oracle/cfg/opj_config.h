/* Hand-written stand-in for grok's cmake-generated opj_config.h
 * (template: ext/grok/src/lib/openjp2/opj_config.h.cmake.in). Test infrastructure only. */
#define OPJ_HAVE_STDINT_H 1
#define OPJ_VERSION_MAJOR 2
#define OPJ_VERSION_MINOR 1
#define OPJ_VERSION_BUILD 1
#define OPENJPEG_PLUGIN_NAME "grok_plugin"
