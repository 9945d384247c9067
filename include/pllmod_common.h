/* Drop-in shim: callers of the reference that include "pllmod_common.h" get the B200 library's
 * declarations for the hot path (see pllb200.h for the reference file:line of each symbol). */
#ifndef PLLB200_SHIM_PLLMOD_COMMON_H_
#define PLLB200_SHIM_PLLMOD_COMMON_H_
#include "pllb200.h"
#endif
