/* Drop-in shim: callers of the reference that include "pll_binary.h" (the binary checkpoint I/O module) get the B200
 * library's declarations (see pllb200.h for the reference file:line of each symbol). */
#ifndef PLLB200_SHIM_PLL_BINARY_H_
#define PLLB200_SHIM_PLL_BINARY_H_
#include "pllb200.h"
#endif
