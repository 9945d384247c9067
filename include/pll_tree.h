/* Drop-in shim: callers of the reference that include "pll_tree.h" get the B200 library's
 * declarations for the hot path (see pllb200.h for the reference file:line of each symbol). */
#ifndef PLLB200_SHIM_PLL_TREE_H_
#define PLLB200_SHIM_PLL_TREE_H_
#include "pllb200.h"
#endif
