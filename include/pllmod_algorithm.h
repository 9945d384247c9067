/* Drop-in shim: callers of the reference that include "pllmod_algorithm.h" (the search / model-optimisation algorithms) get the B200
 * library's declarations (see pllb200.h for the reference file:line of each symbol). */
#ifndef PLLB200_SHIM_PLLMOD_ALGORITHM_H_
#define PLLB200_SHIM_PLLMOD_ALGORITHM_H_
#include "pllb200.h"
#endif
