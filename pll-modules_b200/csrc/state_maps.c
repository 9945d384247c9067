/*
 * state_maps.c - the three stock ASCII -> state-bitmask maps used for tip encoding
 * (reference: libs/libpll/src/maps.c, pll.h:556-559).  Written with C99 designated initialisers:
 *   bin: '0'=1 '1'=2, gap/unknown = both
 *   nt : IUPAC nucleotide codes as 4-bit masks A=1 C=2 G=4 T/U=8, gap/N/X/O/?/. = 15
 *   aa : one bit per residue in the order A R N D C Q E G H I L K M F P S T W Y V,
 *        B={N,D} Z={Q,E} J={I,L}, X * - ? . = any residue
 */
#include "../../include/pllb200.h"

const pll_state_t pll_map_bin[256] = {['-'] = 3, ['.'] = 3, ['0'] = 1, ['1'] = 2, ['?'] = 3};

#define NT2(c, v) [c] = v, [c + 32] = v
const pll_state_t pll_map_nt[256] = {
    ['-'] = 15, ['.'] = 15, ['?'] = 15,
    NT2('A', 1), NT2('B', 14), NT2('C', 2), NT2('D', 13), NT2('G', 4), NT2('H', 11), NT2('K', 12),
    NT2('M', 3), NT2('N', 15), NT2('O', 15), NT2('R', 5), NT2('S', 6), NT2('T', 8), NT2('U', 8),
    NT2('V', 7), NT2('W', 9), NT2('X', 15), NT2('Y', 10)};

#define AA1(c, bit) [c] = (1ull << bit), [c + 32] = (1ull << bit)
#define AAM(c, v) [c] = v, [c + 32] = v
const pll_state_t pll_map_aa[256] = {
    ['*'] = 0xfffff, ['-'] = 0xfffff, ['.'] = 0xfffff, ['?'] = 0xfffff,
    AA1('A', 0), AA1('R', 1), AA1('N', 2), AA1('D', 3), AA1('C', 4), AA1('Q', 5), AA1('E', 6),
    AA1('G', 7), AA1('H', 8), AA1('I', 9), AA1('L', 10), AA1('K', 11), AA1('M', 12), AA1('F', 13),
    AA1('P', 14), AA1('S', 15), AA1('T', 16), AA1('W', 17), AA1('Y', 18), AA1('V', 19),
    AAM('B', 0xc), AAM('Z', 0x60), AAM('J', 0x600), AAM('X', 0xfffff)};
