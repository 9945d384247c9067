#!/bin/bash
# Builds the UNMODIFIED reference (libpll-2 + pll-modules C sources, compiled from where they lie
# under /root/reference) into oracle/_ref/libpll_ref.so. Test infrastructure only: this library is the
# CPU oracle / CPU baseline; nothing in the product links or loads it.
# Recipe follows SURVEY.md Appendix B (direct gcc, no flex/bison, per-file SIMD flags as
# libs/libpll/src/CMakeLists.txt:11-13,162-195).  Outputs go ONLY to oracle/_ref/.
set -e
REF=${REF:-/root/reference}
HERE="$(cd "$(dirname "$0")" && pwd)"
O="$HERE/_ref"
if [ ! -d "$REF/libs/libpll/src" ]; then
  echo "build_ref: $REF not present; keeping prebuilt oracle/_ref (if any)"; exit 0
fi
mkdir -p "$O/obj"
S=$REF/libs/libpll/src ; M=$REF/src
BASE="-O3 -fPIC -D_GNU_SOURCE -DHAVE_SSE3 -DHAVE_AVX -DHAVE_AVX2 -DHAVE_X86INTRIN_H -I$S -I$M -I$M/tree -I$M/optimize -I$M/algorithm -I$M/util -I$M/msa -I$M/binary -w"
cd "$O/obj"
pids=()
cc() { gcc $BASE "$@" & pids+=($!); if [ ${#pids[@]} -ge 8 ]; then wait ${pids[0]}; pids=("${pids[@]:1}"); fi; }
for f in pll models partials likelihood derivatives core_partials core_likelihood core_derivatives core_pmatrix \
         gamma hardware repeats maps list output compress random fasta phylip utree utree_moves rtree \
         parsimony fast_parsimony stepwise utree_svg; do cc -c $S/$f.c -o l_$f.o; done
for f in core_partials core_likelihood core_derivatives core_pmatrix fast_parsimony; do
  cc -msse3       -c $S/${f}_sse.c  -o l_${f}_sse.o
  cc -mavx        -c $S/${f}_avx.c  -o l_${f}_avx.o
  cc -mavx2 -mfma -c $S/${f}_avx2.c -o l_${f}_avx2.o; done
# plain-C parts of parse_utree.y (wraptree/destroy/...), no grammar; extracted by line range into _ref/ only
{ sed -n '22,158p' $S/parse_utree.y; sed -n '270,568p' $S/parse_utree.y; } > utree_nogrammar.c
cc -c utree_nogrammar.c -o l_utree_nogrammar.o
for f in tree/treeinfo tree/pll_tree tree/utree_operations tree/utree_distances tree/utree_constraint \
         tree/rtree_operations tree/tree_hashtable tree/consensus tree/tbe_functions optimize/pll_optimize \
         optimize/opt_algorithms optimize/lbfgsb/lbfgsb optimize/lbfgsb/linesearch optimize/lbfgsb/linpack \
         optimize/lbfgsb/miniCBLAS optimize/lbfgsb/subalgorithms pllmod_common algorithm/pllmod_algorithm \
         algorithm/algo_search algorithm/algo_callback util/models util/models_dna util/models_aa \
         util/models_gt util/models_mult msa/pll_msa binary/pll_binary binary/binary_io_operations; do cc -c $M/$f.c -o m_$(echo $f | tr / _).o; done
printf 'void*pll_utree_parse_newick_string(const char*s){return 0;}\nvoid*pll_utree_split_newick_string(char*s,unsigned t,void*h){return 0;}\n' > stubs.c
cc -c stubs.c -o l_stubs.o
wait
# one shared object; -Bsymbolic so its internal pll_* calls never bind to the product library's symbols
gcc -shared -Wl,-Bsymbolic -o "$O/libpll_ref.so" l_*.o m_*.o -lm
# the same-source harness (oracle/harness.c) compiled against the reference headers + library
if [ -f "$HERE/harness.c" ]; then
  gcc -O2 -fPIC -shared -Wl,-Bsymbolic -DHARNESS_REFERENCE -I$S -I$M -I$M/tree -I$M/optimize -I$M/algorithm -I$M/util -I$M/msa -I$M/binary -w \
      "$HERE/harness.c" -o "$O/libharness_ref.so" -L"$O" -l:libpll_ref.so -Wl,-rpath,'$ORIGIN' -lm -lpthread
fi
echo "build_ref: ok -> $O/libpll_ref.so"
