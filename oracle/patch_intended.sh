#!/bin/sh
# patch_intended.sh REF TMPDIR -- writes P1-P3 patched copies of three reference files into
# TMPDIR (a mktemp directory the Makefile deletes).  Every patch restores something the
# reference itself contains commented out.  Fails if a patch does not apply.
set -e
REF=$1; T=$2
mkdir -p $T/ratio $T/path
# P1
sed 's|owl::common::vec3f brdf = owl::common::vec3f(0.);|owl::common::vec3f p1_wi = normalize(apply_mat(current_si.to_local, L)); owl::common::vec3f brdf = evaluate(p1_wi, current_si.wo_local, current_si.diffuse, current_si.alpha, current_si.emit);|' \
    $REF/cuda_include/ratio/ratio.cuh > $T/ratio/ratio.cuh
grep -q 'p1_wi' $T/ratio/ratio.cuh
# P2
sed -e '/^ *color = wi_local; *$/d' -e '/^ *continue; *$/d' $REF/cuda_include/path/path.cuh > $T/path/path.cuh
test "$(grep -c '^ *continue; *$' $T/path/path.cuh)" = 0
test "$(grep -c '^ *color = wi_local; *$' $REF/cuda_include/path/path.cuh)" = 1
test "$(grep -c '^ *continue; *$' $REF/cuda_include/path/path.cuh)" = 1
# P3
sed -e 's|int n = 4;|int n = shim_shadow_spp;|' -e 's|estimatePathTracing(si, rng, ray, 1)|estimatePathTracing(si, rng, ray, shim_max_depth)|' \
    $REF/cuda_src/deviceCode.cu > $T/deviceCode.cu
grep -q 'shim_shadow_spp' $T/deviceCode.cu
grep -q 'shim_max_depth' $T/deviceCode.cu
