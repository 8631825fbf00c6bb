#!/bin/sh
# rewrite.sh <reference shader dir> <out dir>
# Purely textual GLSL -> C++ rewrites of the reference's brdf.glsl and rand.glsl (pcg2d, prng_seed).
# Output goes to oracle/_ref/ (git-ignored); reference sources are never copied into the repo.
set -e
SRC="$1"; OUT="$2"
LIT='s/\([0-9][0-9]*\.[0-9][0-9]*\(e-\{0,1\}[0-9][0-9]*\)\{0,1\}\)/\1f/g'   # GLSL float literals are fp32
sed -e "$LIT" \
    -e 's/view_dir\.xy \* roughness/xy(view_dir) * roughness/' \
    -e 's/(hemisphere_sample + transformed_view)\.xy/xy(hemisphere_sample + transformed_view)/' \
    "$SRC/brdf.glsl" > "$OUT/brdf.inc"
# only the two live functions of rand.glsl (the rest is dead code)
awk '/^vec2 pcg2d/,/^#endif/' "$SRC/rand.glsl" | grep -v '^#endif' | sed -e "$LIT" \
    -e 's/inout uvec2 seed/uvec2\& seed/' \
    -e 's/return vec2(seed) \*/return to_vec2(seed) */' \
    -e 's/uvec2(pos\.xy)/uvec2(pos)/' \
    > "$OUT/rand.inc"
