#!/bin/bash
# Is the device code of the working tree the same as that of commit $1?  Builds the library of that
# commit in a scratch directory and compares `cuobjdump -sass` of both libraries (anonymous-namespace
# hashes normalised).  No GPU needed.  Used after moving kernels between headers: identical SASS means
# the GPU-validated binary and the refactored one run the same instructions.
#   scripts/sass_identity_check.sh 35ff821
set -e
commit=${1:?commit}
root=$(cd "$(dirname "$0")/.." && pwd)
scratch=$(mktemp -d)
trap 'rm -rf "$scratch"' EXIT
git -C "$root" archive "$commit" or-tools_b200/csrc include | tar -x -C "$scratch"
make -C "$scratch/or-tools_b200/csrc" -j8 > "$scratch/build.log" 2>&1
make -C "$root/or-tools_b200/csrc" -j8 > /dev/null 2>&1
norm() { cuobjdump -sass "$1" 2>/dev/null | sed -E 's/_GLOBAL__N__[0-9a-f]+_[0-9]+_([a-z_]+)_cu_[0-9a-f]+/ANON_\1/g' | grep -vE '^\s*/\* 0x[0-9a-f]+ \*/$'; }
norm "$scratch/or-tools_b200/lib/libpdlp_b200.so" > "$scratch/a.sass"
norm "$root/or-tools_b200/lib/libpdlp_b200.so" > "$scratch/b.sass"
echo "$(wc -l < "$scratch/a.sass") / $(wc -l < "$scratch/b.sass") SASS lines"
if cmp -s "$scratch/a.sass" "$scratch/b.sass"; then echo "device code identical to $commit"; else echo "device code DIFFERS from $commit"; diff "$scratch/a.sass" "$scratch/b.sass" | head -20; exit 1; fi
