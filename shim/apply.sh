#!/bin/sh
# usage: shim/apply.sh /path/to/radix_sort   (a checkout of jgrodzki/radix_sort)
# Copies the shim's files into the crate and applies the insert-only patch to src/radix_sort/mod.rs.
set -e
here=$(cd "$(dirname "$0")" && pwd)
crate=${1:?usage: apply.sh /path/to/radix_sort}
test -f "$crate/src/radix_sort/mod.rs" || { echo "$crate is not a radix_sort checkout" >&2; exit 1; }
cp "$here/Cargo.toml" "$crate/Cargo.toml"
cp "$here/build.rs" "$crate/build.rs"
cp "$here/src/radix_sort/ffi.rs" "$here/src/radix_sort/gpu.rs" "$crate/src/radix_sort/"
patch -p1 -d "$crate" < "$here/src/radix_sort/mod.rs.patch"
echo "applied: build with  B200SORT_LIB_DIR=<dir of libb200sort.so> cargo test --features b200"
