#!/bin/bash
# usage: tools/sass_fn.sh <mangled-function-name> [out-file]   -- SASS of one kernel of libopb200.so, encodings stripped
cd "$(dirname "$0")/.."
out=${2:-/tmp/fn.sass}
cuobjdump -sass -fun "$1" onepose_st_b200/libopb200.so 2>/dev/null | grep -E "^\s+/\*[0-9a-f]{4,5}\*/" | sed 's#/\* 0x[0-9a-f]* \*/##' > "$out"
wc -l "$out"
