#!/bin/bash
# static SASS instruction count + opcode histogram of one kernel in libptx.so:  tools/sass_count.sh k_path_loopILi0E
cuobjdump -sass -fun "$(cuobjdump -elf cuda-pathtracer_b200/libptx.so 2>/dev/null | grep -o "_ZN3ptx[A-Za-z0-9_]*$1[A-Za-z0-9_]*" | head -1)" cuda-pathtracer_b200/libptx.so 2>/dev/null | grep -E '^\s+/\*[0-9a-f]{4}\*/' | sed -E 's/^\s+\/\*[0-9a-f]+\*\/\s+//' | awk '{op=$1; if (op ~ /^@/) op=$2; split(op,a,"."); c[a[1]]++; n++} END {printf "total %d\n", n; for (k in c) printf "%s %d\n", k, c[k] | "sort -k2 -nr | head -25"}'
