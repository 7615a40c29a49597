#!/bin/bash
# usage: tools/sass_hist.sh <mangled-substring> : SASS opcode histogram of one kernel in libsdct.so
cuobjdump -sass simd_dct_b200/libsdct.so | awk -v pat="$1" '/Function :/{on=index($0,pat)>0} on' | grep -E "^\s+/\*[0-9a-f]{4}\*/" | sed -E 's/^\s+\/\*[0-9a-f]{4}\*\/\s+(@!?U?P[0-9T]+\s+)?//' | awk '{print $1}' | sed -E 's/\..*//' | sort | uniq -c | sort -rn
