#!/bin/bash
# usage: tools/sass_hist.sh <object-or-so> <mangled-kernel-substring>   -- static opcode histogram of one kernel
OBJ=$1; PAT=$2
FN=$(cuobjdump -sass $OBJ | grep -oE "Function : [A-Za-z0-9_]*${PAT}[A-Za-z0-9_]*" | head -1 | awk '{print $3}')
echo "kernel: $(echo $FN | c++filt)"
cuobjdump -sass -fun "$FN" $OBJ | grep -oE "^\s+/\*[0-9a-f]+\*/\s+(@!?U?P[0-9T]+ )?[A-Z0-9_.]+" | awk '{print $NF}' | sed -E 's/^(LDS|STS|LDG|STG|LDL|STL)(\.[A-Z0-9.]*)?$/\1\2/; t; s/\..*//' | sort | uniq -c | sort -rn | awk '{printf "%s %s, ", $2, $1} END {print ""}'
