#!/bin/bash
# usage: scripts/sass_info.sh <object or cubin> <kernel-name regex>
# prints registers and the run-length encoded sequence of memory / FP64 / barrier opcodes per matching kernel
f=$1; pat=$2
cuobjdump -res-usage "$f" 2>/dev/null | paste - - | grep -E "$pat" | sed -E 's/.*Function ([^:]*):.*REG:([0-9]+).*SHARED:([0-9]+).*/\1 REG=\2 SHARED=\3/' | while read n rest; do echo "$(echo $n | c++filt | cut -c1-100) $rest"; done
if [ -n "$3" ]; then
for fn in $(cuobjdump -res-usage "$f" 2>/dev/null | grep -oE "Function [^:]*" | awk '{print $2}' | grep -E "$pat" | head -${3}); do
  echo "--- $(echo $fn | c++filt | cut -c1-100)"
  cuobjdump -sass -fun "$fn" "$f" | grep -oE "\b(LDG|STG|DMUL|DADD|DFMA|BAR|WARPSYNC|BRA|LDS|STS|LDGSTS|LDGDEPBAR|DEPBAR|UBLKCP|SYNCS|ATOMS|ATOMG|RED|SHFL|MATCH|VOTE)[A-Z0-9_.]*" | sed -E 's/\..*//' | uniq -c | awk '{printf "%s%s ", ($1>1?$1"x":""), $2} END{print ""}'
done
fi
