#!/usr/bin/env bash
# usage: sass_mix.sh <lib.so> <mangled-name-substring> [top]  -> static opcode histogram of one kernel
cuobjdump -sass "$1" 2>/dev/null | awk -v pat="$2" '/Function :/{f=(index($0,pat)>0)} f{print}' \
 | grep -oE "^\s+/\*[0-9a-f]{4,}\*/\s+(@!?U?P[0-9T]+\s+)?[A-Z0-9_.]+" | awk '{print $NF}' | sed -E 's/^(IMAD\.WIDE|IMAD\.MOV|IMAD\.IADD|IMAD\.X|IMAD\.SHL|IMAD\.HI)[A-Z0-9.]*/\1/; t; s/\..*//' | sort | uniq -c | sort -rn | head -${3:-24}
