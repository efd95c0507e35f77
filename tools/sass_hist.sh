#!/bin/bash
# usage: sass_hist.sh <obj> <function-substring>  -> opcode histogram of that kernel's SASS
cuobjdump -sass "$1" | awk -v pat="$2" '/Function : /{f=($3 ~ pat)} f && /^ +\/\*[0-9a-f]{4}\*\//{op=$2; if(op ~ /^@/) op=$3; sub(/[.;].*/,"",op); print op}' | sort | uniq -c | sort -rn | head -${3:-25}
