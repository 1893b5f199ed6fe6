#!/bin/bash
# usage: tools/sass_kernel.sh <object> <function-name-regex>  -> prints the SASS of the first matching function, one instruction per line
obj=$1; pat=$2
cuobjdump -sass "$obj" | awk -v pat="$pat" '/Function :/{on=($0 ~ pat)} on' | grep -E "^\s+/\*[0-9a-f]{4}\*/" | sed -E 's/^\s+\/\*([0-9a-f]+)\*\/\s+/\1 /; s/\s*\/\*.*$//'
