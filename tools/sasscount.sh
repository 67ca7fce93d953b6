#!/bin/bash
# usage: tools/sasscount.sh <object> [filter]  -- SASS instruction count (and FFMA2 count) per kernel
O=/root/repo/mlt_b200/build/$1.o
cuobjdump -sass $O | grep -E "Function :|^\s+/\*[0-9a-f]{4,5}\*/" | awk '/Function/{if(n)print name, n, f; name=$3; n=0; f=0; next}{n++; if($0 ~ /FFMA2/)f++}END{print name,n,f}' | grep -E "${2:-.}" | c++filt | sed 's/(anonymous namespace):://g'
grep -E "Compiling|Used" /root/repo/mlt_b200/build/$1.ptxas.log | paste - - | grep -E "${2:-.}" | sed -E "s/.*function '([^']*)'.*Used ([0-9]+) registers.*/\2 regs \1/" | c++filt | sed 's/(anonymous namespace):://g'
