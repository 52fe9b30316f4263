#!/bin/bash
# usage: tools/sass_ops.sh <mangled-name-substring>  -> opcode histogram of that kernel in libmargin_b200.so
cuobjdump -sass marginalign_trna_b200/libmargin_b200.so 2>/dev/null | awk -v pat="$1" '
/Function : /{f = index($0, pat) > 0}
f && /^ +\/\*[0-9a-f]+\*\/ /{print}' > /tmp/sass_fn.txt
wc -l < /tmp/sass_fn.txt
sed -E 's/^ +\/\*[0-9a-f]+\*\/ +//' /tmp/sass_fn.txt | sed -E 's/^@!?U?P[0-9T]+ +//' | awk '{split($1,a,"."); c[a[1]]++} END{for(k in c) print c[k],k}' | sort -rn | head -40
