#!/bin/bash
# usage: tools/sass_fn.sh <lib.so> <substring of the mangled kernel name>: prints the SASS instructions of that function
cuobjdump -sass "$1" | awk -v pat="$2" '/Function :/ {on = index($0, pat) > 0} on && /^ +\/\*[0-9a-f]+\*\// {sub(/^ +\/\*[0-9a-f]+\*\/ +/, ""); sub(/ *\/\*.*$/, ""); print}'
