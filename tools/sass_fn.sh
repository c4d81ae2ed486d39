#!/bin/bash
# usage: tools/sass_fn.sh <lib.so> <kernel-name-substring>   -- SASS of one kernel, one instruction per line
cuobjdump -sass "$1" | awk -v pat="$2" '/Function :/{p = index($0, pat) > 0} p' | grep -E "^\s+/\*[0-9a-f]{4}\*/" | sed 's/^\s*//; s/\s*\/\* 0x[0-9a-f]* \*\///'
