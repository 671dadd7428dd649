#!/bin/bash
# SASS listing + opcode histogram of one kernel of an object file.
# usage: scripts/sass_hist.sh <object.o> <function-name regex> [listing-out]
# prints the opcode histogram (predicates stripped); writes the plain listing to $3 when given
OBJ=$1; RE=$2; OUT=${3:-/dev/null}
cuobjdump -sass "$OBJ" | awk -v re="$RE" '
  /Function :/ { p = ($0 ~ re) }
  p && /^[ \t]+\/\*[0-9a-f]{4}\*\// { line=$0; sub(/^[ \t]*\/\*[0-9a-f]+\*\/[ \t]*/, "", line); sub(/[ \t]*\/\*.*$/, "", line); print line }' > /tmp/sass_listing.$$
cp /tmp/sass_listing.$$ "$OUT" 2>/dev/null
echo "instructions: $(wc -l < /tmp/sass_listing.$$)"
sed -E 's/^@!?U?P[0-9T]+ +//' /tmp/sass_listing.$$ | awk '{print $1}' | sed -E 's/\..*$//' | sort | uniq -c | sort -rn
rm -f /tmp/sass_listing.$$
