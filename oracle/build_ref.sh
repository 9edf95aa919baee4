#!/bin/bash
# TEST INFRASTRUCTURE — not part of the product.
#
# Builds oracle/_ref/js: a JavaScript shell made from the SpiderMonkey 1.x C engine
# that the reference vendors under deps/javascriptlint/spidermonkey/src (it is a
# dependency of the reference's lint tool).  The shell lets this container run the
# reference's UNMODIFIED lib/zk-consts.js, lib/jute-buffer.js, lib/zk-buffer.js and
# lib/zk-streams.js (see oracle/ref_js.py) so that the C restatement in
# oracle/zkoracle.c can be pinned against the reference's real behaviour, and so
# that golden fixtures under tests/golden/ can be (re)generated.
#
# The sources are compiled where they lie under /root/reference with plain gcc
# (the reference's own Makefile is not run); every output goes to oracle/_ref/.
# Only usable where /root/reference exists (the build container); the GPU box
# never needs it because it only consumes the committed fixtures.
set -euo pipefail
REF=${ZK_REFERENCE:-/root/reference}
SRC=$REF/deps/javascriptlint/spidermonkey/src
HERE=$(cd "$(dirname "$0")" && pwd)
OUT=$HERE/_ref
OBJ=$OUT/obj
if [ ! -d "$SRC" ]; then echo "build_ref: $SRC not present; skipping" >&2; exit 0; fi
if [ -x "$OUT/js" ] && [ "${1:-}" != "--force" ]; then exit 0; fi
mkdir -p "$OBJ"
CF="-fno-strict-aliasing -O1 -w -std=gnu89 -fcommon -DXP_UNIX -DSVR4 -DSYSV -DHAVE_LOCALTIME_R -UDEBUG -DNDEBUG -I$OBJ -I$SRC"
# generated headers (byte order / keyword table) come from two small generators
gcc $CF -o "$OBJ/jscpucfg" "$SRC/jscpucfg.c" && "$OBJ/jscpucfg" > "$OBJ/jsautocfg.h"
gcc $CF -o "$OBJ/jskwgen"  "$SRC/jskwgen.c"  && "$OBJ/jskwgen"  > "$OBJ/jsautokw.h"
MODS="jsapi jsarena jsarray jsatom jsbool jscntxt jsdate jsdbgapi jsdhash jsdtoa jsemit
      jsexn jsfun jsgc jshash jsinterp jsiter jslock jslog2 jslong jsmath jsnum jsobj
      jsopcode jsparse jsprf jsregexp jsscan jsscope jsscript jsstr jsutil jsxdrapi
      jsxml prmjtime js"
pids=()
for m in $MODS; do
  gcc $CF -c "$SRC/$m.c" -o "$OBJ/$m.o" &
  pids+=($!)
done
for p in "${pids[@]}"; do wait "$p"; done
objs=""; for m in $MODS; do objs="$objs $OBJ/$m.o"; done
gcc -o "$OUT/js" $objs -lm
echo "build_ref: built $OUT/js"
