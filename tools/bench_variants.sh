#!/bin/sh
# Time the output stage of several library variants (tools/build_variant.sh) in one GPU call:
#   sh tools/bench_variants.sh OUT.jsonl NAME1 NAME2 ... [-- extra bench.py arguments]
OUT=$1; shift
: > "$OUT"
NAMES=""
while [ $# -gt 0 ] && [ "$1" != "--" ]; do NAMES="$NAMES $1"; shift; done
[ "$1" = "--" ] && shift
for n in $NAMES; do
  lib=build/lib_$n.so
  [ "$n" = "base" ] && lib=image_codecs_b200/libimage_codecs_b200.so
  IMAGE_CODECS_B200_LIB=$lib python bench.py --no-e2e --no-cpu --steps 10 --warmup 3 "$@" 2>>"$OUT.err" | python -c "
import json,sys
d=json.loads(sys.stdin.readline())
print(json.dumps({'variant':'$n','idct_ms':d['stages']['idct_color_ms'],'entropy_ms':d['stages']['entropy_ms'],'frac':d['roofline']['frac'],'check':d['config'].get('pixel_check')}))" >> "$OUT" || echo "{\"variant\":\"$n\",\"failed\":1}" >> "$OUT"
done
cat "$OUT"
