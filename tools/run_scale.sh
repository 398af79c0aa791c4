#!/bin/sh
# Multi-GPU lines for profiles/: tools/run_scale.sh N OUTDIR   (run under `gpurun --gpus N`)
N=$1; OUT=${2:-gpurun_out}
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node "$N" --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus "$N" "$@"; }
python -m pytest tests -x -q -m gpu -k "multi_device or consumer" > "$OUT/scale_n${N}_tests.log" 2>&1; tail -n 2 "$OUT/scale_n${N}_tests.log"
python bench.py --impl reference --gpus "$N" --steps 3 --warmup 1 > "$OUT/r02_reference_arm_n${N}.json" 2>> "$OUT/scale_n${N}.err"
run --steps 20 --warmup 3 > "$OUT/r02_scale_n${N}.json" 2>> "$OUT/scale_n${N}.err"
run --config c5 --steps 2 --warmup 3 > "$OUT/r02_c5_n${N}.json" 2>> "$OUT/scale_n${N}.err"
python bench.py --impl reference --config c5 --gpus "$N" --steps 2 --warmup 1 > "$OUT/r02_c5_reference_arm_n${N}.json" 2>> "$OUT/scale_n${N}.err"
tail -n 5 "$OUT/scale_n${N}.err"
for f in "$OUT/r02_scale_n${N}.json" "$OUT/r02_c5_n${N}.json"; do python -c "
import json,sys
d=json.load(open('$f')); e=d['e2e']
print('$f', 'value %.0f'%d['value'], 'e2e %.0f'%e['value'], 'ceil %.1f GB/s frac %.3f'%(e['d2h_ceiling_gbs_all_ranks'], e['frac_of_d2h_ceiling']), 'pageable', (e.get('pageable') or {}).get('value'))"; done
