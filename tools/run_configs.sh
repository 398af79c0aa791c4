#!/bin/sh
# One bench line per BASELINE.json config that fits one GPU (C2..C5 shapes), for profiles/.  Run on the GPU box:
#   sh tools/run_configs.sh gpurun_out/configs.jsonl
OUT=${1:-gpurun_out/configs.jsonl}
: > "$OUT"
run() { echo "== $*" >&2; timeout 400 python bench.py --no-cpu --steps 5 --warmup 3 --e2e-steps 1 "$@" >> "$OUT" 2>> "$OUT.err" || echo "{\"failed\": \"$*\"}" >> "$OUT"; }
run --images 1024                                              # C2: 1080p 4:2:0 q90 restart 16
run --images 256 --subsampling 444 --restart 0                 # C3: 4:4:4 without restart markers (self-sync)
run --images 256 --gray 1 --restart 0                          # C3: grayscale without restart markers (self-sync)
run --images 256 --subsampling 444 --restart 16                # 4:4:4 with restart markers (for comparison)
run --images 256 --progressive 1 --restart 0                   # C4: progressive 1080p
run --images 256 --progressive 1 --restart 16                  # C4 with restart markers
run --images 256 --width 3840 --height 2160 --restart 16       # C5 shape: 2160p 4:2:0 (per-GPU shard)
