#!/bin/sh
# One bench line per BASELINE.json config (C1..C5 shapes on one GPU), each with its CPU baseline, for profiles/.
# Run on the GPU box:   sh tools/run_configs.sh gpurun_out/configs.jsonl
OUT=${1:-gpurun_out/configs.jsonl}
: > "$OUT"
run() { echo "== $*" >&2; timeout 900 python bench.py "$@" >> "$OUT" 2>> "$OUT.err" || echo "{\"failed\": \"$*\"}" >> "$OUT"; }
run --config c1                                                              # C1: one 512x512, per-image call, latency
run --config c2                                                              # C2: 1024 x 1080p 4:2:0 q90 restart 16
run --config c3                                                              # C3: 4:4:4 + grayscale without restart markers (self-sync)
run --config c3 --mix-gray 0 --subsampling 444                               # C3, colour half alone
run --config c3 --mix-gray 0 --gray 1                                        # C3, grayscale half alone
run --config c3 --mix-gray 0 --subsampling 444 --restart 16                  # 4:4:4 with restart markers (for comparison)
run --config c3 --mix-gray 0 --subsampling 420                               # 4:2:0 without restart markers: the commonest file type (self-sync + 4:2:0 strip kernel)
run --config c4                                                              # C4: progressive 1080p without restart markers
run --config c4 --images 1024 --steps 2                                      # C4 at the batch size of C2
run --config c4 --images 4096 --steps 2 --no-e2e                             # C4: throughput regime (every scheduler holds several scans)
run --config c4 --restart 16                                                 # C4 with restart markers
run --config c2 --images 256 --subsampling 422 --steps 10                    # N2: 4:2:2 (what most cameras emit)
run --config c5 --total-images 1024 --steps 3                                # C5 shape on one GPU: 1024 x 2160p in chunks of 256
