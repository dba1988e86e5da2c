#!/bin/bash
# Every bench workload at HEAD on one GPU, lines stored under gpurun_out/final/ (copied to profiles/ by hand).
#   bash tools/run_all_benches.sh [tag]
set -u
tag=${1:-final}
out=gpurun_out/$tag
mkdir -p $out
python bench.py --steps 20 --warmup 5 > $out/train.json 2> $out/train.err
for wl in aff_mini_infer aff_tiny_1_5th_infer maskfiner_ud_small_1024x2048_infer; do
  python bench.py --workload $wl --steps 20 --warmup 5 > $out/$wl.json 2> $out/$wl.err
done
python bench.py --workload aff_base_384_infer --steps 20 --warmup 5 > $out/aff_base_384_infer.json 2> $out/aff_base_384_infer.err
python bench.py --workload aff_base_384_infer --batch 64 --steps 20 --warmup 5 --no-cpu-baseline > $out/aff_base_384_infer_b64.json 2> $out/aff_base_384_infer_b64.err
python bench.py --impl reference --steps 2 --warmup 1 > $out/reference_arm.json 2> $out/reference_arm.err
for f in $out/*.json; do python - "$f" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    r = d.get("roofline") or {}
    print(sys.argv[1].split("/")[-1], round(d["value"], 1), d["unit"], "ms/step", round(d["ms_per_step"], 2), "e2e", round(d["e2e"]["value"], 1),
          "roofline", r.get("kernel"), r.get("frac") and round(r["frac"], 3), "traffic", r.get("traffic"), "cpu", (d.get("cpu_baseline") or {}).get("value"))
except Exception as e:
    print(sys.argv[1], "unreadable:", e)
PY
done
