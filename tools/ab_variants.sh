#!/bin/bash
# A/B timing of kernel build variants on one GPU box (tools: MSIK_VARIANT=... python maniskill_b200/build.py).
# usage: bash tools/ab_variants.sh "name[:ENV=V,ENV=V]" ...   ("new" = the product library)
for spec in "$@"; do
  v=${spec%%:*}; envs=""; [ "$spec" != "$v" ] && envs=${spec#*:}
  (
    if [ "$v" != new ]; then export MSIK_LIB=$PWD/maniskill_b200/_build/variants/$v/libmsik.so; fi
    for kv in ${envs//,/ }; do export "$kv"; done
    tag=$(echo "$spec" | tr ':=,' '___')
    python bench.py --steps 300 --warmup 10 --no-extras --no-sweep > gpurun_out/ab_$tag.json 2> gpurun_out/ab_$tag.err || echo FAIL $spec
    python - "$spec" gpurun_out/ab_$tag.json <<'PY'
import json, sys
d = json.load(open(sys.argv[2]))
print(sys.argv[1], "us/step", round(d["ms_per_step"] * 1e3, 2), "frac", round(d["roofline"]["frac"], 4), d["clocks"]["sm_mhz"])
PY
  )
done
