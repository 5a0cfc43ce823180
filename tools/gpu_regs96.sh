#!/bin/bash
cd "${GRAFT_REPO_ROOT:-.}"; mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_fullsize.py -m gpu -x -q -k cfg2 2>&1 | tail -3
L="pl_low_cfg2_w12,pl96,pl96_c4w2,ov_16_6_c4w2_plf"
for sm in 12 13 16 20; do for cw in c4w2 c3w2 c4w3; do L="$L,ov96_${sm}_6_${cw}_plf"; done; done
L="$L,ov96_20_8_c4w2_plf,ov96_20_4_c4w2_plf,ov96_20_6_c5w2_plf"
timeout 600 python tools/bench_stages.py --reps 20 --only $L > gpurun_out/r02_regs96.jsonl 2> gpurun_out/r02_regs96.err; echo rc=$?
python - <<'PY'
import json
for l in open('gpurun_out/r02_regs96.jsonl'):
    try: d=json.loads(l)
    except Exception: continue
    print('%-44s %8.2f us (min %8.2f)'%(d['stage'], d['us_median'] if 'us_median' in d else d.get('us',0), d.get('us_min',0)))
PY
tail -3 gpurun_out/r02_regs96.err
