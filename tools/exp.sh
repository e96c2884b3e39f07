#!/bin/bash
# Development aid: A/B runs of bench.py on one GPU box.  Every argument is "name;ENV=val ENV2=val;path/to/lib.so" (the last two
# may be empty); prints one line per run: name, decode tokens/s, end-to-end tokens/s, roofline fraction.
#   tools/exp.sh "base;;" "pf128k;B2_PF_EARLY=131072;" "nowait;;tools/bin/lib_nowait.so"
# Extra arguments for bench.py go into EXP_ARGS (default: llama3.1:8b, 3 steps, 2 warm-up).
cd "$(dirname "$0")/.."
ARGS=${EXP_ARGS:---steps 3 --warmup 2 --no-cpu-baseline}
for spec in "$@"; do
  IFS=';' read -r name envs lib <<< "$spec"
  libenv=""
  [ -n "$lib" ] && libenv="B2_LIB=$PWD/$lib"
  out=$(env $envs $libenv python bench.py $ARGS 2>/tmp/exp_err.log | tail -1)
  python - "$name" "$out" <<'EOF'
import json, sys
name, out = sys.argv[1], sys.argv[2]
try:
    d = json.loads(out)
    print(f"{name:28s} decode {d['value']:8.1f} tok/s   e2e {d['e2e']['value']:8.1f}   frac {d['roofline']['frac']:.4f}   J/tok {d.get('joules_per_token') or 0:.3f}   sm {d['clocks']['sm_mhz']}", flush=True)
except Exception as ex:
    print(f"{name:28s} FAILED ({ex}): {out[:200]}", flush=True)
EOF
done
