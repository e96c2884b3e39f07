#!/bin/bash
# The reference's run table through the UNCHANGED runner (baseline/_ref = a copy of the replication package, git-ignored) against
# the B200 server: one repetition (7 models x 2 methods x 3 lengths = 42 rows), no cool-down, on 1 GPU with one runner process
# and on N GPUs with N runner processes that all call port 11434.  Writes the run tables and wall times under gpurun_out/.
#   tools/run_sweep.sh 8
set -u
cd "$(dirname "$0")/.."
N=${1:-8}
REPS=${2:-1}
ONLY=${3:-both}
OUT=gpurun_out/r02_sweep
mkdir -p $OUT
ALL="llama3.1:8b,gemma:2b,gemma:7b,phi3:3.8b,qwen2:1.5b,qwen2:7b,mistral:7b"
run() {  # run <shards> <devices>
  local k=$1 devs=$2 work=/tmp/sweep_$1
  rm -rf $work
  local t0=$(date +%s.%N)
  python -m cain_2025_device_remote_llm_energy_rep_pkg_b200.runner.launch --reference-dir baseline/_ref --work-dir $work --name b200 \
      --cooldown-ms 0 --repetitions $REPS --shards $k --serve-devices $devs --preload "$ALL" --seed 0 > $OUT/launch_$k.log 2>&1
  local rc=$?
  local t1=$(date +%s.%N)
  echo "shards=$k devices=$devs rc=$rc total_with_server_start_s=$(python -c "print(round($t1-$t0,1))") $(grep 'sweep wall time' $OUT/launch_$k.log)" | tee -a $OUT/summary.txt
  if [ $k -gt 1 ]; then cp $work/run_table.csv $OUT/run_table_${k}gpu.csv 2>/dev/null; else cp $work/experiments_output/b200/run_table.csv $OUT/run_table_1gpu.csv 2>/dev/null; fi
  cp $work/server.log $OUT/server_$k.log 2>/dev/null
  cp $work/server_stats.jsonl $OUT/server_stats_$k.jsonl 2>/dev/null
  for i in 0 1; do cp $work/shard$i/runner.log $OUT/runner_${k}_shard$i.log 2>/dev/null; done
  tail -3 $work/shard0/runner.log $work/runner.log 2>/dev/null | tail -5 >> $OUT/launch_$k.log
}
[ "$ONLY" = both ] && : > $OUT/summary.txt
if [ "$ONLY" = both ] || [ "$N" -eq 1 ]; then run 1 0; fi
if [ "$N" -gt 1 ]; then run $N 0-$((N-1)); fi
python - <<PY
import csv, json, collections
out = {}
for k in (1, $N):
    try:
        rows = list(csv.DictReader(open("$OUT/run_table_%dgpu.csv" % k)))
    except OSError:
        continue
    empty = collections.Counter(c for r in rows for c, v in r.items() if str(v).strip() == "")
    toks = sum(float(r["nvml__eval_count"] or 0) for r in rows)
    out[k] = {"rows": len(rows), "done": sum(r["__done"] == "DONE" for r in rows), "tokens": toks, "empty_cells_by_column": dict(empty),
              "request_energy_J": round(sum(float(r["nvml__request_energy_J"] or 0) for r in rows), 1)}
print(json.dumps(out))
open("$OUT/summary.json", "w").write(json.dumps(out, indent=1))
PY
