#!/bin/bash
# Scaling sweep on ONE box: ours and the reconstructed reference path at 1/2/4/8 sites, then the aggregation
# micro-benchmark and the low-rank configs at the largest N.  Lines go to gpurun_out/scaling/.
#   bash benchmarks/run_scaling.sh [MAXN] [part]      part = all | main | extra
set -u
MAXN=${1:-8}
PART=${2:-all}
OUT=gpurun_out/scaling
mkdir -p $OUT
run() {   # run N impl extra-args...
  local n=$1; shift; local impl=$1; shift
  local port=$((29600 + RANDOM % 300))
  if [ "$n" = 1 ]; then
    timeout 600 python bench.py --gpus 1 --impl $impl "$@" 2>$OUT/${impl}_$n.err | tail -1 > $OUT/${impl}_$n.json
  else
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port \
      bench.py --gpus $n --impl $impl "$@" 2>$OUT/${impl}_$n.err | tail -1 > $OUT/${impl}_$n.json
  fi
  python - <<PY
import json
try:
    d = json.load(open("$OUT/${impl}_$n.json"))
    print("$impl", d.get("n_gpus"), "value", round(d["value"], 1), "ms/step", round(d["ms_per_step"], 4), "e2e", d.get("e2e", {}).get("value") if isinstance(d.get("e2e"), dict) else None)
except Exception as e:
    print("$impl $n FAILED", e)
PY
}
if [ "$PART" != extra ]; then
  for n in 1 2 4 8; do [ $n -le $MAXN ] && run $n ours; done
  for n in 1 2 4 8; do [ $n -le $MAXN ] && run $n recon-nccl --steps 30 --warmup 3; done
fi
if [ "$PART" != main ]; then
  if [ -z "${SKIP_MICRO:-}" ]; then
    port=$((29600 + RANDOM % 300))
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $MAXN --master-addr 127.0.0.1 --master-port $port \
      benchmarks/agg_microbench.py > $OUT/agg_microbench_$MAXN.jsonl 2>$OUT/agg.err; tail -3 $OUT/agg_microbench_$MAXN.jsonl | cut -c1-200
  fi
  port=$((29600 + RANDOM % 300))
  for cfg in ica-rankdad mlp-powersgd; do
    port=$((29600 + RANDOM % 300))
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $MAXN --master-addr 127.0.0.1 --master-port $port \
      benchmarks/lowrank_bench.py --config $cfg >> $OUT/lowrank_$MAXN.jsonl 2>>$OUT/lowrank.err
  done
  tail -4 $OUT/lowrank_$MAXN.jsonl | cut -c1-250
  port=$((29600 + RANDOM % 300))
  FED_TMP=/tmp/fed_scaling timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $MAXN --master-addr 127.0.0.1 --master-port $port \
    tests/multigpu/mp_federation_check.py 2>$OUT/fed.err | grep federation > $OUT/federation_$MAXN.txt; cat $OUT/federation_$MAXN.txt
fi
