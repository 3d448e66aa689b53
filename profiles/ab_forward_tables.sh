#!/bin/bash
# A/B of the forward kernel's table staging: LSMC_FWD_TABLE_BUFFERS=1 (staged synchronously per period) vs 2 (next period's
# tables fetched with cp.async under this period's arithmetic).  One bench line per variant and shape -> gpurun_out/ab_fwd.jsonl
OUT=gpurun_out/ab_fwd.jsonl
: > $OUT
run() {  # label, buffers, bench args...
  local label=$1 nb=$2; shift 2
  for rep in 1 2; do
    LSMC_FWD_TABLE_BUFFERS=$nb timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --e2e-steps 1 "$@" 2>> gpurun_out/ab_fwd.err \
      | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(json.dumps({'case':'$label','buffers':$nb,'ms':d['ms_per_step'],'phases':d.get('phases_last_step'),'sm_mhz':d['clocks']['sm_mhz']}))" >> $OUT
  done
}
for nb in 1 2; do
  run cfg1 $nb --config 1
  run cfg2 $nb --config 2
  run 125k_600d $nb --config 4 --sims 125000 --days 600
  run 1m_300d $nb --config 4 --days 300
done
cat $OUT
