#!/bin/bash
# usage: profiles/ab_library.sh "<tag> [<tag> ...]" [bench args...]   — A/B of tuning builds (LSMC_BUILD_TAG=<tag> python
# __graft_entry__.py -> storage_b200/lib/liblsmc_b200.<tag>.so) against the default library on the same box, alternating;
# one short line per run -> gpurun_out/ab_library.jsonl
TAGS=$1; shift
ARGS=${@:---config 4 --days 300}
OUT=gpurun_out/ab_library.jsonl
: > $OUT
for rep in $(seq 1 ${REPS:-3}); do
  for lib in default $TAGS; do
    if [ "$lib" = "default" ]; then unset LSMC_LIBRARY; else export LSMC_LIBRARY=$PWD/storage_b200/lib/liblsmc_b200.$lib.so; fi
    timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --e2e-steps 1 $ARGS 2>> gpurun_out/ab_library.err \
      | python -c "import sys,json; d=json.loads(sys.stdin.read()); r=d['roofline']; print(json.dumps({'lib':'$lib','ms':round(d['ms_per_step'],2),'sweep_ms':round(r['avg_launch_ms'],4),'sweep_frac':round(r['frac'],4),'gram':round(d['phases_last_step']['gram_cholesky_ms'],2),'fwd':round(d['phases_last_step']['forward_ms'],2),'bwd':round(d['phases_last_step']['backward_ms'],2),'sm_mhz':d['clocks']['sm_mhz'],'npv':d.get('npv')}))" >> $OUT
  done
done
unset LSMC_LIBRARY
cat $OUT
