#!/bin/bash
# launch list of the default bench command (after it has exited 0 without ncu): ONE ncu pass, bounded launch count
O=gpurun_out; mkdir -p $O; TAG=${1:-r2zb}
timeout 600 python bench.py --steps 2 --warmup 1 --no-cpu-baseline > $O/${TAG}_plain.json 2> $O/${TAG}_plain.err || { echo "plain run failed"; tail -5 $O/${TAG}_plain.err; exit 1; }
cut -c1-200 $O/${TAG}_plain.json
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file $O/${TAG}_launches_default.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > $O/${TAG}_ncu.log 2>&1; echo "ncu rc=$?"
python scripts/summarize_launches.py $O/${TAG}_launches_default.csv > $O/${TAG}_launches_default_summary.txt; head -14 $O/${TAG}_launches_default_summary.txt
