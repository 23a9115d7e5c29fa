#!/bin/bash
# Final round-2 measurement set on one B200 (run under gpurun from the repo root); every bench line lands in
# gpurun_out/ and is copied to profiles/ by hand.  Usage: bash profiles/run_r02_final.sh <tag> [a|b]
tag=${1:-s36}
part=${2:-a}
out=gpurun_out
mkdir -p $out
if [ "$part" = a ]; then
  timeout 900 python -m pytest tests -x -q -m gpu > $out/pytest_gpu_$tag.log 2>&1; tail -3 $out/pytest_gpu_$tag.log
  timeout 120 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -1
  timeout 300 python bench.py > $out/r02_bench_default_$tag.json 2> $out/err_default.log; tail -c 400 $out/r02_bench_default_$tag.json
  timeout 300 python bench.py --impl reference --steps 5 --warmup 3 > $out/r02_bench_reference_$tag.json 2> $out/err_ref.log; tail -c 300 $out/r02_bench_reference_$tag.json
  for w in cfg1 cfg2; do timeout 200 python bench.py --workload $w --steps 200 --warmup 5 > $out/r02_bench_${w}_$tag.json 2> $out/err_$w.log; grep -o '"ms_per_step": [0-9.]*' $out/r02_bench_${w}_$tag.json | head -1; done
  timeout 200 python bench.py --batch 100 --steps 200 --warmup 5 --no-aux > $out/r02_bench_cfg3_B100_$tag.json 2> $out/err_b100.log; grep -o '"ms_per_step": [0-9.]*' $out/r02_bench_cfg3_B100_$tag.json | head -1
  for w in cfg1_fwd cfg3_fwd cfg4 cfg5 metrics; do timeout 300 python bench.py --workload $w --steps 5 --warmup 3 > $out/r02_bench_${w}_$tag.json 2> $out/err_$w.log; grep -o '"ms_per_step": [0-9.]*' $out/r02_bench_${w}_$tag.json | head -1; done
else
  timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/r02_launches_cfg3_B4096_$tag.csv python bench.py --steps 2 --warmup 3 --no-aux > $out/ncu_launches.log 2>&1
  # the report itself is too large to travel back (64 MiB limit): summaries are made here, the report stays on the box
  timeout 600 ncu --set full --clock-control none -c 60 -o /tmp/prof_$tag -f python bench.py --steps 1 --warmup 3 --no-aux --no-graph > $out/ncu_full.log 2>&1
  python profiles/ncu_summary.py /tmp/prof_$tag.ncu-rep > $out/r02_ncu_full_$tag.txt 2>&1
  python profiles/make_traffic_json.py /tmp/prof_$tag.ncu-rep cfg3 4096 > $out/r02_ncu_traffic_$tag.json 2>$out/err_traffic.log
  timeout 400 ncu --set full --clock-control none -c 40 --kernel-name-base demangled --kernel-name regex:"scan_tc_kernel|site_pfm_tc" -o /tmp/prof_interp_$tag -f python bench.py --workload cfg4 --steps 1 --warmup 0 --no-aux > $out/ncu_interp.log 2>&1
  python profiles/ncu_summary.py /tmp/prof_interp_$tag.ncu-rep > $out/r02_ncu_full_interpret_$tag.txt 2>&1
  ls -la /tmp/prof_*.ncu-rep; wc -l $out/r02_ncu_full_$tag.txt $out/r02_ncu_full_interpret_$tag.txt
fi
