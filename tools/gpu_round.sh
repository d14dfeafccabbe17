#!/bin/bash
# One parameterised GPU round (replaces the per-round one-off scripts):  tools/gpu_round.sh <tag> [step ...]
# steps: pytest | bench | smoke | refscripts | launches[:<name>:<bench args>] | ncu:<kernel regex>:<name>:<bench args> ; outputs under gpurun_out/<tag>/
tag=$1; shift; O=gpurun_out/$tag; mkdir -p $O
B="python bench.py --steps 2 --warmup 3 --windows 1 --no-configs --no-sweep --no-cpu-baseline --no-e2e --launch eager"
for s in "$@"; do
  case $s in
    pytest) timeout 2400 python -m pytest tests -m gpu -q -x > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/rc.txt; tail -4 $O/pytest.log;;
    bench) timeout 1500 python bench.py > $O/bench.json 2> $O/bench.err; echo "bench rc=$?" >> $O/rc.txt;;
    smoke) timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?" >> $O/rc.txt;;
    refscripts) timeout 1800 python tools/run_reference_scripts.py --bench --out $O/reference_scripts.txt > $O/refscripts.log 2>&1; echo "refscripts rc=$?" >> $O/rc.txt;;
    launches) timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/launches.csv $B > $O/ncu_launches.log 2>&1; echo "launches rc=$?" >> $O/rc.txt;;
    launches:*) IFS=: read -r _ name args <<< "$s"   # launches:<name>:<bench args>, graph replay as the bench runs it
       timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file $O/launches_$name.csv ${B/--launch eager/} $args > $O/ncu_launches_$name.log 2>&1; echo "launches $name rc=$?" >> $O/rc.txt;;
    ncu:*) IFS=: read -r _ rx name args <<< "$s"
       timeout 900 ncu --set full --clock-control none --import-source on -k regex:$rx -s 4 -c 2 -o $O/$name $B $args > $O/ncu_$name.log 2>&1; echo "ncu $name rc=$?" >> $O/rc.txt;;
    *) echo "unknown step $s";;
  esac
done
cat $O/rc.txt
