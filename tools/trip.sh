#!/bin/bash
# One parameterised GPU trip: tools/trip.sh <tag> [pytest|bench|ncu ...]; logs under gpurun_out/<tag>_*
tag=$1; shift
mkdir -p gpurun_out
for what in "$@"; do
  case $what in
    pytest) timeout 2400 python -m pytest tests -m gpu -q -x --timeout=1500 > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest rc=$?" ;;
    pytestall) timeout 3000 python -m pytest tests -m gpu -q --timeout=1500 > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest rc=$?" ;;
    bench) timeout 900 python bench.py --layers gpurun_out/${tag}_layers.md > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc=$?" ;;
    benchfast) timeout 600 python bench.py --no-extras --no-cpu-baseline --layers gpurun_out/${tag}_layers.md > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc=$?" ;;
    ref) timeout 600 python bench.py --impl reference --steps 2 --warmup 0 > gpurun_out/${tag}_ref.json 2> gpurun_out/${tag}_ref.err; echo "ref rc=$?" ;;
    smoke) timeout 600 python __graft_entry__.py smoke > gpurun_out/${tag}_smoke.log 2>&1; echo "smoke rc=$?" ;;
    launches) timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file gpurun_out/${tag}_launches.csv python bench.py --steps 1 --warmup 3 --lanes 1 --no-graph --no-extras --no-cpu-baseline > gpurun_out/${tag}_ncu.log 2>&1; echo "launches rc=$?" ;;
    kern) timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"${KERN:-stft}" -c 12 --csv --log-file gpurun_out/${tag}_kern.csv python bench.py --steps 1 --warmup 3 --lanes 1 --no-graph --no-extras --no-cpu-baseline > gpurun_out/${tag}_kern.log 2>&1; echo "kern rc=$?"; grep -v "^==" gpurun_out/${tag}_kern.csv | awk -F'","' 'NR>1{print $5, $(NF)}' | tail -6 ;;
    full) timeout 900 ncu --set full --import-source on --clock-control none -k regex:"${KERN:-stft}" -s ${SKIP:-4} -c ${COUNT:-1} -o gpurun_out/${tag}_full -f python bench.py --steps 1 --warmup 3 --lanes 1 --no-graph --no-extras --no-cpu-baseline > gpurun_out/${tag}_full.log 2>&1; echo "full rc=$?"; ncu -i gpurun_out/${tag}_full.ncu-rep --page details > gpurun_out/${tag}_full.txt 2>&1 ;;
    *) echo "unknown step $what" ;;
  esac
done
tail -3 gpurun_out/${tag}_pytest.log 2>/dev/null
head -c 1500 gpurun_out/${tag}_bench.json 2>/dev/null
