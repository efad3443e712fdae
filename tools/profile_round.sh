#!/bin/bash
# One gpurun call's worth of round-end evidence (run from the repo root on a B200 box):
#   bash tools/profile_round.sh <tag>     ->  gpurun_out/<tag>_*  (copy what should be judged into profiles/)
set -x
T=${1:-rX}
python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest_gpu.log 2>&1; tail -2 gpurun_out/${T}_pytest_gpu.log
python bench.py --steps 20 --warmup 3 > gpurun_out/${T}_bench_n1.json 2> gpurun_out/${T}_bench_n1.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${T}_bench_reference_n1.json 2> gpurun_out/${T}_bench_reference_n1.err
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/${T}_launches_bench_steps3.csv python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/${T}_ncu1.log 2>&1
if [ "$2" = "full" ]; then
ncu --set full --clock-control none --import-source on -k regex:k_nn_single_pruned -s 5 -c 1 -o gpurun_out/${T}_prof_nn python bench.py --steps 3 --warmup 3 --no-cpu --insert-scans 0 > gpurun_out/${T}_ncu2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_insert_rank -s 2 -c 2 -o gpurun_out/${T}_prof_rank python bench.py --only insert_batched > gpurun_out/${T}_ncu3.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_insert_commit -s 2 -c 1 -o gpurun_out/${T}_prof_commit python bench.py --only insert_batched > gpurun_out/${T}_ncu4.log 2>&1
fi
cut -c1-300 gpurun_out/${T}_bench_n1.json
