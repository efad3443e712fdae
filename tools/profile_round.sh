#!/bin/bash
# One gpurun call's worth of round-end evidence (run from the repo root on a B200 box):
#   bash tools/profile_round.sh <tag> [full]    ->  gpurun_out/<tag>_*
# With `full`, the ncu captures are exported (profiles/r2_ncu_full_*.csv, copies in gpurun_out/) and summarised into
# profiles/traffic.json ON THE BOX before the bench line is taken, so the line quotes counters of the same sources
# (traffic.json carries a hash of the kernel sources; bench.py quotes it only while the sources are the same). Afterwards,
# in the build container: copy gpurun_out/<tag>_* into profiles/ (r2_* names) and gpurun_out/<tag>_traffic.json to
# profiles/traffic.json.
set -x
T=${1:-rX}
python -m pytest tests -m gpu -q > gpurun_out/${T}_pytest_gpu.log 2>&1; tail -2 gpurun_out/${T}_pytest_gpu.log
python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/${T}_bench_reference_n1.json 2> gpurun_out/${T}_bench_reference_n1.err
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/${T}_launches_bench_steps3.csv python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/${T}_ncu1.log 2>&1
if [ "$2" = "full" ]; then
# the three kernels of one nn_single step (launches 10-12 of tools/nn_ab.py: after the warm-up steps)
AB_STEPS=3 ncu --set full --clock-control none --import-source on -k regex:k_nn_ -s 9 -c 3 -o gpurun_out/${T}_prof_nn python tools/nn_ab.py nn_kernel=0 > gpurun_out/${T}_ncu2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_insert_rank -s 2 -c 2 -o gpurun_out/${T}_prof_rank python bench.py --only insert_batched > gpurun_out/${T}_ncu3.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_insert_commit -s 2 -c 2 -o gpurun_out/${T}_prof_commit python bench.py --only insert_batched > gpurun_out/${T}_ncu4.log 2>&1
# the counters bench.py quotes, regenerated from these captures before the bench line is taken (same source stamp)
R=${T:0:2}
for k in nn:nn_single_step rank:k_insert_rank commit:k_insert_commit; do a=${k%%:*}; b=${k##*:}; ncu -i gpurun_out/${T}_prof_$a.ncu-rep --page raw --csv > profiles/${R}_ncu_full_$b.csv 2>/dev/null; cp profiles/${R}_ncu_full_$b.csv gpurun_out/; done
python tools/ncu_summary.py nn_single_step=profiles/${R}_ncu_full_nn_single_step.csv insert_rank_c5=profiles/${R}_ncu_full_k_insert_rank.csv insert_commit_c5=profiles/${R}_ncu_full_k_insert_commit.csv > gpurun_out/${T}_traffic.json && cp gpurun_out/${T}_traffic.json profiles/traffic.json
# (compute-sanitizer is closed on this GPU pool: `compute-sanitizer --tool racecheck python tools/racecheck_insert.py` where it is open)
fi
python bench.py --steps 20 --warmup 3 > gpurun_out/${T}_bench_n1.json 2> gpurun_out/${T}_bench_n1.err
# the other BASELINE configurations as bench lines (C4: 50 M-point SparseTrees radius sweep; C5: 64 OS1-128 scans inserted + queried)
python bench.py --config c4 --no-cpu > gpurun_out/${T}_c4_n1.json 2> gpurun_out/${T}_c4_n1.err
python bench.py --config c5 > gpurun_out/${T}_c5_n1.json 2> gpurun_out/${T}_c5_n1.err
python tests/bench_configs.py --configs c1,c3 > gpurun_out/${T}_configs_c1_c3.jsonl 2> gpurun_out/${T}_configs.err
AB_MATCH=1 python tools/nn_ab.py nn_kernel=0 > gpurun_out/${T}_match_rate.json 2> gpurun_out/${T}_match_rate.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/${T}_launches_c4_2Mpts.csv python bench.py --config c4 --c4-points 2000000 --no-cpu --steps 1 --warmup 1 > gpurun_out/${T}_ncu5.log 2>&1
cut -c1-300 gpurun_out/${T}_bench_n1.json
