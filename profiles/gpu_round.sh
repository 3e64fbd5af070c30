#!/bin/bash
# One GPU-box pass: parity tests, bench line, ncu launch list, ncu --set full of the two dominant kernels.
# usage (from the repo root, under gpurun): bash profiles/gpu_round.sh <tag>
tag=${1:-r01}
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/${tag}_pytest.log
tail -3 gpurun_out/${tag}_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${tag}_smoke.log 2>&1; echo "smoke exit $?"; tail -2 gpurun_out/${tag}_smoke.log
python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench exit $?"
python bench.py --impl reference --steps 10 --warmup 3 > gpurun_out/${tag}_bench_ref.json 2>> gpurun_out/${tag}_bench.err; echo "ref exit $?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches.csv \
    python bench.py --steps 20 --warmup 3 --no-extras > gpurun_out/${tag}_ncu_launch.log 2>&1; echo "ncu launches exit $?"
ncu --set full --clock-control none --import-source on -k regex:'env_step_kernel|tc_actor_kernel' -s 20 -c 4 \
    -o gpurun_out/${tag}_full_4096 -f python bench.py --steps 20 --warmup 3 --no-extras > gpurun_out/${tag}_ncu_full.log 2>&1; echo "ncu full exit $?"
ncu --set full --clock-control none --import-source on -k regex:'env_step_kernel|tc_actor_kernel' -s 10 -c 2 \
    -o gpurun_out/${tag}_full_1m -f python bench.py --envs 1048576 --steps 6 --warmup 3 --no-extras > gpurun_out/${tag}_ncu_full_1m.log 2>&1; echo "ncu full 1m exit $?"
# launch list of the reference-cadence train step (one policy step + two DDPG updates): profiles/train_step.py
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${tag}_launches_train.csv \
    python profiles/train_step.py 6 > gpurun_out/${tag}_ncu_train.log 2>&1; echo "ncu train launches exit $?"
python profiles/train_step.py 200 > gpurun_out/${tag}_train_step.json 2>&1; echo "train step exit $?"
