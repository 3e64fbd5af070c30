#!/bin/bash
# One GPU-box pass: parity tests, smoke, bench lines, ncu launch lists, ncu --set full of the dominant kernels.
# usage (from the repo root, under gpurun): bash profiles/gpu_round.sh <tag>
tag=${1:-r02}
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/${tag}_pytest.log
tail -3 gpurun_out/${tag}_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${tag}_smoke.log 2>&1; echo "smoke exit $?"; tail -2 gpurun_out/${tag}_smoke.log
python bench.py > gpurun_out/${tag}_bench_n1.json 2> gpurun_out/${tag}_bench.err; echo "bench exit $?"
python bench.py --impl reference --steps 20 --warmup 20 > gpurun_out/${tag}_bench_ref_n1.json 2>> gpurun_out/${tag}_bench.err; echo "ref exit $?"
# launch lists (gpu__time_duration per launch; cold cache, serialised: shares, not absolutes)
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches.csv \
    python bench.py --steps 20 --warmup 3 --no-extras > gpurun_out/${tag}_ncu_launch.log 2>&1; echo "ncu launches exit $?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${tag}_launches_train.csv \
    python profiles/train_step.py 6 > gpurun_out/${tag}_ncu_train.log 2>&1; echo "ncu train launches exit $?"
python profiles/train_step.py 200 > gpurun_out/${tag}_train_step.json 2>&1; echo "train step exit $?"
# --set full: the two kernels of the rollout step at the named configuration and where the state exceeds L2
ncu --set full --clock-control none --import-source on -k regex:'env_step_kernel|actor_kernel' -s 20 -c 4 \
    -o gpurun_out/${tag}_full_4096 -f python bench.py --steps 20 --warmup 3 --no-extras > gpurun_out/${tag}_ncu_full.log 2>&1; echo "ncu full exit $?"
ncu --set full --clock-control none --import-source on -k regex:'env_step_kernel|actor_kernel' -s 10 -c 2 \
    -o gpurun_out/${tag}_full_1m -f python bench.py --envs 1048576 --steps 6 --warmup 3 --no-extras > gpurun_out/${tag}_ncu_full_1m.log 2>&1; echo "ncu full 1m exit $?"
ncu --set full --clock-control none --import-source on -k regex:tct_actor_kernel -s 20 -c 2 \
    -o gpurun_out/${tag}_full_fused -f python profiles/fused_probe.py > gpurun_out/${tag}_ncu_full_fused.log 2>&1; echo "ncu full fused exit $?"
python profiles/step_hist.py > gpurun_out/${tag}_step_hist.txt 2>&1; echo "step hist exit $?"
python profiles/e2e_trace.py > gpurun_out/${tag}_e2e_trace.txt 2>&1; echo "e2e trace exit $?"
python profiles/tct_trace.py 4096 > gpurun_out/${tag}_tct_trace_4096.txt 2>&1; echo "trace exit $?"
python profiles/update_phases.py > gpurun_out/${tag}_update_phases.json 2>&1; echo "phases exit $?"
