M=gpu__time_duration.sum,sm__cycles_active.max,sm__cycles_elapsed.max,smsp__warps_launched.sum,sm__ctas_launched.sum
for mr in 0 100000; do
for cc in all none; do
echo "=== OIS_TCT_MAX_ROWS=$mr cache-control=$cc"
OIS_TCT_MAX_ROWS=$mr ncu --metrics $M --cache-control $cc --clock-control none -k regex:"actor_kernel" -s 10 -c 2 python bench.py --steps 20 --warmup 3 --no-extras 2>&1 | grep -E "gpu__time|sm__cycles|warps_launched|ctas_launched" 
done; done
python profiles/tct_trace.py 4096 2>&1 | grep "all CTAs"
