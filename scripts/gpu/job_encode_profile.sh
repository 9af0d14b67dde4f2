set -x; mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
python bench.py --steps 5 --warmup 3 --windows 4096 > gpurun_out/bench_r1_encode.json 2> gpurun_out/bench_r1_encode.err; echo bench rc=$?
tail -c 3000 gpurun_out/bench_r1_encode.json
CMD="python bench.py --steps 1 --warmup 3 --windows 1024 --no-cpu-baseline"
$CMD > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_r1_encode.csv $CMD > gpurun_out/ncu1.log 2>&1
echo ncu1 rc=$?
$CMD > gpurun_out/plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:encode_windows -s 3 -c 2 -o gpurun_out/prof_r1_encode $CMD > gpurun_out/ncu2.log 2>&1
echo ncu2 rc=$?
ls -la gpurun_out
