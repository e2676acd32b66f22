# GPU-box script: official bench line, reference arm, ncu launch list + full capture of the two hot kernels
mkdir -p gpurun_out
python bench.py > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err; echo bench $?; tail -c 600 gpurun_out/bench_full.json
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo ref $?; tail -c 400 gpurun_out/bench_ref.json
SMALL="--steps 2 --warmup 3 --no-cpu-baseline --width 512 --height 512 --aa 1"
python bench.py $SMALL > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv python bench.py $SMALL > gpurun_out/ncu1.log 2>&1; echo ncu1 $?
python bench.py $SMALL > gpurun_out/plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"gather_kernel|intersect_kernel" -s 8 -c 2 -o gpurun_out/prof_r1 python bench.py $SMALL > gpurun_out/ncu2.log 2>&1; echo ncu2 $?
