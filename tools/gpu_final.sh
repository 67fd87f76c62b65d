mkdir -p gpurun_out
timeout 300 python bench.py > gpurun_out/r1f_bench_c2.json 2> gpurun_out/r1f_bench_c2.err; echo "c2 rc=$?"
for w in c1 c3 c4 c5; do timeout 300 python bench.py --workload $w --no-cpu-baseline > gpurun_out/r1f_bench_$w.json 2> gpurun_out/r1f_bench_$w.err; echo "$w rc=$?"; done
timeout 300 python bench.py --impl reference --steps 2 --warmup 0 > gpurun_out/r1f_bench_reference_c2.json 2> gpurun_out/r1f_bench_reference_c2.err; echo "ref rc=$?"
timeout 1200 bash tools/gpu_profile.sh r1f
