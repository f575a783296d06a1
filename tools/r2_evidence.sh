set -x
python bench.py --steps 20 --warmup 5 > gpurun_out/r2_bench_default.json 2> gpurun_out/r2_bench_default.err; echo bench rc=$?
( time python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/r2_bench_reference_arm.json 2> gpurun_out/r2_bench_reference_arm.err ) 2>&1 | tail -3; echo ref rc=$?
for w in imagenet_raw imagenet_pglz criteo imagenet_packed; do
  ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/r2_launches_$w.csv python bench.py --only --workload $w --steps 2 --warmup 3 --no-cpu --no-graph > /dev/null 2>&1; echo launches $w rc=$?
done
ncu --set full --import-source on --clock-control none -k regex:k_toast_gather -s 4 -c 1 -f -o gpurun_out/r2_gather python bench.py --only --workload imagenet_raw --steps 2 --warmup 3 --no-cpu --no-graph > /dev/null 2>&1; echo rc=$?
ncu --set full --import-source on --clock-control none -k regex:k_toast_gather -s 4 -c 1 -f -o gpurun_out/r2_gather_pglz python bench.py --only --workload imagenet_pglz --steps 2 --warmup 3 --no-cpu --no-graph > /dev/null 2>&1; echo rc=$?
ncu --set full --import-source on --clock-control none -k regex:k_pglz_decode -s 4 -c 1 -f -o gpurun_out/r2_pglz_img_final python bench.py --only --workload imagenet_pglz --steps 2 --warmup 3 --no-cpu --no-graph > /dev/null 2>&1; echo rc=$?
ncu --set full --import-source on --clock-control none -k regex:k_pglz_decode -s 4 -c 1 -f -o gpurun_out/r2_pglz_criteo_final python bench.py --only --workload criteo --steps 2 --warmup 3 --no-cpu --no-graph > /dev/null 2>&1; echo rc=$?
ncu --set full --import-source on --clock-control none -k regex:"k_wide_maps|k_wide_blocks|k_wide_finish" -s 25 -c 5 -f -o gpurun_out/r2_wide_final python bench.py --only --workload imagenet_packed --toast-mib 256 --images-per-buffer 400 --steps 2 --warmup 3 --no-cpu --no-graph > /dev/null 2>&1; echo rc=$?
