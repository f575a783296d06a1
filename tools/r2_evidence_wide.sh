# the wide-decoder part of tools/r2_evidence.sh (after a change that touches only pglz_wide.cuh): bench line, packed launch list, ncu captures
set -x
python bench.py --steps 20 --warmup 5 > gpurun_out/r2_bench_default.json 2> gpurun_out/r2_bench_default.err; echo bench rc=$?
ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/r2_launches_imagenet_packed.csv python bench.py --only --workload imagenet_packed --steps 2 --warmup 3 --no-cpu --no-graph > /dev/null 2>&1; echo launches rc=$?
ncu --set full --import-source on --clock-control none -k regex:"k_wide_maps|k_wide_blocks|k_wide_finish|k_wide_tails" -s 40 -c 8 -f -o gpurun_out/r2_wide_final python bench.py --only --workload imagenet_packed --toast-mib 256 --images-per-buffer 400 --steps 2 --warmup 3 --no-cpu --no-graph > /dev/null 2>&1; echo rc=$?
