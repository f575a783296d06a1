# same-box A/B of library variants under ab/*.so:  bash tools/_ab.sh "A B C" "imagenet_packed" [rounds]
for r in $(seq 1 ${3:-2}); do
for v in $1; do
  for w in $2; do
    CDS_DA_LIB=$PWD/ab/$v.so python bench.py --only --workload $w --steps 10 --warmup 3 --no-cpu 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
st=d['roofline'].get('stages_ms',{})
print('$v $w %.3f stage %.3f blocks %s e2e %.2f'%(d['ms_per_step'],st.get('pglz_stage') or 0,st.get('k_wide_blocks_and_finish (rank 0)'),d['e2e']['ms_per_step']))"
  done
done
done
