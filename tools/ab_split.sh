# A/B of the two fp32 -> bf16 hi/lo split forms on ONE box: integer rounding (-DBSB_SPLIT_INT) vs packed converts.
for v in INT CVT INT CVT; do
  make -C speechseparation_b200/csrc clean > /dev/null
  if [ $v = INT ]; then make -C speechseparation_b200/csrc -j EXTRA=-DBSB_SPLIT_INT > /dev/null 2>&1; else make -C speechseparation_b200/csrc -j > /dev/null 2>&1; fi
  python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-streaming --no-gpu-comparator 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); k=d['kernels']; print('$v', round(d['ms_per_step'],4), {n:v_['ms'] for n,v_ in k.items()}, d['parity']['max_rel_err'])"
done
