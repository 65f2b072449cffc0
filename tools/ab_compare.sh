# A/B timing of alternative builds of libgoofit_b200.so (build/alt*/) against the in-tree one, on the same box
set -e
cp goofit_b200/lib/libgoofit_b200.so /tmp/new.so
for round in 1 2; do
for v in /tmp/new.so alt_builds/*/libgoofit_b200.so; do
  cp $v goofit_b200/lib/libgoofit_b200.so
  python bench.py --no-cpu-baseline --steps 100 > /tmp/o.json 2>/tmp/e.log || tail -3 /tmp/e.log
  python -c "
import json; d=json.load(open('/tmp/o.json')); w=d['workloads']; print('$v', 'dalitz_ms', round(d['roofline']['kernel_ms'],4), 'e2e_ms', round(d['e2e']['ms_per_step'],4), 'dp4_eval', round(w['dp4']['roofline']['eval_kernel_ms'],4), 'td_eval', round(w['tddp4']['roofline']['eval_kernel_ms'],4), 'waves', round(w['dalitz_cache']['kernel_ms'],3), 'dp4_ls_move', round(w['dp4']['lineshape_move']['ms_per_call'],3), 'td_ls_move', round(w['tddp4']['lineshape_move']['ms_per_call'],3), 'dp4_first', round(w['dp4']['e2e']['first_call_s'],4))"
done; done
cp /tmp/new.so goofit_b200/lib/libgoofit_b200.so
