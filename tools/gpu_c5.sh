#!/bin/bash
set -u
O=gpurun_out/${1:-r02f}
mkdir -p $O
for w in C5 C5x; do
  timeout 600 python bench.py --workload $w --no-cpu --no-e2e --steps 5 --warmup 3 > $O/bench_$w.json 2> $O/bench_$w.err; echo "$w rc=$?"
  python -c "
import json;d=json.load(open('$O/bench_$w.json'));print('$w %.1f Mpt/s  ms/step %.3f hbm frac %.3f' % (d['value']/1e6, d['ms_per_step'], d['roofline']['frac'] if d['roofline']['bound']=='hbm' else d['roofline']['hbm']['frac']))"
done
TEQPB200_GERG_ISOPROPS_JETS=1 timeout 900 python bench.py --workload C5x --points 20000 --no-cpu --no-e2e --steps 2 --warmup 1 > $O/bench_C5x_jets.json 2> $O/bench_C5x_jets.err; python -c "
import json;d=json.load(open('$O/bench_C5x_jets.json'));print('C5x with jet-sweep isoprops (20000 pts): %.3f Mpt/s' % (d['value']/1e6))"
tail -3 $O/*.err
