#!/bin/bash
set -u
O=gpurun_out/${1:-r02g}
mkdir -p $O
cp teqp_b200/libteqpb200.so /tmp/lib_orig.so
for u in 1 2 4; do
  cp variants/libteqpb200_u$u.so teqp_b200/libteqpb200.so
  timeout 600 python bench.py --workload C5 --no-cpu --no-e2e --steps 5 --warmup 3 > $O/bench_C5_u$u.json 2> $O/bench_C5_u$u.err; echo "u=$u rc=$?"
  python -c "
import json;d=json.load(open('$O/bench_C5_u$u.json'));print('unroll $u: %.1f Mpt/s' % (d['value']/1e6))"
done
cp /tmp/lib_orig.so teqp_b200/libteqpb200.so
# memcheck of the jet-sweep isoprops on 21 components (crashed with an illegal address in r02f)
TEQPB200_GERG_ISOPROPS_JETS=1 timeout 600 compute-sanitizer --tool memcheck --print-limit 5 python bench.py --workload C5x --points 2000 --no-cpu --no-e2e --steps 1 --warmup 1 > $O/memcheck.log 2>&1; echo "memcheck rc=$?"; grep -E "Invalid|Error|at 0x|by thread|Address" $O/memcheck.log | head -20
