#!/bin/bash
# Round 3, visit B (1 GPU): fixed-point film accumulation - full GPU suite, then same-box A/B of the bench step against the previous library.
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu > gpurun_out/r3_gputests_fixedpoint.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r3_gputests_fixedpoint.log
for round in 1 2; do
for lib in prev new; do
  if [ $lib = prev ]; then export B200PT_LIB=$PWD/scratch/libb200pt_prev.so; else unset B200PT_LIB; fi
  timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/r3_ab_fixedpoint_${lib}_$round.json 2>/dev/null
  python -c "
import json; d=json.loads(open('gpurun_out/r3_ab_fixedpoint_${lib}_$round.json').read().strip().splitlines()[-1])
print('$lib $round', 'value %.4g e2e %.4g ms %.2f nif_ms %.2f trace_ms %.2f frac %.4f' % (d['value'], d['e2e']['value'], d['ms_per_step'], d['config']['nif_ms_per_step'], d['config']['trace_ms_per_step'], d['roofline']['frac']))"
done; done
unset B200PT_LIB
