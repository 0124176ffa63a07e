#!/bin/bash
# runs bench.py (kernel leg only) once per variant library; prints value / kernel ms / checksum
cd "$(dirname "$0")/.."
for so in "$@"; do
  FLASHSDF_LIB=$PWD/point-cloud-signed-distance_b200/lib/variants/$so.so python bench.py --steps 3 --warmup 3 --no-extras --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$so', '%.4g evals/s' % d['value'], 'kernel %.2f ms' % d['roofline']['kernel_ms'], 'e2e %.4g' % d['e2e']['value'], 'checksum %.12g' % d['checksum_cost_sum'])"
done
