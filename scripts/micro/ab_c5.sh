#!/bin/bash
# A/B of writer variants on the large-lattice config (d=25 by default)
for so in scripts/micro/variants/lib_*.so; do
  echo "== $(basename $so)"
  TORIC_B200_LIB=$PWD/$so C5_ONLY_D=${1:-25} python scripts/run_configs.py --only c5 | python -c "
import sys, json
for l in sys.stdin:
    j = json.loads(l)
    if 'perspective_GBps' in j: print('   d=%d p=%.2f chunk=%d  %.0f GB/s' % (j['d'], j['p_error'], j['env_chunk'], j['perspective_GBps']))"
done
