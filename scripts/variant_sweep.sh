#!/bin/bash
# Debug aid (GPU box): run the short bench with each library variant under fewtensorf_b200/variants/ in place of the
# product library and print step time + the stage times that differ.  Usage: scripts/variant_sweep.sh [stage ...]
cd "$(dirname "$0")/.."
cp fewtensorf_b200/libfewtensorf_b200.so /tmp/base.so
for lib in /tmp/base.so fewtensorf_b200/variants/*.so; do
  cp "$lib" fewtensorf_b200/libfewtensorf_b200.so
  python bench.py --steps 30 --warmup 5 --no-cpu --no-render 2>/dev/null | python -c "
import json,sys
b=json.loads([l for l in sys.stdin if l.startswith('{')][0])
st=b['stages']
print('$lib'.split('/')[-1], 'ms/step %.4f' % b['ms_per_step'], ' '.join('%s %.1f' % (k, st[k]['ms']*1e3) for k in sys.argv[1:] if k in st))
" "$@"
done
cp /tmp/base.so fewtensorf_b200/libfewtensorf_b200.so
