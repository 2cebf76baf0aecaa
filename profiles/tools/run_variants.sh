#!/bin/bash
# On the GPU box: bench every variants/*.so (copied over the in-tree library), then restore the default build.
cd "$(dirname "$0")/../.."
cp g.p.u-pathtracer_b200/libgpt_b200.so /tmp/default.so
for f in /tmp/default.so variants/*.so; do
  cp $f g.p.u-pathtracer_b200/libgpt_b200.so
  for rep in 1 2; do
    python bench.py --steps ${STEPS:-8} --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); k=d['roofline']['kernel_class_ms_per_step']; print('$f', round(d['value']), 'ext %.2f shade %.2f' % (k['extend'], k['shade']), 'built-tree', round(d['device_built_tree']['value']))"
  done
done
cp /tmp/default.so g.p.u-pathtracer_b200/libgpt_b200.so
