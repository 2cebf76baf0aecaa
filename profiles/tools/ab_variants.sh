# A/B of library variants built with different -D knobs (g.p.u-pathtracer_b200/_variants/*.so): each is copied over
# libgpt_b200.so and bench.py is run on it.  usage: bash profiles/tools/ab_variants.sh base shade5 ...
cp g.p.u-pathtracer_b200/libgpt_b200.so /tmp/lib_keep.so
for v in "$@"; do
  cp g.p.u-pathtracer_b200/_variants/lib_$v.so g.p.u-pathtracer_b200/libgpt_b200.so
  python bench.py --steps 6 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$v', round(d['value']), d['ms_per_step'], d['roofline'].get('kernel_class_ms_per_step'))"
done
cp /tmp/lib_keep.so g.p.u-pathtracer_b200/libgpt_b200.so
