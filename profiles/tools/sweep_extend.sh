set -e
cd g.p.u-pathtracer_b200
for mb in 8 10 12; do
  touch csrc/gpt_wavefront.cu; make EXTRA=-DWF_MIN_BLOCKS=$mb > /dev/null 2>&1
  cd ..
  for r in 2 4 8 16; do
    v=$(GPT_WF_REFILL_MIN=$r python bench.py --steps 3 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['roofline']['kernel_class_ms_per_step'])")
    c=$(GPT_WF_REFILL_MIN=$r python profiles/tools/cfg4_once.py 5000000 4 2>/dev/null | tail -1)
    echo "mb=$mb refill=$r bench=$v cfg4_5M: $c"
  done
  cd g.p.u-pathtracer_b200
done
