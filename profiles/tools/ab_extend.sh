python -m pytest tests/test_gpu_parity.py tests/test_gpu_build.py -m gpu -x -q 2>&1 | tail -5
for k in pool ww; do
  GPT_WF_EXTEND=$k python bench.py --steps 5 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$k', round(d['value']), d['roofline']['kernel_class_ms_per_step'])"
  GPT_WF_EXTEND=$k python profiles/tools/cfg4_once.py 5000000 4 2>/dev/null | tail -1
done
