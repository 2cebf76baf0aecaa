#!/bin/bash
# round 2, GPU call M (1 GPU): the records of the final build with the rebuilt tree = run_r02_f.sh (GPU suite, frame counters,
# algorithmic bytes, bench both arms, launch list, --set full capture, small frames), plus smoke()
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
bash profiles/tools/run_r02_f.sh > gpurun_out/run_f.log 2>&1
tail -3 gpurun_out/smoke.log gpurun_out/pytest_gpu.log; cat gpurun_out/bench.err | tail -2; cut -c1-600 gpurun_out/bench.json; cat gpurun_out/small_frames.jsonl | cut -c1-420
