#!/bin/bash
# where a replay frame's host time goes (FLIVO_FRAME_TRACE), 1 and 8 sequences in flight, plus the force-search bench line
mkdir -p gpurun_out
for c in 1 8; do
  FLIVO_FRAME_TRACE=1 timeout 600 python scripts/replay_bench.py --sequences 16 --frames 100 --concurrent $c > gpurun_out/l_replay_c$c.json 2> gpurun_out/l_replay_c$c.err; echo "replay conc $c rc=$?"
  cat gpurun_out/l_replay_c$c.json | cut -c1-400; grep "frame trace" gpurun_out/l_replay_c$c.err | head -4
done
timeout 600 python bench.py --force-search 1 --no-replay --no-cpu-baseline > gpurun_out/l_bench_force.json 2> gpurun_out/l_bench_force.err; echo "force rc=$?"; tail -c 300 gpurun_out/l_bench_force.json
