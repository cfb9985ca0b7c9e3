#!/bin/bash
# why do 4 sequences in flight not beat 1?  per-sequence wall-clock phases
mkdir -p gpurun_out
for c in 1 4 8; do
  FLIVO_REPLAY_TRACE=1 timeout 600 python scripts/replay_bench.py --sequences 16 --frames 200 --concurrent $c --mode 1 > gpurun_out/t_replay_c$c.json 2> gpurun_out/t_replay_c$c.err; echo "c$c rc=$?"
  cut -c1-260 gpurun_out/t_replay_c$c.json; grep "replay trace" gpurun_out/t_replay_c$c.err | tail -16
done
