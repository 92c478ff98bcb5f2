mkdir -p gpurun_out
S="--scenes rtow,cornell,rtnw,fox,sponza --reps 3"
L=$PWD/cpu-ray-tracing-implementation_b200
for rep in 1 2; do
for v in "" _b _c _d; do
  B200RT_LIB=$L/libb200rt$v.so python tools/bench_scenes.py $S 2>&1 | sed "s/\"opts\": \[\]/\"opts\": [\"build$v\"]/" >> gpurun_out/r02g_sweep.jsonl
done; done
