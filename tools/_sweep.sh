# scratch: same-box A/B of library builds (tools/bench_scenes.py picks the build up from B200RT_LIB)
mkdir -p gpurun_out
L=$PWD/cpu-ray-tracing-implementation_b200
for rep in 1 2; do for v in "" _r01 _minb6 _minb7 _minb10 _minb12; do
  B200RT_LIB=$L/libb200rt$v.so python tools/bench_scenes.py --scenes rtow,cornell,rtnw,fox,sponza --reps 3 2>&1 | sed "s/\"opts\": \[\]/\"opts\": [\"build$v\"]/" >> gpurun_out/r02aj_builds_same_box.jsonl
done; done
tail -n 3 gpurun_out/r02aj_builds_same_box.jsonl
