# scratch: shared-reciprocal exact vec3 division (round-1 experiment) again, now under the two-pipeline schedule
mkdir -p gpurun_out
L=$PWD/cpu-ray-tracing-implementation_b200
for rep in 1 2; do for v in "" _rcp; do
  B200RT_LIB=$L/libb200rt$v.so python tools/bench_scenes.py --scenes rtow,cornell,rtnw,fox,sponza --reps 2 2>&1 | sed "s/\"opts\": \[\]/\"opts\": [\"build$v\"]/" >> gpurun_out/r02bj_shared_rcp_div.jsonl
done; done
grep -o '"scene": "[a-z]*"\|"opts": [^]]*]\|"mrays_per_s": [0-9.]*' gpurun_out/r02bj_shared_rcp_div.jsonl | paste - - -
