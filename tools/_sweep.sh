# scratch: prefetch of the child that pops next (BVH4)
mkdir -p gpurun_out
L=$PWD/cpu-ray-tracing-implementation_b200
for rep in 1 2; do for v in "" _pf2; do
  B200RT_LIB=$L/libb200rt$v.so python tools/bench_scenes.py --scenes sponza,rtnw --reps 2 2>&1 | sed "s/\"opts\": \[\]/\"opts\": [\"build$v\"]/" >> gpurun_out/r02bd_prefetch_next_pop.jsonl
done; done
grep -o '"scene": "[a-z]*"\|"opts": [^]]*]\|"mrays_per_s": [0-9.]*' gpurun_out/r02bd_prefetch_next_pop.jsonl | paste - - -
