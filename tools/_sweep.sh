# scratch: host SAH with the box-area axis
mkdir -p gpurun_out
for rep in 1 2; do for p in 0 1; do
  B200RT_SAH_AREA=$p python tools/bench_scenes.py --scenes sponza,rtnw,fox,rtow,cornell --reps 2 --builder 0 2>&1 | sed "s/\"opts\": \[\]/\"opts\": [\"area_axis=$p\"]/" >> gpurun_out/r02aw_sah_area_axis.jsonl
done; done
grep -o '"scene": "[a-z]*"\|"opts": [^]]*]\|"mrays_per_s": [0-9.]*\|"build_ms": [0-9.]*\|"sah_cost": [0-9.]*' gpurun_out/r02aw_sah_area_axis.jsonl | paste - - - - -
