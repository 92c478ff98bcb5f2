# scratch: one-colour background finished by k_classify (no miss queue, no miss launch)
mkdir -p gpurun_out
for rep in 1 2; do for p in 0 1; do
  B200RT_CLASSIFY_BG=$p python tools/bench_scenes.py --scenes rtow,cornell,rtnw --reps 2 2>&1 | sed "s/\"opts\": \[\]/\"opts\": [\"classify_bg=$p\"]/" >> gpurun_out/r02bm_classify_solid_background.jsonl
done; done
grep -o '"scene": "[a-z]*"\|"opts": [^]]*]\|"mrays_per_s": [0-9.]*\|"kernel_launches": [0-9]*' gpurun_out/r02bm_classify_solid_background.jsonl | paste - - - -
python -m pytest tests -x -q -m gpu 2>&1 | tail -2
