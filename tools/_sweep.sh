# scratch: solid-colour background added by the extend kernel itself (fused_miss = 2)
mkdir -p gpurun_out
for rep in 1 2; do for p in 0 1; do
  B200RT_FUSE_SOLID_BG=$p python tools/bench_scenes.py --scenes rtow,cornell,rtnw,fox,sponza --reps 2 2>&1 | sed "s/\"opts\": \[\]/\"opts\": [\"fuse_solid_bg=$p\"]/" >> gpurun_out/r02bl_fused_solid_background.jsonl
done; done
grep -o '"scene": "[a-z]*"\|"opts": [^]]*]\|"mrays_per_s": [0-9.]*\|"kernel_launches": [0-9]*' gpurun_out/r02bl_fused_solid_background.jsonl | paste - - - -
python -m pytest tests -x -q -m gpu 2>&1 | tail -2
