# scratch: BVH width on the small scenes with the round-2 kernels
mkdir -p gpurun_out
for o in "" "--opt 12=4" "--opt 12=4 --opt 17=0" "--opt 3=2"; do
  python tools/bench_scenes.py --scenes rtow,cornell,fox --reps 2 $o 2>&1 >> gpurun_out/r02bk_small_scene_width.jsonl
done
python tools/bench_scenes.py --scenes rtnw,sponza --reps 2 --opt 12=2 2>&1 >> gpurun_out/r02bk_small_scene_width.jsonl
grep -o '"scene": "[a-z]*"\|"opts": [^]]*]\|"mrays_per_s": [0-9.]*' gpurun_out/r02bk_small_scene_width.jsonl | paste - - -
