# scratch: BVH4 extend kernel at 20 / 24 warps per SM (B200RT_OPT_EXTEND_BLOCKS = 5 / 6) against the 16-warp default
mkdir -p gpurun_out
for rep in 1 2; do for o in "" "--opt 6=5" "--opt 6=6"; do
  python tools/bench_scenes.py --scenes rtnw,sponza --reps 3 $o 2>&1 >> gpurun_out/r02an_bvh4_occupancy.jsonl
done; done
grep -o '"scene": "[a-z]*", "opts": [^]]*]\|"mrays_per_s": [0-9.]*' gpurun_out/r02an_bvh4_occupancy.jsonl | paste - -
