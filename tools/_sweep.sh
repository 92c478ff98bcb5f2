# scratch: RTNW (final tree): lane refill threshold, with and without the warp-uniform loop
mkdir -p gpurun_out
for o in "" "--opt 5=8" "--opt 5=12" "--opt 5=16" "--opt 5=20" "--opt 5=24" "--opt 17=1 --opt 18=24 --opt 5=16" "--opt 17=1 --opt 18=28 --opt 5=16" "--opt 17=1 --opt 18=16 --opt 5=16" "--opt 17=1 --opt 18=24 --opt 5=24"; do
  python tools/bench_scenes.py --scenes rtnw --reps 2 $o 2>&1 >> gpurun_out/r02ay_rtnw_refill.jsonl
done
grep -o '"scene": "[a-z]*"\|"opts": [^]]*]\|"mrays_per_s": [0-9.]*' gpurun_out/r02ay_rtnw_refill.jsonl | paste - - -
