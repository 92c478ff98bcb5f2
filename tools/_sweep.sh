# scratch: multiply-shift instead of the 64-bit division in k_generate
mkdir -p gpurun_out
python tools/bench_scenes.py --scenes fox,rtow,cornell,sponza --reps 3 2>&1 | sed "s/\"opts\": \[\]/\"opts\": [\"fastdiv\"]/" >> gpurun_out/r02bh_generate_fastdiv.jsonl
grep -o '"scene": "[a-z]*"\|"opts": [^]]*]\|"mrays_per_s": [0-9.]*\|"msamples_per_s": [0-9.]*' gpurun_out/r02bh_generate_fastdiv.jsonl | paste - - - -
python -m pytest tests -x -q -m gpu -k "render or sample_ranges or fisheye or wave or pipelines or converged or group" 2>&1 | tail -3
