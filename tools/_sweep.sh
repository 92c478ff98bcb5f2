# scratch: fisheye early-out, second step (no draw for pixels wholly outside the circle)
mkdir -p gpurun_out
python tools/bench_scenes.py --scenes fox,rtow --reps 3 2>&1 | sed "s/\"opts\": \[\]/\"opts\": [\"fisheye early-out + no draw for dead pixels\"]/" >> gpurun_out/r02ba_fisheye_early_out.jsonl
grep -o '"scene": "[a-z]*"\|"opts": [^]]*]\|"mrays_per_s": [0-9.]*\|"msamples_per_s": [0-9.]*' gpurun_out/r02ba_fisheye_early_out.jsonl | paste - - - -
python -m pytest tests -x -q -m gpu -k "camera or fisheye or fox or converged or render_matches" 2>&1 | tail -3
