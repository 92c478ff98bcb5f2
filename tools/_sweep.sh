# scratch: sanity after zero-initialising WaveParams
mkdir -p gpurun_out
python tools/bench_scenes.py --scenes cornell,rtow --reps 3 2>&1 | grep -o '"scene": "[a-z]*"\|"mrays_per_s": [0-9.]*' | paste - -
python -m pytest tests -x -q -m gpu 2>&1 | tail -2
python bench.py --no-scenes --no-cpu 2>/dev/null | head -c 400
