# scratch: exact agglomerative tree for small scenes
mkdir -p gpurun_out
for rep in 1 2; do for p in 0 1; do
  B200RT_AGGLOMERATIVE=$p python tools/bench_scenes.py --scenes cornell --reps 3 2>&1 | sed "s/\"opts\": \[\]/\"opts\": [\"agglomerative=$p\"]/" >> gpurun_out/r02ax_agglomerative_small_scenes.jsonl
done; done
grep -o '"scene": "[a-z]*"\|"opts": [^]]*]\|"mrays_per_s": [0-9.]*\|"build_ms": [0-9.]*\|"sah_cost": [0-9.]*' gpurun_out/r02ax_agglomerative_small_scenes.jsonl | paste - - - - -
python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "hits_vs_reference or render_matches or edge" 2>&1 | tail -3
