mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_parity.py tests/test_gpu_config_parity.py -m gpu -q -x -k "fox or sponza or gltf" > gpurun_out/r02p_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/r02p_pytest.log
for o in 1 2 0; do python tools/bench_scenes.py --scenes fox,sponza --reps 3 --opt 20=$o >> gpurun_out/r02p_sweep.jsonl 2>&1; done
for l in 1 3 4; do python tools/bench_scenes.py --scenes sponza --reps 2 --opt 3=$l >> gpurun_out/r02p_sweep.jsonl 2>&1; done
tail -3 gpurun_out/r02p_pytest.log
