mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "tuning" > gpurun_out/r02j_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/r02j_pytest.log
S="--scenes rtow,cornell,rtnw,fox,sponza --reps 3"
python tools/bench_scenes.py $S > gpurun_out/r02j_sweep.jsonl 2>&1
python tools/bench_scenes.py $S --opt 21=1 >> gpurun_out/r02j_sweep.jsonl 2>&1
python tools/bench_scenes.py $S --opt 21=1 --opt 5=16 >> gpurun_out/r02j_sweep.jsonl 2>&1
python tools/bench_scenes.py $S --opt 21=1 --pipelines 1 >> gpurun_out/r02j_sweep.jsonl 2>&1
python tools/bench_scenes.py $S --pipelines 1 >> gpurun_out/r02j_sweep.jsonl 2>&1
tail -3 gpurun_out/r02j_pytest.log
