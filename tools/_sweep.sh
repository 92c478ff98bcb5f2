mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/r02ad_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/r02ad_pytest.log
L=$PWD/cpu-ray-tracing-implementation_b200
for rep in 1 2; do for v in "" _r01; do
  B200RT_LIB=$L/libb200rt$v.so python tools/bench_scenes.py --scenes rtow,cornell,rtnw,fox,sponza --reps 3 2>&1 | sed "s/\"opts\": \[\]/\"opts\": [\"build$v\"]/" >> gpurun_out/r02ad_vs_r01.jsonl
done; done
tail -3 gpurun_out/r02ad_pytest.log
