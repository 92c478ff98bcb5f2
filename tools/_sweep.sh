mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/r02x_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/r02x_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02x_smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/r02x_smoke.log
python bench.py > gpurun_out/r02x_bench.json 2> gpurun_out/r02x_bench.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02x_bench_ref.json 2> gpurun_out/r02x_bench_ref.err
tail -3 gpurun_out/r02x_pytest.log; tail -2 gpurun_out/r02x_smoke.log
