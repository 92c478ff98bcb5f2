# scratch: last build of the round
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu > gpurun_out/r02bi_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02bi_pytest.log
python bench.py > gpurun_out/r02bi_bench.json 2> gpurun_out/r02bi_bench.err
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02bi_smoke.log 2>&1
tail -n 3 gpurun_out/r02bi_pytest.log; head -c 300 gpurun_out/r02bi_bench.json; cat gpurun_out/r02bi_smoke.log
