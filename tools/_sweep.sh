# scratch: the head of the branch, built from scratch by __graft_entry__.build(), on a B200
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu > gpurun_out/r02bn_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02bn_pytest.log
python bench.py > gpurun_out/r02bn_bench.json 2> gpurun_out/r02bn_bench.err
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02bn_smoke.log 2>&1
tail -n 3 gpurun_out/r02bn_pytest.log; head -c 300 gpurun_out/r02bn_bench.json; cat gpurun_out/r02bn_smoke.log
