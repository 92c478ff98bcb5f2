# scratch: final round-2 captures (ncu reports are summarised on the box: three of them exceed the 64 MiB that travel back)
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu > gpurun_out/r02as_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02as_pytest.log
python bench.py --impl reference > gpurun_out/r02as_bench_ref.json 2> gpurun_out/r02as_bench_ref.err
python bench.py > gpurun_out/r02as_bench.json 2> gpurun_out/r02as_bench.err
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02as_smoke.log 2>&1
python bench.py --steps 1 --warmup 1 --spp 60 --no-cpu --no-scenes --loop 0 > gpurun_out/r02as_bench_host_loop.json 2> gpurun_out/r02as_bench_host_loop.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/r02as_launches.csv python bench.py --steps 1 --warmup 1 --spp 60 --no-cpu --no-scenes --loop 0 > gpurun_out/r02as_ncu_launches.log 2>&1
python tools/ncu_summary.py launches gpurun_out/r02as_launches.csv > gpurun_out/r02as_launches_summary.txt
for sc in rtow rtnw sponza; do
  spp=20; [ $sc = sponza ] && spp=2; [ $sc = rtnw ] && spp=26
  python tools/profile_run.py --scene $sc --spp $spp > gpurun_out/r02as_profile_run_$sc.log 2>&1 && \
  ncu --set full --clock-control none --import-source on -k regex:k_extend_persistent -s 0 -c 2 -f -o /tmp/r02as_extend_$sc python tools/profile_run.py --scene $sc --spp $spp > gpurun_out/r02as_ncu_$sc.log 2>&1
  for k in 0 1; do python tools/ncu_summary.py kernel /tmp/r02as_extend_$sc.ncu-rep $k > gpurun_out/r02as_extend_${sc}_launch${k}_ncu_full.txt 2>&1; done
done
cp /tmp/r02as_extend_rtnw.ncu-rep gpurun_out/ 2>/dev/null
tail -n 3 gpurun_out/r02as_pytest.log; head -c 300 gpurun_out/r02as_bench.json; du -sh gpurun_out
