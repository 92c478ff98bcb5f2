# scratch: ncu --set full of the shade kernels of wave 0 (RTOW): launch order per wave is miss, lambertian, metal, dielectric
mkdir -p gpurun_out
for k in "1:lambertian" "2:metal" "3:dielectric"; do
  skip=${k%%:*}; name=${k##*:}
  ncu --set full --clock-control none --import-source on -k regex:k_shade -s $skip -c 1 -f -o /tmp/r02bf_shade_$name python tools/profile_run.py --scene rtow --spp 20 > gpurun_out/r02bf_ncu_shade_$name.log 2>&1
  python tools/ncu_summary.py kernel /tmp/r02bf_shade_$name.ncu-rep 0 > gpurun_out/r02bf_shade_${name}_wave0_ncu_full.txt 2>&1
done
head -12 gpurun_out/r02bf_shade_lambertian_wave0_ncu_full.txt
