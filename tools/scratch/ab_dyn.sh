F='import sys,json
out=[]
for l in sys.stdin:
    if l.startswith("{"):
        d=json.loads(l); out.append("%s %.1f" % (d.get("scene"), d.get("mrays_per_s",0)))
print("  ".join(out))'
for ch in 4 8 16; do
  B200RT_NVCC_EXTRA="-DB200_EXT_DYNAMIC=$ch" python cpu-ray-tracing-implementation_b200/build.py --force >/dev/null 2>&1
  echo "DYN$ch (guided)"
  python tools/bench_scenes.py 2>/dev/null | python -c "$F"
  python tools/bench_scenes.py 2>/dev/null | python -c "$F"
done
