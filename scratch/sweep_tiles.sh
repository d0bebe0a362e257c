for td in 8192 12288 16384 20864 32768 53248; do
  echo "n=125000 td=$td"; python bench.py --n-docs 125000 --tile-docs $td --no-cpu-baseline --steps 5 --warmup 3 2>/dev/null | python -c "import sys,json; j=json.loads(sys.stdin.read()); print(j['value'], j['roofline']['kernel_ms_per_launch'], j['roofline']['merge_ms_per_step'], j['index']['n_tiles'])"
done
for td in 12288 16384 32768 53248; do
  echo "n=1000000 td=$td"; python bench.py --tile-docs $td --no-cpu-baseline --steps 3 --warmup 3 2>/dev/null | python -c "import sys,json; j=json.loads(sys.stdin.read()); print(j['value'], j['roofline']['kernel_ms_per_launch'], j['roofline']['merge_ms_per_step'], j['index']['n_tiles'])"
done
