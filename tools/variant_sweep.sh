export TK_VALUE_SECONDS=0.3
cp torchkafka_b200/libtorchkafka_b200.so /tmp/orig.so
for f in variants/*.so; do
  cp $f torchkafka_b200/libtorchkafka_b200.so
  timeout 120 python bench.py --steps 32 --warmup 32 --skip-e2e --no-cpu 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read()); r=d['roofline']
print('$f', round(d['value']/1e6,1), 'Mrec/s  parse_us', round(r['algo_bytes_per_launch']/r['achieved']/1e3,1), 'frac', round(r['frac'],3))"
done
cp /tmp/orig.so torchkafka_b200/libtorchkafka_b200.so
