# per variants/*.so: benchmarks/split_bench.py for the codecs in $CODECS (default gzip)
CODECS=${CODECS:-1}
cp torchkafka_b200/libtorchkafka_b200.so /tmp/orig.so
for f in variants/*.so; do
  cp $f torchkafka_b200/libtorchkafka_b200.so
  timeout 200 python benchmarks/split_bench.py --codecs $CODECS --batch-bytes 16384,65536 --iters 10 --out /tmp/sb.json > /dev/null 2>&1
  python - "$f" <<'PY'
import json, sys
rows = json.load(open('/tmp/sb.json'))
rows = rows["rows"] if isinstance(rows, dict) else rows
print(sys.argv[1], "  ".join("%s/%d: %s" % (r["codec"], r["batch_bytes"], {k: v for k, v in r.items() if "ms" in k}) for r in rows))
PY
done
cp /tmp/orig.so torchkafka_b200/libtorchkafka_b200.so
