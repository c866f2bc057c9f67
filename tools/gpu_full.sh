#!/bin/bash
# full single-GPU check: the whole -m gpu suite + the bench line
TAG=${1:-full}; OUT=gpurun_out; mkdir -p $OUT
timeout 1500 python -m pytest tests -m gpu -q -p no:cacheprovider > $OUT/${TAG}_pytest.log 2>&1; echo "pytest exit $?"; tail -4 $OUT/${TAG}_pytest.log
timeout 900 python bench.py --steps 20 --warmup 5 ${BENCH_ARGS:-} > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err; echo "bench exit $?"
python - <<P
import json
d = json.load(open("$OUT/${TAG}_bench.json"))
k = d["roofline"]["kernels"]
print("value %.1f e2e %.1f (serial %.3f ms) exact %.1f | ms %.4f | frac %.3f | launches %d" % (d["value"], d["e2e"]["value"], d["e2e"]["serial_ms_per_step"], d["exact_mode"]["value"], d["ms_per_step"], d["roofline"]["frac"], d["gpu_launches"]))
print({n: (round(v["ms"], 4), round(v.get("frac", 0), 3)) for n, v in k.items()})
print("cpu:", d.get("cpu_baseline"))
P
tail -3 $OUT/${TAG}_bench.err
