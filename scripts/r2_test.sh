# GPU tests + short bench of the default path.  usage: r2_test.sh TAG [pytest-args]
TAG=${1:-r2_test}
O=gpurun_out/$TAG
mkdir -p $O
timeout 900 python -m pytest tests -x -q -m gpu ${2:-} > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 $O/pytest_gpu.log
timeout 300 python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline ${3:-} > $O/bench.json 2> $O/bench.err; echo "bench rc=$?"
python - <<PY
import json
for l in open("$O/bench.json"):
    if l.startswith("{"):
        d=json.loads(l); r=d["roofline"]
        print("ms/step", round(d["ms_per_step"],3), "Gpix/s", round(d["value"]/1e3,2), "frac", round(r["frac"],4), "alone", {k:round(v["ms_per_step"],3) for k,v in r["kernel_alone"].items()}, "in_schedule", {k:round(v,3) for k,v in r["in_schedule"]["per_kernel_ms_per_step"].items()}, "launches", d["gpu_launches"], "clip49", d.get("clip49"), "dropin", d.get("dropin"))
PY
tail -3 $O/bench.err
