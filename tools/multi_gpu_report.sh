#!/bin/bash
# One multi-GPU box visit: topology, host->device copy matrix, then bench.py at N = 8, 4, 2, 1 (own arm) and the
# CPU reference arm once.  Everything lands in gpurun_out/r02_*.
set -u
N=${1:-8}
bash tools/box_probe.sh > /dev/null 2>&1; cp gpurun_out/box_probe.txt gpurun_out/r02_box_probe_${N}gpu.txt
timeout 300 python tools/h2d_matrix.py --mib 1024 > gpurun_out/r02_h2d_matrix_${N}gpu.json 2> gpurun_out/r02_h2d_matrix.err
port=29600
for n in 8 4 2; do
  [ $n -le $N ] || continue
  port=$((port+1))
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port \
      bench.py --gpus $n --steps 100 --warmup 5 > gpurun_out/r02_configs_n$n.out 2> gpurun_out/r02_configs_n$n.err
  echo "n=$n rc=$?"
  grep '^{' gpurun_out/r02_configs_n$n.out | tail -1 > gpurun_out/r02_configs_n$n.json
done
timeout 600 python bench.py --gpus 1 --steps 100 --warmup 5 > gpurun_out/r02_configs_n1.out 2> gpurun_out/r02_configs_n1.err
echo "n=1 rc=$?"
grep '^{' gpurun_out/r02_configs_n1.out | tail -1 > gpurun_out/r02_configs_n1.json
timeout 300 python bench.py --impl reference --gpus 1 --steps 5 --warmup 1 > gpurun_out/r02_reference_arm.json 2> gpurun_out/r02_reference_arm.err
python - <<'PY'
import json
for n in (1, 2, 4, 8):
    try:
        d = json.load(open(f"gpurun_out/r02_configs_n{n}.json"))
    except Exception as e:
        print(n, "no line", e); continue
    e = d["e2e"]
    print(n, "value", d["value"], "e2e", e["value"], "h2d", e["h2d_only"]["value"], e["h2d_only"]["per_rank"], "frac", e["frac_of_h2d"],
          "scattered", e["scattered"]["value"], "c3", d["config3"].get("ms"), d["config3"].get("efficiency_vs_n1"), d["config3"].get("golden_ok"),
          "c5", {k: v["ms"] for k, v in d["config5"].get("calls", {}).items()}, d["config5"].get("golden_ok"), d.get("multi_device_parity"))
try:
    print("reference", json.load(open("gpurun_out/r02_reference_arm.json"))["value"])
except Exception as e:
    print("reference", e)
PY
