for spec in "C4-ant npg c4_npg_1gpu" "C4-ant trpo c4_trpo_1gpu" "C5-adroit npg c5_npg_1gpu" "C5-adroit dapg c5_dapg_1gpu"; do set -- $spec
  python bench.py --workload $1 --algo $2 --steps 10 --no-cpu-baseline > gpurun_out/r02_bench_$3.json 2>gpurun_out/tmp.err || tail -n 3 gpurun_out/tmp.err
done
bash profiles/probes/ncu_tc_passes.sh > /dev/null 2>&1
python - <<'PY'
import json
for f in ("c4_npg_1gpu","c4_trpo_1gpu","c5_npg_1gpu","c5_dapg_1gpu"):
    j=json.loads(open("gpurun_out/r02_bench_%s.json"%f).read().strip().splitlines()[-1])
    print(f, j.get("value"), j.get("ms_per_step"), (j.get("e2e") or {}).get("value"), j.get("clocks"))
PY
