python bench.py --impl reference > gpurun_out/r02_bench_reference.json 2> gpurun_out/ref.err; tail -n 1 gpurun_out/ref.err
python bench.py > gpurun_out/r02_bench_1gpu.json 2> gpurun_out/b1.err; tail -n 1 gpurun_out/b1.err
for spec in "C1-pointmass npg c1" "C2-swimmer npg c2" "C4-ant npg c4_npg_1gpu" "C4-ant trpo c4_trpo_1gpu" "C5-adroit npg c5_npg_1gpu" "C5-adroit dapg c5_dapg_1gpu"; do set -- $spec
  python bench.py --workload $1 --algo $2 --steps 10 --no-cpu-baseline > gpurun_out/r02_bench_$3.json 2>gpurun_out/tmp.err || tail -n 3 gpurun_out/tmp.err
done
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > /dev/null 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02_ncu_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_l.log 2>&1
bash profiles/probes/ncu_tc_passes.sh > /dev/null 2>&1
python - <<'PY'
import json
for f in ("reference","1gpu","c1","c2","c4_npg_1gpu","c4_trpo_1gpu","c5_npg_1gpu","c5_dapg_1gpu"):
    try:
        j=json.loads(open("gpurun_out/r02_bench_%s.json"%f).read().strip().splitlines()[-1])
        print(f, j.get("value"), j.get("ms_per_step"), (j.get("e2e") or {}).get("value"), j.get("value_configs2"), (j.get("parity") or {}).get("pass"), j.get("clocks"))
    except Exception as e: print(f, "ERR", e)
PY
