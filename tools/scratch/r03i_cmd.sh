B="timeout 100 python bench.py --no-cpu-baseline --steps 5 --warmup 3"
for bs in 0 1; do
  MGL_SW_BLOCKING_SYNC=$bs $B > gpurun_out/r03i_1gpu_allcpus_block$bs.json 2>gpurun_out/r03i_err
  MGL_SW_BLOCKING_SYNC=$bs taskset -c 0-3 $B > gpurun_out/r03i_1gpu_4cpus_block$bs.json 2>>gpurun_out/r03i_err
  MGL_SW_BLOCKING_SYNC=$bs taskset -c 0-1 $B > gpurun_out/r03i_1gpu_2cpus_block$bs.json 2>>gpurun_out/r03i_err
done
python - <<PY
import json,glob
for f in sorted(glob.glob("gpurun_out/r03i_1gpu*.json")):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); e=d["e2e"]; print(f, round(d["value"]), round(e["value"]), round(e["ms_per_step"],2), round(d.get("score_only_e2e",{}).get("value",0)))
    except Exception as x: print(f,"ERR",x)
PY
