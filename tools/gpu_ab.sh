#!/bin/bash
# A/B arms on one box -> gpurun_out/ab.jsonl (one line per arm: variant, workload, kernel ms, step ms, clocks).
# usage: ARMS="name1:ENV1=a ENV2=b|name2:..." WLS="c5 c3" [NVCC="-DX=1" per arm via NVCC_<name>] bash tools/gpu_ab.sh
mkdir -p gpurun_out
IFS="|" read -ra AS <<< "${ARMS}"
for arm in "${AS[@]}"; do
  name="${arm%%:*}"; envs="${arm#*:}"
  nv=""; for kv in $envs; do case "$kv" in NVCC=*) nv="${kv#NVCC=}";; esac; done
  nv="${nv//,/ }"
  if [ "$nv" != "$LAST_NV" ]; then MCOV_NVCC_EXTRA="$nv" python marl_coverage_b200/build.py > /dev/null 2>&1 || { echo "build failed: $name"; continue; }; LAST_NV="$nv"; fi
  for w in ${WLS:-c5}; do
    env $envs timeout 300 python bench.py --workload $w --steps ${STEPS:-100} --warmup 20 --no-cpu --no-extras > gpurun_out/ab_tmp.json 2> gpurun_out/ab_tmp.err
    rc=$?
    python - "$name" "$w" "$envs" $rc <<'PY' | tee -a gpurun_out/ab.jsonl
import json, sys
name, w, envs, rc = sys.argv[1], sys.argv[2], sys.argv[3], int(sys.argv[4])
try:
    d = json.loads(open('gpurun_out/ab_tmp.json').read().strip().splitlines()[-1])
    r = d['roofline']
    print(json.dumps(dict(arm=name, workload=w, flags=envs, kernel_ms=round(r['kernel_ms'], 5), kernel_ms_median=round(r.get('kernel_ms_median', 0), 5),
                          frac=round(r['frac'], 4), ms_per_step=round(d['ms_per_step'], 5), value=d['value'], e2e=d['e2e']['value'],
                          sm_mhz=d['clocks']['sm_mhz'], reasons=d['clocks']['reasons'])))
except Exception as ex:
    print(json.dumps(dict(arm=name, workload=w, flags=envs, rc=rc, error=str(ex), tail=open('gpurun_out/ab_tmp.err').read()[-400:])))
PY
  done
done
if [ -n "$LAST_NV" ]; then python marl_coverage_b200/build.py > /dev/null; fi
