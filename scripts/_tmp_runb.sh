python - <<PY
import json,subprocess,os,sys
for env in ({}, {"RTXNERF_NETR_CTAS":"3"}):
    e=dict(os.environ); e.update(env)
    out=subprocess.run([sys.executable,"bench.py","--no-cpu","--steps","10"],env=e,capture_output=True,text=True).stdout
    d=json.loads(out.strip().splitlines()[-1])
    print(env, d["value"], d["ms_per_step"], {k:round(v["ms"],3) for k,v in d["stages"].items()})
PY
