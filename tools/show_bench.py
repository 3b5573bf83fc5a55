import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(d["value"],d["ms_per_step"],d["e2e"]["ms_per_step"]); print(d["phases_ms"]); print(d["scan_only"])
