import json,sys
d=json.loads(sys.stdin.read()); print(sys.argv[1:], "k", d["config"]["k"], "dens", d["config"]["grid_density"], d["device_ms"], d["restarts"])
