import json,sys
t=sys.stdin.read().strip().splitlines()
for ln in t:
    if ln.startswith("{"):
        d=json.loads(ln); print("value %.2f e2e %.2f frac %s ms/launch %s md5 %s launches %s outside %.3f" % (d["value"], d["e2e"]["value"], d["roofline"]["frac"], d["roofline"]["ms_per_launch"], d["config"]["bitstream_md5"], d["gpu_launches"], d["config"]["share_of_jobs_with_a_window_outside_the_box"])); break
else:
    print("NO JSON", t[-3:])
