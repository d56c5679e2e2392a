"""one-line digest of a bench.py JSON line: python tools/benchline.py FILE (or the line on stdin when no FILE is given)"""
import json
import sys

t = (open(sys.argv[1]).read() if len(sys.argv) > 1 else sys.stdin.read()).strip().splitlines()
for ln in t:
    if ln.startswith("{"):
        d = json.loads(ln)
        print("value %.2f e2e %.2f frac %s ms/launch %s md5 %s launches %s" % (
            d["value"], d["e2e"]["value"], d["roofline"]["frac"], d["roofline"]["ms_per_launch"],
            d["config"]["bitstream_md5"], d["gpu_launches"]))
        break
else:
    print("NO JSON", t[-3:])
