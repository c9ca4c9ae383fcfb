"""Development probe: slice pass 1 time under the HAET_ST_DBG switches (results are garbage for dbg != 0)."""
import os, subprocess, sys, json
for d in sys.argv[1:]:
    e = dict(os.environ); e["HAET_ST_DBG"] = d
    out = subprocess.run([sys.executable, "bench.py", "--steps", "3", "--warmup", "3", "--no-cpu"], env=e, capture_output=True, text=True).stdout
    try:
        j = json.loads(out.strip().splitlines()[-1]); print("dbg", d, "slice_pass1 %.3f ms" % j["stages_ms"]["slice_pass1"], flush=True)
    except Exception as ex:
        print("dbg", d, "failed", ex, out[-300:], flush=True)
