import sys, time, os
sys.path.insert(0, '.')
import numpy as np
import mgl_b200
from mgl_b200 import native, synth
n = 1_000_000
d = synth.generate(2, n, pinned=True)
sw = mgl_b200.MicrosoftSmithWaterman(0); sw._ensure()
cp = int(sys.argv[1]) if len(sys.argv) > 1 else 0
if cp: sw.set_option("chunk_pairs", cp)
if len(sys.argv) > 2: sw.set_option("host_plan", int(sys.argv[2]))
score_only = len(sys.argv) > 3 and sys.argv[3] == "so"
out = (native.pinned_array(n, np.int32), native.pinned_array(n + 1, np.uint64), native.pinned_array(8 * n, np.uint32))
args = (d["arena"], d["ref_off"], d["ref_len"], d["alt_off"], d["alt_len"], d["params"], 1, d["strategies"])
so_out = native.pinned_array(3 * n, np.int32).reshape(n, 3)
call = (lambda: sw.score_arrays(*args, out=so_out)) if score_only else (lambda: sw.align_arrays(*args, out=out))
for k in range(3):
    call()
os.environ["MGL_SW_TIMELINE"] = "1"
t = time.perf_counter(); call(); print("python wall ms", (time.perf_counter() - t) * 1e3)
