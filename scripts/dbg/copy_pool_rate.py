"""Host copy pool throughput (no GPU): GB/s of mb200_debug_parallel_copy for 1..8 concurrent
callers staging 4 MB blocks, the size of a stager slot."""
import sys, threading, time
import numpy as np
sys.path.insert(0, ".")
from miraculix_b200 import api
lib = api.load_library()
blk = 4 << 20
def run(callers, reps=64):
    srcs = [np.random.default_rng(i).integers(0, 255, blk * 8, dtype=np.uint8) for i in range(callers)]
    dsts = [np.zeros(blk, dtype=np.uint8) for _ in range(callers)]
    def w(i):
        for r in range(reps):
            lib.mb200_debug_parallel_copy(dsts[i].ctypes.data, blk, srcs[i].ctypes.data + (r % 8) * blk, blk, blk, 1)
    ts = [threading.Thread(target=w, args=(i,)) for i in range(callers)]
    t0 = time.perf_counter()
    for t in ts: t.start()
    for t in ts: t.join()
    dt = time.perf_counter() - t0
    return callers * reps * blk / dt / 1e9
for c in (1, 2, 4, 8):
    print("callers", c, "GB/s %.1f" % run(c))
