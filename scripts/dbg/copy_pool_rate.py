"""Host copy pool throughput (no GPU): GB/s of mb200_debug_parallel_copy for 1..8 concurrent
callers staging 4 MB blocks (the size of a stager slot) between buffers far larger than the caches
(256 MB per caller and side), i.e. at DRAM speed as in a real staged transfer."""
import sys, threading, time
import numpy as np
sys.path.insert(0, ".")
from miraculix_b200 import api
lib = api.load_library()
blk, nblk = 4 << 20, 64
def run(callers, reps=128):
    srcs = [np.random.default_rng(i).integers(0, 255, blk * nblk, dtype=np.uint8) for i in range(callers)]
    dsts = [np.ones(blk * nblk, dtype=np.uint8) for _ in range(callers)]   # touched: no page faults in the timed part
    def w(i):
        for r in range(reps):
            o = (r % nblk) * blk
            lib.mb200_debug_parallel_copy(dsts[i].ctypes.data + o, blk, srcs[i].ctypes.data + o, blk, blk, 1)
    ts = [threading.Thread(target=w, args=(i,)) for i in range(callers)]
    t0 = time.perf_counter()
    for t in ts: t.start()
    for t in ts: t.join()
    dt = time.perf_counter() - t0
    return callers * reps * blk / dt / 1e9
for c in (1, 2, 4, 8):
    print("callers", c, "GB/s %.1f" % run(c), flush=True)
