"""gather_stack4 at 4096 transitions per launch: GB/s (algorithmic 91 740 B per transition) against the CTA split.
   python scripts/micro/gather_splits_n4096.py [capacity]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
import bench
from dqn_mgsc_zoo_b200 import replay as rp
cap = int(sys.argv[1]) if len(sys.argv) > 1 else 300_000
replay = rp.TransitionReplay(cap, rp.Transition(None, None, None, None, None), np.random.RandomState(0))
src = bench.EpisodeSource(5)
for _ in range(cap):
    replay.add(rp.Transition(*src.next()))
store = replay.frame_store
store.flush()
n = 4096
bufs = [torch.empty((n, 84, 84, 4), dtype=torch.uint8, device="cuda") for _ in range(2)]
ai = torch.empty(n, dtype=torch.int32, device="cuda"); rf = torch.empty(n, device="cuda"); df = torch.empty(n, device="cuda")
rs = np.random.RandomState(1)
sets = [torch.from_numpy(rs.randint(cap, size=n).astype(np.int32)).cuda() for _ in range(24)]
ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for splits in [int(x) for x in os.environ.get("SPLITS", "0 100 0 100 0 100").split()]:
    store.set_gather_splits(splits)
    for i in range(4):
        store.gather_dev(sets[i].data_ptr(), n, bufs[0].data_ptr(), bufs[1].data_ptr(), ai.data_ptr(), rf.data_ptr(), df.data_ptr())
    torch.cuda.synchronize()
    ev0.record()
    for i in range(40):
        store.gather_dev(sets[i % 24].data_ptr(), n, bufs[0].data_ptr(), bufs[1].data_ptr(), ai.data_ptr(), rf.data_ptr(), df.data_ptr())
    ev1.record()
    torch.cuda.synchronize()
    us = ev0.elapsed_time(ev1) / 40 * 1e3
    print("splits %d: %.1f us  %.0f GB/s" % (splits, us, n * bench.GATHER_BYTES_PER_TRANSITION / us / 1e3))
