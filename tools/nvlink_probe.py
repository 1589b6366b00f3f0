import sys, os
sys.path.insert(0, os.path.join(os.getcwd(), "tools"))
import torch
torch.cuda.init()
from nvlink_counters import NvlinkCounters
n = NvlinkCounters(0)
print("nvlink counters:", n.ok, getattr(n, "kind", None), n.err, n.read() if n.ok else None)
