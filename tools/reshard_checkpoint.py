"""Re-slice DistributedFusedAdam checkpoint shards for another world size / pipeline-chunk count.

    python tools/reshard_checkpoint.py --numels numels.json --new-world 4 --new-chunks 8 \
        --out new_rank{rank}.pt old_rank0.pt old_rank1.pt ...

Each input is the torch.save() of one rank's DistributedFusedAdam.state_dict() ({step, fp32_p, fp32_m, fp32_v}), in
rank order; numels.json is the list of parameter sizes in optimizer order ([p.numel() for p in params])."""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "rocm-apex_b200"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("shards", nargs="+")
    ap.add_argument("--numels", required=True)
    ap.add_argument("--new-world", type=int, required=True)
    ap.add_argument("--new-chunks", type=int, required=True)
    ap.add_argument("--old-chunks", type=int, default=None)
    ap.add_argument("--out", required=True, help="output path pattern containing {rank}")
    a = ap.parse_args()
    from apex.contrib.optimizers.distributed_fused_adam import reshard_state_dicts
    numels = json.load(open(a.numels))
    sds = [torch.load(f, map_location="cpu") for f in a.shards]
    for r, sd in enumerate(reshard_state_dicts(sds, numels, a.new_world, a.new_chunks, a.old_chunks)):
        torch.save(sd, a.out.format(rank=r))
        print("wrote", a.out.format(rank=r), "step", sd["step"], "shard elements", sd["fp32_p"].numel())


if __name__ == "__main__":
    main()
