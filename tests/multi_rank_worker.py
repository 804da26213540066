"""torchrun worker of tests/test_gpu_multi.py: one rank per GPU renders its pixel tiles with DistributedRenderer and
rank 0 writes the gathered film; every rank writes its counters. TEST HELPER."""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    out_dir, W, H, sps, steps = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5])
    import ipu_path_trace_b200 as pt
    from ipu_path_trace_b200.distributed import DistributedRenderer
    local = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    r = DistributedRenderer(W, H, pt.synthetic_model(), sps, seed=1, nif_impl=pt.NIF_TCGEN05)
    paths = escaped = segments = 0
    for _ in range(steps):
        r.step()
        c = r.tracer.counters()
        paths += c.paths; escaped += c.escaped; segments += c.path_segments
    img = r.gather()
    json.dump({"rank": r.rank, "items": int(len(r.shard.worklist)), "paths": int(paths), "escaped": int(escaped), "segments": int(segments)},
              open(os.path.join(out_dir, f"rank{r.rank}.json"), "w"))
    if img is not None:
        np.save(os.path.join(out_dir, "film.npy"), img)
    dist.barrier()
    r.tracer.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
