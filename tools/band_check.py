"""torchrun -n N tools/band_check.py: the NCCL latitude-band run (one process per GPU, stage exchange overlapped
with interior work) against the same case on ONE band, on this rank's own rows.  Bar: rel L2 <= 1e-13."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist
from gmcore_b200.model import band_parity

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
r = band_parity(rank, world, local)
if rank == 0:
    ok = r["worst_rel_l2"] <= 1e-13 and r["diag"] <= 1e-13
    print(f"band_check {r['case']}: worst rel L2 {r['worst_rel_l2']:.2e}, diag {r['diag']:.2e} -> {'OK' if ok else 'FAIL'}")
dist.destroy_process_group()
