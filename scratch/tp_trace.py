"""Per-kernel timeline of a tensor-parallel decode step: torchrun --nproc-per-node N scratch/tp_trace.py [8b|70b]
(B200_TRACE=<file> -> <file>.r<rank>; read with scratch/trace_report.py)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

import __graft_entry__ as ge

pkg = ge.load_package()
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl")
box = [pkg.tp_unique_id() if rank == 0 else None]
dist.broadcast_object_list(box, src=0)
model = sys.argv[1] if len(sys.argv) > 1 else "8b"
cfg = pkg.llama31_70b() if model == "70b" else pkg.llama31_8b()
cfg.device, cfg.max_ctx, cfg.seed, cfg.eos_token, cfg.max_seqs = local, 1024, 0, -1, 1
eng = pkg.Engine(cfg, tp_rank=rank, tp_size=world, nccl_id=box[0])
for _ in range(3):
    ms, launches = eng.bench_decode(576, 32)
    dist.barrier()
if rank == 0:
    env = {k: v for k, v in os.environ.items() if k.startswith("B200_") and k != "B200_TRACE"}
    print(f"TPT {model} tp{world} {env}: {ms:.4f} ms/step = {1000 / ms:.1f} tok/s, {launches} launches")
eng.close()
dist.barrier()
dist.destroy_process_group()
