#!/usr/bin/env python
"""pcsim on the GPU: `python tools/pcsim.py listlength -n 1024 -e crc8 -p 8 -w 100000000 -o out` (options as the
reference's pcsim, see polar-codes_b200/pcsim.py). Under torchrun the frames of every job are sharded over the ranks."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import polar_b200_loader  # noqa: E402

polar_b200_loader.load()
from polar_codes_b200 import pcsim  # noqa: E402

if __name__ == "__main__":
    if "RANK" in os.environ and int(os.environ.get("WORLD_SIZE", "1")) > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
        dist.init_process_group("nccl")
    rc = pcsim.main()
    if "RANK" in os.environ and int(os.environ.get("WORLD_SIZE", "1")) > 1:
        import torch.distributed as dist
        dist.destroy_process_group()
    sys.exit(rc)
