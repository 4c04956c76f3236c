"""Debug aid: fill the GPU's free memory with a NaN byte pattern, release it, then run pytest in the same process, so that
every cudaMalloc of the library starts from poisoned memory (a read of something no kernel wrote shows up as a failure
instead of being masked by the leftovers of an earlier, identical run).  usage: poison_run.py <pytest args>"""
import sys
import torch
import pytest

free, _ = torch.cuda.mem_get_info()
chunks = []
left = int(free * 0.9)
while left > 0:
    n = min(left, 8 << 30)
    chunks.append(torch.full((n,), 0xFF, dtype=torch.uint8, device="cuda"))
    left -= n
torch.cuda.synchronize()
del chunks
torch.cuda.empty_cache()
sys.exit(pytest.main(sys.argv[1:]))
