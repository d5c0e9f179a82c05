"""Diagnostic (N ranks under torchrun): where does an end-to-end Hamming step spend its host time,
stage by stage, WITHOUT extra barriers?  Prints one JSON line per rank."""
import json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist
import ssp_b200
from ssp_b200 import synthetic as synth
from ssp_b200.distributed import shard_pairs_contiguous, broadcast_descriptor_bank, gather_matches_resident

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
stream = torch.cuda.Stream(dev)
torch.cuda.set_stream(stream)
eng = ssp_b200.Engine(lr)
eng.set_stream(stream.cuda_stream)
sets = synth.descriptor_bank(21, 50_000, seed=0)
pairs = synth.all_pairs(21)
bank = offs = None
if rank == 0:
    bank_np, offs = ssp_b200.Engine.pack_descriptors(sets)
    pinned = torch.from_numpy(bank_np).pin_memory()
    bank = pinned.numpy()
local_pairs, lo, hi = shard_pairs_contiguous(pairs, rank, world)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
mode = sys.argv[1] if len(sys.argv) > 1 else "knn2_ratio"
log = []
for step in range(6):
    t = [time.perf_counter()]
    flush.zero_()
    b_t, o_all = broadcast_descriptor_bank(bank, offs, 0); t.append(time.perf_counter())
    eng.bind_descriptor_bank(b_t.data_ptr(), o_all, keepalive=b_t); t.append(time.perf_counter())
    local = eng.hamming_match_pairs(local_pairs, mode=mode, fetch=False); t.append(time.perf_counter())
    st = eng.hamming_last_stats()
    out = gather_matches_resident(eng, local.offsets, dst=0); t.append(time.perf_counter())
    log.append({"bcast": (t[1] - t[0]) * 1e3, "bind": (t[2] - t[1]) * 1e3, "match": (t[3] - t[2]) * 1e3,
                "kernel_ms": st["main_kernel_ms"], "gather": (t[4] - t[3]) * 1e3, "total": (t[4] - t[0]) * 1e3})
print(json.dumps({"rank": rank, "steps": [{k: round(v, 2) for k, v in s.items()} for s in log]}), flush=True)
dist.barrier()
dist.destroy_process_group()
