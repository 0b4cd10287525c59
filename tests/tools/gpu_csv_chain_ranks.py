#!/usr/bin/env python3
"""GPU box, under torchrun: the csv -> classify chain sharded over the ranks (NCCL histogram reduce),
checked against the same job run unsharded on rank 0's GPU.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29517 tests/tools/gpu_csv_chain_ranks.py [rows=20000]
"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "agent-tpu_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import model_file as mf  # noqa: E402
import synth  # noqa: E402
from csv_cases import _table  # noqa: E402
from ops import _shard, get_op  # noqa: E402

rows = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
tmp = f"/tmp/b2chain_r{rank}"
os.makedirs(tmp, exist_ok=True)
os.environ["CUDA_VOCAB_PATH"] = synth.write_vocab(os.path.join(tmp, "vocab.txt"))
os.environ["CUDA_MODEL_PATH"] = mf.write_model_file(os.path.join(tmp, "base.b2m"), mf.BERT_BASE)
path = os.path.join(tmp, "table.csv")
with open(path, "wb") as f:
    f.write(_table(rows, 123, "\r\n"))

classify = get_op("map_classify_tpu")
payload = {"source_uri": path, "start_row": 7, "shard_size": rows - 100, "allow_fallback": False}
single = classify(dict(payload))  # no process group yet: the whole job on this rank's GPU
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
t = time.time()
sharded = classify(dict(payload))
wall = time.time() - t
lo, hi = _shard.slice_bounds(single["row_count"], rank, world)
ok = (sharded["label_histogram"] == single["label_histogram"] and sharded["labels"] == single["labels"][lo:hi]
      and sharded["shard"] == {"rank": rank, "world_size": world, "start": lo, "end": hi}
      and sharded["metrics"]["items"] == single["row_count"] and sharded["row_count"] == single["row_count"])
flag = torch.tensor([1 if ok else 0], device="cuda")
dist.all_reduce(flag, op=dist.ReduceOp.MIN)
if rank == 0:
    print(json.dumps({"ranks": world, "rows": single["row_count"], "all_ranks_match_single_gpu": bool(flag.item()),
                      "histogram_sum": sum(sharded["label_histogram"]), "wall_ms_sharded": round(wall * 1e3, 1),
                      "gpu_ms_max_over_ranks": round(sharded["metrics"]["gpu_ms_total"], 1)}))
dist.destroy_process_group()
sys.exit(0 if flag.item() else 1)
