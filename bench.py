#!/usr/bin/env python3
"""bench.py — the map_classify hot path on N B200s, one JSON line per run.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME] [--items I] [--impl ours|reference]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Workloads (BASELINE.json `configs`; the default is the one the headline metric is quoted on):
  classify128       config 3: synthetic sentences of 16-126 word pieces, S = 128, batch 1,024   [default]
  classify128_full  the same with every item filling all 128 positions (the padded-S FLOP count is
                    then exactly the work done: no saving from packed rows)
  classify512       config 4: long documents of 256-510 pieces, S = 512, batch 1,024
  tokenize          config 2: WordPiece tokenize-and-pad of synthetic lines (lines/s)
  csv_chain         config 5: csv_shard -> map_tokenize -> map_classify over a synthetic CSV (rows/s)
  agent_loop        config 1: 1,024 sentences as leased tasks of 32 through the agent's job loop and a local
                    controller (HTTP lease / result protocol of app.py); `--impl reference` drives the CPU port
                    through the same loop

A step is one pass of the hot path over `--items` units per GPU. One process per GPU under torchrun;
units shard as contiguous slices with no data-path collective; the per-job label histogram and
metrics are summed with one NCCL all-reduce per step. `--single-process` (and, under torchrun, an
extra leg on rank 0 reported as `single_process`) drives all N GPUs from ONE process through the
library's own multi-device group (b2_group_*: host thread per device, ncclCommInitAll all-reduce) —
the path the reference's inline agent loop (app.py:283-287) can actually call.

`value` is whole-job throughput with inputs resident in HBM; `e2e` the same through the C-ABI host
call with pinned host buffers (H2D and D2H inside the timed region); `roofline` the dominant kernel
timed live with CUDA events on the library's stream; `cpu_baseline` the CPU oracle on a bounded
sample (a restatement, NOT the reference, whose own CPU path for the op is a stub that classifies
nothing: SURVEY.md section 0 G6).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "agent-tpu_b200"))

import numpy as np  # noqa: E402

BATCH = 1024
TOPK = 5
H, L, I, C = 768, 12, 3072, 16

WORKLOADS = {
    "classify128": dict(kind="classify", S=128, items=8192, min_tok=16, max_tok=126,
                        metric="map_classify items/sec (seq 128)", unit="items/s", config="BASELINE config 3 shape"),
    "classify128_full": dict(kind="classify", S=128, items=8192, min_tok=170, max_tok=190,
                             metric="map_classify items/sec (seq 128)", unit="items/s",
                             config="BASELINE config 3 shape, every item truncated to the full 128 positions"),
    "classify512": dict(kind="classify", S=512, items=4096, min_tok=256, max_tok=510,
                        metric="map_classify items/sec (seq 512)", unit="items/s", config="BASELINE config 4 shape"),
    "tokenize": dict(kind="tokenize", S=128, items=1_000_000, min_tok=16, max_tok=126,
                     metric="map_tokenize lines/sec (WordPiece, max_len 128)", unit="lines/s",
                     config="BASELINE config 2 shape"),
    "csv_chain": dict(kind="csv", S=128, items=262_144, min_tok=16, max_tok=126,
                      metric="csv_shard -> map_tokenize -> map_classify rows/sec (seq 128)", unit="rows/s",
                      config="BASELINE config 5 shape"),
    "agent_loop": dict(kind="loop", S=128, items=1024, min_tok=16, max_tok=126,
                       metric="map_classify_tpu items/sec through the agent job loop (batch 32)", unit="items/s",
                       config="BASELINE config 1 shape: leased tasks of 32 sentences through the lease / result protocol"),
}
# B1: the literal reference op, ops.get_op("map_classify_tpu")(payload) imported from
# /root/reference, measured in the build container — the checkout does not exist on the GPU box.
REFERENCE_STUB = {"ms_per_call": 0.30, "classifies": False, "where": "build container, 8 cores (SURVEY.md section 4.2)",
                  "note": "stub, no classification: _cpu_fallback returns topk [] (ops/map_classify_tpu.py:22-28)"}


def flop_per_item(S: int) -> float:
    """22.35 GFLOP at S = 128, 96.64 at S = 512."""
    return 12 * (24 * S * H * H + 4 * S * S * H) + 2 * H * H + 2 * H * C


def load_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return {"hbm_gbs": float(p["hbm_gbs"]), "bf16_burst": float(p["bf16_tflops"]),
                "bf16_sustained": float(p.get("bf16_tflops_sustained", p["bf16_tflops"])), "source": "measured"}
    except Exception:
        # fallback stated in B200_PROFILING.md
        return {"hbm_gbs": 6650.0, "bf16_burst": 1590.0, "bf16_sustained": 1400.0, "source": "fallback"}


def load_traffic(kernel: str, units: float = None):
    """dram__bytes_read + write per launch from the latest committed `ncu --set full` capture of
    `kernel` (profiles/traffic.json: written by hand after a capture, with the commit it was taken
    at). Kernels whose launch size varies are recorded per unit (line, file byte) and scaled by the
    `units` of this run's launch. None when there is no capture on record — never a literal here."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            t = json.load(f)[kernel]
        src = f"{t['capture']} (ncu --set full at {t['head']}: {t['what']})"
        if "bytes_per_unit" in t:
            return float(t["bytes_per_unit"]) * float(units), src + f", scaled to {units:.0f} {t['unit']}s"
        return float(t["bytes_per_launch"]), src
    except Exception:
        return None, "no ncu capture on record for this kernel"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.rows = []
        self.proc = None
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits", "-lms", "200",
                 "-i", str(index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            parts = [p.strip() for p in line.split(",")]
            if len(parts) >= 7:
                self.rows.append(parts)

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, pw, reasons = [], [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1])); pw.append(float(r[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if sm:  # "under load": samples at or above the median power draw
            cut = float(np.median(pw))
            loaded = [s for s, p in zip(sm, pw) if p >= cut] or sm
            return {"sm_mhz": float(np.median(loaded)), "sm_max_mhz": float(max(mx)), "power_w_max": float(max(pw)),
                    "samples": len(sm), "reasons": sorted(reasons)}
        return {"sm_mhz": None, "sm_max_mhz": None, "reasons": sorted(reasons)}


_POOLS = {}


def sentence_pool(wl: dict, size: int):
    import synth
    key = (wl["min_tok"], wl["max_tok"], size)
    if key not in _POOLS:
        _POOLS[key] = synth.SentenceGenerator().sentences(size, min_tokens=wl["min_tok"], max_tokens=wl["max_tok"])
    return _POOLS[key]


def make_inputs(wl: dict, items: int, steps: int, rank: int):
    """`steps` different item lists of `items` texts each, drawn from one seeded pool."""
    pool = sentence_pool(wl, min(4096, max(items, 64)) if wl["S"] == 128 else min(1024, max(items, 64)))
    out = []
    for s in range(steps):
        shift = (s * 911 + rank * 131) % len(pool)
        rot = pool[shift:] + pool[:shift]
        out.append((rot * (items // len(rot) + 1))[:items])
    return out


def prepare_files(tag: str):
    import model_file as mf
    import synth
    d = os.path.join(tempfile.gettempdir(), f"b2bench_{tag}")
    os.makedirs(d, exist_ok=True)
    vocab = synth.write_vocab(os.path.join(d, "vocab.txt"))
    model = mf.write_model_file(os.path.join(d, "bert_base_seed1234.b2m"), mf.BERT_BASE)
    return model, vocab, d


def write_csv(path: str, rows: int, wl: dict) -> int:
    """Synthetic `id,text,source` table: sentence-case text with commas,
    so every text cell is quoted; 1 in 16 also carries an escaped quote."""
    pool = sentence_pool(wl, 4096)
    cells = []
    for i, s in enumerate(pool):
        if i % 16 == 0:
            s = s.replace(" ", ' "q" ', 1)
        cells.append('"' + s.replace('"', '""') + '"' if any(ch in s for ch in ',"\n') else s)
    with open(path, "w", encoding="utf-8", newline="") as f:
        f.write("id,text,source\n")
        block = []
        for i in range(rows):
            block.append(f"{i},{cells[i % len(cells)]},src{i % 10}\n")
            if len(block) == 65536:
                f.write("".join(block)); block = []
        f.write("".join(block))
    return os.path.getsize(path)


def workload_config(wl_name: str, wl: dict, n: int, world: int, extra: dict = None) -> dict:
    S = wl["S"]
    cfg = {"workload": f"{wl_name}: {wl['config']}; {n} synthetic units per GPU per step "
                       f"({wl['min_tok']}-{wl['max_tok']} word pieces before truncation), BERT-base (12x768, "
                       f"random-init seed 1234, 16 labels), S = {S}, batch {BATCH}",
           "name": wl_name, "seq_len": S, "batch": BATCH, "units_per_step_per_gpu": n,
           "global_units_per_step": n * world, "sharding": f"contiguous slices, dp{world}",
           "l2": "inputs larger than L2: every batch streams > 1 GB of activations (L2 = 126 MB)"}
    cfg.update(extra or {})
    return cfg


# ------------------------------------------------------------------------------------------------
# CPU legs (the only places that touch oracle/)
# ------------------------------------------------------------------------------------------------
def _oracle():
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import torch
    import bert_oracle as bo
    torch.set_num_threads(os.cpu_count() or 1)
    return torch, bo


def cpu_classify_rate(model: str, vocab: str, texts, S: int, budget_s: float, max_items: int, batch: int = 32):
    """items/s of the CPU restatement (tokenizers + fp32 torch), all host threads."""
    torch, bo = _oracle()
    cfg, W = bo.load_model(model)
    done, t0 = 0, time.perf_counter()
    while done < max_items:
        chunk = texts[done % len(texts):][:batch] or texts[:batch]
        ids, mask = bo.tokenize(vocab, chunk, S)
        bo.forward(cfg, W, ids, mask, batch=batch)
        done += len(chunk)
        if time.perf_counter() - t0 > budget_s:
            break
    dt = time.perf_counter() - t0
    return done / dt, done, dt, torch.get_num_threads()


def cpu_tokenize_rate(vocab: str, texts, S: int, budget_s: float):
    """lines/s of tokenizers' BertWordPieceTokenizer.encode_batch (Rust, all host threads) — the oracle
    the GPU ids are bit-exact against — and of the reference's map_tokenize chunker restated
    (oracle/chunk_oracle.py, pure Python like the original, one core): BASELINE.md B3."""
    torch, bo = _oracle()
    import chunk_oracle
    block = texts[:50_000]
    done, t0 = 0, time.perf_counter()
    while time.perf_counter() - t0 < budget_s:
        bo.tokenize(vocab, block, S)
        done += len(block)
    wp = done / (time.perf_counter() - t0)
    done, t0 = 0, time.perf_counter()
    while time.perf_counter() - t0 < budget_s / 2:
        r = chunk_oracle.map_tokenize({"items": block, "chunk_size": 1024})
        assert r["ok"]
        done += len(block)
    ck = done / (time.perf_counter() - t0)
    return wp, ck, os.cpu_count() or 1


def cpu_csv_rate(path: str, rows: int, budget_rows: int):
    """rows/s of the reference's algorithm for read_csv_shard (csv.DictReader over the file, restated in
    oracle/csv_oracle.py; one core — it is a single-threaded C loop): BASELINE.md B4."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import csv_oracle
    n = min(rows, budget_rows)
    t0 = time.perf_counter()
    r = csv_oracle.read_csv_shard({"source_uri": path, "mode": "count", "start_row": 0, "shard_size": n})
    dt = time.perf_counter() - t0
    assert r["ok"] and r["row_count"] == n
    return n / dt, n, dt


def run_reference(args, wl_name: str, wl: dict, rank: int, world: int):
    """`--impl reference`: the CPU path for the same workload on the host cores, rank 0 only.

    The reference's own CPU path for the op (`_cpu_fallback`, ops/map_classify_tpu.py:22-28) returns
    `topk: []` without classifying, its real arithmetic is an absent third-party TFLite / Edge-TPU
    stack, and /root/reference does not exist on the GPU box; so this arm times the CPU oracle port
    on a bounded sample of the workload per step."""
    if rank != 0:
        return
    model, vocab, d = prepare_files("ref")
    S = wl["S"]
    n = args.items or wl["items"]
    total = args.steps + args.warmup
    torch, bo = _oracle()
    cores = torch.get_num_threads()
    extra = None
    if wl["kind"] == "classify":
        per_step = 64 if S == 128 else 16
        texts = make_inputs(wl, per_step * total, 1, 0)[0]
        cfg, W = bo.load_model(model)
        # the same `config` as our arm, key for key: its inputs are a pure function of (workload, items,
        # steps, warmup), so their mean token count is computed here with the oracle tokenizer
        tok = 0
        for t in make_inputs(wl, n, total, 0)[args.warmup:]:
            tok += int(bo.tokenize(vocab, t, S)[1].sum())
        world_cfg = max(1, args.gpus)
        extra = {"collective": "one NCCL all-reduce of int64[16+8] per step" if world_cfg > 1 else "none (1 GPU)",
                 "tokens_per_item_mean": round(tok / (n * args.steps), 1),
                 "token_rows": "packed: pad positions are never materialised (results bit-identical to the padded "
                               "layout); kernel rooflines count real token rows"}

        def step(i):
            chunk = texts[i * per_step:(i + 1) * per_step]
            for lo in range(0, per_step, 32):
                ids, mask = bo.tokenize(vocab, chunk[lo:lo + 32], S)
                bo.forward(cfg, W, ids, mask, batch=32)
        what = f"{per_step} items per step: tokenizers + fp32 torch BERT-base, batch 32"
    elif wl["kind"] == "tokenize":
        per_step = 50_000
        texts = make_inputs(wl, per_step, 1, 0)[0]

        def step(i):
            bo.tokenize(vocab, texts, S)
        what = f"{per_step} lines per step: tokenizers BertWordPieceTokenizer.encode_batch"
    else:
        import csv_oracle
        per_step = 64
        path = os.path.join(d, "ref_sample.csv")
        write_csv(path, per_step * total, wl)
        cfg, W = bo.load_model(model)

        def step(i):
            r = csv_oracle.read_csv_shard({"source_uri": path, "start_row": i * per_step, "shard_size": per_step})
            texts = [row["text"] for row in r["rows"]]
            for lo in range(0, per_step, 32):
                ids, mask = bo.tokenize(vocab, texts[lo:lo + 32], S)
                bo.forward(cfg, W, ids, mask, batch=32)
        what = f"{per_step} rows per step: csv.DictReader shard + tokenizers + fp32 torch BERT-base"
    for i in range(args.warmup):
        step(i)
    t0 = time.perf_counter()
    for i in range(args.steps):
        step(args.warmup + i)
    dt = time.perf_counter() - t0
    value = per_step * args.steps / dt
    line = {
        "impl": "reference", "metric": wl["metric"], "value": value, "unit": wl["unit"], "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(wl_name, wl, n, max(1, args.gpus), extra),
        "cpu_baseline": {"value": value, "unit": wl["unit"], "cores": cores, "kind": "port",
                         "sample": f"{what}; {args.steps} timed steps after {args.warmup} warm-up steps",
                         "reference_stub": REFERENCE_STUB},
        "e2e": {"value": value, "unit": wl["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "the reference's own CPU path is a stub that classifies nothing (SURVEY.md G6); this is the oracle "
                "port on a bounded sample of the same workload",
    }
    emit(line)


# ------------------------------------------------------------------------------------------------
# distributed plumbing
# ------------------------------------------------------------------------------------------------
class Dist:
    def __init__(self, rank: int, local_rank: int, world: int):
        import torch
        self.torch = torch
        self.rank, self.local_rank, self.world = rank, local_rank, world
        torch.cuda.set_device(local_rank)
        self.dev = torch.device("cuda", local_rank)
        self.dist = None
        self.cpu_group = None
        if world > 1:
            import torch.distributed as dist
            dist.init_process_group("nccl", device_id=self.dev)
            self.dist = dist
            self.cpu_group = dist.new_group(backend="gloo")   # host-side waits that must not occupy the GPUs

    def barrier(self):
        if self.dist:
            self.dist.barrier(device_ids=[self.local_rank])

    def cpu_barrier(self):
        if self.dist:
            self.dist.barrier(group=self.cpu_group)

    def _reduce(self, x: float, op) -> float:
        if not self.dist:
            return x
        t = self.torch.tensor([x], dtype=self.torch.float64, device=self.dev)
        self.dist.all_reduce(t, op=op)
        return float(t.item())

    def max(self, x: float) -> float:
        return self._reduce(x, self.dist.ReduceOp.MAX) if self.dist else x

    def sum(self, x: float) -> float:
        return self._reduce(x, self.dist.ReduceOp.SUM) if self.dist else x


def pinned(torch, arr: np.ndarray):
    t = torch.empty(arr.size, dtype=torch.from_numpy(arr[:0].copy()).dtype).pin_memory()
    t.numpy()[:] = arr
    return t


def group_leg(rt, torch, model, vocab, wl, n_dev: int, n_per_gpu: int, steps: int, warmup: int, texts_by_step):
    """All `n_dev` GPUs driven from THIS process through b2_group_classify_text (host buffers in,
    results out, one NCCL all-reduce per step inside the library)."""
    S = wl["S"]
    grp = rt.EngineGroup(model, vocab, devices=list(range(n_dev)), max_batch=BATCH, max_len=S)
    try:
        n = n_per_gpu * n_dev
        jobs = []
        for t in texts_by_step:
            whole = (t * n_dev)[:n] if len(t) < n else t[:n]
            b, o = rt.pack_items(whole)
            jobs.append((pinned(torch, b), pinned(torch, o)))
        o_idx = torch.empty((n, TOPK), dtype=torch.int32).pin_memory()
        o_score = torch.empty((n, TOPK), dtype=torch.float32).pin_memory()
        o_lab = torch.empty(n, dtype=torch.int32).pin_memory()
        o_hist = torch.zeros(C, dtype=torch.int64).pin_memory()
        outs = (o_idx.numpy(), o_score.numpy(), o_lab.numpy(), o_hist.numpy(), None)
        for i in range(min(2, warmup)):
            grp.classify_packed(jobs[i][0].numpy(), jobs[i][1].numpy(), S, TOPK, outputs=outs)
        before = grp.allreduces()
        t0 = time.perf_counter()
        gpu_ms = 0.0
        launches = 0
        for i in range(steps):
            pb, po = jobs[(warmup + i) % len(jobs)]
            r = grp.classify_packed(pb.numpy(), po.numpy(), S, TOPK, outputs=outs)
            assert int(r["histogram"].sum()) == n
            gpu_ms += r["metrics"]["gpu_ms_total"]
            launches += r["metrics"]["kernel_launches"]
        wall = time.perf_counter() - t0
        return {"n_gpus": n_dev, "processes": 1, "value": round(n * steps / (gpu_ms / 1e3), 1),
                "e2e": round(n * steps / wall, 1), "unit": wl["unit"], "global_units_per_step": n,
                "ms_per_step_device": round(gpu_ms / steps, 3), "ms_per_step_wall": round(wall / steps * 1e3, 3),
                "gpu_launches": int(launches), "nccl_allreduces": grp.allreduces() - before,
                "nccl_version": grp.nccl_version,
                "api": "b2_group_classify_text (C ABI): one host thread per device, contiguous slices, one NCCL "
                       "all-reduce of int64[16+8] (+ max of device ms) per step; `value` = units / slowest device's "
                       "kernel time, `e2e` = wall clock of the host call incl. H2D/D2H"}
    finally:
        grp.close()


# ------------------------------------------------------------------------------------------------
# classify workloads
# ------------------------------------------------------------------------------------------------
def run_classify(args, wl_name: str, wl: dict, D: Dist):
    import torch
    from ops import _cuda_runtime as rt
    from ops import _shard
    rank, local_rank, world, dev = D.rank, D.local_rank, D.world, D.dev
    S = wl["S"]
    model, vocab, _ = prepare_files(f"r{rank}")
    peaks = load_peaks()
    n = args.items or wl["items"]
    total_steps = args.warmup + args.steps
    step_texts = make_inputs(wl, n, total_steps, rank)
    packed = [rt.pack_items(t) for t in step_texts]

    if args.single_process:
        if rank == 0:
            sampler = ClockSampler(0)
            leg = group_leg(rt, torch, model, vocab, wl, args.gpus, n, args.steps, args.warmup, step_texts)
            clocks = sampler.stop()
            emit({"metric": wl["metric"], "value": leg["value"], "unit": wl["unit"], "n_gpus": args.gpus,
                  "steps": args.steps, "warmup": args.warmup, "ms_per_step": leg["ms_per_step_device"],
                  "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
                  "data": "synthetic", "config": workload_config(wl_name, wl, n, args.gpus,
                                                                 {"processes": 1, "collective": "NCCL all-reduce inside the library"}),
                  "e2e": {"value": leg["e2e"], "unit": wl["unit"], "api": leg["api"]},
                  "gpu_launches": leg["gpu_launches"], "single_process": leg, "clocks": clocks})
        return

    eng = rt.Engine(model, vocab, device=local_rank, max_batch=BATCH, max_len=S)
    # token lengths of every step's items (for the attention FLOP count): one tokenizer pass, untimed
    len_sq = []
    for b, o in packed:
        _, mask, _ = eng.tokenize_packed(b, o, S)
        ln = mask.sum(1).astype(np.float64)
        len_sq.append(float((ln * ln).sum()))

    # ---- resident inputs / outputs (value) ------------------------------------------------------
    d_in = [(torch.from_numpy(b.copy()).to(dev), torch.from_numpy(o).to(dev)) for b, o in packed]
    d_top_idx = torch.empty((n, TOPK), dtype=torch.int32, device=dev)
    d_top_score = torch.empty((n, TOPK), dtype=torch.float32, device=dev)
    d_labels = torch.empty(n, dtype=torch.int32, device=dev)
    d_hist = torch.zeros(C, dtype=torch.int64, device=dev)
    torch.cuda.synchronize()
    ext = torch.cuda.ExternalStream(eng.stream, device=dev)

    def step_device(i):
        blob, off = d_in[i]
        m = eng.classify_device(blob.data_ptr(), off.data_ptr(), n, blob.numel(), S, TOPK, d_top_idx.data_ptr(),
                                d_top_score.data_ptr(), d_labels.data_ptr(), d_hist.data_ptr())
        # the job's one exchange: label histogram + additive metrics, NCCL sum over ranks
        hist, merged = _shard.reduce_job(d_hist.cpu().numpy(), m, local_rank)
        return m, hist, merged

    # Two passes over the same K steps. Pass 1 is the timed region of `value`: profiling off, as a
    # caller runs it. Pass 2 repeats the steps with per-launch CUDA events on the library's stream
    # (b2_set_profiling) for the per-kernel durations behind `roofline` and `kernels`; those events
    # cost one stream sync per batch and keep both batch lanes on one stream, so they stay out of
    # pass 1. `--profile-run` (the command ncu wraps) runs pass 2 only.
    def timed_pass(profile: bool):
        eng.set_profiling(profile)
        for i in range(args.warmup):
            step_device(i)
        D.barrier(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        kern, launches, tokens, hist = {}, 0, 0, None
        t_wall0 = time.perf_counter()
        e0.record(ext)
        for i in range(args.steps):
            m, hist, merged = step_device(args.warmup + i)
            launches += m["kernel_launches"]
            tokens += m["tokens"]
            for k, v in m.get("kernels", {}).items():
                a = kern.setdefault(k, {"ms": 0.0, "launches": 0})
                a["ms"] += v["ms"]; a["launches"] += v["launches"]
        e1.record(ext)
        D.barrier(); torch.cuda.synchronize()
        wall = (time.perf_counter() - t_wall0) * 1e3
        assert int(hist.sum()) == n * world, "label histogram does not add up to the job's items"
        return e0.elapsed_time(e1), wall, kern, launches, tokens

    sampler = ClockSampler(local_rank) if rank == 0 else None
    if args.profile_run:
        dev_ms, wall_ms, kern, launches, tokens = timed_pass(True)
        prof_ms = dev_ms
    else:
        dev_ms, wall_ms, _, launches, tokens = timed_pass(False)
    clocks = sampler.stop() if sampler else None
    if not args.profile_run:
        prof_ms, _, kern, _, _ = timed_pass(True)
    eng.set_profiling(False)

    dev_ms = D.max(dev_ms)
    wall_ms = D.max(wall_ms)
    ms_per_step = dev_ms / args.steps
    value = n * world / (ms_per_step / 1e3)
    launches_all = int(D.sum(float(launches)))

    if args.profile_run:
        if rank == 0:
            emit({"profile_run": True, "workload": wl_name, "ms_per_step": ms_per_step, "units_per_s": value,
                  "kernels": kern})
        eng.close()
        return

    # ---- e2e: C-ABI host call, pinned host buffers, H2D + D2H inside the timed region -------------
    pin = [(pinned(torch, b), pinned(torch, o)) for b, o in packed]
    o_idx = torch.empty((n, TOPK), dtype=torch.int32).pin_memory()
    o_score = torch.empty((n, TOPK), dtype=torch.float32).pin_memory()
    o_lab = torch.empty(n, dtype=torch.int32).pin_memory()
    o_hist = torch.zeros(C, dtype=torch.int64).pin_memory()
    outs = (o_idx.numpy(), o_score.numpy(), o_lab.numpy(), o_hist.numpy(), None)

    def step_host(i):
        pb, po = pin[i]
        r = eng.classify_packed(pb.numpy(), po.numpy(), S, TOPK, outputs=outs)
        _shard.reduce_job(r["histogram"], r["metrics"], local_rank)
        return r["metrics"]

    for i in range(min(2, args.warmup)):
        step_host(i)
    D.barrier(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    h2d = d2h = 0
    for i in range(args.steps):
        m = step_host(args.warmup + i)
        h2d += m["h2d_bytes"]; d2h += m["d2h_bytes"]
    D.barrier(); torch.cuda.synchronize()
    e2e_s = D.max(time.perf_counter() - t0)
    e2e_value = n * world * args.steps / e2e_s

    # ---- all GPUs from one process (the library's own multi-device group), rank 0 only ----------
    single = None
    if world > 1:
        D.cpu_barrier()                      # every rank's GPU work is finished
        if rank == 0:
            try:
                single = group_leg(rt, torch, model, vocab, wl, world, n, args.steps, args.warmup, step_texts)
            except Exception as e:  # noqa: BLE001 - the torchrun measurement above stands on its own
                single = {"error": f"{type(e).__name__}: {e}"}
        D.cpu_barrier()                      # the other ranks wait here on the host, GPUs idle

    # ---- the op itself (Python list in, JSON-able dict out), one step, informational -------------
    os.environ["CUDA_MODEL_PATH"] = model
    os.environ["CUDA_VOCAB_PATH"] = vocab
    os.environ["B2_DEVICE"] = str(local_rank)
    op_value = None
    packed_rows = bool(eng.info.packed_rows)
    part_slots = (H + 63) // 64
    if world == 1:
        eng.close()
        import ops
        run = ops.get_op("map_classify_tpu")
        run({"items": step_texts[0][:BATCH], "max_len": S})        # loads the handle
        t0 = time.perf_counter()
        r = run({"items": step_texts[-1], "topk": TOPK, "max_len": S})
        op_s = time.perf_counter() - t0
        if "fallback" in r:
            raise RuntimeError("op returned the error envelope: " + r["reason"])
        op_value = len(step_texts[-1]) / op_s

    if rank != 0:
        return

    # ---- per-kernel accounting (rank 0, second pass) ------------------------------------------------
    rows = float(tokens) if packed_rows else float(n * args.steps * S)
    sum_len_sq = sum(len_sq[args.warmup:]) if packed_rows else float(n * args.steps * S * S)
    # the last layer's out-projection / FFN run on the [CLS] rows only (one row per item)
    tail_rows = rows * (L - 1) + float(n * args.steps)
    flop = {"gemm_qkv": 2.0 * rows * 3 * H * H * L, "gemm_attn_out": 2.0 * tail_rows * H * H,
            "gemm_ffn1": 2.0 * tail_rows * I * H, "gemm_ffn2": 2.0 * tail_rows * H * I}
    att_flop = 4.0 * sum_len_sq * H * L          # Q.K^T and P.V over each sequence's own keys
    att_bytes = 4.0 * rows * H * 2 * L           # read Q, K, V, write O (bf16)
    byts = {"embed": rows * (H * 4 + H * 2 + 4.0),
            # with LayerNorm folded into the GEMMs this segment is ln_stats_finalize (12 partial sums in,
            # one (mean, rstd) out per row, once per layer) plus the two [CLS]-row LayerNorms of the tail
            "layernorm": rows * (L - 1) * (part_slots * 8 + 8.0) + 2.0 * n * args.steps * H * 2 * 2}
    kernels = []
    for k, v in kern.items():
        e = {"kernel": k, "ms": round(v["ms"], 3), "launches": v["launches"]}
        if v["ms"] <= 0:
            kernels.append(e); continue
        if k in flop:
            a = flop[k] / v["ms"] / 1e9
            e.update(bound="tensor", achieved=round(a, 1), unit="TFLOP/s", frac=round(a / peaks["bf16_sustained"], 4),
                     frac_of_burst=round(a / peaks["bf16_burst"], 4))
        elif k == "attention":
            gb, tf = att_bytes / v["ms"] / 1e6, att_flop / v["ms"] / 1e9
            if S == 128:
                e.update(bound="hbm", achieved=round(gb, 1), unit="GB/s", frac=round(gb / peaks["hbm_gbs"], 4),
                         tflops=round(tf, 1))
            else:
                e.update(bound="tensor", achieved=round(tf, 1), unit="TFLOP/s",
                         frac=round(tf / peaks["bf16_sustained"], 4), gbs=round(gb, 1))
        elif k in byts:
            a = byts[k] / v["ms"] / 1e6
            e.update(bound="hbm", achieved=round(a, 1), unit="GB/s", frac=round(a / peaks["hbm_gbs"], 4))
            if k == "layernorm":
                e["what"] = "ln_stats_finalize (11 per batch) + the two [CLS]-row LayerNorms; launch-latency sized"
        elif k == "tokenize":
            e.update(bound="latency", achieved=round(n * args.steps / v["ms"] * 1e3 / 1e6, 2), unit="M lines/s",
                     what="vocabulary hash probes (L2-resident table), one warp per line; not DRAM-bound")
        kernels.append(e)
    g_ms = sum(kern[k]["ms"] for k in flop if k in kern)
    g_n = sum(kern[k]["launches"] for k in flop if k in kern)
    g_flop = sum(flop.values())
    achieved = g_flop / g_ms / 1e9 if g_ms > 0 else 0.0
    traffic, traffic_src = load_traffic("gemm2_bf16_tn_kernel")
    roofline = {
        "kernel": "gemm2_bf16_tn_kernel (tcgen05 cta_group::2; QKV, attention-out, FFN1+GELU, FFN2 projections)",
        "bound": "tensor", "achieved": round(achieved, 1), "peak": peaks["bf16_sustained"], "unit": "TFLOP/s",
        "frac": round(achieved / peaks["bf16_sustained"], 4), "frac_of_burst": round(achieved / peaks["bf16_burst"], 4),
        "peak_source": f"{peaks['source']} bf16_tflops_sustained (kernel timed inside a long step)",
        "flop_per_launch": g_flop / g_n if g_n else None, "avg_launch_ms": g_ms / g_n if g_n else None,
        "launches": g_n, "share_of_step": round(g_ms / (sum(v["ms"] for v in kern.values()) or 1.0), 4),
        "measured": "per-launch CUDA events on the library's stream (b2_set_profiling) in a second pass over the "
                    "same K steps; that pass syncs once per batch and runs both batch lanes on one stream, so it is "
                    f"not the timed region of `value` ({prof_ms / args.steps:.2f} ms per step against {ms_per_step:.2f})",
        "traffic": traffic, "traffic_source": traffic_src,
    }
    fpi = flop_per_item(S)
    whole = {"achieved_tflops": round(value / world * fpi / 1e12, 1),
             "frac_of_sustained": round(value / world * fpi / 1e12 / peaks["bf16_sustained"], 4),
             "real_tflops": round((g_flop + att_flop) / args.steps / (ms_per_step / 1e3) / 1e12, 1),
             "note": f"achieved_tflops charges every item the padded-S count ({fpi / 1e9:.2f} GFLOP/item); real_tflops "
                     "counts only the token rows that exist (GEMMs + attention)"}

    cpu = None
    if world == 1:
        rate, done, dt, cores = cpu_classify_rate(model, vocab, step_texts[0], S, budget_s=15.0,
                                                  max_items=2048 if S == 128 else 256, batch=32 if S == 128 else 8)
        cpu = {"value": round(rate, 2), "unit": wl["unit"], "cores": cores, "kind": "port",
               "sample": f"{done} of this workload's items (S={S}) in {dt:.1f} s: tokenizers + fp32 torch "
                         f"BERT-base; not the reference (its CPU path is a stub)",
               "reference_stub": REFERENCE_STUB}

    line = {
        "metric": wl["metric"], "value": round(value, 1), "unit": wl["unit"], "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": round(ms_per_step, 3), "wall_ms_per_step": round(wall_ms / args.steps, 3),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": workload_config(wl_name, wl, n, world, {
            "collective": "one NCCL all-reduce of int64[16+8] per step" if world > 1 else "none (1 GPU)",
            "tokens_per_item_mean": round(tokens / (n * args.steps), 1),
            "token_rows": ("packed: pad positions are never materialised (results bit-identical to the padded "
                           "layout); kernel rooflines count real token rows") if packed_rows else "padded to seq_len"}),
        "e2e": {"value": round(e2e_value, 1), "unit": wl["unit"], "h2d_bytes_per_step": int(h2d // args.steps),
                "d2h_bytes_per_step": int(d2h // args.steps), "api": "b2_classify_text (C ABI), pinned host buffers"},
        "e2e_op": ({"value": round(op_value, 1), "unit": wl["unit"],
                    "api": "ops.get_op('map_classify_tpu')({'items': list[str]}) incl. Python marshalling"}
                   if op_value else None),
        "single_process": single,
        "gpu_launches": launches_all,
        "roofline": roofline, "kernels": kernels, "whole_path": whole,
        "cpu_baseline": cpu, "clocks": clocks,
    }
    emit(line)


# ------------------------------------------------------------------------------------------------
# tokenize workload (config 2)
# ------------------------------------------------------------------------------------------------
def run_tokenize(args, wl_name: str, wl: dict, D: Dist):
    import torch
    from ops import _cuda_runtime as rt
    rank, local_rank, world, dev = D.rank, D.local_rank, D.world, D.dev
    S = wl["S"]
    _, vocab, _ = prepare_files(f"r{rank}")
    peaks = load_peaks()
    n = args.items or wl["items"]
    texts = make_inputs(wl, n, 1, rank)[0]
    blob, off = rt.pack_items(texts)
    eng = rt.Engine(None, vocab, device=local_rank, max_batch=1, max_len=S)
    d_blob, d_off = torch.from_numpy(blob.copy()).to(dev), torch.from_numpy(off).to(dev)
    d_ids = torch.empty((n, S), dtype=torch.int32, device=dev)
    d_mask = torch.empty((n, S), dtype=torch.uint8, device=dev)
    torch.cuda.synchronize()
    ext = torch.cuda.ExternalStream(eng.stream, device=dev)

    def step():
        return eng.tokenize_device(d_blob.data_ptr(), d_off.data_ptr(), n, blob.size, S, d_ids.data_ptr(),
                                   d_mask.data_ptr())
    for _ in range(args.warmup):
        step()
    D.barrier(); torch.cuda.synchronize()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(ext)
    k_ms = 0.0
    for _ in range(args.steps):
        k_ms += step()["gpu_ms_total"]
    e1.record(ext)
    D.barrier(); torch.cuda.synchronize()
    clocks = sampler.stop() if sampler else None
    ms_per_step = D.max(e0.elapsed_time(e1)) / args.steps
    value = n * world / (ms_per_step / 1e3)

    pb, po = pinned(torch, blob), pinned(torch, off)
    h_ids = torch.empty((n, S), dtype=torch.int32).pin_memory()
    h_mask = torch.empty((n, S), dtype=torch.uint8).pin_memory()
    lib = rt.load_library()
    import ctypes as Ct
    m = rt.Metrics(struct_size=Ct.sizeof(rt.Metrics))

    def step_host():
        rc = lib.b2_tokenize(eng.handle, Ct.c_void_p(pb.data_ptr()), Ct.c_void_p(po.data_ptr()), n, S,
                             Ct.c_void_p(h_ids.data_ptr()), Ct.c_void_p(h_mask.data_ptr()), Ct.byref(m))
        if rc:
            raise RuntimeError(lib.b2_last_error().decode())
    step_host()
    D.barrier(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    e2e_steps = max(1, min(args.steps, 5))
    for _ in range(e2e_steps):
        step_host()
    D.barrier()
    e2e_s = D.max(time.perf_counter() - t0)
    if rank != 0:
        return
    alg = float(blob.size + 8 * n + 5 * S * n)       # len(utf8) + 8 + 5*S bytes per line
    gbs = alg / (k_ms / args.steps) / 1e6
    traffic, traffic_src = load_traffic("wordpiece_kernel", n)
    cpu = None
    if world == 1:
        wp, ck, cores = cpu_tokenize_rate(vocab, texts, S, budget_s=10.0)
        cpu = {"value": round(wp, 1), "unit": wl["unit"], "cores": cores, "kind": "port",
               "sample": "50,000-line blocks for 10 s: tokenizers BertWordPieceTokenizer.encode_batch (Rust, all "
                         "threads) — the oracle the GPU ids are bit-exact against; the reference has no WordPiece",
               "b3_reference_map_tokenize": {"value": round(ck, 1), "unit": "lines/s", "cores": 1,
                                             "what": "the reference's fixed-width chunker (ops/map_tokenize.py:13-61) "
                                                     "restated in oracle/chunk_oracle.py, chunk_size 1024, pure Python"}}
    emit({
        "metric": wl["metric"], "value": round(value, 1), "unit": wl["unit"], "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": round(ms_per_step, 3), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8/int32", "data": "synthetic",
        "config": workload_config(wl_name, wl, n, world, {"text_bytes_per_step": int(blob.size)}),
        "e2e": {"value": round(n * world * e2e_steps / e2e_s, 1), "unit": wl["unit"],
                "h2d_bytes_per_step": int(m.h2d_bytes), "d2h_bytes_per_step": int(m.d2h_bytes),
                "api": "b2_tokenize (C ABI), pinned host buffers: jobs of >= 131,072 lines run as 65,536-line chunks on "
                       "three streams (upload, kernel and read-back of successive chunks overlap), which is why this "
                       "figure can exceed `value`, the one-launch kernel over the whole job"},
        "gpu_launches": args.steps * world,
        "roofline": {"kernel": "wordpiece_kernel", "bound": "hbm", "achieved": round(gbs, 1), "peak": peaks["hbm_gbs"],
                     "unit": "GB/s", "frac": round(gbs / peaks["hbm_gbs"], 4), "traffic": traffic,
                     "traffic_source": traffic_src, "algorithmic_bytes_per_launch": alg,
                     "avg_launch_ms": k_ms / args.steps,
                     "limited_by": "not DRAM: per-word greedy longest-match probes of an L2-resident vocabulary hash, "
                                   "one lane per word; report lines/s, the HBM fraction is given only because the "
                                   "contract asks for one"},
        "cpu_baseline": cpu, "clocks": clocks})


# ------------------------------------------------------------------------------------------------
# csv chain workload (config 5)
# ------------------------------------------------------------------------------------------------
def run_csv(args, wl_name: str, wl: dict, D: Dist):
    import mmap

    import torch
    from ops import _cuda_runtime as rt
    from ops import _shard
    rank, local_rank, world, dev = D.rank, D.local_rank, D.world, D.dev
    S = wl["S"]
    model, vocab, d = prepare_files(f"r{rank}")
    n = args.items or wl["items"]                      # rows per GPU per step
    n_dev = args.gpus if args.single_process else 1
    rows = n * (world if not args.single_process else n_dev)
    path = os.path.join(d, f"chain_{rows}.csv")
    if rank == 0 and not (os.path.exists(path) and os.path.getsize(path) > 0):
        write_csv(path + ".tmp", rows, wl)
        os.replace(path + ".tmp", path)
    if world > 1:
        D.cpu_barrier()
        path0 = os.path.join(tempfile.gettempdir(), "b2bench_r0", f"chain_{rows}.csv")
        path = path0
    size = os.path.getsize(path)
    if args.single_process:
        if rank != 0:
            return
        eng = rt.EngineGroup(model, vocab, devices=list(range(n_dev)), max_batch=BATCH, max_len=S)
    else:
        eng = rt.Engine(model, vocab, device=local_rank, max_batch=BATCH, max_len=S)
    with open(path, "rb") as f:
        data = mmap.mmap(f.fileno(), 0, access=mmap.ACCESS_READ)
    sampler = ClockSampler(local_rank) if rank == 0 else None
    opens = []
    idx = None
    for it in range(3):                               # first open allocates the pinned staging
        if idx is not None:
            idx.close()
        t0 = time.perf_counter()
        idx = rt.CsvIndex(eng, data)
        opens.append((time.perf_counter() - t0) * 1e3)
    assert idx.rows == rows, (idx.rows, rows)
    lo, hi = (0, rows) if args.single_process else _shard.slice_bounds(rows, rank, world)

    def step():
        r = idx.classify(1, lo, hi - lo, S, TOPK)
        hist, merged = _shard.reduce_job(r["histogram"], r["metrics"], local_rank)
        return r["metrics"], hist
    for _ in range(min(2, args.warmup)):
        step()
    D.barrier(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    gpu_ms, launches = 0.0, 0
    for _ in range(args.steps):
        m, hist = step()
        gpu_ms += m["gpu_ms_total"]; launches += m["kernel_launches"]
    D.barrier(); torch.cuda.synchronize()
    wall = D.max(time.perf_counter() - t0)
    gpu_ms = D.max(gpu_ms)
    clocks = sampler.stop() if sampler else None
    assert int(np.asarray(hist).sum()) == rows
    launches_all = int(D.sum(float(launches)))
    info = idx.info
    open_ms = min(opens[1:])
    if rank != 0:
        return
    value = rows * args.steps / (gpu_ms / 1e3)
    chain_wall = rows * args.steps / wall
    e2e_value = rows / (open_ms / 1e3 + wall / args.steps)   # one job: open (copy + index) + classify
    peaks = load_peaks()
    index_gbs = size / float(info.index_ms) / 1e6
    traffic, traffic_src = load_traffic("csv_classify_kernel", size)
    cpu = None
    if world == 1 and n_dev == 1:    # the CPU legs run on the one-GPU line only (box time on N GPUs is charged N times)
        rate, done, dt = cpu_csv_rate(path, rows, 2_000_000)
        crate, cdone, cdt, cores = cpu_classify_rate(model, vocab, sentence_pool(wl, 4096), S, budget_s=10.0, max_items=1024)
        chain_cpu = 1.0 / (1.0 / rate + 1.0 / crate)
        cpu = {"value": round(chain_cpu, 2), "unit": wl["unit"], "cores": cores, "kind": "port",
               "sample": f"csv.DictReader over {done} rows in {dt:.1f} s (1 core) + {cdone} rows classified by "
                         f"tokenizers + fp32 torch in {cdt:.1f} s ({cores} threads); rows/s of the two stages in series",
               "b4_reference_read_csv_shard": {"value": round(rate, 1), "unit": "rows/s", "cores": 1,
                                               "what": "the reference's reader (csv.DictReader loop, ops/csv_shard.py:9-26) "
                                                       "restated in oracle/csv_oracle.py, mode count"}}
    emit({
        "metric": wl["metric"], "value": round(value, 1), "unit": wl["unit"],
        "n_gpus": n_dev if args.single_process else world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": round(gpu_ms / args.steps, 3), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u8 (CSV, tokenizer) / bf16 (encoder)", "data": "synthetic",
        "config": workload_config(wl_name, wl, n, n_dev if args.single_process else world, {
            "csv_bytes": size, "csv_rows": rows, "processes": 1 if args.single_process else world,
            "file_copies": "one H2D of the file to the group's first device; text-column slices go device to device"
            if args.single_process else "every rank copies and indexes the whole file, then takes its slice of rows"}),
        "e2e": {"value": round(e2e_value, 1), "unit": wl["unit"], "h2d_bytes_per_step": size,
                "d2h_bytes_per_step": rows * (TOPK * 8 + 4) + C * 8,
                "api": "b2_csv_open (mmap'd file -> HBM, index) + b2_csv_classify / b2_group_csv_classify, results to host",
                "open_ms": round(open_ms, 2), "open_ms_first": round(opens[0], 2),
                "chain_wall_rows_per_s_index_resident": round(chain_wall, 1)},
        "gpu_launches": launches_all + 6,
        "roofline": {"kernel": "csv index (csv_classify_kernel + scans + emit)", "bound": "hbm",
                     "achieved": round(index_gbs, 1), "peak": peaks["hbm_gbs"], "unit": "GB/s",
                     "frac": round(index_gbs / peaks["hbm_gbs"], 4), "traffic": traffic, "traffic_source": traffic_src,
                     "index_ms": round(float(info.index_ms), 3), "algorithmic_bytes": size + 16 * rows,
                     "note": "the chain's time is the encoder's (see the classify128 line for its roofline); this entry "
                             "is the stage this workload adds: file bytes / index kernels' time"},
        "cpu_baseline": cpu, "clocks": clocks})
    idx.close()
    eng.close()


# ------------------------------------------------------------------------------------------------
# agent loop workload (config 1)
# ------------------------------------------------------------------------------------------------
def run_agent_loop(args, wl_name: str, wl: dict, rank: int, reference: bool):
    """A local controller hands out one warm-up task and then `items / 32` tasks of 32 sentences; an agent
    process leases them one at a time, calls `tpu_ops.map_classify_tpu(payload)` and posts the results
    (the protocol of the reference's app.py:161-234, spoken here by tests/tools/agent_loop_standin.py —
    the reference checkout does not exist on the GPU box). Timed on the controller's clock from the
    warm-up task's result to the last result. With --impl reference the agent's `tpu_ops` is the CPU
    port (oracle/loop_shim), so both arms pay the same protocol costs."""
    if rank != 0:
        return
    import threading
    from http.server import BaseHTTPRequestHandler, HTTPServer
    model, vocab, _ = prepare_files("loop")
    n = args.items or wl["items"]
    per_task = 32
    texts = make_inputs(wl, n, 1, 0)[0]
    tasks = [texts[:per_task]] + [texts[i:i + per_task] for i in range(0, n, per_task)]   # [0] = warm-up
    st = {"next": 0, "results": [], "t_first": None, "t_last": None}

    class H(BaseHTTPRequestHandler):
        def log_message(self, *a):
            pass

        def do_POST(self):
            body = json.loads(self.rfile.read(int(self.headers["Content-Length"])))
            if self.path == "/v1/leases":
                if st["next"] >= len(tasks):
                    self.send_response(204); self.end_headers(); return
                i = st["next"]; st["next"] += 1
                out = {"lease_id": f"L{i}", "tasks": [{"id": f"job-{i}", "op": "map_classify_tpu", "job_epoch": 1,
                                                       "payload": {"items": tasks[i], "topk": TOPK, "batch": per_task}}]}
            else:
                now = time.perf_counter()
                if body.get("job_id") == "job-0":
                    st["t_first"] = now
                st["t_last"] = now
                st["results"].append(body)
                out = {"ok": True}
            raw = json.dumps(out).encode()
            self.send_response(200); self.send_header("Content-Type", "application/json")
            self.send_header("Content-Length", str(len(raw))); self.end_headers(); self.wfile.write(raw)

    srv = HTTPServer(("127.0.0.1", 0), H)
    threading.Thread(target=srv.serve_forever, daemon=True).start()
    shim = os.path.join(ROOT, "oracle", "loop_shim") if reference else os.path.join(ROOT, "agent-tpu_b200")
    env = dict(os.environ, CONTROLLER_URL=f"http://127.0.0.1:{srv.server_port}", PYTHONPATH=shim,
               CUDA_MODEL_PATH=model, CUDA_VOCAB_PATH=vocab, B2_DEVICE="0", AGENT_NAME="bench-agent")
    agent = os.path.join(ROOT, "tests", "tools", "agent_loop_standin.py")
    sampler = None if reference else ClockSampler(0)
    p = subprocess.run([sys.executable, agent, "900"], env=env, capture_output=True, text=True, timeout=1200)
    clocks = sampler.stop() if sampler else None
    srv.shutdown()
    done = [r for r in st["results"] if r.get("job_id") != "job-0"]
    bad = [r for r in st["results"] if r.get("status") != "succeeded" or "fallback" in (r.get("result") or {})]
    if len(done) != len(tasks) - 1 or bad:
        raise RuntimeError(f"agent loop: {len(done)} of {len(tasks) - 1} results, {len(bad)} failed: "
                           f"{(bad or [None])[0]} {p.stdout[-500:]} {p.stderr[-500:]}")
    wall = st["t_last"] - st["t_first"]
    value = n / wall
    hist = np.sum([r["result"]["label_histogram"] for r in done], axis=0)
    assert int(hist.sum()) == n
    line = {
        "metric": wl["metric"], "value": round(value, 1), "unit": wl["unit"], "n_gpus": 0 if reference else 1,
        "steps": len(done), "warmup": 1, "ms_per_step": round(wall / len(done) * 1e3, 3), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32" if reference else "bf16", "data": "synthetic",
        "config": workload_config(wl_name, wl, n, 1, {"batch": per_task, "tasks": len(done), "items_per_task": per_task,
                                                      "agent": "tests/tools/agent_loop_standin.py (the reference's "
                                                               "lease / result protocol; its app.py is absent on the GPU box)"}),
        "e2e": {"value": round(value, 1), "unit": wl["unit"],
                "h2d_bytes_per_step": 0 if reference else int(sum(len(t.encode()) for t in tasks[1]) + 8 * (per_task + 1)),
                "d2h_bytes_per_step": 0 if reference else per_task * (TOPK * 8 + 4) + C * 8,
                "api": "controller -> HTTP lease -> agent loop -> tpu_ops.map_classify_tpu(payload) -> HTTP result; a "
                       "step is one leased task; `value` is this same end-to-end figure (there is no device-resident form)"},
        "gpu_launches": 0 if reference else int(sum(r["result"]["metrics"]["kernel_launches"] for r in done)),
        "roofline": None if reference else {
            "kernel": "gemm2_bf16_tn_kernel", "bound": "tensor", "achieved": None, "peak": load_peaks()["bf16_sustained"],
            "unit": "TFLOP/s", "frac": None, "traffic": None,
            "device_ms_per_task": round(float(np.mean([r["result"]["metrics"]["gpu_ms_total"] for r in done])), 3),
            "note": "32-item tasks are launch- and protocol-bound: the kernels' share of a task is device_ms_per_task of "
                    "ms_per_step; see the classify128 line for the kernels' roofline"},
        "cpu_baseline": ({"value": round(value, 2), "unit": wl["unit"], "cores": os.cpu_count(), "kind": "port",
                          "sample": f"all {n} sentences: tokenizers + fp32 torch BERT-base behind the same loop",
                          "reference_stub": REFERENCE_STUB} if reference else None),
        "clocks": clocks,
    }
    if reference:
        line["impl"] = "reference"
    emit(line)


_REAL_STDOUT = None


def claim_stdout() -> None:
    """Keep stdout to the one JSON line: from here on file descriptor 1 points at stderr, so whatever
    a library prints there (NCCL's version banner at NCCL_DEBUG=VERSION or above, for one) cannot land
    in front of it; `emit` writes to the saved descriptor."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def emit(line: dict) -> None:
    data = (json.dumps(line) + "\n").encode()
    fd = _REAL_STDOUT if _REAL_STDOUT is not None else 1
    while data:
        data = data[os.write(fd, data):]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", choices=sorted(WORKLOADS), default="classify128")
    ap.add_argument("--items", type=int, default=0, help="units per GPU per step (0 = the workload's default)")
    ap.add_argument("--impl", choices=("ours", "reference"), default="ours")
    ap.add_argument("--single-process", action="store_true",
                    help="drive all --gpus devices from this one process through b2_group_* (no torchrun)")
    ap.add_argument("--profile-run", action="store_true",
                    help="short run for ncu: skip the e2e, op and CPU-baseline legs (the line is then not a bench value)")
    args = ap.parse_args()
    wl = WORKLOADS[args.workload]
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "ours" and wl["kind"] != "loop" and world != args.gpus and world == 1 and args.gpus > 1 \
            and not args.single_process:
        # launched without torchrun: re-exec under it on one node (before stdout is claimed below — the
        # re-executed ranks must inherit the real stdout for their JSON line)
        os.execvp(sys.executable, [sys.executable, "-m", "torch.distributed.run", "--nnodes=1",
                                   f"--nproc-per-node={args.gpus}", "--master-addr", "127.0.0.1",
                                   "--master-port", "29533", os.path.abspath(__file__)] + sys.argv[1:])
    claim_stdout()
    if args.impl == "reference":
        # torchrun exports OMP_NUM_THREADS=1 to every rank; the CPU arm is to use all the host threads it
        # can (torch is imported only further down, so dropping the variable here is early enough)
        os.environ.pop("OMP_NUM_THREADS", None)
    if wl["kind"] == "loop":
        run_agent_loop(args, args.workload, wl, rank, args.impl == "reference")
        return
    if args.impl == "reference":
        run_reference(args, args.workload, wl, rank, world)
        return
    D = Dist(rank, local_rank, world)
    {"classify": run_classify, "tokenize": run_tokenize, "csv": run_csv}[wl["kind"]](args, args.workload, wl, D)
    if world > 1:
        D.dist.destroy_process_group()


if __name__ == "__main__":
    main()
