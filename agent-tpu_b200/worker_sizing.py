"""Worker capacity profile, GPU-first (reference worker_sizing.py:221-256).

`build_worker_profile()` returns the reference's schema — `cpu`, `gpu`, `tpu`, `workers` with the
same keys — so a controller that reads the reference's profile reads this one unchanged:

  cpu      same sizing rule and the same CPU_* environment knobs (worker_sizing.py:44-124)
  gpu      `{gpu_present, gpu_count, vram_gb, devices[{index, name, total_memory_bytes}],
            max_gpu_workers}` (worker_sizing.py:164-185); devices come from the CUDA runtime through
            libb2classify (b2_get_device_info) and, if the library is not built, from the same
            `nvidia-smi` query the reference uses (worker_sizing.py:139-161). Each device entry also
            carries `sm_count`, `compute_capability` and `l2_bytes` when the library answered.
  tpu      always absent here: this agent drives GPUs. The keys stay to avoid schema drift
            (worker_sizing.py:223-224).

`NVIDIA_VISIBLE_DEVICES=none` hides the GPUs as in the reference (worker_sizing.py:127-136).
"""
from __future__ import annotations

import math
import os
import subprocess
from typing import Any, Callable, Dict, List, Optional, TypeVar

try:
    import psutil
except ImportError:  # pragma: no cover
    psutil = None

T = TypeVar("T")

_NO_GPU: Dict[str, Any] = {"gpu_present": False, "gpu_count": 0, "vram_gb": None, "devices": [], "max_gpu_workers": 0}


def _env(name: str, default: T, cast: Callable[[str], T]) -> T:
    raw = os.getenv(name)
    if raw is None or not str(raw).strip():
        return default
    try:
        return cast(str(raw).strip())
    except Exception:
        return default


def _env_int(name: str, default: int) -> int:
    return _env(name, default, int)


def _env_float(name: str, default: float) -> float:
    return _env(name, default, float)


def _env_bool(name: str, default: bool = False) -> bool:
    raw = os.getenv(name)
    if raw is None:
        return default
    word = str(raw).strip().lower()
    if word in ("1", "true", "yes", "y", "on"):
        return True
    if word in ("0", "false", "no", "n", "off"):
        return False
    return default


def _logical_cores() -> int:
    if psutil is not None:
        try:
            return psutil.cpu_count(logical=True) or 1
        except Exception:
            pass
    return os.cpu_count() or 1


def _detect_cpu() -> Dict[str, Any]:
    """Cores, a pipeline factor (in-flight workers per core) and a soft cap bounded by RAM."""
    total = _logical_cores()
    reserved = min(_env_int("CPU_RESERVED_CORES_CAP", 4), max(_env_int("CPU_RESERVED_CORES_FLOOR", 1), total // 4))
    usable = max(1, total - reserved)
    factor = max(1.0, float(_env_float("CPU_PIPELINE_FACTOR", 4.0)))
    min_workers = max(1, int(_env_int("CPU_MIN_WORKERS", 1)))
    target = int(max(1, math.floor(usable * factor)))
    cap = int(max(min_workers, math.floor(usable * _env_float("CPU_SOFT_CAP_MULTIPLIER", 8.0))))
    if psutil is not None:
        try:
            available = int(getattr(psutil.virtual_memory(), "available", 0) or 0)
            per_worker = _env_int("CPU_PER_WORKER_BYTES", 32 * 1024 * 1024)
            if available > 0 and per_worker > 0:
                cap = max(1, min(cap, int(max(1, available // per_worker))))
        except Exception:
            pass
    return {
        "total_cores": int(total),
        "reserved_cores": int(reserved),
        "usable_cores": int(usable),
        "pipeline_factor": float(factor),
        "target_inflight_workers": int(target),
        "cpu_soft_cap_workers": int(cap),
        "min_cpu_workers": int(min_workers),
        "max_cpu_workers": int(cap),
    }


def _nvidia_visible_devices_allows_gpu() -> bool:
    raw = os.getenv("NVIDIA_VISIBLE_DEVICES")
    return raw is None or str(raw).strip().lower() != "none"


def _parse_nvidia_smi() -> List[Dict[str, Any]]:
    """`name, memory.total` per GPU from nvidia-smi; [] when the tool is missing or fails."""
    try:
        text = subprocess.check_output(
            ["nvidia-smi", "--query-gpu=name,memory.total", "--format=csv,noheader,nounits"],
            stderr=subprocess.STDOUT, text=True)
    except Exception:
        return []
    found: List[Dict[str, Any]] = []
    for index, line in enumerate(text.splitlines()):
        fields = [f.strip() for f in line.strip().split(",")]
        if len(fields) != 2:
            continue
        try:
            mib = float(fields[1])
        except ValueError:
            continue
        found.append({"index": index, "name": fields[0], "total_memory_bytes": int(mib * 1024 * 1024)})
    return found


def _cuda_devices() -> List[Dict[str, Any]]:
    """Devices as the CUDA runtime reports them (through the C-ABI library); [] if unavailable."""
    try:
        from ops._cuda_runtime import device_infos
        return device_infos()
    except Exception:
        return []


def _cuda_interconnect() -> Dict[str, Any]:
    """Peer-access / NVLink facts from the CUDA runtime (through the C-ABI library); {} if unavailable."""
    try:
        from ops._cuda_runtime import interconnect
        return interconnect()
    except Exception:
        return {}


def _detect_gpu() -> Dict[str, Any]:
    if not _nvidia_visible_devices_allows_gpu():
        return dict(_NO_GPU)
    from_cuda = _cuda_devices()
    devices = from_cuda or _parse_nvidia_smi()
    if not devices:
        return dict(_NO_GPU)
    biggest = max(int(d.get("total_memory_bytes", 0) or 0) for d in devices)
    gpu = {
        "gpu_present": True,
        "gpu_count": len(devices),
        "vram_gb": round(biggest / float(1024 ** 3), 2) if biggest > 0 else None,
        "devices": devices,
        "max_gpu_workers": len(devices),  # one worker process per GPU
    }
    link = _cuda_interconnect() if from_cuda and len(devices) > 1 else None
    if link and link.get("gpu_count") == len(devices):
        # beyond the reference's schema (worker_sizing.py:164-185): whether one agent process can spread a
        # job over all GPUs with device-to-device slices (ops/_cuda_runtime.EngineGroup)
        gpu["interconnect"] = link
    return gpu


def _detect_tpu() -> Dict[str, Any]:
    hinted = (str(os.getenv("JAX_PLATFORM_NAME", "")).strip().lower() == "tpu"
              or os.getenv("TPU_NAME") is not None or os.getenv("TPU_TYPE") is not None)
    if _env_bool("TPU_DISABLED", False):
        hinted = False
    return {"tpu_present": False, "tpu_kind": ("hinted" if hinted else None), "devices": [], "max_tpu_workers": 0}


def build_worker_profile() -> Dict[str, Any]:
    cpu = _detect_cpu()
    gpu = _detect_gpu()
    tpu = _detect_tpu()
    if _env_bool("TPU_ONLY", False):  # kept for schema parity: caps CPU and hides GPUs
        cpu["cpu_soft_cap_workers"] = cpu["max_cpu_workers"] = cpu["min_cpu_workers"] = 1
        gpu = dict(_NO_GPU)
    total = max(1, int(cpu.get("cpu_soft_cap_workers", 1) or 1) + int(gpu.get("max_gpu_workers", 0) or 0)
                + int(tpu.get("max_tpu_workers", 0) or 0))
    return {
        "cpu": cpu,
        "gpu": gpu,
        "tpu": tpu,
        "workers": {"max_total_workers": int(total), "current_workers": 0},
    }
