"""NVLink topology discovery for the process manager.

The reference assigns GPUs to ranks blindly (``gpu_ids[rank]`` or ``rank % count``,
ref:src/nbdistributed/process_manager.py:104-115).  The b200 backend moves data with direct
peer loads/stores, so every pair of ranks must be P2P-reachable and, for the advertised
bandwidth, connected through NVLink/NVSwitch.  ``discover()`` reports the peer matrix (NVML
first, torch's ``can_device_access_peer`` as a second source); ``check_assignment()`` is what
``ProcessManager.start_workers`` calls before spawning workers with ``--backend b200`` and
what ``%dist_status`` prints.
"""
from __future__ import annotations

from typing import Any, Dict, List, Optional, Sequence


def _nvml_matrix(indices: Sequence[int]) -> Optional[Dict[str, Any]]:
    try:
        import pynvml as nv
    except Exception:
        return None
    try:
        nv.nvmlInit()
    except Exception:
        return None
    try:
        handles = [nv.nvmlDeviceGetHandleByIndex(i) for i in indices]
        gpus = []
        for i, h in zip(indices, handles):
            name = nv.nvmlDeviceGetName(h)
            pci = nv.nvmlDeviceGetPciInfo(h).busId
            links = to_switch = 0
            for link in range(getattr(nv, "NVML_NVLINK_MAX_LINKS", 18)):
                try:
                    if nv.nvmlDeviceGetNvLinkState(h, link) == nv.NVML_FEATURE_ENABLED:
                        links += 1
                        try:
                            if nv.nvmlDeviceGetNvLinkRemoteDeviceType(h, link) == nv.NVML_NVLINK_DEVICE_TYPE_SWITCH:
                                to_switch += 1
                        except Exception:
                            pass
                except Exception:
                    break
            gpus.append({"index": i, "name": name.decode() if isinstance(name, bytes) else name,
                         "pci": pci.decode() if isinstance(pci, bytes) else pci, "nvlinks_up": links,
                         "nvswitch_links": to_switch})
        n = len(indices)
        nvlink = [[a == b for b in range(n)] for a in range(n)]
        p2p = [[a == b for b in range(n)] for a in range(n)]
        for a in range(n):
            for b in range(n):
                if a == b:
                    continue
                try:
                    nvlink[a][b] = nv.nvmlDeviceGetP2PStatus(handles[a], handles[b], nv.NVML_P2P_CAPS_INDEX_NVLINK) == nv.NVML_P2P_STATUS_OK
                    ok_r = nv.nvmlDeviceGetP2PStatus(handles[a], handles[b], nv.NVML_P2P_CAPS_INDEX_READ) == nv.NVML_P2P_STATUS_OK
                    ok_w = nv.nvmlDeviceGetP2PStatus(handles[a], handles[b], nv.NVML_P2P_CAPS_INDEX_WRITE) == nv.NVML_P2P_STATUS_OK
                    p2p[a][b] = ok_r and ok_w
                except Exception:
                    nvlink[a][b] = False
                    p2p[a][b] = False
                # Found on the 8xB200 box: inside a container nvmlDeviceGetP2PStatus answers "not
                # supported" for pairs that `nvidia-smi topo -m` shows as NV18 and that CUDA maps
                # peer to peer.  Two GPUs that both hang off an NVSwitch reach each other over
                # NVLink; for plain reachability CUDA's own answer is the authority.
                if not nvlink[a][b] and gpus[a]["nvswitch_links"] > 0 and gpus[b]["nvswitch_links"] > 0:
                    nvlink[a][b] = True
                if not p2p[a][b]:
                    p2p[a][b] = nvlink[a][b] or _cuda_peer(indices[a], indices[b])
        return {"source": "nvml", "gpus": gpus, "p2p": p2p, "nvlink": nvlink}
    finally:
        try:
            nv.nvmlShutdown()
        except Exception:
            pass


def _cuda_peer(a: int, b: int) -> bool:
    try:
        import torch

        return bool(torch.cuda.can_device_access_peer(a, b))
    except Exception:
        return False


def _torch_matrix(indices: Sequence[int]) -> Dict[str, Any]:
    import torch

    n = len(indices)
    gpus = [{"index": i, "name": torch.cuda.get_device_name(i), "pci": None, "nvlinks_up": None, "nvswitch_links": None}
            for i in indices]
    p2p = [[a == b or bool(torch.cuda.can_device_access_peer(indices[a], indices[b])) for b in range(n)]
           for a in range(n)]
    return {"source": "torch", "gpus": gpus, "p2p": p2p, "nvlink": None}


def discover(gpu_ids: Optional[Sequence[int]] = None) -> Dict[str, Any]:
    """Peer matrix of the GPUs ``gpu_ids`` (default: all visible).

    Returns ``{"source", "gpus": [...], "p2p": NxN bool, "nvlink": NxN bool | None,
    "all_p2p": bool, "all_nvlink": bool | None}``; ``{"gpus": []}`` on a machine without GPUs.
    """
    import torch

    if not torch.cuda.is_available():
        return {"source": "none", "gpus": [], "p2p": [], "nvlink": None, "all_p2p": False, "all_nvlink": None}
    indices = list(gpu_ids) if gpu_ids is not None else list(range(torch.cuda.device_count()))
    info = _nvml_matrix(indices) or _torch_matrix(indices)
    n = len(indices)
    info["all_p2p"] = all(info["p2p"][a][b] for a in range(n) for b in range(n))
    info["all_nvlink"] = None if info["nvlink"] is None else all(info["nvlink"][a][b] for a in range(n) for b in range(n))
    return info


def rank_gpu_map(num_processes: int, gpu_ids: Optional[Sequence[int]], device_count: int) -> List[int]:
    """The reference's rank -> GPU rule (process_manager.py:104-115 with the worker's fallback
    worker.py:141-143): explicit ids cycle, otherwise rank % device_count."""
    if gpu_ids:
        return [gpu_ids[r] if r < len(gpu_ids) else gpu_ids[r % len(gpu_ids)] for r in range(num_processes)]
    return [r % max(device_count, 1) for r in range(num_processes)]


def check_assignment(num_processes: int, gpu_ids: Optional[Sequence[int]] = None,
                     topo: Optional[Dict[str, Any]] = None, device_count: Optional[int] = None) -> Dict[str, Any]:
    """Validate a rank->GPU assignment for the b200 backend.

    Raises ``ValueError`` when two ranks share a GPU or a pair is not P2P-reachable (the
    backend would fail at connect time anyway; failing in the coordinator gives a readable
    error before any worker is spawned).  Returns ``{"map", "warnings"}``; a pair without
    NVLink is a warning (P2P over PCIe works, slowly).
    """
    if topo is None:
        topo = discover()
    if device_count is None:
        device_count = len(topo["gpus"])
    if num_processes > 8:
        raise ValueError("the b200 backend drives one NVSwitch node: at most 8 ranks")
    mapping = rank_gpu_map(num_processes, gpu_ids, device_count)
    if len(set(mapping)) != len(mapping):
        raise ValueError(f"ranks must own distinct GPUs for backend 'b200', got rank->gpu {mapping}")
    index_of = {g["index"]: k for k, g in enumerate(topo["gpus"])}
    warnings: List[str] = []
    for a in mapping:
        if a not in index_of:
            raise ValueError(f"GPU {a} is not visible (have {sorted(index_of)})")
    for i, a in enumerate(mapping):
        for b in mapping[i + 1:]:
            ka, kb = index_of[a], index_of[b]
            if not (topo["p2p"][ka][kb] and topo["p2p"][kb][ka]):
                raise ValueError(f"GPU {a} and GPU {b} cannot access each other's memory (no P2P): "
                                 "backend 'b200' cannot run on this pair")
            if topo["nvlink"] is not None and not (topo["nvlink"][ka][kb] and topo["nvlink"][kb][ka]):
                warnings.append(f"GPU {a} <-> GPU {b}: P2P without NVLink (PCIe path, expect low bandwidth)")
    return {"map": mapping, "warnings": warnings}


def format_matrix(topo: Dict[str, Any]) -> str:
    """Text block for %dist_status: NV = NVLink peer, PP = P2P without NVLink, -- = unreachable."""
    gpus = topo["gpus"]
    if not gpus:
        return "no CUDA GPUs"
    head = "        " + " ".join(f"GPU{g['index']:<2}" for g in gpus)
    rows = [head]
    for a, ga in enumerate(gpus):
        cells = []
        for b in range(len(gpus)):
            if a == b:
                cells.append(" X   ")
            elif topo["nvlink"] is not None and topo["nvlink"][a][b]:
                cells.append(" NV  ")
            elif topo["p2p"][a][b]:
                cells.append(" PP  ")
            else:
                cells.append(" --  ")
        rows.append(f"GPU{ga['index']:<2}   " + "".join(cells))
    rows.append(f"(source: {topo['source']}; all P2P: {topo.get('all_p2p')}; all NVLink: {topo.get('all_nvlink')})")
    return "\n".join(rows)
