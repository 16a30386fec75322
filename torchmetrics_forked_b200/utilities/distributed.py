"""State synchronisation across ranks.

`gather_all_tensors` keeps the reference's contract (utilities/distributed.py:97-147: a list with one
tensor per rank, uneven shapes allowed) and is what user supplied ``dist_sync_fn`` replacements look
like.  The default path of ``Metric._sync_dist`` does not use it: sum states are packed into one flat
buffer per dtype and reduced with ONE all-reduce, list ("cat") states are exchanged with ONE size
all-gather plus ONE data all-gather per state — no barrier, no gather-then-sum.
Works with the NCCL backend (CUDA tensors, NVLink/NVSwitch) and with gloo (CPU tests).
"""
from typing import Any, Dict, List, Optional, Sequence

import torch
import torch.distributed as dist
import torch.nn.functional as F  # noqa: N812
from torch import Tensor


def _world(group: Optional[Any]) -> Any:
    return dist.group.WORLD if group is None else group


def gather_all_tensors(result: Tensor, group: Optional[Any] = None) -> List[Tensor]:
    """All-gather one tensor from every rank; shapes may differ in any dimension (same ndim)."""
    group = _world(group)
    result = result.contiguous()
    world = dist.get_world_size(group)
    if result.ndim == 0:
        out = [torch.zeros_like(result) for _ in range(world)]
        dist.all_gather(out, result, group=group)
        return out
    shape = torch.tensor(result.shape, device=result.device, dtype=torch.int64)
    shapes = [torch.zeros_like(shape) for _ in range(world)]
    dist.all_gather(shapes, shape, group=group)
    shapes_cpu = torch.stack(shapes).cpu()
    biggest = shapes_cpu.max(dim=0).values
    if bool((shapes_cpu == biggest).all()):
        out = [torch.zeros_like(result) for _ in range(world)]
        dist.all_gather(out, result, group=group)
        return out
    pad: List[int] = []
    for have, want in zip(reversed(result.shape), reversed(biggest.tolist())):
        pad.extend([0, int(want) - int(have)])
    padded = F.pad(result, pad)
    out = [torch.zeros_like(padded) for _ in range(world)]
    dist.all_gather(out, padded, group=group)
    return [o[tuple(slice(int(d)) for d in shp)] for o, shp in zip(out, shapes_cpu.tolist())]


def packed_all_reduce_sum(tensors: Sequence[Tensor], group: Optional[Any] = None) -> List[Tensor]:
    """Sum every tensor over all ranks with one all-reduce per dtype present.

    The inputs are left untouched (``unsync`` must be able to restore them, metric.py:523-550); the
    outputs are views into the reduced flat buffer.
    """
    group = _world(group)
    out: List[Optional[Tensor]] = [None] * len(tensors)
    by_dtype: Dict[Any, List[int]] = {}
    for i, t in enumerate(tensors):
        by_dtype.setdefault((t.dtype, t.device), []).append(i)
    for idxs in by_dtype.values():
        flat = _pack([tensors[i] for i in idxs])
        dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
        offset = 0
        for i in idxs:
            n = tensors[i].numel()
            out[i] = flat[offset : offset + n].view(tensors[i].shape)
            offset += n
    return out  # type: ignore[return-value]


_PACK_TABLES: Dict[Any, Tensor] = {}


def _pack(states: Sequence[Tensor]) -> Tensor:
    """The states of one dtype as one flat buffer.  CUDA states of 4- or 8-byte elements go through
    ``tmb_pack_states`` (one launch for any number of states, the address table cached per set of states); a single
    state is just cloned; CPU tensors (gloo tests) are concatenated by torch."""
    if len(states) == 1:
        return states[0].reshape(-1).clone()
    first = states[0]
    if not first.is_cuda or first.element_size() not in (4, 8) or not all(s.is_contiguous() for s in states):
        return torch.cat([s.reshape(-1) for s in states])
    from torchmetrics_forked_b200 import _native as nat

    key = (tuple(s.data_ptr() for s in states), tuple(s.numel() for s in states), first.dtype, first.device.index)
    table = _PACK_TABLES.get(key)
    offsets = [0]
    for s in states:
        offsets.append(offsets[-1] + s.numel() * s.element_size())
    if table is None:
        if len(_PACK_TABLES) > 256:
            _PACK_TABLES.clear()
        table = torch.tensor(offsets + [s.data_ptr() for s in states], dtype=torch.int64, device=first.device)
        _PACK_TABLES[key] = table
    flat = torch.empty(offsets[-1] // first.element_size(), dtype=first.dtype, device=first.device)
    with nat.on_device(first.device):
        rc = nat.lib().tmb_pack_states(nat.ptr(table), len(states), offsets[-1], nat.ptr(flat), nat.stream_ptr(first.device))
    nat.check(rc, "tmb_pack_states")
    return flat


_DTYPES = [torch.float32, torch.float16, torch.bfloat16, torch.float64, torch.int64, torch.int32, torch.int16,
           torch.int8, torch.uint8, torch.bool]
_MAX_TRAIL = 6
_WIRE = {torch.bool: torch.uint8, torch.int16: torch.float16}  # dtype of a list state -> dtype on the wire


def all_gather_cat(
    local: Optional[Tensor], device: torch.device, group: Optional[Any] = None, extra_word: Optional[int] = None
):
    """Rank-ordered concatenation along dim 0 of every rank's ``local`` (``None`` = this rank has no rows).

    One small all-gather exchanges (rows, trailing shape, dtype); one all-gather moves the data. Returns
    ``None`` when no rank has rows.  Trailing shapes must agree between the ranks that have data.
    ``extra_word``: a small integer that rides in the size exchange; the call then returns ``(result, OR of every
    rank's word)`` (the metric's data-error bits travel this way).
    """
    if extra_word is None:
        return _all_gather_cat(local, device, group, 0)[0]
    return _all_gather_cat(local, device, group, int(extra_word))


def _all_gather_cat(local: Optional[Tensor], device: torch.device, group: Optional[Any], word: int):
    group = _world(group)
    world = dist.get_world_size(group)
    meta = torch.zeros(4 + _MAX_TRAIL, dtype=torch.int64)
    meta[3 + _MAX_TRAIL] = word
    if local is not None:
        if local.ndim == 0:
            local = local.unsqueeze(0)
        local = local.contiguous()
        trail = list(local.shape[1:])
        if len(trail) > _MAX_TRAIL:
            raise ValueError(f"list states with more than {_MAX_TRAIL + 1} dimensions are not supported")
        meta[0] = local.shape[0]
        meta[1] = _DTYPES.index(local.dtype)
        meta[2] = len(trail)
        for k, d in enumerate(trail):
            meta[3 + k] = d
    else:
        meta[1] = -1
    meta = meta.to(device)
    metas = torch.empty(world * meta.numel(), dtype=torch.int64, device=device)
    dist.all_gather_into_tensor(metas, meta, group=group)
    metas_cpu = metas.view(world, -1).cpu()
    word_or = 0
    for r in range(world):
        word_or |= int(metas_cpu[r, 3 + _MAX_TRAIL])
    metas_cpu = metas_cpu[:, : 3 + _MAX_TRAIL]
    owners = [r for r in range(world) if int(metas_cpu[r, 1]) >= 0]
    if not owners:
        return None, word_or
    ref = metas_cpu[owners[0]]
    for r in owners[1:]:
        if not torch.equal(metas_cpu[r, 1:], ref[1:]):
            raise RuntimeError("list states must have the same dtype and trailing shape on every rank to be synced")
    dtype = _DTYPES[int(ref[1])]
    trail = [int(d) for d in ref[3 : 3 + int(ref[2])]]
    rows = [int(metas_cpu[r, 0]) if r in owners else 0 for r in range(world)]
    most = max(rows)
    if local is None:
        local = torch.empty((0, *trail), dtype=dtype, device=device)
    if local.shape[0] < most:
        send = torch.empty((most, *trail), dtype=dtype, device=device)
        send[: local.shape[0]] = local
    else:
        send = local
    # NCCL moves a fixed list of dtypes; an all-gather only copies bits, so the others travel as a same-sized one
    wire_dtype = _WIRE.get(dtype)
    wire = send.view(wire_dtype) if wire_dtype is not None else send
    recv = torch.empty((world * most, *trail), dtype=wire.dtype, device=device)
    dist.all_gather_into_tensor(recv, wire, group=group)
    if wire_dtype is not None:
        recv = recv.view(dtype)
    if all(r == most for r in rows):
        return recv, word_or
    recv = recv.view(world, most, *trail)
    return torch.cat([recv[r, : rows[r]] for r in range(world)], dim=0), word_or
