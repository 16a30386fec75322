"""Label-sharded compute of the exact-mode scalar metrics (AUROC / AveragePrecision, ``thresholds=None``) over
several GPUs — the K4 path for "cat" states.

The reference synchronises such a metric by all-gathering every rank's scores and targets (metric.py:423-453,
utilities/distributed.py:97-147: 12 bytes per label-sample to every rank) and then lets EVERY rank sort EVERY label
(roc.py:345-355): N GPUs do N times the work of one.  Here ``compute()`` deals the labels out to the ranks instead:

1. one small all-gather of (local row count, data-error word);
2. every rank turns its local rows into 32-bit sort keys and its pack kernel stores the keys of label ``l`` straight
   into the key buffer of the rank that owns ``l`` — symmetric memory, plain stores over NVLink / NVSwitch, 4 bytes per
   label-sample, no staging copy and no separate all-to-all (``tmb_clf_pack_keys_sharded``);
3. a barrier, then every rank sorts and reduces only its own ``ceil(L / world)`` labels (``tmb_clf_sorted_scores``);
4. one all-gather of the ``[L, 4]`` float64 reductions.

An explicit ``metric.sync()`` still performs the literal all-gather (``metric.preds`` is then the rank-ordered
concatenation, as in the reference).  ``TMB_EXACT_SHARDED=0`` switches this path off; ``TMB_EXACT_EXCHANGE=nccl``
replaces step 2's remote stores by a local pack + ``all_to_all_single`` (also the fallback when symmetric memory
cannot be set up).
"""
import os
from typing import Any, Dict, List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist
from torch import Tensor

from torchmetrics_forked_b200 import _native as nat

_HDR_ALIGN = 256


class ShardedChunks(list):
    """The local per-update buffers of a list state plus what the label-sharded compute needs to know about the other
    ranks.  ``kernels.clf_curve_scores`` recognises it and runs the sharded exchange instead of a local sort."""

    def __init__(self, chunks: Sequence[Tensor], info: "ShardInfo") -> None:
        super().__init__(chunks)
        self.shard_info = info


class ShardInfo:
    def __init__(self, group: Any, rows: List[int], num_labels: int, device: torch.device) -> None:
        self.group, self.rows, self.num_labels, self.device = group, rows, num_labels, device


def enabled() -> bool:
    return os.environ.get("TMB_EXACT_SHARDED", "1") != "0"


def exchange_rows_and_errors(n_local: int, err_word: int, device: torch.device, group: Any) -> Tuple[List[int], int]:
    """One all-gather of (rows, error bits) per rank -> (rows of every rank, OR of the error words)."""
    world = dist.get_world_size(group)
    mine = torch.tensor([n_local, err_word], dtype=torch.int64, device=device)
    every = torch.empty(world * 2, dtype=torch.int64, device=device)
    dist.all_gather_into_tensor(every, mine, group=group)
    every = every.view(world, 2).cpu()
    err = 0
    for r in range(world):
        err |= int(every[r, 1])
    return [int(v) for v in every[:, 0]], err


class _Exchange:
    """The symmetric key buffer of one (process group, device): allocated once, grown when a compute needs more."""

    def __init__(self, group: Any, device: torch.device) -> None:
        self.group, self.device = group, device
        self.buf: Optional[Tensor] = None
        self.hdl: Any = None
        self.ptr_table: Optional[Tensor] = None

    def ensure(self, nbytes: int) -> None:
        if self.buf is not None and self.buf.numel() >= nbytes:
            return
        import torch.distributed._symmetric_memory as symm

        grain = 64 << 20
        cap = (nbytes + grain - 1) // grain * grain
        self.buf = None  # drop the old allocation first
        self.hdl = None
        buf = symm.empty(cap, dtype=torch.uint8, device=self.device)
        self.hdl = symm.rendezvous(buf, self.group)
        self.buf = buf
        self.ptr_table = None

    def peer_tables(self, hdr: int) -> Tensor:
        """int64 [2, world] on the device: row 0 = key-buffer addresses, row 1 = n_valid addresses."""
        if self.ptr_table is None or self.ptr_table[2] != hdr:
            ptrs = [int(p) for p in self.hdl.buffer_ptrs]
            t = torch.tensor([[p + hdr for p in ptrs], ptrs], dtype=torch.int64, device=self.device)
            self.ptr_table = (t, None, hdr)
        return self.ptr_table[0]

    def barrier(self) -> None:
        self.hdl.barrier()


_EXCHANGES: Dict[Any, _Exchange] = {}
_SYMM_BROKEN = False
LAST = {"exchange": None, "labels_per_rank": None, "rows": None}  # what the most recent sharded compute did (diagnostics)


def _exchange_for(group: Any, device: torch.device) -> _Exchange:
    key = (id(group), device.index)
    ex = _EXCHANGES.get(key)
    if ex is None:
        ex = _Exchange(group, device)
        _EXCHANGES[key] = ex
    return ex


def _chunk_table(ps: List[Tensor], ts: List[Tensor], device: torch.device) -> Tensor:
    starts = [0]
    for p in ps:
        starts.append(starts[-1] + p.shape[0])
    return torch.tensor(starts + [p.data_ptr() for p in ps] + [t.data_ptr() for t in ts], dtype=torch.int64,
                        device=device)


def sharded_curve_scores(
    ps: List[Tensor], ts: List[Tensor], info: ShardInfo, ignore_index: Optional[int], pos_label: int, want_ap: bool,
    class_targets: bool,
) -> Tensor:  # fmt: skip
    """Global ``[L, 4]`` float64 reductions (same rows as ``kernels.clf_curve_scores``) from every rank's LOCAL chunks."""
    global _SYMM_BROKEN
    group = dist.group.WORLD if info.group is None else info.group
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    dev, L = info.device, info.num_labels
    rows = info.rows
    n_total, n_local, row_offset = sum(rows), rows[rank], sum(rows[:rank])
    per = (L + world - 1) // world                      # labels per rank (block distribution)
    mine = max(0, min(per, L - rank * per))            # labels this rank owns
    out_all = torch.zeros(world * (per + 1), 4, dtype=torch.float64, device=dev)
    if n_total == 0:
        return out_all[:L]
    has_ign = int(ignore_index is not None)
    err = torch.zeros(1, dtype=torch.int32, device=dev)
    hdr = (per * 8 + _HDR_ALIGN - 1) // _HDR_ALIGN * _HDR_ALIGN
    lib = nat.lib()
    pcode = nat.float_code(ps[0]) if ps else 0
    tcode = nat.label_code(ts[0]) if ts else 0
    use_symm = os.environ.get("TMB_EXACT_EXCHANGE", "symm") != "nccl" and not _SYMM_BROKEN
    keys_view: Optional[Tensor] = None
    nvalid_view: Optional[Tensor] = None
    with nat.on_device(dev):
        st = nat.stream_ptr(dev)
        if use_symm:
            ex = _exchange_for(group, dev)
            try:
                ex.ensure(hdr + per * n_total * 4)
            except Exception as e:  # noqa: BLE001  (no peer access / no symmetric allocator: every rank fails alike)
                _SYMM_BROKEN, use_symm = True, False
                if rank == 0:
                    print(f"torchmetrics_forked_b200: symmetric memory unavailable ({e!r}); using all_to_all", flush=True)
        if use_symm:
            ex.buf[:hdr].zero_()                        # this rank's n_valid counters
            ex.barrier()                                # ... zeroed everywhere, and the previous compute is done with the buffer
            if n_local > 0:
                table = _chunk_table(ps, ts, dev)
                peers = ex.peer_tables(hdr)
                rc = lib.tmb_clf_pack_keys_sharded(
                    nat.ptr(table), len(ps), pcode, tcode, n_local, L, has_ign, int(ignore_index) if has_ign else 0,
                    int(pos_label), int(class_targets), peers[0].data_ptr(), peers[1].data_ptr(), world, per, n_total,
                    row_offset, nat.ptr(err), st,
                )  # fmt: skip
                nat.check(rc, "tmb_clf_pack_keys_sharded")
            ex.barrier()                                # every rank's keys have landed
            keys_view = ex.buf[hdr : hdr + per * n_total * 4]
            nvalid_view = ex.buf[:hdr]
        else:
            # local pack into [world * per, n_local] label-major keys, then one all_to_all of whole label blocks
            local_keys = torch.empty(world * per * max(n_local, 1), dtype=torch.int32, device=dev)
            local_nv = torch.zeros(world * per, dtype=torch.int64, device=dev)
            if n_local > 0:
                table = _chunk_table(ps, ts, dev)
                self_ptrs = torch.tensor([[local_keys.data_ptr()], [local_nv.data_ptr()]], dtype=torch.int64, device=dev)
                rc = lib.tmb_clf_pack_keys_sharded(
                    nat.ptr(table), len(ps), pcode, tcode, n_local, L, has_ign, int(ignore_index) if has_ign else 0,
                    int(pos_label), int(class_targets), self_ptrs[0].data_ptr(), self_ptrs[1].data_ptr(), 1, world * per,
                    n_local, 0, nat.ptr(err), st,
                )  # fmt: skip
                nat.check(rc, "tmb_clf_pack_keys_sharded")
            recv = torch.empty(per * n_total, dtype=torch.int32, device=dev)
            dist.all_to_all_single(recv, local_keys[: world * per * n_local], [per * r for r in rows],
                                   [per * n_local] * world, group=group)
            blocks, at = [], 0
            for r in range(world):                       # [per, rows[r]] per source -> [per, n_total]
                blocks.append(recv[at : at + per * rows[r]].view(per, rows[r]))
                at += per * rows[r]
            keys_view = torch.cat(blocks, dim=1).contiguous().view(torch.uint8).view(-1)
            dist.all_reduce(local_nv, group=group)
            nvalid_view = local_nv[rank * per : (rank + 1) * per].contiguous().view(torch.uint8)
        LAST.update(exchange="symmetric-memory remote stores" if use_symm else "nccl all_to_all", labels_per_rank=per,
                    rows=list(rows))
        send = out_all[rank * (per + 1) : (rank + 1) * (per + 1)]
        if mine > 0:
            nbytes = int(lib.tmb_clf_sorted_scores_scratch_bytes(n_total, mine))
            scratch = torch.empty(max(nbytes, 1), dtype=torch.uint8, device=dev)
            scores = torch.empty(mine, 4, dtype=torch.float64, device=dev)
            rc = lib.tmb_clf_sorted_scores(
                keys_view.data_ptr(), nvalid_view.data_ptr(), n_total, mine, int(not has_ign), int(want_ap),
                nat.ptr(scratch), scratch.numel(), nat.ptr(scores), st,
            )
            nat.check(rc, "tmb_clf_sorted_scores")
            send[:mine] = scores
        send[per, 0] = err.to(torch.float64)[0]        # a negative score seen by this rank's pack poisons the result
        gathered = torch.empty_like(out_all)
        dist.all_gather_into_tensor(gathered, send.contiguous(), group=group)
    gathered = gathered.view(world, per + 1, 4)
    bad = (gathered[:, per, 0] != 0).any()
    out = gathered[:, :per].reshape(world * per, 4)[:L]
    return torch.where(bad, torch.full_like(out, float("nan")), out)
