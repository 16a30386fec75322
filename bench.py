#!/usr/bin/env python
"""bench.py — throughput of the classification-metric update path on B200 (driver contract).

    python bench.py --gpus N --steps K --warmup W            # our CUDA path (one JSON line on stdout)
    python bench.py --impl reference --steps K --warmup W    # the reference's own CPU path (oracle/_ref), all host threads
    torchrun ... bench.py --gpus N ...                        # N > 1: one rank per GPU, NCCL; adds the sync parity
                                                              # prologue and the strong-scaling block

Headline workload (BASELINE.json configs[1]): `MulticlassConfusionMatrix(1000)`, one `update` = bf16 logits
[2^22, 1000] + int64 targets per GPU (8.42 GB, far larger than the 126 MB L2, so no flush is needed),
samples independent -> sharded by sample, weak scaling, one `compute()` (packed all-reduce sync) closes the
timed region.  `value` = samples/s over all GPUs with inputs resident in HBM; `e2e` = the same through the
public Metric API from pinned host buffers (H2D inside the timed region, accuracy read back every step).
`extras` carries the other BASELINE configs (1, 3, 4, 5) measured the same way on rank 0 at N = 1, each beside the
UNMODIFIED reference run on the same GPU (`ref_cuda_ms`, torch's stock CUDA kernels; `oracle/_ref`).  At N > 1 the run
starts with an untimed parity prologue of the state sync (`sync_parity`) and `extras.strong` holds the strong-scaling
measurement of configs 2, 3 and 4 (the global batch BASELINE.json names, sharded over the N ranks, sync included).
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent
for _p in (str(ROOT), str(ROOT / "oracle")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

C2_ROWS, C2_CLASSES = 1 << 22, 1000
METRIC = "samples/s per update, MulticlassConfusionMatrix(C=1000), bf16 logits [2^22,1000] per GPU"


# ------------------------------------------------------------------------------------------ helpers
def measured_peak_gbs():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        try:
            return float(json.loads(p.read_text())["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:  # noqa: BLE001
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def recorded_traffic(kernel: str):
    """DRAM bytes per launch from the committed `ncu --set full` summary, if one exists for this kernel."""
    p = ROOT / "profiles" / "traffic.json"
    if p.exists():
        try:
            return json.loads(p.read_text()).get(kernel)
        except Exception:  # noqa: BLE001
            return None
    return None


class ClockSampler:
    """SM clock / throttle reasons sampled WHILE the timed region runs: NVML in a thread every ~5 ms (the default
    timed region is only tens of milliseconds long, which an `nvidia-smi -lms` child process mostly misses), with
    `nvidia-smi` as the fall-back when NVML cannot be loaded."""

    FIELDS = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self.proc, self.thread = index, [], None, None
        self.nvml, self.handle, self.stop = None, None, threading.Event()
        self.sm, self.mx, self.reason_bits = [], [], 0

    def __enter__(self):
        try:
            import pynvml

            pynvml.nvmlInit()
            phys = self.index
            visible = os.environ.get("CUDA_VISIBLE_DEVICES")
            if visible:  # NVML numbers the physical devices
                ids = [v.strip() for v in visible.split(",") if v.strip()]
                if self.index < len(ids) and ids[self.index].isdigit():
                    phys = int(ids[self.index])
            self.handle = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.nvml = pynvml
            self.mx.append(float(pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM)))
            self.thread = threading.Thread(target=self._poll_nvml, daemon=True)
            self.thread.start()
            return self
        except Exception:  # noqa: BLE001
            self.nvml = None
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except Exception:  # noqa: BLE001
            self.proc = None
        return self

    def _poll_nvml(self):
        n = self.nvml
        while not self.stop.is_set():
            try:
                self.sm.append(float(n.nvmlDeviceGetClockInfo(self.handle, n.NVML_CLOCK_SM)))
                self.reason_bits |= int(n.nvmlDeviceGetCurrentClocksThrottleReasons(self.handle))
            except Exception:  # noqa: BLE001
                break
            self.stop.wait(0.005)

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def __exit__(self, *exc):
        if self.nvml is not None:
            self.stop.set()
            if self.thread is not None:
                self.thread.join(timeout=1)
            return
        if self.proc is not None:
            time.sleep(0.15)
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:  # noqa: BLE001
                self.proc.kill()

    def summary(self):
        if self.nvml is not None:
            n = self.nvml
            flags = {"hw_slowdown": n.nvmlClocksThrottleReasonHwSlowdown,
                     "hw_thermal_slowdown": n.nvmlClocksThrottleReasonHwThermalSlowdown,
                     "sw_thermal_slowdown": n.nvmlClocksThrottleReasonSwThermalSlowdown,
                     "sw_power_cap": n.nvmlClocksThrottleReasonSwPowerCap}
            reasons = sorted(k for k, bit in flags.items() if self.reason_bits & bit)
            return {"sm_mhz": statistics.median(self.sm) if self.sm else None,
                    "sm_max_mhz": max(self.mx) if self.mx else None, "reasons": reasons, "samples": len(self.sm),
                    "source": "nvml"}
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "", 1).isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "", 1).isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows if len(r) >= 6 for n, v in zip(names, r[2:6]) if v.lower() == "active"})
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm), "source": "nvidia-smi"}


def dist_setup(n_gpus: int):
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    return rank, world, local


def max_over_ranks(x: float, world: int, device) -> float:
    if world == 1:
        return x
    import torch.distributed as dist

    t = torch.tensor([x], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def barrier(world: int):
    if world > 1:
        import torch.distributed as dist

        dist.barrier()
    torch.cuda.synchronize()


def make_cfg2(rows: int, device, seed: int):
    """Config 2 inputs: bf16 logits with +4 on the target column for 70 % of the rows (diagonal heavy)."""
    g = torch.Generator(device=device).manual_seed(seed)
    target = torch.randint(C2_CLASSES, (rows,), generator=g, device=device)
    preds = torch.empty(rows, C2_CLASSES, dtype=torch.bfloat16, device=device)
    chunk = 1 << 18
    for s in range(0, rows, chunk):
        e = min(rows, s + chunk)
        x = torch.randn(e - s, C2_CLASSES, generator=g, device=device)
        hot = torch.rand(e - s, generator=g, device=device) < 0.7
        x[hot, target[s:e][hot]] += 4
        preds[s:e] = x.bfloat16()
    return preds, target


def timed_loop(fn, steps: int, warmup: int, world: int, device, tail=None):
    """W untimed calls, then K timed calls between barrier+sync fences. Returns (total ms max-over-ranks,
    mean per-call ms on this rank from per-step events)."""
    for _ in range(warmup):
        fn()
    barrier(world)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
    end = torch.cuda.Event(enable_timing=True)
    ev[0].record()
    for i in range(steps):
        fn()
        ev[i + 1].record()
    if tail is not None:
        tail()
    end.record()
    barrier(world)
    total = ev[0].elapsed_time(end)
    per_step = [ev[i].elapsed_time(ev[i + 1]) for i in range(steps)]
    return max_over_ranks(total, world, device), sum(per_step) / len(per_step)


# ------------------------------------------------------------------------------------------ reference arms
def host_threads() -> int:
    """Give torch's intra-op pool every host core this process may use.  `torch.distributed.run` exports
    OMP_NUM_THREADS=1 to its workers, which would leave the CPU arm on ONE thread."""
    try:
        n = len(os.sched_getaffinity(0))
    except AttributeError:
        n = os.cpu_count() or 1
    torch.set_num_threads(max(n, 1))
    return torch.get_num_threads()


_REF = {"tried": False, "module": None}


def reference_module():
    """The unmodified reference package from `oracle/_ref` (staged by oracle/build_ref.py), or None."""
    if not _REF["tried"]:
        _REF["tried"] = True
        try:
            import ref_loader

            _REF["module"] = ref_loader.load_staged_reference()
        except Exception as e:  # noqa: BLE001
            print(f"reference package not importable: {e!r}", file=sys.stderr)
    return _REF["module"]


def cpu_confmat_rate(rows: int, steps: int, warmup: int, budget_s: float = 25.0):
    """The reference's CPU update path on a bounded sample of config 2: the unmodified
    `MulticlassConfusionMatrix(1000).update` when `oracle/_ref` is present (kind "reference"), else the op-for-op
    ATen port of it (kind "port")."""
    cores = host_threads()
    g = torch.Generator().manual_seed(1234)
    target = torch.randint(C2_CLASSES, (rows,), generator=g)
    preds = torch.randn(rows, C2_CLASSES, generator=g)
    hot = torch.rand(rows, generator=g) < 0.7
    preds[hot, target[hot]] += 4
    preds = preds.bfloat16()
    ref = reference_module()
    if ref is not None:
        metric = ref.classification.MulticlassConfusionMatrix(C2_CLASSES, validate_args=False)
        kind = "reference"

        def step():
            metric.update(preds, target)
    else:
        import torch_cpu as ref_port

        state = torch.zeros(C2_CLASSES, C2_CLASSES, dtype=torch.long)
        kind = "port"

        def step():
            state.add_(ref_port.multiclass_confusion_matrix_update(preds, target, C2_CLASSES, validate_args=False))
    for _ in range(warmup):
        step()
    t0 = time.perf_counter()
    done = 0
    for _ in range(steps):
        step()
        done += 1
        if time.perf_counter() - t0 > budget_s:
            break
    dt = time.perf_counter() - t0
    return rows * done / dt, dt / done * 1e3, done, kind, cores


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", 0))
    if rank != 0:
        return
    rows = args.cpu_rows
    rate, ms, done, kind, cores = cpu_confmat_rate(rows, args.steps, max(args.warmup, 1), budget_s=120.0)
    print(f"reference arm: kind={kind}, {cores} host threads (OMP_NUM_THREADS was "
          f"{os.environ.get('OMP_NUM_THREADS')!r}), {ms:.2f} ms per {rows} rows", file=sys.stderr)
    what = ("the unmodified reference (oracle/_ref: torchmetrics 1.3.0dev, MulticlassConfusionMatrix.update)"
            if kind == "reference" else "op-for-op torch-CPU port of the reference (oracle/torch_cpu.py; oracle/_ref absent)")
    sample = f"{done} updates of bf16 logits [{rows},1000] + int64 targets (config 2 rows, bounded sample)"
    line = {
        "impl": "reference", "metric": METRIC, "value": rate, "unit": "samples/s", "n_gpus": args.gpus,
        "steps": done, "warmup": max(args.warmup, 1), "ms_per_step": ms, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": "cfg2: MulticlassConfusionMatrix(1000).update, validate_args=False", "rows_per_step": rows,
                   "classes": C2_CLASSES, "note": f"reference algorithm on host cores: {what}", "host_threads": cores},
        "cpu_baseline": {"value": rate, "unit": "samples/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": rate, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


def time_on_device(fn, reps: int, warmup: int = 1) -> float:
    """Mean CUDA-event milliseconds of `fn` over `reps` calls after `warmup` calls."""
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def ref_cuda_update_ms(make_metric, inputs, reps: int = 3):
    """The unmodified reference metric on the SAME device and inputs: ms per `update`, validate_args False / True.
    Returns {} when `oracle/_ref` is absent.  `make_metric(ref_module, validate_args)` builds the reference object."""
    ref = reference_module()
    if ref is None:
        return {}
    out = {}
    for key, validate in (("ref_cuda_ms", False), ("ref_cuda_ms_validate_args", True)):
        try:
            m = make_metric(ref, validate).to(inputs[0].device)
            out[key] = time_on_device(lambda m=m: m.update(*inputs), reps)
            del m
        except Exception as e:  # noqa: BLE001
            out[key] = None
            out[key + "_error"] = repr(e)[:200]
        torch.cuda.empty_cache()
    return out


# ------------------------------------------------------------------------------------------ extras (other configs)
def extras_block(device, steps: int, warmup: int, peak: float):
    """BASELINE configs 3, 5 and 1 on one GPU: same timing rules, reported beside the headline."""
    import torchmetrics_forked_b200 as tm
    from torchmetrics_forked_b200.classification import (
        BinaryPrecisionRecallCurve,
        MulticlassAccuracy,
        MulticlassConfusionMatrix,
        MulticlassJaccardIndex,
    )

    out = {}
    g = torch.Generator(device=device).manual_seed(4321)

    def record(name, fn, samples, algo_bytes, preds_bytes, launches, ref=None):
        _, ms = timed_loop(fn, steps, warmup, 1, device)
        out[name] = {
            "ms_per_update": ms, "samples_per_s": samples / ms * 1e3, "preds_gbs": preds_bytes / ms / 1e6,
            "algo_gbs": algo_bytes / ms / 1e6, "roofline_frac": algo_bytes / ms / 1e6 / peak,
            "launches_per_update": launches,
        }
        if ref is not None:  # the unmodified reference on the same GPU and the same tensors (torch's stock kernels)
            out[name].update(ref_cuda_update_ms(*ref))
            if out[name].get("ref_cuda_ms"):
                out[name]["speedup_vs_ref_cuda"] = out[name]["ref_cuda_ms"] / ms

    # config 3: BinaryPrecisionRecallCurve(thresholds=200), 2^28 fp32 logits
    n3 = 1 << 28
    logits = torch.randn(n3, generator=g, device=device) * 3
    for tdt, tag, tb in ((torch.int64, "cfg3_int64_targets", 8), (torch.uint8, "cfg3_uint8_targets", 1)):
        tgt = torch.randint(2, (n3,), generator=g, device=device).to(tdt)
        m = BinaryPrecisionRecallCurve(thresholds=200, validate_args=False).to(device)
        ref3 = (lambda R, v: R.classification.BinaryPrecisionRecallCurve(thresholds=200, validate_args=v),
                (logits, tgt), 2) if tdt == torch.int64 else None
        record(tag, lambda m=m, tgt=tgt: m.update(logits, tgt), n3, n3 * (4 + tb), n3 * 4, 4, ref=ref3)
        del tgt, m
    probs = torch.rand(n3, generator=g, device=device)
    tgt = torch.randint(2, (n3,), generator=g, device=device)
    m = BinaryPrecisionRecallCurve(thresholds=200, validate_args=False).to(device)
    record("cfg3_probs_int64_targets", lambda: m.update(probs, tgt), n3, n3 * 12, n3 * 4, 4)
    del probs, m
    # K1b on the same data: binary stat scores / accuracy (one strict threshold, four counters)
    from torchmetrics_forked_b200.classification import (
        BinaryAccuracy,
        MultilabelPrecisionRecallCurve,
        MultilabelStatScores,
        MulticlassStatScores,
    )

    m = BinaryAccuracy(validate_args=False).to(device)
    record("k1b_binary_accuracy_2^28_logits", lambda: m.update(logits, tgt), n3, n3 * 12, n3 * 4, 4)
    del logits, tgt, m
    torch.cuda.empty_cache()
    # multilabel: [2^21, 100] fp32 logits + int64 targets, stat scores and binned curve (T = 200)
    nl, L = 1 << 21, 100
    ml_p = torch.randn(nl, L, generator=g, device=device)
    ml_t = torch.randint(2, (nl, L), generator=g, device=device)
    m = MultilabelStatScores(L, validate_args=False).to(device)
    record("k1b_multilabel_stat_scores_2^21x100", lambda: m.update(ml_p, ml_t), nl, nl * L * 12, nl * L * 4, 4)
    m = MultilabelPrecisionRecallCurve(L, thresholds=200, validate_args=False).to(device)
    record("k2_multilabel_curve_T200_2^21x100", lambda: m.update(ml_p, ml_t), nl, nl * L * 12, nl * L * 4, 4)
    del ml_p, ml_t, m
    torch.cuda.empty_cache()
    # row f4: MulticlassAccuracy + MulticlassConfusionMatrix in one MetricCollection at the config-2 row shape: one
    # pass over the logits feeds both (the bytes counted are ONE read of the batch)
    from torchmetrics_forked_b200 import MetricCollection
    from torchmetrics_forked_b200.classification import MulticlassAccuracy
    from torchmetrics_forked_b200.classification import MulticlassConfusionMatrix as _MCM

    r4 = 1 << 20
    p4, t4 = make_cfg2(r4, device, 55)
    coll = MetricCollection([MulticlassAccuracy(C2_CLASSES, average="micro", validate_args=False),
                             _MCM(C2_CLASSES, validate_args=False)]).to(device)
    coll.update(p4, t4)  # the first update discovers the compute groups; sharing starts with the second
    record("f4_collection_accuracy+confmat_2^20_rows", lambda: coll.update(p4, t4), r4, r4 * 2008, r4 * 2000, 3)
    del p4, t4, coll
    torch.cuda.empty_cache()
    # row f3: top-5 accuracy at the config-2 row shape (bf16 [2^20, 1000])

    r5 = 1 << 20
    p5, t5 = make_cfg2(r5, device, 77)
    m = MulticlassAccuracy(C2_CLASSES, top_k=5, average="micro", validate_args=False).to(device)
    record("f3_top5_accuracy_2^20_rows", lambda: m.update(p5, t5), r5, r5 * 2008, r5 * 2000, 2)
    del p5, m
    torch.cuda.empty_cache()
    p5f = torch.randn(r5 // 2, C2_CLASSES, generator=g, device=device)
    t5f = t5[: r5 // 2].contiguous()
    m = MulticlassAccuracy(C2_CLASSES, top_k=5, average="micro", validate_args=False).to(device)
    record("f3_top5_accuracy_fp32_2^19_rows", lambda: m.update(p5f, t5f), r5 // 2, (r5 // 2) * 4008, (r5 // 2) * 4000, 2)
    del p5f, t5f, t5, m
    torch.cuda.empty_cache()
    # row f2: multiclass binned curves, fused row softmax + binning (MulticlassAUROC(thresholds=200) update)
    from torchmetrics_forked_b200.classification import MulticlassAUROC

    for rows_mc, C_mc in ((1 << 21, 100), (1 << 24, 10), (1 << 18, 1000)):
        mc_p = torch.randn(rows_mc, C_mc, generator=g, device=device)
        mc_t = torch.randint(C_mc, (rows_mc,), generator=g, device=device)
        m = MulticlassAUROC(C_mc, thresholds=200, validate_args=False).to(device)
        record(f"f2_multiclass_curve_T200_{rows_mc}x{C_mc}", lambda: m.update(mc_p, mc_t), rows_mc,
               rows_mc * (C_mc * 4 + 8), rows_mc * C_mc * 4, 4)
        del mc_p, mc_t, m
        torch.cuda.empty_cache()
    # row f2, exact mode: MulticlassAUROC(thresholds=None) — update = conditional row softmax (read + write), compute =
    # sort + scan of all classes
    try:
        for rows_s, C_s in ((1 << 24, 10),):  # narrow rows: one thread per row
            ms_ = MulticlassAUROC(C_s, thresholds=None, average="macro", validate_args=False).to(device)
            sp = torch.randn(rows_s, C_s, generator=g, device=device)
            st_ = torch.randint(C_s, (rows_s,), generator=g, device=device)

            def upd_s():
                ms_.reset()
                ms_.update(sp, st_)

            record(f"f2_multiclass_exact_update_{rows_s}x{C_s}", upd_s, rows_s, rows_s * (C_s * 8 + 8), rows_s * C_s * 4, 2)
            del ms_, sp, st_
            torch.cuda.empty_cache()
        rows_e, C_e = 1 << 21, 100
        m = MulticlassAUROC(C_e, thresholds=None, average="macro", validate_args=False).to(device)
        mc_p = torch.randn(rows_e, C_e, generator=g, device=device)
        mc_t = torch.randint(C_e, (rows_e,), generator=g, device=device)

        def upd():
            m.reset()
            m.update(mc_p, mc_t)

        record(f"f2_multiclass_exact_update_{rows_e}x{C_e}", upd, rows_e, rows_e * (C_e * 8 + 8), rows_e * C_e * 4, 2)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        m.compute()          # warm-up (allocator, first-call set-up)
        m._computed = None   # drop the cached result so that the timed call really computes
        e0.record()
        val = m.compute()
        e1.record()
        torch.cuda.synchronize()
        out[f"f2_multiclass_exact_update_{rows_e}x{C_e}"]["compute_ms"] = e0.elapsed_time(e1)
        out[f"f2_multiclass_exact_update_{rows_e}x{C_e}"]["auroc"] = float(val)
        del m, mc_p, mc_t
    except Exception as e:  # noqa: BLE001
        out["f2_multiclass_exact"] = {"error": repr(e)}
    torch.cuda.empty_cache()
    # config 2's kernel under atomic contention: the state is updated with one RED.64 per row, so what matters is how
    # many rows hit the same cell.  (a) a collapsed model: every row lands in ONE cell; (b) all mass on the diagonal of
    # 10 classes.  Beside them the bench data (70 % of the rows spread over the 1000 diagonal cells).
    r7 = 1 << 20
    p7, t7 = make_cfg2(r7, device, 31)
    m = _MCM(C2_CLASSES, validate_args=False).to(device)
    record("cfg2_contention_bench_data_2^20_rows", lambda: m.update(p7, t7), r7, r7 * 2008, r7 * 2000, 1)
    p7[:, 7] = 30.0
    t7.fill_(7)
    record("cfg2_contention_all_rows_one_cell", lambda: m.update(p7, t7), r7, r7 * 2008, r7 * 2000, 1)
    t7 = torch.randint(10, (r7,), generator=g, device=device)
    p7.copy_(torch.randn(r7, C2_CLASSES, generator=g, device=device).bfloat16())
    p7[torch.arange(r7, device=device), t7] += 12
    record("cfg2_contention_diagonal_of_10_classes", lambda: m.update(p7, t7), r7, r7 * 2008, r7 * 2000, 1)
    base_ms = out["cfg2_contention_bench_data_2^20_rows"]["ms_per_update"]
    for k in ("cfg2_contention_all_rows_one_cell", "cfg2_contention_diagonal_of_10_classes"):
        out[k]["vs_bench_data"] = out[k]["ms_per_update"] / base_ms
    del p7, t7, m
    torch.cuda.empty_cache()
    # config 2 rows through StatScores (three C-bin histograms instead of the C x C matrix)
    r2 = 1 << 20
    p2, t2 = make_cfg2(r2, device, 99)
    m = MulticlassStatScores(C2_CLASSES, average=None, validate_args=False).to(device)
    record("cfg2_stat_scores_2^20_rows", lambda: m.update(p2, t2), r2, r2 * 2008, r2 * 2000, 2)
    del p2, t2, m
    torch.cuda.empty_cache()

    # config 5: MulticlassJaccardIndex(150), logits [16,150,512,512]
    for dt, tag, eb in ((torch.float32, "cfg5_fp32", 4), (torch.bfloat16, "cfg5_bf16", 2)):
        preds = torch.randn(16, 150, 512, 512, generator=g, device=device).to(dt)
        tgt = torch.randint(150, (16, 512, 512), generator=g, device=device)
        m = MulticlassJaccardIndex(150, validate_args=False).to(device)
        px = 16 * 512 * 512
        record(tag, lambda m=m, preds=preds, tgt=tgt: m.update(preds, tgt), px, px * (150 * eb + 8), px * 150 * eb, 1,
               ref=(lambda R, v: R.classification.MulticlassJaccardIndex(150, validate_args=v), (preds, tgt)))
        del preds, tgt, m
    torch.cuda.empty_cache()

    # config 4: MultilabelAUROC(100, thresholds=None), 2^24 samples over 8 updates of 2^21; exact compute()
    try:
        from torchmetrics_forked_b200.classification import MultilabelAUROC

        n4, L4, upd = 1 << 21, 100, 8
        m = MultilabelAUROC(L4, thresholds=None, validate_args=False).to(device)
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(upd + 3)]
        batches = []
        for _ in range(upd):
            batches.append((torch.randn(n4, L4, generator=g, device=device),
                            torch.randint(2, (n4, L4), generator=g, device=device)))
        # epoch 0 warms torch's caching allocator (the stored scores are 839 MB per update); epoch 1 is timed
        for p, t in batches:
            m.update(p, t)
        m.compute()
        m.reset()
        torch.cuda.synchronize()
        for i, (p, t) in enumerate(batches):
            ev[i].record()
            m.update(p, t)
        ev[upd].record()
        del batches
        value = m.compute()
        ev[upd + 1].record()
        torch.cuda.synchronize()
        upd_ms = ev[0].elapsed_time(ev[upd]) / upd
        cmp_ms = ev[upd].elapsed_time(ev[upd + 1])
        ls = n4 * upd * L4
        pr = torch.rand(n4, L4, generator=g, device=device)  # inputs that already are probabilities: one copy pass
        tr = torch.randint(2, (n4, L4), generator=g, device=device)
        mp = MultilabelAUROC(L4, thresholds=None, validate_args=False).to(device)

        def upd_probs():
            mp.reset()
            mp.update(pr, tr)

        _, probs_ms = timed_loop(upd_probs, 5, 2, 1, device)
        del pr, tr, mp
        out["cfg4_multilabel_auroc_exact"] = {
            "update_probs_input_ms": probs_ms, "update_probs_input_roofline_frac": n4 * L4 * 8 / probs_ms / 1e6 / peak,
            "update_ms": upd_ms, "update_algo_gbs": n4 * L4 * 8 / upd_ms / 1e6,
            "update_roofline_frac": n4 * L4 * 8 / upd_ms / 1e6 / peak,
            "compute_ms": cmp_ms, "compute_label_samples_per_s": ls / cmp_ms * 1e3,
            "compute_gbs_lower_bound_24B": ls * 24 / cmp_ms / 1e6, "compute_gbs_radix_model_74B": ls * 74 / cmp_ms / 1e6,
            "auroc": float(value), "samples": n4 * upd, "labels": L4,
        }
        del m, value
        torch.cuda.empty_cache()
    except Exception as e:  # noqa: BLE001
        out["cfg4_multilabel_auroc_exact"] = {"error": repr(e)}
        torch.cuda.empty_cache()
    # the unmodified reference on the same GPU for config 4: 8 updates (sigmoid + append) and the exact compute()
    # (dim_zero_cat, then a Python loop of 100 x (argsort, gathers, cumsum))
    ref = reference_module()
    if ref is not None and "error" not in out.get("cfg4_multilabel_auroc_exact", {"error": 1}):
        try:
            n4, L4, upd = 1 << 21, 100, 8
            g4 = torch.Generator(device=device).manual_seed(4321)
            rm = ref.classification.MultilabelAUROC(L4, thresholds=None, validate_args=False).to(device)
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
            batches = [(torch.randn(n4, L4, generator=g4, device=device),
                        torch.randint(2, (n4, L4), generator=g4, device=device)) for _ in range(upd)]
            rm.update(*batches[0])
            rm.reset()
            torch.cuda.synchronize()
            ev[0].record()
            for p, t in batches:
                rm.update(p, t)
            ev[1].record()
            del batches
            rv = rm.compute()
            ev[2].record()
            torch.cuda.synchronize()
            o4 = out["cfg4_multilabel_auroc_exact"]
            o4["ref_cuda_update_ms"] = ev[0].elapsed_time(ev[1]) / upd
            o4["ref_cuda_compute_ms"] = ev[1].elapsed_time(ev[2])
            o4["ref_cuda_auroc"] = float(rv)
            o4["compute_speedup_vs_ref_cuda"] = o4["ref_cuda_compute_ms"] / o4["compute_ms"]
            del rm, rv
        except Exception as e:  # noqa: BLE001
            out["cfg4_multilabel_auroc_exact"]["ref_cuda_error"] = repr(e)[:200]
        torch.cuda.empty_cache()

    # config 1: MetricCollection([MulticlassAccuracy, MulticlassConfusionMatrix]), 100 000 fp32 probs (latency bound)
    for C in (5, 10):
        preds = torch.randn(100_000, C, generator=g, device=device).softmax(-1)
        tgt = torch.randint(C, (100_000,), generator=g, device=device)
        coll = tm.MetricCollection(
            [MulticlassAccuracy(C, average="micro", validate_args=False), MulticlassConfusionMatrix(C, validate_args=False)]
        ).to(device)
        record(f"cfg1_C{C}_collection", lambda coll=coll, preds=preds, tgt=tgt: coll.update(preds, tgt), 100_000,
               2 * 100_000 * (4 * C + 8), 2 * 100_000 * 4 * C, 3,
               ref=(lambda R, v, C=C: R.MetricCollection([R.classification.MulticlassAccuracy(C, average="micro", validate_args=v),
                                                          R.classification.MulticlassConfusionMatrix(C, validate_args=v)]),
                    (preds, tgt), 10))
        # the same update captured once in a CUDA graph and replayed: what is left when Python + launch overhead go
        try:
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                coll.update(preds, tgt)
            torch.cuda.current_stream().wait_stream(side)
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                coll.update(preds, tgt)
            record(f"cfg1_C{C}_collection_cuda_graph", graph.replay, 100_000, 100_000 * (4 * C + 8), 100_000 * 4 * C, 3)
            del graph
        except Exception as e:  # noqa: BLE001
            out[f"cfg1_C{C}_collection_cuda_graph"] = {"error": repr(e)}
    return out


# ------------------------------------------------------------------------------------------ N > 1: sync parity + strong scaling
def _rank_auroc_f64(scores, y):
    """AUROC of one label as the float64 rank statistic (ties share their average rank) — independent of K3."""
    s = scores.double()
    _, inv, cnt = torch.unique(s, return_inverse=True, return_counts=True)
    rank = (torch.cumsum(cnt, 0) - cnt + (cnt + 1) / 2.0)[inv]
    n_pos = int((y == 1).sum())
    u = rank[y == 1].sum() - n_pos * (n_pos + 1) / 2
    return float(u / (n_pos * (s.numel() - n_pos)))


def sync_parity_prologue(rank: int, world: int, device) -> dict:
    """Untimed proof that the packed state sync (K4) returns what the reference's `_sync_dist` would
    (reference metric.py:423-453, utilities/distributed.py:97-147, tests/unittests/bases/test_ddp.py:33-76).

    Every rank updates from its own seeded shard (uneven row counts), then EVERY rank regenerates all shards and
    checks, bit for bit: (sum path) `MulticlassConfusionMatrix(1000).compute()` and the four `MulticlassStatScores`
    states against `argmax` + `bincount` of all shards; (cat path) `MultilabelAUROC(thresholds=None)` — inside
    `sync_context()` the gathered `preds` / `target` against the rank-ordered concatenation, one rank without any
    rows taking part, and the AUROC against a float64 rank statistic (1e-6).  `compute()` must leave the local
    state as it was.  Raises on any difference (the run then exits non-zero)."""
    import torch.distributed as dist

    from torchmetrics_forked_b200.classification import (
        MulticlassConfusionMatrix,
        MulticlassStatScores,
        MultilabelAUROC,
    )

    C = C2_CLASSES
    rows_of = lambda r: (1 << 14) + 640 * r  # noqa: E731  (uneven shards)
    shard = lambda r: make_cfg2(rows_of(r), device, 9000 + r)  # noqa: E731
    checks = {}

    # ---- sum path: one packed all-reduce per metric
    preds, target = shard(rank)
    cm = MulticlassConfusionMatrix(C, validate_args=False).to(device)
    ss = MulticlassStatScores(C, average=None, validate_args=False).to(device)
    cm.update(preds, target)
    ss.update(preds, target)
    local = cm.confmat.clone()
    got_cm = cm.compute()
    got_ss = ss.compute()  # [C, 5]: tp, fp, tn, fn, support
    want = torch.zeros(C, C, dtype=torch.int64, device=device)
    for r in range(world):
        p, t = shard(r)
        want += torch.bincount(t * C + p.argmax(1), minlength=C * C).reshape(C, C)
    tp = want.diag()
    fp, fn = want.sum(0) - tp, want.sum(1) - tp
    tn = want.sum() - (tp + fp + fn)
    checks["sum_confmat_bit_exact"] = bool(torch.equal(got_cm, want))
    checks["sum_stat_scores_bit_exact"] = bool(torch.equal(got_ss, torch.stack([tp, fp, tn, fn, tp + fn], dim=1)))
    checks["sum_unsync_restores_local"] = bool(torch.equal(cm.confmat, local))
    del preds, target, want

    # ---- cat path: uneven shards, two updates per rank, the last rank of a second metric has no rows at all
    L = 6

    def ml_shard(r):
        g = torch.Generator(device=device).manual_seed(7000 + r)
        n1, n2 = 3000 + 517 * r, 1000 + 3 * r
        return [(torch.randn(n, L, generator=g, device=device), torch.randint(2, (n, L), generator=g, device=device))
                for n in (n1, n2)]

    au = MultilabelAUROC(L, thresholds=None, average="none", validate_args=False).to(device)
    part = MultilabelAUROC(L, thresholds=None, average="none", validate_args=False).to(device)
    for p, t in ml_shard(rank):
        au.update(p, t)
        if rank != world - 1:
            part.update(p, t)
    every = [ml_shard(r) for r in range(world)]
    all_p = torch.cat([p.sigmoid() for sh in every for p, _ in sh])
    all_t = torch.cat([t for sh in every for _, t in sh])
    some_p = torch.cat([p.sigmoid() for sh in every[:-1] for p, _ in sh])
    with au.sync_context():
        checks["cat_preds_rank_ordered_bit_exact"] = bool(torch.equal(au.preds, all_p))
        checks["cat_target_rank_ordered"] = bool(torch.equal(au.target.to(torch.int64), all_t))
    checks["cat_unsync_restores_list"] = isinstance(au.preds, list) and len(au.preds) == 2
    with part.sync_context():
        checks["cat_rank_without_rows"] = bool(torch.equal(part.preds, some_p))
    got = au.compute()
    worst = max(abs(float(got[l]) - _rank_auroc_f64(all_p[:, l], all_t[:, l])) for l in range(L))
    checks["cat_auroc_within_1e-6"] = worst < 1e-6
    checks["cat_auroc_max_abs_diff"] = worst

    ok = torch.tensor([int(all(v for k, v in checks.items() if isinstance(v, bool)))], device=device)
    dist.all_reduce(ok, op=dist.ReduceOp.MIN)
    checks["all_ranks"] = bool(ok.item())
    if not checks["all_ranks"]:
        raise AssertionError(f"rank {rank}: state-sync parity failed: {checks}")
    return checks


def _instrument_sync(metric):
    """CUDA events around every `_sync_dist` of `metric` (the collectives + pack / unpack of one state sync)."""
    spans, orig = [], metric._sync_dist

    def timed(*a, **k):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        orig(*a, **k)
        e1.record()
        spans.append((e0, e1))

    metric.__dict__["_sync_dist"] = timed
    return spans


def _epoch_ms(metric, batches, world, device, synced: bool, reps: int = 3):
    """ms of one epoch = update over `batches` + one compute(), max over ranks; sync ms inside it (0 if local)."""
    spans = _instrument_sync(metric) if synced else []
    best, best_sync, value = None, 0.0, None
    for i in range(reps + 1):  # first epoch = warm-up (allocator, communicator)
        metric.reset()
        del spans[:]
        barrier(world)
        e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        e0.record()
        for b in batches:
            metric.update(*b)
        e1.record()
        value = metric.compute()
        e2.record()
        barrier(world)
        total = max_over_ranks(e0.elapsed_time(e2), world, device)
        upd = max_over_ranks(e0.elapsed_time(e1), world, device)
        sync = max_over_ranks(sum(a.elapsed_time(b) for a, b in spans), world, device) if spans else 0.0
        if i > 0 and (best is None or total < best[0]):
            best, best_sync = (total, upd), sync
    return {"epoch_ms": best[0], "update_ms_total": best[1], "compute_ms": best[0] - best[1], "sync_ms": best_sync}, value


def strong_block(rank: int, world: int, device, steps: int) -> dict:
    """Strong scaling of BASELINE.json's configs 2, 3, 4: the named global batch sharded by sample over the N ranks,
    one `compute()` (state sync included) per epoch.  Beside each, the same epoch on ONE GPU (every rank runs the
    whole problem locally with `sync_on_compute=False`; max over ranks) — `efficiency` = t_1 / (N * t_N)."""
    from torchmetrics_forked_b200.classification import (
        BinaryPrecisionRecallCurve,
        MulticlassConfusionMatrix,
        MultilabelAUROC,
    )

    out = {}
    k = max(3, min(steps, 10))

    def both(name, make_metric, full_batches, shard_batches, units):
        one, v1 = _epoch_ms(make_metric(False).to(device), full_batches, world, device, synced=False)
        many, vn = _epoch_ms(make_metric(True).to(device), shard_batches, world, device, synced=True)
        out[name] = {"n_gpus": world, "units_per_epoch_global": units,
                     "one_gpu": one, "sharded": many, "efficiency": one["epoch_ms"] / (world * many["epoch_ms"]),
                     "speedup": one["epoch_ms"] / many["epoch_ms"],
                     "units_per_s_sharded": units / many["epoch_ms"] * 1e3}
        return v1, vn

    # config 2: [2^22, 1000] bf16 per update over all ranks -> 2^22 / N rows per rank; 8 MB packed all-reduce
    rows = C2_ROWS
    preds, target = make_cfg2(rows, device, 1234)  # every rank holds the same global batch; rank r takes slice r
    lo, hi = rank * (rows // world), (rank + 1) * (rows // world)
    full = [(preds, target)] * k
    mine = [(preds[lo:hi], target[lo:hi])] * k
    v1, vn = both("cfg2_confmat_2^22_rows_per_update",
                  lambda sync: MulticlassConfusionMatrix(C2_CLASSES, validate_args=False, sync_on_compute=sync),
                  full, mine, rows * k)
    out["cfg2_confmat_2^22_rows_per_update"]["sharded_equals_one_gpu"] = bool(torch.equal(v1, vn))
    del preds, target, full, mine, v1, vn
    torch.cuda.empty_cache()

    # config 3: 2^28 fp32 logits per update -> 2^28 / N per rank; 6.4 kB all-reduce
    n3 = 1 << 28
    g = torch.Generator(device=device).manual_seed(4321)
    logits = torch.randn(n3, generator=g, device=device) * 3
    tgt = torch.randint(2, (n3,), generator=g, device=device)
    lo, hi = rank * (n3 // world), (rank + 1) * (n3 // world)
    v1, vn = both("cfg3_binary_curve_T200_2^28_per_update",
                  lambda sync: BinaryPrecisionRecallCurve(thresholds=200, validate_args=False, sync_on_compute=sync),
                  [(logits, tgt)] * k, [(logits[lo:hi], tgt[lo:hi])] * k, n3 * k)
    out["cfg3_binary_curve_T200_2^28_per_update"]["sharded_equals_one_gpu"] = bool(
        all(torch.equal(a, b) for a, b in zip(v1, vn)))
    del logits, tgt, v1, vn
    torch.cuda.empty_cache()

    # config 4: MultilabelAUROC exact, 2^24 samples x 100 labels in 8 updates -> 2^21 / N rows per rank and update
    n4, L4, upd = 1 << 21, 100, 8
    batches = [(torch.randn(n4, L4, generator=g, device=device), torch.randint(2, (n4, L4), generator=g, device=device))
               for _ in range(upd)]
    lo, hi = rank * (n4 // world), (rank + 1) * (n4 // world)
    v1, vn = both("cfg4_multilabel_auroc_exact_2^24x100",
                  lambda sync: MultilabelAUROC(L4, thresholds=None, validate_args=False, sync_on_compute=sync),
                  batches, [(p[lo:hi], t[lo:hi]) for p, t in batches], n4 * upd * L4)
    out["cfg4_multilabel_auroc_exact_2^24x100"]["auroc_abs_diff_sharded_vs_one_gpu"] = abs(float(v1) - float(vn))
    from torchmetrics_forked_b200 import _sharded

    out["cfg4_multilabel_auroc_exact_2^24x100"]["compute_path"] = dict(_sharded.LAST, rows=None)
    del batches
    torch.cuda.empty_cache()
    return out


# ------------------------------------------------------------------------------------------ our arm
def run_ours(args):
    rank, world, local = dist_setup(args.gpus)
    device = torch.device("cuda", local)
    from torchmetrics_forked_b200.classification import MulticlassConfusionMatrix

    rows = args.rows
    parity = sync_parity_prologue(rank, world, device) if world > 1 else None
    preds, target = make_cfg2(rows, device, 1234 + rank)
    metric = MulticlassConfusionMatrix(C2_CLASSES, validate_args=False).to(device)
    peak, peak_src = measured_peak_gbs()

    def step():
        metric.update(preds, target)

    result = {}

    def tail():  # one compute() per epoch of updates: packed all-reduce of the 8 MB state when world > 1
        result["confmat"] = metric.compute()

    # untimed: one update + compute() so that the NCCL communicator and the allocator are warm before the W warm-up
    # steps (compute() does not change the local state; the extra update is accounted for in the check below)
    step()
    tail()
    with ClockSampler(local) as clocks:
        total_ms, kernel_ms = timed_loop(step, args.steps, args.warmup, world, device, tail=tail)
    clk = clocks.summary()
    samples = rows * world * args.steps
    value = samples / total_ms * 1e3
    algo_bytes = rows * (2 * C2_CLASSES + 8)
    achieved = algo_bytes / kernel_ms / 1e6

    # sanity: every sample of every update landed in the state exactly once
    seen = int(result["confmat"].sum().item())
    expect = rows * world * (args.steps + args.warmup + 1)
    assert seen == expect, f"confusion matrix holds {seen} samples, expected {expect}"

    # the unmodified reference on the same GPU and tensors (rank 0 of a 1-GPU run; torch's stock argmax + bincount)
    ref_cuda = {}
    if rank == 0 and world == 1 and not args.no_extras:
        ref_cuda = ref_cuda_update_ms(
            lambda R, v: R.classification.MulticlassConfusionMatrix(C2_CLASSES, validate_args=v), (preds, target))

    # end to end through the public API from pinned host memory
    metric.reset()
    e2e_steps = min(args.steps, 10)
    h_preds = torch.empty(rows, C2_CLASSES, dtype=torch.bfloat16, pin_memory=True)
    h_target = torch.empty(rows, dtype=torch.int64, pin_memory=True)
    h_preds.copy_(preds)
    h_target.copy_(target)
    del preds, target
    torch.cuda.empty_cache()

    def e2e_step():
        p = h_preds.to(device, non_blocking=True)
        t = h_target.to(device, non_blocking=True)
        metric.update(p, t)
        cm = metric.confmat
        return float((cm.diagonal().sum() / cm.sum()).item())  # running accuracy read back every step

    e2e_ms, _ = timed_loop(e2e_step, e2e_steps, 1, world, device)
    e2e_value = rows * world * e2e_steps / e2e_ms * 1e3

    line = {
        "metric": METRIC, "value": value, "unit": "samples/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": "cfg2: MulticlassConfusionMatrix(1000).update(bf16 logits, int64 targets), validate_args=False;"
                   " one compute() (state sync) per timed region", "rows_per_gpu_per_step": rows, "classes": C2_CLASSES,
                   "l2": "inputs (8.4 GB/update) larger than L2, no flush", "parallelism": f"samples sharded over {world} GPU(s)"},
        "preds_gbs": rows * world * args.steps * 2 * C2_CLASSES / total_ms / 1e6,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": recorded_traffic("k1_rows_bf16"), "peak_source": peak_src,
                     "kernel": "k1 rows (bf16, C=1000)", "kernel_ms": kernel_ms, "algorithmic_bytes": algo_bytes,
                     "frac_of_nominal_8000": achieved / 8000.0},
        "e2e": {"value": e2e_value, "unit": "samples/s", "h2d_bytes_per_step": rows * (2 * C2_CLASSES + 8),
                "d2h_bytes_per_step": 4, "steps": e2e_steps, "ms_per_step": e2e_ms / e2e_steps},
        "gpu_launches": args.steps,
        "clocks": clk,
    }
    del h_preds, h_target
    torch.cuda.empty_cache()
    if parity is not None:
        line["sync_parity"] = parity["all_ranks"]
        line["sync_parity_checks"] = parity
    if world > 1 and not args.no_extras:
        try:
            strong = strong_block(rank, world, device, args.steps)
        except Exception as e:  # noqa: BLE001
            strong = {"error": repr(e)}
        line["extras"] = {"strong": strong}
    if rank == 0 and world == 1:
        rate, ms, done, kind, cores = cpu_confmat_rate(args.cpu_rows, 1000, 1, budget_s=args.cpu_budget)
        what = "the unmodified reference, oracle/_ref" if kind == "reference" else "op-for-op torch-CPU port of the reference"
        line["cpu_baseline"] = {
            "value": rate, "unit": "samples/s", "cores": cores, "kind": kind,
            "sample": f"{done} updates of bf16 logits [{args.cpu_rows},1000] ({what}, {ms:.1f} ms each)",
        }
        if not args.no_extras:
            try:
                line["extras"] = extras_block(device, max(3, min(args.steps, 10)), max(args.warmup, 3), peak)
            except Exception as e:  # noqa: BLE001
                line["extras"] = {"error": repr(e)}
            line["extras"]["cfg2_confmat_2^22_rows"] = dict(
                ms_per_update=kernel_ms, roofline_frac=achieved / peak, **ref_cuda,
                **({"speedup_vs_ref_cuda": ref_cuda["ref_cuda_ms"] / kernel_ms} if ref_cuda.get("ref_cuda_ms") else {}))
    if rank == 0:
        emit(line)
    if world > 1:
        import torch.distributed as dist

        dist.destroy_process_group()


_JSON_FD = None


def emit(line) -> None:
    """The one JSON line, on the process's original stdout."""
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_JSON_FD, data)


def main():
    # stdout carries the JSON line and nothing else: native libraries write there too (NCCL prints its version
    # line to stdout when NCCL_DEBUG is set, as it is on the GPU boxes), so file descriptor 1 is pointed at stderr
    # for the run and the line goes to a duplicate of the original
    global _JSON_FD
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--rows", type=int, default=C2_ROWS, help="rows per GPU per update (default: config 2)")
    ap.add_argument("--cpu-rows", type=int, default=1 << 16, help="rows of the bounded CPU sample")
    ap.add_argument("--cpu-budget", type=float, default=15.0)
    ap.add_argument("--no-extras", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
