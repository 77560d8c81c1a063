"""bench.py -- throughput of the HEAT MF_CCL training hot path on B200 (see DESIGN.md "Measurement").

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME]

A "step" is one training epoch (Engine::train_one_epoch) over the workload's synthetic click data:
LR schedule, every sample through the fused sampler+forward+backward+SGD kernel, zero_grad.  At N=1
the workload is BASELINE.json configs[2], the configuration the metric is quoted on: Amazon-Books
shape (52 643 users, 91 599 items, 2 380 730 interactions), dim 64, 16 negatives, max_his 100.  At
N>1 every rank owns one such user shard (users split by id range, weak scaling), the item table and
W0 are replicated and their deltas all-reduced over NCCL every `sync_interval` samples.

One JSON line on stdout (rank 0).  `value` = samples/s with everything resident in HBM (device time,
CUDA events on the library's stream, max over ranks); `e2e` = the same epoch through the C ABI with
HOST buffers (dataset + tables uploaded from pinned memory, trained, tables copied back) inside the
timed region; `roofline` = the fused kernel's algorithmic bytes over its CUDA-event time against
the measured HBM peak (and against the measured L2 row-gather peak, the bound that applies while
the tables fit L2); `cpu_baseline` = the unmodified reference C++ (oracle/_ref) on the host cores.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

D, K, MAX_HIS = 64, 16, 100
METRIC, UNIT = "train_samples_per_sec", "samples/s"


def algorithmic_bytes(clicks, masks, d=D, k=K):
    """B(D,K,H) = 4D(4(2+K)+H) + 12 + 4H per sample, H = masks[user]."""
    h = masks[clicks[:, 0].astype(np.int64), 0].astype(np.float64)
    return float(np.sum(4 * d * (4 * (2 + k) + h) + 12 + 4 * h))


def ncu_traffic(kernel_file):
    """dram__bytes_read.sum + dram__bytes_write.sum of one launch from the committed ncu --set full capture."""
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", kernel_file)
    try:
        import csv
        with open(path) as f:
            rows = list(csv.reader(f))
        head, unit, val = rows[0], rows[1], rows[2]
        scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        tot = 0.0
        for k in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
            i = head.index(k)
            tot += float(val[i].replace(",", "")) * scale[unit[i]]
        return tot
    except Exception:
        return None


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            j = json.load(f)
        return float(j["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.rows, self.proc, self.index = [], None, index

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def __exit__(self, *a):
        if self.proc:
            time.sleep(0.15)
            self.proc.terminate()
            self.t.join(timeout=2)

    def summary(self):
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
            except (ValueError, IndexError):
                continue
            for n, v in zip(names, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons),
                "samples": len(sm)}


def make_workload(name, rank=0):
    """Seeded synthetic click data of the named shape; rank r gets its own user shard (seed 2022+r)."""
    from heat_b200 import synth
    d = synth.generate_named(name, max_his=MAX_HIS, seed=2022 + rank)
    tables = synth.init_tables(d.num_users, d.num_items, D, seed=2022)   # item table / W0 replicated
    if rank:
        tables = (synth.init_tables(d.num_users, d.num_items, D, seed=2022 + rank)[0], tables[1], tables[2])
    return d, tables


def cpu_reference_run(d, tables, n_samples, epochs, threads=0):
    """The reference's own C++ per-sample code (oracle/_ref, compiled unmodified from
    /root/reference) -- or the C restatement when that library is absent -- over the first
    n_samples samples, upstream OpenMP loop, all host threads.  Returns (seconds per epoch list,
    kind, threads, mean loss)."""
    import oracle
    clicks = np.ascontiguousarray(d.clicks[:n_samples])
    uw, iw, w0 = (t.copy() for t in tables)
    if oracle.RefOracle.available():
        o = oracle.RefOracle(D, K, clicks, d.his_items, d.masks, uw, iw, w0)
        kind, nthr = "reference", o.max_threads()
        run = lambda: o.train_epoch(nthreads=threads, sampler=0)          # its own mt19937 sampler
    else:
        o = oracle.COracle(D, K, clicks, d.his_items, d.masks, uw, iw, w0)
        kind, nthr = "port", os.cpu_count() or 1
        run = lambda: o.train_epoch(nthreads=threads)
    if threads:
        nthr = threads
    secs, loss = [], 0.0
    for _ in range(epochs):
        t0 = time.perf_counter()
        loss = run()
        secs.append(time.perf_counter() - t0)
    o.close()
    return secs, kind, nthr, loss


def cpu_baseline(d, tables, budget_s=12.0):
    """Bounded sample: calibrate on 100k samples, then one pass sized for ~budget_s of CPU work."""
    n0 = min(100_000, d.train_size)
    s0, kind, nthr, _ = cpu_reference_run(d, tables, n0, 1)
    rate = n0 / s0[0]
    n = int(min(d.train_size, max(n0, rate * budget_s)))
    secs, kind, nthr, loss = cpu_reference_run(d, tables, n, 1)
    return {"value": n / secs[0], "unit": UNIT, "cores": nthr, "kind": kind,
            "sample": f"first {n} of {d.train_size} samples, 1 pass, {secs[0]:.2f} s, OpenMP schedule(dynamic,512)",
            "loss": loss}


def run_reference_arm(args, rank, world):
    """--impl reference: the reference's CPU implementation of the path on this box's host cores."""
    if rank != 0:
        return
    d, tables = make_workload(args.workload)
    n0 = min(100_000, d.train_size)
    s0, kind, nthr, _ = cpu_reference_run(d, tables, n0, 1)
    n = int(min(d.train_size, max(n0, (n0 / s0[0]) * args.ref_step_seconds)))
    secs, kind, nthr, loss = cpu_reference_run(d, tables, n, args.warmup + args.steps)
    timed = secs[args.warmup:]
    value = n * len(timed) / sum(timed)
    sample = f"each step = first {n} of {d.train_size} samples, upstream OpenMP loop, {nthr} threads"
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * sum(timed) / len(timed),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args, d, world),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": nthr, "kind": kind, "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0, "loss": loss}
    print(json.dumps(line), file=_JSON_OUT, flush=True)


def workload_config(args, d, world):
    return {"workload": f"{args.workload}-shaped synthetic MF_CCL (BASELINE.json configs[2])",
            "users_per_gpu": d.num_users, "items": d.num_items, "samples_per_gpu": d.train_size,
            "emb_dim": D, "num_negs": K, "max_his": MAX_HIS, "optimizer": "sgd+clip (persistent grad rows)",
            "parallelism": f"user-sharded x{world}" if world > 1 else "single GPU",
            "l2": "flushed between timed steps (256 MiB write); within a step the 74 MB of tables stay L2-resident by design"}


def run_gpu_arm(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    from heat_b200 import abi

    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    d, tables = make_workload(args.workload, rank)
    N = d.train_size

    # pinned host copies: the buffers the reference frontend would own (uint64 ids, fp32 tables)
    def pin(a):
        p = abi.pinned_empty(a.shape, a.dtype)
        p[...] = a
        return p
    clicks, his, masks = pin(d.clicks), pin(d.his_items), pin(d.masks)
    uw, iw, w0 = (pin(t) for t in tables)
    init = [t.copy() for t in tables]

    ctx = abi.Context(D, K, d.num_users, d.num_items, clip_val=1.0, l_r=1e-2, milestones=(10,), seed=2022,
                      sample_offset=rank * N, device=local_rank, update_mode=args.update_mode,
                      warps_per_block=args.warps, max_blocks=args.max_blocks, chunk=args.chunk, engine=args.engine)
    ctx.set_dataset(clicks, his, masks)
    ctx.bind_tables(uw, iw, w0)
    if world > 1:
        uid = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            uid.copy_(torch.from_numpy(abi.Context.comm_unique_id()))
        dist.broadcast(uid, 0)
        ctx.comm_init(rank, world, uid.cpu().numpy())
        ctx.set_sync_interval(args.sync_interval)
    exchange = {0: None, 1: "NCCL all-reduce of the item/W0 delta", 2: "fused peer-memory exchange kernel (CUDA IPC over NVLink)"}[ctx.comm_mode()]

    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    def flush_l2():
        flush.add_(1)
        torch.cuda.synchronize()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def reduce_max(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- resident arm: tables and dataset stay in HBM; a step = heat_train_epoch
    losses = []
    for _ in range(args.warmup):
        flush_l2(); losses.append(ctx.train_epoch())
    launches0 = ctx.launch_count()
    barrier()
    epoch_ms, kernel_ms = [], []
    wall0 = time.perf_counter()
    with ClockSampler(local_rank) as clk:
        for _ in range(args.steps):
            flush_l2()
            losses.append(ctx.train_epoch())
            epoch_ms.append(ctx.last_epoch_ms()); kernel_ms.append(ctx.last_kernel_ms())
        barrier()
    wall = time.perf_counter() - wall0
    launches = ctx.launch_count() - launches0
    total_ms = reduce_max(float(np.sum(epoch_ms)))
    value = world * N * args.steps / (total_ms * 1e-3)

    # ---- end-to-end arm: host buffers in, trained host tables out, every step
    h2d = clicks.nbytes + his.nbytes + masks.nbytes + uw.nbytes + iw.nbytes + w0.nbytes
    d2h = uw.nbytes + iw.nbytes + w0.nbytes + 8
    e2e_steps = max(3, min(args.steps, 10))

    def e2e_step():
        ctx.set_dataset(clicks, his, masks)      # uint64 host arrays -> device int32
        ctx.upload_tables()                      # host tables -> device
        loss = ctx.train_epoch()
        ctx.sync_to_host()                       # trained tables -> the caller's host arrays
        return loss
    for _ in range(2):
        e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        flush_l2()
        e2e_step()
    barrier()
    e2e_s = reduce_max(time.perf_counter() - t0)
    e2e_value = world * N * e2e_steps / e2e_s

    # ---- evaluation leg ("eval ms" of BASELINE.json's metric): U.I^T on tcgen05 (3xTF32) fused with the masked top-50
    #      and Recall/NDCG/HitRate, on the tables the epochs above trained; every rank evaluates its own user shard.
    ev = None
    if not args.no_eval:
        ev_ms, ev_wall = [], []
        res = None
        for i in range(2 + 5):
            flush_l2()
            t0 = time.perf_counter()
            res, _ = ctx.eval_metrics(d.train_indptr, d.train_items, d.test_indptr, d.test_items)
            if i >= 2:
                ev_wall.append((time.perf_counter() - t0) * 1e3); ev_ms.append(ctx.last_eval_ms())
        ev_ms_max = reduce_max(float(np.mean(ev_ms)))
        flops = 2.0 * d.num_users * d.num_items * D
        tf32_peak = 1125.0                                   # TFLOP/s dense, nominal (half the bf16 rate; no measured TF32 peak on file)
        ev = {"ms": ev_ms_max, "wall_ms_with_csr_upload": float(np.mean(ev_wall)), "users_per_gpu": int(d.num_users),
              "items": int(d.num_items), "k": 50, "metrics": res,
              "roofline": {"bound": "tensor", "achieved": 3 * flops / (ev_ms_max * 1e-3) / 1e12, "peak": tf32_peak, "unit": "TFLOP/s",
                           "frac": 3 * flops / (ev_ms_max * 1e-3) / 1e12 / tf32_peak,
                           "note": "executed TF32 flops (3 passes of 2*U*I*D: A_lo.B_hi + A_hi.B_lo + A_hi.B_hi); fp32-equivalent "
                                   "work is a third of it; includes the TF32 split of both tables and the metrics kernel",
                           "fp32_equivalent_tflops": flops / (ev_ms_max * 1e-3) / 1e12}}
    ctx.close()

    # ---- throughput mode, reported beside the headline: the warp-per-sample kernel on a full grid (thousands of
    #      samples in flight).  Faster, but its Hogwild dynamics leave the reference's Recall@20,
    #      so it is not the configuration `value` is quoted on.
    thr = None
    if args.engine == 0 and not args.no_throughput_mode:
        for t, t0_ in zip((uw, iw, w0), init):
            t[...] = t0_
        ctx2 = abi.Context(D, K, d.num_users, d.num_items, clip_val=1.0, l_r=1e-2, milestones=(10,), seed=2022,
                           sample_offset=rank * N, device=local_rank, engine=1)
        ctx2.set_dataset(clicks, his, masks)
        ctx2.bind_tables(uw, iw, w0)
        if world > 1:
            uid = torch.zeros(128, dtype=torch.uint8, device="cuda")
            if rank == 0:
                uid.copy_(torch.from_numpy(abi.Context.comm_unique_id()))
            dist.broadcast(uid, 0)
            ctx2.comm_init(rank, world, uid.cpu().numpy())
            ctx2.set_sync_interval(args.sync_interval)
        for _ in range(args.warmup):
            flush_l2(); ctx2.train_epoch()
        barrier()
        ms2, k2 = [], []
        for _ in range(args.steps):
            flush_l2(); ctx2.train_epoch()
            ms2.append(ctx2.last_epoch_ms()); k2.append(ctx2.last_kernel_ms())
        barrier()
        tot2 = reduce_max(float(np.sum(ms2)))
        thr = {"value": world * N * args.steps / (tot2 * 1e-3), "unit": UNIT, "ms_per_step": tot2 / args.steps,
               "kernel": "heat::train_kernel (warp per sample, full grid)", "kernel_ms": float(np.mean(k2)),
               "recall_parity": "outside the 0.005 tolerance (profiles/r01_recall_parity.md)"}
        ctx2.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank != 0:
        return

    # ---- roofline of the fused training kernel (dominant: one launch per epoch at N=1)
    alg = algorithmic_bytes(d.clicks, d.masks)
    k_ms = float(np.mean(kernel_ms))
    peak, peak_src = measured_peaks()
    achieved = alg / (k_ms * 1e-3) / 1e9
    table_bytes = 2 * (d.num_users + d.num_items) * D * 4
    l2_peak = abi.probe_l2_bandwidth(local_rank, table_bytes, 200)
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": args.ncu_traffic if args.ncu_traffic is not None else ncu_traffic(
                    "r01_train_cta_kernel_ncu_raw.csv" if args.engine != 1 else "r01_train_kernel_ncu_raw.csv"),
                "traffic_source": "profiles/ ncu --set full capture of the same bench command (bytes per launch)", "peak_source": peak_src,
                "kernel": "heat::train_cta_kernel" if args.engine != 1 else "heat::train_kernel",
                "kernel_ms": k_ms, "algorithmic_bytes_per_launch": alg,
                "note": "weights+grads (%.0f MB) are L2-resident: the bound that applies is the L2 row "
                        "gather/scatter rate, measured by heat_probe_l2_bandwidth on the same footprint" % (table_bytes / 1e6),
                "l2_peak": l2_peak, "l2_frac": achieved / l2_peak}
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": total_ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args, d, world), "epoch_s": total_ms / args.steps / 1e3,
            "clocks": clk.summary(), "gpu_launches": int(launches),
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                    "steps": e2e_steps, "ms_per_step": 1e3 * e2e_s / e2e_steps,
                    "what": "heat_set_dataset + heat_upload_tables + heat_train_epoch + heat_sync_to_host from pinned host buffers"},
            "roofline": roofline, "loss_first_last": [losses[0], losses[-1]], "wall_s_timed_region": wall,
            "update_mode": args.update_mode, "sync_interval": args.sync_interval if world > 1 else None, "exchange": exchange,
            "engine": {0: "auto: CTA-per-sample kernel, parity-grade concurrency", 1: "warp-per-sample kernel, full grid",
                       2: "CTA-per-sample kernel"}[args.engine]}
    if ev is not None:
        line["eval"] = ev
    if thr is not None:
        thr["l2_frac"] = alg / (thr["kernel_ms"] * 1e-3) / 1e9 / l2_peak
        line["throughput_mode"] = thr
    if not args.no_cpu_baseline:
        d0, t0_ = (d, init)
        line["cpu_baseline"] = cpu_baseline(d0, t0_, args.cpu_seconds)
    print(json.dumps(line), file=_JSON_OUT, flush=True)


_JSON_OUT = sys.stdout


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="amazon-book")
    ap.add_argument("--update-mode", type=int, default=0)
    ap.add_argument("--warps", type=int, default=0)
    ap.add_argument("--max-blocks", type=int, default=0)
    ap.add_argument("--chunk", type=int, default=0)
    ap.add_argument("--engine", type=int, default=0, help="0 auto (CTA-per-sample kernel at parity-grade concurrency), 1 warp-per-sample, 2 CTA-per-sample")
    ap.add_argument("--no-throughput-mode", action="store_true", help="skip the extra warp-per-sample full-grid measurement")
    ap.add_argument("--sync-interval", type=int, default=0, help="samples per GPU between item-delta all-reduces (0 = per epoch)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-eval", action="store_true", help="skip the evaluation leg")
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--ref-step-seconds", type=float, default=4.0)
    ap.add_argument("--ncu-traffic", type=float, default=None, help="dram bytes per launch from profiles/ (ncu --set full)")
    args = ap.parse_args()
    # stdout carries exactly ONE JSON line: the plugin prints the reference's progress lines with std::cout
    # (cf_c.modules.test.test_out), so fd 1 points at stderr while the arms run and the line goes to the saved fd
    global _JSON_OUT
    sys.stdout.flush()
    _JSON_OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    if args.impl == "reference":
        run_reference_arm(args, rank, world)
        return
    if args.gpus > 1 and world == 1:
        raise SystemExit("launch N>1 with: python -m torch.distributed.run --nnodes=1 --nproc-per-node N "
                         "--master-addr 127.0.0.1 --master-port P bench.py --gpus N ...")
    run_gpu_arm(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
