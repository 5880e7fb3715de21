#!/usr/bin/env python3
"""Benchmark of the poll path: samples/s parsed -> line protocol (BASELINE.json `metric`).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl kgm|reference] [--config config2]

A *step* is one pass of the whole hot path (tokenize -> join -> scan -> encode; by default
all of it inside the one fused kernel k_poll, `--pipeline staged` runs the four staged
kernels) over one batch of synthetic replayed nvidia-smi text.  At N=1 the workload is BASELINE.json
`configs[1]`: 64 nodes x 8 GPUs x 60 ticks = 3840 documents, 30720 samples, 2048 pods.
At N>1 (torchrun, one rank per GPU) every rank holds its own 64-node shards ("weak"
scaling; hosts are the partition unit) and the step additionally gathers the
encoded shards and their sizes on one GPU: k_poll writes them straight into slots on the
step's gathering rank over NVLink peer memory, NCCL carries a 4-byte all-reduce per step
(fallback without CUDA IPC: one NCCL all-gather of fixed-size slots).

Printed JSON (one line, rank 0):
  value   whole-job samples/s with the batches already resident in HBM: CUDA events around
          K steps (barrier + synchronize on both sides, every step finished inside), max
          over ranks; every rank rotates over 4 distinct resident batches (240 MB > L2)
          instead of flushing, each in its own context, queued asynchronously
  e2e     the same through kgm_encode_batch with pinned HOST buffers: H2D of the texts and
          D2H of lines/offsets/statuses inside the timed region (wall clock around the call)
  roofline  the dominant kernel (k_poll: the whole fused pass; k_encode when staged):
          algorithmic bytes per launch / its CUDA-event time with the GPU to itself
          (roofline.launch_ms_source says where that time was taken)
  cpu_baseline  the reference (gpu.py main(), else the port) on one host core over a bounded
          prefix of the workload (N=1, rank 0)

`--impl reference` times the reference's CPU implementation of the path on all host cores, every
document of the configs[1] workload per step: the unmodified gpu.py main() under
oracle/ref_harness.py (oracle/make_ref.py puts a git-ignored copy of the script into
oracle/_ref/, which travels to the GPU box); the port oracle/poll_oracle.py only if that copy
is missing.
"""
from __future__ import annotations

import argparse
import json

import numpy as np
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "k8s-gpu-monitor_b200")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

METRIC = "samples/sec parsed->line protocol"


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, device):
        self.device, self.rows, self.proc = device, [], None

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.device), "--query-gpu=" + self.Q,
                 "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def __exit__(self, *exc):
        if self.proc:
            time.sleep(0.15)
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()

    def summary(self):
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
            except Exception:
                continue
            for name, v in zip(names, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "reasons": sorted(reasons),
                "samples": len(sm)}


def _replay_of_nodes(replay, seed, node_ids, ticks, tick0=0, **node_kw):
    """A replay of the given hosts (global host numbers), all ticks of a host together."""
    snaps = [replay.make_node(seed, n, **node_kw) for n in node_ids]
    doc_node, doc_tick, gt, at = [], [], [], []
    for i, (n, snap) in enumerate(zip(node_ids, snaps)):
        for t in range(ticks):
            a, b2 = replay.make_texts(seed, n, snap, tick0 + t)
            doc_node.append(i); doc_tick.append(tick0 + t); gt.append(a); at.append(b2)
    return replay.Replay(seed, snaps, np.asarray(doc_node, np.int32), np.asarray(doc_tick, np.int32), gt, at)


def build_workload(config, rank, world=1, variant=0):
    """(replay, fleet entries, shard info) of this rank.
      config1/2  every rank its own fleet (weak scaling; `rank` numbers the fleet)
      config3    ONE fleet of 1000 hosts, 40 of the 600 ticks, hosts partitioned over the ranks by
                 kgm/shard.py assign_hosts (strong scaling)
      config4    12 500 hosts, one window of 4 ticks (`variant` = which window); the ranks' hosts by
                 assign_hosts; the hour is 900 such windows, streamed: encoded, handed on, discarded
      config5    skewed fleet (Zipf pods per host, short-lived pods, MIG hosts), hosts partitioned by
                 assign_hosts with their text bytes as weights"""
    from kgm import replay, shard
    from kgm.fleet import node_entry
    info = None
    if config in ("config1", "config2"):
        kw = dict(replay.CONFIGS[config])
        rp = replay.make_replay(seed=0, node_base=rank * kw["nodes"], **kw)
    elif config == "config3":
        kw = dict(replay.CONFIGS["config3"])
        nodes, distinct, span = kw.pop("nodes"), 4, 40
        kw.pop("ticks")
        names = ["k8s-node-gpu-%05d" % n for n in range(nodes)]
        owner = shard.assign_hosts(names, [1] * nodes, world)
        mine = [n for n in range(nodes) if owner[n] == rank]
        base = _replay_of_nodes(replay, 0, mine, distinct, **kw)
        # tick-major inside the batch; 4 distinct ticks replicated to 40 so that host-side
        # generation stays in seconds
        order = [i * distinct + (t % distinct) for t in range(span) for i in range(len(mine))]
        rp = replay.Replay(base.seed, base.nodes, base.doc_node[order],
                           np.asarray([t for t in range(span) for _ in mine], np.int32),
                           [base.gpu_texts[d] for d in order], [base.apps_texts[d] for d in order])
        info = {"hosts_total": nodes, "hosts_this_rank": len(mine), "partition": "assign_hosts, equal weights"}
    elif config == "config4":
        nodes, window = 12500, 4
        names = ["k8s-node-gpu-%05d" % n for n in range(nodes)]
        owner = shard.assign_hosts(names, [1] * nodes, 8)          # always the 8-GPU partition: rank r takes part r,
        mine = [n for n in range(nodes) if owner[n] == rank]       # so a rank's work does not change with the world size
        rp = _replay_of_nodes(replay, 0, mine, window, tick0=variant * window, gpus=8, containers=8,
                              procs_per_container=4)
        rp.node_ids, rp.window, rp.tick0 = mine, window, variant * window      # (for the device-side generator)
        info = {"hosts_total": nodes, "hosts_this_rank": len(mine), "window_ticks": window,
                "windows_per_hour": 3600 // window, "partition": "assign_hosts, equal weights (one eighth per rank)"}
    elif config == "config5":
        full = replay.make_skewed_replay(seed=5, hosts=240, epochs=2, ticks_per_epoch=6, max_containers=160)
        hosts = sorted({s.name for s in full.nodes})
        weight = {h: 0 for h in hosts}
        for d in range(full.n_docs):
            weight[full.nodes[int(full.doc_node[d])].name] += len(full.gpu_texts[d]) + len(full.apps_texts[d])
        owner = dict(zip(hosts, shard.assign_hosts(hosts, [weight[h] for h in hosts], world)))
        keep = [d for d in range(full.n_docs) if owner[full.nodes[int(full.doc_node[d])].name] == rank]
        used = sorted({int(full.doc_node[d]) for d in keep})
        remap = {n: i for i, n in enumerate(used)}
        rp = replay.Replay(full.seed, [full.nodes[n] for n in used],
                           np.asarray([remap[int(full.doc_node[d])] for d in keep], np.int32),
                           full.doc_tick[keep], [full.gpu_texts[d] for d in keep], [full.apps_texts[d] for d in keep])
        loads = [sum(w for h, w in weight.items() if owner[h] == r) for r in range(world)]
        info = {"hosts_total": len(hosts), "hosts_this_rank": sum(1 for h in hosts if owner[h] == rank),
                "partition": "assign_hosts, text bytes per host as weights",
                "input_bytes_max_over_mean": max(loads) / (sum(loads) / world),
                "largest_host_share": max(weight.values()) / sum(weight.values())}
    else:
        raise ValueError(config)
    entries = [node_entry(s.name, s.ps_text, s.stats_text, s.top, s.inspect, s.smi_list) for s in rp.nodes]
    return rp, entries, info


# ---------------------------------------------------------------------------------------
# CPU legs (the only places bench.py touches oracle/)
# ---------------------------------------------------------------------------------------
_RP = None


def _cpu_poll():
    """The CPU implementation the two CPU legs time: the reference's own main() (gpu.py:90-256)
    under oracle/ref_harness.py when the script is there (/root/reference in the build container,
    the copy oracle/make_ref.py put into oracle/_ref/ on the GPU box), else the port."""
    from oracle import poll_oracle, ref_harness
    if ref_harness.available():
        return ref_harness.run_reference, "reference", "gpu.py main() unmodified, I/O seams stubbed (oracle/ref_harness.py)"
    return poll_oracle.poll, "port", "oracle/poll_oracle.py (CPython port of gpu.py main(); oracle/_ref/gpu.py not found)"


def _cpu_docs(idx):
    poll = _cpu_poll()[0]
    n = 0
    for d in idx:
        s = _RP.nodes[int(_RP.doc_node[d])]
        lines, err = poll(s.name, _RP.gpu_texts[d], _RP.apps_texts[d], s.ps_text, s.stats_text, s.top, s.inspect,
                          s.smi_list)
        n += len(lines)
    return n


def cpu_baseline_single_core(rp, gpu_result, budget_s=12.0):
    """The reference on ONE host core over a bounded prefix of the same workload (only the
    reference call is timed); doubles as a parity spot check of the GPU result on those
    documents."""
    poll, kind, what = _cpu_poll()
    spent = 0.0
    docs = samples = mismatches = 0
    for d in range(rp.n_docs):
        s = rp.nodes[int(rp.doc_node[d])]
        t0 = time.perf_counter()
        lines, err = poll(s.name, rp.gpu_texts[d], rp.apps_texts[d], s.ps_text, s.stats_text, s.top, s.inspect,
                          s.smi_list)
        spent += time.perf_counter() - t0
        docs += 1
        samples += rp.gpu_texts[d].count("\n") - 1
        if gpu_result is not None and (gpu_result.error(d) != err or gpu_result.lines(d) != lines):
            mismatches += 1
        if spent > budget_s:
            break
    return {"value": samples / spent if spent else 0.0, "unit": "samples/s", "cores": 1, "kind": kind,
            "sample": "first %d of %d documents of the same workload (%.1f s), %s, 1 process"
                      % (docs, rp.n_docs, spent, what),
            "parity_docs_checked": docs if gpu_result is not None else 0,
            "parity_mismatches": mismatches}


def run_reference_arm(args):
    """`--impl reference`: the reference's CPU implementation of the path on all host cores, every
    document of the workload in every step."""
    global _RP
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    import multiprocessing as mp
    rp, _, _ = build_workload(args.config, 0)
    _RP = rp
    _, kind, what = _cpu_poll()
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    # the whole configs[1] batch per step; of the larger configs a bounded sample (10 000 documents)
    sample_docs = min(rp.n_docs, 10000)
    per_doc_samples = rp.n_samples() / rp.n_docs
    chunks = [list(range(i, sample_docs, cores * 4)) for i in range(cores * 4)]
    ctx = mp.get_context("fork")
    times = []
    with ctx.Pool(cores) as pool:
        for step in range(args.warmup + args.steps):
            t0 = time.perf_counter()
            pool.map(_cpu_docs, chunks, chunksize=1)
            dt = time.perf_counter() - t0
            if step >= args.warmup:
                times.append(dt)
    ms = 1e3 * sum(times) / len(times)
    value = sample_docs * per_doc_samples / (ms / 1e3)
    sample = ("all %d documents of the workload per step" % rp.n_docs if sample_docs == rp.n_docs else
              "%d of %d documents per step" % (sample_docs, rp.n_docs))
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "samples/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8",
        "data": "synthetic",
        "config": {"workload": workload_name(args.config), "docs_per_step": sample_docs,
                   "docs_per_gpu": rp.n_docs, "samples_per_gpu": rp.n_samples()},
        "cpu_baseline": {"value": value, "unit": "samples/s", "cores": cores, "kind": kind,
                         "sample": "%s, %s, %d processes" % (sample, what, cores)},
        "e2e": {"value": value, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))
    return 0


def workload_name(config):
    return {"config1": "configs[0]: 1 node, 8 GPUs, 32 processes, 1 tick",
            "config2": "configs[1]: 64-node fleet batch, 512 GPUs, 2048 pods, 60 ticks at 1 Hz "
                       "(3840 documents, 30720 samples) per GPU",
            "config3": "configs[2] streamed: one batch of 1k nodes, 8k GPUs, 40k processes x 40 of the 600 ticks "
                       "(40000 documents, 320000 samples; 4 distinct ticks replicated), hosts partitioned over the ranks",
            "config4": "configs[3] streamed: 12.5k nodes, 100k GPUs, windows of 4 ticks (900 per hour; 3 distinct "
                       "windows resident per rank, visited in turn), one eighth of the hosts per rank",
            "config5": "configs[4] stress: 240 hosts, Zipf(1.2) pods per host up to 160, pods replaced between "
                       "epochs, every third host MIG (utilization [N/A], idle pods see MIG uuids), hosts "
                       "partitioned over the ranks by text bytes"}[config]


# ---------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------
ROTATE = 4          # resident batches per rank, visited round-robin: 4 x 60 MB > 126 MB of L2
                    # (config3: one batch of 750 MB is larger than L2 by itself)


def bind_near_gpu(torch, local_rank):
    """Several ranks on one box: run on the cores of the GPU's own NUMA node, so that the pinned host buffers
    of the end-to-end leg are allocated in the memory its PCIe link reaches without crossing sockets.
    Returns what was done (for the bench line); does nothing when the box does not say."""
    try:
        p = torch.cuda.get_device_properties(local_rank)
        bdf = "%04x:%02x:%02x.0" % (p.pci_domain_id, p.pci_bus_id, p.pci_device_id)
        node = int(open("/sys/bus/pci/devices/%s/numa_node" % bdf).read())
        if node < 0:
            return {"gpu": bdf, "numa_node": node, "bound": False}
        cpus = set()
        for part in open("/sys/devices/system/node/node%d/cpulist" % node).read().strip().split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return {"gpu": bdf, "numa_node": node, "bound": False}
        os.sched_setaffinity(0, cpus)
        return {"gpu": bdf, "numa_node": node, "bound": True, "cpus": len(cpus)}
    except Exception as e:                           # no sysfs, no such attribute: leave the process where it is
        return {"bound": False, "why": type(e).__name__}


def run_kgm_arm(args):
    import torch
    import torch.distributed as dist
    from kgm import monitor, shard

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    # stdout carries the one JSON line only: whatever libraries print there goes to stderr
    json_fd = os.dup(1)
    os.dup2(2, 1)
    torch.cuda.set_device(local_rank)
    numa = bind_near_gpu(torch, local_rank) if world > 1 and os.environ.get("KGM_BENCH_NUMA", "1") == "1" else None
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    device = torch.device("cuda", local_rank)

    # ---- ROTATE distinct batches of the workload per rank, each resident in its own context ----
    # resident batches per rank, visited in turn: configs[1] 4 fleets (4 x 60 MB > L2); the large
    # configs are larger than L2 by themselves (config4: 3 distinct windows of the hour)
    sharded = args.config in ("config3", "config4", "config5")
    rotate = {"config3": 2, "config4": 3, "config5": 4}.get(args.config, ROTATE)
    mons, batches, rps = [], [], []
    shard_info = None
    # configs[1] on several GPUs: a whole step lands on ONE gathering GPU ((N - 1) x 46 MB through its
    # NVLink ingress: 0.36 ms at N = 8, longer than three steps' kernels), so N steps, bound for N different
    # GPUs, have to be in flight for the transfers to stay off the critical path: N contexts per rank,
    # over the same `distinct` fleets
    distinct = rotate
    if not sharded and world > rotate:
        rotate = min(world, 8)
    built = {}
    device_gen = None
    for v in range(rotate):
        if sharded:
            rp, entries, shard_info = build_workload(args.config, rank, world, variant=v if args.config == "config4" else 0)
        else:
            if v % distinct not in built:
                built[v % distinct] = build_workload(args.config, rank * distinct + v % distinct)
            rp, entries, _ = built[v % distinct]
        mon = monitor.Monitor(local_rank)
        mon.set_pipeline({"fused": 0, "staged": 1, "fused-generic": 2, "fused-v1": 3, "fused-v2": 4}[args.pipeline])
        mon.load_fleet(entries)
        batch = mon.make_batch(rp.gpu_texts, rp.apps_texts, rp.doc_node, rp.doc_tick, pinned=True)
        if args.config == "config4":
            # the resident window comes from the device-side generator (kgm_replay_fill): same bytes as the
            # host generator's (tests/test_replay_gpu.py; sizes re-checked here), never on the host
            from kgm import replay_device
            if device_gen is None:
                device_gen = replay_device.DeviceReplay(rp.nodes, rp.node_ids, 0, device=local_rank)
            win = device_gen.window(rp.tick0, rp.window)
            if (win.gpu_bytes, win.apps_bytes, win.n_docs) != (batch.gpu_bytes, batch.apps_bytes, batch.n_docs):
                raise SystemExit("device generator and host generator disagree on the window's sizes")
            win.upload(mon)
        else:
            mon.upload(batch)
        mons.append(mon); batches.append(batch); rps.append(rp)
    samples_step = [rp.n_samples() for rp in rps]
    sizes = []
    for mon in mons:
        r = mon.encode_resident()                    # sizes the workspace
        sizes.append((int(r.n_bytes), int(r.n_lines)))
    n_bytes, n_lines = sizes[0]
    gather = (shard.ShardGather(max(s[0] for s in sizes) + (1 << 20), device,
                                 push=os.environ.get("KGM_SHARD_PUSH", "1") == "1",
                                 rotate_root=os.environ.get("KGM_GATHER_ROTATE", "1") == "1", contexts=rotate)
              if world > 1 else None)

    def sync_all():
        torch.cuda.synchronize(device)
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize(device)

    keys = ("tokenize_ms", "join_ms", "scan_ms", "encode_ms", "total_ms")
    acc = {k: [] for k in keys}
    done_samples = 0

    host_s = {"encode_call": 0.0, "gather_call": 0.0, "finish_wait": 0.0}

    # Steps are queued without waiting for their results: every context has at most one
    # kgm_encode_resident_async in flight, finished (sizes, overflow check, kernel times)
    # when the rotation comes back to it, so the host runs up to `rotate` steps ahead of the
    # GPU.  With more than one rank the shard is handed to the gather on the device: the
    # sizes travel in the 16-byte header in front of the lines.
    use_async = gather is None or gather.mode in ("slots", "push")
    ext = [torch.cuda.ExternalStream(m.stream_handle, device=device) for m in mons]
    inflight = [None] * rotate                       # (timed, event after which the output buffer is free)
    # Each context has its own stream, so the persistent kernel of the next step starts on
    # the SMs the previous one has already left (its last wave of documents is only partly
    # full): consecutive launches overlap at their tails, as they would in a service that
    # streams batches.  The duration of a launch that has the GPU to itself -- what the
    # roofline is computed from -- is therefore measured in a synchronous pass right after the
    # timed region.  KGM_BENCH_SERIALIZE=1 instead chains the steps of the timed region one
    # after the other.
    serialize = os.environ.get("KGM_BENCH_SERIALIZE") == "1"
    last_launch = [None]
    if gather is not None and gather.mode == "slots":
        for m in mons:
            assert gather.stride <= m.output_capacity, (gather.stride, m.output_capacity)

    def finish(v):
        nonlocal done_samples
        if inflight[v] is None:
            return
        timed, fin = inflight[v]
        inflight[v] = None
        mons[v].encode_finish()
        if fin is not None:
            ext[v].wait_event(fin)                   # the shard has left the output buffer
        if timed:
            for k, val in zip(keys, mons[v].kernel_times()):
                acc[k].append(val)
            done_samples += samples_step[v]

    fake_push = None
    if os.environ.get("KGM_BENCH_FAKE_PUSH") and gather is None:
        fake_push = (torch.cuda.Stream(device),
                     [torch.empty(((n_bytes + (1 << 20)) // 16) * 16, dtype=torch.uint8, device=device) for _ in range(rotate)])

    def step(i, timed):
        nonlocal done_samples
        v = i % rotate
        mon = mons[v]
        t0 = time.perf_counter()
        if use_async:
            finish(v)
            if timed:
                host_s["finish_wait"] += time.perf_counter() - t0
            if serialize and last_launch[0] is not None:
                ext[v].wait_event(last_launch[0])    # k_poll launches of different contexts do not overlap
            push = gather is not None and gather.mode == "push"
            if push:                                 # one library call: wait, encode straight into the slot on the
                r = gather.push_step(mon)            # step's gathering rank, tell it the shard has landed
            else:
                r = mon.encode_resident_async()
                if fake_push is not None:            # (experiment) the copy traffic of a multi-GPU step, on one GPU
                    side, bufs = fake_push
                    side.wait_stream(ext[v])
                    with torch.cuda.stream(side):
                        src = shard.tensor_of_device_buffer(int(r.lines) - 16, bufs[v].numel(), device)
                        bufs[v].copy_(src, non_blocking=True)
                    ext[v].wait_stream(side) if os.environ.get("KGM_BENCH_FAKE_PUSH") == "2" else None
            if serialize:
                last_launch[0] = torch.cuda.Event()
                last_launch[0].record(ext[v])
            t1 = time.perf_counter()
            fin = None
            if gather is not None and not push:
                fin = gather.start_device(int(r.lines), ext[v])
            inflight[v] = (timed, fin)
            if timed:
                host_s["encode_call"] += t1 - t0
                host_s["gather_call"] += time.perf_counter() - t1
            return
        r = mon.encode_resident()                    # device time per kernel from the library's CUDA events
        t1 = time.perf_counter()
        local = shard.tensor_of_device_buffer(r.lines, max(int(r.n_bytes), 1), device)
        gather.start(local, int(r.n_bytes), int(r.n_lines))
        if timed:
            host_s["encode_call"] += t1 - t0
            host_s["gather_call"] += time.perf_counter() - t1
            for k, val in zip(keys, mon.kernel_times()):
                acc[k].append(val)
            done_samples += samples_step[v]

    def drain():
        for v in range(rotate):
            finish(v)
        if gather is not None and gather.mode == "push":
            for m in mons:
                m.set_output_buffer(0, 0)            # synchronous calls write the context's own buffer again

    n_warm = max(args.warmup, rotate)                # every context runs at least once before the timed region
    for i in range(n_warm):
        step(i, False)
    drain()
    concat_check = None
    pack_ms = None
    if gather:
        gather.finish()
        # one untimed check that the gathering rank really holds the rank-major concatenation
        import zlib
        last = (n_warm - 1) % rotate
        mons[last].encode_resident()
        mine = zlib.crc32(mons[last].fetch().body)
        crcs = [None] * world
        dist.all_gather_object(crcs, mine)
        holder = gather.last_root if gather.mode == "push" else 0      # the rank the last step was gathered on
        if rank == holder:
            off = gather.shard_offsets()
            whole = gather.body().cpu().numpy().tobytes()
            concat_check = all(zlib.crc32(whole[int(off[r]):int(off[r + 1])]) == crcs[r] for r in range(world))
            # what closing the gaps between the slots costs when it has the GPU to itself
            pe = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
            gather.pack()
            pe[0].record(gather.stream)
            for _ in range(3):
                gather.pack()
            pe[1].record(gather.stream)
            gather.stream.synchronize()
            pack_ms = pe[0].elapsed_time(pe[1]) / 3
        seen = [None] * world
        dist.all_gather_object(seen, (concat_check, pack_ms))
        concat_check, pack_ms = seen[holder]
    # clocks / throttle reasons: ONE nvidia-smi (rank 0), started well before the timed region so
    # that its start-up does not fall into it; it keeps sampling through the region and through
    # an untimed repetition of the same steps right after it (the region itself is a few
    # milliseconds, shorter than one sampling period)
    clocks = ClockSampler(local_rank) if rank == 0 else None
    if clocks:
        clocks.__enter__()
        time.sleep(0.6)
    sync_all()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    wall0 = time.perf_counter()
    ev0.record()
    for i in range(args.steps):
        step(n_warm + i, True)
    drain()
    if gather:
        gather.finish()
    ev1.record()
    torch.cuda.synchronize(device)
    sync_all()
    wall_ms = 1e3 * (time.perf_counter() - wall0)
    region_ms = ev0.elapsed_time(ev1)
    step_ms = region_ms / args.steps
    n_extra = int(min(4000, max(20, 500.0 / max(step_ms, 1e-3))))   # ~0.5 s of the same steps, untimed, under the sampler
    if world > 1:                                    # (every rank must run the same number of steps)
        t = torch.tensor([n_extra], dtype=torch.int64, device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        n_extra = int(t.item())
    for extra in range(n_extra):
        step(n_warm + args.steps + extra, False)
    drain()
    if gather:
        gather.finish()
    if clocks:
        clocks.__exit__(None, None, None)
    mean = {k: sum(v) / len(v) for k, v in acc.items()}
    if not serialize:
        # the launches of the timed region overlap, so their own event times include waiting
        # for SMs: the duration of a launch that has the GPU to itself is measured right after,
        # on the same resident batches, one synchronous call at a time
        iso = {k: [] for k in keys}
        for i in range(max(8, min(args.steps, 40))):
            mons[i % rotate].encode_resident()
            if i >= 4:
                for k, val in zip(keys, mons[i % rotate].kernel_times()):
                    iso[k].append(val)
        in_region = mean
        mean = {k: sum(v) / len(v) for k, v in iso.items()}
    last_t = mons[0].timing()
    fused = bool(last_t["fused"])
    kernel_name = mons[0].last_kernel() if fused else "k_encode"
    launches_per_step = int(last_t["launches"])
    slow_docs = int(last_t["slow_docs"])

    # ---- configs[3] streamed: window after window of the hour, generated on the device, encoded, ----
    # ---- checksummed (sizes) and discarded; untimed by the contract's bracket, reported on its own ----
    stream_info = None
    if args.config == "config4" and device_gen is not None and args.stream_windows > 0:
        mon = mons[0]
        if gather is not None and gather.mode == "push":
            mon.set_output_buffer(0, 0)
        torch.cuda.synchronize(device)
        t0 = time.perf_counter()
        done_w = tot_lines = tot_bytes = tot_samples = 0
        win_ticks = rps[0].window
        for wdx in range(args.stream_windows):
            win = device_gen.window(wdx * win_ticks, win_ticks)
            win.upload(mon)
            r = mon.encode_resident()
            tot_lines += int(r.n_lines); tot_bytes += int(r.n_bytes)
            tot_samples += samples_step[0]           # every window has the same hosts and ticks
            done_w += 1
            if time.perf_counter() - t0 > args.stream_seconds:
                break
        torch.cuda.synchronize(device)
        secs = time.perf_counter() - t0
        agg = torch.tensor([float(tot_samples), float(tot_lines), float(tot_bytes), float(done_w)], dtype=torch.float64,
                           device=device)
        tmax = torch.tensor([secs], dtype=torch.float64, device=device)
        if world > 1:
            dist.all_reduce(agg, op=dist.ReduceOp.SUM)
            dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        stream_info = {"windows_per_rank": done_w, "of_windows_per_hour": 3600 // win_ticks,
                       "samples": int(agg[0].item()), "lines": int(agg[1].item()), "bytes": int(agg[2].item()),
                       "seconds": float(tmax[0].item()), "samples_per_s": float(agg[0].item() / tmax[0].item()),
                       "what": "per window: kgm_replay_offsets + kgm_replay_fill (texts generated in HBM), "
                               "kgm_upload_batch from device memory, kgm_encode_resident; sizes summed as the "
                               "checksum, output discarded; wall clock, synchronous, one context"}
        # the first window back into the context for the end-to-end leg's sizes
        device_gen.window(rps[0].tick0, win_ticks).upload(mon)
        mon.encode_resident()

    # ---- end to end through the C ABI with host buffers ----
    e2e_times = []
    e2e_warm = max(args.warmup, rotate)              # every context's first call allocates its pinned result buffers
    for i in range(e2e_warm + args.steps):
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        mons[i % rotate].encode_raw(batches[i % rotate], want_host_copy=True)
        dt = time.perf_counter() - t0
        if i >= e2e_warm:
            e2e_times.append(dt)
    e2e_ms = 1e3 * sum(e2e_times) / len(e2e_times)
    b0 = batches[0]
    h2d = b0.gpu_bytes + b0.apps_bytes + 2 * 8 * (b0.n_docs + 1) + 2 * 4 * b0.n_docs
    d2h = n_bytes + 8 * (n_lines + 1) + 8 * (b0.n_docs + 1) + 4 * b0.n_docs + 48
    host_result = None
    if rank == 0 and world == 1:
        mons[0].encode_resident()
        host_result = mons[0].fetch()
    # what the D2H copy of one result alone costs on this box (the e2e floor)
    pin = torch.empty(n_bytes, dtype=torch.uint8).pin_memory()
    src = torch.empty(n_bytes, dtype=torch.uint8, device=device)
    for _ in range(3):
        pin.copy_(src, non_blocking=True)
    torch.cuda.synchronize(device)
    t0 = time.perf_counter()
    for _ in range(5):
        pin.copy_(src, non_blocking=True)
    torch.cuda.synchronize(device)
    d2h_only_ms = 1e3 * (time.perf_counter() - t0) / 5

    # ---- max over ranks ----
    if world > 1:
        t = torch.tensor([step_ms, e2e_ms, mean["encode_ms"]], dtype=torch.float64, device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        step_ms, e2e_ms, enc_ms = (float(x) for x in t.tolist())
        tot = torch.tensor([done_samples, samples_step[0]], dtype=torch.float64, device=device)
        dist.all_reduce(tot)
        samples_total, samples_one_step = float(tot[0].item()), float(tot[1].item())
    else:
        enc_ms = mean["encode_ms"]
        samples_total, samples_one_step = float(done_samples), float(samples_step[0])

    if rank == 0:
        peak, peak_src = peaks()
        algo_bytes = b0.gpu_bytes + b0.apps_bytes + n_bytes             # per launch, this rank
        achieved = algo_bytes / (enc_ms * 1e-3) / 1e9
        traffic = None
        tpath = os.path.join(ROOT, "profiles", {"k_poll": "poll_traffic.json", "k_poll2": "poll2_traffic.json"}.get(
            kernel_name, "encode_traffic.json"))
        if os.path.isfile(tpath):
            try:
                tj = json.load(open(tpath))
                # (measured on the configs[1] batch: only that workload's line carries it)
                traffic = tj.get("dram_bytes_per_launch") if args.config == "config2" else None
            except Exception:
                traffic = None
        line = {
            "metric": METRIC, "value": samples_total / (step_ms * args.steps * 1e-3), "unit": "samples/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": step_ms,
            "higher_is_better": True, "scaling": "strong" if args.config in ("config3", "config5") else "weak",
            "vs_baseline": None, "dtype": "u8",
            "data": "synthetic",
            "config": {"workload": workload_name(args.config), "docs_per_gpu": b0.n_docs, "shard": shard_info,
                       "samples_per_gpu": samples_step[0], "lines_per_gpu": n_lines,
                       "input_bytes_per_gpu": b0.gpu_bytes + b0.apps_bytes,
                       "output_bytes_per_gpu": n_bytes,
                       "l2": "%d resident batches per GPU (%d distinct fleets) visited round-robin (%d MB working set "
                             "> 126 MB L2), no flush inside the timed bracket" % (rotate, distinct, rotate * algo_bytes // 1000000),
                       "parallelism": ("hosts sharded over %d ranks, no data-path collective inside the poll path; "
                                       "per step every rank's encoded shard and its {bytes, lines} header are "
                                       "gathered (%s; inside the timed bracket, queued on the device behind the "
                                       "step's kernel and overlapping the next step's); rank-major concatenation "
                                       "of the gathered slots verified against every rank's own result: %s; "
                                       "packing the slots into one contiguous buffer (kgm_concat_shards) is done "
                                       "on demand, not per step: %s ms when it has the GPU to itself"
                                       % (world, {"slots": "one fixed-size NCCL all-gather straight out of the "
                                                           "library's output buffers into fixed-stride slots, sizes "
                                                           "never visit the host",
                                                  "push": "k_poll itself writes every shard into its fixed-stride slot "
                                                          "on the step's gathering rank (%s) over NVLink peer memory "
                                                          "(CUDA IPC) while it encodes; NCCL carries one 4-byte "
                                                          "all-reduce per step as the landed / slot-free signal"
                                                          % ("step k's whole concatenation lands on rank k mod N, so "
                                                             "the ingress is spread over all NVLink ports"
                                                             if gather.rotate_root else "always rank 0"),
                                                  "p2p": "NCCL send/recv"}[gather.mode], concat_check,
                                          "%.3f" % pack_ms if pack_ms is not None else "n/a"))
                       if world > 1 else "1 rank"},
            "timing": "CUDA events around %d back-to-back steps (barrier + synchronize on both sides), max over "
                      "ranks; steps are queued asynchronously, every context on its own stream and finished when "
                      "the rotation returns to it%s; kernel_ms from the library's own events on its stream%s"
                      % (args.steps, ", launches chained one after the other" if serialize else "",
                         "" if serialize else " in a synchronous pass after the region"),
            "kernel_ms": {k: round(mean[k], 5) for k in keys},
            "host_ms_per_step_rank0": {k: round(1e3 * v / args.steps, 4) for k, v in host_s.items()},
            "wall_ms_region": wall_ms,
            "pipeline_frac_of_peak": algo_bytes / (mean["total_ms"] * 1e-3) / 1e9 / peak,
            "pipeline": "fused (%s)" % kernel_name if fused else "staged (k_tokenize, k_join, k_scan, k_encode)",
            "slow_docs_per_step": slow_docs,
            "fused_ctas_per_sm": int(last_t.get("fused_ctas_per_sm", 0)),
            "docs_in_flight_per_sm": int(last_t.get("docs_in_flight_per_sm", 0)),
            "roofline": {"bound": "hbm", "kernel": kernel_name, "achieved": achieved,
                         "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                         "algorithmic_bytes_per_launch": algo_bytes, "peak_source": peak_src,
                         "frac_of_nominal_8TBs": achieved / 8000.0,
                         "launch_ms": enc_ms,
                         "launch_ms_source": "library CUDA events around the kernel on its stream, " + (
                             "timed region (launches chained)" if serialize else
                             "synchronous pass right after the timed region (in the region consecutive launches "
                             "overlap at their tails)"),
                         "achieved_in_region": algo_bytes / (step_ms * 1e-3) / 1e9 if world == 1 else None},
            "e2e": {"value": samples_one_step / (e2e_ms * 1e-3), "unit": "samples/s",
                    "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "ms_per_step": e2e_ms,
                    "d2h_copy_alone_ms": d2h_only_ms,
                    "note": "kgm_encode_batch with pinned host buffers: chunked H2D / k_poll / D2H overlap; "
                            "bounded below by the D2H copy of the result over PCIe (d2h_copy_alone_ms)"},
            "gpu_launches": launches_per_step * args.steps,
            "stream": stream_info,
            "numa": numa,
            "clocks": clocks.summary(),
        }
        if world > 1:
            line["gather_mode"] = gather.mode
            line["concat_verified"] = bool(concat_check)
            line["pack_ms"] = pack_ms
        if world == 1 and not args.no_cpu:
            line["cpu_baseline"] = cpu_baseline_single_core(rps[0], host_result)
        os.write(json_fd, (json.dumps(line) + "\n").encode())
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    for mon in mons:
        mon.close()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=int(os.environ.get("WORLD_SIZE", "1")))
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="kgm", choices=["kgm", "reference"])
    ap.add_argument("--config", default="config2", choices=["config1", "config2", "config3", "config4", "config5"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--stream-windows", type=int, default=900,
                    help="config4: windows of the hour streamed after the timed region (0 = none)")
    ap.add_argument("--stream-seconds", type=float, default=20.0, help="config4: stop streaming after this long")
    ap.add_argument("--pipeline", default="fused", choices=["fused", "staged", "fused-generic", "fused-v1", "fused-v2"])
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    if args.impl == "reference":
        return run_reference_arm(args)
    return run_kgm_arm(args)


if __name__ == "__main__":
    sys.exit(main())
