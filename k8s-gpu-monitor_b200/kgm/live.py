"""Live single-node producer: one poll the way the reference takes it, encoded by the library.

``gpu.py`` shells out for every text it needs (``gpu.py:92-116,177,184,221-224``) and POSTs each
line (``gpu.py:145,218,252``).  ``LivePoller`` issues the same commands -- same argv, same shell
pipeline -- through an injectable ``command(args) -> str`` / ``popen(cmd) -> str``, builds the
node's fleet entry with the reference's own parsing rules (kgm/fleet.py) and hands the two
nvidia-smi texts to the library.  It restores drop-in use on one node (gpu_monitor.yaml); it is
not a performance path.  Without docker / nvidia-smi (the build and GPU boxes have neither) it
is driven with canned texts, as tests/test_live_cpu.py does.
"""
from __future__ import annotations

import os
import subprocess

from .fleet import node_entry

# argv exactly as the reference builds them (gpu.py:58-73: csvheaderargs / commandtodictdict)
PS_ARGS = ["docker", "ps", "--format", "'{{.ID}},{{.Image}},{{.Ports}}'"]
STATS_ARGS = ["docker", "stats", "--no-stream", "--format", "'{{.Container}},{{.MemUsage}},{{.CPUPerc}}'"]
GPU_ARGS = ["nvidia-smi", "--format=csv",
            "--query-gpu=gpu_uuid,utilization.gpu,utilization.memory,memory.total,memory.used,memory.free"]
APPS_ARGS = ["nvidia-smi", "--format=csv", "--query-compute-apps=pid,process_name,gpu_uuid,used_memory"]
SMI_LIST_FMT = "docker exec -i %s nvidia-smi -L |cut -d: -f3|cut -d\\) -f1"   # gpu.py:221


def _command(args):                                   # gpu.py:36-37
    return subprocess.check_output(args).decode()


def _popen(cmd):                                      # gpu.py:221-223
    pipe = os.popen(cmd)
    try:
        return pipe.read()
    finally:
        pipe.close()


class LivePoller:
    def __init__(self, monitor, command=None, popen=None, node_name=None, sink=None):
        self.mon = monitor
        self.command = command or _command
        self.popen = popen or _popen
        # gpu.py:131 re-reads NODE_NAME inside main(); '%s' prints None when it is unset
        self.node_name = node_name
        self.sink = sink                             # callable(bytes): e.g. one batched POST to /write

    def collect(self):
        """The eight inputs of one poll, in the order the reference reads them."""
        ps_text = self.command(PS_ARGS)
        stats_text = self.command(STATS_ARGS)
        gpu_csv = self.command(GPU_ARGS)
        apps_csv = self.command(APPS_ARGS)
        top, inspect, smi_list = {}, {}, {}
        for row in stats_text.split("\n"):           # container ids in docker stats order (gpu.py:101-103,174)
            if not row:
                continue
            cid = row.split(",")[0][1:]              # the literal leading quote of the unparsed format (gpu.py:101)
            if cid in top:
                continue
            top[cid] = self.command(["docker", "top", cid, "-eo", "pid"])
            inspect[cid] = self.command(["docker", "inspect", cid])
            smi_list[cid] = self.popen(SMI_LIST_FMT % cid)
        name = self.node_name if self.node_name is not None else os.environ.get("NODE_NAME")
        return dict(node_name=name, gpu_csv=gpu_csv, apps_csv=apps_csv, ps_text=ps_text, stats_text=stats_text,
                    top=top, inspect=inspect, smi_list=smi_list)

    def poll(self):
        """One poll: (lines, exception class name or None), as gpu.main() would have POSTed."""
        x = self.collect()
        entry = node_entry(x["node_name"], x["ps_text"], x["stats_text"], x["top"], x["inspect"], x["smi_list"])
        self.mon.load_fleet([entry])
        res = self.mon.encode([x["gpu_csv"]], [x["apps_csv"]], [0])
        if self.sink is not None and res.n_bytes:
            self.sink(res.body)
        return res.lines(0), res.error(0)
