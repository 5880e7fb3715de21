"""TEST INFRASTRUCTURE ONLY -- CPU statement of the stage-3 window reductions.

PARITY UNPINNED: the reference has no counterpart for this stage (``gpu.py`` emits raw
per-poll lines, ``gpu.py:136-143,206-216``, and never aggregates), so there is nothing in
it to pin this file against.  What is pinned is the *input* of the reductions: the samples
are read back from the very lines ``oracle/poll_oracle.py`` emits (itself pinned to the
reference), plus the ``power.draw`` cell of an extended ``--query-gpu`` text parsed with the
reference's own rules (``csv.reader``, ``lstrip`` on the key, last row wins: ``gpu.py:40-55``,
``re.sub('\\D','',cell)``: ``gpu.py:133``).  The reductions themselves are plain Python.

    windows(docs, window_ticks) -> {(kind, entity, gpu_id, window): group}

``docs`` = iterable of ``(node_index, tick, lines, gpu_csv, container_index_of)`` in batch
order, where ``container_index_of`` maps a container id of that node to its global index.
Only ``tests/`` may import this module.
"""
from __future__ import annotations

import csv
import io
import re

from oracle.poll_oracle import digits_of

SAT = 0xFFFFFFFE
NONE_ID = 0xFFFF
METRICS = ("gpu_util", "mem", "power")
_NODE = re.compile(r"^node_gpu,host=.*?,gpu_id=([^ ,]*) gpu_util=([^,]*),mem_util=[^,]*,mem_used=([^,]*),")
_POD = re.compile(r"^pod_gpu,host=.*?,namespace=.*?,gpu_id=([^,]*),containerid=([^,]*),pid_t=.*"
                  r"\",pid=\"(.*?)\",gpu_id_v=.*,gpu_util=([^,]*),used_memory=([^,]*)$")


def _value(digits: str):
    """int(digits) saturating at 2^32-2; '' -> missing."""
    return None if digits == "" else min(int(digits), SAT)


def power_by_short_id(gpu_csv: str):
    """short id -> power value of that gpu's values row (last row with the key, gpu.py:54)."""
    stream = io.StringIO(gpu_csv)
    stream.readline()                                     # gpu.py:67-68
    table = {}
    for cells in csv.reader(stream):
        table[cells[0].lstrip()] = cells                  # gpu.py:108, 54
    out = {}
    for i, cells in enumerate(table.values()):            # gpu.py:119-120
        out[str(i)] = _value(digits_of(cells[6])) if len(cells) > 6 else None
    return out


def samples_of_doc(lines, gpu_csv, node_index, container_index_of):
    """(kind, entity, gpu_id, {metric: value-or-None}) per emitted line, in line order."""
    power = power_by_short_id(gpu_csv) if lines else {}
    out = []
    for line in lines:
        m = _NODE.match(line)
        if m:
            sid, util, mem = m.groups()
            out.append((0, node_index, int(sid), {"gpu_util": _value(digits_of(util)), "mem": _value(digits_of(mem)),
                                                  "power": power.get(sid)}))
            continue
        m = _POD.match(line)
        assert m, line
        sid, cid, pid, util, mem = m.groups()
        gid = NONE_ID if sid == "None" else int(sid)
        idle = pid == "-"
        vals = {"gpu_util": _value(digits_of(util)), "mem": _value(digits_of(mem)),
                "power": None if idle else power.get(sid)}
        out.append((1, container_index_of[cid], gid, vals))
    return out


def windows(docs, window_ticks):
    groups = {}
    for d, (node_index, tick, lines, gpu_csv, container_index_of) in enumerate(docs):
        for pos, (kind, entity, gid, vals) in enumerate(samples_of_doc(lines, gpu_csv, node_index,
                                                                        container_index_of)):
            key = (kind, entity, gid, tick // window_ticks)
            g = groups.setdefault(key, {"count": 0, "node": node_index,
                                        **{m: {"count": 0, "min": None, "max": None, "sum": 0, "last": None,
                                               "order": None} for m in METRICS}})
            g["count"] += 1
            for m in METRICS:
                v = vals[m]
                if v is None:
                    continue
                a = g[m]
                a["count"] += 1
                a["min"] = v if a["min"] is None else min(a["min"], v)
                a["max"] = v if a["max"] is None else max(a["max"], v)
                a["sum"] += v
                if a["order"] is None or (tick, d, pos) > a["order"]:
                    a["order"], a["last"] = (tick, d, pos), v
    for g in groups.values():
        for m in METRICS:
            a = g[m]
            a["mean"] = a["sum"] / a["count"] if a["count"] else 0.0
    return dict(sorted(groups.items()))


def lines(groups, host_of_node, cid_of_container, window_ticks, t0_ns=0, tick_ns=10 ** 9):
    """The groups of windows() in line form (PARITY UNPINNED like the rest of this file):
    node_gpu_w / pod_gpu_w measurements with explicit nanosecond timestamps (the window's last
    tick), mirroring what the dashboards compute in InfluxQL over node_gpu / pod_gpu fields
    (gpu.py:136-143,206-216)."""
    out = []
    for (kind, entity, gid, window), g in groups.items():
        fields = []
        for m, name in (("gpu_util", "gpu_util"), ("mem", "mem_used"), ("power", "power")):
            a = g[m]
            if a["count"]:
                fields += ["%s_min=%d" % (name, a["min"]), "%s_max=%d" % (name, a["max"]),
                           "%s_mean=%s" % (name, repr(a["mean"])), "%s_last=%d" % (name, a["last"])]
        fields.append("samples=%di" % g["count"])
        gid_s = "None" if gid == NONE_ID else str(gid)
        ts = t0_ns + ((window + 1) * window_ticks - 1) * tick_ns
        host = host_of_node[g["node"]]
        head = ("node_gpu_w,host=%s,gpu_id=%s" % (host, gid_s) if kind == 0 else
                "pod_gpu_w,host=%s,containerid=%s,gpu_id=%s" % (host, cid_of_container[entity], gid_s))
        out.append("%s %s %d" % (head, ",".join(fields), ts))
    return out
