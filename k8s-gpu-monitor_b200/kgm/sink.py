"""Sink side of the path: what leaves towards InfluxDB.

The reference POSTs one un-timestamped line per request (``gpu.py:145,218,252``).  A replayed
fleet batch needs more: explicit timestamps (or all ticks of a replay collapse to the server's
arrival time), bodies of many lines, and the additive window measurements of stage 3 in line
form.  Everything here is OFF the parity path: the raw ``node_gpu`` / ``pod_gpu`` bytes the
library emits are never changed unless the caller asks for a timestamp suffix.

    with_timestamps(result, ts_ns_per_doc)          raw lines + " <ns>"   (InfluxDB line protocol's optional 3rd part)
    write_bodies(body, max_bytes, compress)        line-aligned chunks for POST /write, optionally gzip'd
    window_lines(groups, fleet, window_ticks, ...) node_gpu_w / pod_gpu_w lines from kgm_window_reduce's groups
"""
from __future__ import annotations

import gzip

import numpy as np

METRICS = (("gpu_util", "gpu_util"), ("mem", "mem_used"), ("power", "power"))


def with_timestamps(result, ts_ns_per_doc) -> bytes:
    """Every line of document d gets `` <ts_ns_per_doc[d]>`` appended (nanoseconds since the
    epoch, the precision InfluxDB assumes by default).  ``result`` is a HostResult."""
    body = np.frombuffer(result.body, np.uint8)
    line_off = np.asarray(result.line_off, np.int64)
    doc_line_off = np.asarray(result.doc_line_off, np.int64)
    n_lines = int(line_off.shape[0]) - 1
    if n_lines <= 0:
        return b""
    ts = [b" %d" % int(t) for t in ts_ns_per_doc]
    doc_of_line = np.repeat(np.arange(len(ts)), np.diff(doc_line_off))
    add = np.asarray([len(t) for t in ts], np.int64)[doc_of_line]
    new_off = np.concatenate([[0], np.cumsum(np.diff(line_off) + add)])
    out = np.empty(int(new_off[-1]), np.uint8)
    # the line without its '\n', the suffix, the '\n'
    src_len = np.diff(line_off) - 1
    idx = np.repeat(new_off[:-1] - line_off[:-1], src_len) + np.concatenate(
        [np.arange(a, a + n) for a, n in zip(line_off[:-1], src_len)]) if n_lines else np.zeros(0, np.int64)
    out[idx] = body[np.concatenate([np.arange(a, a + n) for a, n in zip(line_off[:-1], src_len)])]
    for i in range(n_lines):
        t = ts[int(doc_of_line[i])]
        p = int(new_off[i] + src_len[i])
        out[p:p + len(t)] = np.frombuffer(t, np.uint8)
        out[p + len(t)] = 10
    return out.tobytes()


def write_bodies(body: bytes, max_bytes: int = 4 << 20, compress: bool = True):
    """Cuts a '\\n'-terminated body into chunks of whole lines of at most ``max_bytes`` (a
    longer single line becomes its own chunk) -> [(payload, headers)], gzip'd if asked."""
    out, pos, n = [], 0, len(body)
    while pos < n:
        end = min(n, pos + max_bytes)
        if end < n:
            cut = body.rfind(b"\n", pos, end)
            end = cut + 1 if cut >= pos else body.find(b"\n", end) + 1 or n
        chunk = body[pos:end]
        if compress:
            out.append((gzip.compress(chunk, 1), {"Content-Encoding": "gzip"}))
        else:
            out.append((chunk, {}))
        pos = end
    return out


def _fields(g, i):
    parts = []
    for key, name in METRICS:
        if int(g[key + "_count"][i]) == 0:
            continue
        parts += ["%s_min=%d" % (name, g[key + "_min"][i]), "%s_max=%d" % (name, g[key + "_max"][i]),
                  "%s_mean=%r" % (name, float(g[key + "_mean"][i])), "%s_last=%d" % (name, g[key + "_last"][i])]
    parts.append("samples=%di" % g["count"][i])
    return ",".join(parts)


def window_lines(groups, host_of_node, cid_of_container, window_ticks, t0_ns=0, tick_ns=10 ** 9):
    """Stage 3 as additive measurements (never mixed into node_gpu / pod_gpu):
        node_gpu_w,host=H,gpu_id=K <fields> <ts>
        pod_gpu_w,host=H,containerid=C,gpu_id=G <fields> <ts>
    fields = <metric>_{min,max,mean,last} for gpu_util, mem_used, power (those with samples) and
    samples=<lines in the group>i; ts = the window's last tick.  ``groups`` is what
    Monitor.window_reduce returns; gpu_id 0xffff is the idle branch's None (gpu.py:226)."""
    out = []
    for i in range(int(groups["n"])):
        gid = int(groups["gpu_id"][i])
        gid_s = "None" if gid == 0xFFFF else str(gid)
        ts = t0_ns + ((int(groups["window"][i]) + 1) * window_ticks - 1) * tick_ns
        host = host_of_node[int(groups["node"][i])]
        if int(groups["kind"][i]) == 0:
            head = "node_gpu_w,host=%s,gpu_id=%s" % (host, gid_s)
        else:
            head = "pod_gpu_w,host=%s,containerid=%s,gpu_id=%s" % (host, cid_of_container[int(groups["container"][i])],
                                                                    gid_s)
        out.append("%s %s %d" % (head, _fields(groups, i), ts))
    return out
