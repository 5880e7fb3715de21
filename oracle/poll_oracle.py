"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference poll path.

This module restates what one call of the reference's ``main()`` (``gpu.py:90-256``)
computes from the texts its subprocesses return, as a pure function:

    poll(node_name, gpu_csv, apps_csv, ps_text, stats_text, top, inspect, smi_list)
        -> (lines, error_class_name_or_None)

It is the *checker* for the CUDA path: only ``tests/``, ``__graft_entry__.smoke()`` and
``bench.py``'s CPU-baseline / ``--impl reference`` legs may import it.  The product
(``k8s-gpu-monitor_b200/``) never does, and has no CPU fallback.

Parity pin: the reference has no tests or golden vectors of its own,
so this restatement is pinned by differential runs against the unmodified script --
``tests/test_oracle.py`` (in the build container, where ``/root/reference`` exists) and the
committed fixtures ``tests/golden/*.json.gz`` produced from the reference by
``tests/golden/make_golden.py``.  The stdlib pieces the reference leans on (``csv.reader``,
``re.sub('\\D')``, ``int``, float ``*`` ``/``, ``repr(float)``, ``json.loads``, ``str.split``)
are used here as-is, i.e. pinned to this image's CPython 3.12.

Each step cites the reference lines it follows.
"""
from __future__ import annotations

import csv
import io
import json
import re

_NON_DIGIT = re.compile(r"\D")


def digits_of(cell: str) -> str:
    """``re.sub('\\D', '', cell)`` -- gpu.py:133-134,138-143,204-205."""
    return _NON_DIGIT.sub("", cell)


def table_from_csv(text, columns, key, *, drop_first_line=False, cell_filters=None):
    """``commandtodictdict`` + ``csvtodictdict`` (gpu.py:40-55, 62-73).

    Rows come from ``csv.reader`` (default dialect) over the text, optionally after one
    blind ``readline()`` (gpu.py:67-68).  Cells are paired with ``columns`` and truncated
    to the shorter (gpu.py:49); the key column(s) are popped (KeyError when missing,
    gpu.py:51,53); a repeated key keeps its first position and takes the later row's
    values (dict assignment, gpu.py:54).
    """
    cell_filters = cell_filters or {}
    stream = io.StringIO(text)
    if drop_first_line:
        stream.readline()
    table = {}
    for cells in csv.reader(stream):
        record = {}
        for name, cell in zip(columns, cells):
            fn = cell_filters.get(name)
            record[name] = fn(cell) if fn else cell
        if isinstance(key, str):
            k = record.pop(key)
        else:
            k = tuple(record.pop(part) for part in key)
        table[k] = record
    return table


def user_of_pod(podname: str) -> str:
    """gpu.py:196-199 (and 230-233)."""
    marker = "turing-gpu-notebook-"
    if marker in podname:
        return podname.split(marker)[1].rsplit("-", 2)[0]
    return podname.split("-")[1]


def labels_of(inspect_text: str):
    """``getContainer(c)['Config']['Labels']`` -- gpu.py:86-88,184."""
    squeezed = inspect_text.replace("\n", "").replace(" ", "")
    return json.loads(squeezed)[0]["Config"]["Labels"]


def node_line(host, short_id, row):
    """gpu.py:133-143.  Raises exactly where the reference would."""
    total = int(digits_of(row["memory.total"]))
    used = int(digits_of(row["memory.used"]))
    ratio = used * 100.0 / total
    return "".join((
        "node_gpu,host=", str(host), ",gpu_id=", short_id,
        " gpu_util=", digits_of(row["used_gpu"]),
        ",mem_util=", repr(ratio),
        ",mem_used=", digits_of(row["memory.used"]),
        ",mem_total=", digits_of(row["memory.total"]),
        ",node=\"", str(host), "\",gid=", short_id,
        ",mem_free=", digits_of(row["memory.free"])))


def pod_line(host, namespace, gpu_id, cid, pid, gpu_util, user, procname, podname, used_memory):
    """gpu.py:206-216 / 240-250 (same template for the busy and the idle branch)."""
    return "".join((
        "pod_gpu,host=", str(host), ",namespace=", namespace, ",gpu_id=", str(gpu_id),
        ",containerid=", cid, ",pid_t=", pid, ",gpu_util_t=", str(gpu_util),
        ",user_name_t=", user,
        " procressname=\"", procname, "\",podname=\"", podname, "\",pid=\"", pid,
        "\",gpu_id_v=\"", str(gpu_id), "\",node=\"", str(host), "\",user_name=\"", user,
        "\",gpu_util=", str(gpu_util), ",used_memory=", str(used_memory)))


def _poll(emit, node_name, gpu_csv, apps_csv, ps_text, stats_text, top, inspect, smi_list):
    strip1 = lambda s: s[1:]
    lstrip = lambda s: s.lstrip()
    # gpu.py:92-116 -- the four tables, in this order (an exception in a later parse
    # happens after the earlier ones succeeded, but before any line is emitted).
    ps = table_from_csv(ps_text, ["ID", "Image", "Ports"], "ID", cell_filters={"ID": strip1})
    stats = table_from_csv(stats_text, ["Container", "MemUsage", "CPUPerc"], "Container",
                           cell_filters={"Container": strip1})
    gpus = table_from_csv(gpu_csv,
                          ["gpu_uuid", "utilization.gpu", "utilization.memory",
                           "memory.total", "memory.used", "memory.free"],
                          "gpu_uuid", drop_first_line=True, cell_filters={"gpu_uuid": lstrip})
    procs = table_from_csv(apps_csv, ["pid", "process_name", "gpu_uuid", "used_memory"],
                           ["pid", "gpu_uuid"], drop_first_line=True,
                           cell_filters={"gpu_uuid": lstrip})

    # gpu.py:119-128 -- uuid -> "0","1",... in first-seen order; utilization.gpu -> used_gpu
    short = {uuid: str(i) for i, uuid in enumerate(gpus)}
    by_short = {}
    for uuid, row in gpus.items():
        row["used_gpu"] = row.pop("utilization.gpu")
        by_short[short[uuid]] = row

    # gpu.py:131-146 -- node lines
    for sid, row in by_short.items():
        emit(node_line(node_name, sid, row))

    # gpu.py:151 -- process table re-keyed by short id (KeyError on an unknown uuid)
    procs = {(pid, short[uuid]): row for (pid, uuid), row in procs.items()}

    # gpu.py:172 -- container set and order come from `docker stats`; ps must know each id
    for cid in stats:
        ps[cid]
    # gpu.py:174-253
    for cid in stats:
        pids = top[cid].split("\n")[1:-1]                       # gpu.py:177
        mine = [(k, row) for k, row in procs.items() if k[0] in pids]   # gpu.py:178
        namespace = podname = user = "-"                        # gpu.py:181-183
        labels = labels_of(inspect[cid])                        # gpu.py:184
        gate = labels["mlgpu"] if "mlgpu" in labels.keys() else "False"   # gpu.py:185-187
        if gate != "mlgpu":
            continue
        if mine:                                                # gpu.py:190-219
            for (pid, sid), row in mine:
                labels = labels_of(inspect[cid])                # gpu.py:192
                if "io.kubernetes.pod.name" in labels.keys():
                    podname = labels["io.kubernetes.pod.name"]
                    namespace = labels["io.kubernetes.pod.namespace"]
                    user = user_of_pod(podname)
                procname = row["process_name"].replace(" ", "")
                used_memory = digits_of(row["used_memory"])
                gpu_util = digits_of(by_short[sid].get("used_gpu"))
                emit(pod_line(node_name, namespace, sid, cid, pid, gpu_util, user,
                              procname, podname, used_memory))
        else:                                                   # gpu.py:220-253
            for uuid in str(smi_list.get(cid, "")).split():
                sid = short.get(uuid)
                if "io.kubernetes.pod.name" in labels.keys():
                    podname = labels["io.kubernetes.pod.name"]
                    namespace = labels["io.kubernetes.pod.namespace"]
                    user = user_of_pod(podname)
                emit(pod_line(node_name, namespace, sid, cid, "-", 0, user, "-", podname, 0))


def poll(node_name, gpu_csv, apps_csv, ps_text, stats_text, top, inspect, smi_list):
    """One poll.  Returns ``(lines, err)``: the lines the reference would have POSTed, in
    order, and the class name of the exception it would have died with (or ``None``).
    Lines emitted before the exception persist."""
    lines = []
    err = None
    try:
        _poll(lines.append, node_name, gpu_csv, apps_csv, ps_text, stats_text,
              top, inspect, smi_list)
    except Exception as exc:
        err = type(exc).__name__
    return lines, err


def body_of(lines) -> bytes:
    """The sink format shared by the oracle and the library: every line + ``\\n``."""
    return "".join(line + "\n" for line in lines).encode()
