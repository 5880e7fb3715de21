"""Hand-built and generated poll inputs shared by the CPU and GPU parity tests.

Each case is a dict with the eight inputs of one reference poll (see
``oracle/poll_oracle.poll``).  ``tests/golden/make_golden.py`` runs the unmodified
reference on them and freezes its output into ``tests/golden/polls.json.gz``.
"""
from __future__ import annotations

import json

from kgm import replay

HDR_G = replay.GPU_HEADER
HDR_A = replay.APPS_HEADER


def inspect_json(labels, indent=4):
    return json.dumps([{"Id": "deadbeef", "Config": {"Image": "x", "Labels": labels}}],
                      indent=indent)


def _docker(containers):
    """containers: list of (cid, pids:list[str], labels|None|raw str, visible:list[str])."""
    ps = "".join("'%s,img:%d,0.0.0.0:80->80/tcp'\n" % (c[0], i) for i, c in enumerate(containers))
    stats = "".join("'%s,1MiB / 2GiB,0.5%%'\n" % c[0] for c in containers)
    top = {c[0]: "PID\n" + "".join(p + "\n" for p in c[1]) for c in containers}
    insp = {c[0]: (c[2] if isinstance(c[2], str) else inspect_json(c[2])) for c in containers}
    smi = {c[0]: "".join(u + "\n" for u in c[3]) for c in containers}
    return dict(ps_text=ps, stats_text=stats, top=top, inspect=insp, smi_list=smi)


def case(name, node_name, gpu_rows, app_rows, containers, **over):
    d = dict(name=name, node_name=node_name,
             gpu_csv=HDR_G + "".join(r + "\n" for r in gpu_rows),
             apps_csv=HDR_A + "".join(r + "\n" for r in app_rows))
    d.update(_docker(containers))
    d.update(over)
    return d


ML = {"mlgpu": "mlgpu"}


def k8s(pod, ns, gated=True, **extra):
    d = {"io.kubernetes.pod.name": pod, "io.kubernetes.pod.namespace": ns}
    if gated:
        d["mlgpu"] = "mlgpu"
    d.update(extra)
    return d


G_A = "GPU-aaaa, 49 %, 28 %, 16160 MiB, 15524 MiB, 636 MiB"
G_B = "GPU-bbbb, 0 %, 0 %, 16160 MiB, 0 MiB, 16160 MiB"
P_1 = "34443, /opt/conda/envs/py2/bin/python, GPU-aaaa, 15524 MiB"
P_2 = "71169, python3 train.py, GPU-aaaa, 419 MiB"
C_1 = ("16ee8a182726", ["34443", "71169", "1"],
       k8s("turing-gpu-notebook-alice-bob-5d9f-xk2p", "10007089-name"), [])
C_2 = ("76a082ecc41d", ["5"], k8s("jupyter-q-0", "ns2"), ["GPU-bbbb", "GPU-zzzz"])
C_3 = ("ffffffffffff", [], {}, [])


def hand_cases():
    out = []
    add = out.append
    # SURVEY.md Appendix B
    add(case("appendix_b", "node-a", [G_A, G_B], [P_1, P_2], [C_1, C_2, C_3]))
    add(case("node_name_unset", None, [G_A, G_B], [P_1, P_2], [C_1, C_2, C_3]))
    # mem_util repr pins (used,total)
    pairs = [(1, 183359), (183359, 183359), (1, 3), (7, 9), (61120, 81920), (1, 1000001),
             (12345, 16160), (2, 3), (13, 81920), (0, 5), (5, 1), (123456789, 7),
             (1, 10000000), (1, 100000), (3, 1000), (99999999999, 3), (1, 99999999999),
             (123456789012345678, 1), (1, 123456789012345678), (17, 1024), (100, 3),
             (200, 3), (1, 7), (22, 7), (1844674407370955161, 3), (9007199254740993, 1)]
    add(case("mem_util_reprs", "n",
             ["GPU-%04d, %d %%, 1 %%, %d MiB, %d MiB, 7 MiB" % (i, i % 101, t, u)
              for i, (u, t) in enumerate(pairs)], [], []))
    # digit-strip decode
    add(case("digit_strip", "n",
             ["GPU-x, 1.5 %, 28 %, 1e3 MiB, -9 MiB, 0012 MiB",
              "GPU-y, [N/A], [N/A], 16160 MiB, 007 MiB, [N/A]",
              "GPU-z,49,28,100,50,50"],
             ["11, [Not Found], GPU-x, [N/A]", "12, a b  c , GPU-y, 1,5 MiB", "13,x,GPU-z,0042MiB"],
             [("c1", ["11", "12", "13"], k8s("app-u1-0", "ns"), [])]))
    # duplicate keys: last values, first position
    add(case("dup_keys", "n",
             [G_A, G_B, "GPU-aaaa, 7 %, 1 %, 100 MiB, 50 MiB, 50 MiB"],
             [P_1, "5, other, GPU-bbbb, 1 MiB", "34443, renamed proc, GPU-aaaa, 77 MiB",
              "34443, on b, GPU-bbbb, 3 MiB"],
             [("c1", ["34443", "5"], k8s("app-u1-0", "ns"), [])]))
    # pid in two containers -> a line under each; pid string compare is exact
    add(case("pid_twice", "n", [G_A], [P_1, P_2],
             [("c1", ["34443"], k8s("a-b-c", "ns1"), []),
              ("c2", [" 34443", "71169", "34443"], k8s("d-e-f", "ns2"), []),
              ("c3", ["034443", "3444", "344430"], k8s("g-h-i", "ns3"), ["GPU-aaaa"])]))
    # gate variants, containers without k8s labels
    add(case("gates", "n", [G_A, G_B], [P_1, P_2, "9, z, GPU-bbbb, 4 MiB"],
             [("c1", ["34443"], {"mlgpu": "no"}, ["GPU-aaaa"]),
              ("c2", ["71169"], dict(ML), []),
              ("c3", ["9"], k8s("x-y", "ns", gated=False), []),
              ("c4", [], dict(ML), ["GPU-bbbb", "GPU-aaaa", "GPU-aaaa"]),
              ("c5", [], dict(ML), []),
              ("c6", ["9"], k8s("turing-gpu-notebook-bob", "ns"), []),
              ("c7", [], k8s("turing-gpu-notebook-a-b-c-d-e", "n-s"), ["GPU-bbbb"])]))
    # spaces are stripped from the inspect JSON before parsing
    add(case("label_spaces", "n", [G_A], [P_1],
             [("c1", ["34443"], k8s("my pod-u 1-x", "name space"), [])]))
    # empty tables
    add(case("no_gpus", "n", [], [], [C_2]))
    add(case("no_procs_no_containers", "n", [G_A, G_B], [], []))
    add(case("only_headers_no_newline", "n", [], [], [], gpu_csv="header", apps_csv="h"))
    add(case("empty_texts", "n", [], [], [C_2], gpu_csv="", apps_csv=""))
    add(case("no_final_newline", "n", [], [], [C_1],
             gpu_csv=HDR_G + G_A + "\n" + G_B, apps_csv=HDR_A + P_1 + "\n" + P_2))
    add(case("crlf", "n", [], [], [C_1],
             gpu_csv=HDR_G.replace("\n", "\r\n") + G_A + "\r\n" + G_B + "\r\n",
             apps_csv=HDR_A + P_1 + "\r\n" + P_2 + "\r\n"))
    add(case("extra_cells", "n", [G_A + ", extra, more"], [P_1 + ", x, y"], [C_1]))
    add(case("big_ids", "n",
             ["GPU-%02d, %d %%, 0 %%, 100 MiB, %d MiB, 1 MiB" % (i, i, i) for i in range(13)],
             ["%d, p%d, GPU-%02d, %d MiB" % (100 + i, i, i % 13, i) for i in range(40)],
             [("c%d" % j, [str(100 + i) for i in range(40) if i % 5 == j],
               k8s("pod-user%d-x" % j, "ns%d" % j), []) for j in range(5)]))
    # ---- inputs on which the reference raises ----
    add(case("err_blank_gpu_line", "n", [G_A, "", G_B], [P_1], [C_1]))
    add(case("err_blank_apps_line", "n", [G_A], [P_1, "", P_2], [C_1]))
    add(case("err_trailing_blank", "n", [G_A, G_B, ""], [P_1], [C_1]))
    add(case("err_gpu_one_cell", "n", [G_A, "GPU-bbbb"], [P_1], [C_1]))
    add(case("err_gpu_two_cells", "n", [G_A, "GPU-bbbb, 5 %", G_B.replace("bbbb", "cccc")],
             [P_1], [C_1]))
    add(case("err_gpu_three_cells", "n", [G_A, "GPU-bbbb, 5 %, 6 %"], [P_1], [C_1]))
    add(case("err_gpu_four_cells", "n", [G_A, "GPU-bbbb, 5 %, 6 %, 100 MiB"], [P_1], [C_1]))
    add(case("err_gpu_five_cells", "n", [G_A, "GPU-bbbb, 5 %, 6 %, 100 MiB, 50 MiB"], [P_1], [C_1]))
    add(case("ok_short_row_overridden", "n", [G_A, "GPU-bbbb, 5 %", G_B], [P_1], [C_1]))
    add(case("err_mem_total_na", "n", [G_A, "GPU-bbbb, 5 %, 6 %, [N/A], 50 MiB, 1 MiB"], [P_1], [C_1]))
    add(case("err_mem_used_na", "n", [G_A, "GPU-bbbb, 5 %, 6 %, 100 MiB, [N/A], 1 MiB"], [P_1], [C_1]))
    add(case("err_zero_total", "n", ["GPU-bbbb, 5 %, 6 %, 0 MiB, 50 MiB, 1 MiB", G_A], [P_1], [C_1]))
    add(case("err_unknown_uuid", "n", [G_A, G_B], [P_1, "77, mig proc, MIG-1234, 5 MiB"], [C_1]))
    add(case("err_comma_in_procname", "n", [G_A], ["34443, python a,b, GPU-aaaa, 5 MiB"], [C_1]))
    add(case("err_apps_two_cells", "n", [G_A], [P_1, "5, name"], [C_1]))
    add(case("err_apps_three_cells_matched", "n", [G_A], [P_1, "71169, name, GPU-aaaa", P_2.replace("71169", "8")],
             [("c1", ["34443", "71169", "8"], k8s("a-b-c", "ns"), [])]))
    add(case("ok_apps_three_cells_unmatched", "n", [G_A], [P_1, "5, name, GPU-aaaa"], [C_1]))
    add(case("err_stats_not_in_ps", "n", [G_A], [P_1], [C_1, C_2],
             ps_text="'16ee8a182726,img,'\n"))
    add(case("err_labels_null", "n", [G_A, G_B], [P_1],
             [C_1, ("cnull", [], None, []), C_2]))
    add(case("err_pod_no_dash_busy", "n", [G_A], [P_1, P_2],
             [C_2, ("c9", ["34443"], k8s("nodash", "ns"), []), C_1]))
    add(case("err_pod_no_dash_idle", "n", [G_A], [P_1],
             [("c9", [], k8s("nodash", "ns"), ["GPU-aaaa"]), C_1]))
    add(case("ok_pod_no_dash_silent", "n", [G_A], [P_1],
             [("c9", [], k8s("nodash", "ns"), []), ("c8", ["1"], k8s("nodash", "ns", gated=False), []), C_1]))
    add(case("err_no_namespace_label", "n", [G_A], [P_1],
             [("c9", ["34443"], {"mlgpu": "mlgpu", "io.kubernetes.pod.name": "a-b"}, [])]))
    add(case("err_bad_json", "n", [G_A], [P_1], [C_1, ("cbad", [], "not json", [])]))
    add(case("err_no_config", "n", [G_A], [P_1], [("cbad", [], '[{"Id": 1}]', []), C_1]))
    add(case("err_quoted_first_cell", "n", ['"GPU-aaaa, 49 %', G_B], [P_1], [C_1]))
    add(case("ok_quote_inside", "n", [G_A], ['34443, say "hi" there, GPU-aaaa, 5 MiB'], [C_1]))
    add(case("ok_utf8_process_name", "n", [G_A], ["34443, /home/ren\u00e9/\u6a21\u578b/py thon, GPU-aaaa, 15524 MiB", P_2], [C_1]))
    add(case("quoted_first_cell_closed", "n", ['"GPU-aaaa", 49 %, 28 %, 16160 MiB, 15524 MiB, 636 MiB', G_B], [P_1], [C_1]))
    add(case("mem_20_digits", "n", ["GPU-aaaa, 49 %, 28 %, 16160 MiB, 12345678901234567890 MiB, 636 MiB"], [P_1], [C_1]))
    add(case("err_stray_cr", "n", ["GPU-aaaa, 4\r9 %, 28 %, 16160 MiB, 15524 MiB, 636 MiB"], [P_1], [C_1]))
    add(case("tabs_and_ws", "n", ["\t GPU-aaaa,\t49 %, 28 %, 16160 MiB , 15524 MiB, 636 MiB "],
             ["34443,\tpy thon\t, \x0b GPU-aaaa, 1 MiB"], [C_1]))
    return out


def replay_cases():
    """Generated documents: the config-1 shape plus idle / ungated / odd-size variants."""
    out = []
    shapes = [
        ("cfg1", dict(gpus=8, containers=8, procs_per_container=4), 6),
        ("idle", dict(gpus=8, containers=12, procs_per_container=2, idle_every=3, ungated_every=4), 6),
        ("wide", dict(gpus=16, containers=40, procs_per_container=1, idle_every=7), 4),
        ("tiny", dict(gpus=1, containers=1, procs_per_container=1), 4),
        ("big", dict(gpus=4, containers=3, procs_per_container=30), 3),
    ]
    for tag, kw, nticks in shapes:
        rp = replay.make_replay(seed=7, nodes=2, ticks=nticks, **kw)
        for d in range(rp.n_docs):
            snap = rp.nodes[rp.doc_node[d]]
            out.append(dict(name="replay_%s_n%d_t%d" % (tag, rp.doc_node[d], rp.doc_tick[d]),
                            node_name=snap.name, gpu_csv=rp.gpu_texts[d],
                            apps_csv=rp.apps_texts[d], ps_text=snap.ps_text,
                            stats_text=snap.stats_text, top=snap.top, inspect=snap.inspect,
                            smi_list=snap.smi_list))
    return out


def all_cases():
    return hand_cases() + replay_cases()


INPUT_KEYS = ("node_name", "gpu_csv", "apps_csv", "ps_text", "stats_text", "top", "inspect",
              "smi_list")


def inputs_of(c):
    return tuple(c[k] for k in INPUT_KEYS)
